// srt_damping_setup.h -- host-side (run once per call) staging for the damping kernels (srt_damping.cu):
//   * the SM -> MAG rotation of the run's date and time as one 3x3 matrix, and the geomagnetic longitude of Greenwich
//     (what the reference evaluates per row through the xform_double chain: SM_TO_MAG.f95 = T4^-1, T3^-1, T2^-1, T1,
//     T5; GEO_TO_MAG.f95 = T5; coord_transforms.cpp:195-210) -- functions of time only;
//   * the 50-node Gauss-Legendre table as the reference builds it at run time (n = 50 is not among the tabulated
//     orders of damping/src/gauss_legendre.c:182-210, so gauss_legendre_tbl :436-538 runs with eps = 1e-10).
// Plain C++ with libm: floating-point parity with the reference is by tolerance here (see srt_damping.cu).
#pragma once
#include <float.h>
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "srt_damping_row.h"

namespace srt_host {

namespace dmp {
const double PI_ = 3.141592653589793238462643;
const double DEGRAD = PI_ / 180.0;
// rotation about axis 0 / 1 / 2 by `a` in the convention of Rotate_{x,y,z}.f95
inline void rotate(int axis, double a, const double x[3], double o[3]) {
  const double c = cos(a), s = sin(a);
  const int i = (axis + 1) % 3, j = (axis + 2) % 3;
  o[axis] = x[axis];
  if (axis == 1) {  // y: x' = x c + z s ; z' = z c - x s
    o[0] = x[0] * c + x[2] * s;
    o[2] = x[2] * c - x[0] * s;
  } else {          // x: y' = y c + z s ; z' = z c - y s     z: x' = x c + y s ; y' = y c - x s
    o[i] = x[i] * c + x[j] * s;
    o[j] = x[j] * c - x[i] * s;
  }
}
struct TimeAngles {
  double theta;            // T1: Greenwich sidereal angle
  double epsilon, lamdas;  // T2: obliquity, ecliptic longitude of the Sun
  double phi, lamda;       // T5: latitude / longitude of the dipole axis
  double psi, mu;          // T3, T4
};
inline TimeAngles time_angles(const int itime[2]) {
  TimeAngles t;
  const int iyr = itime[0] / 1000, iday = itime[0] - iyr * 1000;
  const double ut = itime[1] / 3600000.0;
  const double rmjd = 45.0 + (double)(iyr - 1859) * 365.0 + ((double)((iyr - 1861) / 4) + 1.0) + (double)iday - 1.0 + ut / 24.0;
  const double t0 = (rmjd - 51544.5) / 36525.0;  // T0.f95
  t.theta = (100.461 + 36000.770 * t0 + 15.04107 * ut) * DEGRAD;  // T1.f95
  t.epsilon = (23.439 - 0.013 * t0) * DEGRAD;                     // T2.f95
  const double m = (357.528 + 35999.05 * t0 + 0.04107 * ut) * DEGRAD;
  const double cgamma = 280.46 + 36000.772 * t0 + 0.04107 * ut;
  t.lamdas = (cgamma + (1.915 - 0.0048 * t0) * sin(m) + 0.02 * sin(2.0 * m)) * DEGRAD;
  const double factor = (rmjd - 46066.0) / 365.25;  // T5.f95 / Get_q_c.f95
  t.phi = (78.8 + 4.283e-2 * factor) * DEGRAD;
  t.lamda = (289.1 - 1.413e-2 * factor) * DEGRAD;
  // Get_q_c.f95: the dipole axis in GSE
  double q_g[3] = {cos(t.phi) * cos(t.lamda), cos(t.phi) * sin(t.lamda), sin(t.phi)}, a[3], b[3], q_c[3];
  rotate(2, -t.theta, q_g, a);    // T1 inverse
  rotate(0, t.epsilon, a, b);     // T2 forward
  rotate(2, t.lamdas, b, q_c);
  t.psi = (q_c[2] == 0.0) ? -copysign(PI_ / 2.0, q_c[1]) : -atan(q_c[1] / q_c[2]);  // T3.f95
  t.mu = -atan(q_c[0] / sqrt(q_c[1] * q_c[1] + q_c[2] * q_c[2]));                    // T4.f95
  return t;
}
inline void t5_forward(const TimeAngles& t, const double x[3], double o[3]) {  // GEO -> MAG
  double a[3];
  rotate(2, t.lamda, x, a);
  rotate(1, t.phi - PI_ / 2.0, a, o);
}
}  // namespace dmp

// x_mag = M x_sm: SM_TO_MAG.f95 applied to the unit vectors (M row-major)
inline void sm_to_mag_matrix(const int itime[2], double M[9]) {
  const dmp::TimeAngles t = dmp::time_angles(itime);
  for (int c = 0; c < 3; ++c) {
    double x[3] = {0, 0, 0}, a[3], b[3], d[3], e[3], f[3], o[3];
    x[c] = 1.0;
    dmp::rotate(1, -t.mu, x, a);        // T4 inverse
    dmp::rotate(0, -t.psi, a, b);       // T3 inverse
    dmp::rotate(2, -t.lamdas, b, d);    // T2 inverse: rotate_z(-lamdas) then rotate_x(-epsilon)
    dmp::rotate(0, -t.epsilon, d, e);
    dmp::rotate(2, t.theta, e, f);      // T1 forward
    dmp::t5_forward(t, f, o);           // T5 forward
    for (int r = 0; r < 3; ++r) M[3 * r + c] = o[r];
  }
}

// longitude [deg] in MAG of the point at geographic latitude 51.48, longitude 0 (coord_transforms.cpp:199-207)
inline double greenwich_mag_lon_deg(const int itime[2]) {
  const dmp::TimeAngles t = dmp::time_angles(itime);
  const double D2R = 3.14159265358979311599796346854418516159 / 180.0, R2D = 1.0 / D2R * 1.0;
  const double lat = D2R * 51.48, lon = D2R * 0.0;
  double a[3] = {1.0 * cos(lat) * cos(lon), 1.0 * cos(lat) * sin(lon), 1.0 * sin(lat)}, b[3];
  dmp::t5_forward(t, a, b);
  return (180.0 / 3.14159265358979311599796346854418516159) * atan2(b[1], b[0]);
}

// damping/src/gauss_legendre.c:436-538 for n < 1024: Newton iteration on the Legendre recurrence from the asymptotic
// initial guess, at least one step, until both the abscissa and the weight move by less than eps; x[0] is the
// smallest positive root.  ltbl[k] of the reference is 1 - 1/k.
inline void gauss_legendre_tbl(int n, double* x, double* w, double eps) {
  const double PI = 3.1415926535897932384626433832795028841971693993751;
  const int m = (n + 1) >> 1;
  const double t0 = (1.0 - (1.0 - 1.0 / (double)n) / (8.0 * (double)n * (double)n));
  const double t1 = 1.0 / (4.0 * (double)n + 2.0);
  for (int i = 1; i <= m; i++) {
    double x0 = cos(PI * (double)((i << 2) - 1) * t1) * t0, x1 = x0, w0 = 0.0, w1 = 0.0, dx = DBL_MAX, dw = DBL_MAX;
    int j = 0;
    do {
      double P_1 = 1.0, P0 = x0, P_2;
      for (int k = 2; k <= n; k++) {
        P_2 = P_1;
        P_1 = P0;
        const double t2 = x0 * P_1;
        P0 = t2 + (1.0 - 1.0 / (double)k) * (t2 - P_2);
      }
      const double dpdx = ((x0 * P0 - P_1) * (double)n) / (x0 * x0 - 1.0);
      x1 = x0 - P0 / dpdx;
      w1 = 2.0 / ((1.0 - x1 * x1) * dpdx * dpdx);
      if (j == 0) w0 = 2.0 / ((1.0 - x0 * x0) * dpdx * dpdx);
      dx = x0 - x1;
      dw = w0 - w1;
      x0 = x1;
      w0 = w1;
      j++;
    } while ((fabs(dx) > eps || fabs(dw) > eps) && j < 100);
    x[(m - 1) - (i - 1)] = x1;
    w[(m - 1) - (i - 1)] = w1;
  }
}

// The per-call constants of the kernels and the part of the CRRES table they read: Lvec[nL], MLTvec[nM], then
// J_int[nE][nL * nM] of the energy channels of the AE level in use, in file order (psd_model.cpp:158-215).
// Returns 0, -1 (table cannot be read), -2 (bad AE level / arguments).
inline int damp_setup(const char* crres_table, int nspec, const double* qs, const double* ms, double Kp, double AE_level, int yearday, int msec,
                      int include_geom_factor, DampSetup& S, std::vector<double>& tab, char* err, size_t errcap) {
  memset(&S, 0, sizeof S);
  FILE* f = fopen(crres_table, "rb");
  if (!f) { snprintf(err, errcap, "cannot open CRRES table %s", crres_table); return -1; }
  int hdr[4];
  if (fread(hdr, 4, 4, f) != 4 || hdr[0] != 0x53525443) { fclose(f); snprintf(err, errcap, "%s is not a CRRES table", crres_table); return -1; }
  const int n = hdr[1], nL = hdr[2], nM = hdr[3];
  const size_t per = 2 + (size_t)nL + nM + (size_t)nL * nM;
  std::vector<double> raw(per * n);
  const size_t got = fread(raw.data(), 8, raw.size(), f);
  fclose(f);
  if (got != raw.size()) { snprintf(err, errcap, "%s is truncated", crres_table); return -1; }
  gauss_legendre_tbl(50, S.glx, S.glw, 1e-10);
  int itime[2] = {yearday, msec};
  sm_to_mag_matrix(itime, S.sm2mag);
  S.greenwich_lon_deg = greenwich_mag_lon_deg(itime);
  S.ut_hr = msec / 1000.0 / 60.0;
  S.Kp = Kp;
  S.AE_target = round(AE_level);
  if (S.AE_target < 1 || S.AE_target > 3) { snprintf(err, errcap, "AE level must round to 1, 2 or 3"); return -2; }
  S.include_geom = include_geom_factor ? 1 : 0;
  S.nspec = nspec;
  for (int i = 0; i < nspec; ++i) { S.qs[i] = qs[i]; S.ms[i] = ms[i]; }
  S.nL = nL; S.nM = nM; S.nE = 0;
  std::vector<double> J;
  tab.clear();
  for (int k = 0; k < n; ++k) {
    const double* e = raw.data() + per * k;
    if (e[1] != S.AE_target) continue;
    if (S.nE == 0) tab.assign(e + 2, e + 2 + nL + nM);
    if (S.nE < 4) {
      S.logE[S.nE++] = log(e[0]);
      J.insert(J.end(), e + 2 + nL + nM, e + per);
    }
  }
  if (S.nE < 2) { snprintf(err, errcap, "CRRES table has no entries for AE level %g", S.AE_target); return -1; }
  tab.insert(tab.end(), J.begin(), J.end());
  return 0;
}

}  // namespace srt_host
