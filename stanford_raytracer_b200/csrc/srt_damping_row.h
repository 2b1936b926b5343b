// srt_damping_row.h -- the per-row arithmetic of the damping kernels (srt_damping.cu), written once as host/device code:
// nvcc compiles it into the kernels; g++ compiles the same text into tests/csrc/damping_probe.cpp, which runs a row
// sequentially on the host so that the logic can be checked against the reference binary without a GPU and the part of a
// GPU-vs-reference difference that comes from the two math libraries can be told from a difference of algorithm.
#pragma once
#include <float.h>
#include <math.h>
#include <stddef.h>
#include <string.h>
#if defined(__CUDACC__)
#define DMP_HD __host__ __device__
#else
#define DMP_HD
#endif

// damping/include/consts.h
#define DMP_PI 3.14159265358979311599796346854418516159
#define DMP_Q_EL 1.602E-19
#define DMP_M_EL 9.1E-31
#define DMP_EPS0 8.854E-12
#define DMP_C 2.997956376932163e+08
#define DMP_R_E 6378000.0

struct DampSetup {
  double glx[25], glw[25];  // Gauss-Legendre nodes / weights for n = 50 as gauss_legendre_tbl builds them
  double sm2mag[9];         // SM -> MAG rotation of the run's date and time (row-major)
  double greenwich_lon_deg; // longitude of (lat 51.48, lon 0) in MAG, degrees (MLT reference)
  double ut_hr;             // the reference's "fractional hours": ms / 1000 / 60
  double Kp, AE_target;
  int include_geom, nspec;
  double qs[4], ms[4];
  int nL, nM, nE;           // CRRES table: bins and energy channels of the AE level in use
  double logE[4];
};

struct PsdState {  // psd_model after set_params
  double L_sh, L_pp, n_fit, An_fit;
};

#if defined(DMP_TEST_NOISE)
// test hook (tests/csrc/damping_probe.cpp only): a factor 1 +- amp, the sign a hash of the arguments -- a stand-in for
// the last-place differences between two math libraries, to measure how well a row's result is determined at all
extern double dmp_test_noise_amp;
extern unsigned long long dmp_test_noise_seed;
inline double dmp_noise(double a, double b) {
  unsigned long long x, y;
  memcpy(&x, &a, 8); memcpy(&y, &b, 8);
  unsigned long long z = (x * 0x9e3779b97f4a7c15ull) ^ (y + dmp_test_noise_seed);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull; z = (z ^ (z >> 27)) * 0x94d049bb133111ebull; z ^= z >> 31;
  return 1.0 + ((z & 1ull) ? dmp_test_noise_amp : -dmp_test_noise_amp);
}
#define DMP_NOISE(f, a, b) ((f) * dmp_noise(a, b))
#else
#define DMP_NOISE(f, a, b) (f)
#endif
// the transcendentals of the density model: with the test hook, each result carries a last-place error of its own, which is
// what separates two math libraries -- note that the suprathermal density a/v^4 - b/v^5 + c/v^6 cancels around
// v = b/a = 1.7e7 m/s and the blended density is exp(mean of logs), so an error of one place in a pow or log is many
// places of the density there
#define DMP_POW(x, y) DMP_NOISE(pow((x), (y)), (x), (y))
#define DMP_LOG(x) DMP_NOISE(log(x), (x), 1.0)
#define DMP_EXP(x) DMP_NOISE(exp(x), (x), 2.0)

DMP_HD inline double psd_suprathermal(double vperp, double vpar) {  // psd_model.cpp:259-278
  const double a = 4.9e5, b = 8.3e14, c = 5.4e23, v0 = 1;
  double v = 100. * sqrt(vperp * vperp + vpar * vpar + v0);
  double f = (a / DMP_POW(v, 4.0) - b / DMP_POW(v, 5.0) + c / DMP_POW(v, 6.0));
  return f * pow(100., 6.0);
}
DMP_HD inline double psd_crres(const PsdState& s, double vperp, double vpar) {  // :281-305
  const double v0 = 1;
  double v = 100. * sqrt(vperp * vperp + vpar * vpar + v0);
  double f = s.An_fit / DMP_POW(v, s.n_fit);
  return f * pow(100., 6.0);
}
DMP_HD inline double psd_hybrid(const PsdState& s, double vperp, double vpar) {  // :308-341
  if (s.L_pp - s.L_sh > 1) return psd_suprathermal(vperp, vpar);
  if (s.L_sh - s.L_pp > 1) return psd_crres(s, vperp, vpar);
  double f_polar = psd_suprathermal(vperp, vpar);
  double f_crres = psd_crres(s, vperp, vpar);
  double w_polar = exp(5 * (s.L_pp - s.L_sh)) / (1 + exp(5 * (s.L_pp - s.L_sh)));
  double w_crres = exp(5 * (s.L_sh - s.L_pp)) / (1 + exp(5 * (s.L_sh - s.L_pp)));
  return DMP_EXP((DMP_LOG(f_polar) * w_polar + DMP_LOG(f_crres) * w_crres) / (w_polar + w_crres));
}

DMP_HD inline double bessel_jn(int n, double x) {
  // jn(n, x) for n in -2..2: J_{-n} = (-1)^n J_n
  int an = n < 0 ? -n : n;
  double v = an == 0 ? j0(x) : (an == 1 ? j1(x) : jn(2, x));
  return (n < 0 && (an & 1)) ? -v : v;
}

struct Integrand {  // integrand::initialize, integrand.cpp:7-45
  double kperp, kpar, w, wch, n_sq, ct, st, st_sq, R, L, P, S;
};

// integrand::evaluate_t (integrand.cpp:48-95)
DMP_HD inline double integrand_t(const Integrand& q, const PsdState& psd, double t) {
  const double eps = DBL_EPSILON, DEL = 1e-8;
  const double vnorm = (1 - t + eps) / (t + eps);
  const double vperp = vnorm * DMP_C;
  double suminteg = 0;
  // the density gradient with respect to v_perp does not depend on the resonance only through vpar: per m below
  for (int m = -1; m <= 1; m++) {
    const double arg = q.kperp * vperp / q.wch;
    const double Jm = bessel_jn(m, arg);
    const double Jm_m1 = bessel_jn(m + 1, arg);  // the reference's names (swapped)
    const double Jm_p1 = bessel_jn(m - 1, arg);
    const double vpar = (q.w - m * q.wch) / q.kpar;
    // fG1 / fG2 (integrand.cpp:99-160) evaluate the same two central differences
    double d = DEL * fabs(vperp);
    if (d < 10 * eps) d = 10 * eps;
    const double df_dvperp = (psd_hybrid(psd, vperp + d, vpar) - psd_hybrid(psd, vperp - d, vpar)) / (2 * d);
    d = DEL * fabs(vpar);
    if (d < 10 * eps) d = 10 * eps;
    const double df_dvpar = (psd_hybrid(psd, vperp, vpar + d) - psd_hybrid(psd, vperp, vpar - d)) / (2 * d);
    const double G1 = df_dvperp - (q.kpar / q.w) * (vpar * df_dvperp - vperp * df_dvpar);
    const double G2 = Jm * (df_dvpar - ((double)m * q.wch + eps) / (q.w * vperp + eps) * (vpar * df_dvperp - vperp * df_dvpar));
    const double dJ = Jm_p1 - Jm_m1;
    suminteg += (G1 * ((q.P - q.n_sq * q.st_sq) * (2 * (q.L - q.n_sq) * vperp * Jm_p1 * Jm_p1 + 2 * vperp * (q.R - q.n_sq) * Jm_m1 * Jm_m1 +
                                                   q.n_sq * q.st_sq * vperp * (dJ * dJ)) -
                       q.n_sq * q.ct * q.st * (2 * vpar * Jm * (Jm_p1 * (q.R - q.n_sq) + Jm_m1 * (q.L - q.n_sq)) + q.n_sq * q.ct * q.st * vperp * (dJ * dJ))) +
                 G2 * (4 * vpar * Jm * ((q.L - q.n_sq) * (q.R - q.n_sq) + q.n_sq * q.st_sq * (q.S - q.n_sq)) -
                       2 * q.n_sq * q.ct * q.st * ((q.R - q.n_sq) * vperp * Jm_m1 + (q.L - q.n_sq) * vperp * Jm_p1)));
  }
  const double ret = -2 * DMP_PI * DMP_PI * ((DMP_Q_EL * DMP_Q_EL / DMP_M_EL / DMP_EPS0) / (q.w * fabs(q.kpar))) * suminteg * vperp;
  return ((1 + eps) / (t * t + eps)) * (DMP_C * ret);
}

DMP_HD inline double bulge(double akp, double amlt) {  // bulge.cpp:8-83
  const double ahour_rad = 0.26179939, ahrrad = 2.6179939e-1;
  double centroid = 47.0 / (akp + 3.9) + 11.3;
  double x = amlt - centroid;
  if (x < -12.0) x += 24.0;
  if (x > 12.0) x -= 24.0;
  double absx = fabs(x) * ahrrad;
  double along = amlt * ahour_rad + 1.5707963;
  double salong = sin(along);
  double b1 = 0.043 * salong - 0.4589;
  double b2 = -0.361 * salong + 5.7464;
  return (b1 * akp + b2) * (1.0 + exp(-1.5 * absx * absx + 0.08 * absx - 0.7));
}

// CRRES_fit_params (psd_model.cpp:113-256): nearest MLT and L bins, log J_perp of the energy channels of the AE level
// in use, straight-line least squares in log-log (the reference solves it by Householder QR: the same minimiser)
DMP_HD inline void crres_fit(const DampSetup& c_damp, const double* tab, double L, double MLT, double& n_fit, double& An_fit) {
  const int nL = c_damp.nL, nM = c_damp.nM, nE = c_damp.nE;
  const double* Lv = tab;
  const double* Mv = tab + nL;
  const double* J = tab + nL + nM;  // [nE][nL * nM], column-major per channel
  int row = 0, col = 0;
  double best = fabs(Mv[0] - MLT);
  for (int i = 1; i < nM; ++i) { double d = fabs(Mv[i] - MLT); if (d < best) { best = d; row = i; } }
  best = fabs(Lv[0] - L);
  for (int i = 1; i < nL; ++i) { double d = fabs(Lv[i] - L); if (d < best) { best = d; col = i; } }
  double sx = 0, sy = 0, sxx = 0, sxy = 0;
  for (int e = 0; e < nE; ++e) {
    double y = log(J[(size_t)e * nL * nM + col + (size_t)nL * row]);
    double x = c_damp.logE[e];
    sx += x; sy += y; sxx += x * x; sxy += x * y;
  }
  const double xm = sx / nE, ym = sy / nE;
  double sdd = 0, sdy = 0;
  for (int e = 0; e < nE; ++e) {
    double y = log(J[(size_t)e * nL * nM + col + (size_t)nL * row]);
    double dx = c_damp.logE[e] - xm;
    sdd += dx * dx;
    sdy += dx * (y - ym);
  }
  const double p1 = sdy / sdd, p0 = ym - p1 * xm;
  const double J0 = exp(p0), m = -1.0 * p1;
  n_fit = 2 * m + 2;
  An_fit = 2 * J0 / pow((0.5 * (6.25e11) * DMP_M_EL), m - 1);
}

// Everything of damping_foust.cpp:97-205 that precedes the quadrature, for the row r (20 doubles; rp = the previous row of
// the ray): fills the integrand and density state, returns the quantities the row's factor needs afterwards.
struct DampRow {
  Integrand q;
  PsdState psd;
  double w, n, a, b, k[3], kmag, vg[3], dist, geom;
};
DMP_HD inline void damp_row_setup(const DampSetup& S, const double* crres, const double* r, const double* rp, double w, DampRow& o) {
  const double pos[3] = {r[1], r[2], r[3]};
  double pm[3];
  for (int i = 0; i < 3; ++i) pm[i] = S.sm2mag[3 * i] * pos[0] + S.sm2mag[3 * i + 1] * pos[1] + S.sm2mag[3 * i + 2] * pos[2];
  // cart_to_pol_d (Cart_to_pol.f95)
  const double rho = sqrt(pm[0] * pm[0] + pm[1] * pm[1]);
  const double rad = sqrt(rho * rho + pm[2] * pm[2]);
  const double lat_m = atan2(pm[2], rho);
  double lon_m = atan2(pm[1], pm[0]);
  lon_m = lon_m + (1.0 - copysign(1.0, lon_m)) * 3.141592653589793238462643;
  o.geom = rad * cos(lat_m) / DMP_R_E;
  o.w = w;
  if (!rp) return;
  const double B0[3] = {r[13], r[14], r[15]}, nv[3] = {r[10], r[11], r[12]};
  o.vg[0] = r[7]; o.vg[1] = r[8]; o.vg[2] = r[9];
  // local L-shell (damping_foust.cpp:104-107)
  const double lat = atan(pos[2] / sqrt(pos[0] * pos[0] + pos[1] * pos[1]));
  const double rr = sqrt(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
  const double cl = cos(lat);
  o.psd.L_sh = rr / (cl * cl) / DMP_R_E;
  // MLT (coord_transforms.cpp:195-210; lon in radians against a longitude in degrees, as the reference does)
  const double mlt = fmod(S.ut_hr + (lon_m - S.greenwich_lon_deg) / 15.0, 24.0);
  o.psd.L_pp = bulge(S.Kp, mlt);
  crres_fit(S, crres, o.psd.L_sh, mlt, o.psd.n_fit, o.psd.An_fit);
  const double B0mag = sqrt(B0[0] * B0[0] + B0[1] * B0[1] + B0[2] * B0[2]);
  Integrand& q = o.q;
  q.wch = DMP_Q_EL * B0mag / DMP_M_EL;
  q.w = w;
  for (int i = 0; i < 3; ++i) o.k[i] = nv[i] * w / DMP_C;
  o.kmag = sqrt(o.k[0] * o.k[0] + o.k[1] * o.k[1] + o.k[2] * o.k[2]);
  const double Bh[3] = {B0[0] / B0mag, B0[1] / B0mag, B0[2] / B0mag};
  q.kpar = o.k[0] * Bh[0] + o.k[1] * Bh[1] + o.k[2] * Bh[2];
  const double kp[3] = {o.k[0] - q.kpar * Bh[0], o.k[1] - q.kpar * Bh[1], o.k[2] - q.kpar * Bh[2]};
  q.kperp = sqrt(kp[0] * kp[0] + kp[1] * kp[1] + kp[2] * kp[2]);
  const double theta = atan2(q.kperp, q.kpar);
  q.st = sin(theta);
  q.ct = cos(theta);
  q.st_sq = q.st * q.st;
  const double sin_th_sq = q.st * q.st, cos_th_sq = q.ct * q.ct;
  o.n = sqrt(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]);
  q.n_sq = o.n * o.n;
  double R = 1., L = 1., P = 1.;
  for (int j = 0; j < S.nspec; ++j) {
    const double wps2 = r[16 + j] * (S.qs[j] * S.qs[j]) / (S.ms[j] * DMP_EPS0);
    const double whs = S.qs[j] * B0mag / S.ms[j];
    R -= wps2 / (w * (w + whs));
    L -= wps2 / (w * (w - whs));
    P -= wps2 / (w * w);
  }
  const double Sx = (R + L) / 2.;
  o.a = Sx * sin_th_sq + P * cos_th_sq;
  o.b = R * L * sin_th_sq + P * Sx * (1 + cos_th_sq);
  q.R = R; q.L = L; q.P = P; q.S = Sx;
  const double dp[3] = {pos[0] - rp[1], pos[1] - rp[2], pos[2] - rp[3]};
  o.dist = sqrt(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
}
// damping_foust.cpp:208-229 after the quadrature: the row's factor exp(-dist * ki_along_vg * 2)
DMP_HD inline double damp_row_factor(const DampRow& o, double Di) {
  const double ki = -(o.w / DMP_C) * (0.5) * (1. / (4. * o.n * (2. * o.a * o.n * o.n - o.b))) * Di;
  const double vgn = sqrt(o.vg[0] * o.vg[0] + o.vg[1] * o.vg[1] + o.vg[2] * o.vg[2]);
  const double ki_along_vg = ki * ((o.k[0] * o.vg[0] + o.k[1] * o.vg[1] + o.k[2] * o.vg[2]) / (o.kmag * vgn));
  return exp(-o.dist * ki_along_vg * 2.0);
}
