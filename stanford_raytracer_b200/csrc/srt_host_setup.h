// srt_host_setup.h -- host-side (run once) staging of an SrtModel.
//
// Replaces, on the host, what the reference does at start-up or needlessly on every
// funcPlasmaParams call:
//   * readinput            fortran/ngo_dens_model.f95:29-160   (Ngo config cards)
//   * SM_TO_GSM_d chain    xform_double/{T4,Get_q_c,T0,T1,T2,POL_TO_CART,Rotate_*}.f95
//                          -- a function of time only; the reference re-evaluates it
//                          2-3 times per plasma call
//   * interp setup         fortran/interp_dens_model_adapter.f95:52-134 (srt_grid_setup.h)
//   * RECALC_08            tsyganenko/geopack2008.for:486-1195 (srt_igrf_setup.h)
// Plain C++ (no CUDA types) so that g++ can build it for the CPU-side tests too.
#pragma once
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <vector>

#include "srt_listread.h"
#include "srt_physics.h"

namespace srt_host {

// ---- Fortran list-directed numeric input: srt_listread.h (one READ statement = one ListReader::read)

// ---- dipole tilt (xform_double chain), FP64 with the shared elementary functions
inline void rot_x(double a, const double v[3], double o[3]) {
  o[0] = v[0];
  o[1] = v[1] * srt_cos(a) + v[2] * srt_sin(a);
  o[2] = v[2] * srt_cos(a) - v[1] * srt_sin(a);
}
inline void rot_z(double a, const double v[3], double o[3]) {
  o[0] = v[0] * srt_cos(a) + v[1] * srt_sin(a);
  o[1] = v[1] * srt_cos(a) - v[0] * srt_sin(a);
  o[2] = v[2];
}
// mu of xform_double/T4.f95:13 for itime = (yyyyddd, ms of day)
inline double dipole_tilt(int yearday, int msec) {
  const double degrad = 3.141592653589793238462643 / 180.0;
  const int iyr = yearday / 1000, iday = yearday - iyr * 1000;
  const double ut = msec / 3600000.0;
  const double fracday = ut / 24.0;
  // Get_q_c.f95:19-25 (dfloat) and T0.f95:19-21 (float of small integers: exact) give the same rmjd
  const double rmjd = 45.0 + (double)(iyr - 1859) * 365.0 + ((double)((iyr - 1861) / 4) + 1.0) + (double)(iday)-1.0 + fracday;
  const double factor = (rmjd - 46066.0) / 365.25;
  const double phi = (78.8 + 4.283e-2 * factor) * degrad;
  const double lamda = (289.1 - 1.413e-2 * factor) * degrad;
  const double coslat = srt_cos(phi);
  double q_g[3] = {1.0 * coslat * srt_cos(lamda), 1.0 * coslat * srt_sin(lamda), 1.0 * srt_sin(phi)};  // POL_TO_CART.f95
  const double t0 = (rmjd - 51544.5) / 36525.0;                                                         // T0.f95:22
  // T1 inverse: rotate_z(-theta)
  const double theta = (100.461 + 36000.770 * t0 + 15.04107 * ut) * degrad;
  double tmp[3], tmp2[3], q_c[3];
  rot_z(-1.0 * theta, q_g, tmp);
  // T2 forward: rotate_x(epsilon) then rotate_z(lamdas)
  const double epsilon = (23.439 - 0.013 * t0) * degrad;
  const double m = (357.528 + 35999.05 * t0 + 0.04107 * ut) * degrad;
  const double cgamma = 280.46 + 36000.772 * t0 + 0.04107 * ut;
  const double lamdas = (cgamma + (1.915 - 0.0048 * t0) * srt_sin(m) + 0.02 * srt_sin(2.0 * m)) * degrad;
  rot_x(epsilon, tmp, tmp2);
  rot_z(lamdas, tmp2, q_c);
  return -srt_atan(q_c[0] / srt_sqrt(q_c[1] * q_c[1] + q_c[2] * q_c[2]));
}

inline void set_species_invariants(SrtModel& M) {
  double maxq = srt_fabs(M.qs[0]), minm = M.ms[0];
  for (int i = 0; i < SRT_NSPEC; ++i) {
    M.q2[i] = M.qs[i] * M.qs[i];
    if (i < M.nspec) {
      if (srt_fabs(M.qs[i]) > maxq) maxq = srt_fabs(M.qs[i]);
      if (M.ms[i] < minm) minm = M.ms[i];
    }
  }
  M.fs_q2 = maxq * maxq;
  M.fs_den = minm * SRT_EPS0;
}

inline void set_time(SrtModel& M, int yearday, int msec) {
  const double mu = dipole_tilt(yearday, msec);
  srt_sincos(mu, M.sin_mu, M.cos_mu);
}

// ---- Ngo config cards -> SrtNgo.  Returns 0, or a negative error:
//   -1 cannot open, -2 truncated/garbled cards, -3 too many ducts
inline int load_ngo_config(const char* path, SrtModel& M) {
  ListReader rd;
  if (!rd.open(path)) return -1;
  SrtNgo& g = M.ngo;
  memset(&g, 0, sizeof g);
  // default-real literals widened to double (Q14)
  g.pi_f = (double)3.141592653589793f;
  g.r0 = (double)6370.f;
  const double radgra = 180. / g.pi_f;
  const double grarad = 1. / radgra;
  double v[16];
  if (rd.read(4, v) != 4) return -2;  // INTERA NUMRES NSUPPR SPELAT
  double latitu = 0.0;
  for (;;) {  // satellite cards until DISTRE <= -1
    if (rd.read(2, v) != 2) return -2;
    latitu = v[1];
    if (v[0] <= -1.) break;
  }
  if (rd.read(10, v) != 10) return -2;  // NUM KSKIP MODE KOUNT KDUCTS KTAPE REFALT DSRRNG DSRLAT DSDENS
  g.num = (int)v[0];
  g.kducts = (int)v[4];
  const double dsrrng = v[7], dsrlat = v[8], dsdens = v[9];
  if (g.num < 2 || g.num > 4) return -2;
  if (rd.read(5, v) != 5) return -2;  // EGFEQ THERM HM ABSB RELB
  const double therm = v[1];
  if (rd.read(5, v) != 5) return -2;  // RBASE ANE0 ALPHA0(2:4)
  g.rbase = v[0]; g.ane0 = v[1]; g.alpha0[2] = v[2]; g.alpha0[3] = v[3]; g.alpha0[4] = v[4];
  if (rd.read(5, v) != 5) return -2;  // RZERO SCBOT RSTOP RDIV HMIN
  g.rzero = v[0]; g.scbot = v[1];
  if (g.kducts != 0) {
    if (rd.read(5, v) != 5) return -2;  // LK EXPK DDK RCONSN SCR
    g.lk = v[0]; g.expk = v[1]; g.ddk = v[2]; g.rconsn = v[3]; g.scr = v[4];
    if (g.kducts > SRT_MAX_DUCTS) return -3;
    for (int k = 2; k <= g.kducts; ++k) {
      if (rd.read(12, v) != 12) return -2;
      g.l0[k] = v[0]; g.def[k] = v[1]; g.dd[k] = v[2];
      g.rducln[k] = v[3]; g.hl2n[k] = v[4] * v[4];
      g.rducun[k] = v[5]; g.hu2n[k] = v[6] * v[6];
      g.rducls[k] = v[7]; g.hl2s[k] = v[8] * v[8];
      g.rducus[k] = v[9]; g.hu2s[k] = v[10] * v[10];
      g.sidedu[k] = v[11];
    }
  }
  if (rd.read(8, v) != 8) return -2;  // profile card (values unused: plotting only)
  // loop invariants of dens (ngo_dens_model.f95:181-184, :221)
  const double rb7370 = g.rbase / (double)7370.f;
  g.sh[2] = (double)1.150600f * therm * rb7370 * rb7370;
  g.sh[3] = g.sh[2] / 4.;
  g.sh[4] = g.sh[3] / 4.;
  g.d2k_2 = (g.ddk * g.ddk) * 2.0;
  // ane0 renormalised so that ani(1) = DSDENS at (DSRRNG R_E, DSRLAT deg) (:120-123)
  double ani[SRT_NSPEC];
  srt_ngo_dens(g, dsrrng * g.r0, (90.00 - dsrlat) * grarad, latitu, ani);
  g.ane0 = g.ane0 * dsdens / ani[0];
  // species of the adapter (ngo_dens_model_adapter.f95:135-140)
  M.density = SRT_DENS_NGO;
  M.nspec = 4;
  const double sgn[4] = {-1.0, 1.0, 1.0, 1.0};
  for (int i = 0; i < 4; ++i) M.qs[i] = 1.602e-19 * sgn[i];
  M.ms[0] = 9.10938188e-31;
  M.ms[1] = 1.6726e-27;
  M.ms[2] = 4.0 * 1.6726e-27;
  M.ms[3] = 16.0 * 1.6726e-27;
  M.del = 1.0e-4;  // delSP, raytracer_driver.f95:251
  set_species_invariants(M);
  return 0;
}

}  // namespace srt_host
