// srt_physics.h -- the per-ray arithmetic of the hot path, written once as
// __host__ __device__ inline code.  nvcc compiles it into the sm_100a kernels
// (srt_kernels.cu); g++ compiles the same text for the host-side model setup
// (ane0 normalisation, srt_host_setup.cpp) and for tests/csrc/hostemu.cpp, a
// test-only CPU build used to check the de-duplicated evaluation order against
// the oracle without a GPU.  Compile with contraction OFF (nvcc -fmad=false,
// g++ -ffp-contract=off): every result is then identical on both sides.
//
// What is restated here (reference file:line in each function):
//   fortran/raytracer.f95         dispersion_relation, stix_parameters, dF/dk, dF/dw,
//                                 dF/dx, evalrhs, solve_dispersion_relation,
//                                 is_right_handed, rk4/rk45 stage algebra
//   fortran/ngo_dens_model*.f95   Ngo density + adapter
//   fortran/bmodel_dipole.f95, fortran/util.f95, xform_double/Rotate_y.f95
//
// What is different from the reference: evaluations at bit-identical arguments are
// done once: per RHS 7 plasma evaluations instead of 8, 9 Stix
// parameter sets instead of 14; the plasma-only parts of stix_parameters (wps2, wcs)
// are computed once per plasma evaluation.  Operation order inside every expression
// is the reference's, so de-duplication does not change a single bit.
#pragma once
#include "srt_math.h"
#include "srt_model.h"
#include "srt_ts05.h"

// fortran/constants.f95:4-12
#define SRT_EPS0 8.854187817e-12
#define SRT_PI 3.141592653589793238462643
#define SRT_R_E 6371.2e3
// C = sqrt(1/EPS0/MU0) with MU0 = PI*4e-7, evaluated in FP64 exactly like the
// Fortran parameter expression (constants.f95:6-7)
SRT_HD double srt_c_light() { return srt_sqrt(1.0 / SRT_EPS0 / (SRT_PI * 4e-7)); }
#define SRT_R2D (180.0 / SRT_PI)


// How device code reaches the model: the kernels read it from __constant__ memory
// (operands fold into FP64 instructions as c[bank][offset]); host builds carry a pointer.
#if defined(__CUDACC__)
__constant__ SrtModel c_srt_model;
#endif
struct SrtModelRef {
#if defined(__CUDA_ARCH__)
  SRT_HD const SrtModel& get() const { return c_srt_model; }
#else
  const SrtModel* p;
  SRT_HD const SrtModel& get() const { return *p; }
#endif
};

struct SrtPlasma {
  double Ns[SRT_NSPEC];
  double B0[3];
};

// plasma-only invariants of stix_parameters / dispersion_relation
struct SrtPlasmaDerived {
  double wps2[SRT_NSPEC];  // raytracer.f95:92
  double wcs[SRT_NSPEC];   // raytracer.f95:93
  double B0[3];
  double B0sq;             // dot_product(B0,B0)
  double fs_thr;           // right-hand side of the free-space test, raytracer.f95:65
};

struct SrtStix {
  double S, D, P, R, L;
};

SRT_HD double srt_dot3(const double a[3], const double b[3]) {
  // dot_product() as gfortran inlines it: s = 0; s = s + a(i)*b(i)
  double s = 0.0;
  s = s + a[0] * b[0];
  s = s + a[1] * b[1];
  s = s + a[2] * b[2];
  return s;
}

// ------------------------------------------------------------------ geometry
// fortran/util.f95:109-122
SRT_HD void srt_cart_to_sph(const double x[3], double p[3]) {
  double s = 0.0;
  s = s + x[0] * x[0];
  s = s + x[1] * x[1];
  s = s + x[2] * x[2];
  p[0] = srt_sqrt(s);
  p[1] = srt_atan2(x[1], x[0]);
  if (p[0] != 0.0) p[2] = srt_acos((x[2]) / (p[0]));
  else p[2] = 0.0;
}

// rotate_y_d (xform_double/Rotate_y.f95:7-9) with precomputed cos/sin
SRT_HD void srt_rot_y(double c, double s, const double a[3], double b[3]) {
  b[0] = a[0] * c + a[2] * s;
  b[1] = a[1];
  b[2] = a[2] * c - a[0] * s;
}

// ------------------------------------------------------------------ Ngo density
// fortran/ngo_dens_model.f95:165-353 (`dens`), live outputs only: ani(1:4).
// z1 = geocentric radius seen by the model [km], z2 = colatitude [rad],
// latitu = latitude [deg] (used by the duct branches).
SRT_HDP void srt_ngo_dens(const SrtNgo& g, double z1, double z2, double latitu, double ani[SRT_NSPEC]) {
  double sinz2, cosz2;
  srt_sincos(z2, sinz2, cosz2);
  double sinz22 = sinz2 * sinz2;
  double gph = g.rbase * (1.0 - g.rbase / z1);
  double exnor2 = srt_exp(-gph / g.sh[2]);
  double exnor3 = exnor2 * exnor2 * exnor2 * exnor2;
  double exnor4 = exnor3 * exnor3 * exnor3 * exnor3;
  double qi2 = g.alpha0[2] * exnor2, qi3 = 0.0, qi4 = 0.0;
  double q = 0.0;
  q = q + qi2;
  if (g.num >= 3) { qi3 = g.alpha0[3] * exnor3; q = q + qi3; }
  if (g.num >= 4) { qi4 = g.alpha0[4] * exnor4; q = q + qi4; }
  double alpha2 = qi2 / q, alpha3 = 0.0, alpha4 = 0.0;
  if (g.num >= 3) alpha3 = qi3 / q;
  if (g.num >= 4) alpha4 = qi4 / q;
  double anr = srt_sqrt(q);
  double arg = (z1 - g.rzero) / g.scbot;
  if (!(arg < 13.0)) arg = 13.0;
  double exarg = srt_exp(-arg * arg);
  double anli = 1.0 - exarg;
  double l = z1 / (g.r0 * sinz22);
  double a1 = g.ane0 * anr;
  a1 = a1 * anli;
  if (g.kducts != 0) {
    double deltal = l - g.lk;
    if (!(deltal < 0.0)) {
      double argl = deltal * deltal / g.d2k_2;
      if (!(argl < 80.00)) argl = 80.00;
      double f = srt_exp(-argl);
      double trm = srt_pow((g.rconsn / z1), g.expk);
      double argr = (z1 - g.rconsn) / g.scr;
      if (!(argr < 12.50)) argr = 12.5;
      double fr = srt_exp(-argr * argr);
      double trmodl = trm + (1. - trm) * fr;
      double anlk = f + trmodl * (1.0 - f);
      a1 = a1 * anlk;
    }
    if (g.kducts != 1) {
      // ducts: ngo_dens_model.f95:240-337
      int kinit = 2;
      bool to250 = false;
      if (!(g.l0[2] > 0.0)) {
        kinit = 3;
        double dl = l + g.l0[2];
        if (!(dl * g.sidedu[2] >= 0.0)) dl = 0;
        double delk = -g.l0[2] - (g.lk + g.ddk) + g.dd[2] / 2;
        double critl = (g.lk + g.ddk) + srt_fmod(delk, g.dd[2]);
        if (!(l <= critl)) {
          double argl = 2.0 * g.pi_f * dl / g.dd[2];
          double delnl = (g.def[2] / 2.) * (1. + srt_cos(argl));
          bool lower = (latitu <= 0 && z1 <= g.rducus[2]) || (latitu >= 0 && z1 <= g.rducun[2]);
          bool apply = true, plain = false;
          double delr = 0.0, arglr = 0.0;
          if (!lower) {
            if (latitu >= 0) delr = z1 - g.rducun[2];
            if (latitu <= 0) delr = z1 - g.rducus[2];
            if (latitu <= 0) arglr = delr * delr / g.hu2s[2];
            if (latitu >= 0) arglr = delr * delr / g.hu2n[2];
            if (arglr >= 75.0) apply = false;
          } else {
            if ((latitu <= 0 && z1 >= g.rducls[2]) || (latitu >= 0 && z1 >= g.rducln[2])) plain = true;
            else {
              if (latitu >= 0) delr = z1 - g.rducln[2];
              if (latitu <= 0) delr = z1 - g.rducls[2];
              if (latitu <= 0) arglr = delr * delr / g.hl2s[2];
              if (latitu >= 0) arglr = delr * delr / g.hl2n[2];
              if (arglr >= 75.0) apply = false;
            }
          }
          if (apply) {
            if (!plain) delnl = delnl * srt_exp(-arglr);
            double anl = 1.0 + delnl;
            a1 = a1 * anl;
          }
        }
        if (g.kducts == 2) to250 = true;
      }
      if (!to250) {
        for (int kd = kinit; kd <= g.kducts; ++kd) {
          double dl = l - g.l0[kd];
          if (!(dl * g.sidedu[kd] >= 0.0)) dl = 0;
          double d2 = g.dd[kd] * g.dd[kd];
          double argl = dl * dl / (d2 * 2.0);
          if (argl > 80.0) continue;
          double delnl = g.def[kd] * srt_exp(-argl);
          bool lower = (latitu >= 0 && z1 <= g.rducun[kd]) || (latitu <= 0 && z1 <= g.rducus[kd]);
          bool plain = false;
          double delr = 0.0, arglr = 0.0;
          if (!lower) {
            if (latitu >= 0) delr = z1 - g.rducun[kd];
            if (latitu <= 0) delr = z1 - g.rducus[kd];
            if (latitu >= 0) arglr = delr * delr / g.hu2n[kd];
            if (latitu <= 0) arglr = delr * delr / g.hu2s[kd];
            if (arglr >= 75.0) continue;
          } else {
            if ((latitu >= 0 && z1 >= g.rducln[kd]) || (latitu <= 0 && z1 >= g.rducls[kd])) plain = true;
            else {
              if (latitu >= 0) delr = z1 - g.rducln[kd];
              if (latitu <= 0) delr = z1 - g.rducls[kd];
              if (latitu >= 0) arglr = delr * delr / g.hl2n[kd];
              if (latitu <= 0) arglr = delr * delr / g.hl2s[kd];
              if (arglr >= 75.0) continue;
            }
          }
          if (!plain) delnl = delnl * srt_exp(-arglr);
          double anl = 1.0 + delnl;
          a1 = a1 * anl;
        }
      }
    }
  }
  ani[0] = a1;
  ani[1] = a1 * alpha2;
  ani[2] = (g.num >= 3) ? a1 * alpha3 : 0.0;
  ani[3] = (g.num >= 4) ? a1 * alpha4 : 0.0;
}

// ------------------------------------------------------------------ magnetic field
// Centred dipole in SM: bmodel_cartesian (fortran/bmodel_dipole.f95:20-48) followed by
// spherical_to_cartesian_vec (fortran/util.f95:126-142, matmul(transpose(A),p) with
// the accumulator starting at 0).  p = cartesian_to_spherical(x); sph/cph, sth/cth
// are sin/cos of p(3) (polar angle) and p(2) (azimuth).
SRT_HDP void srt_dipole_sm(const double p[3], double sph, double cph, double sth, double cth, double B[3]) {
  double Bo = .312 / 10000.0;
  double R = p[0] / SRT_R_E;
  double Bor3 = Bo * (srt_pow(R, -3.0));
  double Brad = -2.0 * Bor3 * cph;
  double Bth = -Bor3 * sph;
  // A(1,1)=ct*sp A(2,1)=-st A(3,1)=ct*cp | A(1,2)=st*sp A(2,2)=ct A(3,2)=st*cp | A(1,3)=cp A(2,3)=0 A(3,3)=-sp
  double s;
  s = 0.0; s = s + (cth * sph) * Brad; s = s + (-sth) * 0.0; s = s + (cth * cph) * Bth; B[0] = s;
  s = 0.0; s = s + (sth * sph) * Brad; s = s + (cth) * 0.0;  s = s + (sth * cph) * Bth; B[1] = s;
  s = 0.0; s = s + (cph) * Brad;       s = s + (0.0) * 0.0;  s = s + (-sph) * Bth;      B[2] = s;
}

// IGRF_GSW_08 in fp32; defined in srt_igrf.h
SRT_HDN void srt_igrf_gsw(const SrtIgrf& ig, float xgsw, float ygsw, float zgsw, float& hx, float& hy, float& hz);

// The field block both adapters share (ngo_dens_model_adapter.f95:142-202,
// interp_dens_model_adapter.f95:208-267) with use_tsyganenko = 0:
// base field in GSM [nT] rounded to real*4 (Q3), + 0.0 (fp32), *1e-9, rotated to SM.
// For the dipole the caller passes the spherical coordinates of x it already has.
template <int USE_IGRF>
SRT_HD void srt_field_sm(SrtModelRef MR, const double x[3], const double p[3], double sph, double cph, double sth, double cth,
                         double B0[3]) {
  const SrtModel& M = MR.get();
  float bx, by, bz;
  if (USE_IGRF & 1) {
    double xg[3];
    srt_rot_y(M.cos_mu, -M.sin_mu, x, xg);  // SM_TO_GSM_d: rotate_y(-mu)
    srt_igrf_gsw(M.igrf, (float)(xg[0] / SRT_R_E), (float)(xg[1] / SRT_R_E), (float)(xg[2] / SRT_R_E), bx, by, bz);
  } else {
    double Bsm[3], Bg[3];
    srt_dipole_sm(p, sph, cph, sth, cth, Bsm);
    srt_rot_y(M.cos_mu, -M.sin_mu, Bsm, Bg);
    bx = (float)(1.0e9 * Bg[0]);
    by = (float)(1.0e9 * Bg[1]);
    bz = (float)(1.0e9 * Bg[2]);
  }
  float tx = 0.0f, ty = 0.0f, tz = 0.0f;
  if (USE_IGRF & 2) {  // use_tsyganenko = 1: T04_s at the GSM position in R_E (adapter :183-187)
    double xg[3];
    srt_rot_y(M.cos_mu, -M.sin_mu, x, xg);
    srt_ts05(M.ts, (float)(xg[0] / SRT_R_E), (float)(xg[1] / SRT_R_E), (float)(xg[2] / SRT_R_E), tx, ty, tz);
  }
  double Bt[3];
  Bt[0] = (double)(bx + tx) * 1.0e-9;
  Bt[1] = (double)(by + ty) * 1.0e-9;
  Bt[2] = (double)(bz + tz) * 1.0e-9;
  srt_rot_y(M.cos_mu, M.sin_mu, Bt, B0);  // GSM_TO_SM_d: rotate_y(+mu)
}

// gridded log-density (srt_tricubic.h)
SRT_HDN void srt_grid_density(const SrtGrid& g, const double x[3], double Ns[SRT_NSPEC]);
// scattered samples, moving least squares (modelnum 4)
#include "srt_scattered.h"

// ------------------------------------------------------------------ funcPlasmaParams
// DENS = SrtDensityKind, USE_IGRF = the field code: bit 0 IGRF instead of the dipole, bit 1 TS05 external field added.
//   Ngo:  fortran/ngo_dens_model_adapter.f95:63-207
//   grid: fortran/interp_dens_model_adapter.f95:148-269
//   scattered samples: fortran/scattered_interp_dens_model_adapter.f95:214-370 (the same field block)
template <int DENS, int USE_IGRF>
SRT_HDP void srt_plasma(SrtModelRef MR, const double x[3], SrtPlasma& o) {
  SRT_COUNT(SRT_T_PLASMA);
  const SrtModel& M = MR.get();
  double p[3] = {0.0, 0.0, 0.0};
  double sph = 0.0, cph = 0.0, sth = 0.0, cth = 0.0;
  if (DENS == SRT_DENS_NGO || !(USE_IGRF & 1)) {
    srt_cart_to_sph(x, p);
    srt_sincos(p[2], sph, cph);
  }
  if (DENS == SRT_DENS_NGO) {
    const double d2r = 2.0 * SRT_PI / 360.0;
    double L;
    if (SRT_R_E * (sph * sph) != 0.0) L = p[0] / (SRT_R_E * (sph * sph));
    else L = 0.0;
    double lam = 90.0 - (p[2] * 360.0 / 2.0 / SRT_PI);  // degrees
    double lamr = d2r * lam;
    double cl = srt_cos(lamr);
    double r = M.ngo.r0 * L * (cl * cl);
    double ani[SRT_NSPEC];
    srt_ngo_dens(M.ngo, r, d2r * (90.0 - lam), lam, ani);
    o.Ns[0] = 1.0e6 * ani[0];
    o.Ns[1] = 1.0e6 * ani[1];
    o.Ns[2] = 1.0e6 * ani[2];
    o.Ns[3] = 1.0e6 * ani[3];
  } else if (DENS == SRT_DENS_SCAT) {
    srt_scat_density(M.scat, x, o.Ns);
  } else {
    srt_grid_density(M.grid, x, o.Ns);
  }
  if (!(USE_IGRF & 1)) srt_sincos(p[1], sth, cth);
  srt_field_sm<USE_IGRF>(MR, x, p, sph, cph, sth, cth, o.B0);
}

// ------------------------------------------------------------------ Stix / dispersion
// plasma-only parts of stix_parameters (raytracer.f95:92-93) and of the free-space
// test (raytracer.f95:65)
SRT_HDS void srt_plasma_derive(SrtModelRef MR, const SrtPlasma& pl, SrtPlasmaDerived& d) {
  SRT_COUNT(SRT_T_DERIVE);
  const SrtModel& M = MR.get();
  d.B0[0] = pl.B0[0]; d.B0[1] = pl.B0[1]; d.B0[2] = pl.B0[2];
  d.B0sq = srt_dot3(pl.B0, pl.B0);
  double B0mag = srt_sqrt(d.B0sq);
  double maxNs = pl.Ns[0];
  // wps2 = (Ns*qs**2/ms/EPS0), wcs = (qs*B0mag)/ms: the divisions of the four species are independent (srt_divn)
  double num[SRT_NSPEC], den[SRT_NSPEC], t1[SRT_NSPEC], eps0[SRT_NSPEC], qb[SRT_NSPEC];
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) {
    num[i] = pl.Ns[i] * M.q2[i];
    den[i] = M.ms[i];
    eps0[i] = SRT_EPS0;
    qb[i] = (M.qs[i] * B0mag);
    if (i > 0 && pl.Ns[i] > maxNs) maxNs = pl.Ns[i];
  }
  srt_divn(num, den, t1);
  srt_divn(t1, eps0, d.wps2);
  srt_divn(qb, den, d.wcs);
  d.fs_thr = 100.0 * srt_sqrt((maxNs * M.fs_q2)) / M.fs_den;
}

// stix_parameters, raytracer.f95:96-100
SRT_HDS void srt_stix(const SrtPlasmaDerived& d, double w, SrtStix& s) {
  SRT_COUNT(SRT_T_STIX);
  double sr = 0.0, sl = 0.0, sp = 0.0;
  // the twelve quotients are independent of each other (srt_divn); the sums run in index order as Fortran's sum()
  double dr[SRT_NSPEC], dl[SRT_NSPEC], dp[SRT_NSPEC], qr[SRT_NSPEC], ql[SRT_NSPEC], qp[SRT_NSPEC];
  double w2 = w * w;
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) {
    dr[i] = (w * (w + d.wcs[i]));
    dl[i] = (w * (w - d.wcs[i]));
    dp[i] = w2;
  }
  srt_divn(d.wps2, dr, qr);
  srt_divn(d.wps2, dl, ql);
  srt_divn(d.wps2, dp, qp);
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) sr = sr + qr[i];
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) sl = sl + ql[i];
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) sp = sp + qp[i];
  s.R = 1.0 - sr;
  s.L = 1.0 - sl;
  s.P = 1.0 - sp;
  s.S = 1.0 / 2.0 * (s.R + s.L);
  s.D = 1.0 / 2.0 * (s.R - s.L);
}

// dispersion_relation after the Stix parameters are known (raytracer.f95:50-68)
SRT_HDS double srt_F(const double n[3], double w, const SrtPlasmaDerived& d, const SrtStix& s) {
  SRT_COUNT(SRT_T_FTAIL);
  double nmag2 = srt_dot3(n, n);
  double nb = srt_dot3(n, d.B0);
  double cos2phi = (nb * nb) / (nmag2 * d.B0sq);
  double sin2phi = 1.0 - cos2phi;
  double A = s.S * sin2phi + s.P * cos2phi;
  double RL = s.R * s.L;
  double B = RL * sin2phi + s.P * s.S * (1.0 + cos2phi);
  if (w > d.fs_thr) return -nmag2 + 1.0;
  return A * (nmag2 * nmag2) - B * nmag2 + RL * s.P;
}

SRT_HD double srt_fd_step(double del, double v) {
  double a = del * srt_fabs(v);
  return (a > del) ? a : del;  // max(del*abs(v), del)
}

// dispersion_relation_dFdk (raytracer.f95:118-155) given the derived plasma at x and
// the Stix set at w.  Components of n that are not perturbed are k(j)*C/w on both
// sides ((k(j) +- d*0) == k(j)).
SRT_HDN void srt_dFdk(const double k[3], double w, const SrtPlasmaDerived& d, const SrtStix& s, double out[3]) {
  const double C = srt_c_light();
  const double del = 1.0e-8;
  double n0[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) n0[j] = srt_div(k[j] * C, w);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double dd = srt_fd_step(del, k[i]);
    double np[3] = {n0[0], n0[1], n0[2]};
    double nn[3] = {n0[0], n0[1], n0[2]};
    np[i] = (k[i] + dd) * C / w;
    nn[i] = (k[i] - dd) * C / w;
    out[i] = srt_div(srt_F(np, w, d, s) - srt_F(nn, w, d, s), dd) / 2.0;
  }
}

// dispersion_relation_dFdw (raytracer.f95:172-198)
SRT_HDN double srt_dFdw(const double k[3], double w, const SrtPlasmaDerived& d) {
  const double C = srt_c_light();
  double dd = srt_fd_step(1.0e-8, w);
  double wp = w + dd, wm = w - dd;
  double np[3], nn[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    np[j] = srt_div(k[j] * C, wp);
    nn[j] = srt_div(k[j] * C, wm);
  }
  SrtStix sp, sm;
  srt_stix(d, wp, sp);
  srt_stix(d, wm, sm);
  return (srt_F(np, wp, d, sp) - srt_F(nn, wm, d, sm)) / dd / 2.0;
}

// one side of a dF/dx difference: plasma at xs, Stix at w, F(n)
template <int DENS, int USE_IGRF>
SRT_HDN double srt_F_at(SrtModelRef MR, const double xs[3], const double n[3], double w) {
  const SrtModel& M = MR.get();
  SrtPlasma pl;
  SrtPlasmaDerived d;
  SrtStix s;
  srt_plasma<DENS, USE_IGRF>(MR, xs, pl);
  srt_plasma_derive(MR, pl, d);
  srt_stix(d, w, s);
  return srt_F(n, w, d, s);
}

// dispersion_relation_dFdx (raytracer.f95:215-263)
template <int DENS, int USE_IGRF>
SRT_HD void srt_dFdx(SrtModelRef MR, const double k[3], double w, const double x[3], double out[3]) {
  const SrtModel& M = MR.get();
  const double C = srt_c_light();
  double n[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) n[j] = srt_div(k[j] * C, w);
#pragma unroll 1
  for (int i = 0; i < 3; ++i) {
    double dd = srt_fd_step(M.del, x[i]);
    double xp[3] = {x[0], x[1], x[2]}, xn[3] = {x[0], x[1], x[2]};
    xp[i] = x[i] + dd;
    xn[i] = x[i] - dd;
    double Fp = srt_F_at<DENS, USE_IGRF>(MR, xp, n, w);
    double Fn = srt_F_at<DENS, USE_IGRF>(MR, xn, n, w);
    out[i] = srt_div(Fp - Fn, dd) / 2.0;
  }
}

// centre-point quantities that several consumers share at one (x, w)
struct SrtCentre {
  SrtPlasma pl;
  SrtPlasmaDerived d;
  SrtStix s;  // Stix set at (x, w)
};
template <int DENS, int USE_IGRF>
SRT_HDN void srt_centre(SrtModelRef MR, const double x[3], double w, SrtCentre& c) {
  const SrtModel& M = MR.get();
  srt_plasma<DENS, USE_IGRF>(MR, x, c.pl);
  srt_plasma_derive(MR, c.pl, c.d);
  srt_stix(c.d, w, c.s);
}

// raytracer_evalrhs (raytracer.f95:282-314): rhs[0:3] = -(dfdk/dfdw), rhs[3:6] = dfdx/dfdw
// (rhs(7) = 0 is implicit).
template <int DENS, int USE_IGRF>
SRT_HDN void srt_evalrhs(SrtModelRef MR, const double x[3], const double k[3], double w, double rhs[6]) {
  const SrtModel& M = MR.get();
  SrtCentre c;
  srt_centre<DENS, USE_IGRF>(MR, x, w, c);
  double dfdk[3], dfdx[3];
  srt_dFdk(k, w, c.d, c.s, dfdk);
  double dfdw = srt_dFdw(k, w, c.d);
  srt_dFdx<DENS, USE_IGRF>(MR, k, w, x, dfdx);
#pragma unroll
  for (int i = 0; i < 3; ++i) rhs[i] = -srt_div(dfdk[i], dfdw);
#pragma unroll
  for (int i = 0; i < 3; ++i) rhs[3 + i] = srt_div(dfdx[i], dfdw);
}

// ------------------------------------------------------------------ root solve + handedness
struct SrtCplx {
  double re, im;
};
SRT_HD SrtCplx srt_cdiv(SrtCplx a, SrtCplx b) {
  // Smith's algorithm, the order GCC emits for Fortran complex division
  SrtCplx r;
  if (srt_fabs(b.re) < srt_fabs(b.im)) {
    double ratio = srt_div(b.re, b.im);
    double div = (b.re * ratio) + b.im;
    r.re = srt_div((a.re * ratio) + a.im, div);
    r.im = srt_div((a.im * ratio) - a.re, div);
  } else {
    double ratio = srt_div(b.im, b.re);
    double div = (b.im * ratio) + b.re;
    r.re = srt_div((a.im * ratio) + a.re, div);
    r.im = srt_div(a.im - (a.re * ratio), div);
  }
  return r;
}
SRT_HD SrtCplx srt_cmul(SrtCplx a, SrtCplx b) {
  SrtCplx r;
  r.re = a.re * b.re - a.im * b.im;
  r.im = a.re * b.im + a.im * b.re;
  return r;
}

// is_right_handed (raytracer.f95:373-405).  The reference takes the right singular
// vector of the smallest singular value of the 3x3 wave matrix M (entries rounded to
// real*4, Q8; phi in degrees fed to cos/sin, Q9) from LAPACK ZGESVD and tests the
// sign of the phase lead of E_y over E_x.  M = T Mr T^H with T = diag(1, i, 1) and Mr
// real symmetric [[a, D, b], [D, c, 0], [b, 0, e]], so that vector is T*v (times a
// unit phase) with v the eigenvector of Mr whose eigenvalue is smallest in
// magnitude, and the reference's angle has the sign of v1*v2.  Mr is diagonalised by
// cyclic Jacobi rotations.
SRT_HDE bool srt_is_right_handed(double n2, double phi_deg, double S, double D, double P) {
  SRT_COUNT(SRT_T_HAND);
  double sphi, cphi;
  srt_sincos(phi_deg, sphi, cphi);
  double d0 = (double)(float)(S - n2 * (cphi * cphi));
  double d1 = (double)(float)(S - n2);
  double d2 = (double)(float)(P - n2 * (sphi * sphi));
  double a01 = (double)(float)(D);
  double a02 = (double)(float)(n2 * cphi * sphi);
  double a12 = 0.0;
  double v[3][3] = {{1.0, 0.0, 0.0}, {0.0, 1.0, 0.0}, {0.0, 0.0, 1.0}};
  double dg[3] = {d0, d1, d2};
  double off[3] = {a01, a02, a12};  // (0,1), (0,2), (1,2)
  for (int sweep = 0; sweep < 40; ++sweep) {
    double sm = srt_fabs(off[0]) + srt_fabs(off[1]) + srt_fabs(off[2]);
    if (sm == 0.0) break;
#pragma unroll
    for (int pair = 0; pair < 3; ++pair) {
      const int p = (pair == 2) ? 1 : 0;
      const int q = (pair == 0) ? 1 : 2;
      const int r = 3 - p - q;  // the third index
      double apq = off[pair];
      double g = 100.0 * srt_fabs(apq);
      if (sweep > 3 && srt_fabs(dg[p]) + g == srt_fabs(dg[p]) && srt_fabs(dg[q]) + g == srt_fabs(dg[q])) {
        off[pair] = 0.0;
      } else if (apq != 0.0) {
        double h = dg[q] - dg[p];
        double t;
        if (srt_fabs(h) + g == srt_fabs(h)) {
          t = apq / h;
        } else {
          double theta = 0.5 * h / apq;
          t = 1.0 / (srt_fabs(theta) + srt_sqrt(1.0 + theta * theta));
          if (theta < 0.0) t = -t;
        }
        double c = 1.0 / srt_sqrt(1.0 + t * t);
        double s = t * c;
        dg[p] = dg[p] - t * apq;
        dg[q] = dg[q] + t * apq;
        off[pair] = 0.0;
        // rows/cols r: a_rp, a_rq
        const int irp = (r < p) ? (r + p - 1) : (p + r - 1);  // index of (min,max): (0,1)->0 (0,2)->1 (1,2)->2
        const int irq = (r < q) ? (r + q - 1) : (q + r - 1);
        double arp = off[irp], arq = off[irq];
        off[irp] = c * arp - s * arq;
        off[irq] = s * arp + c * arq;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          double vip = v[i][p], viq = v[i][q];
          v[i][p] = c * vip - s * viq;
          v[i][q] = s * vip + c * viq;
        }
      }
    }
  }
  int im = 0;
  if (srt_fabs(dg[1]) < srt_fabs(dg[im])) im = 1;
  if (srt_fabs(dg[2]) < srt_fabs(dg[im])) im = 2;
  double v1 = (im == 0) ? v[0][0] : (im == 1 ? v[0][1] : v[0][2]);
  double v2 = (im == 0) ? v[1][0] : (im == 1 ? v[1][1] : v[1][2]);
  return !(v1 * v2 < 0.0);
}

// solve_dispersion_relation (raytracer.f95:408-502) with the plasma at x (derived) and
// the Stix set at (x, w) supplied.  Returns the complex root selected by `root`
// (1: k1, else k2), i.e. what raytracer_run picks at :685-689 / :828-836.
SRT_HDE SrtCplx srt_solve_root(const double k[3], double w, const SrtPlasmaDerived& d, const SrtStix& s, int root) {
  SRT_COUNT(SRT_T_ROOT);
  const double C = srt_c_light();
  double kb = srt_dot3(k, d.B0);
  double cos2phi = (kb * kb) / (srt_dot3(k, k) * d.B0sq);
  double sin2phi = 1.0 - cos2phi;
  double phi = srt_acos(srt_sqrt(cos2phi)) * SRT_R2D;
  double A = s.S * sin2phi + s.P * cos2phi;
  double B = s.R * s.L * sin2phi + s.P * s.S * (1.0 + cos2phi);
  double disc = B * B - 4.0 * A * s.R * s.L * s.P;
  SrtCplx sq;
  srt_csqrt(disc, 0.0, sq.re, sq.im);
  SrtCplx twoA = {2.0 * A, 0.0};
  SrtCplx a1 = {B + sq.re, sq.im}, a2 = {B - sq.re, -sq.im};
  SrtCplx nsq1 = srt_cdiv(a1, twoA);
  SrtCplx nsq2 = srt_cdiv(a2, twoA);
  SrtCplx n1, n2;
  srt_csqrt(nsq1.re, nsq1.im, n1.re, n1.im);
  srt_csqrt(nsq2.re, nsq2.im, n2.re, n2.im);
  bool swap = false;
  if (n1.re > 0.0) swap = srt_is_right_handed(nsq1.re, phi, s.S, s.D, s.P);
  // k1 = w*n1/C, k2 = w*n2/C; swapped when root 1 is the right-handed mode
  bool want_n1 = (root == 1) ? !swap : swap;
  SrtCplx n = want_n1 ? n1 : n2;
  SrtCplx kk = {w * n.re / C, srt_div(w * n.im, C)};
  return kk;
}

// ------------------------------------------------------------------ the per-ray integrator
// raytracer_run (raytracer.f95:609-995) as an explicit state machine: srt_ray_init
// records row 1, then each srt_ray_attempt is one pass through the `do` loop at
// :748-993 (one accepted step, one rejected / cone-refined attempt, or termination).

#define SRT_ROW_DOUBLES 20  // t, pos3, vprel3, vgrel3, n3, B03, Ns4 (driver:1197-1217 minus constants)

struct SrtRunParams {
  double dt0, dtmax, maxerr, minalt, tmax;
  int maxsteps, root, fixedstep, outputper;
};

enum { SRT_ATTEMPT_RETRY = 0, SRT_ATTEMPT_ACCEPTED = 1, SRT_ATTEMPT_DONE = 2 };

struct SrtRayState {
  double x[3], k[3], w;  // state vector x(1:7)
  double t, dt;
  double wl;             // the local `w` of raytracer_run: 0 until the first acceptance (Q1)
  double vg[3];          // vgrel of the last recorded row (stop condition 3)
  double dfdk[3], dfdw;  // at (x, k, w): recorded for vgrel, reused by stage 1
  double rhs1[6];        // raytracer_evalrhs at the current point, kept across rejections
  int have_rhs1;
  int nstep, lastrefinedown, stopcond;
  int fixed_cone_exit;   // Q6 side channel
  unsigned nrej, ncone;
};

// fills row[] for the current state (:710-741 and :937-986)
SRT_HDE void srt_fill_row(const SrtRayState& st, const SrtPlasma& pl, bool propagating, double row[SRT_ROW_DOUBLES]) {
  const double C = srt_c_light();
  double n[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) n[i] = srt_div(st.k[i] * C, st.w);
  row[0] = st.t;
#pragma unroll
  for (int i = 0; i < 3; ++i) row[1 + i] = st.x[i];
  if (propagating) {
    double nn = srt_dot3(n, n);
#pragma unroll
    for (int i = 0; i < 3; ++i) row[4 + i] = srt_div(n[i], nn);
#pragma unroll
    for (int i = 0; i < 3; ++i) row[7 + i] = -srt_div(srt_div(st.dfdk[i], st.dfdw), C);
  } else {
#pragma unroll
    for (int i = 0; i < 3; ++i) { row[4 + i] = 0.0; row[7 + i] = 0.0; }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) row[10 + i] = n[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) row[13 + i] = pl.B0[i];
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) row[16 + i] = pl.Ns[i];
}

// raytracer.f95:661-742.  dir0 is inout (field-aligned convention when all zero).
template <int DENS, int USE_IGRF>
SRT_HDN void srt_ray_init(SrtModelRef MR, const SrtRunParams& P, const double pos0[3], double dir0[3], double w0, SrtRayState& st,
                         double row[SRT_ROW_DOUBLES]) {
  const SrtModel& M = MR.get();
  SrtCentre c;
  srt_centre<DENS, USE_IGRF>(MR, pos0, w0, c);
  if (dir0[0] == 0 && dir0[1] == 0 && dir0[2] == 0) {
    double p[3], sth, cth, sph, cph;
    srt_cart_to_sph(pos0, p);
    srt_sincos(p[1], sth, cth);
    srt_sincos(p[2], sph, cph);
    // A as in util.f95:156-158; cartesian_to_spherical_vec = matmul(A, B0)
    double A[3][3] = {{cth * sph, sth * sph, cph}, {-sth, cth, 0.0}, {cth * cph, sth * cph, -sph}};
    double Bs[3], Bc[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      double s = 0.0;
      s = s + A[i][0] * c.pl.B0[0]; s = s + A[i][1] * c.pl.B0[1]; s = s + A[i][2] * c.pl.B0[2];
      Bs[i] = s;
    }
    Bs[0] = srt_fabs(Bs[0]);
    // spherical_to_cartesian_vec = matmul(transpose(A), Bs)
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      double s = 0.0;
      s = s + A[0][i] * Bs[0]; s = s + A[1][i] * Bs[1]; s = s + A[2][i] * Bs[2];
      Bc[i] = s;
    }
    double nb = srt_sqrt(srt_dot3(Bc, Bc));
#pragma unroll
    for (int i = 0; i < 3; ++i) dir0[i] = Bc[i] / nb;
  }
  SrtCplx k0 = srt_solve_root(dir0, w0, c.d, c.s, P.root);
#pragma unroll
  for (int i = 0; i < 3; ++i) st.x[i] = pos0[i];
#pragma unroll
  // real(k0mag*dir0), raytracer.f95:691: COMPLEX * REAL is the full complex product with dir0 as (d, +0.0) (gcc does not shortcut
  // an operand whose imaginary part is the constant 0.0 while signed zeros are honoured), so the real part carries -im*0.0:
  // only the sign of a zero (evanescent start: k0 purely imaginary) or a NaN imaginary part can show.  Found by the
  // comparison with the reference's translated source (tests/test_raytracer_ref_cpu.py).
  for (int i = 0; i < 3; ++i) st.k[i] = k0.re * dir0[i] - k0.im * 0.0;
  st.w = w0;
  st.dt = P.dt0;
  st.t = 0.0;
  st.wl = 0.0;
  st.lastrefinedown = 0;
  st.nstep = 1;
  st.stopcond = 0;
  st.fixed_cone_exit = 0;
  st.nrej = 0;
  st.ncone = 0;
  st.have_rhs1 = 0;
  srt_dFdk(st.k, w0, c.d, c.s, st.dfdk);
  st.dfdw = srt_dFdw(st.k, w0, c.d);
  const double C = srt_c_light();
  double n[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) n[i] = srt_div(st.k[i] * C, w0);
  bool prop = srt_dot3(n, n) > 0;
  srt_fill_row(st, c.pl, prop, row);
#pragma unroll
  for (int i = 0; i < 3; ++i) st.vg[i] = row[7 + i];
}

// raytracer_stopconditions (raytracer.f95:324-353)
SRT_HD int srt_stopconditions(const SrtRayState& st, const SrtRunParams& P) {
  if (srt_sqrt(srt_dot3(st.x, st.x)) < P.minalt) return 1;
  if (srt_sqrt(srt_dot3(st.k, st.k)) == 0.0) return 2;
  if (srt_sqrt(srt_dot3(st.vg, st.vg)) > 1.0 + 1e-2) return 3;
  if (st.dt < (double)1e-14f) return 5;  // single-precision literal (Q2)
  if (st.nstep >= P.maxsteps) return 6;
  return 0;
}

// rk45 tableau (raytracer.f95:8-27), evaluated as the same FP64 quotients
#define SRT_A21 (1.0 / 4.0)
#define SRT_A31 (3.0 / 32.0)
#define SRT_A32 (9.0 / 32.0)
#define SRT_A41 (1932.0 / 2197.0)
#define SRT_A42 (-7200.0 / 2197.0)
#define SRT_A43 (7296.0 / 2197.0)
#define SRT_A51 (439.0 / 216.0)
#define SRT_A52 (-8.0)
#define SRT_A53 (3680.0 / 513.0)
#define SRT_A54 (-845.0 / 4104.0)
#define SRT_A61 (-8.0 / 27.0)
#define SRT_A62 (2.0)
#define SRT_A63 (-3544.0 / 2565.0)
#define SRT_A64 (1859.0 / 4104.0)
#define SRT_A65 (-11.0 / 40.0)
#define SRT_B41 (25.0 / 216.0)
#define SRT_B42 (0.0)
#define SRT_B43 (1408.0 / 2565.0)
#define SRT_B44 (2197.0 / 4104.0)
#define SRT_B45 (-1.0 / 5.0)
#define SRT_B46 (0.0)
#define SRT_B51 (16.0 / 135.0)
#define SRT_B52 (0.0)
#define SRT_B53 (6656.0 / 12825.0)
#define SRT_B54 (28561.0 / 56430.0)
#define SRT_B55 (-9.0 / 50.0)
#define SRT_B56 (2.0 / 55.0)

// One pass through the loop body of raytracer_run (:748-993).
template <int DENS, int USE_IGRF>
SRT_HD int srt_ray_attempt(SrtModelRef MR, const SrtRunParams& P, SrtRayState& st, double row[SRT_ROW_DOUBLES]) {
  const SrtModel& M = MR.get();
  if (st.t >= P.tmax) { st.stopcond = 0; return SRT_ATTEMPT_DONE; }  // :749
  st.stopcond = srt_stopconditions(st, P);                           // :756
  if (st.stopcond != 0) return SRT_ATTEMPT_DONE;

  // stage 1 = raytracer_evalrhs at the current point.  dFdk and dFdw there are the
  // ones recorded at the end of the previous step (same k, w, x); only dF/dx is new.
  if (!st.have_rhs1) {
    double dfdx[3];
    srt_dFdx<DENS, USE_IGRF>(MR, st.k, st.w, st.x, dfdx);
#pragma unroll
    for (int i = 0; i < 3; ++i) st.rhs1[i] = -(st.dfdk[i] / st.dfdw);
#pragma unroll
    for (int i = 0; i < 3; ++i) st.rhs1[3 + i] = dfdx[i] / st.dfdw;
    st.have_rhs1 = 1;
  }
  const double dt = st.dt;
  double xs[6] = {st.x[0], st.x[1], st.x[2], st.k[0], st.k[1], st.k[2]};
  double e2x[6];  // est2(1:6)
  SrtCentre c2;   // plasma / derived / Stix(w) at est2(1:3) = the new point
  bool have_c2 = false;
  double k1[6];
#pragma unroll
  for (int i = 0; i < 6; ++i) k1[i] = dt * st.rhs1[i];

  if (P.fixedstep == 0) {
    double k2[6], k3[6], k4[6], k5[6], k6[6], tmp[6], r[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + (SRT_A21 * k1[i]);
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k2[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + (SRT_A31 * k1[i] + SRT_A32 * k2[i]);
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k3[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + (SRT_A41 * k1[i] + SRT_A42 * k2[i] + SRT_A43 * k3[i]);
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k4[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + (SRT_A51 * k1[i] + SRT_A52 * k2[i] + SRT_A53 * k3[i] + SRT_A54 * k4[i]);
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k5[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i)
      tmp[i] = xs[i] + (SRT_A61 * k1[i] + SRT_A62 * k2[i] + SRT_A63 * k3[i] + SRT_A64 * k4[i] + SRT_A65 * k5[i]);
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k6[i] = dt * r[i];
    double e1x[6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
      e1x[i] = xs[i] + (SRT_B41 * k1[i] + SRT_B42 * k2[i] + SRT_B43 * k3[i] + SRT_B44 * k4[i] + SRT_B45 * k5[i] + SRT_B46 * k6[i]);
#pragma unroll
    for (int i = 0; i < 6; ++i)
      e2x[i] = xs[i] + (SRT_B51 * k1[i] + SRT_B52 * k2[i] + SRT_B53 * k3[i] + SRT_B54 * k4[i] + SRT_B55 * k5[i] + SRT_B56 * k6[i]);
    // error estimate (:778-788)
    double sa = 0.0, sb = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) sa = sa + srt_fabs(e1x[3 + i] - e2x[3 + i]);
#pragma unroll
    for (int i = 0; i < 3; ++i) sb = sb + srt_fabs(e2x[3 + i]);
    double err = sa / sb;
    if (st.wl != 0.0) {
      // dFdk at est1 and est2 with the local w (== x(7) after the first acceptance)
      SrtCentre c1;
      double g1[3], g2[3];
      srt_centre<DENS, USE_IGRF>(MR, e1x, st.wl, c1);
      srt_dFdk(e1x + 3, st.wl, c1.d, c1.s, g1);
      // the plasma at est2 is also the plasma of the new point (:819): kept for the re-projection below
      srt_centre<DENS, USE_IGRF>(MR, e2x, st.wl, c2);
      srt_dFdk(e2x + 3, st.wl, c2.d, c2.s, g2);
      have_c2 = (st.wl == st.w);
      double sc = 0.0, sd = 0.0;
#pragma unroll
      for (int i = 0; i < 3; ++i) sc = sc + srt_fabs(g1[i] - g2[i]);
#pragma unroll
      for (int i = 0; i < 3; ++i) sd = sd + srt_fabs(g2[i]);
      double e2 = sc / sd;
      if (e2 > err || err != err) err = e2;  // MAX() as gfortran <= 8 expands it
    } else {
      // Q1: w == 0 makes both dFdk estimates NaN; MAX(a1, NaN) keeps a1 unless a1 is NaN
      if (err != err) err = srt_nan();
    }
    if (err > P.maxerr) {  // :790-798
      st.dt = 0.8 * st.dt;
      st.lastrefinedown = 1;
      st.nrej++;
      return SRT_ATTEMPT_RETRY;
    }
    if (st.lastrefinedown == 0 && err < P.maxerr / 100.0 && st.dt * 1.25 < P.dtmax) {  // :804-812
      st.dt = st.dt * 1.25;
      st.lastrefinedown = 0;
    }
  } else {
    // rk4 (:504-532)
    double k2[6], k3[6], k4[6], tmp[6], r[6];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + 1.0 / 2.0 * k1[i];
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k2[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + 1.0 / 2.0 * k2[i];
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k3[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) tmp[i] = xs[i] + k3[i];
    srt_evalrhs<DENS, USE_IGRF>(MR, tmp, tmp + 3, st.w, r);
#pragma unroll
    for (int i = 0; i < 6; ++i) k4[i] = dt * r[i];
#pragma unroll
    for (int i = 0; i < 6; ++i) e2x[i] = xs[i] + 1.0 / 6.0 * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
  }
  const double dtincr = dt;
  // :821 `w = est2(7)`: the local w of raytracer_run is set as soon as an attempt passes the error test -- before the
  // resonance-cone test -- so an attempt that is cone-refined before the first acceptance already ends Q1
  st.wl = st.w;
  // re-project |k| onto the dispersion surface at the new point (:819-836)
  if (!have_c2) srt_centre<DENS, USE_IGRF>(MR, e2x, st.w, c2);
  SrtCplx ksel = srt_solve_root(e2x + 3, st.w, c2.d, c2.s, P.root);
  // k = kmag*(k/sqrt(dot_product(k,k))) in complex arithmetic (Q18)
  double kk = 0.0;
#pragma unroll
  for (int i = 0; i < 3; ++i) kk = kk + (e2x[3 + i] * e2x[3 + i] - (-0.0) * 0.0);
  SrtCplx nrm;
  srt_csqrt(kk, 0.0, nrm.re, nrm.im);
  SrtCplx kn[3];
  double imsum = 0.0;
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    SrtCplx ki = {e2x[3 + i], 0.0};
    kn[i] = srt_cmul(ksel, srt_cdiv(ki, nrm));
    imsum = imsum + kn[i].im * kn[i].im;
  }
  if (imsum > 0.0) {  // :891-906
    if (P.fixedstep == 0) {
      st.dt = st.dt / 2.0;
      st.lastrefinedown = 1;
      st.ncone++;
      return SRT_ATTEMPT_RETRY;
    } else {
      st.fixed_cone_exit = 1;
      st.stopcond = 0;
      return SRT_ATTEMPT_DONE;  // :904
    }
  }
  // accept (:909-913)
#pragma unroll
  for (int i = 0; i < 3; ++i) st.x[i] = e2x[i];
#pragma unroll
  for (int i = 0; i < 3; ++i) st.k[i] = kn[i].re;
  st.lastrefinedown = 0;
  st.t = st.t + dtincr;
  st.nstep = st.nstep + 1;
  st.have_rhs1 = 0;
  // group velocity and recorded plasma at the new point (:916-922)
  srt_dFdk(st.k, st.w, c2.d, c2.s, st.dfdk);
  st.dfdw = srt_dFdw(st.k, st.w, c2.d);
  srt_fill_row(st, c2.pl, true, row);
#pragma unroll
  for (int i = 0; i < 3; ++i) st.vg[i] = row[7 + i];
  return SRT_ATTEMPT_ACCEPTED;
}
