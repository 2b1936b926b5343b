// oracle/orc_raytracer.h -- TEST INFRASTRUCTURE ONLY.  CPU restatement of module
// `raytracer` (fortran/raytracer.f95), statement by statement, with the
// reference's call structure kept (nothing de-duplicated): 8 plasma + 14
// dispersion evaluations per raytracer_evalrhs.  PARITY UNPINNED (see
// orc_common.h).
#pragma once
#include "orc_common.h"

namespace orc {

// sum()/dot_product() as gfortran inlines them: accumulator starts at 0
inline double dot3(const double a[3], const double b[3]) {
  double s = 0.0;
  for (int i = 0; i < 3; ++i) s = s + a[i] * b[i];
  return s;
}

// fortran/raytracer.f95:81-102
inline void stix_parameters(double w, const Plasma& p, double B0mag, double& S, double& D, double& P, double& R, double& L) {
  double wps2[4], wcs[4];
  for (int i = 0; i < p.nspec; ++i) {
    wps2[i] = (p.Ns[i] * (p.qs[i] * p.qs[i]) / p.ms[i] / EPS0);  // :92
    wcs[i] = ((p.qs[i] * B0mag) / p.ms[i]);                      // :93
  }
  double sr = 0.0, sl = 0.0, sp = 0.0;
  for (int i = 0; i < p.nspec; ++i) sr = sr + wps2[i] / (w * (w + wcs[i]));
  for (int i = 0; i < p.nspec; ++i) sl = sl + wps2[i] / (w * (w - wcs[i]));
  for (int i = 0; i < p.nspec; ++i) sp = sp + wps2[i] / (w * w);
  R = 1.0 - sr;  // :96
  L = 1.0 - sl;  // :97
  P = 1.0 - sp;  // :98
  S = 1.0 / 2.0 * (R + L);  // :99
  D = 1.0 / 2.0 * (R - L);  // :100
}

// fortran/raytracer.f95:41-72
inline double dispersion_relation(const double n[3], double w, const Plasma& p) {
  double nmag2 = dot3(n, n);
  double cos2phi = (dot3(n, p.B0) * dot3(n, p.B0)) / (dot3(n, n) * dot3(p.B0, p.B0));  // :52
  double sin2phi = 1.0 - cos2phi;
  double S, D, P, R, L;
  stix_parameters(w, p, m_sqrt(dot3(p.B0, p.B0)), S, D, P, R, L);
  double A = S * sin2phi + P * cos2phi;                    // :60
  double B = R * L * sin2phi + P * S * (1.0 + cos2phi);    // :61
  double maxNs = p.Ns[0], maxq = fabs(p.qs[0]), minm = p.ms[0];
  for (int i = 1; i < p.nspec; ++i) {
    if (p.Ns[i] > maxNs) maxNs = p.Ns[i];
    if (fabs(p.qs[i]) > maxq) maxq = fabs(p.qs[i]);
    if (p.ms[i] < minm) minm = p.ms[i];
  }
  // :65 -- the square root covers only the numerator (Q20)
  if (w > 100.0 * m_sqrt((maxNs * (maxq * maxq))) / (minm * EPS0)) {
    return -nmag2 + 1.0;
  }
  return A * (nmag2 * nmag2) - B * nmag2 + R * L * P;  // :68
}

inline double fmax_del(double a, double b) { return (a > b) ? a : b; }  // max(del*abs(v), del)

// fortran/raytracer.f95:118-155
inline void dispersion_relation_dFdk(const double k[3], double w, const double x[3], double del, Model& m, double out[3]) {
  Plasma p;
  m.plasma(x, p);
  const double C = C_light();
  for (int i = 0; i < 3; ++i) {
    double d = fmax_del(del * fabs(k[i]), del);
    double dk[3] = {0.0, 0.0, 0.0};
    for (int j = 0; j < 3; ++j) dk[j] = d * ((j == i) ? 1.0 : 0.0);
    double np[3], nn[3];
    for (int j = 0; j < 3; ++j) {
      np[j] = (k[j] + dk[j]) * C / w;
      nn[j] = (k[j] - dk[j]) * C / w;
    }
    out[i] = (dispersion_relation(np, w, p) - dispersion_relation(nn, w, p)) / d / 2.0;
  }
}

// fortran/raytracer.f95:172-198
inline double dispersion_relation_dFdw(const double k[3], double w, const double x[3], double del, Model& m) {
  Plasma p;
  m.plasma(x, p);
  const double C = C_light();
  double d = fmax_del(del * fabs(w), del);
  double np[3], nn[3];
  for (int j = 0; j < 3; ++j) {
    np[j] = k[j] * C / (w + d);
    nn[j] = k[j] * C / (w - d);
  }
  return (dispersion_relation(np, (w + d), p) - dispersion_relation(nn, (w - d), p)) / d / 2.0;
}

// fortran/raytracer.f95:215-263
inline void dispersion_relation_dFdx(const double k[3], double w, const double x[3], double del, Model& m, double out[3]) {
  const double C = C_light();
  double n[3];
  for (int j = 0; j < 3; ++j) n[j] = k[j] * C / w;
  Plasma p;
  for (int i = 0; i < 3; ++i) {
    double d = fmax_del(del * fabs(x[i]), del);
    double xp[3], xn[3];
    for (int j = 0; j < 3; ++j) {
      double dx = d * ((j == i) ? 1.0 : 0.0);
      xp[j] = x[j] + dx;
      xn[j] = x[j] - dx;
    }
    m.plasma(xp, p);
    double Fp = dispersion_relation(n, w, p);
    m.plasma(xn, p);
    double Fn = dispersion_relation(n, w, p);
    out[i] = (Fp - Fn) / d / 2.0;
  }
}

// fortran/raytracer.f95:282-314
inline void raytracer_evalrhs(const double args[7], double del, Model& m, double rhs[7]) {
  const double* x = args;
  const double* k = args + 3;
  double w = args[6];
  double dfdk[3], dfdx[3];
  dispersion_relation_dFdk(k, w, x, 1.0e-8, m, dfdk);
  double dfdw = dispersion_relation_dFdw(k, w, x, 1.0e-8, m);
  dispersion_relation_dFdx(k, w, x, del, m, dfdx);
  for (int i = 0; i < 3; ++i) rhs[i] = -(dfdk[i] / dfdw);
  for (int i = 0; i < 3; ++i) rhs[3 + i] = dfdx[i] / dfdw;
  rhs[6] = 0.0;
}

// fortran/raytracer.f95:324-353
inline int raytracer_stopconditions(const double pos[3], const double k[3], const double vgrel[3], double dt, int nstep, int maxsteps, double minalt) {
  if (m_sqrt(dot3(pos, pos)) < minalt) return 1;
  if (m_sqrt(dot3(k, k)) == 0.0) return 2;
  if (m_sqrt(dot3(vgrel, vgrel)) > 1.0 + 1e-2) return 3;  // 1.0_DP+1e-2_DP
  if (dt < (double)1e-14f) return 5;  // single-precision literal (Q2)
  if (nstep >= maxsteps) return 6;
  return 0;
}

// 3x3 complex SVD by one-sided (Hestenes) Jacobi: stands in for LAPACK ZGESVD
// (fortran/blas.f95:180-221) -- the only thing used from it is the right singular
// vector of the smallest singular value, up to a unit phase.
inline void csvd3_smallest_right_vector(const cplx Min[3][3], cplx vout[3], double sv[3]) {
  cplx U[3][3], V[3][3];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      U[i][j] = Min[i][j];
      V[i][j] = {i == j ? 1.0 : 0.0, 0.0};
    }
  for (int sweep = 0; sweep < 60; ++sweep) {
    double off = 0.0;
    for (int p = 0; p < 2; ++p)
      for (int q = p + 1; q < 3; ++q) {
        double alpha = 0, beta = 0;
        cplx gamma = {0, 0};
        for (int i = 0; i < 3; ++i) {
          alpha += U[i][p].re * U[i][p].re + U[i][p].im * U[i][p].im;
          beta += U[i][q].re * U[i][q].re + U[i][q].im * U[i][q].im;
          // conj(U[i][p]) * U[i][q]
          gamma.re += U[i][p].re * U[i][q].re + U[i][p].im * U[i][q].im;
          gamma.im += U[i][p].re * U[i][q].im - U[i][p].im * U[i][q].re;
        }
        double g = ::sqrt(gamma.re * gamma.re + gamma.im * gamma.im);
        if (g == 0.0 || g <= 1e-300) continue;
        double lim = ::sqrt(alpha) * ::sqrt(beta);
        if (g <= 4e-16 * lim) continue;
        off += g / (lim > 0 ? lim : 1.0);
        cplx ph = {gamma.re / g, -gamma.im / g};  // e^{-i phi}
        double zeta = (beta - alpha) / (2.0 * g);
        double t = (zeta >= 0 ? 1.0 : -1.0) / (fabs(zeta) + ::sqrt(1.0 + zeta * zeta));
        double c = 1.0 / ::sqrt(1.0 + t * t), s = c * t;
        for (int i = 0; i < 3; ++i) {
          cplx up = U[i][p], uq = c_mul(ph, U[i][q]);
          U[i][p] = {c * up.re - s * uq.re, c * up.im - s * uq.im};
          U[i][q] = {s * up.re + c * uq.re, s * up.im + c * uq.im};
          cplx vp = V[i][p], vq = c_mul(ph, V[i][q]);
          V[i][p] = {c * vp.re - s * vq.re, c * vp.im - s * vq.im};
          V[i][q] = {s * vp.re + c * vq.re, s * vp.im + c * vq.im};
        }
      }
    if (off == 0.0) break;
  }
  int imin = 0;
  for (int j = 0; j < 3; ++j) {
    double s2 = 0;
    for (int i = 0; i < 3; ++i) s2 += U[i][j].re * U[i][j].re + U[i][j].im * U[i][j].im;
    sv[j] = ::sqrt(s2);
    if (sv[j] < sv[imin]) imin = j;
  }
  for (int i = 0; i < 3; ++i) vout[i] = V[i][imin];
}

// fortran/raytracer.f95:355-371 -- cmplx() without kind= rounds to single (Q8);
// phi arrives in DEGREES and is fed to cos/sin as radians (Q9).
inline void form_dispersion_matrix(cplx M[3][3], double n2, double phi, double S, double D, double P) {
  auto f = [](double v) { return (double)(float)v; };
  double cphi = m_cos(phi), sphi = m_sin(phi);
  M[0][0] = {f(S - n2 * (cphi * cphi)), 0.0};
  M[0][1] = {0.0, f(-D)};
  M[0][2] = {f(n2 * cphi * sphi), 0.0};
  M[1][0] = {0.0, f(D)};
  M[1][1] = {f(S - n2), 0.0};
  M[1][2] = {0.0, 0.0};
  M[2][0] = {f(n2 * cphi * sphi), 0.0};
  M[2][1] = {0.0, 0.0};
  M[2][2] = {f(P - n2 * (sphi * sphi)), 0.0};
}

// fortran/raytracer.f95:373-405
inline bool is_right_handed(double n2, double phi, double S, double D, double P) {
  cplx M[3][3], v[3];
  double sv[3];
  form_dispersion_matrix(M, n2, phi, S, D, P);
  csvd3_smallest_right_vector(M, v, sv);
  // E = VT(3,:) = conjg(v)
  double E0[3], E90[3];
  for (int i = 0; i < 3; ++i) {
    cplx E = {v[i].re, -v[i].im};
    E0[i] = E.re;     // real(E*(1,0))
    E90[i] = -E.im;   // real(E*(0,1))
  }
  // the sign decision is made with libm-independent atan2 (both libm's agree on sign)
  double angle = ::atan2(E90[1], E90[0]) - ::atan2(E0[1], E0[0]);
  if (angle > PI) angle = angle - 2.0 * PI;
  else if (angle < -PI) angle = angle + 2.0 * PI;
  return !(angle < 0.0);
}

// fortran/raytracer.f95:408-502
inline void solve_dispersion_relation(const double k[3], double w, const double x[3], cplx& k1, cplx& k2, Model& m) {
  Plasma p;
  m.plasma(x, p);
  const double C = C_light();
  double cos2phi = (dot3(k, p.B0) * dot3(k, p.B0)) / (dot3(k, k) * dot3(p.B0, p.B0));
  double sin2phi = 1.0 - cos2phi;
  double phi = m_acos(m_sqrt(cos2phi)) * R2D;
  double B0mag = m_sqrt(dot3(p.B0, p.B0));
  double S, D, P, R, L;
  stix_parameters(w, p, B0mag, S, D, P, R, L);
  double A = S * sin2phi + P * cos2phi;
  double B = R * L * sin2phi + P * S * (1.0 + cos2phi);
  cplx disc = {B * B - 4.0 * A * R * L * P, 0.0};
  cplx sq = c_sqrt(disc);
  cplx twoA = {2.0 * A, 0.0};
  cplx nsq1 = c_div({B + sq.re, sq.im}, twoA);
  cplx nsq2 = c_div({B - sq.re, -sq.im}, twoA);
  cplx n1 = c_sqrt(nsq1);
  cplx n2 = c_sqrt(nsq2);
  k1 = {w * n1.re / C, w * n1.im / C};
  k2 = {w * n2.re / C, w * n2.im / C};
  if (n1.re > 0.0 && is_right_handed(nsq1.re, phi, S, D, P)) {
    k1 = {w * n2.re / C, w * n2.im / C};
    k2 = {w * n1.re / C, w * n1.im / C};
  }
}

static const double rk45_c[6] = {0.0, 1.0 / 4.0, 3.0 / 8.0, 12.0 / 13.0, 1.0, 1.0 / 2.0};
static const double rk45_a2[1] = {1.0 / 4.0};
static const double rk45_a3[2] = {3.0 / 32.0, 9.0 / 32.0};
static const double rk45_a4[3] = {1932.0 / 2197.0, -7200.0 / 2197.0, 7296.0 / 2197.0};
static const double rk45_a5[4] = {439.0 / 216.0, -8.0, 3680.0 / 513.0, -845.0 / 4104.0};
static const double rk45_a6[5] = {-8.0 / 27.0, 2.0, -3544.0 / 2565.0, 1859.0 / 4104.0, -11.0 / 40.0};
static const double rk45_b4[6] = {25.0 / 216.0, 0.0, 1408.0 / 2565.0, 2197.0 / 4104.0, -1.0 / 5.0, 0.0};
static const double rk45_b5[6] = {16.0 / 135.0, 0.0, 6656.0 / 12825.0, 28561.0 / 56430.0, -9.0 / 50.0, 2.0 / 55.0};

// fortran/raytracer.f95:504-532
inline void rk4(const double x[7], double del, double dt, Model& m, double out[7]) {
  double k1[7], k2[7], k3[7], k4[7], tmp[7], r[7];
  raytracer_evalrhs(x, del, m, r);
  for (int i = 0; i < 7; ++i) k1[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + 1.0 / 2.0 * k1[i];
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k2[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + 1.0 / 2.0 * k2[i];
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k3[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + k3[i];
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k4[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) out[i] = x[i] + 1.0 / 6.0 * (k1[i] + 2.0 * k2[i] + 2.0 * k3[i] + k4[i]);
}

// fortran/raytracer.f95:534-596
inline void rk45(const double x[7], double del, double dt, Model& m, double out4[7], double out5[7]) {
  double k1[7], k2[7], k3[7], k4[7], k5[7], k6[7], tmp[7], r[7];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i];
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k1[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + (rk45_a2[0] * k1[i]);
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k2[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + (rk45_a3[0] * k1[i] + rk45_a3[1] * k2[i]);
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k3[i] = dt * r[i];
  for (int i = 0; i < 7; ++i) tmp[i] = x[i] + (rk45_a4[0] * k1[i] + rk45_a4[1] * k2[i] + rk45_a4[2] * k3[i]);
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k4[i] = dt * r[i];
  for (int i = 0; i < 7; ++i)
    tmp[i] = x[i] + (rk45_a5[0] * k1[i] + rk45_a5[1] * k2[i] + rk45_a5[2] * k3[i] + rk45_a5[3] * k4[i]);
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k5[i] = dt * r[i];
  for (int i = 0; i < 7; ++i)
    tmp[i] = x[i] + (rk45_a6[0] * k1[i] + rk45_a6[1] * k2[i] + rk45_a6[2] * k3[i] + rk45_a6[3] * k4[i] + rk45_a6[4] * k5[i]);
  raytracer_evalrhs(tmp, del, m, r);
  for (int i = 0; i < 7; ++i) k6[i] = dt * r[i];
  for (int i = 0; i < 7; ++i)
    out4[i] = x[i] + (rk45_b4[0] * k1[i] + rk45_b4[1] * k2[i] + rk45_b4[2] * k3[i] + rk45_b4[3] * k4[i] + rk45_b4[4] * k5[i] +
                      rk45_b4[5] * k6[i]);
  for (int i = 0; i < 7; ++i)
    out5[i] = x[i] + (rk45_b5[0] * k1[i] + rk45_b5[1] * k2[i] + rk45_b5[2] * k3[i] + rk45_b5[3] * k4[i] + rk45_b5[4] * k5[i] +
                      rk45_b5[5] * k6[i]);
}

// util.f95:109-122
inline void cartesian_to_spherical(const double x[3], double p[3]) {
  double s = 0.0;
  for (int i = 0; i < 3; ++i) s = s + x[i] * x[i];
  p[0] = m_sqrt(s);
  p[1] = m_atan2(x[1], x[0]);
  if (p[0] != 0.0) p[2] = m_acos((x[2]) / (p[0]));
  else p[2] = 0.0;
}
// the matrix A of util.f95:138-140 / 156-158, filled column-major by reshape
inline void sph_matrix(double theta, double phi, double A[3][3]) {
  double ct = m_cos(theta), st = m_sin(theta), cp = m_cos(phi), sp = m_sin(phi);
  // reshape((/ a1..a9 /),(/3,3/)): A(1,1)=a1 A(2,1)=a2 A(3,1)=a3 A(1,2)=a4 ...
  A[0][0] = ct * sp; A[1][0] = -st; A[2][0] = ct * cp;
  A[0][1] = st * sp; A[1][1] = ct;  A[2][1] = st * cp;
  A[0][2] = cp;      A[1][2] = 0.0; A[2][2] = -sp;
}
// util.f95:126-142  matmul(transpose(A), p)
inline void spherical_to_cartesian_vec(const double p[3], double theta, double phi, double out[3]) {
  double A[3][3];
  sph_matrix(theta, phi, A);
  for (int i = 0; i < 3; ++i) {
    double s = 0.0;
    for (int j = 0; j < 3; ++j) s = s + A[j][i] * p[j];
    out[i] = s;
  }
}
// util.f95:148-162  matmul(A, p)
inline void cartesian_to_spherical_vec(const double p[3], double theta, double phi, double out[3]) {
  double A[3][3];
  sph_matrix(theta, phi, A);
  for (int i = 0; i < 3; ++i) {
    double s = 0.0;
    for (int j = 0; j < 3; ++j) s = s + A[i][j] * p[j];
    out[i] = s;
  }
}

// One recorded row (driver :1197-1217 writes these columns)
struct Row {
  double t, pos[3], vprel[3], vgrel[3], n[3], B0[3], qs[4], ms[4], Ns[4], nus[4];
};

struct RunResult {
  std::vector<Row> rows;
  int stopcond = 0;
  long n_rejected = 0;   // err > maxerr retries (:790-798)
  long n_cone = 0;       // resonance-cone refinements (:891-899)
  int fixed_cone_exit = 0;  // Q6 side channel
  double dt_final = 0;
};

// fortran/raytracer.f95:609-995
inline void raytracer_run(RunResult& out, const double pos0[3], double dir0[3], double w0, double dt0, double dtmax, double maxerr,
                          int maxsteps, double minalt, int root, double tmax, int fixedstep, double del, Model& m) {
  const double C = C_light();
  Plasma ptmp;
  double p[3];
  if (dir0[0] == 0 && dir0[1] == 0 && dir0[2] == 0) {  // :661-674
    m.plasma(pos0, ptmp);
    cartesian_to_spherical(pos0, p);
    double Bs[3], Bc[3];
    cartesian_to_spherical_vec(ptmp.B0, p[1], p[2], Bs);
    Bs[0] = fabs(Bs[0]);
    spherical_to_cartesian_vec(Bs, p[1], p[2], Bc);
    double nb = m_sqrt(dot3(Bc, Bc));
    for (int i = 0; i < 3; ++i) dir0[i] = Bc[i] / nb;
  }
  cplx k1mag, k2mag, k0mag;
  solve_dispersion_relation(dir0, w0, pos0, k1mag, k2mag, m);  // :683
  k0mag = (root == 1) ? k1mag : k2mag;
  double x[7], est1[7], est2[7];
  for (int i = 0; i < 3; ++i) x[i] = pos0[i];
  for (int i = 0; i < 3; ++i) x[3 + i] = k0mag.re * dir0[i] - k0mag.im * 0.0;  // real(k0mag*dir0): the full complex product with (dir0, +0.0)
  x[6] = w0;
  double dt = dt0, t = 0.0;
  int lastrefinedown = 0;
  double dfdk[3], dfdw;
  dispersion_relation_dFdk(x + 3, w0, x, 1.0e-8, m, dfdk);  // :700
  dfdw = dispersion_relation_dFdw(x + 3, w0, x, 1.0e-8, m);
  m.plasma(x, ptmp);  // :706
  Row r;
  auto fill_plasma = [&](Row& rr) {
    for (int i = 0; i < 3; ++i) rr.B0[i] = ptmp.B0[i];
    for (int i = 0; i < 4; ++i) {
      rr.qs[i] = ptmp.qs[i]; rr.ms[i] = ptmp.ms[i]; rr.Ns[i] = ptmp.Ns[i]; rr.nus[i] = ptmp.nus[i];
    }
  };
  for (int i = 0; i < 3; ++i) r.pos[i] = x[i];
  r.t = t;
  for (int i = 0; i < 3; ++i) r.n[i] = x[3 + i] * C / w0;
  if (dot3(r.n, r.n) > 0) {  // :725
    double nn = dot3(r.n, r.n);
    for (int i = 0; i < 3; ++i) r.vprel[i] = r.n[i] / nn;
    for (int i = 0; i < 3; ++i) r.vgrel[i] = -(dfdk[i] / dfdw) / C;
  } else {
    for (int i = 0; i < 3; ++i) r.vprel[i] = r.vgrel[i] = 0.0;
  }
  fill_plasma(r);
  out.rows.push_back(r);
  out.stopcond = 0;
  int nstep = 1;
  double w = 0.0;  // Q1: uninitialised local, zero under -finit-local-zero
  for (;;) {
    if (t >= tmax) { out.stopcond = 0; break; }  // :749
    const Row& last = out.rows.back();
    out.stopcond = raytracer_stopconditions(x, x + 3, last.vgrel, dt, nstep, maxsteps, minalt);  // :756
    if (out.stopcond != 0) break;
    double dtincr;
    if (fixedstep == 0) {
      rk45(x, del, dt, m, est1, est2);  // :772
      dtincr = dt;
      double d1[3], d2[3];
      dispersion_relation_dFdk(est1 + 3, w, est1, 1.0e-8, m, d1);  // :778
      dispersion_relation_dFdk(est2 + 3, w, est2, 1.0e-8, m, d2);
      double sa = 0.0, sb = 0.0, sc = 0.0, sd = 0.0;
      for (int i = 0; i < 3; ++i) sa = sa + fabs(est1[3 + i] - est2[3 + i]);
      for (int i = 0; i < 3; ++i) sb = sb + fabs(est2[3 + i]);
      for (int i = 0; i < 3; ++i) sc = sc + fabs(d1[i] - d2[i]);
      for (int i = 0; i < 3; ++i) sd = sd + fabs(d2[i]);
      double e1 = sa / sb, e2 = sc / sd;
      // MAX(a1,a2) as gfortran <= 8 expands it: m=a1; if (a2 > m .or. isnan(m)) m = a2  (Q1)
      double err = e1;
      if (e2 > err || err != err) err = e2;
      if (err > maxerr) {  // :790
        dt = 0.8 * dt;
        lastrefinedown = 1;
        out.n_rejected++;
        continue;
      }
      if (lastrefinedown == 0 && err < maxerr / 100.0 && dt * 1.25 < dtmax) {  // :804
        dt = dt * 1.25;
        lastrefinedown = 0;
      }
    } else {
      rk4(x, del, dt, m, est2);  // :815
      dtincr = dt;
    }
    double cur_pos[3] = {est2[0], est2[1], est2[2]};
    cplx k[3] = {{est2[3], 0.0}, {est2[4], 0.0}, {est2[5], 0.0}};
    w = est2[6];
    double kre[3] = {k[0].re, k[1].re, k[2].re};
    solve_dispersion_relation(kre, w, cur_pos, k1mag, k2mag, m);  // :826
    cplx ksel = (root == 1) ? k1mag : k2mag;
    // k = kmag*(k/sqrt(dot_product(k,k)))  (complex; dot_product conjugates its first argument)
    cplx dp = {0.0, 0.0};
    for (int i = 0; i < 3; ++i) {
      cplx cj = {k[i].re, -k[i].im};
      cplx pr = c_mul(cj, k[i]);
      dp = {dp.re + pr.re, dp.im + pr.im};
    }
    cplx nrm = c_sqrt(dp);
    for (int i = 0; i < 3; ++i) k[i] = c_mul(ksel, c_div(k[i], nrm));
    double imsum = 0.0;
    for (int i = 0; i < 3; ++i) imsum = imsum + k[i].im * k[i].im;
    if (imsum > 0.0) {  // :891
      if (fixedstep == 0) {
        dt = dt / 2.0;
        lastrefinedown = 1;
        out.n_cone++;
        continue;
      } else {
        out.fixed_cone_exit = 1;
        out.dt_final = dt;
        return;  // :904 (stopcond stays 0, Q6)
      }
    }
    for (int i = 0; i < 7; ++i) x[i] = est2[i];  // :909
    for (int i = 0; i < 3; ++i) x[3 + i] = k[i].re;
    lastrefinedown = 0;
    t = t + dtincr;
    nstep = nstep + 1;
    dispersion_relation_dFdk(x + 3, w, x, 1.0e-8, m, dfdk);  // :916
    dfdw = dispersion_relation_dFdw(x + 3, w, x, 1.0e-8, m);
    m.plasma(x, ptmp);
    for (int i = 0; i < 3; ++i) r.pos[i] = x[i];
    r.t = t;
    for (int i = 0; i < 3; ++i) r.n[i] = x[3 + i] * C / w;
    {
      double nn = dot3(r.n, r.n);
      for (int i = 0; i < 3; ++i) r.vprel[i] = r.n[i] / nn;
    }
    for (int i = 0; i < 3; ++i) r.vgrel[i] = -(dfdk[i] / dfdw) / C;
    fill_plasma(r);
    out.rows.push_back(r);
  }
  out.dt_final = dt;
}

}  // namespace orc
