// srt_damping.cu -- Landau / cyclotron damping along the ray (config 5): the reference's damping post-processor
// damping/src/damping_foust.cpp:15-236 ("mode 1") on the GPU, fed with the trajectory rows of the tracer.
//
// What the reference does per recorded row ii >= 1 of a ray (single thread, rays one after another):
//   * local L-shell from the SM position (:104-107), position -> geomagnetic dipole coordinates by the xform_double
//     chain (:114-115), magnetic local time (coord_transforms.cpp:195-210), plasmapause L at that MLT from Kp
//     (bulge.cpp:8-83), a power-law fit of the CRRES suprathermal electron fluxes at (L, MLT, AE)
//     (psd_model.cpp:113-256: nearest bins, log-log least squares over the four energy channels);
//   * k, its components along / across B0, the Stix parameters of the cold plasma (:118-182);
//   * D_i = 50-node Gauss-Legendre quadrature over t in (0,1), v_perp = c (1-t)/t, of the hot-plasma integrand
//     (integrand.cpp:48-160): three resonances m = -1, 0, 1, Bessel functions J_{m-1}, J_m, J_{m+1} of
//     k_perp v_perp / w_ce, and the velocity-space gradients of the phase-space density by central differences
//     (8 density evaluations per resonance: pow / log / exp);
//   * k_i from D_i (:210), projected on the group velocity (:214), power[ii] = geom * power[ii-1] * exp(-2 dist k_i).
// Rows are independent except for the running product, so here: ONE WARP PER ROW for everything up to the row's
// factor exp(-2 dist k_i) -- the 50 quadrature nodes are spread over the lanes (2 per lane), the row's scalar set-up
// is computed redundantly by the lanes -- then ONE THREAD PER RAY multiplies the factors in row order, exactly as
// the reference's loop does.  Bound: FP64 pipe (about 5 400 transcendental calls per row, no memory traffic to
// speak of: 160 B in, 8 B out per row).
//
// Parity: floating point against the reference's own C++ (oracle/_ref/damping, glibc libm, Eigen): the integrand
// differentiates the density with a relative step of 1e-8, which amplifies the last-bit differences between CUDA's
// and glibc's pow/log/exp to ~1e-8 in k_i; tests/test_gpu_damping.py states the tolerance (1e-6 relative on the
// power while it is above the reference's own 1e-6 threshold).  The reference's quirks are kept: milliseconds / 1000 /
// 60 as "hours" and a longitude in radians minus one in degrees in MLT (coord_transforms.cpp:200,209), R_E = 6378 km,
// EPS0 = 8.854e-12, M_EL = 9.1e-31 (damping/include/consts.h), the swapped names of J_{m+1} and J_{m-1}
// (integrand.cpp:64-65).  AE is taken as the number the code expects (the .mat stores it as uint8).
#include <cuda_runtime.h>
#include <float.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/srt.h"
#include "srt_damping_setup.h"

namespace {

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
__constant__ DampSetup c_damp;

struct PsdState {  // psd_model after set_params
  double L_sh, L_pp, n_fit, An_fit;
};

__device__ __forceinline__ double psd_suprathermal(double vperp, double vpar) {  // psd_model.cpp:259-278
  const double a = 4.9e5, b = 8.3e14, c = 5.4e23, v0 = 1;
  double v = 100. * sqrt(vperp * vperp + vpar * vpar + v0);
  double f = (a / pow(v, 4.0) - b / pow(v, 5.0) + c / pow(v, 6.0));
  return f * pow(100., 6.0);
}
__device__ __forceinline__ double psd_crres(const PsdState& s, double vperp, double vpar) {  // :281-305
  const double v0 = 1;
  double v = 100. * sqrt(vperp * vperp + vpar * vpar + v0);
  double f = s.An_fit / pow(v, s.n_fit);
  return f * pow(100., 6.0);
}
__device__ double psd_hybrid(const PsdState& s, double vperp, double vpar) {  // :308-341
  if (s.L_pp - s.L_sh > 1) return psd_suprathermal(vperp, vpar);
  if (s.L_sh - s.L_pp > 1) return psd_crres(s, vperp, vpar);
  double f_polar = psd_suprathermal(vperp, vpar);
  double f_crres = psd_crres(s, vperp, vpar);
  double w_polar = exp(5 * (s.L_pp - s.L_sh)) / (1 + exp(5 * (s.L_pp - s.L_sh)));
  double w_crres = exp(5 * (s.L_sh - s.L_pp)) / (1 + exp(5 * (s.L_sh - s.L_pp)));
  return exp((log(f_polar) * w_polar + log(f_crres) * w_crres) / (w_polar + w_crres));
}

__device__ __forceinline__ double bessel_jn(int n, double x) {
  // jn(n, x) for n in -2..2: J_{-n} = (-1)^n J_n
  int an = n < 0 ? -n : n;
  double v = an == 0 ? j0(x) : (an == 1 ? j1(x) : jn(2, x));
  return (n < 0 && (an & 1)) ? -v : v;
}

struct Integrand {  // integrand::initialize, integrand.cpp:7-45
  double kperp, kpar, w, wch, n_sq, ct, st, st_sq, R, L, P, S;
};

// integrand::evaluate_t (integrand.cpp:48-95)
__device__ double integrand_t(const Integrand& q, const PsdState& psd, double t) {
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

__device__ double bulge(double akp, double amlt) {  // bulge.cpp:8-83
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
__device__ void crres_fit(const double* __restrict__ tab, double L, double MLT, double& n_fit, double& An_fit) {
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

// One warp per row.  rows: [nrows][20] (t, pos, vprel, vgrel, n, B0, Ns4); first[ray] = index of the ray's first row;
// ray_of_row by binary search.  factor[row] = exp(-dist * ki_along_vg * 2) (1 for the first row of a ray);
// geom[row] = r cos(lat) / R_E in MAG (for the geometric factor).
__global__ void srt_damp_row_kernel(long long nrows, long long nrays, const double* __restrict__ rows, const long long* __restrict__ first,
                                    const double* __restrict__ w_of_ray, const double* __restrict__ crres, double* __restrict__ factor,
                                    double* __restrict__ geom) {
  const int lane = threadIdx.x & 31;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long row = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; row < nrows; row += nwarps) {
    // ray of this row: last ray with first[ray] <= row
    long long lo = 0, hi = nrays - 1;
    while (lo < hi) {
      long long mid = (lo + hi + 1) >> 1;
      if (first[mid] <= row) lo = mid; else hi = mid - 1;
    }
    const long long ray = lo;
    const double* r = rows + row * 20;
    const double pos[3] = {r[1], r[2], r[3]};
    // geomagnetic dipole coordinates (sm_to_mag_d, cart_to_pol_d)
    double pm[3];
    for (int i = 0; i < 3; ++i) pm[i] = c_damp.sm2mag[3 * i] * pos[0] + c_damp.sm2mag[3 * i + 1] * pos[1] + c_damp.sm2mag[3 * i + 2] * pos[2];
    const double rho = sqrt(pm[0] * pm[0] + pm[1] * pm[1]);
    const double rad = sqrt(rho * rho + pm[2] * pm[2]);
    const double lat_m = atan2(pm[2], rho);
    double lon_m = atan2(pm[1], pm[0]);
    lon_m = lon_m + (1.0 - copysign(1.0, lon_m)) * 3.141592653589793238462643;
    if (lane == 0) geom[row] = rad * cos(lat_m) / DMP_R_E;
    if (row == first[ray]) {
      if (lane == 0) factor[row] = 1.0;
      continue;
    }
    const double w = w_of_ray[ray];
    const double* rp = r - 20;
    const double B0[3] = {r[13], r[14], r[15]}, nv[3] = {r[10], r[11], r[12]}, vg[3] = {r[7], r[8], r[9]};
    // local L-shell (damping_foust.cpp:104-107)
    const double lat = atan(pos[2] / sqrt(pos[0] * pos[0] + pos[1] * pos[1]));
    const double rr = sqrt(pos[0] * pos[0] + pos[1] * pos[1] + pos[2] * pos[2]);
    const double cl = cos(lat);
    PsdState psd;
    psd.L_sh = rr / (cl * cl) / DMP_R_E;
    // MLT (coord_transforms.cpp:195-210; lon in radians against a longitude in degrees, as the reference does)
    const double mlt = fmod(c_damp.ut_hr + (lon_m - c_damp.greenwich_lon_deg) / 15.0, 24.0);
    psd.L_pp = bulge(c_damp.Kp, mlt);
    crres_fit(crres, psd.L_sh, mlt, psd.n_fit, psd.An_fit);
    const double B0mag = sqrt(B0[0] * B0[0] + B0[1] * B0[1] + B0[2] * B0[2]);
    Integrand q;
    q.wch = DMP_Q_EL * B0mag / DMP_M_EL;
    q.w = w;
    const double k[3] = {nv[0] * w / DMP_C, nv[1] * w / DMP_C, nv[2] * w / DMP_C};
    const double kmag = sqrt(k[0] * k[0] + k[1] * k[1] + k[2] * k[2]);
    const double Bh[3] = {B0[0] / B0mag, B0[1] / B0mag, B0[2] / B0mag};
    q.kpar = k[0] * Bh[0] + k[1] * Bh[1] + k[2] * Bh[2];
    const double kp[3] = {k[0] - q.kpar * Bh[0], k[1] - q.kpar * Bh[1], k[2] - q.kpar * Bh[2]};
    q.kperp = sqrt(kp[0] * kp[0] + kp[1] * kp[1] + kp[2] * kp[2]);
    const double theta = atan2(q.kperp, q.kpar);
    q.st = sin(theta);
    q.ct = cos(theta);
    q.st_sq = q.st * q.st;
    const double sin_th_sq = q.st * q.st, cos_th_sq = q.ct * q.ct;
    const double n = sqrt(nv[0] * nv[0] + nv[1] * nv[1] + nv[2] * nv[2]);
    q.n_sq = n * n;
    double R = 1., L = 1., P = 1.;
    for (int j = 0; j < c_damp.nspec; ++j) {
      const double wps2 = r[16 + j] * (c_damp.qs[j] * c_damp.qs[j]) / (c_damp.ms[j] * DMP_EPS0);
      const double whs = c_damp.qs[j] * B0mag / c_damp.ms[j];
      R -= wps2 / (w * (w + whs));
      L -= wps2 / (w * (w - whs));
      P -= wps2 / (w * w);
    }
    const double S = (R + L) / 2., a = S * sin_th_sq + P * cos_th_sq, b = R * L * sin_th_sq + P * S * (1 + cos_th_sq);
    q.R = R; q.L = L; q.P = P; q.S = S;
    // gauss_legendre(50, f, 0, 1) (gauss_legendre.c:267-287): s += w[i] * (f(B + A x[i]) + f(B - A x[i])), i = 0..24
    const double A = 0.5, Bc = 0.5;
    double fp = 0.0, fm = 0.0;
    if (lane < 25) {
      const double Ax = A * c_damp.glx[lane];
      fp = integrand_t(q, psd, Bc + Ax);
      fm = integrand_t(q, psd, Bc - Ax);
    }
    double term = (lane < 25) ? c_damp.glw[lane] * (fp + fm) : 0.0;
    double s = 0.0;
    for (int i = 0; i < 25; ++i) s += __shfl_sync(0xffffffffu, term, i);  // the reference's order of summation
    const double Di = A * s;
    const double ki = -(w / DMP_C) * (0.5) * (1. / (4. * n * (2. * a * n * n - b))) * Di;
    const double vgn = sqrt(vg[0] * vg[0] + vg[1] * vg[1] + vg[2] * vg[2]);
    const double ki_along_vg = ki * ((k[0] * vg[0] + k[1] * vg[1] + k[2] * vg[2]) / (kmag * vgn));
    const double dp[3] = {pos[0] - rp[1], pos[1] - rp[2], pos[2] - rp[3]};
    const double dist = sqrt(dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2]);
    if (lane == 0) factor[row] = exp(-dist * ki_along_vg * 2.0);
  }
}

// power[ii] = geom_fact * power[ii-1] * exp(...) in row order (damping_foust.cpp:229): one thread per ray
__global__ void srt_damp_scan_kernel(long long nrays, long long nrows, const long long* __restrict__ first, const double* __restrict__ factor,
                                     const double* __restrict__ geom, double* __restrict__ power) {
  for (long long ray = (long long)blockIdx.x * blockDim.x + threadIdx.x; ray < nrays; ray += (long long)gridDim.x * blockDim.x) {
    const long long b = first[ray], e = (ray + 1 < nrays) ? first[ray + 1] : nrows;
    if (e <= b) continue;
    double d = 1.0;
    power[b] = d;
    const double g0 = geom[b];  // r_init cos(lat_init) with r_init in Earth radii
    for (long long i = b + 1; i < e; ++i) {
      const double gf = c_damp.include_geom ? g0 / geom[i] : 1.0;
      d = gf * d * factor[i];
      power[i] = d;
    }
  }
}

std::mutex g_damp_mutex;
std::vector<double> g_crres_raw;  // the whole table file
int g_crres_hdr[4] = {0, 0, 0, 0};
std::string g_crres_path;
thread_local char g_damp_err[256] = "";

int load_crres(const char* path) {
  if (!g_crres_raw.empty() && g_crres_path == path) return SRT_OK;
  FILE* f = fopen(path, "rb");
  if (!f) { snprintf(g_damp_err, sizeof g_damp_err, "cannot open CRRES table %s", path); return SRT_E_IO; }
  int hdr[4];
  if (fread(hdr, 4, 4, f) != 4 || hdr[0] != 0x53525443) { fclose(f); return SRT_E_IO; }
  const size_t per = 2 + (size_t)hdr[2] + hdr[3] + (size_t)hdr[2] * hdr[3];
  std::vector<double> raw(per * hdr[1]);
  if (fread(raw.data(), 8, raw.size(), f) != raw.size()) { fclose(f); return SRT_E_IO; }
  fclose(f);
  g_crres_raw.swap(raw);
  memcpy(g_crres_hdr, hdr, sizeof hdr);
  g_crres_path = path;
  return SRT_OK;
}

}  // namespace

extern "C" {

const char* srt_damping_last_error(void) { return g_damp_err; }

int srt_damping_foust(int64_t nrays, const int64_t* first_row, int64_t nrows, const double* rows, const double* w, int nspec, const double* qs,
                      const double* ms, double Kp, double AE_level, int yearday, int msec, int include_geom_factor, const char* crres_table,
                      double* power) {
  if (nrays < 0 || nrows < 0 || nspec < 1 || nspec > 4 || (nrows > 0 && (!first_row || !rows || !w || !qs || !ms || !power)) || !crres_table) return SRT_E_ARG;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
    snprintf(g_damp_err, sizeof g_damp_err, "no CUDA device");
    return SRT_E_NODEVICE;
  }
  if (nrows == 0) return SRT_OK;
  std::lock_guard<std::mutex> lk(g_damp_mutex);
  int rc = load_crres(crres_table);
  if (rc) return rc;
  DampSetup S;
  memset(&S, 0, sizeof S);
  srt_host::gauss_legendre_tbl(50, S.glx, S.glw, 1e-10);
  int itime[2] = {yearday, msec};
  srt_host::sm_to_mag_matrix(itime, S.sm2mag);
  S.greenwich_lon_deg = srt_host::greenwich_mag_lon_deg(itime);
  S.ut_hr = msec / 1000.0 / 60.0;
  S.Kp = Kp;
  S.AE_target = round(AE_level);
  if (S.AE_target < 1 || S.AE_target > 3) { snprintf(g_damp_err, sizeof g_damp_err, "AE level must round to 1, 2 or 3"); return SRT_E_ARG; }
  S.include_geom = include_geom_factor ? 1 : 0;
  S.nspec = nspec;
  for (int i = 0; i < nspec; ++i) { S.qs[i] = qs[i]; S.ms[i] = ms[i]; }
  // the energy channels of this AE level, in file order (psd_model.cpp:158-215)
  const int n = g_crres_hdr[1], nL = g_crres_hdr[2], nM = g_crres_hdr[3];
  const size_t per = 2 + (size_t)nL + nM + (size_t)nL * nM;
  std::vector<double> tab;
  S.nL = nL; S.nM = nM; S.nE = 0;
  std::vector<double> J;
  for (int k = 0; k < n; ++k) {
    const double* e = g_crres_raw.data() + per * k;
    if (e[1] != S.AE_target) continue;
    if (S.nE == 0) tab.assign(e + 2, e + 2 + nL + nM);
    if (S.nE < 4) {
      S.logE[S.nE++] = log(e[0]);
      J.insert(J.end(), e + 2 + nL + nM, e + per);
    }
  }
  if (S.nE < 2) { snprintf(g_damp_err, sizeof g_damp_err, "CRRES table has no entries for AE level %g", S.AE_target); return SRT_E_IO; }
  tab.insert(tab.end(), J.begin(), J.end());

  double *d_rows = nullptr, *d_w = nullptr, *d_tab = nullptr, *d_factor = nullptr, *d_geom = nullptr, *d_power = nullptr;
  long long* d_first = nullptr;
  cudaError_t e = cudaSuccess;
#define DMP_CK(call) do { if (e == cudaSuccess) e = (call); } while (0)
  DMP_CK(cudaMalloc(&d_rows, (size_t)nrows * 20 * 8));
  DMP_CK(cudaMalloc(&d_w, (size_t)(nrays ? nrays : 1) * 8));
  DMP_CK(cudaMalloc(&d_first, (size_t)(nrays ? nrays : 1) * 8));
  DMP_CK(cudaMalloc(&d_tab, tab.size() * 8));
  DMP_CK(cudaMalloc(&d_factor, (size_t)nrows * 8));
  DMP_CK(cudaMalloc(&d_geom, (size_t)nrows * 8));
  DMP_CK(cudaMalloc(&d_power, (size_t)nrows * 8));
  DMP_CK(cudaMemcpy(d_rows, rows, (size_t)nrows * 20 * 8, cudaMemcpyHostToDevice));
  DMP_CK(cudaMemcpy(d_w, w, (size_t)nrays * 8, cudaMemcpyHostToDevice));
  DMP_CK(cudaMemcpy(d_first, first_row, (size_t)nrays * 8, cudaMemcpyHostToDevice));
  DMP_CK(cudaMemcpy(d_tab, tab.data(), tab.size() * 8, cudaMemcpyHostToDevice));
  DMP_CK(cudaMemcpyToSymbol(c_damp, &S, sizeof S));
  if (e == cudaSuccess) {
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long warps = nrows;
    int blocks = (int)((warps * 32 + 255) / 256 < (long long)sms * 8 ? (warps * 32 + 255) / 256 : (long long)sms * 8);
    srt_damp_row_kernel<<<blocks, 256>>>(nrows, nrays, d_rows, d_first, d_w, d_tab, d_factor, d_geom);
    int sb = (int)((nrays + 127) / 128 < (long long)sms * 8 ? (nrays + 127) / 128 : (long long)sms * 8);
    srt_damp_scan_kernel<<<sb < 1 ? 1 : sb, 128>>>(nrays, nrows, d_first, d_factor, d_geom, d_power);
    DMP_CK(cudaGetLastError());
    DMP_CK(cudaMemcpy(power, d_power, (size_t)nrows * 8, cudaMemcpyDeviceToHost));
  }
#undef DMP_CK
  cudaFree(d_rows); cudaFree(d_w); cudaFree(d_first); cudaFree(d_tab); cudaFree(d_factor); cudaFree(d_geom); cudaFree(d_power);
  if (e != cudaSuccess) { snprintf(g_damp_err, sizeof g_damp_err, "CUDA: %s", cudaGetErrorString(e)); return SRT_E_CUDA; }
  return SRT_OK;
}

// The same for the rays of a batch result (rows in ray order: ray 0's rows, then ray 1's, ...).  first_row[nrays] and
// power[srt_result_total_rows] are caller-allocated.
int srt_result_damping(const srt_result* r, const srt_model* m, const double* w0, double Kp, double AE_level, int yearday, int msec,
                       int include_geom_factor, const char* crres_table, int64_t* first_row, double* power) {
  if (!r || !m || !w0 || !first_row || !power) return SRT_E_ARG;
  const int64_t nrays = srt_result_nrays(r), total = srt_result_total_rows(r);
  int nspec = 0;
  double qs[4], ms[4], nus[4];
  int rc = srt_model_species(m, &nspec, qs, ms, nus);
  if (rc) return rc;
  std::vector<double> rows((size_t)total * SRT_ROW_DOUBLES);
  int64_t at = 0;
  for (int64_t i = 0; i < nrays; ++i) {
    const double* p = nullptr;
    int64_t n = 0;
    if ((rc = srt_result_ray_view(r, i, &p, &n))) return rc;
    first_row[i] = at;
    if (n) memcpy(rows.data() + (size_t)at * SRT_ROW_DOUBLES, p, (size_t)n * SRT_ROW_DOUBLES * sizeof(double));
    at += n;
  }
  if (at != total) return SRT_E_ARG;
  return srt_damping_foust(nrays, first_row, total, rows.data(), w0, nspec, qs, ms, Kp, AE_level, yearday, msec, include_geom_factor, crres_table, power);
}

}  // extern "C"
