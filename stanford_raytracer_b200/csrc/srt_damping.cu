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
#include "srt_damping_row.h"
#include "srt_damping_setup.h"

namespace {

__constant__ DampSetup c_damp;

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
    const bool is_first = (row == first[ray]);
    DampRow R;
    damp_row_setup(c_damp, crres, r, is_first ? nullptr : r - 20, w_of_ray[ray], R);
    if (lane == 0) geom[row] = R.geom;
    if (is_first) {
      if (lane == 0) factor[row] = 1.0;
      continue;
    }
    // gauss_legendre(50, f, 0, 1) (gauss_legendre.c:267-287): s += w[i] * (f(B + A x[i]) + f(B - A x[i])), i = 0..24
    const double A = 0.5, Bc = 0.5;
    double fp = 0.0, fm = 0.0;
    if (lane < 25) {
      const double Ax = A * c_damp.glx[lane];
      fp = integrand_t(R.q, R.psd, Bc + Ax);
      fm = integrand_t(R.q, R.psd, Bc - Ax);
    }
    const double term = (lane < 25) ? c_damp.glw[lane] * (fp + fm) : 0.0;
    double s = 0.0;
    for (int i = 0; i < 25; ++i) s += __shfl_sync(0xffffffffu, term, i);  // the reference's order of summation
    if (lane == 0) factor[row] = damp_row_factor(R, A * s);
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
thread_local char g_damp_err[256] = "";

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
  DampSetup S;
  std::vector<double> tab;
  int rc = srt_host::damp_setup(crres_table, nspec, qs, ms, Kp, AE_level, yearday, msec, include_geom_factor, S, tab, g_damp_err, sizeof g_damp_err);
  if (rc == -1) return SRT_E_IO;
  if (rc) return SRT_E_ARG;

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
