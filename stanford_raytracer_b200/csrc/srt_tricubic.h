// srt_tricubic.h -- gridded log-density by Lekien-Marsden tricubic interpolation:
// tricubic_interpolate_at (tricubic-for/libtricubic.f95:796-933) as called once per species from
// fortran/interp_dens_model_adapter.f95:190-206, followed by Ns = exp(.) (:209).
//
// What the reference does per query and species: three O(n) maxloc scans for the cell, a gather of
// 8 corners x 8 arrays, a dense 64x64 matmul with Amat, a 64-term polynomial.  What this does:
//   * cell index from the regular axes (one division + a fix-up against the stored node
//     coordinates, so the comparison `0 <= xi - x(ind)` is the reference's);
//   * the 64 coefficients x 4 species are PRECOMPUTED per cell (SrtGrid::coef, 2 KiB per cell, built once by
//     srt_coef_kernel with srt_tricubic_cell_coef_ext below, the same accumulation order as the reference's
//     matmul), so a query is one 2 KiB gather served from L2 plus the polynomial;
//   * the table covers the EXTENDED lattice (nx+1)(ny+1)(nz+1): the cells touching or outside the grid edge too,
//     with the sticky clamp flags applied when their coefficients are built.  This matters for
//     speed, not only for completeness: about 1 % of the config-3 rays turn NaN and are then stepped to maxsteps
//     by the reference's controller (no stop condition catches a NaN); a NaN position locates to cell (0,0,0), and
//     on the on-the-fly path (a 64 x 64 product per species and query) each such ray held up its whole block in
//     every round;
//   * without a table (SrtGrid::coef == NULL: SRT_GRID_COEF=0, CPU test build) every query takes the on-the-fly path.
// Both paths give the same bits (tests/test_grid_cpu.py).
#pragma once
#include "srt_physics.h"
#include "srt_amat.h"

#if defined(__CUDA_ARCH__)
__constant__ signed char c_srt_amat[64][64] = {SRT_AMAT_ROWS};
#define SRT_AMAT(r, c) c_srt_amat[r][c]
#else
#if defined(__CUDACC__)
__constant__ signed char c_srt_amat[64][64] = {SRT_AMAT_ROWS};
#endif
static const signed char h_srt_amat[64][64] = {SRT_AMAT_ROWS};
#define SRT_AMAT(r, c) h_srt_amat[r][c]
#endif

// index (1-based, as maxloc returns it; 0 when the mask is all false) of the last node with
// 0 <= xi - x(ind)   (libtricubic.f95:835-840).  xs ascending, n >= 1.
SRT_HD int srt_grid_locate(const double* xs, int n, double minx, double delx, double xi) {
  double g = (xi - minx) / delx;
  int i;
  if (!(g > -1.0)) i = -1;
  else if (!(g < (double)n)) i = n - 1;
  else i = (int)g;  // truncation; g in (-1, n)
  if (i > n - 1) i = n - 1;
  while (i + 1 < n && 0.0 <= (xi - xs[i + 1])) ++i;
  while (i >= 0 && !(0.0 <= (xi - xs[i]))) --i;
  return i + 1;
}

// a = matmul(Amat, x) for one species (libtricubic.f95:719): per row, accumulation from 0 in
// ascending column order; exact zeros of Amat are skipped (adding +-0 to the running sum changes
// nothing for finite data, Q13).  xv = the 64 stacked, scaled corner data of tricubic_get_coeff.
SRT_HD double srt_amat_row(int r, const double xv[64]) {
  double s = 0.0;
  for (int c = 0; c < 64; ++c) {
    const int a = SRT_AMAT(r, c);
    if (a != 0) s = s + (double)a * xv[c];
  }
  return s;
}

// the scaled corner datum q (0..7: f, fx, fy, fz, fxy, fxz, fyz, fxyz) of libtricubic.f95:891-899
SRT_HD double srt_corner_scale(int q, double v, double dx, double dy, double dz) {
  switch (q) {
    case 0: return v;
    case 1: return v * dx;
    case 2: return v * dy;
    case 3: return v * dz;
    case 4: return v * dx * dy;
    case 5: return v * dx * dz;
    case 6: return v * dy * dz;
    default: return v * dx * dy * dz;
  }
}

// Coefficients of the interior cell whose lower corner is node (i0, j0, k0) (0-based), species s:
// out[t], t = 16*i + 4*j + k  <-  a(i + 4j + 16k).  No clamp can fire for an interior cell.
SRT_HD void srt_tricubic_cell_coef(const SrtGrid& g, int i0, int j0, int k0, int s, double out[64]) {
  double xv[64];
  for (int l = 0; l < 8; ++l) {
    const int ii = i0 + (l & 1), jj = j0 + ((l >> 1) & 1), kk = k0 + ((l >> 2) & 1);
    const double* rec = g.nodes + ((((size_t)kk * g.ny + jj) * g.nx + ii) * 32);
    for (int q = 0; q < 8; ++q) xv[8 * q + l] = srt_corner_scale(q, rec[4 * q + s], g.delx, g.dely, g.delz);
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      for (int k = 0; k < 4; ++k) out[16 * i + 4 * j + k] = srt_amat_row(i + 4 * j + 16 * k, xv);
}

// x**n for n = 0..3 as libgfortran's pow_r8_i4 evaluates it (square-and-multiply, Q24)
SRT_HD void srt_powers3(double x, double p[4]) {
  p[0] = 1.0;
  p[1] = x;
  p[2] = x * x;
  p[3] = x * (x * x);
}

// The 64 stacked, scaled corner data of the cell whose lower corner is node (is0, js0, ks0) (1-BASED, as maxloc returns
// them: 0 = below the first node, n = at or beyond the last), with the clamp logic of libtricubic.f95:859-921: an
// out-of-range corner index is clamped and raises a flag that zeroes the derivative terms across that axis; the flags
// are sticky across the 8 corners (Q12).  For an interior cell no clamp fires.
SRT_HD void srt_tricubic_xv(const SrtGrid& g, int is0, int js0, int ks0, int s, double xv[64]) {
  int flagi = 0, flagj = 0, flagk = 0;
  for (int l = 0; l < 8; ++l) {
    int istmp = is0 + (l & 1), jstmp = js0 + ((l >> 1) & 1), kstmp = ks0 + ((l >> 2) & 1);  // 1-based
    if (istmp < 1) { istmp = 1; flagi = 1; }
    if (istmp > g.nx) { istmp = g.nx; flagi = 1; }
    if (jstmp < 1) { jstmp = 1; flagj = 1; }
    if (jstmp > g.ny) { jstmp = g.ny; flagj = 1; }
    if (kstmp < 1) { kstmp = 1; flagk = 1; }
    if (kstmp > g.nz) { kstmp = g.nz; flagk = 1; }
    const double* rec = g.nodes + ((((size_t)(kstmp - 1) * g.ny + (jstmp - 1)) * g.nx + (istmp - 1)) * 32);
    for (int q = 0; q < 8; ++q) xv[8 * q + l] = srt_corner_scale(q, rec[4 * q + s], g.delx, g.dely, g.delz);
    if (flagi == 1) { xv[8 + l] = 0.0; xv[32 + l] = 0.0; xv[40 + l] = 0.0; xv[56 + l] = 0.0; }
    if (flagj == 1) { xv[16 + l] = 0.0; xv[32 + l] = 0.0; xv[48 + l] = 0.0; xv[56 + l] = 0.0; }
    if (flagk == 1) { xv[24 + l] = 0.0; xv[40 + l] = 0.0; xv[48 + l] = 0.0; xv[56 + l] = 0.0; }
  }
}

// Coefficients of ANY cell of the extended lattice is0 in [0, nx], js0 in [0, ny], ks0 in [0, nz] (1-based lower
// corner as srt_grid_locate returns it), species s: out[t], t = 16*i + 4*j + k  <-  a(i + 4j + 16k).  They depend on
// the cell only, not on the query point, so edge and outside cells are tabulated like interior ones; for an interior
// cell this is srt_tricubic_cell_coef(g, is0 - 1, js0 - 1, ks0 - 1, s, out).
SRT_HD void srt_tricubic_cell_coef_ext(const SrtGrid& g, int is0, int js0, int ks0, int s, double out[64]) {
  double xv[64];
  srt_tricubic_xv(g, is0, js0, ks0, s, xv);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      for (int k = 0; k < 4; ++k) out[16 * i + 4 * j + k] = srt_amat_row(i + 4 * j + 16 * k, xv);
}

// record of cell (is0, js0, ks0) in the extended coefficient table [nz+1][ny+1][nx+1][64 terms][4 species]
SRT_HD size_t srt_grid_cell_index(const SrtGrid& g, int is0, int js0, int ks0) {
  return ((size_t)ks0 * (size_t)(g.ny + 1) + (size_t)js0) * (size_t)(g.nx + 1) + (size_t)is0;
}
SRT_HD size_t srt_grid_cell_count(const SrtGrid& g) { return (size_t)(g.nx + 1) * (size_t)(g.ny + 1) * (size_t)(g.nz + 1); }

// the on-the-fly path: tricubic_interpolate_at for one species with the clamp logic of
// libtricubic.f95:859-921 (flags are sticky across the 8 corners, Q12)
SRT_HDN double srt_tricubic_slow(const SrtGrid& g, int is0, int js0, int ks0, double xil, double yil, double zil, int s) {
  double xv[64];
  srt_tricubic_xv(g, is0, js0, ks0, s, xv);
  double xp[4], yp[4], zp[4];
  srt_powers3(xil, xp);
  srt_powers3(yil, yp);
  srt_powers3(zil, zp);
  double v = 0.0;
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      for (int k = 0; k < 4; ++k) {
        double cont = srt_amat_row(i + 4 * j + 16 * k, xv) * xp[i] * yp[j] * zp[k];
        v = v + cont;
      }
  return v;
}

#if defined(__CUDA_ARCH__) && defined(SRT_TEAM_TU)
extern __shared__ unsigned long long srt_stage_smem[];  // the team kernel's dynamic shared memory (srt_team.cuh)
#endif

// Ns(1:4) = exp(tricubic(log-density)) at x (SM, metres)
SRT_HDN void srt_grid_density(const SrtGrid& g, const double x[3], double Ns[SRT_NSPEC]) {
  SRT_COUNT(SRT_T_TRICUBIC);
  const int is0 = srt_grid_locate(g.xs, g.nx, g.minx, g.delx, x[0]);
  const int js0 = srt_grid_locate(g.ys, g.ny, g.miny, g.dely, x[1]);
  const int ks0 = srt_grid_locate(g.zs, g.nz, g.minz, g.delz, x[2]);
  const bool ini = is0 >= 1 && is0 < g.nx, inj = js0 >= 1 && js0 < g.ny, ink = ks0 >= 1 && ks0 < g.nz;
  double xil = 0.0, yil = 0.0, zil = 0.0;
  if (ini) xil = (x[0] - g.xs[is0 - 1]) / g.delx;
  if (inj) yil = (x[1] - g.ys[js0 - 1]) / g.dely;
  if (ink) zil = (x[2] - g.zs[ks0 - 1]) / g.delz;
  double v[SRT_NSPEC] = {0.0, 0.0, 0.0, 0.0};
#if defined(__CUDA_ARCH__) && defined(SRT_TEAM_TU)
  // ---- Team kernel: cooperative fetch of the cell record (+4-7 % on config 3, profiles/r2a_c3_age.txt).
  // The 8 lanes of a team evaluate points a finite-difference step apart, i.e. almost always in the same cell, and
  // each of them walks the cell's whole 2 KiB record with its own 128 LDG.128.  Here the team fetches the record of
  // the centre lane's cell together, one i-slab (64 doubles) at a time: 4 coalesced LDG.128 per lane and slab into a
  // 528-byte shared-memory row per team (aliased onto the compaction scratch, which is only live between block
  // barriers), then every lane of that cell reads the slab by broadcast.  Lanes in another cell (a stencil point
  // across a cell face, est1 vs est2) or on the edge path fall through to the per-lane code below.  Same operands in
  // the same order: same bits.  Every caller in this translation unit reaches this point with whole teams.
  {
    const int lane_ = threadIdx.x & 31, tbase_ = lane_ & ~7, r_ = lane_ & 7;
    const unsigned tmask_ = 0xffu << tbase_;
    const long long c_ = g.coef ? (long long)srt_grid_cell_index(g, is0, js0, ks0) : -1ll;
    const long long c0_ = __shfl_sync(tmask_, c_, tbase_ + 6);
    if (c0_ >= 0) {  // team-uniform
      double* stage = (double*)srt_stage_smem + (size_t)(threadIdx.x >> 3) * 66;  // 66: the 4 teams of a warp on distinct banks
      const double2* src = (const double2*)(g.coef + (size_t)c0_ * 256);
      double xp[4], yp[4], zp[4];
      srt_powers3(xil, xp);
      srt_powers3(yil, yp);
      srt_powers3(zil, zp);
#pragma unroll 1
      for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int q = 0; q < 4; ++q) ((double2*)stage)[q * 8 + r_] = __ldg(src + i * 32 + q * 8 + r_);
        __syncwarp(tmask_);
        if (c_ == c0_) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              const double2 a01 = ((const double2*)stage)[2 * (4 * j + k)];
              const double2 a23 = ((const double2*)stage)[2 * (4 * j + k) + 1];
              const double as[4] = {a01.x, a01.y, a23.x, a23.y};
#pragma unroll
              for (int s = 0; s < SRT_NSPEC; ++s) v[s] = v[s] + as[s] * xp[i] * yp[j] * zp[k];
            }
          }
        }
        __syncwarp(tmask_);
      }
    }
    if (c0_ >= 0 && c_ == c0_) {
      for (int s = 0; s < SRT_NSPEC; ++s) Ns[s] = (s < g.nspec) ? srt_exp(v[s]) : 0.0;
      return;
    }
  }
#endif
  if (g.coef) {
    // every cell of the extended lattice is tabulated (edge and outside cells with their clamped corners), so a
    // query outside the grid -- or a ray whose position has become NaN: all three locates return 0 and the local
    // offsets stay 0 -- costs the same as an interior one
    const double* a = g.coef + srt_grid_cell_index(g, is0, js0, ks0) * 256;
    double xp[4], yp[4], zp[4];
    srt_powers3(xil, xp);
    srt_powers3(yil, yp);
    srt_powers3(zil, zp);
#pragma unroll 1
    for (int i = 0; i < 4; ++i) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int t = 16 * i + 4 * j + k;
#if defined(__CUDA_ARCH__)
          const double2 a01 = __ldg((const double2*)(a + 4 * t));
          const double2 a23 = __ldg((const double2*)(a + 4 * t + 2));
          const double as[4] = {a01.x, a01.y, a23.x, a23.y};
#else
          const double* as = a + 4 * t;
#endif
#pragma unroll
          for (int s = 0; s < SRT_NSPEC; ++s) v[s] = v[s] + as[s] * xp[i] * yp[j] * zp[k];
        }
      }
    }
  } else {
    for (int s = 0; s < g.nspec; ++s) v[s] = srt_tricubic_slow(g, is0, js0, ks0, xil, yil, zil, s);
  }
  for (int s = 0; s < SRT_NSPEC; ++s) Ns[s] = (s < g.nspec) ? srt_exp(v[s]) : 0.0;
}
