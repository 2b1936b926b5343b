// srt_scattered.h -- modelnum 4: density from scattered samples by moving least squares, written once as
// __host__ __device__ code like the rest of srt_physics.h (which includes this file).
//
// Reference: fortran/scattered_interp_dens_model_adapter.f95:241-269 (funcPlasmaParams: lsinterp on the log-densities
// when dot_product(x,x) > R_E**2, then exp; zero inside the Earth), fortran/lsinterp_mod.f95:185-235 (etainv,
// coswindow) and :255-449 (lsinterp), fortran/blas.f95:40-62,96-110,131-139 (dposv / dgemm / dgemv of the vendored
// lapack-3.2.1: dpotf2, dpotrs -> dtrsm, reference BLAS loops).
//
// What is different from the reference, and why it cannot be otherwise:
//   * The reference keeps the samples in a pointer-linked kd-tree (fortran/kdtree_mod.f95) that it fills in the order of
//     `randperm`, i.e. of the compiler's unseeded random_number (adapter :127-151).  A radius query returns its points
//     in the tree's traversal order (kdtree_mod.f95:160-188: right subtree, left subtree, node), so the ORDER of the
//     neighbours -- and with it the last bits of every sum over them -- differs from run to run of the reference itself.
//     The SET of neighbours is what the reference defines: all samples with dot_product(p-x, p-x) < radius**2.
//   * Here the samples are sorted once into a uniform cell list (srt_scattered_setup.h): a query walks the cells that
//     overlap [x-radius, x+radius] (z, then y, then one contiguous run of records along x), applies the reference's
//     own test to every candidate, and therefore finds the same set; its order is the cell order, fixed for a model.
//     No recursion, no pointers, 64-byte records read in runs: the form a GPU wants.
//   * Nothing per neighbour is stored: the reference materialises r(:), dinv(:), E(:,:) and aa(:); here the neighbours
//     are walked three times (window sums; normal matrix; interpolation weights) and the row of E is recomputed, with
//     the reference's expressions, so that memory per query is the packed J x J matrix only (J <= 20 monomials).
// Given the same neighbour order every operation below is the reference's, in its order (tests/test_scattered_cpu.py
// feeds the oracle's lsinterp this order and compares bit for bit).
#pragma once

#define SRT_SCAT_MAXJ 20  // monomials of degree <= 3 in three dimensions (lsinterp_mod.f95:73-95)
#define SRT_SCAT_REC 8    // doubles per sample: x, y, z, log-density of species 1..4, nearest-neighbour distance
// out of line, private to each translation unit that includes srt_physics.h
#if defined(__CUDACC__)
#define SRT_SCAT_FN static __host__ __device__ __noinline__
#else
#define SRT_SCAT_FN inline __attribute__((noinline))
#endif

// lsinterp_mod.f95:228-235
SRT_HD double srt_scat_coswindow(double r, double radius) { return 0.5 + 0.5 * srt_cos(r * 2.0 * SRT_PI / radius / 2.0); }

// lsinterp_mod.f95:185-222
SRT_HD double srt_scat_etainv(double r, double radius, double hin, int exact) {
  const double eps = 5.0e-16;
  if (exact == 1) {
    const double h = hin, s = radius;
    const double q = r / h;
    return ((1.0 + eps) / (srt_exp(q * q) - 1.0 + eps)) * (0.5 + 0.5 * srt_cos(r * 2.0 * SRT_PI / s / 2.0));
  }
  const double h = hin / 4.0, s = radius, alpha = 1.1;
  return srt_exp(-srt_pow((r + radius * eps) / h, alpha)) * (0.5 + 0.5 * srt_cos(r * 2.0 * SRT_PI / s / 2.0));
}

// Walk over the candidate samples of a query: the records of every cell overlapping the cube [x - radius, x + radius].
struct SrtScatWalk {
  int lox, hix, loy, hiy, loz, hiz;
  int cy, cz;
  int cur, end;
};

SRT_HD int srt_scat_cell_of(double t, int n) {  // t = (coordinate - origin) * inv_cell; NaN -> 0
  if (!(t > 0.0)) return 0;
  if (t >= (double)n) return n - 1;
  return (int)t;
}

SRT_HD void srt_scat_row(const SrtScat& s, SrtScatWalk& w) {
  const long long row = ((long long)w.cz * s.ncy + w.cy) * s.ncx;
  w.cur = s.cell_start[row + w.lox];
  w.end = s.cell_start[row + w.hix + 1];
}

SRT_HD void srt_scat_walk_begin(const SrtScat& s, const double x[3], SrtScatWalk& w) {
  const double r = s.radius;
  w.lox = srt_scat_cell_of((x[0] - r - s.ox) * s.inv_cell, s.ncx);
  w.hix = srt_scat_cell_of((x[0] + r - s.ox) * s.inv_cell, s.ncx);
  w.loy = srt_scat_cell_of((x[1] - r - s.oy) * s.inv_cell, s.ncy);
  w.hiy = srt_scat_cell_of((x[1] + r - s.oy) * s.inv_cell, s.ncy);
  w.loz = srt_scat_cell_of((x[2] - r - s.oz) * s.inv_cell, s.ncz);
  w.hiz = srt_scat_cell_of((x[2] + r - s.oz) * s.inv_cell, s.ncz);
  w.cy = w.loy;
  w.cz = w.loz;
  srt_scat_row(s, w);
}

// next sample inside the sphere (the test of kdtree_mod.f95:167-168): its record, or NULL when the walk is over; d2 = |p - x|^2
SRT_HD const double* srt_scat_walk_next(const SrtScat& s, const double x[3], double r2, SrtScatWalk& w, double& d2) {
  for (;;) {
    while (w.cur < w.end) {
      const double* p = s.pts + (size_t)w.cur * SRT_SCAT_REC;
      ++w.cur;
      const double a = p[0] - x[0], b = p[1] - x[1], c = p[2] - x[2];
      const double sm = a * a + b * b + c * c;
      if (sm < r2) { d2 = sm; return p; }
    }
    if (w.cy < w.hiy) ++w.cy;
    else if (w.cz < w.hiz) { w.cy = w.loy; ++w.cz; }
    else return nullptr;
    srt_scat_row(s, w);
  }
}

// the weight of one neighbour as lsinterp forms it (:345-356): dinv (after the square root) and, when scaled, sqrt(k)
SRT_HD void srt_scat_weight(const SrtScat& s, double eta, double& dinv, double& sqk) {
  dinv = 0.5 * eta;
  sqk = 1.0;
  if (s.scaled == 1) {
    const double k = dinv / (dinv + 1.0);
    dinv = 1.0 + dinv;
    sqk = srt_sqrt(k);
  }
  dinv = srt_sqrt(dinv);
}

// row of E for one neighbour (:358-372): monomials of (p - x), times dinv, times sqrt(k) when scaled
SRT_HD void srt_scat_erow(const SrtScat& s, const double* p, const double x[3], double dinv, double sqk, double e[SRT_SCAT_MAXJ]) {
  double pw[3][4];
  for (int k = 0; k < 3; ++k) {  // x**n as libgfortran's pow_r8_i4 (Q24)
    const double d = p[k] - x[k];
    pw[k][0] = 1.0;
    pw[k][1] = d;
    pw[k][2] = d * d;
    pw[k][3] = d * (d * d);
  }
  for (int j = 0; j < s.nmono; ++j) {
    double v = 1.0;
    for (int k = 0; k < 3; ++k) {
      const int m = s.mono[3 * j + k];
      if (m != 0) v = v * pw[k][m];
    }
    v = dinv * v;
    if (s.scaled == 1) v = sqk * v;
    e[j] = v;
  }
}

// lsinterp (lsinterp_mod.f95:255-449) at x: fi[0..nspec) and the status (0 ok, 1 solve failed, 2 too few points)
SRT_SCAT_FN int srt_scat_lsinterp(const SrtScat& s, const double x[3], double fi[SRT_NSPEC]) {
  const int J = s.nmono;
  const double radius = s.radius, r2 = radius * radius;
  for (int q = 0; q < SRT_NSPEC; ++q) fi[q] = 0.0;
  // ---- walk 1: number of points in the sphere, the windowed mean of their nearest-neighbour distances (:312-316)
  SrtScatWalk w;
  double d2;
  int I = 0;
  double sw = 0.0, swd = 0.0;
  srt_scat_walk_begin(s, x, w);
  while (const double* p = srt_scat_walk_next(s, x, r2, w, d2)) {
    const double cw = srt_scat_coswindow(srt_sqrt(d2), radius);
    swd = swd + cw * p[7];
    sw = sw + cw;
    ++I;
  }
  if (I < J) return 2;
  const double avgdist = swd / sw;
  const double h = s.local_scale * avgdist;
  // ---- walk 2: A = E' E over the points whose weight exceeds 1e-16 (:328-379); all points if fewer than J pass (:333-337)
  double A[SRT_SCAT_MAXJ * (SRT_SCAT_MAXJ + 1) / 2];  // upper triangle, column by column: A(j,l), j <= l, at l(l+1)/2 + j
  double e[SRT_SCAT_MAXJ], b[SRT_SCAT_MAXJ];
  bool all = false;
  for (;;) {
    for (int t = 0; t < J * (J + 1) / 2; ++t) A[t] = 0.0;
    int kept = 0;
    srt_scat_walk_begin(s, x, w);
    while (const double* p = srt_scat_walk_next(s, x, r2, w, d2)) {
      const double eta = srt_scat_etainv(srt_sqrt(d2), radius, h, s.exact);
      if (!all && !(eta > 1.0e-16)) continue;
      ++kept;
      double dinv, sqk;
      srt_scat_weight(s, eta, dinv, sqk);
      srt_scat_erow(s, p, x, dinv, sqk, e);
      for (int l = 0; l < J; ++l)
        for (int j = 0; j <= l; ++j) A[l * (l + 1) / 2 + j] = A[l * (l + 1) / 2 + j] + e[j] * e[l];
    }
    if (all || kept >= J) break;
    all = true;
  }
  // ---- dposv 'U' (blas.f95:54): dpotf2, then U' y = c, U a = y with c = (1, 0, ..., 0)
  for (int j = 0; j < J; ++j) {
    const int cj = j * (j + 1) / 2;
    double dot = 0.0;
    for (int k = 0; k < j; ++k) dot = dot + A[cj + k] * A[cj + k];
    double ajj = A[cj + j] - dot;
    if (ajj <= 0.0 || ajj != ajj) return 1;
    ajj = srt_sqrt(ajj);
    A[cj + j] = ajj;
    if (j + 1 < J) {
      if (j > 0)
        for (int l = j + 1; l < J; ++l) {
          const int cl = l * (l + 1) / 2;
          double t = 0.0;
          for (int i = 0; i < j; ++i) t = t + A[cl + i] * A[cj + i];
          A[cl + j] = A[cl + j] + (-1.0) * t;
        }
      const double rec = 1.0 / ajj;
      for (int l = j + 1; l < J; ++l) A[l * (l + 1) / 2 + j] = rec * A[l * (l + 1) / 2 + j];
    }
  }
  for (int i = 0; i < J; ++i) b[i] = (i == 0) ? 1.0 : 0.0;
  for (int i = 0; i < J; ++i) {
    const int ci = i * (i + 1) / 2;
    double t = b[i];
    for (int k = 0; k < i; ++k) t = t - A[ci + k] * b[k];
    b[i] = t / A[ci + i];
  }
  for (int k = J - 1; k >= 0; --k) {
    const int ck = k * (k + 1) / 2;
    if (b[k] != 0.0) {
      b[k] = b[k] / A[ck + k];
      for (int i = 0; i < k; ++i) b[i] = b[i] - b[k] * A[ck + i];
    }
  }
  // ---- walk 3: aa = (E a) * dinv (* sqrt(k)), fi = dot_product(aa, values) (:392-402)
  srt_scat_walk_begin(s, x, w);
  while (const double* p = srt_scat_walk_next(s, x, r2, w, d2)) {
    const double eta = srt_scat_etainv(srt_sqrt(d2), radius, h, s.exact);
    if (!all && !(eta > 1.0e-16)) continue;
    double dinv, sqk;
    srt_scat_weight(s, eta, dinv, sqk);
    srt_scat_erow(s, p, x, dinv, sqk, e);
    double y = 0.0;
    for (int j = 0; j < J; ++j)
      if (b[j] != 0.0) y = y + b[j] * e[j];
    double aa = y * dinv;
    if (s.scaled == 1) aa = aa * sqk;
    for (int q = 0; q < s.nspec; ++q) fi[q] = fi[q] + aa * p[3 + q];
  }
  return 0;
}

// Ns(1:4) at x (SM, metres): adapter :245-262.  A failed interpolation leaves fi = 0, hence exp(0) = 1 m^-3, as in the reference.
SRT_SCAT_FN void srt_scat_density(const SrtScat& s, const double x[3], double Ns[SRT_NSPEC]) {
  double fi[SRT_NSPEC] = {0.0, 0.0, 0.0, 0.0};
  const bool outside = x[0] * x[0] + x[1] * x[1] + x[2] * x[2] > SRT_R_E * SRT_R_E;
  if (outside) srt_scat_lsinterp(s, x, fi);
  for (int q = 0; q < SRT_NSPEC; ++q) Ns[q] = (outside && q < s.nspec) ? srt_exp(fi[q]) : 0.0;
}
