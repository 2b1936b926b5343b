// oracle/orc_grid.h -- TEST INFRASTRUCTURE ONLY (CPU oracle).  PARITY UNPINNED.
//
// Restatement of the gridded-density adapter (modelnum 3) and of libtricubic:
//   interp_dens_model_adapter::setup            fortran/interp_dens_model_adapter.f95:52-134
//   interp_dens_model_adapter::funcPlasmaParams fortran/interp_dens_model_adapter.f95:148-269
//   tricubic_interpolate_at                     tricubic-for/libtricubic.f95:796-933
//   tricubic_get_coeff / _stacked               :638-656, :715-720   (a = matmul(Amat, x))
//   tricubic_eval                               :658-695
//   point2xyz                                   :594-636
//   tricubic_compute_finite_difference_derivatives :722-793
// Arrays keep the Fortran shape f(nspec, nx, ny, nz) (species fastest).  Follows the reference
// statement by statement: three maxloc scans per query, 8-corner gather with sticky clamp flags
// (Q12), dense 64x64 matmul in ascending column order (Q13), x**n by square-and-multiply (Q24).
#pragma once
#include "orc_common.h"
#include "orc_models.h"
#include "orc_igrf.h"
#include "orc_amat.h"

namespace orc {

// libgfortran's _gfortran_pow_r8_i4 for n >= 0
inline double pow_r8_i4(double a, int n) {
  double pow = 1.0, x = a;
  unsigned u = (unsigned)n;
  if (n != 0) {
    for (;;) {
      if (u & 1) pow *= x;
      u >>= 1;
      if (u) x *= x;
      else break;
    }
  }
  return pow;
}

inline void point2xyz(int p, int& x, int& y, int& z) {
  switch (p) {
    case 0: x = 0; y = 0; z = 0; break;
    case 1: x = 1; y = 0; z = 0; break;
    case 2: x = 0; y = 1; z = 0; break;
    case 3: x = 1; y = 1; z = 0; break;
    case 4: x = 0; y = 0; z = 1; break;
    case 5: x = 1; y = 0; z = 1; break;
    case 6: x = 0; y = 1; z = 1; break;
    case 7: x = 1; y = 1; z = 1; break;
    default: x = 0; y = 0; z = 0;
  }
}

inline void tricubic_get_coeff_stacked(double a[64], const double x[64], bool dense) {
  for (int i = 0; i < 64; ++i) {
    double s = 0.0;
    for (int j = 0; j < 64; ++j) {
      if (!dense && Amat[i][j] == 0.0) continue;
      s = s + Amat[i][j] * x[j];
    }
    a[i] = s;
  }
}

inline void tricubic_get_coeff(double a[64], const double f[8], const double dfdx[8], const double dfdy[8], const double dfdz[8],
                               const double d2fdxdy[8], const double d2fdxdz[8], const double d2fdydz[8], const double d3fdxdydz[8], bool dense) {
  double x[64];
  for (int i = 0; i < 8; ++i) {
    x[0 + i] = f[i];
    x[8 + i] = dfdx[i];
    x[16 + i] = dfdy[i];
    x[24 + i] = dfdz[i];
    x[32 + i] = d2fdxdy[i];
    x[40 + i] = d2fdxdz[i];
    x[48 + i] = d2fdydz[i];
    x[56 + i] = d3fdxdydz[i];
  }
  tricubic_get_coeff_stacked(a, x, dense);
}

inline double tricubic_eval(const double a[64], double x, double y, double z, int derx, int dery, int derz) {
  double r = 0.0;
  for (int i = derx; i <= 3; ++i)
    for (int j = dery; j <= 3; ++j)
      for (int k = derz; k <= 3; ++k) {
        double cont = a[(i + 4 * j + 16 * k)] * (pow_r8_i4(x, i - derx)) * (pow_r8_i4(y, j - dery)) * (pow_r8_i4(z, k - derz));
        for (int w = 0; w <= derx - 1; ++w) cont = cont * (i - w);
        for (int w = 0; w <= dery - 1; ++w) cont = cont * (j - w);
        for (int w = 0; w <= derz - 1; ++w) cont = cont * (k - w);
        r = r + cont;
      }
  return r;
}

// one species' 3-D array view f(:,:,:) of a (nspec, nx, ny, nz) array
struct View3 {
  const double* base;
  int nspec, nx, ny, nz, s;
  double operator()(int i, int j, int k) const {  // 1-based
    return base[(((size_t)(k - 1) * ny + (j - 1)) * nx + (i - 1)) * nspec + s];
  }
};

// maxloc((/ (ind, ind=1,n) /), mask = 0.0 <= (xi - x)): 0 when the mask is all false
inline int maxloc_le(const std::vector<double>& x, double xi) {
  int best = 0;
  for (int ind = 1; ind <= (int)x.size(); ++ind)
    if (0.0 <= (xi - x[ind - 1])) best = ind;  // values ind ascend, so the last true wins
  return best;
}

inline double tricubic_interpolate_at(double xi, double yi, double zi, const View3& f, const std::vector<double>& x, const std::vector<double>& y,
                                      const std::vector<double>& z, const View3& dfdx, const View3& dfdy, const View3& dfdz, const View3& d2fdxdy,
                                      const View3& d2fdxdz, const View3& d2fdydz, const View3& d3fdxdydz, double dx, double dy, double dz, int derx,
                                      int dery, int derz, bool dense) {
  int flagi = 0, flagj = 0, flagk = 0;
  const int nx = f.nx, ny = f.ny, nz = f.nz;
  int is0 = maxloc_le(x, xi), js0 = maxloc_le(y, yi), ks0 = maxloc_le(z, zi);
  double xil, yil, zil;
  if (is0 >= 1 && is0 < (int)x.size()) xil = (xi - x[is0 - 1]) / dx; else xil = 0.0;
  if (js0 >= 1 && js0 < (int)y.size()) yil = (yi - y[js0 - 1]) / dy; else yil = 0.0;
  if (ks0 >= 1 && ks0 < (int)z.size()) zil = (zi - z[ks0 - 1]) / dz; else zil = 0.0;
  double fval[8], dfdxval[8], dfdyval[8], dfdzval[8], d2fdxdyval[8], d2fdxdzval[8], d2fdydzval[8], d3fdxdydzval[8];
  for (int l = 1; l <= 8; ++l) {
    int ioffset, joffset, koffset;
    point2xyz(l - 1, ioffset, joffset, koffset);
    int istmp = is0 + ioffset, jstmp = js0 + joffset, kstmp = ks0 + koffset;
    if (istmp < 1) { istmp = 1; flagi = 1; }
    if (istmp > nx) { istmp = nx; flagi = 1; }
    if (jstmp < 1) { jstmp = 1; flagj = 1; }
    if (jstmp > ny) { jstmp = ny; flagj = 1; }
    if (kstmp < 1) { kstmp = 1; flagk = 1; }
    if (kstmp > nz) { kstmp = nz; flagk = 1; }
    const int m = l - 1;
    fval[m] = f(istmp, jstmp, kstmp);
    dfdxval[m] = dfdx(istmp, jstmp, kstmp) * dx;
    dfdyval[m] = dfdy(istmp, jstmp, kstmp) * dy;
    dfdzval[m] = dfdz(istmp, jstmp, kstmp) * dz;
    d2fdxdyval[m] = d2fdxdy(istmp, jstmp, kstmp) * dx * dy;
    d2fdxdzval[m] = d2fdxdz(istmp, jstmp, kstmp) * dx * dz;
    d2fdydzval[m] = d2fdydz(istmp, jstmp, kstmp) * dy * dz;
    d3fdxdydzval[m] = d3fdxdydz(istmp, jstmp, kstmp) * dx * dy * dz;
    if (flagi == 1) { dfdxval[m] = 0.0; d2fdxdyval[m] = 0.0; d2fdxdzval[m] = 0.0; d3fdxdydzval[m] = 0.0; }
    if (flagj == 1) { dfdyval[m] = 0.0; d2fdxdyval[m] = 0.0; d2fdydzval[m] = 0.0; d3fdxdydzval[m] = 0.0; }
    if (flagk == 1) { dfdzval[m] = 0.0; d2fdxdzval[m] = 0.0; d2fdydzval[m] = 0.0; d3fdxdydzval[m] = 0.0; }
  }
  double a[64];
  tricubic_get_coeff(a, fval, dfdxval, dfdyval, dfdzval, d2fdxdyval, d2fdxdzval, d2fdydzval, d3fdxdydzval, dense);
  double r = tricubic_eval(a, xil, yil, zil, derx, dery, derz);
  r = r / pow_r8_i4(dx, derx);
  r = r / pow_r8_i4(dy, dery);
  r = r / pow_r8_i4(dz, derz);
  return r;
}

// tricubic_compute_finite_difference_derivatives for ONE species' 3-D arrays, stored [nz][ny][nx]
struct Arr3 {
  std::vector<double> v;
  int nx = 0, ny = 0, nz = 0;
  void init(int a, int b, int c) { nx = a; ny = b; nz = c; v.assign((size_t)a * b * c, 0.0); }
  double& operator()(int i, int j, int k) { return v[((size_t)(k - 1) * ny + (j - 1)) * nx + (i - 1)]; }
  double operator()(int i, int j, int k) const { return v[((size_t)(k - 1) * ny + (j - 1)) * nx + (i - 1)]; }
};

// d(src)/d(axis) with the one-sided ends of libtricubic.f95:738-741 etc.
inline void fd_axis(const Arr3& src, Arr3& dst, int axis, double h) {
  const int nx = src.nx, ny = src.ny, nz = src.nz;
  for (int k = 1; k <= nz; ++k)
    for (int j = 1; j <= ny; ++j)
      for (int i = 1; i <= nx; ++i) {
        int c = axis == 0 ? i : (axis == 1 ? j : k);
        int n = axis == 0 ? nx : (axis == 1 ? ny : nz);
        auto at = [&](int cc) { return axis == 0 ? src(cc, j, k) : (axis == 1 ? src(i, cc, k) : src(i, j, cc)); };
        if (c == 1) dst(i, j, k) = (at(2) - at(1)) / h;
        else if (c == n) dst(i, j, k) = (at(n) - at(n - 1)) / h;
        else dst(i, j, k) = (at(c + 1) - at(c - 1)) / 2.0 / h;
      }
}

inline void tricubic_compute_finite_difference_derivatives(const Arr3& f, double dx, double dy, double dz, Arr3& dfdx, Arr3& dfdy, Arr3& dfdz,
                                                           Arr3& d2fdxdy, Arr3& d2fdxdz, Arr3& d2fdydz, Arr3& d3fdxdydz) {
  const int nx = f.nx, ny = f.ny, nz = f.nz;
  Arr3* all[7] = {&dfdx, &dfdy, &dfdz, &d2fdxdy, &d2fdxdz, &d2fdydz, &d3fdxdydz};
  for (Arr3* a : all) a->init(nx, ny, nz);  // "= 0.0_8" branches
  if (nx > 2) fd_axis(f, dfdx, 0, dx);
  if (ny > 2) fd_axis(f, dfdy, 1, dy);
  if (nz > 2) fd_axis(f, dfdz, 2, dz);
  if (nx > 2 && ny > 2) fd_axis(dfdy, d2fdxdy, 0, dx);
  if (nx > 2 && nz > 2) fd_axis(dfdz, d2fdxdz, 0, dx);
  if (ny > 2 && nz > 2) fd_axis(dfdz, d2fdydz, 1, dy);
  if (nx > 2 && ny > 2 && nz > 2) fd_axis(d2fdydz, d3fdxdydz, 0, dx);
}

// fortran/interp_dens_model_adapter.f95
struct GridModel : Model {
  int computederivatives = 0, nspec = 0, nx = 0, ny = 0, nz = 0;
  double minx, maxx, miny, maxy, minz, maxz, delx, dely, delz;
  std::vector<double> qs, ms;
  std::vector<double> x, y, z;
  // (nspec, nx, ny, nz), species fastest
  std::vector<double> f, dfdx, dfdy, dfdz, d2fdxdy, d2fdxdz, d2fdydz, d3fdxdydz;
  FieldBlock field;
  void* field_block_ptr() override { return &field; }
  bool dense_matmul = false;

  int setup(const char* filename) {
    ListReader rd;
    if (!rd.load(filename)) return -1;
    double h[6];
    if (!rd.read(5, h)) return -2;
    computederivatives = (int)h[0]; nspec = (int)h[1]; nx = (int)h[2]; ny = (int)h[3]; nz = (int)h[4];
    if (nspec < 1 || nspec > 4 || nx < 1 || ny < 1 || nz < 1) return -2;
    if (!rd.read(6, h)) return -2;
    minx = h[0]; maxx = h[1]; miny = h[2]; maxy = h[3]; minz = h[4]; maxz = h[5];
    qs.resize(nspec); ms.resize(nspec);
    if (!rd.read(nspec, qs.data())) return -2;
    if (!rd.read(nspec, ms.data())) return -2;
    delx = (maxx - minx) / (nx - 1.0);
    dely = (maxy - miny) / (ny - 1.0);
    delz = (maxz - minz) / (nz - 1.0);
    x.resize(nx); y.resize(ny); z.resize(nz);
    for (int ind = 0; ind < nx; ++ind) x[ind] = ind * delx + minx;
    for (int ind = 0; ind < ny; ++ind) y[ind] = ind * dely + miny;
    for (int ind = 0; ind < nz; ++ind) z[ind] = ind * delz + minz;
    const size_t n = (size_t)nspec * nx * ny * nz;
    f.assign(n, 0.0);
    // the "slow read": one READ per node (a fresh record each), species innermost
    for (int l = 0; l < nz; ++l)
      for (int k = 0; k < ny; ++k)
        for (int j = 0; j < nx; ++j)
          if (!rd.read(nspec, &f[(((size_t)l * ny + k) * nx + j) * nspec])) return -2;
    std::vector<double>* der[7] = {&dfdx, &dfdy, &dfdz, &d2fdxdy, &d2fdxdz, &d2fdydz, &d3fdxdydz};
    for (auto* d : der) d->assign(n, 0.0);
    if (computederivatives == 1) {
      for (auto* d : der)
        if (!rd.read((int)n, d->data())) return -2;  // whole-array READ, array element order
    } else {
      for (int ind = 0; ind < nspec; ++ind) {
        Arr3 fs, o[7];
        fs.init(nx, ny, nz);
        for (size_t c = 0; c < (size_t)nx * ny * nz; ++c) fs.v[c] = f[c * nspec + ind];
        tricubic_compute_finite_difference_derivatives(fs, delx, dely, delz, o[0], o[1], o[2], o[3], o[4], o[5], o[6]);
        for (int q = 0; q < 7; ++q)
          for (size_t c = 0; c < (size_t)nx * ny * nz; ++c) (*der[q])[c * nspec + ind] = o[q].v[c];
      }
    }
    return 0;
  }

  void plasma(const double xx[3], Plasma& o) override {
    ++n_plasma_calls;
    o.nspec = nspec;
    for (int i = 0; i < 4; ++i) { o.qs[i] = 0.0; o.Ns[i] = 0.0; o.ms[i] = 0.0; o.nus[i] = 0.0; }
    for (int ind = 0; ind < nspec; ++ind) {
      auto view = [&](const std::vector<double>& a) { return View3{a.data(), nspec, nx, ny, nz, ind}; };
      o.Ns[ind] = tricubic_interpolate_at(xx[0], xx[1], xx[2], view(f), x, y, z, view(dfdx), view(dfdy), view(dfdz), view(d2fdxdy), view(d2fdxdz),
                                          view(d2fdydz), view(d3fdxdydz), delx, dely, delz, 0, 0, 0, dense_matmul);
    }
    for (int ind = 0; ind < nspec; ++ind) o.Ns[ind] = m_exp(o.Ns[ind]);
    for (int ind = 0; ind < nspec; ++ind) { o.qs[ind] = qs[ind]; o.ms[ind] = ms[ind]; o.nus[ind] = 0.0; }
    field.eval(xx, o.B0);
  }
};

}  // namespace orc
