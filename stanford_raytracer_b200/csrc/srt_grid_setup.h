// srt_grid_setup.h -- host loader of the gridded-density file of modelnum 3:
// interp_dens_model_adapter::setup (fortran/interp_dens_model_adapter.f95:52-134; file written by
// fortran/gcpm_dens_model_buildgrid.f95:302-327) and, when the file carries no derivatives,
// tricubic_compute_finite_difference_derivatives (tricubic-for/libtricubic.f95:722-793).
//
// File: `compder nspec nx ny nz` / `minx maxx miny maxy minz maxz` /
// qs(1:nspec) / ms(1:nspec) / f(s,i,j,k), s fastest then x, y, z -- natural log of the number density
// [m^-3] -- then, iff compder == 1, the seven derivative arrays in the same order.
//
// Output layout (SrtGrid::nodes): [nz][ny][nx][8 arrays][4 species], one 256-byte record per node.
#pragma once
#include <math.h>

#include "srt_host_setup.h"

namespace srt_host {

struct HostGrid {
  std::vector<double> nodes, xs, ys, zs;
};

// one derivative pass of libtricubic.f95:738-790 along `axis` (0: x, 1: y, 2: z) for array slot
// `src` -> slot `dst` of every node record, species s
inline void fd_pass(std::vector<double>& nodes, int nx, int ny, int nz, int s, int src, int dst, int axis, double h, bool enabled) {
  const int n[3] = {nx, ny, nz};
  auto at = [&](int i, int j, int k, int q) -> double& { return nodes[((((size_t)k * ny + j) * nx + i) * 8 + q) * 4 + s]; };
  for (int k = 0; k < nz; ++k)
    for (int j = 0; j < ny; ++j)
      for (int i = 0; i < nx; ++i) {
        if (!enabled) { at(i, j, k, dst) = 0.0; continue; }
        int c[3] = {i, j, k};
        int lo[3] = {i, j, k}, hi[3] = {i, j, k};
        const int m = n[axis];
        double v;
        if (c[axis] == 0) {
          hi[axis] = 1;
          v = (at(hi[0], hi[1], hi[2], src) - at(lo[0], lo[1], lo[2], src)) / h;
        } else if (c[axis] == m - 1) {
          lo[axis] = m - 2;
          v = (at(hi[0], hi[1], hi[2], src) - at(lo[0], lo[1], lo[2], src)) / h;
        } else {
          lo[axis] = c[axis] - 1;
          hi[axis] = c[axis] + 1;
          v = (at(hi[0], hi[1], hi[2], src) - at(lo[0], lo[1], lo[2], src)) / 2.0 / h;
        }
        at(i, j, k, dst) = v;
      }
}

// Returns 0, -1 cannot open, -2 garbled / truncated, -3 unsupported (nspec > 4).
// host_derivatives = false leaves the derivative slots of a file without derivatives at zero: the library computes
// them on the device (srt_fd_kernel, srt_api.cu); the CPU test build keeps the host passes below.
inline int load_grid(const char* path, SrtModel& M, HostGrid& hg, bool host_derivatives = true, int* compder_out = nullptr) {
  // every READ of `setup` (interp_dens_model_adapter.f95:59-118) is one ListReader::read: list-directed, starting at a record
  ListReader in;
  if (!in.open(path)) return -1;
  double v[6] = {0, 0, 0, 0, 0, 0};
  if (in.read(5, v) != 5) return -2;  // computederivatives, nspec, nx, ny, nz
  const int compder = (int)v[0], nspec = (int)v[1], nx = (int)v[2], ny = (int)v[3], nz = (int)v[4];
  if (nspec < 1 || nx < 1 || ny < 1 || nz < 1) return -2;
  if (nspec > SRT_NSPEC) return -3;
  if (in.read(6, v) != 6) return -2;
  const double minx = v[0], maxx = v[1], miny = v[2], maxy = v[3], minz = v[4], maxz = v[5];
  memset(&M.grid, 0, sizeof M.grid);
  for (int i = 0; i < SRT_NSPEC; ++i) { M.qs[i] = 0.0; M.ms[i] = 1.0; }  // absent species: zero charge, zero density
  if (in.read(nspec, M.qs) != nspec) return -2;
  if (in.read(nspec, M.ms) != nspec) return -2;
  SrtGrid& g = M.grid;
  g.nx = nx; g.ny = ny; g.nz = nz; g.nspec = nspec;
  g.minx = minx; g.miny = miny; g.minz = minz;
  // "equivalent to linspace" (:90-95)
  g.delx = (maxx - minx) / ((double)nx - 1.0);
  g.dely = (maxy - miny) / ((double)ny - 1.0);
  g.delz = (maxz - minz) / ((double)nz - 1.0);
  hg.xs.resize(nx); hg.ys.resize(ny); hg.zs.resize(nz);
  for (int i = 0; i < nx; ++i) hg.xs[i] = (double)i * g.delx + minx;
  for (int i = 0; i < ny; ++i) hg.ys[i] = (double)i * g.dely + miny;
  for (int i = 0; i < nz; ++i) hg.zs[i] = (double)i * g.delz + minz;
  const size_t nnodes = (size_t)nx * ny * nz;
  hg.nodes.assign(nnodes * 32, 0.0);
  // f: one READ per node (:102-108)
  for (size_t node = 0; node < nnodes; ++node) {
    double val[SRT_NSPEC] = {0, 0, 0, 0};
    if (in.read(nspec, val) != nspec) return -2;
    for (int s = 0; s < nspec; ++s) hg.nodes[(node * 8 + 0) * 4 + s] = val[s];
  }
  // the seven derivative arrays: one READ each of a whole array (:112-118)
  if (compder == 1) {
    std::vector<double> arr(nnodes * (size_t)nspec, 0.0);
    for (int q = 1; q < 8; ++q) {
      if (in.read((long)arr.size(), arr.data()) != (long)arr.size()) return -2;
      for (size_t node = 0; node < nnodes; ++node)
        for (int s = 0; s < nspec; ++s) hg.nodes[(node * 8 + q) * 4 + s] = arr[node * nspec + s];
    }
  }
  if (compder_out) *compder_out = compder;
  if (compder != 1 && host_derivatives) {
    for (int s = 0; s < nspec; ++s) {
      fd_pass(hg.nodes, nx, ny, nz, s, 0, 1, 0, g.delx, nx > 2);                      // df/dx
      fd_pass(hg.nodes, nx, ny, nz, s, 0, 2, 1, g.dely, ny > 2);                      // df/dy
      fd_pass(hg.nodes, nx, ny, nz, s, 0, 3, 2, g.delz, nz > 2);                      // df/dz
      fd_pass(hg.nodes, nx, ny, nz, s, 2, 4, 0, g.delx, nx > 2 && ny > 2);            // d2f/dxdy from df/dy
      fd_pass(hg.nodes, nx, ny, nz, s, 3, 5, 0, g.delx, nx > 2 && nz > 2);            // d2f/dxdz from df/dz
      fd_pass(hg.nodes, nx, ny, nz, s, 3, 6, 1, g.dely, ny > 2 && nz > 2);            // d2f/dydz from df/dz
      fd_pass(hg.nodes, nx, ny, nz, s, 6, 7, 0, g.delx, nx > 2 && ny > 2 && nz > 2);  // d3f/dxdydz from d2f/dydz
    }
  }
  M.density = SRT_DENS_GRID;
  M.nspec = nspec;
  M.del = 1.0e-6;  // delDP, raytracer_driver.f95:252
  set_species_invariants(M);
  return 0;
}

}  // namespace srt_host
