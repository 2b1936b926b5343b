// srt_scattered_setup.h -- host-side set-up of modelnum 4 (scattered samples), the role of `setup` in
// fortran/scattered_interp_dens_model_adapter.f95:63-212: read the sample file, drop repeated positions, find every
// sample's distance to its nearest other sample and the largest of those (maxnearest), and lay the samples out for the
// device code of srt_scattered.h.
//
// File (adapter :88-130, list-directed): nspec minx maxx miny maxy minz maxz / qs(1:nspec) / ms(1:nspec) / then one READ
// of pos(3), Ns(nspec) per sample until a READ fails (end of file or a bad value).  The values are log-densities.
//
// Differences from the reference (all forced by its use of an unseeded random permutation, adapter :141):
//   * of several samples at one position the reference keeps whichever its permutation inserts first; here the first in
//     the file is kept;
//   * the sample that happens to become the root of the reference's kd-tree gets the nearest distance 0 (kdtree_nearest,
//     kdtree_mod.f95:397-418, starts from best = root and nothing can beat distance 0); here every sample gets its true
//     nearest distance, which is what the others get in the reference too.
// Instead of a kd-tree the set-up uses two uniform cell lists: a fine one (about one sample per cell) for the
// nearest-neighbour pass, and the query list (cell edge = radius / 2, bounded in size) the device code walks.
#pragma once
#include <algorithm>
#include <cmath>
#include <numeric>
#include <vector>

#include "srt_listread.h"
#include "srt_physics.h"

namespace srt_host {

struct HostScat {
  std::vector<double> pts;      // [npts][8], sorted by query cell
  std::vector<int> cell_start;  // [cells + 1]
  std::vector<int> file_index;  // [npts]: position of each kept record in the file (0-based), for tests
};

struct ScatOptions {
  double window_scale;        // --scattered_interp_window_scale
  int order;                  // --scattered_interp_order
  int exact;                  // --scattered_interp_exact
  int scaled;                 // 0 in the driver (driver:722)
  double local_window_scale;  // --scattered_interp_local_window_scale
};

// cell list over a bounding box: origin, 1/edge, dims; the cell of a coordinate is srt_scat_cell_of, as in the query
struct ScatCells {
  double o[3], inv;
  int n[3];
  long long cells() const { return (long long)n[0] * n[1] * n[2]; }
  long long of(const double* p) const {
    const int cx = srt_scat_cell_of((p[0] - o[0]) * inv, n[0]);
    const int cy = srt_scat_cell_of((p[1] - o[1]) * inv, n[1]);
    const int cz = srt_scat_cell_of((p[2] - o[2]) * inv, n[2]);
    return ((long long)cz * n[1] + cy) * n[0] + cx;
  }
};

inline ScatCells scat_make_cells(const double lo[3], const double hi[3], double edge, long long max_cells) {
  ScatCells c;
  if (!(edge > 0.0) || !std::isfinite(edge)) edge = 1.0;
  for (;;) {
    long long total = 1;
    for (int k = 0; k < 3; ++k) {
      const double span = (hi[k] - lo[k]) / edge;
      c.n[k] = span < 2047.0 ? (int)span + 1 : 2048;
      c.o[k] = lo[k];
      total *= c.n[k];
    }
    c.inv = 1.0 / edge;
    if (total <= max_cells) return c;
    edge *= 1.26;  // halves the cell count
  }
}

// counting sort of the samples (records of 8 doubles) by cell; order inside a cell = order of `recs`
inline void scat_sort(const ScatCells& c, const std::vector<double>& recs, const std::vector<int>& idx, std::vector<double>& out,
                      std::vector<int>& idx_out, std::vector<int>& start) {
  const size_t n = recs.size() / SRT_SCAT_REC;
  start.assign((size_t)c.cells() + 1, 0);
  std::vector<long long> cell(n);
  for (size_t i = 0; i < n; ++i) { cell[i] = c.of(&recs[i * SRT_SCAT_REC]); ++start[(size_t)cell[i] + 1]; }
  for (size_t q = 0; q + 1 < start.size(); ++q) start[q + 1] += start[q];
  std::vector<int> fill(start.begin(), start.end() - 1);
  out.resize(recs.size());
  idx_out.resize(n);
  for (size_t i = 0; i < n; ++i) {
    const int d = fill[(size_t)cell[i]]++;
    std::copy(&recs[i * SRT_SCAT_REC], &recs[i * SRT_SCAT_REC] + SRT_SCAT_REC, &out[(size_t)d * SRT_SCAT_REC]);
    idx_out[d] = idx[i];
  }
}

// 0 ok, -1 cannot open, -2 bad file, -3 not supported (nspec > 4, order outside 0..3)
inline int load_scattered(const char* path, const ScatOptions& opt, SrtModel& M, HostScat& hs) {
  ListReader in;
  if (!in.open(path)) return -1;
  double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  if (in.read(7, v) != 7) return -2;  // nspec, minx, maxx, miny, maxy, minz, maxz (:88-89; the bounds are not used afterwards)
  const int nspec = (int)v[0];
  if (nspec < 1) return -2;
  if (nspec > SRT_NSPEC) return -3;
  if (opt.order < 0 || opt.order > 3) return -3;
  for (int i = 0; i < SRT_NSPEC; ++i) { M.qs[i] = 0.0; M.ms[i] = 1.0; }  // absent species: zero charge, zero density
  if (in.read(nspec, M.qs) != nspec) return -2;
  if (in.read(nspec, M.ms) != nspec) return -2;
  // the samples (:106-131): each READ starts on a new record; the first READ that fails ends the list
  std::vector<double> recs;
  double rec[SRT_SCAT_REC] = {0, 0, 0, 0, 0, 0, 0, 1.0};  // the extra value slot starts at 1.0 (:149)
  for (;;) {
    double item[3 + SRT_NSPEC];
    for (int k = 0; k < 3 + nspec; ++k) item[k] = rec[k];  // null values keep what pos / Ns held
    if (in.read(3 + nspec, item) != 3 + nspec) break;
    for (int k = 0; k < 3 + nspec; ++k) rec[k] = item[k];
    if (!(std::isfinite(rec[0]) && std::isfinite(rec[1]) && std::isfinite(rec[2]))) return -2;
    recs.insert(recs.end(), rec, rec + SRT_SCAT_REC);
  }
  size_t n = recs.size() / SRT_SCAT_REC;
  if (n == 0) return -2;
  // repeated positions (:143-155: distance to the nearest sample already in the tree == 0): the first in the file stays
  std::vector<int> order(n);
  std::iota(order.begin(), order.end(), 0);
  auto pos_less = [&](int a, int b) {
    const double* p = &recs[(size_t)a * SRT_SCAT_REC];
    const double* q = &recs[(size_t)b * SRT_SCAT_REC];
    for (int k = 0; k < 3; ++k) {
      if (p[k] < q[k]) return true;
      if (q[k] < p[k]) return false;
    }
    return a < b;
  };
  std::sort(order.begin(), order.end(), pos_less);
  std::vector<char> keep(n, 1);
  for (size_t i = 1; i < n; ++i) {
    const double* p = &recs[(size_t)order[i - 1] * SRT_SCAT_REC];
    const double* q = &recs[(size_t)order[i] * SRT_SCAT_REC];
    if (p[0] == q[0] && p[1] == q[1] && p[2] == q[2]) keep[order[i]] = 0;  // sorted by file index inside a run of equals
  }
  std::vector<double> uniq;
  std::vector<int> uidx;
  for (size_t i = 0; i < n; ++i)
    if (keep[i]) {
      uniq.insert(uniq.end(), &recs[i * SRT_SCAT_REC], &recs[i * SRT_SCAT_REC] + SRT_SCAT_REC);
      uidx.push_back((int)i);
    }
  recs.swap(uniq);
  n = recs.size() / SRT_SCAT_REC;
  double lo[3], hi[3];
  for (int k = 0; k < 3; ++k) lo[k] = hi[k] = recs[k];
  for (size_t i = 0; i < n; ++i)
    for (int k = 0; k < 3; ++k) {
      lo[k] = std::min(lo[k], recs[i * SRT_SCAT_REC + k]);
      hi[k] = std::max(hi[k], recs[i * SRT_SCAT_REC + k]);
    }
  // ---- nearest other sample of every sample outside the Earth (:160-191), by rings of a fine cell list
  double maxnearest = 0.0;
  {
    const double vol = std::max(hi[0] - lo[0], 1e-300) * std::max(hi[1] - lo[1], 1e-300) * std::max(hi[2] - lo[2], 1e-300);
    const double edge = std::cbrt(vol / (double)n);
    const ScatCells fc = scat_make_cells(lo, hi, edge, (long long)std::max<size_t>(4 * n, 4096));
    const double fedge = 1.0 / fc.inv;
    std::vector<double> sorted;
    std::vector<int> sidx, start;
    scat_sort(fc, recs, uidx, sorted, sidx, start);
    const int reach = std::max(fc.n[0], std::max(fc.n[1], fc.n[2]));
    for (size_t i = 0; i < n; ++i) {
      double* p = &sorted[i * SRT_SCAT_REC];
      if (!(p[0] * p[0] + p[1] * p[1] + p[2] * p[2] >= SRT_R_E * SRT_R_E)) continue;  // keeps 1.0 (:162)
      const int cx = srt_scat_cell_of((p[0] - fc.o[0]) * fc.inv, fc.n[0]);
      const int cy = srt_scat_cell_of((p[1] - fc.o[1]) * fc.inv, fc.n[1]);
      const int cz = srt_scat_cell_of((p[2] - fc.o[2]) * fc.inv, fc.n[2]);
      double best = HUGE_VAL;
      for (int ring = 0; ring <= reach; ++ring) {
        for (int z = cz - ring; z <= cz + ring; ++z) {
          if (z < 0 || z >= fc.n[2]) continue;
          for (int y = cy - ring; y <= cy + ring; ++y) {
            if (y < 0 || y >= fc.n[1]) continue;
            const bool shell_zy = (z == cz - ring || z == cz + ring || y == cy - ring || y == cy + ring);
            for (int x = cx - ring; x <= cx + ring; ++x) {
              if (x < 0 || x >= fc.n[0]) continue;
              if (!shell_zy && x != cx - ring && x != cx + ring) { x = cx + ring - 1; continue; }  // interior: jump to the far face
              const size_t cell = ((size_t)z * fc.n[1] + y) * fc.n[0] + x;
              for (int q = start[cell]; q < start[cell + 1]; ++q) {
                if ((size_t)q == i) continue;
                const double* o = &sorted[(size_t)q * SRT_SCAT_REC];
                const double a = o[0] - p[0], b = o[1] - p[1], c = o[2] - p[2];
                const double d2 = a * a + b * b + c * c;
                if (d2 < best) best = d2;
              }
            }
          }
        }
        // everything not yet visited lies more than ring * edge away along some axis
        const double safe = (double)ring * fedge * (1.0 - 1e-9);
        if (best <= safe * safe) break;
      }
      const double dist = (best == HUGE_VAL) ? 0.0 : std::sqrt(best);
      p[7] = dist;
      if (dist > maxnearest) maxnearest = dist;
    }
    recs.swap(sorted);
    uidx.swap(sidx);
  }
  // ---- the query cell list
  SrtScat& s = M.scat;
  memset(&s, 0, sizeof s);
  s.npts = (int)n;
  s.nspec = nspec;
  s.order = opt.order; s.exact = opt.exact; s.scaled = opt.scaled;
  s.maxnearest = maxnearest;
  s.window_scale = opt.window_scale;
  s.radius = maxnearest * opt.window_scale;  // adapter :247
  s.local_scale = opt.local_window_scale;
  int J = 0;
  for (int a = 0; a <= opt.order; ++a)  // tabular_monomials (lsinterp_mod.f95:73-95): first exponent slowest
    for (int b = 0; a + b <= opt.order; ++b)
      for (int c = 0; a + b + c <= opt.order; ++c) {
        s.mono[3 * J + 0] = (unsigned char)a; s.mono[3 * J + 1] = (unsigned char)b; s.mono[3 * J + 2] = (unsigned char)c;
        ++J;
      }
  s.nmono = J;
  const ScatCells qc = scat_make_cells(lo, hi, s.radius / 2.0, (long long)std::max<size_t>(8 * n, 4096));
  s.ox = qc.o[0]; s.oy = qc.o[1]; s.oz = qc.o[2];
  s.inv_cell = qc.inv;
  s.ncx = qc.n[0]; s.ncy = qc.n[1]; s.ncz = qc.n[2];
  // order inside a cell: the file's
  {
    std::vector<int> by_file(n);
    std::iota(by_file.begin(), by_file.end(), 0);
    std::sort(by_file.begin(), by_file.end(), [&](int a, int b) { return uidx[a] < uidx[b]; });
    std::vector<double> tmp(recs.size());
    std::vector<int> tidx(n);
    for (size_t i = 0; i < n; ++i) {
      std::copy(&recs[(size_t)by_file[i] * SRT_SCAT_REC], &recs[(size_t)by_file[i] * SRT_SCAT_REC] + SRT_SCAT_REC, &tmp[i * SRT_SCAT_REC]);
      tidx[i] = uidx[by_file[i]];
    }
    scat_sort(qc, tmp, tidx, hs.pts, hs.file_index, hs.cell_start);
  }
  s.pts = hs.pts.data();
  s.cell_start = hs.cell_start.data();
  M.density = SRT_DENS_SCAT;
  M.nspec = nspec;
  M.del = 1.0e-6;  // delDP, raytracer_driver.f95:1176
  set_species_invariants(M);
  return 0;
}

}  // namespace srt_host
