// tests/csrc/hostemu.cpp -- TEST-ONLY CPU build of the device physics
// (stanford_raytracer_b200/csrc/srt_physics.h).  It lets the CPU suite check, without a
// GPU, that the de-duplicated evaluation order of the kernels reproduces the oracle's
// faithful call structure bit for bit.  Not shipped, not a fallback: libsrt.so has no
// CPU path.
#include "../../stanford_raytracer_b200/csrc/srt_physics.h"
#include "../../stanford_raytracer_b200/csrc/srt_igrf.h"
#include "../../stanford_raytracer_b200/csrc/srt_tricubic.h"
#include "../../stanford_raytracer_b200/csrc/srt_host_setup.h"
#ifdef EMU_HAVE_GRID
#include "../../stanford_raytracer_b200/csrc/srt_grid_setup.h"
#include "../../stanford_raytracer_b200/csrc/srt_igrf_setup.h"
#include "../../stanford_raytracer_b200/csrc/srt_scattered_setup.h"
#endif

#if defined(SRT_TALLY)
thread_local long long srt_tally[SRT_T_COUNT];
// per-attempt tallies of the current thread's last emu_trace_tally call: [accepted | rejected+cone][SRT_T_COUNT], attempts of each kind
static thread_local long long g_tally_sum[2][SRT_T_COUNT], g_tally_n[2];
template <int D, int I>
static void run_tally(const SrtModel& M, const SrtRunParams& P, const double* pos0, double* dir0, double w0) {
  SrtRayState st;
  double row[SRT_ROW_DOUBLES];
  srt_ray_init<D, I>(SrtModelRef{&M}, P, pos0, dir0, w0, st, row);
  for (;;) {
    long long before[SRT_T_COUNT];
    memcpy(before, srt_tally, sizeof before);
    const bool nan_state = st.x[0] != st.x[0] && st.k[0] != st.k[0];  // fast-forwarded on the GPU: not integrated work
    int r = srt_ray_attempt<D, I>(SrtModelRef{&M}, P, st, row);
    if (r == SRT_ATTEMPT_DONE) break;
    if (nan_state) continue;
    const int kind = (r == SRT_ATTEMPT_ACCEPTED) ? 0 : 1;
    for (int i = 0; i < SRT_T_COUNT; ++i) g_tally_sum[kind][i] += srt_tally[i] - before[i];
    ++g_tally_n[kind];
  }
}
#endif

template <int D, int I>
static long run_one(const SrtModel& M, const SrtRunParams& P, const double* pos0, double* dir0, double w0, double* rows, long max_rows,
                    int* stopcond, long* nrej, long* ncone, int* cone_exit) {
  SrtRayState st;
  double row[SRT_ROW_DOUBLES];
  srt_ray_init<D, I>(SrtModelRef{&M}, P, pos0, dir0, w0, st, row);
  long n = 0;
  if (n < max_rows) memcpy(rows + SRT_ROW_DOUBLES * n, row, sizeof row);
  ++n;
  for (;;) {
    int r = srt_ray_attempt<D, I>(SrtModelRef{&M}, P, st, row);
    if (r == SRT_ATTEMPT_DONE) break;
    if (r == SRT_ATTEMPT_ACCEPTED) {
      if (n < max_rows) memcpy(rows + SRT_ROW_DOUBLES * n, row, sizeof row);
      ++n;
    }
  }
  *stopcond = st.stopcond;
  *nrej = st.nrej;
  *ncone = st.ncone;
  *cone_exit = st.fixed_cone_exit;
  return n;
}

// whole ray, endpoint only (the CPU-baseline leg of bench.py: "deduplicated" mode)
template <int D, int I>
static long run_summary(const SrtModel& M, const SrtRunParams& P, const double* pos0, double* dir0, double w0, double* last_row, int* stopcond,
                        long* nrej) {
  SrtRayState st;
  double row[SRT_ROW_DOUBLES], keep[SRT_ROW_DOUBLES];
  srt_ray_init<D, I>(SrtModelRef{&M}, P, pos0, dir0, w0, st, keep);
  long n = 1;
  for (;;) {
    int r = srt_ray_attempt<D, I>(SrtModelRef{&M}, P, st, row);
    if (r == SRT_ATTEMPT_DONE) break;
    if (r == SRT_ATTEMPT_ACCEPTED) { memcpy(keep, row, sizeof keep); ++n; }
  }
  memcpy(last_row, keep, sizeof keep);
  *stopcond = st.stopcond;
  *nrej = (long)st.nrej + (long)st.ncone;
  return n;
}

template <int D, int I>
static void plasma_n(const SrtModel& M, long n, const double* x, double* out) {
  for (long i = 0; i < n; ++i) {
    SrtPlasma p;
    srt_plasma<D, I>(SrtModelRef{&M}, x + 3 * i, p);
    double* o = out + 19 * i;
    for (int s = 0; s < 4; ++s) { o[s] = M.qs[s]; o[4 + s] = p.Ns[s]; o[8 + s] = s < M.nspec ? M.ms[s] : 0.0; o[12 + s] = 0.0; }
    for (int c = 0; c < 3; ++c) o[16 + c] = p.B0[c];
  }
}
// model -> template arguments <D, I>: D = density kind, I = field code (bit 0 IGRF, bit 1 TS05 external field)
#define EMU_DISPATCH(M, CALL)                                                                      \
  do {                                                                                             \
    const int f_ = ((M).use_igrf ? 1 : 0) | ((M).ts.on ? 2 : 0);                                   \
    if ((M).density == SRT_DENS_NGO) {                                                             \
      constexpr int D = 0;                                                                         \
      if (f_ == 0) { constexpr int I = 0; CALL; } else if (f_ == 1) { constexpr int I = 1; CALL; } \
      else if (f_ == 2) { constexpr int I = 2; CALL; } else { constexpr int I = 3; CALL; }         \
    } else if ((M).density == SRT_DENS_SCAT) {                                                     \
      constexpr int D = 2;                                                                         \
      if (f_ == 0) { constexpr int I = 0; CALL; } else if (f_ == 1) { constexpr int I = 1; CALL; } \
      else if (f_ == 2) { constexpr int I = 2; CALL; } else { constexpr int I = 3; CALL; }         \
    } else {                                                                                       \
      constexpr int D = 1;                                                                         \
      if (f_ == 0) { constexpr int I = 0; CALL; } else if (f_ == 1) { constexpr int I = 1; CALL; } \
      else if (f_ == 2) { constexpr int I = 2; CALL; } else { constexpr int I = 3; CALL; }         \
    }                                                                                              \
  } while (0)
extern "C" {
void* emu_model_ngo(const char* cfg, int yearday, int msec, int use_igrf) {
  SrtModel* M = new SrtModel();
  memset(M, 0, sizeof *M);
  if (srt_host::load_ngo_config(cfg, *M) != 0) { delete M; return nullptr; }
  srt_host::set_time(*M, yearday, msec);
  M->use_igrf = use_igrf;
#ifdef EMU_HAVE_GRID
  srt_host::igrf_recalc(yearday, msec, M->igrf);
#endif
  return M;
}
#ifdef EMU_HAVE_GRID
void* emu_model_grid(const char* path, int yearday, int msec, int use_igrf) {
  SrtModel* M = new SrtModel();
  memset(M, 0, sizeof *M);
  srt_host::HostGrid* hg = new srt_host::HostGrid();
  if (srt_host::load_grid(path, *M, *hg) != 0) { delete M; delete hg; return nullptr; }
  M->grid.nodes = hg->nodes.data();
  M->grid.xs = hg->xs.data(); M->grid.ys = hg->ys.data(); M->grid.zs = hg->zs.data();
  M->grid.coef = nullptr;
  if (use_igrf & 2) {  // bit 1: precompute the per-cell coefficients the way srt_coef_kernel does
    use_igrf &= 1;
    const SrtGrid& g = M->grid;
    const size_t ncx = g.nx + 1, ncy = g.ny + 1, ncells = srt_grid_cell_count(g);
    std::vector<double>* coef = new std::vector<double>(ncells * 256, 0.0);
    for (size_t cell = 0; cell < ncells; ++cell)
      for (int sp = 0; sp < g.nspec; ++sp) {
        double a[64];
        srt_tricubic_cell_coef_ext(g, (int)(cell % ncx), (int)((cell / ncx) % ncy), (int)(cell / (ncx * ncy)), sp, a);
        for (int t = 0; t < 64; ++t) (*coef)[cell * 256 + 4 * t + sp] = a[t];
      }
    M->grid.coef = coef->data();
  }
  srt_host::set_time(*M, yearday, msec);
  M->use_igrf = use_igrf;
  srt_host::igrf_recalc(yearday, msec, M->igrf);
  return M;  // hg leaks with the model: test-only
}
#endif
#ifdef EMU_HAVE_GRID
// modelnum 4 through the product's own set-up (csrc/srt_scattered_setup.h)
void* emu_model_scattered(const char* path, int yearday, int msec, int use_igrf, double window_scale, int order, int exact, int scaled,
                          double local_window_scale) {
  SrtModel* M = new SrtModel();
  memset(M, 0, sizeof *M);
  srt_host::HostScat* hs = new srt_host::HostScat();
  const srt_host::ScatOptions opt = {window_scale, order, exact, scaled, local_window_scale};
  if (srt_host::load_scattered(path, opt, *M, *hs) != 0) { delete M; delete hs; return nullptr; }
  srt_host::set_time(*M, yearday, msec);
  M->use_igrf = use_igrf;
  srt_host::igrf_recalc(yearday, msec, M->igrf);
  return M;  // hs leaks with the model: test-only
}
// the samples as set-up laid them out: records [npts][8] in cell order; info = npts, nmono, ncx, ncy, ncz; dinfo = maxnearest, radius
long emu_scattered_state(void* m, double* recs, long cap, int* info, double* dinfo) {
  const SrtScat& s = ((const SrtModel*)m)->scat;
  info[0] = s.npts; info[1] = s.nmono; info[2] = s.ncx; info[3] = s.ncy; info[4] = s.ncz;
  dinfo[0] = s.maxnearest; dinfo[1] = s.radius;
  if (recs && cap >= (long)s.npts * SRT_SCAT_REC) memcpy(recs, s.pts, sizeof(double) * (size_t)s.npts * SRT_SCAT_REC);
  return s.npts;
}
// the neighbours of x in the order the device code walks them: indices into the records above
long emu_scattered_neighbours(void* m, const double* x, int* out, long cap) {
  const SrtScat& s = ((const SrtModel*)m)->scat;
  SrtScatWalk w;
  double d2;
  long n = 0;
  srt_scat_walk_begin(s, x, w);
  while (const double* p = srt_scat_walk_next(s, x, s.radius * s.radius, w, d2)) {
    if (n < cap) out[n] = (int)((p - s.pts) / SRT_SCAT_REC);
    ++n;
  }
  return n;
}
// lsinterp alone: fi[4] and the status
int emu_scattered_lsinterp(void* m, const double* x, double* fi) { return srt_scat_lsinterp(((const SrtModel*)m)->scat, x, fi); }
#endif
#ifdef EMU_HAVE_GRID
// what load_grid read: dims = nspec nx ny nz; out[q][node][s] for the eight arrays q (f and the seven derivative arrays)
long emu_grid_nodes(void* m, int* dims, double* qs_ms, double* out, long cap) {
  const SrtModel* M = (const SrtModel*)m;
  const SrtGrid& g = M->grid;
  dims[0] = g.nspec; dims[1] = g.nx; dims[2] = g.ny; dims[3] = g.nz;
  const long nodes = (long)g.nx * g.ny * g.nz;
  if (8 * nodes * g.nspec > cap) return -1;
  for (int s = 0; s < g.nspec; ++s) { qs_ms[s] = M->qs[s]; qs_ms[g.nspec + s] = M->ms[s]; }
  for (int q = 0; q < 8; ++q)
    for (long node = 0; node < nodes; ++node)
      for (int s = 0; s < g.nspec; ++s) out[(q * nodes + node) * g.nspec + s] = g.nodes[(node * 8 + q) * 4 + s];
  return nodes;
}
#endif
// a sequence of list-directed READs through the product's reader (csrc/srt_listread.h), as gfio_read_counts of
// oracle/ref_stubs/gfortran_io_shim.c does through libgfortran: -> READs completed, status[r] = 0 / -1 end of file / -2 bad value
int emu_read_counts(const char* path, const int* counts, int nreads, double* out, int* status) {
  srt_host::ListReader rd;
  if (!rd.open(path)) return -1;
  int done = 0;
  for (int r = 0; r < nreads; ++r) {
    const long got = rd.read(counts[r], out);
    status[r] = got == counts[r] ? 0 : (int)got;
    if (status[r] != 0) break;
    out += counts[r];
    ++done;
  }
  return done;
}
// use_tsyganenko = 1 with the given parameters; the tilt argument as finish_model (csrc/srt_api.cu) sets it
void emu_model_tsyganenko(void* m, const double* parmod, int yearday, int msec) {
  SrtModel* M = (SrtModel*)m;
  for (int i = 0; i < 10; ++i) M->ts.parmod[i] = (float)parmod[i];
  M->ts.on = 1;
#ifdef EMU_HAVE_GRID
  SrtIgrf tmp;
  srt_host::igrf_recalc(yearday, msec, tmp, &M->ts.ps);
#endif
}
#ifdef EMU_HAVE_GRID
// the IGRF state the device code reads (SrtIgrf: G, H, REC and the GEO->GSW matrix) and the device statement of IGRF_GSW_08
void emu_model_igrf(void* m, float* a9, float* ghrec) {
  const SrtIgrf& ig = ((SrtModel*)m)->igrf;
  const float a[9] = {ig.a11, ig.a21, ig.a31, ig.a12, ig.a22, ig.a32, ig.a13, ig.a23, ig.a33};
  for (int i = 0; i < 9; ++i) a9[i] = a[i];
  for (int i = 0; i < 105; ++i) { ghrec[i] = ig.g[i + 1]; ghrec[105 + i] = ig.h[i + 1]; ghrec[210 + i] = ig.rec[i + 1]; }
}
void emu_igrf_gsw(void* m, float x, float y, float z, float* b) { srt_igrf_gsw(((SrtModel*)m)->igrf, x, y, z, b[0], b[1], b[2]); }
float emu_model_ts_ps(void* m) { return ((SrtModel*)m)->ts.ps; }
#endif
void emu_model_free(void* m) { delete (SrtModel*)m; }
void emu_model_get(void* m, double* out /* ane0, cos_mu, sin_mu, del */) {
  SrtModel* M = (SrtModel*)m;
  out[0] = M->ngo.ane0; out[1] = M->cos_mu; out[2] = M->sin_mu; out[3] = M->del;
}

void emu_plasma(void* m, long n, const double* x, double* out) {
  const SrtModel& M = *(SrtModel*)m;
  EMU_DISPATCH(M, (plasma_n<D, I>(M, n, x, out)));
}
int emu_is_right_handed(double n2, double phi, double S, double D, double P) { return srt_is_right_handed(n2, phi, S, D, P) ? 1 : 0; }

void emu_is_right_handed_n(long n, const double* n2, const double* phi, const double* S, const double* D, const double* P, int* out) {
  for (long i = 0; i < n; ++i) out[i] = srt_is_right_handed(n2[i], phi[i], S[i], D[i], P[i]) ? 1 : 0;
}

#if defined(SRT_TALLY)
// out[2][SRT_T_COUNT + 1]: per kind (accepted, rejected) the summed counts of this ray's attempts and the number of attempts
void emu_trace_tally(void* m, const double* pos0, double* dir0, double w0, double dt0, double dtmax, double maxerr, int maxsteps, double minalt, int root,
                     double tmax, int fixedstep, long long* out) {
  const SrtModel& M = *(SrtModel*)m;
  SrtRunParams P = {dt0, dtmax, maxerr, minalt, tmax, maxsteps, root, fixedstep, 1};
  memset(g_tally_sum, 0, sizeof g_tally_sum);
  g_tally_n[0] = g_tally_n[1] = 0;
  EMU_DISPATCH(M, (run_tally<D, I>(M, P, pos0, dir0, w0)));
  for (int k = 0; k < 2; ++k) {
    for (int i = 0; i < SRT_T_COUNT; ++i) out[k * (SRT_T_COUNT + 1) + i] = g_tally_sum[k][i];
    out[k * (SRT_T_COUNT + 1) + SRT_T_COUNT] = g_tally_n[k];
  }
}
int emu_tally_slots(void) { return SRT_T_COUNT; }
#endif

long emu_trace_summary(void* m, const double* pos0, double* dir0, double w0, double dt0, double dtmax, double maxerr, int maxsteps, double minalt,
                       int root, double tmax, int fixedstep, double* last_row, int* stopcond, long* nrej) {
  const SrtModel& M = *(SrtModel*)m;
  SrtRunParams P = {dt0, dtmax, maxerr, minalt, tmax, maxsteps, root, fixedstep, 1};
  EMU_DISPATCH(M, return (run_summary<D, I>(M, P, pos0, dir0, w0, last_row, stopcond, nrej)));
  return -1;
}

long emu_trace(void* m, const double* pos0, double* dir0, double w0, double dt0, double dtmax, double maxerr, int maxsteps, double minalt,
               int root, double tmax, int fixedstep, double* rows, long max_rows, int* stopcond, long* nrej, long* ncone, int* cone_exit) {
  const SrtModel& M = *(SrtModel*)m;
  SrtRunParams P = {dt0, dtmax, maxerr, minalt, tmax, maxsteps, root, fixedstep, 1};
  EMU_DISPATCH(M, return (run_one<D, I>(M, P, pos0, dir0, w0, rows, max_rows, stopcond, nrej, ncone, cone_exit)));
  return -1;
}
}
