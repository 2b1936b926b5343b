// oracle/orc_api.cpp -- TEST INFRASTRUCTURE ONLY.  C entry points over the CPU
// oracle for ctypes (tests/, bench.py cpu_baseline / --impl reference, smoke()).
// PARITY UNPINNED (see orc_common.h).
#include "orc_common.h"
#include "orc_raytracer.h"
#include "orc_models.h"
#ifdef ORC_HAVE_GRID
#include "orc_grid.h"
#include "orc_scattered.h"
#endif

using namespace orc;

extern "C" {

// row layout shared with the product (include/srt.h): 20 doubles
//   t, pos[3], vprel[3], vgrel[3], n[3], B0[3], Ns[4]
#define ORC_ROW 20

void* orc_model_ngo(const char* cfg, int yearday, int msec, int use_igrf, int faithful_cost) {
  NgoModel* m = new NgoModel();
  if (m->st.readinput(cfg) != 0) { delete m; return nullptr; }
  m->del = 1.0e-4;  // delSP, driver:251
  m->field.tilt.init(yearday, msec);
  m->field.tilt.cache = !faithful_cost;
  m->field.use_igrf = use_igrf;
#ifdef ORC_HAVE_GRID
  if (!orc_attach_igrf(m->field, yearday, msec, faithful_cost)) { delete m; return nullptr; }
#else
  if (use_igrf) { delete m; return nullptr; }
#endif
  return (Model*)m;
}

#ifdef ORC_HAVE_GRID
void* orc_model_grid(const char* gridfile, int yearday, int msec, int use_igrf, int faithful_cost) {
  GridModel* m = new GridModel();
  if (m->setup(gridfile) != 0) { delete m; return nullptr; }
  m->del = 1.0e-6;  // delDP, driver:252
  m->dense_matmul = faithful_cost != 0;
  m->field.tilt.init(yearday, msec);
  m->field.tilt.cache = !faithful_cost;
  m->field.use_igrf = use_igrf;
  if (!orc_attach_igrf(m->field, yearday, msec, faithful_cost)) { delete m; return nullptr; }
  return (Model*)m;
}
#endif

#ifdef ORC_HAVE_GRID
// modelnum 4.  perm[nperm]: the insertion order of the samples (0-based file indices; nperm = 0: file order), the part of the
// reference's set-up its source leaves to an unseeded random_number; root_quirk: see orc_scattered.h
void* orc_model_scattered(const char* file, int yearday, int msec, int use_igrf, double window_scale, int order, int exact, int scaled,
                          double local_window_scale, const long* perm, long nperm, int root_quirk) {
  ScatteredModel* m = new ScatteredModel();
  std::vector<long> pv(perm, perm + (perm ? nperm : 0));
  if (m->setup(file, pv, root_quirk != 0) != 0) { delete m; return nullptr; }
  m->window_scale = window_scale; m->order = order; m->exact = exact; m->scaled = scaled; m->local_window_scale = local_window_scale;
  m->del = 1.0e-6;  // delDP, driver:1176
  m->field.tilt.init(yearday, msec);
  m->field.use_igrf = use_igrf;
  if (!orc_attach_igrf(m->field, yearday, msec, 0)) { delete m; return nullptr; }
  return (Model*)m;
}
// out: maxnearest, samples kept, duplicates ignored, nearest distance stored with the root sample
void orc_scattered_info(void* mv, double* out) {
  ScatteredModel* m = (ScatteredModel*)(Model*)mv;
  out[0] = m->maxnearest; out[1] = (double)m->nkept; out[2] = (double)m->nduplicates; out[3] = m->tree ? m->tree->val.back() : -1.0;
}
// radius query at x: the neighbours in the tree's traversal order, recs[cap][3 + nspec + 1]; returns how many there are
long orc_scattered_search(void* mv, const double* x, double* recs, long cap) {
  ScatteredModel* m = (ScatteredModel*)(Model*)mv;
  std::vector<KdTree*> found;
  m->search(x, found);
  const int w = 3 + m->nspec + 1;
  for (long i = 0; i < (long)found.size() && i < cap; ++i) {
    for (int k = 0; k < 3; ++k) recs[i * w + k] = found[i]->point[k];
    for (int k = 0; k <= m->nspec; ++k) recs[i * w + 3 + k] = found[i]->val[k];
  }
  return (long)found.size();
}
// the three BLAS / LAPACK calls of lsinterp as restated (column-major): compared with the vendored lapack-3.2.1 itself, translated
// into oracle/_ref/libdposv_ref.so
int orc_dposv_upper(int n, double* a, double* b) {
  std::vector<double> A(a, a + (size_t)n * n), B(b, b + n);
  const int info = dposv_upper(n, A, B);
  std::copy(A.begin(), A.end(), a);
  std::copy(B.begin(), B.end(), b);
  return info;
}
void orc_mat_trans_mat_mult(int m, int n, int k, const double* a, const double* b, double* c) { mat_trans_mat_mult(m, n, k, a, b, c); }
void orc_mat_vec_mult(int m, int n, const double* a, const double* x, double* y) { mat_vec_mult(m, n, a, x, y); }
long orc_scattered_impose_order(void* mv, long n, const double* pts) { return ((ScatteredModel*)(Model*)mv)->impose_order(n, pts); }
int orc_scattered_interpolate(void* mv, const double* x, double* fi) { return ((ScatteredModel*)(Model*)mv)->interpolate(x, fi); }
// lsinterp on neighbour arrays given in the caller's order: pts[I][3], vals[I][nval]
int orc_lsinterp_found(const double* x, long I, const double* pts, const double* vals, int nval, double radius, int order, int exact, int scaled,
                       double local_window_scale, double* fi) {
  std::vector<const double*> fp(I), fv(I);
  for (long i = 0; i < I; ++i) { fp[i] = pts + 3 * i; fv[i] = vals + (size_t)nval * i; }
  return lsinterp_found(x, fp, fv, nval, radius, order, exact, scaled, local_window_scale, fi);
}
#endif

// use_tsyganenko = 1: parmod = Pdyn, Dst, ByIMF, BzIMF, W1..W6; t04_s = address of ts05_ref_t04_s in oracle/_ref/libts05_ref*.so
void orc_model_tsyganenko(void* mv, const double* parmod, void* t04_s) {
  Model* m = (Model*)mv;
  FieldBlock* f = (FieldBlock*)m->field_block_ptr();
  for (int i = 0; i < 10; ++i) f->parmod[i] = (float)parmod[i];  // real(parmod)
  f->t04_s = (void (*)(const float*, float, float, float, float, float*))t04_s;
}
double orc_model_psi(void* mv) { return ((FieldBlock*)((Model*)mv)->field_block_ptr())->psi; }

#ifdef ORC_HAVE_GRID
// the oracle's Geopack state after TSY_RECALC, in the layout of COMMON /GEOPACK1/ (34 words) and /GEOPACK2/ (G, H, REC: 3 x 105),
// and IGRF_GSW_08 with it: compared with the reference's own geopack2008.for as translated into oracle/_ref
void orc_model_geopack(void* mv, float* geopack1, float* geopack2) {
  const Geopack& c = *(const Geopack*)((FieldBlock*)((Model*)mv)->field_block_ptr())->igrf_ctx;
  const float* w = &c.ST0;
  for (int i = 0; i < 34; ++i) geopack1[i] = w[i];
  for (int i = 0; i < 105; ++i) { geopack2[i] = c.G[i + 1]; geopack2[105 + i] = c.H[i + 1]; geopack2[210 + i] = c.REC[i + 1]; }
}
void orc_model_igrf_gsw(void* mv, float x, float y, float z, float* b) {
  const Geopack& c = *(const Geopack*)((FieldBlock*)((Model*)mv)->field_block_ptr())->igrf_ctx;
  IGRF_GSW_08(c, x, y, z, b[0], b[1], b[2]);
}
#endif

// `dens` of the Ngo model with the arguments funcPlasmaParams hands it (z(1), z(2), latitu) and the state readinput left:
// compared with the reference's own ngo_dens_model.f95 as translated into oracle/_ref/libngo_ref*.so
void orc_ngo_dens(void* mv, double z1, double z2, double latitu, double* ani) {
  NgoState& st = ((NgoModel*)(Model*)mv)->st;
  st.z[1] = z1; st.z[2] = z2; st.latitu = latitu;
  st.dens();
  for (int i = 0; i < 4; ++i) ani[i] = st.ani[i + 1];
}
void orc_ngo_state(void* mv, double* out) {
  const NgoState& st = ((NgoModel*)(Model*)mv)->st;
  const double s[12] = {st.ane0, st.rbase, st.r0, (double)st.num, (double)st.kducts, st.therm, st.rzero, st.scbot, st.lk, st.expk, st.ddk, st.pi};
  for (int i = 0; i < 12; ++i) out[i] = s[i];
}

// SM <-> GSM of the adapters (xform_double chain restated in orc_models.h): compared with the reference's own
// xform_double/*.f95 as translated into oracle/_ref/libxform_ref*.so
void orc_sm_to_gsm(void* mv, const double* x, double* out) { ((FieldBlock*)((Model*)mv)->field_block_ptr())->tilt.sm_to_gsm(x, out); }
void orc_gsm_to_sm(void* mv, const double* x, double* out) { ((FieldBlock*)((Model*)mv)->field_block_ptr())->tilt.gsm_to_sm(x, out); }

// Hooks for the reference's integrator as translated into oracle/_ref/librt_ref.so (f90_to_cpp.py): funcPlasmaParams = this
// oracle's adapter for the model chosen with orc_hook_model, csvd = the oracle's stand-in for ZGESVD (row 3 of VT = the conjugate
// of the right singular vector of the smallest singular value; the other rows are not used by the caller).
static thread_local Model* g_hook_model = nullptr;
void orc_hook_model(void* mv) { g_hook_model = (Model*)mv; }
void orc_hook_plasma(const double* x, double* qs, double* Ns, double* ms, double* nus, double* B0, int* nspec) {
  Plasma p;
  g_hook_model->plasma(x, p);
  *nspec = p.nspec;
  for (int i = 0; i < p.nspec; ++i) { qs[i] = p.qs[i]; Ns[i] = p.Ns[i]; ms[i] = p.ms[i]; nus[i] = p.nus[i]; }
  for (int i = 0; i < 3; ++i) B0[i] = p.B0[i];
}
void orc_hook_csvd(const double* m_re_im, double* vt_re_im) {
  cplx M[3][3], v[3];
  double sv[3];
  for (int j = 0; j < 3; ++j) for (int i = 0; i < 3; ++i) M[i][j] = {m_re_im[2 * (i + 3 * j)], m_re_im[2 * (i + 3 * j) + 1]};
  csvd3_smallest_right_vector(M, v, sv);
  for (int k = 0; k < 18; ++k) vt_re_im[k] = 0.0;
  for (int j = 0; j < 3; ++j) { vt_re_im[2 * (2 + 3 * j)] = v[j].re; vt_re_im[2 * (2 + 3 * j) + 1] = -v[j].im; }
}

#ifdef ORC_HAVE_GRID
// the gridded model's state (interpStateData of interp_dens_model_adapter.f95) for the reference's translated adapter + libtricubic:
// dims = nspec nx ny nz, del3, and pointers to qs, ms, x, y, z, f, dfdx, dfdy, dfdz, d2fdxdy, d2fdxdz, d2fdydz, d3fdxdydz (Fortran order)
// x**n of tricubic_eval as the oracle evaluates it (orc_grid.h), for the comparison with libgcc's __powidf2 (what gfortran's
// __builtin_powi becomes) and libgfortran's pow_r8_i4
double orc_pow_r8_i4(double x, int n) { return pow_r8_i4(x, n); }
void orc_grid_state(void* mv, int* dims, double* del3, double** ptrs) {
  GridModel* g = (GridModel*)(Model*)mv;
  dims[0] = g->nspec; dims[1] = g->nx; dims[2] = g->ny; dims[3] = g->nz;
  del3[0] = g->delx; del3[1] = g->dely; del3[2] = g->delz;
  std::vector<double>* v[13] = {&g->qs, &g->ms, &g->x, &g->y, &g->z, &g->f, &g->dfdx, &g->dfdy, &g->dfdz, &g->d2fdxdy, &g->d2fdxdz, &g->d2fdydz, &g->d3fdxdydz};
  for (int i = 0; i < 13; ++i) ptrs[i] = v[i]->data();
}
#endif

void orc_model_free(void* m) { delete (Model*)m; }

double orc_model_del(void* m) { return ((Model*)m)->del; }

// funcPlasmaParams at n points: out[n][19] = qs4, Ns4, ms4, nus4, B03
void orc_plasma(void* mv, long n, const double* x, double* out) {
  Model* m = (Model*)mv;
  for (long i = 0; i < n; ++i) {
    Plasma p;
    m->plasma(x + 3 * i, p);
    double* o = out + 19 * i;
    for (int s = 0; s < 4; ++s) { o[s] = p.qs[s]; o[4 + s] = p.Ns[s]; o[8 + s] = p.ms[s]; o[12 + s] = p.nus[s]; }
    for (int c = 0; c < 3; ++c) o[16 + c] = p.B0[c];
  }
}

double orc_dispersion_relation(void* mv, const double* n, double w, const double* x) {
  Plasma p;
  ((Model*)mv)->plasma(x, p);
  return dispersion_relation(n, w, p);
}
void orc_stix(void* mv, double w, const double* x, double* out5) {
  Plasma p;
  ((Model*)mv)->plasma(x, p);
  stix_parameters(w, p, m_sqrt(dot3(p.B0, p.B0)), out5[0], out5[1], out5[2], out5[3], out5[4]);
}
void orc_evalrhs(void* mv, const double* args7, double* rhs7) {
  Model* m = (Model*)mv;
  raytracer_evalrhs(args7, m->del, *m, rhs7);
}
void orc_solve(void* mv, const double* k, double w, const double* x, double* k1k2 /*re,im,re,im*/) {
  cplx k1, k2;
  solve_dispersion_relation(k, w, x, k1, k2, *(Model*)mv);
  k1k2[0] = k1.re; k1k2[1] = k1.im; k1k2[2] = k2.re; k1k2[3] = k2.im;
}
int orc_is_right_handed(double n2, double phi, double S, double D, double P) { return is_right_handed(n2, phi, S, D, P) ? 1 : 0; }
void orc_is_right_handed_n(long n, const double* n2, const double* phi, const double* S, const double* D, const double* P, int* out) {
  for (long i = 0; i < n; ++i) out[i] = is_right_handed(n2[i], phi[i], S[i], D[i], P[i]) ? 1 : 0;
}
// the float-rounded matrix of form_dispersion_matrix, for the numpy SVD cross-check: out[3][3][2]
void orc_form_dispersion_matrix(double n2, double phi, double S, double D, double P, double* out) {
  cplx M[3][3];
  form_dispersion_matrix(M, n2, phi, S, D, P);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { out[(i * 3 + j) * 2] = M[i][j].re; out[(i * 3 + j) * 2 + 1] = M[i][j].im; }
}

struct OrcTrace {
  RunResult r;
  double dir0[3];
  long plasma_calls;
};

// one raytracer_run; returns an opaque handle
void* orc_trace(void* mv, const double* pos0, const double* dir0, double w0, double dt0, double dtmax, double maxerr, int maxsteps,
                double minalt, int root, double tmax, int fixedstep) {
  Model* m = (Model*)mv;
  OrcTrace* t = new OrcTrace();
  for (int i = 0; i < 3; ++i) t->dir0[i] = dir0[i];
  long before = m->n_plasma_calls;
  raytracer_run(t->r, pos0, t->dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep, m->del, *m);
  t->plasma_calls = m->n_plasma_calls - before;
  return t;
}
long orc_trace_nrows(void* tv) { return (long)((OrcTrace*)tv)->r.rows.size(); }
int orc_trace_stopcond(void* tv) { return ((OrcTrace*)tv)->r.stopcond; }
long orc_trace_nrejected(void* tv) { return ((OrcTrace*)tv)->r.n_rejected; }
long orc_trace_ncone(void* tv) { return ((OrcTrace*)tv)->r.n_cone; }
long orc_trace_plasma_calls(void* tv) { return ((OrcTrace*)tv)->plasma_calls; }
int orc_trace_fixed_cone_exit(void* tv) { return ((OrcTrace*)tv)->r.fixed_cone_exit; }
void orc_trace_dir0(void* tv, double* d) { for (int i = 0; i < 3; ++i) d[i] = ((OrcTrace*)tv)->dir0[i]; }
void orc_trace_rows(void* tv, double* out) {
  OrcTrace* t = (OrcTrace*)tv;
  for (size_t i = 0; i < t->r.rows.size(); ++i) {
    const Row& r = t->r.rows[i];
    double* o = out + ORC_ROW * i;
    o[0] = r.t;
    for (int c = 0; c < 3; ++c) { o[1 + c] = r.pos[c]; o[4 + c] = r.vprel[c]; o[7 + c] = r.vgrel[c]; o[10 + c] = r.n[c]; o[13 + c] = r.B0[c]; }
    for (int s = 0; s < 4; ++s) o[16 + s] = r.Ns[s];
  }
}
// species constants of row 0 (qs4, ms4, nus4)
void orc_trace_species(void* tv, double* out12) {
  const Row& r = ((OrcTrace*)tv)->r.rows[0];
  for (int s = 0; s < 4; ++s) { out12[s] = r.qs[s]; out12[4 + s] = r.ms[s]; out12[8 + s] = r.nus[s]; }
}
void orc_trace_free(void* tv) { delete (OrcTrace*)tv; }

// batch over rays for the timed CPU baseline: returns total accepted steps (rows-1) and
// writes the last row + counts per ray.  Rays [first, first+count) of the arrays.
long orc_trace_batch_summary(void* mv, long nrays, const double* pos0, const double* dir0, const double* w0, double dt0, double dtmax,
                             double maxerr, int maxsteps, double minalt, int root, double tmax, int fixedstep, double* last_row,
                             long* nrows, int* stopcond, long* nrejected) {
  Model* m = (Model*)mv;
  long acc = 0;
  for (long i = 0; i < nrays; ++i) {
    RunResult r;
    double d[3] = {dir0[3 * i], dir0[3 * i + 1], dir0[3 * i + 2]};
    raytracer_run(r, pos0 + 3 * i, d, w0[i], dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep, m->del, *m);
    acc += (long)r.rows.size() - 1;
    if (nrows) nrows[i] = (long)r.rows.size();
    if (stopcond) stopcond[i] = r.stopcond;
    if (nrejected) nrejected[i] = r.n_rejected + r.n_cone;
    if (last_row) {
      const Row& q = r.rows.back();
      double* o = last_row + ORC_ROW * i;
      o[0] = q.t;
      for (int c = 0; c < 3; ++c) { o[1 + c] = q.pos[c]; o[4 + c] = q.vprel[c]; o[7 + c] = q.vgrel[c]; o[10 + c] = q.n[c]; o[13 + c] = q.B0[c]; }
      for (int s = 0; s < 4; ++s) o[16 + s] = q.Ns[s];
    }
  }
  return acc;
}

}  // extern "C"
