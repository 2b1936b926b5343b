// tests/csrc/damping_probe.cpp -- TEST-ONLY: the host set-up of the damping kernels (csrc/srt_damping_setup.h) next to the
// stand-ins of the Fortran transforms the reference build links (oracle/ref_stubs/xform_d.cpp), for tests/test_damping_cpu.py.
#define DMP_TEST_NOISE 1
#include "../../stanford_raytracer_b200/csrc/srt_damping_setup.h"
double dmp_test_noise_amp = 0.0;
unsigned long long dmp_test_noise_seed = 0;
extern "C" {
void sm_to_mag_d_(int* itime, double* x_in, double* x_out);
void geo_to_mag_d_(int* itime, double* x_in, double* x_out);
// the GPU path's arithmetic (csrc/srt_damping_row.h) run sequentially on the host with glibc: same inputs and outputs as
// srt_damping_foust
// test hook: relative perturbations applied to the fitted density parameters of every row (conditioning measurement)
static double g_perturb[3] = {0.0, 0.0, 0.0};
void probe_set_noise(double amp, unsigned long long seed) { dmp_test_noise_amp = amp; dmp_test_noise_seed = seed; }
void probe_set_perturbation(double n_fit_rel, double An_fit_rel, double L_pp_rel) { g_perturb[0] = n_fit_rel; g_perturb[1] = An_fit_rel; g_perturb[2] = L_pp_rel; }
int probe_damping(long nrays, const long* first, long nrows, const double* rows, const double* w, int nspec, const double* qs, const double* ms,
                  double Kp, double AE, int yearday, int msec, int geom, const char* table, double* power) {
  DampSetup S;
  std::vector<double> tab;
  char err[256];
  int rc = srt_host::damp_setup(table, nspec, qs, ms, Kp, AE, yearday, msec, geom, S, tab, err, sizeof err);
  if (rc) return rc;
  for (long ray = 0; ray < nrays; ++ray) {
    const long b = first[ray], e = ray + 1 < nrays ? first[ray + 1] : nrows;
    double d = 1.0, g0 = 0.0;
    for (long i = b; i < e; ++i) {
      DampRow R;
      damp_row_setup(S, tab.data(), rows + 20 * i, i == b ? nullptr : rows + 20 * (i - 1), w[ray], R);
      if (i == b) { g0 = R.geom; power[i] = 1.0; continue; }
      R.psd.n_fit *= 1.0 + g_perturb[0];
      R.psd.An_fit *= 1.0 + g_perturb[1];
      R.psd.L_pp *= 1.0 + g_perturb[2];
      double s = 0.0;
      for (int j = 0; j < 25; ++j) {
        const double Ax = 0.5 * S.glx[j];
        s += S.glw[j] * (integrand_t(R.q, R.psd, 0.5 + Ax) + integrand_t(R.q, R.psd, 0.5 - Ax));
      }
      const double gf = S.include_geom ? g0 / R.geom : 1.0;
      d = gf * d * damp_row_factor(R, 0.5 * s);
      power[i] = d;
    }
  }
  return 0;
}
void probe_gl(int n, double* x, double* w) { srt_host::gauss_legendre_tbl(n, x, w, 1e-10); }
void probe_sm2mag(int yearday, int msec, double* M) { int it[2] = {yearday, msec}; srt_host::sm_to_mag_matrix(it, M); }
double probe_greenwich(int yearday, int msec) { int it[2] = {yearday, msec}; return srt_host::greenwich_mag_lon_deg(it); }
void probe_ref_sm2mag(int yearday, int msec, const double* x, double* o) { int it[2] = {yearday, msec}; double xi[3] = {x[0], x[1], x[2]}; sm_to_mag_d_(it, xi, o); }
void probe_ref_geo2mag(int yearday, int msec, const double* x, double* o) { int it[2] = {yearday, msec}; double xi[3] = {x[0], x[1], x[2]}; geo_to_mag_d_(it, xi, o); }
}
