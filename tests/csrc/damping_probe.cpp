// tests/csrc/damping_probe.cpp -- TEST-ONLY: the host set-up of the damping kernels (csrc/srt_damping_setup.h) next to the
// stand-ins of the Fortran transforms the reference build links (oracle/ref_stubs/xform_d.cpp), for tests/test_damping_cpu.py.
#include "../../stanford_raytracer_b200/csrc/srt_damping_setup.h"
extern "C" {
void sm_to_mag_d_(int* itime, double* x_in, double* x_out);
void geo_to_mag_d_(int* itime, double* x_in, double* x_out);
void probe_gl(int n, double* x, double* w) { srt_host::gauss_legendre_tbl(n, x, w, 1e-10); }
void probe_sm2mag(int yearday, int msec, double* M) { int it[2] = {yearday, msec}; srt_host::sm_to_mag_matrix(it, M); }
double probe_greenwich(int yearday, int msec) { int it[2] = {yearday, msec}; return srt_host::greenwich_mag_lon_deg(it); }
void probe_ref_sm2mag(int yearday, int msec, const double* x, double* o) { int it[2] = {yearday, msec}; double xi[3] = {x[0], x[1], x[2]}; sm_to_mag_d_(it, xi, o); }
void probe_ref_geo2mag(int yearday, int msec, const double* x, double* o) { int it[2] = {yearday, msec}; double xi[3] = {x[0], x[1], x[2]}; geo_to_mag_d_(it, xi, o); }
}
