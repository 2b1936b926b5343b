// oracle/orc_common.h -- TEST INFRASTRUCTURE ONLY (CPU oracle).
//
// This directory is a CPU restatement of the reference's algorithm for the hot
// path.  It is the checker: only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may build, load or run it.
// The product (stanford_raytracer_b200/, libsrt.so) never includes or links it.
//
// PARITY: the reference ships no golden vectors for this path and no Fortran compiler
// exists in this environment.
//   PINNED by the reference's own source, bit for bit -- the hot path of configs 1-4 (Ngo and
//   gridded density, dipole / IGRF / TS05): the reference's sources are translated at build time
//   (f77_to_cpp.py for the FORTRAN 77-style files, f90_to_cpp.py over ref_stubs/f90rt.h for
//   raytracer.f95, both adapters, libtricubic.f95, util.f95, bmodel_dipole.f95, constants.f95
//   -> _ref/*.so) and the restatements here are compared with them, down to whole rays through
//   the translated raytracer_run (tests/test_*_ref_cpu.py, test_ts05_cpu.py).  f90rt.h states
//   the array semantics a compiler would decide.
//   The interp adapter's setup (grid file reader, node coordinates, finite-difference
//   derivatives) is translated too and leaves the state orc_models.h builds, bit for bit.
//   csvd = zgesvd of the LAPACK 3.2.1 the reference vendors: translated as well
//   (lapack_to_cpp.py -> _ref/liblapack_ref.so) and hooked under the translated is_right_handed,
//   it gives this restatement's answers (20 000 matrices, whole rays on both roots).
//   The driver's file formats are not restated here at all: the product's writer and reader
//   are compared with gfortran's own run-time library executing the driver's READ / WRITE
//   statements (ref_stubs/gfortran_io_shim.c, tests/test_formats_gfortran.py).
//
// Elementary functions: by default the oracle uses the same srt_math.h layer
// as the device (so GPU == oracle is a bitwise statement).  Compile with
// -DORC_LIBM to route them through glibc's libm instead, which is what the
// reference Fortran links; tests use that build to measure how far a different
// libm moves the results.
#pragma once
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <complex>
#include <string>
#include <vector>

#include "../stanford_raytracer_b200/csrc/srt_math.h"

namespace orc {

#ifdef ORC_LIBM
inline double m_exp(double x) { return ::exp(x); }
inline double m_log(double x) { return ::log(x); }
inline double m_pow(double x, double y) { return ::pow(x, y); }
inline double m_sin(double x) { return ::sin(x); }
inline double m_cos(double x) { return ::cos(x); }
inline double m_atan(double x) { return ::atan(x); }
inline double m_atan2(double y, double x) { return ::atan2(y, x); }
inline double m_acos(double x) { return ::acos(x); }
inline double m_fmod(double x, double y) { return ::fmod(x, y); }
inline void m_csqrt(double re, double im, double& ore, double& oim) {
  std::complex<double> r = std::sqrt(std::complex<double>(re, im));
  ore = r.real(); oim = r.imag();
}
#else
inline double m_exp(double x) { return srt_exp(x); }
inline double m_log(double x) { return srt_log(x); }
inline double m_pow(double x, double y) { return srt_pow(x, y); }
inline double m_sin(double x) { return srt_sin(x); }
inline double m_cos(double x) { return srt_cos(x); }
inline double m_atan(double x) { return srt_atan(x); }
inline double m_atan2(double y, double x) { return srt_atan2(y, x); }
inline double m_acos(double x) { return srt_acos(x); }
inline double m_fmod(double x, double y) { return srt_fmod(x, y); }
inline void m_csqrt(double re, double im, double& ore, double& oim) { srt_csqrt(re, im, ore, oim); }
#endif
inline double m_sqrt(double x) { return ::sqrt(x); }  // IEEE, correctly rounded everywhere

// fortran/constants.f95:4-12
static const double EPS0 = 8.854187817e-12;
static const double PI = 3.141592653589793238462643;
static const double MU0 = PI * 4e-7;
// C = sqrt(1/EPS0/MU0) evaluated in FP64 (constants.f95:7); not pasted.
inline double C_light() { static const double c = ::sqrt(1.0 / EPS0 / MU0); return c; }
static const double R_E = 6371.2e3;
static const double D2R = PI / 180.0;
static const double R2D = 180.0 / PI;

// minimal complex type with the operation order gfortran emits under
// -fcx-fortran-rules (no NaN recovery; Smith's division)
struct cplx {
  double re, im;
};
inline cplx c_mul(cplx a, cplx b) { return {a.re * b.re - a.im * b.im, a.re * b.im + a.im * b.re}; }
inline cplx c_div(cplx a, cplx b) {
  // Smith's algorithm as in GCC's expand_complex_div_wide
  if (fabs(b.re) < fabs(b.im)) {
    double ratio = b.re / b.im;
    double div = (b.re * ratio) + b.im;
    return {((a.re * ratio) + a.im) / div, ((a.im * ratio) - a.re) / div};
  } else {
    double ratio = b.im / b.re;
    double div = (b.im * ratio) + b.re;
    return {((a.im * ratio) + a.re) / div, (a.im - (a.re * ratio)) / div};
  }
}
inline cplx c_sqrt(cplx a) {
  cplx r;
  m_csqrt(a.re, a.im, r.re, r.im);
  return r;
}

struct Plasma {
  int nspec;
  double qs[4], Ns[4], ms[4], nus[4], B0[3];
};

// The funcPlasmaParams callback of fortran/raytracer.f95:121-129
struct Model {
  virtual ~Model() {}
  virtual void plasma(const double x[3], Plasma& p) = 0;
  virtual void* field_block_ptr() { return nullptr; }  // FieldBlock of the adapter (orc_models.h)
  double del;  // dF/dx step: delSP (1e-4) for modelnum 1, delDP (1e-6) for modelnum 3
  long n_plasma_calls = 0;
};

}  // namespace orc
