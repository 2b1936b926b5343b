// Test-only probe: exposes the srt_math.h elementary functions over arrays so
// tests/test_math.py can compare them with mpmath.  Not part of the product.
#include "../../stanford_raytracer_b200/csrc/srt_math.h"
extern "C" {
#define UN(name) void probe_##name(const double* x, double* y, long n) { for (long i = 0; i < n; ++i) y[i] = srt_##name(x[i]); }
UN(exp) UN(log) UN(sin) UN(cos) UN(atan) UN(acos)
void probe_pow(const double* x, const double* p, double* y, long n) { for (long i = 0; i < n; ++i) y[i] = srt_pow(x[i], p[i]); }
void probe_atan2(const double* a, const double* b, double* y, long n) { for (long i = 0; i < n; ++i) y[i] = srt_atan2(a[i], b[i]); }
void probe_fmod(const double* a, const double* b, double* y, long n) { for (long i = 0; i < n; ++i) y[i] = srt_fmod(a[i], b[i]); }
void probe_csqrt(const double* re, const double* im, double* ore, double* oim, long n) { for (long i = 0; i < n; ++i) srt_csqrt(re[i], im[i], ore[i], oim[i]); }
}
