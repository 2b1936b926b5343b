// Host build of the kernels' TS05 field (csrc/srt_ts05.h is host/device code) for the CPU tests: the same statements
// the GPU runs, compared with the reference's own source as translated into oracle/_ref (tests/test_ts05_cpu.py).
#include "srt_ts05.h"

extern "C" void ts05_probe_extern(const double* par, double ps, double x, double y, double z, double* out) {
  Ts05Parts o;
  for (int i = 0; i < 3; ++i) o.b[i] = o.cf[i] = o.t1[i] = o.t2[i] = o.src[i] = o.prc[i] = o.r11[i] = o.r21[i] = 0.0;
  ts05_extern(par[0], par[1], par[2], par[3], par + 4, ps, x, y, z, o);
  const double* src[8] = {o.b, o.cf, o.t1, o.t2, o.src, o.prc, o.r11, o.r21};
  for (int k = 0; k < 8; ++k)
    for (int i = 0; i < 3; ++i) out[3 * k + i] = src[k][i];
}
extern "C" void ts05_probe_t04(const float* parmod, float ps, float x, float y, float z, float* b) {
  SrtTs05 m;
  m.on = 1;
  for (int i = 0; i < 10; ++i) m.parmod[i] = parmod[i];
  m.ps = ps;
  srt_ts05(m, x, y, z, b[0], b[1], b[2]);
}
