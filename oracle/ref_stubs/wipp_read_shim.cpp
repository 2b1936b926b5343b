// TEST INFRASTRUCTURE (oracle/): a C entry point over the REFERENCE'S OWN ray-file reader, read_rayfile of
// damping/src/wipp_fileutils.cpp (compiled where it lies into oracle/_ref/libwipp_read.so by oracle/Makefile `ref`), so that the
// product's output.ray writer is checked by the reference's consumer of that format.
#include <damping.h>

extern "C" {
// head[0..3] = number of rays in the file, rows of ray `raynum`, nspec, stopcond; head[4] = w; qs_ms[2 * nspec];
// rows[nrows][16 + 2 * nspec] = time, pos3, vprel3, vgrel3, n3, B03, Ns, nus.  Returns rows written, -1 without the ray.
long wipp_read_ray(const char* path, int raynum, double* head, double* qs_ms, double* rows, long cap_rows) {
  std::map<int, rayF> all = read_rayfile(path);
  head[0] = (double)all.size();
  if (!all.count(raynum)) return -1;
  rayF& r = all[raynum];
  int ns = (int)r.nspec;
  head[1] = (double)r.time.size(); head[2] = r.nspec; head[3] = r.stopcond; head[4] = r.w;
  for (int s = 0; s < ns; ++s) { qs_ms[s] = r.qs[s]; qs_ms[ns + s] = r.ms[s]; }
  long n = 0;
  for (size_t i = 0; i < r.time.size() && n < cap_rows; ++i, ++n) {
    double* o = rows + n * (16 + 2 * ns);
    o[0] = r.time[i];
    for (int c = 0; c < 3; ++c) { o[1 + c] = r.pos[i][c]; o[4 + c] = r.vprel[i][c]; o[7 + c] = r.vgrel[i][c]; o[10 + c] = r.n[i][c]; o[13 + c] = r.B0[i][c]; }
    for (int s = 0; s < ns; ++s) { o[16 + s] = r.Ns[i][s]; o[16 + ns + s] = r.nus[i][s]; }
  }
  return n;
}
}
