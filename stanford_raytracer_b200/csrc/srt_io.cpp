// srt_io.cpp -- the two text formats of the reference driver, host side, no CUDA.
//
//   input  : inputraysfile, one list-directed READ of `pos0(3) dir0(3) w` per ray
//            (fortran/raytracer_driver.f95:1146), reading stops at the first record that fails;
//   output : per recorded row `(2i10)`, `(17es24.15e3)`, `(i10)`, then qs, ms, Ns, nus as `(4es24.15e3)` records
//            of nspec values each and a newline (fortran/raytracer_driver.f95:1197-1217).
//
// At GPU rates the writer is the wall-clock bottleneck of a drop-in run (822 bytes of text per row against
// 160 bytes of doubles), so formatting is exact *and* parallel: rays are cut into chunks of about equal row count,
// worker threads format chunks into private buffers, the calling thread writes them in ray order.
//
// ES24.15E3 is printf's %.15E with a three-digit exponent: gfortran prints the correctly rounded (round-half-even on
// the exact binary value) 16-digit decimal, the same digits std::to_chars(..., scientific, 15) produces (Ryu printf).
#include <atomic>
#include <cctype>
#include <charconv>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#include "../../include/srt.h"
#include "srt_listread.h"

namespace {

inline void put_es24(double x, char* out) {
  // 24 columns, right-justified
  if (std::isnan(x)) { memcpy(out, "                     NaN", 24); return; }
  if (std::isinf(x)) { memcpy(out, x > 0 ? "                Infinity" : "               -Infinity", 24); return; }
  char buf[40];
  auto r = std::to_chars(buf, buf + sizeof buf, x, std::chars_format::scientific, 15);
  // [-]d.ddddddddddddddde[+-]dd[d]
  const char* e = buf;
  while (e < r.ptr && *e != 'e') ++e;
  const int mant = (int)(e - buf);          // 17 or 18 with sign
  const char sign = e[1];
  int ex = 0;
  for (const char* p = e + 2; p < r.ptr; ++p) ex = ex * 10 + (*p - '0');
  if (ex > 999) { memset(out, '*', 24); return; }
  const int len = mant + 5;                 // E, sign, 3 digits
  const int pad = 24 - len;
  memset(out, ' ', pad);
  memcpy(out + pad, buf, mant);
  char* q = out + pad + mant;
  q[0] = 'E';
  q[1] = sign;
  q[2] = (char)('0' + ex / 100);
  q[3] = (char)('0' + (ex / 10) % 10);
  q[4] = (char)('0' + ex % 10);
}

inline void put_i10(long long v, char* out) {
  // Fortran I10: right-justified, ten asterisks when it does not fit
  char buf[32];
  int n = snprintf(buf, sizeof buf, "%lld", v);
  if (n > 10) { memset(out, '*', 10); return; }
  memset(out, ' ', 10 - n);
  memcpy(out + 10 - n, buf, n);
}

}  // namespace

extern "C" {

int srt_format_es24(const double* v, int64_t n, char* out) {
  if (n < 0 || (n > 0 && (!v || !out))) return SRT_E_ARG;
  for (int64_t i = 0; i < n; ++i) put_es24(v[i], out + 24 * i);
  return SRT_OK;
}

int64_t srt_row_text_bytes(int nspec) { return 10 + 10 + 17 * 24 + 10 + 4 * (int64_t)nspec * 24 + 1; }

int64_t srt_format_rows(int64_t raynum, int stopcond, const double* rows, int64_t nrows, double w, int nspec, const double* qs, const double* ms,
                        const double* nus, char* out) {
  if (nrows < 0 || nspec < 0 || nspec > 4 || (nrows > 0 && (!rows || !out)) || (nspec > 0 && (!qs || !ms || !nus))) return SRT_E_ARG;
  const int64_t stride = srt_row_text_bytes(nspec);
  // the columns that do not change along a ray are formatted once
  char head[20], wtxt[24], nstxt[10], qtxt[4 * 24], mtxt[4 * 24], ntxt[4 * 24];
  put_i10(raynum, head);
  put_i10(stopcond, head + 10);
  put_es24(w, wtxt);
  put_i10(nspec, nstxt);
  for (int s = 0; s < nspec; ++s) { put_es24(qs[s], qtxt + 24 * s); put_es24(ms[s], mtxt + 24 * s); put_es24(nus[s], ntxt + 24 * s); }
  for (int64_t i = 0; i < nrows; ++i) {
    const double* r = rows + i * SRT_ROW_DOUBLES;
    char* o = out + i * stride;
    memcpy(o, head, 20); o += 20;
    for (int c = 0; c < 16; ++c, o += 24) put_es24(r[c], o);     // time, pos, vprel, vgrel, n, B0
    memcpy(o, wtxt, 24); o += 24;
    memcpy(o, nstxt, 10); o += 10;
    memcpy(o, qtxt, 24 * nspec); o += 24 * nspec;
    memcpy(o, mtxt, 24 * nspec); o += 24 * nspec;
    for (int s = 0; s < nspec; ++s, o += 24) put_es24(r[16 + s], o);  // Ns
    memcpy(o, ntxt, 24 * nspec); o += 24 * nspec;
    *o = '\n';
  }
  return nrows * stride;
}

int srt_result_write_text(const srt_result* res, const srt_model* m, const double* w0, const char* path, int64_t first_raynum, int append,
                          int nthreads) {
  if (!res || !m || !w0 || !path) return SRT_E_ARG;
  int nspec = 0;
  double qs[4], ms[4], nus[4];
  int rc = srt_model_species(m, &nspec, qs, ms, nus);
  if (rc) return rc;
  const int64_t nrays = srt_result_nrays(res);
  const int64_t stride = srt_row_text_bytes(nspec);
  FILE* f = fopen(path, append ? "ab" : "wb");
  if (!f) return SRT_E_IO;
  if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
  if (nthreads <= 0) nthreads = 1;
  // chunks of consecutive rays, ~4 MiB of text each
  const int64_t rows_per_chunk = (4 << 20) / stride + 1;
  std::vector<int64_t> cut{0};
  for (int64_t r = 0, acc = 0; r < nrays; ++r) {
    acc += srt_result_nrows(res, r);
    if (acc >= rows_per_chunk || r == nrays - 1) { cut.push_back(r + 1); acc = 0; }
  }
  const int64_t nchunks = (int64_t)cut.size() - 1;
  std::vector<std::vector<char>> text((size_t)nchunks);
  std::vector<std::atomic<int>> ready((size_t)nchunks);
  for (auto& a : ready) a.store(0);
  std::atomic<int64_t> next{0};
  std::atomic<int64_t> written{0};   // chunks already on disk: bounds the text held in memory
  std::atomic<int> failed{0};
  const int64_t window = 4 * (int64_t)nthreads + 4;
  auto worker = [&]() {
    for (;;) {
      const int64_t c = next.fetch_add(1);
      if (c >= nchunks) return;
      while (c - written.load() > window && !failed.load()) std::this_thread::yield();
      int64_t rows = 0;
      for (int64_t r = cut[c]; r < cut[c + 1]; ++r) rows += srt_result_nrows(res, r);
      std::vector<char> buf((size_t)(rows * stride));
      char* o = buf.data();
      for (int64_t r = cut[c]; r < cut[c + 1]; ++r) {
        const double* rr = nullptr;
        int64_t nr = 0;
        if (srt_result_ray_view(res, r, &rr, &nr) != SRT_OK) { failed.store(1); nr = 0; }
        int64_t w = srt_format_rows(first_raynum + r, srt_result_stopcond(res, r), rr, nr, w0[r], nspec, qs, ms, nus, o);
        if (w < 0) { failed.store(1); w = 0; }
        o += w;
      }
      buf.resize((size_t)(o - buf.data()));
      text[(size_t)c].swap(buf);
      ready[(size_t)c].store(1, std::memory_order_release);
    }
  };
  std::vector<std::thread> pool;
  for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker);
  int status = SRT_OK;
  for (int64_t c = 0; c < nchunks; ++c) {
    while (!ready[(size_t)c].load(std::memory_order_acquire)) std::this_thread::yield();
    std::vector<char>& b = text[(size_t)c];
    if (!b.empty() && fwrite(b.data(), 1, b.size(), f) != b.size()) { status = SRT_E_IO; failed.store(1); }
    std::vector<char>().swap(b);
    written.store(c + 1);
  }
  for (auto& t : pool) t.join();
  if (fclose(f) != 0) status = SRT_E_IO;
  if (failed.load() && status == SRT_OK) status = SRT_E_IO;
  return status;
}

int srt_read_inputrays(const char* path, double** pos0, double** dir0, double** w0, int64_t* nrays) {
  if (!path || !pos0 || !dir0 || !w0 || !nrays) return SRT_E_ARG;
  // The driver's loop `read(infile, *, iostat=status) pos0, dir0, w` (raytracer_driver.f95:1146) under gfortran's list-directed
  // rules (srt_listread.h).  pos0, dir0 and w are variables of the main program: an item given a null value keeps what the
  // previous READ left (zeros at first); the first READ that does not complete -- a token that is not a number, or the end of
  // the file -- ends the loop.
  srt_host::ListReader rd;
  if (!rd.open(path)) return SRT_E_IO;
  std::vector<double> vals;
  double cur[7] = {0, 0, 0, 0, 0, 0, 0};
  while (rd.read(7, cur) == 7) vals.insert(vals.end(), cur, cur + 7);
  const int64_t n = (int64_t)(vals.size() / 7);
  *pos0 = (double*)malloc(sizeof(double) * 3 * (size_t)(n ? n : 1));
  *dir0 = (double*)malloc(sizeof(double) * 3 * (size_t)(n ? n : 1));
  *w0 = (double*)malloc(sizeof(double) * (size_t)(n ? n : 1));
  if (!*pos0 || !*dir0 || !*w0) { free(*pos0); free(*dir0); free(*w0); return SRT_E_NOMEM; }
  for (int64_t i = 0; i < n; ++i) {
    for (int c = 0; c < 3; ++c) { (*pos0)[3 * i + c] = vals[(size_t)(7 * i + c)]; (*dir0)[3 * i + c] = vals[(size_t)(7 * i + 3 + c)]; }
    (*w0)[i] = vals[(size_t)(7 * i + 6)];
  }
  *nrays = n;
  return SRT_OK;
}

void srt_free(void* p) { free(p); }

}  // extern "C"
