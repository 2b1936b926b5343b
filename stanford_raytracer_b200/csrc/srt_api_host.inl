// srt_api_host.inl -- host-buffer entry point (srt_trace_batch), result accessors and
// the FP64 peak probe.  Included at the end of srt_api.cu.

namespace {

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return SRT_OK;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMalloc(&p, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc");
    cap = bytes;
    return SRT_OK;
  }
};
struct PinBuf {
  void* p = nullptr;
  size_t cap = 0;
  int reserve(size_t bytes) {
    if (bytes <= cap) return SRT_OK;
    if (p) cudaFreeHost(p);
    p = nullptr;
    cap = 0;
    cudaError_t e = cudaMallocHost(&p, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMallocHost");
    cap = bytes;
    return SRT_OK;
  }
};

// scratch kept between calls (one process = one GPU); guarded by g_batch_mutex
std::mutex g_batch_mutex;
DevBuf g_rays, g_log, g_sorted, g_summary, g_counters, g_nlogged, g_offsets;
PinBuf g_rays_h, g_summary_h, g_small_h;
cudaStream_t g_stream = nullptr;

// Pinned host buffers for trajectory rows are expensive to create (page-locking GBs), so
// srt_result_free() parks the buffer here and the next srt_trace_batch() reuses it.
std::mutex g_pool_mutex;
double* g_pool_ptr = nullptr;
size_t g_pool_bytes = 0;
int pinned_rows_alloc(double** out, size_t bytes, size_t* got_bytes) {
  {
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    if (g_pool_ptr && g_pool_bytes >= bytes) {
      *out = g_pool_ptr;
      *got_bytes = g_pool_bytes;
      g_pool_ptr = nullptr;
      g_pool_bytes = 0;
      return SRT_OK;
    }
  }
  cudaError_t e = cudaMallocHost((void**)out, bytes);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMallocHost(rows)");
  *got_bytes = bytes;
  return SRT_OK;
}
void pinned_rows_release(double* p, size_t bytes) {
  if (!p) return;
  std::lock_guard<std::mutex> lk(g_pool_mutex);
  if (bytes > g_pool_bytes) {
    if (g_pool_ptr) cudaFreeHost(g_pool_ptr);
    g_pool_ptr = p;
    g_pool_bytes = bytes;
  } else {
    cudaFreeHost(p);
  }
}

size_t log_budget_rows(size_t want_rows) {
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) free_b = (size_t)8 << 30;
  // the log and its sorted copy share 60 % of what is free right now (cached scratch counts as free)
  size_t avail = (free_b + g_log.cap + g_sorted.cap) / 10 * 6;
  size_t per_row = (SRT_LOG_WORDS + SRT_ROW_DOUBLES) * sizeof(double);
  size_t max_rows = avail / per_row;
  const char* env = getenv("SRT_LOG_ROWS");  // test hook: force a small log to exercise the re-run path
  if (env) max_rows = (size_t)strtoull(env, nullptr, 10);
  return want_rows < max_rows ? want_rows : max_rows;
}

}  // namespace

extern "C" {

int srt_trace_batch(const srt_model* m, int64_t nrays, const double* pos0, double* dir0, const double* w0, const srt_run_params* p,
                    srt_result** out) {
  if (!m || !p || !out || nrays < 0 || (nrays > 0 && (!pos0 || !dir0 || !w0))) return SRT_E_ARG;
  if (nrays > 0xffffffffll) return SRT_E_ARG;
  *out = nullptr;
  int rc = ensure_device();
  if (rc) return rc;
  std::lock_guard<std::mutex> batch_lock(g_batch_mutex);
  if (!g_stream) CK(cudaStreamCreateWithFlags(&g_stream, cudaStreamNonBlocking));
  srt_result* res = new (std::nothrow) srt_result();
  if (!res) return SRT_E_NOMEM;
  res->nrays = nrays;
  res->outputper = p->outputper < 1 ? 1 : p->outputper;
  res->summary.assign((size_t)nrays * SRT_SUMMARY_WORDS, 0.0);
  res->offsets.assign((size_t)nrays, 0);
  res->nheld.assign((size_t)nrays, 0);
  if (nrays == 0) { *out = res; return SRT_OK; }
  SrtRunParams P;
  memcpy(&P, p, sizeof P);
  P.outputper = res->outputper;

#define FAIL(code) do { int c_ = (code); srt_result_free(res); return c_; } while (0)
#define CKR(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) FAIL(cuda_fail(e_, #call)); } while (0)

  // rows a ray can hold at most (Q15): maxsteps rows, every outputper-th is kept
  const size_t max_rows_per_ray = (size_t)((P.maxsteps < 1 ? 1 : P.maxsteps) + res->outputper - 1) / res->outputper + 1;

  // work list of the current wave (ray ids); first wave = all rays
  std::vector<int64_t> todo((size_t)nrays);
  for (int64_t i = 0; i < nrays; ++i) todo[(size_t)i] = i;
  // host rows are gathered wave by wave
  struct Chunk { double* first; uint64_t second; size_t bytes; };
  std::vector<Chunk> chunks;  // (pinned buffer, rows, capacity)
  std::vector<uint32_t> chunk_of((size_t)nrays, 0);

  // A full log drops every ray still in flight, so after a wave that dropped rays the next waves are
  // limited to what is guaranteed to fit (or to what just completed, if that is more).
  size_t wave_limit = (size_t)nrays;
  std::vector<int64_t> pending;  // rays not yet launched in any wave + rays to re-run
  pending.swap(todo);
  while (!pending.empty()) {
    const size_t n = pending.size() < wave_limit ? pending.size() : wave_limit;
    todo.assign(pending.begin(), pending.begin() + n);
    pending.erase(pending.begin(), pending.begin() + n);
    if ((rc = g_rays.reserve(n * 7 * sizeof(double)))) FAIL(rc);
    if ((rc = g_rays_h.reserve(n * 7 * sizeof(double)))) FAIL(rc);
    if ((rc = g_summary.reserve(n * SRT_SUMMARY_WORDS * sizeof(double)))) FAIL(rc);
    if ((rc = g_summary_h.reserve(n * SRT_SUMMARY_WORDS * sizeof(double)))) FAIL(rc);
    if ((rc = g_counters.reserve(8 * sizeof(unsigned long long)))) FAIL(rc);
    if ((rc = g_nlogged.reserve(n * sizeof(unsigned)))) FAIL(rc);
    if ((rc = g_offsets.reserve(n * sizeof(unsigned long long)))) FAIL(rc);
    if ((rc = g_small_h.reserve(n * (sizeof(unsigned) + sizeof(unsigned long long)) + 64))) FAIL(rc);
    size_t cap_rows = log_budget_rows(n * max_rows_per_ray);
    if (cap_rows < max_rows_per_ray && cap_rows < n * max_rows_per_ray) {
      // a log that cannot hold one worst-case ray may still hold the actual ones; only give up at zero
      if (cap_rows == 0) FAIL(SRT_E_NOMEM);
    }
    if ((rc = g_log.reserve(cap_rows * SRT_LOG_WORDS * sizeof(double)))) FAIL(rc);

    // AoS -> SoA staging in pinned memory, one H2D copy
    double* hr = (double*)g_rays_h.p;
    for (size_t s = 0; s < n; ++s) {
      const int64_t r = todo[s];
      hr[0 * n + s] = pos0[3 * r + 0]; hr[1 * n + s] = pos0[3 * r + 1]; hr[2 * n + s] = pos0[3 * r + 2];
      hr[3 * n + s] = dir0[3 * r + 0]; hr[4 * n + s] = dir0[3 * r + 1]; hr[5 * n + s] = dir0[3 * r + 2];
      hr[6 * n + s] = w0[r];
    }
    CKR(cudaMemcpyAsync(g_rays.p, hr, n * 7 * sizeof(double), cudaMemcpyHostToDevice, g_stream));
    SrtBatch B;
    B.nrays = (long long)n;
    B.rays = (double*)g_rays.p;
    B.log = (double*)g_log.p;
    B.log_capacity = cap_rows;
    B.summary = (double*)g_summary.p;
    B.counters = (unsigned long long*)g_counters.p;
    B.nlogged = (unsigned*)g_nlogged.p;
    {
      std::lock_guard<std::mutex> lk(g_model_mutex);
      if ((rc = launch_trace(const_cast<srt_model*>(m), P, B, g_stream))) FAIL(rc);
      CKR(cudaStreamSynchronize(g_stream));
    }
    unsigned long long counters[8];
    unsigned* nlogged_h = (unsigned*)g_small_h.p;
    unsigned long long* offsets_h = (unsigned long long*)((char*)g_small_h.p + ((n * sizeof(unsigned) + 63) / 64) * 64);
    CKR(cudaMemcpyAsync(counters, g_counters.p, sizeof counters, cudaMemcpyDeviceToHost, g_stream));
    CKR(cudaMemcpyAsync(nlogged_h, g_nlogged.p, n * sizeof(unsigned), cudaMemcpyDeviceToHost, g_stream));
    CKR(cudaMemcpyAsync(g_summary_h.p, g_summary.p, n * SRT_SUMMARY_WORDS * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
    CKR(cudaMemcpyAsync(hr, g_rays.p, n * 7 * sizeof(double), cudaMemcpyDeviceToHost, g_stream));  // dir0 is inout
    CKR(cudaStreamSynchronize(g_stream));
    res->attempts += counters[2];

    // offsets of the rays that completed in this wave
    uint64_t kept_rows = 0;
    std::vector<int64_t> again;
    const double* sh = (const double*)g_summary_h.p;
    for (size_t s = 0; s < n; ++s) {
      offsets_h[s] = kept_rows;
      const int flags = (int)sh[s * SRT_SUMMARY_WORDS + 23];
      if (flags & SRT_FLAG_DROPPED) { again.push_back(todo[s]); continue; }
      kept_rows += nlogged_h[s];
    }
    if (!again.empty()) {
      const size_t guaranteed = cap_rows / max_rows_per_ray > 0 ? cap_rows / max_rows_per_ray : 1;
      if (again.size() == n && n <= guaranteed) FAIL(SRT_E_NOMEM);  // even a guaranteed-size wave made no progress
      const size_t completed = n - again.size();
      wave_limit = completed > guaranteed ? completed : guaranteed;
    }
    double* chunk = nullptr;
    if (kept_rows) {
      if ((rc = g_sorted.reserve(kept_rows * SRT_ROW_DOUBLES * sizeof(double)))) FAIL(rc);
      CKR(cudaMemcpyAsync(g_offsets.p, offsets_h, n * sizeof(unsigned long long), cudaMemcpyHostToDevice, g_stream));
      unsigned long long nlog = counters[0] < cap_rows ? counters[0] : cap_rows;
      int blocks = g_num_sms * 8;
      srt_scatter_kernel<<<blocks, 256, 0, g_stream>>>((const double*)g_log.p, nlog, (const unsigned long long*)g_offsets.p,
                                                      (const unsigned*)g_nlogged.p, (double*)g_sorted.p);
      g_launches++;
      CKR(cudaGetLastError());
      size_t chunk_bytes = 0;
      if ((rc = pinned_rows_alloc(&chunk, kept_rows * SRT_ROW_DOUBLES * sizeof(double), &chunk_bytes))) FAIL(rc);
      chunks.push_back({chunk, kept_rows, chunk_bytes});
      CKR(cudaMemcpyAsync(chunk, g_sorted.p, kept_rows * SRT_ROW_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost, g_stream));
      CKR(cudaStreamSynchronize(g_stream));
    }
    for (size_t s = 0; s < n; ++s) {
      const int64_t r = todo[s];
      const int flags = (int)sh[s * SRT_SUMMARY_WORDS + 23];
      if (flags & SRT_FLAG_DROPPED) continue;
      memcpy(&res->summary[(size_t)r * SRT_SUMMARY_WORDS], sh + s * SRT_SUMMARY_WORDS, SRT_SUMMARY_WORDS * sizeof(double));
      res->nheld[(size_t)r] = nlogged_h[s];
      res->offsets[(size_t)r] = offsets_h[s];
      chunk_of[(size_t)r] = (uint32_t)(chunks.empty() ? 0 : chunks.size() - 1);
      res->accepted += (uint64_t)sh[s * SRT_SUMMARY_WORDS + 20] - 1;
      dir0[3 * r + 0] = hr[3 * n + s]; dir0[3 * r + 1] = hr[4 * n + s]; dir0[3 * r + 2] = hr[5 * n + s];
    }
    pending.insert(pending.end(), again.begin(), again.end());
  }
  // one contiguous pinned buffer when there was a single wave; otherwise concatenate
  if (chunks.size() == 1) {
    res->rows = chunks[0].first;
    res->rows_bytes = chunks[0].bytes;
    res->total_rows = chunks[0].second;
  } else if (chunks.size() > 1) {
    uint64_t total = 0;
    std::vector<uint64_t> base(chunks.size());
    for (size_t c = 0; c < chunks.size(); ++c) { base[c] = total; total += chunks[c].second; }
    double* all = nullptr;
    size_t all_bytes = 0;
    if ((rc = pinned_rows_alloc(&all, total * SRT_ROW_DOUBLES * sizeof(double), &all_bytes))) FAIL(rc);
    for (size_t c = 0; c < chunks.size(); ++c) {
      memcpy(all + base[c] * SRT_ROW_DOUBLES, chunks[c].first, chunks[c].second * SRT_ROW_DOUBLES * sizeof(double));
      pinned_rows_release(chunks[c].first, chunks[c].bytes);
    }
    for (int64_t r = 0; r < nrays; ++r) res->offsets[(size_t)r] += base[chunk_of[(size_t)r]];
    res->rows = all;
    res->rows_bytes = all_bytes;
    res->total_rows = total;
  }
#undef FAIL
#undef CKR
  *out = res;
  return SRT_OK;
}

int64_t srt_result_nrays(const srt_result* r) { return r ? r->nrays : 0; }
int64_t srt_result_total_rows(const srt_result* r) { return r ? (int64_t)r->total_rows : 0; }
int64_t srt_result_accepted_steps(const srt_result* r) { return r ? (int64_t)r->accepted : 0; }
int64_t srt_result_attempts(const srt_result* r) { return r ? (int64_t)r->attempts : 0; }
static bool ray_ok(const srt_result* r, int64_t ray) { return r && ray >= 0 && ray < r->nrays; }
int64_t srt_result_nrows(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? r->nheld[(size_t)ray] : SRT_E_ARG; }
int64_t srt_result_nsteps(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? (int64_t)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 20] : SRT_E_ARG; }
int srt_result_stopcond(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? (int)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 21] : SRT_E_ARG; }
int srt_result_flags(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? ((int)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 23] & SRT_FLAG_FIXED_CONE_EXIT) : SRT_E_ARG; }
int srt_result_ray_view(const srt_result* r, int64_t ray, const double** rows, int64_t* nrows) {
  if (!ray_ok(r, ray) || !rows || !nrows) return SRT_E_ARG;
  *nrows = r->nheld[(size_t)ray];
  *rows = r->rows ? r->rows + r->offsets[(size_t)ray] * SRT_ROW_DOUBLES : nullptr;
  return SRT_OK;
}
int srt_result_copy_ray(const srt_result* r, int64_t ray, double* rows) {
  const double* src;
  int64_t n;
  int rc = srt_result_ray_view(r, ray, &src, &n);
  if (rc) return rc;
  if (n && !rows) return SRT_E_ARG;
  if (n) memcpy(rows, src, (size_t)n * SRT_ROW_DOUBLES * sizeof(double));
  return SRT_OK;
}
int srt_result_summary(const srt_result* r, double* last_row, int64_t* nsteps, int64_t* nrejected, int32_t* stopcond) {
  if (!r) return SRT_E_ARG;
  for (int64_t i = 0; i < r->nrays; ++i) {
    const double* s = &r->summary[(size_t)i * SRT_SUMMARY_WORDS];
    if (last_row) memcpy(last_row + i * SRT_ROW_DOUBLES, s, SRT_ROW_DOUBLES * sizeof(double));
    if (nsteps) nsteps[i] = (int64_t)s[20];
    if (stopcond) stopcond[i] = (int32_t)s[21];
    if (nrejected) nrejected[i] = (int64_t)s[22];
  }
  return SRT_OK;
}
void srt_result_free(srt_result* r) {
  if (!r) return;
  pinned_rows_release(r->rows, r->rows_bytes);
  delete r;
}

int srt_fp64_peak(int fma, int iters, double* gflops) {
  if (!gflops || iters < 1) return SRT_E_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  double* d = nullptr;
  CK(cudaMalloc(&d, 8));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  const int blocks = g_num_sms * 8, threads = 256;
  for (int rep = 0; rep < 2; ++rep) {  // first pass warms up
    cudaEventRecord(e0);
    if (fma) srt_fp64_peak_kernel<1><<<blocks, threads>>>(iters, 1.0, d);
    else srt_fp64_peak_kernel<0><<<blocks, threads>>>(iters, 1.0, d);
    g_launches++;
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
  }
  float ms = 0;
  cudaEventElapsedTime(&ms, e0, e1);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "srt_fp64_peak");
  *gflops = 2.0 * 8.0 * (double)iters * blocks * threads / (ms * 1e-3) / 1e9;
  return SRT_OK;
}

}  // extern "C"
