// srt_api_host.inl -- host-buffer entry point (srt_trace_batch: one call for the whole ray file, spread over every
// active device), result objects and accessors, the FP64 peak probe.  Included at the end of srt_api.cu.
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace {

// ---- host memory for trajectory rows.  Page-locking GBs is expensive, so buffers are pooled between calls.
// Two kinds: cudaMallocHost (default) and POSIX shared memory registered with CUDA (srt_set_rows_shm): the rows of a
// rank then land, by the D2H copy itself, in a segment the gathering rank of a multi-process run maps -- the host-side
// gather of SURVEY.md 8(e) without another copy.
struct HostRows {
  double* p = nullptr;
  size_t bytes = 0;
  bool shm = false;
  char name[80] = "";
};
std::mutex g_pool_mutex;
std::vector<HostRows> g_pool;
std::string g_shm_prefix;
std::atomic<unsigned> g_shm_seq{0};

void host_rows_destroy(HostRows& h) {
  if (!h.p) return;
  if (h.shm) {
    cudaHostUnregister(h.p);
    munmap(h.p, h.bytes);
    shm_unlink(h.name);
  } else {
    cudaFreeHost(h.p);
  }
  h = HostRows();
}

int host_rows_alloc(size_t bytes, HostRows* out) {
  std::string prefix;
  {
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    prefix = g_shm_prefix;
    const bool want_shm = !prefix.empty();
    int best = -1;
    for (size_t i = 0; i < g_pool.size(); ++i)
      if (g_pool[i].shm == want_shm && g_pool[i].bytes >= bytes && (best < 0 || g_pool[i].bytes < g_pool[(size_t)best].bytes)) best = (int)i;
    if (best >= 0) {
      *out = g_pool[(size_t)best];
      g_pool.erase(g_pool.begin() + best);
      return SRT_OK;
    }
  }
  HostRows h;
  h.bytes = bytes;
  if (prefix.empty()) {
    cudaError_t e = cudaMallocHost((void**)&h.p, bytes);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMallocHost(rows)");
  } else {
    h.shm = true;
    snprintf(h.name, sizeof h.name, "/%s.%d.%u", prefix.c_str(), (int)getpid(), g_shm_seq++);
    int fd = shm_open(h.name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0 || ftruncate(fd, (off_t)bytes) != 0) {
      if (fd >= 0) { close(fd); shm_unlink(h.name); }
      snprintf(g_cuda_err, sizeof g_cuda_err, "shm_open/ftruncate(%s, %zu bytes) failed", h.name, bytes);
      return SRT_E_CUDA;
    }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_POPULATE, fd, 0);
    close(fd);
    if (p == MAP_FAILED) {
      shm_unlink(h.name);
      snprintf(g_cuda_err, sizeof g_cuda_err, "mmap(%s, %zu bytes) failed", h.name, bytes);
      return SRT_E_CUDA;
    }
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) {
      munmap(p, bytes);
      shm_unlink(h.name);
      return cuda_fail(e, "cudaHostRegister(rows)");
    }
    h.p = (double*)p;
  }
  *out = h;
  return SRT_OK;
}

void host_rows_release(HostRows& h) {
  if (!h.p) return;
  std::lock_guard<std::mutex> lk(g_pool_mutex);
  g_pool.push_back(h);
  h = HostRows();
  // keep the pool small: beyond 2 buffers per device in use, the smallest goes
  const size_t keep = 2 * (g_ctx.size() ? g_ctx.size() : 1) + 2;
  while (g_pool.size() > keep) {
    size_t s = 0;
    for (size_t i = 1; i < g_pool.size(); ++i)
      if (g_pool[i].bytes < g_pool[s].bytes) s = i;
    host_rows_destroy(g_pool[s]);
    g_pool.erase(g_pool.begin() + (long)s);
  }
}

void pool_clear() {
  std::lock_guard<std::mutex> lk(g_pool_mutex);
  for (auto& h : g_pool) host_rows_destroy(h);
  g_pool.clear();
}
struct PoolAtExit { ~PoolAtExit() { for (auto& h : g_pool) if (h.shm && h.p) shm_unlink(h.name); } } g_pool_at_exit;  // no segment outlives the process

}  // namespace

struct RowChunk {
  HostRows mem;
  uint64_t rows = 0;
};

struct srt_result {
  int64_t nrays = 0;
  int outputper = 1;
  std::vector<double> summary;     // [nrays][SRT_SUMMARY_WORDS]
  std::vector<uint32_t> nheld;     // rows held per ray
  std::vector<uint32_t> chunk_of;  // which chunk holds the rows of a ray
  std::vector<uint64_t> offsets;   // first row of each ray inside its chunk
  std::mutex chunks_mutex;
  std::vector<RowChunk> chunks;    // one per (device, wave) that returned rows
  std::atomic<uint64_t> total_rows{0}, accepted{0}, attempts{0}, fastforwarded{0};
  int ndevices = 1;
};

namespace {

size_t log_budget_rows(DevCtx* c, size_t want_rows) {
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) free_b = (size_t)8 << 30;
  // the log and its sorted copy share 60 % of what is free right now (cached scratch counts as free)
  size_t avail = (free_b + c->log.cap + c->sorted.cap) / 10 * 6;
  size_t per_row = (SRT_LOG_WORDS + SRT_ROW_DOUBLES) * sizeof(double);
  size_t max_rows = avail / per_row;
  const char* env = getenv("SRT_LOG_ROWS");  // test hook: force a small log to exercise the re-run path
  if (env) max_rows = (size_t)strtoull(env, nullptr, 10);
  return want_rows < max_rows ? want_rows : max_rows;
}

// The rays ids[0..n) of the caller's arrays on device c: waves of (H2D, persistent trace kernel, (ray,row) ordering,
// D2H into host row memory), results written into res at the rays' own indices.  Runs in the calling thread.
int trace_on_device(DevCtx* c, srt_model* m, const std::vector<int64_t>& ids, const double* pos0, double* dir0, const double* w0,
                    const SrtRunParams& P, srt_result* res) {
  CK(cudaSetDevice(c->dev));
  std::lock_guard<std::mutex> batch_lock(c->batch_mutex);
  if (!c->stream) CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
  cudaStream_t st = c->stream;
  int rc;
  // rows a ray can hold at most (Q15): maxsteps rows, every outputper-th is kept
  const size_t max_rows_per_ray = (size_t)((P.maxsteps < 1 ? 1 : P.maxsteps) + res->outputper - 1) / res->outputper + 1;
  // A full log drops every ray still in flight, so after a wave that dropped rays the next waves are
  // limited to what is guaranteed to fit (or to what just completed, if that is more).
  size_t wave_limit = ids.size();
  std::vector<int64_t> pending(ids), todo;
  while (!pending.empty()) {
    const size_t n = pending.size() < wave_limit ? pending.size() : wave_limit;
    todo.assign(pending.begin(), pending.begin() + (long)n);
    pending.erase(pending.begin(), pending.begin() + (long)n);
    if ((rc = c->rays.reserve(n * 7 * sizeof(double)))) return rc;
    if ((rc = c->rays_h.reserve(n * 7 * sizeof(double)))) return rc;
    if ((rc = c->summary.reserve(n * SRT_SUMMARY_WORDS * sizeof(double)))) return rc;
    if ((rc = c->summary_h.reserve(n * SRT_SUMMARY_WORDS * sizeof(double)))) return rc;
    if ((rc = c->counters.reserve(8 * sizeof(unsigned long long)))) return rc;
    if ((rc = c->nlogged.reserve(n * sizeof(unsigned)))) return rc;
    if ((rc = c->offsets.reserve(n * sizeof(unsigned long long)))) return rc;
    if ((rc = c->small_h.reserve(n * (sizeof(unsigned) + sizeof(unsigned long long)) + 128))) return rc;
    size_t cap_rows = log_budget_rows(c, n * max_rows_per_ray);
    if (cap_rows == 0) return SRT_E_NOMEM;  // a log that cannot hold one worst-case ray may still hold the actual ones
    if ((rc = c->log.reserve(cap_rows * SRT_LOG_WORDS * sizeof(double)))) return rc;

    // AoS -> SoA staging in pinned memory, one H2D copy
    double* hr = (double*)c->rays_h.p;
    for (size_t s = 0; s < n; ++s) {
      const int64_t r = todo[s];
      hr[0 * n + s] = pos0[3 * r + 0]; hr[1 * n + s] = pos0[3 * r + 1]; hr[2 * n + s] = pos0[3 * r + 2];
      hr[3 * n + s] = dir0[3 * r + 0]; hr[4 * n + s] = dir0[3 * r + 1]; hr[5 * n + s] = dir0[3 * r + 2];
      hr[6 * n + s] = w0[r];
    }
    CK(cudaMemcpyAsync(c->rays.p, hr, n * 7 * sizeof(double), cudaMemcpyHostToDevice, st));
    SrtBatch B;
    B.nrays = (long long)n;
    B.rays = (double*)c->rays.p;
    B.log = (double*)c->log.p;
    B.log_capacity = cap_rows;
    B.summary = (double*)c->summary.p;
    B.counters = (unsigned long long*)c->counters.p;
    B.nlogged = (unsigned*)c->nlogged.p;
    B.order_mul = 1;
    B.shadow = 1;
    if ((rc = launch_trace(c, m, P, B, st))) return rc;
    unsigned long long* counters = (unsigned long long*)((char*)c->small_h.p);
    unsigned* nlogged_h = (unsigned*)((char*)c->small_h.p + 64);
    unsigned long long* offsets_h = (unsigned long long*)((char*)c->small_h.p + 64 + ((n * sizeof(unsigned) + 63) / 64) * 64);
    CK(cudaMemcpyAsync(counters, c->counters.p, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(nlogged_h, c->nlogged.p, n * sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(c->summary_h.p, c->summary.p, n * SRT_SUMMARY_WORDS * sizeof(double), cudaMemcpyDeviceToHost, st));
    CK(cudaMemcpyAsync(hr, c->rays.p, n * 7 * sizeof(double), cudaMemcpyDeviceToHost, st));  // dir0 is inout
    CK(cudaStreamSynchronize(st));
    res->attempts += counters[2];
    res->fastforwarded += counters[5];

    // offsets of the rays that completed in this wave
    uint64_t kept_rows = 0;
    std::vector<int64_t> again;
    const double* sh = (const double*)c->summary_h.p;
    for (size_t s = 0; s < n; ++s) {
      offsets_h[s] = kept_rows;
      const int flags = (int)sh[s * SRT_SUMMARY_WORDS + 23];
      if (flags & SRT_FLAG_DROPPED) { again.push_back(todo[s]); continue; }
      kept_rows += nlogged_h[s];
    }
    if (!again.empty()) {
      const size_t guaranteed = cap_rows / max_rows_per_ray > 0 ? cap_rows / max_rows_per_ray : 1;
      if (again.size() == n && n <= guaranteed) return SRT_E_NOMEM;  // even a guaranteed-size wave made no progress
      const size_t completed = n - again.size();
      wave_limit = completed > guaranteed ? completed : guaranteed;
    }
    uint32_t chunk_index = 0;
    if (kept_rows) {
      if ((rc = c->sorted.reserve(kept_rows * SRT_ROW_DOUBLES * sizeof(double)))) return rc;
      CK(cudaMemcpyAsync(c->offsets.p, offsets_h, n * sizeof(unsigned long long), cudaMemcpyHostToDevice, st));
      unsigned long long nlog = counters[0] < cap_rows ? counters[0] : cap_rows;
      srt_scatter_kernel<<<c->num_sms * 8, 256, 0, st>>>((const double*)c->log.p, nlog, (const unsigned long long*)c->offsets.p,
                                                         (const unsigned*)c->nlogged.p, (double*)c->sorted.p);
      g_launches++;
      CK(cudaGetLastError());
      RowChunk ch;
      if ((rc = host_rows_alloc(kept_rows * SRT_ROW_DOUBLES * sizeof(double), &ch.mem))) return rc;
      ch.rows = kept_rows;
      cudaError_t e = cudaMemcpyAsync(ch.mem.p, c->sorted.p, kept_rows * SRT_ROW_DOUBLES * sizeof(double), cudaMemcpyDeviceToHost, st);
      if (e == cudaSuccess) e = cudaStreamSynchronize(st);
      if (e != cudaSuccess) { host_rows_release(ch.mem); return cuda_fail(e, "D2H(rows)"); }
      std::lock_guard<std::mutex> lk(res->chunks_mutex);
      chunk_index = (uint32_t)res->chunks.size();
      res->chunks.push_back(ch);
      res->total_rows += kept_rows;
    }
    uint64_t acc = 0;
    for (size_t s = 0; s < n; ++s) {
      const int64_t r = todo[s];
      const int flags = (int)sh[s * SRT_SUMMARY_WORDS + 23];
      if (flags & SRT_FLAG_DROPPED) continue;
      memcpy(&res->summary[(size_t)r * SRT_SUMMARY_WORDS], sh + s * SRT_SUMMARY_WORDS, SRT_SUMMARY_WORDS * sizeof(double));
      res->nheld[(size_t)r] = nlogged_h[s];
      res->offsets[(size_t)r] = offsets_h[s];
      res->chunk_of[(size_t)r] = chunk_index;
      acc += (uint64_t)sh[s * SRT_SUMMARY_WORDS + 20] - 1;
      dir0[3 * r + 0] = hr[3 * n + s]; dir0[3 * r + 1] = hr[4 * n + s]; dir0[3 * r + 2] = hr[5 * n + s];
    }
    res->accepted += acc;
    pending.insert(pending.end(), again.begin(), again.end());
  }
  return SRT_OK;
}

}  // namespace

extern "C" {

int srt_set_rows_shm(const char* prefix) {
  if (prefix) {
    for (const char* q = prefix; *q; ++q)
      if (*q == '/' || q - prefix > 40) return SRT_E_ARG;
  }
  const std::string want = prefix ? prefix : "";
  {
    std::lock_guard<std::mutex> lk(g_pool_mutex);
    if (g_shm_prefix == want) return SRT_OK;
  }
  pool_clear();  // pooled buffers of the other kind are of no use any more
  std::lock_guard<std::mutex> lk(g_pool_mutex);
  g_shm_prefix = want;
  return SRT_OK;
}

int srt_trace_batch(const srt_model* m, int64_t nrays, const double* pos0, double* dir0, const double* w0, const srt_run_params* p,
                    srt_result** out) {
  if (!m || !p || !out || nrays < 0 || (nrays > 0 && (!pos0 || !dir0 || !w0))) return SRT_E_ARG;
  if (nrays > 0xffffffffll) return SRT_E_ARG;
  *out = nullptr;
  std::vector<DevCtx*> devs;
  int rc = active_devices(devs);
  if (rc) return rc;
  srt_result* res = new (std::nothrow) srt_result();
  if (!res) return SRT_E_NOMEM;
  res->nrays = nrays;
  res->outputper = p->outputper < 1 ? 1 : p->outputper;
  res->summary.assign((size_t)nrays * SRT_SUMMARY_WORDS, 0.0);
  res->offsets.assign((size_t)nrays, 0);
  res->nheld.assign((size_t)nrays, 0);
  res->chunk_of.assign((size_t)nrays, 0);
  if (nrays == 0) { *out = res; return SRT_OK; }
  SrtRunParams P;
  memcpy(&P, p, sizeof P);
  P.outputper = res->outputper;
  srt_model* mm = const_cast<srt_model*>(m);

  // Rays are independent: ray i goes to device (i + i / ndev) mod ndev -- neighbouring launch points,
  // which have similar lifetimes, spread over all devices, and the deal rotates so that a lattice whose fastest index
  // has ndev values is still mixed (same rule as shard.py) -- one host thread per device, no exchange between them;
  // every device writes the results of its rays at their own indices: the result is in input order whatever ran where.
  const size_t ndev = (size_t)nrays < devs.size() ? (size_t)nrays : devs.size();
  res->ndevices = (int)ndev;
  int prev_dev = 0;
  cudaGetDevice(&prev_dev);
  std::vector<std::vector<int64_t>> ids(ndev);
  for (size_t d = 0; d < ndev; ++d) ids[d].reserve((size_t)nrays / ndev + 1);
  for (int64_t i = 0; i < nrays; ++i) ids[(size_t)((i + i / (int64_t)ndev) % (int64_t)ndev)].push_back(i);
  std::vector<int> rcs(ndev, SRT_OK);
  std::vector<std::string> errs(ndev);
  if (ndev == 1) {
    rcs[0] = trace_on_device(devs[0], mm, ids[0], pos0, dir0, w0, P, res);
    errs[0] = g_cuda_err;
  } else {
    std::vector<std::thread> th;
    for (size_t d = 0; d < ndev; ++d)
      th.emplace_back([&, d]() {
        rcs[d] = trace_on_device(devs[d], mm, ids[d], pos0, dir0, w0, P, res);
        errs[d] = g_cuda_err;  // thread-local: handed to the caller below
      });
    for (auto& t : th) t.join();
  }
  cudaSetDevice(prev_dev);
  for (size_t d = 0; d < ndev; ++d)
    if (rcs[d]) {
      snprintf(g_cuda_err, sizeof g_cuda_err, "device %d: %s", devs[d]->dev, errs[d].c_str());
      srt_result_free(res);
      return rcs[d];
    }
  *out = res;
  return SRT_OK;
}

int64_t srt_result_nrays(const srt_result* r) { return r ? r->nrays : 0; }
int64_t srt_result_total_rows(const srt_result* r) { return r ? (int64_t)r->total_rows.load() : 0; }
int64_t srt_result_accepted_steps(const srt_result* r) { return r ? (int64_t)r->accepted.load() : 0; }
int64_t srt_result_attempts(const srt_result* r) { return r ? (int64_t)r->attempts.load() : 0; }
int64_t srt_result_fastforwarded_steps(const srt_result* r) { return r ? (int64_t)r->fastforwarded.load() : 0; }
int srt_result_devices(const srt_result* r) { return r ? r->ndevices : 0; }
static bool ray_ok(const srt_result* r, int64_t ray) { return r && ray >= 0 && ray < r->nrays; }
int64_t srt_result_nrows(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? r->nheld[(size_t)ray] : SRT_E_ARG; }
int64_t srt_result_nsteps(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? (int64_t)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 20] : SRT_E_ARG; }
int srt_result_stopcond(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? (int)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 21] : SRT_E_ARG; }
int srt_result_flags(const srt_result* r, int64_t ray) { return ray_ok(r, ray) ? ((int)r->summary[(size_t)ray * SRT_SUMMARY_WORDS + 23] & SRT_FLAG_FIXED_CONE_EXIT) : SRT_E_ARG; }
int srt_result_ray_view(const srt_result* r, int64_t ray, const double** rows, int64_t* nrows) {
  if (!ray_ok(r, ray) || !rows || !nrows) return SRT_E_ARG;
  *nrows = r->nheld[(size_t)ray];
  *rows = nullptr;
  if (*nrows && r->chunk_of[(size_t)ray] < r->chunks.size())
    *rows = r->chunks[r->chunk_of[(size_t)ray]].mem.p + r->offsets[(size_t)ray] * SRT_ROW_DOUBLES;
  return SRT_OK;
}
int srt_result_copy_ray(const srt_result* r, int64_t ray, double* rows) {
  const double* src;
  int64_t n;
  int rc = srt_result_ray_view(r, ray, &src, &n);
  if (rc) return rc;
  if (n && (!rows || !src)) return SRT_E_ARG;
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
int srt_result_chunk_count(const srt_result* r) { return r ? (int)r->chunks.size() : 0; }
int srt_result_chunk_info(const srt_result* r, int chunk, char* shm_name, int name_capacity, int64_t* nrows, const double** rows) {
  if (!r || chunk < 0 || (size_t)chunk >= r->chunks.size()) return SRT_E_ARG;
  const RowChunk& c = r->chunks[(size_t)chunk];
  if (shm_name && name_capacity > 0) snprintf(shm_name, (size_t)name_capacity, "%s", c.mem.shm ? c.mem.name : "");
  if (nrows) *nrows = (int64_t)c.rows;
  if (rows) *rows = c.mem.p;
  return SRT_OK;
}
int srt_result_ray_location(const srt_result* r, int64_t nrays, int32_t* chunk, int64_t* first_row, int64_t* nrows) {
  if (!r || nrays != r->nrays) return SRT_E_ARG;
  for (int64_t i = 0; i < nrays; ++i) {
    if (chunk) chunk[i] = (int32_t)r->chunk_of[(size_t)i];
    if (first_row) first_row[i] = (int64_t)r->offsets[(size_t)i];
    if (nrows) nrows[i] = (int64_t)r->nheld[(size_t)i];
  }
  return SRT_OK;
}
void srt_result_free(srt_result* r) {
  if (!r) return;
  for (auto& c : r->chunks) host_rows_release(c.mem);
  delete r;
}

int srt_fp64_peak(int fma, int iters, double* gflops) {
  if (!gflops || iters < 1) return SRT_E_ARG;
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  double* d = nullptr;
  CK(cudaMalloc(&d, 8));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  cudaError_t e = cudaEventCreate(&e0);
  if (e == cudaSuccess) e = cudaEventCreate(&e1);
  const int blocks = c->num_sms * 8, threads = 256;
  float best_ms = 0.f;
  for (int rep = 0; rep < 4 && e == cudaSuccess; ++rep) {  // first pass warms up; best of the next three
    e = cudaEventRecord(e0);
    if (fma) srt_fp64_peak_kernel<1><<<blocks, threads>>>(iters, 1.0, d);
    else srt_fp64_peak_kernel<0><<<blocks, threads>>>(iters, 1.0, d);
    g_launches++;
    if (e == cudaSuccess) e = cudaEventRecord(e1);
    if (e == cudaSuccess) e = cudaEventSynchronize(e1);
    float ms = 0.f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, e0, e1);
    if (rep > 0 && e == cudaSuccess && (best_ms == 0.f || ms < best_ms)) best_ms = ms;
  }
  if (e0) cudaEventDestroy(e0);
  if (e1) cudaEventDestroy(e1);
  cudaFree(d);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "srt_fp64_peak");
  if (!(best_ms > 0.f)) return cuda_fail(cudaErrorUnknown, "srt_fp64_peak: no time measured");
  *gflops = 2.0 * 8.0 * (double)iters * blocks * threads / (best_ms * 1e-3) / 1e9;
  return SRT_OK;
}

// releases every cached scratch buffer of the library (device scratch, pinned staging, pooled row buffers)
int srt_release_scratch(void) {
  pool_clear();
  std::lock_guard<std::mutex> lk(g_dev_mutex);
  int prev = 0;
  if (cudaGetDevice(&prev) != cudaSuccess) return SRT_OK;
  for (auto& c : g_ctx) {
    std::lock_guard<std::mutex> bl(c->batch_mutex);
    cudaSetDevice(c->dev);
    c->rays.release(); c->log.release(); c->sorted.release(); c->summary.release(); c->counters.release(); c->nlogged.release(); c->offsets.release();
    c->rays_h.release(); c->summary_h.release(); c->small_h.release();
  }
  cudaSetDevice(prev);
  return SRT_OK;
}

}  // extern "C"
