// srt_api.cu -- the C ABI of include/srt.h over the sm_100a kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo (build.py).
// This unit holds the one-thread-per-ray trace kernel: 256-thread blocks (one per SM, <= 255 registers)
// that start every attempt together (SRT_BLOCK_SYNC), elementary functions inlined (SRT_INLINE_MATH in
// build.py).  The team-of-8 kernel is compiled separately (srt_team_tu.cu).
#ifndef SRT_BLOCK_THREADS
#define SRT_BLOCK_THREADS 256
#endif
#ifndef SRT_THREAD_MIN_BLOCKS
#define SRT_THREAD_MIN_BLOCKS 1
#endif
#ifndef SRT_BLOCK_SYNC
#define SRT_BLOCK_SYNC 1
#endif
#include "srt_kernels.cuh"
#include "srt_team_tu.h"
#include "srt_host_setup.h"
#include "srt_grid_setup.h"
#include "srt_igrf_setup.h"
#include "../../include/srt.h"

#include <atomic>
#include <mutex>
#include <new>
#include <vector>

static_assert(SRT_ROW_DOUBLES == 20, "row layout");
static_assert(sizeof(srt_run_params) == sizeof(SrtRunParams), "params layout");

namespace {

thread_local char g_cuda_err[512] = "";
std::atomic<long long> g_launches{0};
std::mutex g_model_mutex;  // c_srt_model is one __constant__ object: launches with different models are serialised
int g_device = -1;
int g_num_sms = 0;

int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_cuda_err, sizeof g_cuda_err, "%s: %s", what, cudaGetErrorString(e));
  return SRT_E_CUDA;
}
#define CK(call)                                           \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call);    \
  } while (0)

int ensure_device() {
  if (g_device >= 0) return SRT_OK;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    snprintf(g_cuda_err, sizeof g_cuda_err, "no CUDA device: %s", cudaGetErrorString(e));
    return SRT_E_NODEVICE;
  }
  int dev = 0;
  CK(cudaGetDevice(&dev));
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, dev));
  g_device = dev;
  g_num_sms = prop.multiProcessorCount;
  return SRT_OK;
}

typedef void (*trace_kernel_t)(const SrtRunParams, const SrtBatch);
typedef void (*eval_kernel_t)(long long, const double*, double*);

trace_kernel_t pick_trace(const SrtModel& M) {
  if (M.density == SRT_DENS_NGO) return M.use_igrf ? srt_trace_kernel<SRT_DENS_NGO, 1> : srt_trace_kernel<SRT_DENS_NGO, 0>;
  return M.use_igrf ? srt_trace_kernel<SRT_DENS_GRID, 1> : srt_trace_kernel<SRT_DENS_GRID, 0>;
}
eval_kernel_t pick_eval(const SrtModel& M) {
  if (M.density == SRT_DENS_NGO) return M.use_igrf ? srt_eval_kernel<SRT_DENS_NGO, 1> : srt_eval_kernel<SRT_DENS_NGO, 0>;
  return M.use_igrf ? srt_eval_kernel<SRT_DENS_GRID, 1> : srt_eval_kernel<SRT_DENS_GRID, 0>;
}

}  // namespace

struct srt_model {
  SrtModel M;
  double nus[SRT_NSPEC];
  // device copies of the grid (modelnum 3)
  double* d_nodes = nullptr;
  double* d_axes = nullptr;
  double* d_coef = nullptr;  // per-cell tricubic coefficients
  int blocks_per_sm[2] = {0, 0};  // [thread-per-ray, team-of-8]
  int grid[2] = {0, 0};
};

struct srt_result {
  int64_t nrays = 0;
  int outputper = 1;
  std::vector<double> summary;    // [nrays][SRT_SUMMARY_WORDS]
  std::vector<uint64_t> offsets;  // first row of each ray in `rows`
  std::vector<uint32_t> nheld;    // rows held per ray
  double* rows = nullptr;         // pinned host memory, [total_rows][20]
  size_t rows_bytes = 0;          // capacity of `rows`
  uint64_t total_rows = 0;
  uint64_t accepted = 0, attempts = 0;
};

namespace {

// 0 = one thread per ray, 1 = team of 8 lanes per ray.  SRT_KERNEL=thread|team overrides; by default
// the team form is used while the batch cannot fill the machine with one thread per ray.
int pick_mode(const srt_model* m, long long nrays) {
  const char* env = getenv("SRT_KERNEL");
  if (env && !strcmp(env, "thread")) return 0;
  if (env && !strcmp(env, "team")) return 1;
  // Since the team form steps block-synchronously with 6 equal rounds per attempt it is the faster one at every
  // batch size measured (profiles/r1c_modes.txt: 34.7 M vs 32.9 M ray-steps/s at 40 k rays, 26.9 M vs 7.6 M at
  // 10 k, config-3 shape 30.8 M vs 23.2 M), so it is the default; SRT_KERNEL=thread selects the other form.
  (void)m; (void)nrays;
  return 1;
}

// persistent launch geometry of a batch: mode (0 thread / 1 team), block, grid = one resident wave at most
int launch_config(srt_model* m, long long nrays, int* mode_out, int* block_out, int* grid_out, int* blocks_per_sm_out) {
  if (!m->grid[0]) {
    int nb = 0;
    CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, pick_trace(m->M), SRT_BLOCK_THREADS, 0));
    if (nb < 1) nb = 1;
    m->blocks_per_sm[0] = nb;
    m->grid[0] = nb * g_num_sms;
  }
  const int mode = pick_mode(m, nrays);
  int block = SRT_BLOCK_THREADS, nb = m->blocks_per_sm[0];
  if (mode) {
    block = srt_team_block_threads(nrays, g_num_sms);
    CK(srt_team_occupancy(m->M, block, &nb));
    if (nb < 1) nb = 1;
  }
  const long long lanes_per_ray = mode ? 8 : 1;
  long long blocks_needed = (nrays * lanes_per_ray + block - 1) / block;
  long long resident = (long long)nb * g_num_sms;
  int grid = (int)(blocks_needed < resident ? blocks_needed : resident);
  if (grid < 1) grid = 1;
  *mode_out = mode; *block_out = block; *grid_out = grid;
  if (blocks_per_sm_out) *blocks_per_sm_out = nb;
  return SRT_OK;
}

// upload the model and launch the persistent trace kernel on `stream`; caller holds g_model_mutex
// multiplier of the queue permutation (srt_queue_ray): the integer nearest nrays/golden ratio that is coprime to nrays
unsigned long long ray_order_mul(long long nrays) {
  const char* e = getenv("SRT_RAY_ORDER");
  if ((e && atoi(e) == 0) || nrays < 4) return 1ull;
  unsigned long long n = (unsigned long long)nrays, mul = (unsigned long long)((double)nrays * 0.6180339887498949);
  for (;; ++mul) {
    unsigned long long a = n, b = mul % n;
    while (b) { unsigned long long t = a % b; a = b; b = t; }
    if (a == 1ull && mul % n != 0ull) return mul % n;
  }
}

int launch_trace(srt_model* m, const SrtRunParams& P, const SrtBatch& B0, cudaStream_t stream) {
  SrtBatch B = B0;
  B.order_mul = ray_order_mul(B.nrays);
  { const char* e = getenv("SRT_TEAM_COMPACT"); B.compact = (e && atoi(e) == 0) ? 0 : 1; }
  { const char* e = getenv("SRT_TEAM_PHASE_ALIGN"); B.phase_align = (e && atoi(e) != 0) ? 1 : 0; }  // off until measured (open item)
  int mode, block, grid;
  int rc = launch_config(m, B.nrays, &mode, &block, &grid, nullptr);
  if (rc) return rc;
  CK(cudaMemsetAsync(B.counters, 0, 8 * sizeof(unsigned long long), stream));
  if (mode) {
    CK(srt_team_launch(m->M, P, B, block, grid, stream));
  } else {
    CK(cudaMemcpyToSymbolAsync(c_srt_model, &m->M, sizeof(SrtModel), 0, cudaMemcpyHostToDevice, stream));
    pick_trace(m->M)<<<grid, block, 0, stream>>>(P, B);
  }
  g_launches++;
  CK(cudaGetLastError());
  return SRT_OK;
}

int finish_model(srt_model* mm, int yearday, int msec, int use_igrf, int use_tsyganenko) {
  if (use_tsyganenko != 0) return SRT_E_UNSUPPORTED;
  if (use_igrf != 0 && use_igrf != 1) return SRT_E_ARG;
  srt_host::set_time(mm->M, yearday, msec);
  mm->M.use_igrf = use_igrf;
  srt_host::igrf_recalc(yearday, msec, mm->M.igrf);
  for (int i = 0; i < SRT_NSPEC; ++i) mm->nus[i] = 0.0;
  return SRT_OK;
}

}  // namespace

extern "C" {

int srt_init(int device) {
  if (device >= 0) {
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) {
      snprintf(g_cuda_err, sizeof g_cuda_err, "cudaSetDevice(%d): %s", device, cudaGetErrorString(e));
      return SRT_E_NODEVICE;
    }
    g_device = -1;
  }
  return ensure_device();
}

const char* srt_strerror(int code) {
  switch (code) {
    case SRT_OK: return "ok";
    case SRT_E_NODEVICE: return "no usable CUDA device (libsrt has no CPU path)";
    case SRT_E_IO: return "cannot open or parse model file";
    case SRT_E_ARG: return "invalid argument";
    case SRT_E_UNSUPPORTED: return "unsupported option (use_tsyganenko must be 0)";
    case SRT_E_CUDA: return "CUDA runtime error";
    case SRT_E_NOMEM: return "trajectory log too small for a single ray";
  }
  return "unknown error";
}
const char* srt_last_cuda_error(void) { return g_cuda_err; }
int64_t srt_launch_count(void) { return g_launches.load(); }

int srt_model_ngo(srt_model** m, const char* cfg, int yearday, int msec, int use_igrf, int use_tsyganenko) {
  if (!m || !cfg) return SRT_E_ARG;
  *m = nullptr;
  srt_model* mm = new (std::nothrow) srt_model();
  if (!mm) return SRT_E_NOMEM;
  memset(&mm->M, 0, sizeof mm->M);
  int rc = srt_host::load_ngo_config(cfg, mm->M);
  if (rc != 0) { delete mm; return rc == -3 ? SRT_E_UNSUPPORTED : SRT_E_IO; }
  rc = finish_model(mm, yearday, msec, use_igrf, use_tsyganenko);
  if (rc) { delete mm; return rc; }
  *m = mm;
  return SRT_OK;
}

int srt_model_grid(srt_model** m, const char* path, int yearday, int msec, int use_igrf, int use_tsyganenko) {
  if (!m || !path) return SRT_E_ARG;
  *m = nullptr;
  int rc = ensure_device();
  if (rc) return rc;
  srt_model* mm = new (std::nothrow) srt_model();
  if (!mm) return SRT_E_NOMEM;
  memset(&mm->M, 0, sizeof mm->M);
  srt_host::HostGrid hg;
  rc = srt_host::load_grid(path, mm->M, hg);
  if (rc != 0) { delete mm; return rc == -3 ? SRT_E_UNSUPPORTED : SRT_E_IO; }
  rc = finish_model(mm, yearday, msec, use_igrf, use_tsyganenko);
  if (rc) { delete mm; return rc; }
  size_t nb = hg.nodes.size() * sizeof(double);
  size_t ax = (hg.xs.size() + hg.ys.size() + hg.zs.size()) * sizeof(double);
  const SrtGrid& hgd = mm->M.grid;
  // coefficient table over the extended lattice (edge and outside cells included), srt_tricubic.h
  const size_t ncells = (hgd.nx > 1 && hgd.ny > 1 && hgd.nz > 1) ? srt_grid_cell_count(hgd) : 0;
  if (cudaMalloc(&mm->d_nodes, nb) != cudaSuccess || cudaMalloc(&mm->d_axes, ax) != cudaSuccess ||
      (ncells && cudaMalloc(&mm->d_coef, ncells * 256 * sizeof(double)) != cudaSuccess)) {
    rc = cuda_fail(cudaGetLastError(), "cudaMalloc(grid)");
    srt_model_free(mm);
    return rc;
  }
  cudaMemcpy(mm->d_nodes, hg.nodes.data(), nb, cudaMemcpyHostToDevice);
  cudaMemcpy(mm->d_axes, hg.xs.data(), hg.xs.size() * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(mm->d_axes + hg.xs.size(), hg.ys.data(), hg.ys.size() * 8, cudaMemcpyHostToDevice);
  cudaMemcpy(mm->d_axes + hg.xs.size() + hg.ys.size(), hg.zs.data(), hg.zs.size() * 8, cudaMemcpyHostToDevice);
  mm->M.grid.nodes = mm->d_nodes;
  mm->M.grid.coef = nullptr;
  mm->M.grid.xs = mm->d_axes;
  mm->M.grid.ys = mm->d_axes + hg.xs.size();
  mm->M.grid.zs = mm->d_axes + hg.xs.size() + hg.ys.size();
  if (ncells) {
    // per-cell tricubic coefficients, once (SRT_GRID_COEF=0 keeps the on-the-fly path: test hook)
    const char* env = getenv("SRT_GRID_COEF");
    if (!(env && !strcmp(env, "0"))) {
      long long threads = (long long)ncells * 4;
      int blocks = (int)((threads + 127) / 128 < (long long)g_num_sms * 16 ? (threads + 127) / 128 : (long long)g_num_sms * 16);
      srt_coef_kernel<<<blocks, 128>>>(mm->M.grid, mm->d_coef);
      g_launches++;
      cudaDeviceSynchronize();
      mm->M.grid.coef = mm->d_coef;
    }
  }
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) { srt_model_free(mm); return cuda_fail(e, "grid upload"); }
  *m = mm;
  return SRT_OK;
}

void srt_model_free(srt_model* m) {
  if (!m) return;
  if (m->d_nodes) cudaFree(m->d_nodes);
  if (m->d_axes) cudaFree(m->d_axes);
  if (m->d_coef) cudaFree(m->d_coef);
  delete m;
}

int srt_model_species(const srt_model* m, int* nspec, double* qs, double* ms, double* nus) {
  if (!m) return SRT_E_ARG;
  if (nspec) *nspec = m->M.nspec;
  for (int i = 0; i < m->M.nspec; ++i) {
    if (qs) qs[i] = m->M.qs[i];
    if (ms) ms[i] = m->M.ms[i];
    if (nus) nus[i] = m->nus[i];
  }
  return SRT_OK;
}

int srt_model_eval(const srt_model* m, int64_t n, const double* x, double* out) {
  if (!m || n < 0 || (n > 0 && (!x || !out))) return SRT_E_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  if (n == 0) return SRT_OK;
  double *dx = nullptr, *dout = nullptr;
  CK(cudaMalloc(&dx, n * 3 * sizeof(double)));
  if (cudaMalloc(&dout, n * 19 * sizeof(double)) != cudaSuccess) { cudaFree(dx); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
  cudaMemcpy(dx, x, n * 3 * sizeof(double), cudaMemcpyHostToDevice);
  {
    std::lock_guard<std::mutex> lk(g_model_mutex);
    cudaMemcpyToSymbol(c_srt_model, &m->M, sizeof(SrtModel));
    int blocks = (int)((n + 127) / 128);
    if (blocks > g_num_sms * 8) blocks = g_num_sms * 8;
    pick_eval(m->M)<<<blocks, 128>>>(n, dx, dout);
    g_launches++;
    cudaDeviceSynchronize();
  }
  cudaMemcpy(out, dout, n * 19 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(dx);
  cudaFree(dout);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_fail(e, "srt_model_eval");
  return SRT_OK;
}

int srt_kernel_info(const srt_model* m, int64_t nrays, int* mode, int* regs, int* local_bytes, int* blocks_per_sm, int* grid, int* block) {
  if (!m) return SRT_E_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  srt_model* mm = const_cast<srt_model*>(m);
  int md, blk, grd, nb;
  rc = launch_config(mm, nrays, &md, &blk, &grd, &nb);
  if (rc) return rc;
  cudaFuncAttributes fa;
  if (md) CK(srt_team_attributes(m->M, blk, &fa));
  else CK(cudaFuncGetAttributes(&fa, pick_trace(m->M)));
  if (mode) *mode = md;
  if (regs) *regs = fa.numRegs;
  if (local_bytes) *local_bytes = (int)fa.localSizeBytes;
  if (blocks_per_sm) *blocks_per_sm = nb;
  if (grid) *grid = grd;
  if (block) *block = blk;
  return SRT_OK;
}

int srt_trace_batch_device(const srt_model* m, int64_t nrays, double* d_rays, const srt_run_params* p, double* d_log, int64_t log_capacity,
                           double* d_summary, uint64_t* d_counters, void* stream) {
  if (!m || !p || nrays < 0 || !d_rays || !d_summary || !d_counters || log_capacity < 0 || (log_capacity > 0 && !d_log)) return SRT_E_ARG;
  int rc = ensure_device();
  if (rc) return rc;
  SrtRunParams P;
  memcpy(&P, p, sizeof P);
  SrtBatch B;
  B.nrays = nrays;
  B.rays = d_rays;
  B.log = d_log;
  B.log_capacity = (unsigned long long)log_capacity;
  B.summary = d_summary;
  B.counters = (unsigned long long*)d_counters;
  B.nlogged = nullptr;
  B.order_mul = 1;  // set by launch_trace
  std::lock_guard<std::mutex> lk(g_model_mutex);
  return launch_trace(const_cast<srt_model*>(m), P, B, (cudaStream_t)stream);
}

}  // extern "C"

#include "srt_api_host.inl"
