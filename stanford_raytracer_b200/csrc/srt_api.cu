// srt_api.cu -- the C ABI of include/srt.h over the sm_100a kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false -lineinfo (build.py).
// This unit holds the one-thread-per-ray trace kernel: 256-thread blocks (one per SM, <= 255 registers)
// that start every attempt together (SRT_BLOCK_SYNC), elementary functions inlined (SRT_INLINE_MATH in
// build.py).  The team-of-8 kernel is compiled separately (srt_team_tu.cu).
//
// Devices.  The library drives every CUDA device selected by srt_init() / srt_init_devices(): one context per
// device (stream, scratch buffers, the state of that device's __constant__ model copies), one replica of each
// model per device (grid arrays, built lazily), one host thread per device inside srt_trace_batch().
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
#include "srt_scattered_setup.h"
#include "srt_igrf_setup.h"
#include "../../include/srt.h"

#include <atomic>
#include <map>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

static_assert(SRT_ROW_DOUBLES == 20, "row layout");
static_assert(sizeof(srt_run_params) == sizeof(SrtRunParams), "params layout");

namespace {

thread_local char g_cuda_err[512] = "";
std::atomic<long long> g_launches{0};
std::atomic<uint64_t> g_model_serial{0};

int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_cuda_err, sizeof g_cuda_err, "%s: %s", what, cudaGetErrorString(e));
  return SRT_E_CUDA;
}
#define CK(call)                                           \
  do {                                                     \
    cudaError_t e_ = (call);                               \
    if (e_ != cudaSuccess) return cuda_fail(e_, #call);    \
  } while (0)

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
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
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
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// Everything the library keeps per CUDA device.
struct DevCtx {
  int dev = -1;
  int num_sms = 0;
  cudaStream_t stream = nullptr;  // the host-buffer path's stream on this device
  // The model is read by the kernels from ONE __constant__ object per translation unit and device ([0] this unit:
  // thread kernel + eval kernel, [1] the team unit).  `resident` is the serial of the model it holds; `readers` are
  // events recorded after every launch that reads it.  A launch with the same model uploads nothing; a launch with
  // another model first makes its stream wait for every earlier reader, so a kernel in flight on any stream never
  // sees the symbol change under it.
  std::mutex sym_mutex;
  uint64_t resident[2] = {0, 0};
  std::vector<cudaEvent_t> readers[2];
  // scratch of srt_trace_batch on this device (one host-buffer batch at a time per device)
  std::mutex batch_mutex;
  DevBuf rays, log, sorted, summary, counters, nlogged, offsets;
  PinBuf rays_h, summary_h, small_h;
  // device address of g_eval_static on this device (srt_model_eval_noalloc); taken once, by cudaGetSymbolAddress
  double* eval_static = nullptr;
};

// Scratch of srt_model_eval_noalloc: the staged points ([n][3]) and results ([n][19]) of one call, in static device
// memory, so that the call allocates nothing.  Every device that loads the library holds these 2.75 MiB until exit.
// In host code this name is a stand-in, not an address: the kernels get the pointer cudaGetSymbolAddress returns.
#define SRT_EVAL_STATIC_MAX_POINTS 16384
static __device__ __align__(256) double g_eval_static[(size_t)SRT_EVAL_STATIC_MAX_POINTS * 22];

std::mutex g_dev_mutex;
std::vector<std::unique_ptr<DevCtx>> g_ctx;  // every device ever used by this process
std::vector<DevCtx*> g_active;               // the devices srt_trace_batch() spreads a batch over

DevCtx* ctx_for_device(int dev) {
  std::lock_guard<std::mutex> lk(g_dev_mutex);
  for (auto& c : g_ctx)
    if (c->dev == dev) return c.get();
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) return nullptr;
  std::unique_ptr<DevCtx> c(new DevCtx());
  c->dev = dev;
  c->num_sms = prop.multiProcessorCount;
  g_ctx.push_back(std::move(c));
  return g_ctx.back().get();
}

// the context of the calling thread's current device; SRT_E_NODEVICE without one
int current_ctx(DevCtx** out) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    snprintf(g_cuda_err, sizeof g_cuda_err, "no CUDA device: %s", cudaGetErrorString(e));
    return SRT_E_NODEVICE;
  }
  int dev = 0;
  CK(cudaGetDevice(&dev));
  DevCtx* c = ctx_for_device(dev);
  if (!c) return cuda_fail(cudaGetLastError(), "cudaGetDeviceProperties");
  *out = c;
  return SRT_OK;
}

// the devices a host-buffer batch runs on: what srt_init*/ selected, else the current device
int active_devices(std::vector<DevCtx*>& out) {
  {
    std::lock_guard<std::mutex> lk(g_dev_mutex);
    out = g_active;
  }
  if (!out.empty()) return SRT_OK;
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  out.push_back(c);
  return SRT_OK;
}

typedef void (*trace_kernel_t)(const SrtRunParams, const SrtBatch);
typedef void (*eval_kernel_t)(long long, const double*, double*);

// kernels are instantiated per density kind and field code (bit 0 IGRF, bit 1 TS05 external field)
int density_slot(const SrtModel& M) { return M.density == SRT_DENS_NGO ? 0 : (M.density == SRT_DENS_SCAT ? 2 : 1); }
trace_kernel_t pick_trace(const SrtModel& M) {
  static const trace_kernel_t k[3][4] = {
      {srt_trace_kernel<SRT_DENS_NGO, 0>, srt_trace_kernel<SRT_DENS_NGO, 1>, srt_trace_kernel<SRT_DENS_NGO, 2>, srt_trace_kernel<SRT_DENS_NGO, 3>},
      {srt_trace_kernel<SRT_DENS_GRID, 0>, srt_trace_kernel<SRT_DENS_GRID, 1>, srt_trace_kernel<SRT_DENS_GRID, 2>, srt_trace_kernel<SRT_DENS_GRID, 3>},
      {srt_trace_kernel<SRT_DENS_SCAT, 0>, srt_trace_kernel<SRT_DENS_SCAT, 1>, srt_trace_kernel<SRT_DENS_SCAT, 2>, srt_trace_kernel<SRT_DENS_SCAT, 3>}};
  return k[density_slot(M)][(M.use_igrf ? 1 : 0) | (M.ts.on ? 2 : 0)];
}
eval_kernel_t pick_eval(const SrtModel& M) {
  static const eval_kernel_t k[3][4] = {
      {srt_eval_kernel<SRT_DENS_NGO, 0>, srt_eval_kernel<SRT_DENS_NGO, 1>, srt_eval_kernel<SRT_DENS_NGO, 2>, srt_eval_kernel<SRT_DENS_NGO, 3>},
      {srt_eval_kernel<SRT_DENS_GRID, 0>, srt_eval_kernel<SRT_DENS_GRID, 1>, srt_eval_kernel<SRT_DENS_GRID, 2>, srt_eval_kernel<SRT_DENS_GRID, 3>},
      {srt_eval_kernel<SRT_DENS_SCAT, 0>, srt_eval_kernel<SRT_DENS_SCAT, 1>, srt_eval_kernel<SRT_DENS_SCAT, 2>, srt_eval_kernel<SRT_DENS_SCAT, 3>}};
  return k[density_slot(M)][(M.use_igrf ? 1 : 0) | (M.ts.on ? 2 : 0)];
}

// ---- set-up kernel of the gridded model: tricubic_compute_finite_difference_derivatives (libtricubic.f95:722-793).
// One pass = array slot `src` -> slot `dst` of every node record along `axis` (0: x, 1: y, 2: z): one-sided
// differences at the two ends, centred `(f(i+1) - f(i-1)) / 2 / h` inside; `enabled` = the axis (and, for the mixed
// derivatives, every axis involved) has more than two nodes, else the slot is zero.  One thread per (node, species);
// a pass reads only slot `src` and writes only slot `dst`, so there is no hazard inside a launch.
__global__ void srt_fd_kernel(double* __restrict__ nodes, int nx, int ny, int nz, int nspec, int src, int dst, int axis, double h, int enabled) {
  const long long total = (long long)nx * ny * nz * 4;
  for (long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
    const int s = (int)(id & 3);
    if (s >= nspec) continue;
    const long long node = id >> 2;
    const int i = (int)(node % nx), j = (int)((node / nx) % ny), k = (int)(node / ((long long)nx * ny));
    double* rec = nodes + node * 32;
    if (!enabled) { rec[4 * dst + s] = 0.0; continue; }
    const int c = axis == 0 ? i : (axis == 1 ? j : k);
    const int m = axis == 0 ? nx : (axis == 1 ? ny : nz);
    const long long stride = (axis == 0 ? 1ll : (axis == 1 ? (long long)nx : (long long)nx * ny)) * 32;
    double v;
    if (c == 0) v = (rec[stride + 4 * src + s] - rec[4 * src + s]) / h;
    else if (c == m - 1) v = (rec[4 * src + s] - rec[-stride + 4 * src + s]) / h;
    else v = (rec[stride + 4 * src + s] - rec[-stride + 4 * src + s]) / 2.0 / h;
    rec[4 * dst + s] = v;
  }
}

// Self-test of srt_divn (srt_math.h) against the compiler's own IEEE division: groups of four pseudo-random operand
// pairs -- random bit patterns, i.e. every exponent incl. subnormals, infinities and NaNs, plus zeros, equal operands,
// powers of two and mantissas of all ones -- one group in eight with an operand that forces the group onto the slow path.
__device__ __forceinline__ unsigned long long srt_mix64(unsigned long long z) {
  z += 0x9e3779b97f4a7c15ull;
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
  return z ^ (z >> 31);
}
__global__ void srt_divtest_kernel(long long ngroups, unsigned long long seed, unsigned long long* mismatches) {
  unsigned long long bad = 0;
  for (long long g = blockIdx.x * (long long)blockDim.x + threadIdx.x; g < ngroups; g += (long long)gridDim.x * blockDim.x) {
    double a[4], b[4], q[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      unsigned long long ua = srt_mix64(seed + 8ull * (unsigned long long)g + 2 * i), ub = srt_mix64(seed + 8ull * (unsigned long long)g + 2 * i + 1);
      const unsigned kind = (unsigned)(srt_mix64(ua ^ ub) & 63ull);
      if (kind < 40) {  // moderate exponents (what the ray tracer divides): 2^-300 .. 2^300
        ua = (ua & 0x800fffffffffffffull) | ((unsigned long long)(723 + (ua >> 52) % 600) << 52);
        ub = (ub & 0x800fffffffffffffull) | ((unsigned long long)(723 + (ub >> 52) % 600) << 52);
      } else if (kind == 40) ub = ua;                                   // equal operands
      else if (kind == 41) ua &= 0xfff0000000000000ull;                 // power of two numerator
      else if (kind == 42) ub |= 0x000fffffffffffffull;                 // divisor mantissa all ones
      else if (kind == 43) ua = 0ull;                                   // zero numerator
      else if (kind == 44) ub = (ub & 0x8000000000000000ull);           // zero divisor
      else if (kind == 45) ua |= 0x000fffffffffffffull;
      // else: raw random bit patterns (any exponent, NaN, Inf, subnormal)
      a[i] = __longlong_as_double((long long)ua);
      b[i] = __longlong_as_double((long long)ub);
    }
    srt_divn(a, b, q);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const double r = a[i] / b[i];
      const bool same = (__double_as_longlong(r) == __double_as_longlong(q[i])) || (r != r && q[i] != q[i]);
      bad += same ? 0ull : 1ull;
    }
  }
  if (bad) atomicAdd(mismatches, bad);
}

}  // namespace

// device copies of the grid arrays of a model (modelnum 3) on one device
struct GridReplica {
  double* d_nodes = nullptr;
  double* d_axes = nullptr;
  double* d_coef = nullptr;  // per-cell tricubic coefficients (NULL: on-the-fly path, SRT_GRID_COEF=0)
  // modelnum 4: the sample records sorted by cell, and the first record of every cell
  double* d_scat_pts = nullptr;
  int* d_scat_cells = nullptr;
};

struct srt_model {
  SrtModel M;  // grid pointers are filled per device at launch time
  uint64_t serial = 0;
  double nus[SRT_NSPEC];
  // host copy of the grid file: f only (derivative slots zero) unless the file carried derivatives
  std::vector<double> h_nodes, h_axes;
  srt_host::HostScat h_scat;  // modelnum 4: host copy of the sample records and the cell list
  bool file_has_derivatives = false;
  bool use_coef = true;
  std::mutex replica_mutex;
  std::map<int, GridReplica> replicas;  // by device
  std::map<int, int> blocks_per_sm;     // thread-per-ray kernel, by device
};

namespace {

// 0 = one thread per ray, 1 = team of 8 lanes per ray.  SRT_KERNEL=thread|team overrides.
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
int launch_config(DevCtx* c, srt_model* m, long long nrays, int* mode_out, int* block_out, int* grid_out, int* blocks_per_sm_out) {
  const int mode = pick_mode(m, nrays);
  int block = SRT_BLOCK_THREADS, nb = 1;
  if (mode) {
    block = srt_team_block_threads(nrays, c->num_sms);
    if ((m->M.ts.on || m->M.density == SRT_DENS_SCAT) && block > 512) block = 512;  // the kernels with the external field, and those of the scattered-sample model, exist in the 128-register budget only
    CK(srt_team_occupancy(m->M, block, &nb));
  } else {
    std::lock_guard<std::mutex> lk(m->replica_mutex);
    auto it = m->blocks_per_sm.find(c->dev);
    if (it == m->blocks_per_sm.end()) {
      CK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, pick_trace(m->M), SRT_BLOCK_THREADS, 0));
      m->blocks_per_sm[c->dev] = nb;
    } else {
      nb = it->second;
    }
  }
  if (nb < 1) nb = 1;
  const long long lanes_per_ray = mode ? 8 : 1;
  long long blocks_needed = (nrays * lanes_per_ray + block - 1) / block;
  long long resident = (long long)nb * c->num_sms;
  int grid = (int)(blocks_needed < resident ? blocks_needed : resident);
  if (grid < 1) grid = 1;
  *mode_out = mode; *block_out = block; *grid_out = grid;
  if (blocks_per_sm_out) *blocks_per_sm_out = nb;
  return SRT_OK;
}

// multiplier of the queue permutation (srt_queue_ray): the integer nearest nrays/golden ratio that is coprime to nrays
unsigned long long ray_order_mul(long long nrays) {
  if (nrays < 4) return 1ull;
  unsigned long long n = (unsigned long long)nrays, mul = (unsigned long long)((double)nrays * 0.6180339887498949);
  for (;; ++mul) {
    unsigned long long a = n, b = mul % n;
    while (b) { unsigned long long t = a % b; a = b; b = t; }
    if (a == 1ull && mul % n != 0ull) return mul % n;
  }
}

// The grid arrays of `m` on device c->dev (current device of the caller): upload of the node records, the
// finite-difference derivative passes and the per-cell coefficient table, once per device.
int ensure_replica(DevCtx* c, srt_model* m, GridReplica* out) {
  *out = GridReplica();
  if (m->M.density != SRT_DENS_GRID && m->M.density != SRT_DENS_SCAT) return SRT_OK;
  std::lock_guard<std::mutex> lk(m->replica_mutex);
  auto it = m->replicas.find(c->dev);
  if (it != m->replicas.end()) { *out = it->second; return SRT_OK; }
  GridReplica r;
  if (m->M.density == SRT_DENS_SCAT) {
    const size_t pb = m->h_scat.pts.size() * sizeof(double), cb = m->h_scat.cell_start.size() * sizeof(int);
    if (cudaMalloc(&r.d_scat_pts, pb) != cudaSuccess || cudaMalloc(&r.d_scat_cells, cb) != cudaSuccess) {
      int rc = cuda_fail(cudaGetLastError(), "cudaMalloc(scattered samples)");
      if (r.d_scat_pts) cudaFree(r.d_scat_pts);
      return rc;
    }
    cudaError_t e = cudaMemcpy(r.d_scat_pts, m->h_scat.pts.data(), pb, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(r.d_scat_cells, m->h_scat.cell_start.data(), cb, cudaMemcpyHostToDevice);
    if (e != cudaSuccess) {
      cudaFree(r.d_scat_pts); cudaFree(r.d_scat_cells);
      return cuda_fail(e, "upload of the scattered samples");
    }
    m->replicas[c->dev] = r;
    *out = r;
    return SRT_OK;
  }
  const SrtGrid& g = m->M.grid;
  const size_t nb = m->h_nodes.size() * sizeof(double), ax = m->h_axes.size() * sizeof(double);
  const size_t ncells = (m->use_coef && g.nx > 1 && g.ny > 1 && g.nz > 1) ? srt_grid_cell_count(g) : 0;
  if (cudaMalloc(&r.d_nodes, nb) != cudaSuccess || cudaMalloc(&r.d_axes, ax) != cudaSuccess ||
      (ncells && cudaMalloc(&r.d_coef, ncells * 256 * sizeof(double)) != cudaSuccess)) {
    int rc = cuda_fail(cudaGetLastError(), "cudaMalloc(grid)");
    if (r.d_nodes) cudaFree(r.d_nodes);
    if (r.d_axes) cudaFree(r.d_axes);
    return rc;
  }
  cudaMemcpy(r.d_nodes, m->h_nodes.data(), nb, cudaMemcpyHostToDevice);
  cudaMemcpy(r.d_axes, m->h_axes.data(), ax, cudaMemcpyHostToDevice);
  const long long threads = (long long)g.nx * g.ny * g.nz * 4;
  const int blocks = (int)((threads + 255) / 256 < (long long)c->num_sms * 16 ? (threads + 255) / 256 : (long long)c->num_sms * 16);
  if (!m->file_has_derivatives) {
    // libtricubic.f95:738-790: df/dx, df/dy, df/dz, then the mixed ones from the first derivatives
    const bool X = g.nx > 2, Y = g.ny > 2, Z = g.nz > 2;
    const struct { int src, dst, axis; double h; bool on; } pass[7] = {
        {0, 1, 0, g.delx, X}, {0, 2, 1, g.dely, Y}, {0, 3, 2, g.delz, Z}, {2, 4, 0, g.delx, X && Y},
        {3, 5, 0, g.delx, X && Z}, {3, 6, 1, g.dely, Y && Z}, {6, 7, 0, g.delx, X && Y && Z}};
    for (int q = 0; q < 7; ++q) {
      srt_fd_kernel<<<blocks, 256>>>(r.d_nodes, g.nx, g.ny, g.nz, g.nspec, pass[q].src, pass[q].dst, pass[q].axis, pass[q].h, pass[q].on ? 1 : 0);
      g_launches++;
    }
  }
  SrtGrid gd = g;
  gd.nodes = r.d_nodes;
  gd.coef = nullptr;
  gd.xs = r.d_axes;
  gd.ys = r.d_axes + g.nx;
  gd.zs = r.d_axes + g.nx + g.ny;
  if (ncells) {
    long long ct = (long long)ncells * 4;
    int cb = (int)((ct + 127) / 128 < (long long)c->num_sms * 16 ? (ct + 127) / 128 : (long long)c->num_sms * 16);
    srt_coef_kernel<<<cb, 128>>>(gd, r.d_coef);
    g_launches++;
  }
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    cudaFree(r.d_nodes); cudaFree(r.d_axes);
    if (r.d_coef) cudaFree(r.d_coef);
    return cuda_fail(e, "grid set-up");
  }
  m->replicas[c->dev] = r;
  *out = r;
  return SRT_OK;
}

// SrtModel as device c->dev sees it
int device_model(DevCtx* c, srt_model* m, SrtModel* out) {
  *out = m->M;
  if (m->M.density == SRT_DENS_GRID) {
    GridReplica r;
    int rc = ensure_replica(c, m, &r);
    if (rc) return rc;
    out->grid.nodes = r.d_nodes;
    out->grid.coef = r.d_coef;
    out->grid.xs = r.d_axes;
    out->grid.ys = r.d_axes + m->M.grid.nx;
    out->grid.zs = r.d_axes + m->M.grid.nx + m->M.grid.ny;
  } else if (m->M.density == SRT_DENS_SCAT) {
    GridReplica r;
    int rc = ensure_replica(c, m, &r);
    if (rc) return rc;
    out->scat.pts = r.d_scat_pts;
    out->scat.cell_start = r.d_scat_cells;
  }
  return SRT_OK;
}

// Makes `Md` (serial `serial`) the model of translation unit `tu` on device c->dev for a launch on `stream`.
// Caller holds c->sym_mutex until it has called symbol_reader() for its launch.
int symbol_prepare(DevCtx* c, int tu, uint64_t serial, const SrtModel& Md, cudaStream_t stream) {
  if (c->resident[tu] == serial) return SRT_OK;
  for (cudaEvent_t ev : c->readers[tu]) {
    CK(cudaStreamWaitEvent(stream, ev, 0));
    cudaEventDestroy(ev);  // released once it has completed; the wait above stays valid
  }
  c->readers[tu].clear();
  if (tu == 0) CK(cudaMemcpyToSymbolAsync(c_srt_model, &Md, sizeof(SrtModel), 0, cudaMemcpyHostToDevice, stream));
  else CK(srt_team_upload_model(Md, stream));
  c->resident[tu] = serial;
  // the upload itself must be over before a launch on ANOTHER stream with the same model may read the symbol
  cudaEvent_t up;
  CK(cudaEventCreateWithFlags(&up, cudaEventDisableTiming));
  CK(cudaEventRecord(up, stream));
  c->readers[tu].push_back(up);
  return SRT_OK;
}
int symbol_reader(DevCtx* c, int tu, cudaStream_t stream) {
  // a launch on `stream` reads the symbol: later uploads wait for it; launches on other streams wait for the upload
  if (!c->readers[tu].empty()) CK(cudaStreamWaitEvent(stream, c->readers[tu].front(), 0));
  if (c->readers[tu].size() > 64) {  // long runs with one model: drop the events that have completed
    std::vector<cudaEvent_t> keep;
    keep.push_back(c->readers[tu].front());
    for (size_t i = 1; i < c->readers[tu].size(); ++i) {
      if (cudaEventQuery(c->readers[tu][i]) == cudaSuccess) cudaEventDestroy(c->readers[tu][i]);
      else keep.push_back(c->readers[tu][i]);
    }
    c->readers[tu].swap(keep);
  }
  return SRT_OK;
}
int symbol_launched(DevCtx* c, int tu, cudaStream_t stream) {
  cudaEvent_t ev;
  CK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  CK(cudaEventRecord(ev, stream));
  c->readers[tu].push_back(ev);
  return SRT_OK;
}

// upload the model (if it is not the resident one) and launch the persistent trace kernel on `stream`
int launch_trace(DevCtx* c, srt_model* m, const SrtRunParams& P, const SrtBatch& B0, cudaStream_t stream) {
  SrtBatch B = B0;
  B.order_mul = ray_order_mul(B.nrays);
  {
    const char* env = getenv("SRT_TEAM_SHADOW");  // A/B hook: 0 = no shadow teams
    B.shadow = !env ? 1 : (env[0] == '0' ? 0 : (env[0] == '4' ? 4 : 1));  // 4: a primary and three shadows where the live rays fit four times
  }
  int mode, block, grid;
  int rc = launch_config(c, m, B.nrays, &mode, &block, &grid, nullptr);
  if (rc) return rc;
  SrtModel Md;
  if ((rc = device_model(c, m, &Md))) return rc;
  CK(cudaMemsetAsync(B.counters, 0, 8 * sizeof(unsigned long long), stream));
  std::lock_guard<std::mutex> lk(c->sym_mutex);
  const int tu = mode ? 1 : 0;
  if ((rc = symbol_prepare(c, tu, m->serial, Md, stream))) return rc;
  if ((rc = symbol_reader(c, tu, stream))) return rc;
  if (mode) {
    CK(srt_team_launch(Md, P, B, block, grid, stream));
  } else {
    pick_trace(Md)<<<grid, block, 0, stream>>>(P, B);
    CK(cudaGetLastError());
  }
  g_launches++;
  return symbol_launched(c, tu, stream);
}

int finish_model(srt_model* mm, int yearday, int msec, int use_igrf, int use_tsyganenko) {
  if (use_tsyganenko != 0 && use_tsyganenko != 1) return SRT_E_ARG;
  if (use_igrf != 0 && use_igrf != 1) return SRT_E_ARG;
  srt_host::set_time(mm->M, yearday, msec);
  mm->M.use_igrf = use_igrf;
  // use_tsyganenko = 1: T04_s with the parameters of srt_model_tsyganenko() (all zero until then, like the
  // reference's state before its --tsyganenko_* options are read) and the `PSI` the adapters pass (Q25)
  memset(&mm->M.ts, 0, sizeof mm->M.ts);
  mm->M.ts.on = use_tsyganenko;
  srt_host::igrf_recalc(yearday, msec, mm->M.igrf, &mm->M.ts.ps);
  for (int i = 0; i < SRT_NSPEC; ++i) mm->nus[i] = 0.0;
  mm->serial = ++g_model_serial;
  return SRT_OK;
}

}  // namespace

extern "C" {

int srt_init_devices(const int* devices, int ndevices) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n == 0) {
    snprintf(g_cuda_err, sizeof g_cuda_err, "no CUDA device: %s", cudaGetErrorString(e));
    return SRT_E_NODEVICE;
  }
  std::vector<int> want;
  if (ndevices <= 0 || !devices) {
    for (int d = 0; d < n; ++d) want.push_back(d);
  } else {
    for (int i = 0; i < ndevices; ++i) {
      if (devices[i] < 0 || devices[i] >= n) {
        snprintf(g_cuda_err, sizeof g_cuda_err, "device %d of %d", devices[i], n);
        return SRT_E_NODEVICE;
      }
      bool dup = false;
      for (int d : want) dup = dup || d == devices[i];
      if (!dup) want.push_back(devices[i]);
    }
  }
  std::vector<DevCtx*> act;
  for (int d : want) {
    DevCtx* c = ctx_for_device(d);
    if (!c) return cuda_fail(cudaGetLastError(), "cudaGetDeviceProperties");
    act.push_back(c);
  }
  CK(cudaSetDevice(want[0]));
  std::lock_guard<std::mutex> lk(g_dev_mutex);
  g_active = act;
  return SRT_OK;
}

int srt_init(int device) {
  if (device >= 0) return srt_init_devices(&device, 1);
  // keep the calling thread's current device
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_dev_mutex);
  g_active.assign(1, c);
  return SRT_OK;
}

int srt_device_count(void) {
  std::lock_guard<std::mutex> lk(g_dev_mutex);
  return (int)g_active.size();
}

const char* srt_strerror(int code) {
  switch (code) {
    case SRT_OK: return "ok";
    case SRT_E_NODEVICE: return "no usable CUDA device (libsrt has no CPU path)";
    case SRT_E_IO: return "cannot open or parse model file";
    case SRT_E_ARG: return "invalid argument";
    case SRT_E_UNSUPPORTED: return "unsupported option or state";
    case SRT_E_CUDA: return "CUDA runtime error";
    case SRT_E_NOMEM: return "trajectory log too small for a single ray";
    case SRT_E_TOOBIG: return "more points than the static scratch of srt_model_eval_noalloc holds (16384)";
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
  std::vector<DevCtx*> devs;
  int rc = active_devices(devs);
  if (rc) return rc;
  srt_model* mm = new (std::nothrow) srt_model();
  if (!mm) return SRT_E_NOMEM;
  memset(&mm->M, 0, sizeof mm->M);
  srt_host::HostGrid hg;
  int compder = 0;
  rc = srt_host::load_grid(path, mm->M, hg, /*host_derivatives=*/false, &compder);
  if (rc != 0) { delete mm; return rc == -3 ? SRT_E_UNSUPPORTED : SRT_E_IO; }
  rc = finish_model(mm, yearday, msec, use_igrf, use_tsyganenko);
  if (rc) { delete mm; return rc; }
  mm->file_has_derivatives = (compder == 1);
  mm->h_nodes.swap(hg.nodes);
  mm->h_axes = hg.xs;
  mm->h_axes.insert(mm->h_axes.end(), hg.ys.begin(), hg.ys.end());
  mm->h_axes.insert(mm->h_axes.end(), hg.zs.begin(), hg.zs.end());
  {  // per-cell tricubic coefficients unless SRT_GRID_COEF=0 (test hook: the on-the-fly path)
    const char* env = getenv("SRT_GRID_COEF");
    mm->use_coef = !(env && !strcmp(env, "0"));
  }
  // replicas on every active device now, so that set-up is not part of the first batch
  int prev = 0;
  cudaGetDevice(&prev);
  for (DevCtx* c : devs) {
    GridReplica r;
    if (cudaSetDevice(c->dev) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "cudaSetDevice"); break; }
    if ((rc = ensure_replica(c, mm, &r))) break;
  }
  cudaSetDevice(prev);
  if (rc) { srt_model_free(mm); return rc; }
  *m = mm;
  return SRT_OK;
}

int srt_model_scattered(srt_model** m, const char* path, int yearday, int msec, int use_igrf, int use_tsyganenko, double window_scale, int order,
                        int exact, int scaled, double local_window_scale) {
  if (!m || !path) return SRT_E_ARG;
  *m = nullptr;
  std::vector<DevCtx*> devs;
  int rc = active_devices(devs);
  if (rc) return rc;
  srt_model* mm = new (std::nothrow) srt_model();
  if (!mm) return SRT_E_NOMEM;
  memset(&mm->M, 0, sizeof mm->M);
  const srt_host::ScatOptions opt = {window_scale, order, exact, scaled, local_window_scale};
  rc = srt_host::load_scattered(path, opt, mm->M, mm->h_scat);
  if (rc != 0) { delete mm; return rc == -3 ? SRT_E_UNSUPPORTED : SRT_E_IO; }
  rc = finish_model(mm, yearday, msec, use_igrf, use_tsyganenko);
  if (rc) { delete mm; return rc; }
  mm->M.scat.pts = nullptr;  // filled per device at launch time
  mm->M.scat.cell_start = nullptr;
  int prev = 0;
  cudaGetDevice(&prev);
  for (DevCtx* c : devs) {
    GridReplica r;
    if (cudaSetDevice(c->dev) != cudaSuccess) { rc = cuda_fail(cudaGetLastError(), "cudaSetDevice"); break; }
    if ((rc = ensure_replica(c, mm, &r))) break;
  }
  cudaSetDevice(prev);
  if (rc) { srt_model_free(mm); return rc; }
  *m = mm;
  return SRT_OK;
}

int srt_model_scattered_info(const srt_model* m, int64_t* npoints, double* maxnearest, double* radius, int* nmono, int* cells /*[3]*/) {
  if (!m || m->M.density != SRT_DENS_SCAT) return SRT_E_ARG;
  const SrtScat& s = m->M.scat;
  if (npoints) *npoints = s.npts;
  if (maxnearest) *maxnearest = s.maxnearest;
  if (radius) *radius = s.radius;
  if (nmono) *nmono = s.nmono;
  if (cells) { cells[0] = s.ncx; cells[1] = s.ncy; cells[2] = s.ncz; }
  return SRT_OK;
}

int64_t srt_model_scattered_samples(const srt_model* m, double* recs, int64_t capacity) {
  if (!m || m->M.density != SRT_DENS_SCAT) return SRT_E_ARG;
  const int64_t n = (int64_t)m->h_scat.pts.size();
  if (recs) {
    if (capacity < n) return SRT_E_ARG;
    memcpy(recs, m->h_scat.pts.data(), sizeof(double) * (size_t)n);
  }
  return n / SRT_SCAT_REC;
}

void srt_model_free(srt_model* m) {
  if (!m) return;
  int prev = 0;
  cudaGetDevice(&prev);
  for (auto& kv : m->replicas) {
    cudaSetDevice(kv.first);
    // a kernel still reading this model's arrays must finish first
    cudaDeviceSynchronize();
    if (kv.second.d_nodes) cudaFree(kv.second.d_nodes);
    if (kv.second.d_axes) cudaFree(kv.second.d_axes);
    if (kv.second.d_coef) cudaFree(kv.second.d_coef);
    if (kv.second.d_scat_pts) cudaFree(kv.second.d_scat_pts);
    if (kv.second.d_scat_cells) cudaFree(kv.second.d_scat_cells);
  }
  if (!m->replicas.empty()) cudaSetDevice(prev);
  delete m;
}

int srt_model_tsyganenko(srt_model* m, const double* parmod) {
  if (!m || !parmod) return SRT_E_ARG;
  for (int i = 0; i < 10; ++i) m->M.ts.parmod[i] = (float)parmod[i];  // real(parmod), adapter :184
  m->M.ts.on = 1;
  m->serial = ++g_model_serial;  // the copies of the model on the devices are stale
  return SRT_OK;
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
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  if (n == 0) return SRT_OK;
  srt_model* mm = const_cast<srt_model*>(m);
  SrtModel Md;
  if ((rc = device_model(c, mm, &Md))) return rc;
  double *dx = nullptr, *dout = nullptr;
  CK(cudaMalloc(&dx, n * 3 * sizeof(double)));
  if (cudaMalloc(&dout, n * 19 * sizeof(double)) != cudaSuccess) { cudaFree(dx); return cuda_fail(cudaGetLastError(), "cudaMalloc"); }
  cudaStream_t st = nullptr;  // legacy default stream: ordered after the copies below
  cudaMemcpy(dx, x, n * 3 * sizeof(double), cudaMemcpyHostToDevice);
  {
    std::lock_guard<std::mutex> lk(c->sym_mutex);
    rc = symbol_prepare(c, 0, mm->serial, Md, st);
    if (!rc) rc = symbol_reader(c, 0, st);
    if (!rc) {
      int blocks = (int)((n + 127) / 128);
      if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
      pick_eval(Md)<<<blocks, 128, 0, st>>>(n, dx, dout);
      g_launches++;
      rc = symbol_launched(c, 0, st);
    }
  }
  if (!rc) {
    cudaError_t e = cudaMemcpy(out, dout, n * 19 * sizeof(double), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) rc = cuda_fail(e, "srt_model_eval");
  }
  cudaFree(dx);
  cudaFree(dout);
  return rc;
}

int64_t srt_model_eval_noalloc_max_points(void) { return SRT_EVAL_STATIC_MAX_POINTS; }

// srt_model_eval without any allocation: points and results are staged in g_eval_static (shared by every call on the
// device; two calls that overlap on one GPU race, so the whole call runs under the device's batch mutex and is
// synchronous).  Nothing in the array has to start at zero: the kernel writes all 19 words of each of the n points
// before they are copied out, and words beyond n are never read.
int srt_model_eval_noalloc(const srt_model* m, int64_t n, const double* x, double* out) {
  if (!m || n < 0 || (n > 0 && (!x || !out))) return SRT_E_ARG;
  if (n > SRT_EVAL_STATIC_MAX_POINTS) {
    snprintf(g_cuda_err, sizeof g_cuda_err, "srt_model_eval_noalloc: %lld points exceed the static scratch of %d points; split the call or use srt_model_eval",
             (long long)n, SRT_EVAL_STATIC_MAX_POINTS);
    return SRT_E_TOOBIG;
  }
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  if (n == 0) return SRT_OK;
  srt_model* mm = const_cast<srt_model*>(m);
  SrtModel Md = mm->M;
  if (mm->M.density == SRT_DENS_GRID) {
    // the grid arrays of a model are placed on a device by its first ordinary call there (ensure_replica allocates)
    std::lock_guard<std::mutex> lk(mm->replica_mutex);
    auto it = mm->replicas.find(c->dev);
    if (it == mm->replicas.end()) {
      snprintf(g_cuda_err, sizeof g_cuda_err, "srt_model_eval_noalloc: the gridded model is not yet on device %d; one srt_model_eval or trace call places it", c->dev);
      return SRT_E_UNSUPPORTED;
    }
    Md.grid.nodes = it->second.d_nodes;
    Md.grid.coef = it->second.d_coef;
    Md.grid.xs = it->second.d_axes;
    Md.grid.ys = it->second.d_axes + mm->M.grid.nx;
    Md.grid.zs = it->second.d_axes + mm->M.grid.nx + mm->M.grid.ny;
  } else if (mm->M.density == SRT_DENS_SCAT) {
    std::lock_guard<std::mutex> lk(mm->replica_mutex);
    auto it = mm->replicas.find(c->dev);
    if (it == mm->replicas.end()) {
      snprintf(g_cuda_err, sizeof g_cuda_err, "srt_model_eval_noalloc: the scattered-sample model is not yet on device %d; one srt_model_eval or trace call places it", c->dev);
      return SRT_E_UNSUPPORTED;
    }
    Md.scat.pts = it->second.d_scat_pts;
    Md.scat.cell_start = it->second.d_scat_cells;
  }
  std::lock_guard<std::mutex> bl(c->batch_mutex);
  if (!c->eval_static) {  // once per device, with that device current (current_ctx)
    double* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void**)&p, g_eval_static);
    if (e != cudaSuccess || !p) return cuda_fail(e, "cudaGetSymbolAddress(g_eval_static)");
    c->eval_static = p;
  }
  double* dx = c->eval_static;
  double* dout = c->eval_static + (size_t)SRT_EVAL_STATIC_MAX_POINTS * 3;
  cudaStream_t st = nullptr;  // legacy default stream: copies and kernel are ordered
  CK(cudaMemcpy(dx, x, n * 3 * sizeof(double), cudaMemcpyHostToDevice));
  std::lock_guard<std::mutex> lk(c->sym_mutex);
  if (c->resident[0] != mm->serial) {
    // as symbol_prepare, without creating an event: the upload is synchronous and the kernel has finished (copy below)
    // before the mutex is released, so nobody needs an event to wait for either
    for (cudaEvent_t ev : c->readers[0]) {
      CK(cudaStreamWaitEvent(st, ev, 0));
      cudaEventDestroy(ev);
    }
    c->readers[0].clear();
    CK(cudaMemcpyToSymbol(c_srt_model, &Md, sizeof(SrtModel), 0, cudaMemcpyHostToDevice));
    c->resident[0] = mm->serial;
  } else if ((rc = symbol_reader(c, 0, st))) {
    return rc;
  }
  int blocks = (int)((n + 127) / 128);
  if (blocks > c->num_sms * 8) blocks = c->num_sms * 8;
  pick_eval(Md)<<<blocks, 128, 0, st>>>(n, dx, dout);
  CK(cudaGetLastError());
  g_launches++;
  cudaError_t e = cudaMemcpy(out, dout, n * 19 * sizeof(double), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) return cuda_fail(e, "srt_model_eval_noalloc");
  return SRT_OK;
}

int srt_selftest_div(int64_t ngroups, uint64_t seed, uint64_t* mismatches) {
  if (!mismatches || ngroups < 0) return SRT_E_ARG;
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  unsigned long long* d = nullptr;
  CK(cudaMalloc(&d, 8));
  cudaMemset(d, 0, 8);
  srt_divtest_kernel<<<c->num_sms * 8, 256>>>(ngroups, seed, d);
  g_launches++;
  unsigned long long h = 0;
  cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) return cuda_fail(e, "srt_selftest_div");
  *mismatches = h;
  return SRT_OK;
}

int srt_kernel_info(const srt_model* m, int64_t nrays, int* mode, int* regs, int* local_bytes, int* blocks_per_sm, int* grid, int* block) {
  if (!m) return SRT_E_ARG;
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
  if (rc) return rc;
  srt_model* mm = const_cast<srt_model*>(m);
  int md, blk, grd, nb;
  rc = launch_config(c, mm, nrays, &md, &blk, &grd, &nb);
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
  DevCtx* c = nullptr;
  int rc = current_ctx(&c);
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
  B.shadow = 1;     // set by launch_trace
  return launch_trace(c, const_cast<srt_model*>(m), P, B, (cudaStream_t)stream);
}

}  // extern "C"

#include "srt_api_host.inl"
