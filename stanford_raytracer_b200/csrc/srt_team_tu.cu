// srt_team_tu.cu -- translation unit of the team-of-8 trace kernel.
//
// The two forms of the integrator want different code generation (measured, profiles/r1_variants.txt):
//   one thread per ray : 256-thread blocks stepping block-synchronously, elementary functions inlined
//   team of 8 per ray  : up to 1024-thread blocks stepping their rounds block-synchronously (2.1x over
//                        unsynchronised 64-thread blocks); the launch picks a smaller block when the batch is
//                        too small to give every SM one
// so each lives in its own translation unit with its own copy of the model in __constant__ memory.
#define SRT_TEAM_TU 1
#ifndef SRT_TEAM_BLOCK_SYNC
#define SRT_TEAM_BLOCK_SYNC 1
#endif
#define SRT_BLOCK_THREADS 512
#define SRT_THREAD_MIN_BLOCKS 1
#include "srt_kernels.cuh"
#include "srt_team.cuh"
#include "srt_team_tu.h"

#define SRT_TEAM_MAXT_SMALL 512  // 255-register (block 256) and 168-register (block 320) builds measured slower on config 2 (profiles/r1i_register_budget.txt)

typedef void (*team_kernel_t)(const SrtRunParams, const SrtBatch);

// Two register budgets of the same kernel: MAXT = 512 (128 registers; best when rays are scarce and the latency of
// one ray matters, e.g. config 2 and every tail) and MAXT = 1024 (64 registers, 32 warps per SM; +8 % when the
// batch fills the machine).  profiles/r1c_variants.txt.
template <int MAXT>
static team_kernel_t pick_team_t(const SrtModel& M) {
  if (M.density == SRT_DENS_NGO) return M.use_igrf ? srt_trace_team_kernel<SRT_DENS_NGO, 1, MAXT> : srt_trace_team_kernel<SRT_DENS_NGO, 0, MAXT>;
  return M.use_igrf ? srt_trace_team_kernel<SRT_DENS_GRID, 1, MAXT> : srt_trace_team_kernel<SRT_DENS_GRID, 0, MAXT>;
}
// (a third budget, 768 threads at 80 registers, measured 5-7 % slower than 1024 at 64 on the 40 000-ray, 100 000-ray and config-3
// cases: profiles/r2h_damping_ab768_traffic.txt)
// use_tsyganenko = 1 (field code bit 1): the TS05 external field is ~100 x the work of the rest of a plasma evaluation and wants
// registers, so these kernels exist in the 128-register budget only (launch_config keeps their blocks at <= 512 threads)
static team_kernel_t pick_team_ts(const SrtModel& M) {
  if (M.density == SRT_DENS_NGO) return M.use_igrf ? srt_trace_team_kernel<SRT_DENS_NGO, 3, SRT_TEAM_MAXT_SMALL> : srt_trace_team_kernel<SRT_DENS_NGO, 2, SRT_TEAM_MAXT_SMALL>;
  return M.use_igrf ? srt_trace_team_kernel<SRT_DENS_GRID, 3, SRT_TEAM_MAXT_SMALL> : srt_trace_team_kernel<SRT_DENS_GRID, 2, SRT_TEAM_MAXT_SMALL>;
}
// modelnum 4 (scattered samples): a plasma evaluation is a moving-least-squares fit over the neighbours of the point, far
// heavier than the rest of the attempt, and keeps a packed normal matrix per lane: the 128-register budget only
static team_kernel_t pick_team_scat(const SrtModel& M) {
  static const team_kernel_t k[4] = {srt_trace_team_kernel<SRT_DENS_SCAT, 0, SRT_TEAM_MAXT_SMALL>, srt_trace_team_kernel<SRT_DENS_SCAT, 1, SRT_TEAM_MAXT_SMALL>,
                                     srt_trace_team_kernel<SRT_DENS_SCAT, 2, SRT_TEAM_MAXT_SMALL>, srt_trace_team_kernel<SRT_DENS_SCAT, 3, SRT_TEAM_MAXT_SMALL>};
  return k[(M.use_igrf ? 1 : 0) | (M.ts.on ? 2 : 0)];
}
static team_kernel_t pick_team(const SrtModel& M, int block) {
  if (M.density == SRT_DENS_SCAT) return pick_team_scat(M);
  if (M.ts.on) return pick_team_ts(M);
  return block > SRT_TEAM_MAXT_SMALL ? pick_team_t<1024>(M) : pick_team_t<SRT_TEAM_MAXT_SMALL>(M);
}

// block size for a batch: the largest power of two <= 1024 that still gives every SM at least one block
int srt_team_block_threads(long long nrays, int num_sms) {
  long long lanes_per_sm = (nrays * 8) / (num_sms > 0 ? num_sms : 1);
  int block = 128;
  while (block * 2 <= 1024 && block * 2 <= lanes_per_sm) block *= 2;
  const char* env = getenv("SRT_TEAM_BLOCK");  // test / tuning hook
  if (env && atoi(env) >= 32 && atoi(env) <= 1024) block = atoi(env) / 32 * 32;
  return block;
}

static size_t team_smem_bytes(int block) { return (size_t)SRT_PACK_WORDS * 8 * (size_t)(block / 2); }

cudaError_t srt_team_occupancy(const SrtModel& M, int block, int* blocks_per_sm) {
  cudaError_t e = cudaFuncSetAttribute(pick_team(M, block), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)team_smem_bytes(1024));
  if (e != cudaSuccess) return e;
  return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, pick_team(M, block), block, team_smem_bytes(block));
}

cudaError_t srt_team_attributes(const SrtModel& M, int block, cudaFuncAttributes* fa) { return cudaFuncGetAttributes(fa, pick_team(M, block)); }

// the model of the next launches of this translation unit's kernels (its own __constant__ copy)
cudaError_t srt_team_upload_model(const SrtModel& M, cudaStream_t stream) {
  return cudaMemcpyToSymbolAsync(c_srt_model, &M, sizeof(SrtModel), 0, cudaMemcpyHostToDevice, stream);
}

cudaError_t srt_team_launch(const SrtModel& M, const SrtRunParams& P, const SrtBatch& B, int block, int grid, cudaStream_t stream) {
  cudaError_t e = cudaFuncSetAttribute(pick_team(M, block), cudaFuncAttributeMaxDynamicSharedMemorySize, (int)team_smem_bytes(1024));
  if (e != cudaSuccess) return e;
  pick_team(M, block)<<<grid, block, team_smem_bytes(block), stream>>>(P, B);
  return cudaGetLastError();
}
