// srt_kernels.cuh -- sm_100a kernels of the hot path.
//
//   srt_trace_kernel    persistent one-thread-per-ray integrator (raytracer_run,
//                       fortran/raytracer.f95:609-995) with a warp-aggregated ray queue:
//                       a lane whose ray terminated pulls the next ray index at the top
//                       of the loop, so a finished ray never idles its lane while the
//                       rest of the warp keeps stepping; a rejected attempt keeps its
//                       lane (it retries with the smaller dt, :790-798).
//   srt_scatter_kernel  orders the append log by (ray, row) for the host gather.
//   srt_eval_kernel     funcPlasmaParams at n points (the dumpmodel parity surface).
//   srt_fp64_peak_kernel  FP64 pipe micro-benchmark: the roofline denominator.
//
// Bound: FP64 (non-tensor) pipe; nothing here is a dense contraction, so no tensor
// cores, no TMA.  HBM traffic = 168 B per recorded row + 56 B per ray in, 192 B per ray
// out; the model lives in __constant__ memory (c_srt_model).
// Gridded model: + one 2 KiB coefficient record per distinct cell a query lands in (L2-resident gathers).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "srt_physics.h"
#include "srt_igrf.h"
#include "srt_tricubic.h"

#define SRT_LOG_WORDS 21      // header word + SRT_ROW_DOUBLES
#define SRT_SUMMARY_WORDS 24  // last row + nsteps, stopcond, rejected, flags
#define SRT_FLAG_CONE_EXIT 1
#define SRT_FLAG_DROPPED 2    // log full: the ray was abandoned and must be re-run

struct SrtBatch {
  long long nrays;
  double* rays;                     // SoA [7][nrays]
  double* log;                      // [capacity][SRT_LOG_WORDS]
  unsigned long long log_capacity;
  double* summary;                  // [nrays][SRT_SUMMARY_WORDS]
  unsigned long long* counters;     // [0] log rows, [1] accepted steps (incl. [5]), [2] attempts (incl. [5]), [3] dropped rays, [4] ray queue head,
                                    // [5] steps of all-NaN states taken by the fast-forward (srt_nan_fastforward), team kernel: [6] attempts that were a shadow team's (computed alongside the attempt they retry), [7] max over blocks of rounds
  unsigned* nlogged;                // optional [nrays]: rows this ray put in the log
  unsigned long long order_mul;     // queue position q is ray (q * order_mul) mod nrays; coprime to nrays (1 = input order)
  int shadow;                       // team kernel: free team slots run shadow attempts once the queue is empty (srt_team.cuh): 0 none, 1 pairs, 4 a primary
                                    // and three shadows where the live rays fit four times; set by launch_trace
};

// Queue position -> ray index.  A batch of the size of one machine fill (config 2) is handed out once: in input order
// the long-lived rays (neighbours in frequency) would all land in the same few blocks and the rest of the SMs would
// idle through the tail.  A golden-ratio stride spreads every run of consecutive queue positions evenly over the
// input order.  Which thread integrates which ray has no effect on any result.
__device__ __forceinline__ long long srt_queue_ray(const SrtBatch& B, long long q) {
  return (long long)(((unsigned long long)q * B.order_mul) % (unsigned long long)B.nrays);
}

__device__ __forceinline__ void srt_log_row(const SrtBatch& B, long long ray, int rowidx, const double row[SRT_ROW_DOUBLES], bool& dropped) {
  if (!B.log) return;  // summary-only run: nothing is recorded, nothing is dropped
  unsigned long long slot = atomicAdd(&B.counters[0], 1ull);
  if (slot >= B.log_capacity) { dropped = true; return; }
  double* dst = B.log + slot * SRT_LOG_WORDS;
  unsigned long long hdr = (unsigned long long)(unsigned)ray | ((unsigned long long)(unsigned)rowidx << 32);
  dst[0] = __longlong_as_double((long long)hdr);
#pragma unroll
  for (int i = 0; i < SRT_ROW_DOUBLES; ++i) dst[1 + i] = row[i];
}

// NaN fast-forward.  A state whose six components are all NaN stays so, and the reference steps it on to maxsteps or
// tmax (tests/test_nan_rays_cpu.py pins this on the oracle; ~1 % of the config-3 rays, 5 % of its accepted steps):
// the error estimate is NaN, so `err > maxerr` and the growth test are false and dt stays; the cone test `imsum > 0`
// is false, so every attempt is accepted; no stop condition but `t >= tmax` and `nstep >= maxsteps` can fire (every
// norm is NaN; dt can only have grown since the `dt < 1e-14` test of this attempt passed).  Every further row is the row
// just recorded -- evaluated at this NaN position from this NaN wave vector -- with t advanced by dt, so the remaining
// steps are taken here as one addition each (same bits: t = t + dt sequentially) instead of a full RK45 step each.
// Called by the lane that writes the ray's rows, right after it recorded `row` for step nstep.  Returns the number
// of steps taken; they are booked in counters[5] as well as in [1]/[2], so that rates can be quoted without them.
__device__ __forceinline__ unsigned srt_nan_fastforward(const SrtRunParams& P, const SrtBatch& B, long long ray, int outputper, double row[SRT_ROW_DOUBLES],
                                                        double& t, double dt, int& nstep, unsigned& nlog, bool& dropped) {
  unsigned n = 0;
  while (!dropped && !(t >= P.tmax) && !(nstep >= P.maxsteps)) {
    t = t + dt;
    nstep = nstep + 1;
    ++n;
    row[0] = t;
    const int ri = nstep - 1;
    if (ri % outputper == 0) {
      srt_log_row(B, ray, ri / outputper, row, dropped);
      if (!dropped) ++nlog;
    }
  }
  return n;
}
__device__ __forceinline__ bool srt_all_nan6(const double x[3], const double k[3]) {
  return x[0] != x[0] && x[1] != x[1] && x[2] != x[2] && k[0] != k[0] && k[1] != k[1] && k[2] != k[2];
}

template <int DENS, int USE_IGRF>
__global__ void __launch_bounds__(SRT_BLOCK_THREADS, SRT_THREAD_MIN_BLOCKS) srt_trace_kernel(const __grid_constant__ SrtRunParams P, const __grid_constant__ SrtBatch B) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const SrtModelRef MR = SrtModelRef();
  SrtRayState st;
  double row[SRT_ROW_DOUBLES];
  bool active = false, exhausted = false, dropped = false;
  long long slot = -1, ray = -1;
  unsigned nlog = 0;
  unsigned long long my_accepted = 0, my_attempts = 0, my_ff = 0;
  const int outputper = P.outputper < 1 ? 1 : P.outputper;

  for (;;) {
    // ---- refill: every idle lane of the warp takes the next ray with ONE atomic per warp
    const unsigned need = __ballot_sync(full, !active && !exhausted);
    if (need) {
      const int leader = __ffs(need) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(&B.counters[4], (unsigned long long)__popc(need));
      base = __shfl_sync(full, base, leader);
      if (!active && !exhausted) {
        slot = (long long)(base + __popc(need & ((1u << lane) - 1u)));
        if (slot >= B.nrays) {
          exhausted = true;
        } else {
          slot = srt_queue_ray(B, slot);
          ray = slot;
          // coalesced SoA loads: neighbouring lanes hold neighbouring rays
          double pos0[3] = {B.rays[0 * B.nrays + slot], B.rays[1 * B.nrays + slot], B.rays[2 * B.nrays + slot]};
          double dir0[3] = {B.rays[3 * B.nrays + slot], B.rays[4 * B.nrays + slot], B.rays[5 * B.nrays + slot]};
          double w0 = B.rays[6 * B.nrays + slot];
          srt_ray_init<DENS, USE_IGRF>(MR, P, pos0, dir0, w0, st, row);
          B.rays[3 * B.nrays + slot] = dir0[0];
          B.rays[4 * B.nrays + slot] = dir0[1];
          B.rays[5 * B.nrays + slot] = dir0[2];
          active = true;
          dropped = false;
          nlog = 0;
          srt_log_row(B, ray, 0, row, dropped);
          if (!dropped) nlog = 1;
        }
      }
    }
#if defined(SRT_BLOCK_SYNC)
    // block-synchronous attempts: every warp of the block starts its next attempt together, so the warps
    // walk the (instruction-cache-sized) step code in phase and share the fetched lines
    if (__syncthreads_or(active ? 1 : 0) == 0) break;
#else
    if (__ballot_sync(full, active) == 0u) break;  // queue empty and every lane finished
#endif

    if (active) {
      int r = SRT_ATTEMPT_DONE;
      if (!dropped) {
        double newrow[SRT_ROW_DOUBLES];
        r = srt_ray_attempt<DENS, USE_IGRF>(MR, P, st, newrow);
        if (r != SRT_ATTEMPT_DONE) ++my_attempts;
        if (r == SRT_ATTEMPT_ACCEPTED) {
          ++my_accepted;
#pragma unroll
          for (int i = 0; i < SRT_ROW_DOUBLES; ++i) row[i] = newrow[i];
          const int rowidx = st.nstep - 1;  // 0-based index of the row just recorded
          if (rowidx % outputper == 0) {
            srt_log_row(B, ray, rowidx / outputper, row, dropped);
            if (!dropped) ++nlog;
          }
          if (srt_all_nan6(st.x, st.k)) {
            const unsigned nff = srt_nan_fastforward(P, B, ray, outputper, row, st.t, st.dt, st.nstep, nlog, dropped);
            my_accepted += nff; my_attempts += nff; my_ff += nff;
          }
        }
      }
      if (r == SRT_ATTEMPT_DONE) {
        double* s = B.summary + slot * SRT_SUMMARY_WORDS;
#pragma unroll
        for (int i = 0; i < SRT_ROW_DOUBLES; ++i) s[i] = row[i];
        s[20] = (double)st.nstep;
        s[21] = (double)st.stopcond;
        s[22] = (double)(st.nrej + st.ncone);
        s[23] = (double)((st.fixed_cone_exit ? SRT_FLAG_CONE_EXIT : 0) | (dropped ? SRT_FLAG_DROPPED : 0));
        if (B.nlogged) B.nlogged[slot] = dropped ? 0u : nlog;
        if (dropped) atomicAdd(&B.counters[3], 1ull);
        active = false;
      }
    }
  }
  // per-warp totals -> two atomics per warp
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    my_accepted += __shfl_down_sync(full, my_accepted, o);
    my_attempts += __shfl_down_sync(full, my_attempts, o);
    my_ff += __shfl_down_sync(full, my_ff, o);
  }
  if (lane == 0) {
    atomicAdd(&B.counters[1], my_accepted);
    atomicAdd(&B.counters[2], my_attempts);
    if (my_ff) atomicAdd(&B.counters[5], my_ff);
  }
}

#if !defined(SRT_TEAM_TU)
// log[slot] -> sorted[offsets[ray] + row]; entries of dropped rays are skipped
__global__ void srt_scatter_kernel(const double* __restrict__ log, unsigned long long nlog, const unsigned long long* __restrict__ offsets,
                                   const unsigned* __restrict__ nlogged, double* __restrict__ sorted) {
  // one warp per log entry: 21 words in, 20 words out, coalesced both ways
  const unsigned long long warp = (blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  const unsigned long long nwarps = ((unsigned long long)gridDim.x * blockDim.x) >> 5;
  for (unsigned long long e = warp; e < nlog; e += nwarps) {
    const double* src = log + e * SRT_LOG_WORDS;
    unsigned long long hdr = (unsigned long long)__double_as_longlong(src[0]);
    unsigned ray = (unsigned)(hdr & 0xffffffffull), rowidx = (unsigned)(hdr >> 32);
    long long slot = (long long)ray;
    if (nlogged[slot] == 0u) continue;
    double* dst = sorted + (offsets[slot] + rowidx) * SRT_ROW_DOUBLES;
    if (lane < SRT_ROW_DOUBLES) dst[lane] = src[1 + lane];
  }
}

template <int DENS, int USE_IGRF>
__global__ void srt_eval_kernel(long long n, const double* __restrict__ x, double* __restrict__ out) {
  const SrtModelRef MR = SrtModelRef();
  const SrtModel& M = MR.get();
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    double xx[3] = {x[3 * i], x[3 * i + 1], x[3 * i + 2]};
    SrtPlasma p;
    srt_plasma<DENS, USE_IGRF>(MR, xx, p);
    double* o = out + 19 * i;
#pragma unroll
    for (int s = 0; s < 4; ++s) {
      o[s] = M.qs[s];
      o[4 + s] = p.Ns[s];
      o[8 + s] = s < M.nspec ? M.ms[s] : 0.0;
      o[12 + s] = 0.0;
    }
#pragma unroll
    for (int c = 0; c < 3; ++c) o[16 + c] = p.B0[c];
  }
}

// One-shot set-up of the gridded model: tricubic coefficients of every cell of the extended lattice
// (tricubic_get_coeff, libtricubic.f95:638-656,715-720, with the corner clamps of :859-921 for edge and outside
// cells) -> coef[cell][64 terms][4 species].
// One thread per (cell, species); the 64 x 64 int8 Amat sits in __constant__ memory and every
// thread of a warp reads the same entry (broadcast).
__global__ void srt_coef_kernel(const __grid_constant__ SrtGrid g, double* __restrict__ coef) {
  const long long ncx = g.nx + 1, ncy = g.ny + 1;
  const long long total = (long long)srt_grid_cell_count(g) * 4;
  for (long long id = blockIdx.x * (long long)blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
    const int s = (int)(id & 3);
    const long long cell = id >> 2;
    double* dst = coef + cell * 256 + s;
    if (s >= g.nspec) {
      for (int t = 0; t < 64; ++t) dst[4 * t] = 0.0;
      continue;
    }
    const int is0 = (int)(cell % ncx), js0 = (int)((cell / ncx) % ncy), ks0 = (int)(cell / (ncx * ncy));
    double a[64];
    srt_tricubic_cell_coef_ext(g, is0, js0, ks0, s, a);
    for (int t = 0; t < 64; ++t) dst[4 * t] = a[t];
  }
}

// FP64 pipe peak: 8 independent accumulator chains per thread
template <int FMA>
__global__ void srt_fp64_peak_kernel(int iters, double seed, double* out) {
  double a0 = seed + threadIdx.x, a1 = a0 + 1, a2 = a0 + 2, a3 = a0 + 3, a4 = a0 + 4, a5 = a0 + 5, a6 = a0 + 6, a7 = a0 + 7;
  const double m = 0.999999999, c = 1e-9;
  for (int i = 0; i < iters; ++i) {
    if (FMA) {
      a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
      a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
    } else {
      a0 = __dadd_rn(__dmul_rn(a0, m), c); a1 = __dadd_rn(__dmul_rn(a1, m), c);
      a2 = __dadd_rn(__dmul_rn(a2, m), c); a3 = __dadd_rn(__dmul_rn(a3, m), c);
      a4 = __dadd_rn(__dmul_rn(a4, m), c); a5 = __dadd_rn(__dmul_rn(a5, m), c);
      a6 = __dadd_rn(__dmul_rn(a6, m), c); a7 = __dadd_rn(__dmul_rn(a7, m), c);
    }
  }
  double s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
  if (s == 123.456) out[0] = s;  // keeps the loop alive; never true
}
#endif  // !SRT_TEAM_TU
