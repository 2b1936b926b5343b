// srt_team.cuh -- the cooperative ("team") form of the per-ray integrator: 8 lanes per ray.
//
// Why: a ray is a serial ODE, and one thread per ray leaves the FP64 pipe idle whenever
// the batch is small or a few long rays are left (config 2: 10^4 rays, a third of which
// run 2*10^4 steps): one accepted RK45 step is a chain of 43 plasma evaluations.  The
// 7-point stencil of raytracer_evalrhs (x, x+-dx, x+-dy, x+-dz; raytracer.f95:137,190,
// 241-261) is embarrassingly parallel, so a team of 8 lanes evaluates it in ONE plasma-
// evaluation latency: lanes 0-5 take the six displaced points, lanes 6 and 7 the centre
// (for the w+d and w-d Stix sets).  An accepted step becomes 6 rounds (stages 2-6, then
// error estimate + re-projection + the next step's stage-1 dF/dx stencil in one round)
// instead of 43 sequential evaluations.
//
// Layout of a team: lane r < 6 OWNS component r of the state vector (x,y,z,kx,ky,kz): its
// stage increments k1..k6, est1/est2 and the RK algebra for that component live only
// there (6x fewer registers than the replicated form); scalars (t, dt, w, counters) are
// replicated.  Every round is "plasma at my point -> derived -> Stix at my w", which is
// the same instruction stream for all lanes of the warp whatever phase their team is in;
// only the short tails differ per phase.  All arithmetic is the arithmetic of
// srt_physics.h in the same order, so results are bit-identical to the thread-per-ray
// kernel and to the oracle (checked by tests/test_gpu_parity.py).
#pragma once
#include "srt_kernels.cuh"

enum { SRT_PH_IDLE = 0, SRT_PH_S1DX = 1, SRT_PH_STAGE = 2, SRT_PH_EST = 3 };

struct SrtTeamCtx {
  unsigned tmask;  // lanes of my team
  int tbase;       // first lane of my team
  int r;           // my role 0..7
};

__device__ __forceinline__ double team_get(const SrtTeamCtx& T, double v, int src_role) { return __shfl_sync(T.tmask, v, T.tbase + src_role); }

// gathers the 6 owned components into a full vector on every lane
__device__ __forceinline__ void team_gather6(const SrtTeamCtx& T, double mine, double out[6]) {
#pragma unroll
  for (int j = 0; j < 6; ++j) out[j] = team_get(T, mine, j);
}

__device__ __forceinline__ void team_bcast_derived(const SrtTeamCtx& T, const SrtPlasmaDerived& in, int src_role, SrtPlasmaDerived& out) {
#pragma unroll
  for (int i = 0; i < SRT_NSPEC; ++i) {
    out.wps2[i] = team_get(T, in.wps2[i], src_role);
    out.wcs[i] = team_get(T, in.wcs[i], src_role);
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) out.B0[i] = team_get(T, in.B0[i], src_role);
  out.B0sq = team_get(T, in.B0sq, src_role);
  out.fs_thr = team_get(T, in.fs_thr, src_role);
}

// the (i, sign) perturbed refractive-index vector of dispersion_relation_dFdk (raytracer.f95:139-153)
__device__ __forceinline__ void team_dfdk_n(const double k[3], double w, int axis, int minus, double n[3], double& dk_axis) {
  const double C = srt_c_light();
#pragma unroll
  for (int j = 0; j < 3; ++j) n[j] = srt_div(k[j] * C, w);
  double dk = srt_fd_step(1.0e-8, axis == 0 ? k[0] : (axis == 1 ? k[1] : k[2]));
  double kc = axis == 0 ? k[0] : (axis == 1 ? k[1] : k[2]);
  double v = (minus ? (kc - dk) : (kc + dk)) * C / w;
  if (axis == 0) n[0] = v;
  else if (axis == 1) n[1] = v;
  else n[2] = v;
  dk_axis = dk;
}

// compaction scratch: 24 eight-byte words per lane slot, blockDim/2 lane slots (dynamic shared memory)
#define SRT_PACK_WORDS 24
extern __shared__ unsigned long long srt_pack_smem[];
__shared__ int srt_pack_count;

template <int DENS, int USE_IGRF, int MAXT>
__global__ void __launch_bounds__(MAXT, 1) srt_trace_team_kernel(const __grid_constant__ SrtRunParams P, const __grid_constant__ SrtBatch B) {
  const unsigned full = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const SrtModelRef MR = SrtModelRef();
  const SrtModel& M = MR.get();
  SrtTeamCtx T;
  T.r = lane & 7;
  T.tbase = lane & ~7;
  T.tmask = 0xffu << T.tbase;
  const int r = T.r;
  const int axis = (r < 6) ? (r >> 1) : 0;  // stencil axis of lanes 0-5
  const int minus = r & 1;                  // + side on even lanes, - side on odd lanes
  const double C = srt_c_light();
  const int outputper = P.outputper < 1 ? 1 : P.outputper;
  const bool adaptive = (P.fixedstep == 0);

  // ---- replicated scalars
  int phase = SRT_PH_IDLE, stage = 0;
  bool exhausted = false, dropped = false, attempt_start = false, finish_now = false;
  long long slot = -1;
  double w = 0.0, wl = 0.0, t = 0.0, dt = 0.0, dfdw_rec = 0.0;
  double vg[3] = {0.0, 0.0, 0.0};
  int nstep = 0, lastrefinedown = 0, stopcond = 0, cone_exit = 0;
  unsigned nrej = 0, ncone = 0, nlog = 0;
  unsigned long long my_accepted = 0, my_attempts = 0, my_ff = 0;
  // ---- owned component (meaningful on lanes 0-5; lanes 6,7 mirror component r-6 harmlessly)
  const int comp = (r < 6) ? r : (r - 6);
  double xc = 0.0;          // state component
  double dfdk_c = 0.0;      // recorded dF/dk component (lanes 0-2)
  double rhs1c = 0.0;       // stage-1 rhs component, kept across rejections
  double kk1 = 0.0, kk2 = 0.0, kk3 = 0.0, kk4 = 0.0, kk5 = 0.0, kk6 = 0.0;
  double tmpc = 0.0;        // my component of the current evaluation point
  double e1c = 0.0, e2c = 0.0;

  unsigned sync_iter = 0;
  int cap_teams = (int)(blockDim.x >> 3);  // team slots of this block that may hold a live ray
  (void)sync_iter; (void)cap_teams;
  // ---- shadow teams (block-synchronous kernel, adaptive runs).  Once the ray queue is empty a block has more team slots
  // than rays, and a long ray is a serial chain of attempts of which about one in four fails at the resonance-cone test
  // (raytracer.f95:891-899: dt -> dt/2, retry from the same state; config 2: 6 439 of the 26 477 attempts of a
  // 20 000-step ray).  The compaction step therefore seats each live ray in an EVEN team slot (the primary) and leaves
  // the odd slot next to it to a shadow team, which runs the SAME attempt from the same state with the step the retry
  // after a cone refinement would use.  If the primary's attempt then ends in a cone refinement to exactly that dt
  // (compared bit for bit), the retry has already been computed: the shadow's attempt IS the retry (same state, same dt,
  // same instructions), its outcome is taken and the failed attempt cost no extra rounds.  In every other case the
  // shadow's work is dropped.  After each estimate round both teams take the state of the live one and the shadow sets up
  // its next guess.  Pure scheduling of identical arithmetic: no result changes (tests/test_gpu_parity.py,
  // test_gpu_c3_long.py and the bench's parity check run whole batches in this mode).
  // Round 2, later: with B.shadow >= 4 a ray gets THREE shadows (a whole warp per ray) as soon as the live rays fit four times:
  // the shadow of depth q runs the retry after q failures of the kind last seen (0.8^q dt, or dt / 2^q after the growth guess),
  // and the chain of valid shadows is resolved in the estimate round (srt_team_est.inl).  tools/shadow_predict.cpp replays this.
  int group = 1;               // block-uniform: team slots are groups of 1, 2 or 4 (a primary and its shadows)
  int gq = 0, gbase = 0;       // my team's index in its group; first lane of the group
  unsigned gmask = full;       // lanes of my group
  bool shadow_ready = false;   // group-uniform: the shadows run the current attempt alongside the primary
  int live_is_S = 0, did_est = 0;  // live_is_S: index in the group of the team whose state is the ray's after this estimate round
  int shadow_kind = 2;         // shadow: the kind of failure whose retries the shadows are running
  int err_small = 1;           // the most recent error estimate of this ray was below maxerr / 100 (the growth test's condition)
  int last_fail = 0;           // kind of this ray's most recent failed attempt: 1 error test (dt -> 0.8 dt), 2 cone test (dt -> dt / 2)
  double dt_p = 0.0;           // shadow: the primary's dt (the stop test at the attempt start is the primary's)
  unsigned long long my_spec = 0;
  auto set_group = [&](int g) {
    group = g;
    gq = g == 4 ? (lane >> 3) : (g == 2 ? ((lane >> 3) & 1) : 0);
    gbase = g == 4 ? 0 : (lane & 16);
    gmask = g == 4 ? 0xffffffffu : (0xffffu << (lane & 16));
  };
  (void)last_fail; (void)err_small; (void)shadow_kind; (void)group; (void)shadow_ready; (void)live_is_S; (void)did_est; (void)dt_p; (void)gq; (void)gmask; (void)gbase; (void)my_spec; (void)set_group;
  for (;;) {
    // ================= refill: idle teams take the next ray; one atomic per warp
#if defined(SRT_TEAM_BLOCK_SYNC)
    // Phase alignment: an attempt is `cyc` rounds (stages 2..6 + estimate; fixed step: stages 2..4 + estimate) and
    // the first wave of a block starts at round 0, so every team's estimate round falls on a multiple of cyc for as
    // long as no team joins in between.  A team that refilled at another round would run its estimate branch while its
    // warp neighbours run a stage branch, in every round from then on (measured on config 3: +25 % at 1 000 steps per
    // ray, +34 % at 4 000, profiles/r2a_c3_age.txt).  So an idle team takes its next ray only at a multiple of cyc:
    // its stage-1 stencil round coincides with the others' estimate round and the block stays in phase (at most cyc-1
    // idle rounds per ray).  Pure scheduling: no result changes.
    const bool refill_round = (sync_iter % (adaptive ? 6u : 4u)) == 0u;
#else
    const bool refill_round = true;
#endif
    const bool want = (phase == SRT_PH_IDLE) && !exhausted && refill_round;
    const unsigned need = __ballot_sync(full, want && r == 0);
    if (need) {
      const int leader = __ffs(need) - 1;
      unsigned long long base = 0;
      if (lane == leader) base = atomicAdd(&B.counters[4], (unsigned long long)__popc(need));
      base = __shfl_sync(full, base, leader);
      // every lane of a wanting team computes its team's rank from the leader-lane bits
      if (want) {
        slot = (long long)(base + __popc(need & ((1u << T.tbase) - 1u)));
        if (slot >= B.nrays) {
          exhausted = true;
        } else {
          slot = srt_queue_ray(B, slot);
          double pos0[3] = {B.rays[0 * B.nrays + slot], B.rays[1 * B.nrays + slot], B.rays[2 * B.nrays + slot]};
          double dir0[3] = {B.rays[3 * B.nrays + slot], B.rays[4 * B.nrays + slot], B.rays[5 * B.nrays + slot]};
          double w0 = B.rays[6 * B.nrays + slot];
          SrtRayState st;
          double row0[SRT_ROW_DOUBLES];
          srt_ray_init<DENS, USE_IGRF>(MR, P, pos0, dir0, w0, st, row0);  // replicated on the 8 lanes (once per ray)
          xc = (comp == 0) ? st.x[0] : (comp == 1) ? st.x[1] : (comp == 2) ? st.x[2] : (comp == 3) ? st.k[0] : (comp == 4) ? st.k[1] : st.k[2];
          dfdk_c = (comp == 0) ? st.dfdk[0] : (comp == 1) ? st.dfdk[1] : st.dfdk[2];
          dfdw_rec = st.dfdw;
          w = st.w; wl = 0.0; t = 0.0; dt = st.dt;
          vg[0] = st.vg[0]; vg[1] = st.vg[1]; vg[2] = st.vg[2];
          nstep = 1; lastrefinedown = 0; stopcond = 0; cone_exit = 0; nrej = 0; ncone = 0;
          dropped = false; nlog = 0; finish_now = false;
          if (r == 0) {
            B.rays[3 * B.nrays + slot] = dir0[0];
            B.rays[4 * B.nrays + slot] = dir0[1];
            B.rays[5 * B.nrays + slot] = dir0[2];
            srt_log_row(B, slot, 0, row0, dropped);
            double* sm = B.summary + slot * SRT_SUMMARY_WORDS;  // the last recorded row lives in the summary
#pragma unroll
            for (int i = 0; i < SRT_ROW_DOUBLES; ++i) sm[i] = row0[i];
          }
          dropped = __shfl_sync(T.tmask, (int)dropped, T.tbase) != 0;
          if (!dropped) nlog = 1;
          phase = SRT_PH_S1DX;
          attempt_start = true;
        }
      }
    }
    bool is_shadow = group > 1 && gq != 0;
#if defined(SRT_TEAM_BLOCK_SYNC)
    // block-synchronous rounds: the warps of a block walk the round's code in phase and share instruction lines
    // live teams of this block; a team waiting for its refill round (phase alignment) still has the queue to look at
    const int live = __syncthreads_count(((phase != SRT_PH_IDLE || !exhausted) && r == 0 && !is_shadow) ? 1 : 0);
    if (live == 0) break;
    // ---- compaction.  Once the ray queue is empty, finished rays leave holes: a block whose few live teams are
    // spread over all its warps still issues every warp.  When at most half of the team slots in use are live, the
    // live teams are re-packed into the lowest warps through shared memory (pure data movement: results unchanged);
    // the emptied warps then only attend the barrier.  Config 2 spends most of its time in this regime.
    const int nteams = (int)(blockDim.x >> 3);
    const bool regroup = adaptive && B.shadow >= 4 && group == 2 && live * 4 <= nteams;  // pairs may become quads before the live rays have halved
    if ((++sync_iter & 31u) == 0u && (live * 2 * group <= cap_teams || regroup) && cap_teams > 4) {
      if (__syncthreads_or(exhausted ? 1 : 0)) {
        unsigned long long* pk = srt_pack_smem;
        const int nslots = (int)(blockDim.x >> 1);  // lane slots available: half a block
        if (threadIdx.x == 0) srt_pack_count = 0;
        __syncthreads();
        if (phase != SRT_PH_IDLE && !is_shadow) {
          int idx = 0;
          if (r == 0) idx = atomicAdd(&srt_pack_count, 1);
          idx = __shfl_sync(T.tmask, idx, T.tbase);
          const int o = idx * 8 + r;
          const double dv[20] = {w, wl, t, dt, dfdw_rec, vg[0], vg[1], vg[2], xc, dfdk_c, rhs1c, kk1, kk2, kk3, kk4, kk5, kk6, tmpc, e1c, e2c};
#pragma unroll
          for (int i = 0; i < 20; ++i) pk[i * nslots + o] = (unsigned long long)__double_as_longlong(dv[i]);
          pk[20 * nslots + o] = (unsigned long long)slot;
          const unsigned flags = (dropped ? 1u : 0u) | (attempt_start ? 2u : 0u) | (finish_now ? 4u : 0u) | (cone_exit ? 8u : 0u) |
                                 ((unsigned)lastrefinedown << 4) | ((unsigned)phase << 8) | ((unsigned)stage << 12) | ((unsigned)stopcond << 16);
          pk[21 * nslots + o] = (unsigned long long)flags | ((unsigned long long)(unsigned)nstep << 32);
          pk[22 * nslots + o] = (unsigned long long)nrej | ((unsigned long long)ncone << 32);
          pk[23 * nslots + o] = (unsigned long long)nlog;
        }
        __syncthreads();
        const int nlive = srt_pack_count;
        // the live rays fit twice (nlive <= cap_teams / 2): adaptive runs seat them as (primary, shadow) pairs, or, where they fit
        // four times and B.shadow asks for it, as a primary and three shadows
        const int want_group = !(adaptive && B.shadow != 0) ? 1 : ((B.shadow >= 4 && nlive * 4 <= nteams) ? 4 : (nlive * 2 <= nteams ? 2 : 1));
        set_group(want_group);
        const int seat = (int)(threadIdx.x >> 3);
        const int src_team = seat / want_group;
        if (src_team < nlive && gq == 0) {
          const int o = src_team * 8 + r;
          double dv[20];
#pragma unroll
          for (int i = 0; i < 20; ++i) dv[i] = __longlong_as_double((long long)pk[i * nslots + o]);
          w = dv[0]; wl = dv[1]; t = dv[2]; dt = dv[3]; dfdw_rec = dv[4]; vg[0] = dv[5]; vg[1] = dv[6]; vg[2] = dv[7];
          xc = dv[8]; dfdk_c = dv[9]; rhs1c = dv[10]; kk1 = dv[11]; kk2 = dv[12]; kk3 = dv[13]; kk4 = dv[14]; kk5 = dv[15]; kk6 = dv[16];
          tmpc = dv[17]; e1c = dv[18]; e2c = dv[19];
          slot = (long long)pk[20 * nslots + o];
          const unsigned long long f = pk[21 * nslots + o];
          const unsigned flags = (unsigned)f;
          dropped = flags & 1u; attempt_start = flags & 2u; finish_now = flags & 4u; cone_exit = (flags >> 3) & 1u;
          lastrefinedown = (int)((flags >> 4) & 0xfu); phase = (int)((flags >> 8) & 0xfu); stage = (int)((flags >> 12) & 0xfu);
          stopcond = (int)((flags >> 16) & 0xfu);
          nstep = (int)(unsigned)(f >> 32);
          const unsigned long long g = pk[22 * nslots + o];
          nrej = (unsigned)g; ncone = (unsigned)(g >> 32);
          nlog = (unsigned)pk[23 * nslots + o];
        } else {
          phase = SRT_PH_IDLE;
          exhausted = true;
        }
        last_fail = 0;
        shadow_ready = false;  // the shadows join at the end of their primary's next estimate round
        did_est = 0;
        cap_teams = ((want_group * nlive) + 3) & ~3;  // whole warps
        is_shadow = group > 1 && gq != 0;
        __syncthreads();
      }
    }
#else
    if (__ballot_sync(full, phase != SRT_PH_IDLE) == 0u) break;
#endif
    if (phase != SRT_PH_IDLE) {
      // ================= top of an attempt: raytracer.f95:749-763
      if (attempt_start) {
        attempt_start = false;
        double xs[6];
        team_gather6(T, xc, xs);
        int done = 0;
        if (dropped || finish_now) { done = 1; }
        else if (t >= P.tmax) { stopcond = 0; done = 1; }
        else {
          int sc = 0;
          if (srt_sqrt(srt_dot3(xs, xs)) < P.minalt) sc = 1;
          else if (srt_sqrt(srt_dot3(xs + 3, xs + 3)) == 0.0) sc = 2;
          else if (srt_sqrt(srt_dot3(vg, vg)) > 1.0 + 1e-2) sc = 3;
          else if ((is_shadow ? dt_p : dt) < (double)1e-14f) sc = 5;  // a shadow ends when its primary does
          else if (nstep >= P.maxsteps) sc = 6;
          stopcond = sc;
          if (sc != 0) done = 1;
        }
        if (done) {
          if (r == 0 && !is_shadow) {
            double* s = B.summary + slot * SRT_SUMMARY_WORDS;
            s[20] = (double)nstep;
            s[21] = (double)stopcond;
            s[22] = (double)(nrej + ncone);
            s[23] = (double)((cone_exit ? SRT_FLAG_CONE_EXIT : 0) | (dropped ? SRT_FLAG_DROPPED : 0));
            if (B.nlogged) B.nlogged[slot] = dropped ? 0u : nlog;
            if (dropped) atomicAdd(&B.counters[3], 1ull);
          }
          phase = SRT_PH_IDLE;
        } else if (phase == SRT_PH_STAGE) {
          // retry after a rejection: stage 1 is unchanged, only dt differs
          kk1 = dt * rhs1c;
          tmpc = adaptive ? xc + (SRT_A21 * kk1) : xc + 1.0 / 2.0 * kk1;
          stage = 2;
        }
      }
    }

    if (phase != SRT_PH_IDLE) {
      // ================= the round: plasma at my point, derived quantities, Stix at my w
      double y[6];  // the point this round differentiates at (x,k)
      double e1[6], e2[6];
      if (phase == SRT_PH_S1DX) team_gather6(T, xc, y);
      else if (phase == SRT_PH_STAGE) team_gather6(T, tmpc, y);
      else {
        team_gather6(T, e1c, e1);
        team_gather6(T, e2c, e2);
#pragma unroll
        for (int j = 0; j < 6; ++j) y[j] = (r == 6 && adaptive) ? e1[j] : e2[j];
      }
      double pt[3] = {y[0], y[1], y[2]};
      double dx_axis = 0.0;
      // In the error-estimate round only two centre points are needed (lane 6: est1, lane 7: est2), so lanes 0-5
      // already evaluate the six displaced points around est2 = the NEXT attempt's stage-1 dF/dx stencil
      // (raytracer.f95:819: the new point is est2(1:3)); used if the step is accepted, discarded otherwise.
      if (r < 6) {
        const double xi = axis == 0 ? y[0] : (axis == 1 ? y[1] : y[2]);
        dx_axis = srt_fd_step(M.del, xi);
        const double v = minus ? xi - dx_axis : xi + dx_axis;
        if (axis == 0) pt[0] = v;
        else if (axis == 1) pt[1] = v;
        else pt[2] = v;
      }
      const double dw = srt_fd_step(1.0e-8, w);
      double wr = w;  // the frequency my Stix set is evaluated at
      if (phase == SRT_PH_STAGE) wr = (r == 6) ? w + dw : (r == 7 ? w - dw : w);
      else if (phase == SRT_PH_EST) wr = (r >= 6 && wl != 0.0) ? wl : w;
      SrtPlasma pl;
      SrtPlasmaDerived pd;
      SrtStix sx;
      srt_plasma<DENS, USE_IGRF>(MR, pt, pl);
      srt_plasma_derive(MR, pl, pd);
      srt_stix(pd, wr, sx);

      if (phase == SRT_PH_S1DX) {
        // ---- stage 1: dF/dx at the current point; dF/dk, dF/dw are the recorded ones
        double n[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) n[j] = srt_div(y[3 + j] * C, w);
        const double FA = srt_F(n, w, pd, sx);
        const int j2 = 2 * (comp % 3);
        const double Fp = team_get(T, FA, j2), Fn = team_get(T, FA, j2 + 1);
        const double ddx = team_get(T, dx_axis, j2);
        if (comp < 3) rhs1c = -srt_div(dfdk_c, dfdw_rec);
        else rhs1c = srt_div(srt_div(Fp - Fn, ddx) / 2.0, dfdw_rec);
        kk1 = dt * rhs1c;
        tmpc = adaptive ? xc + (SRT_A21 * kk1) : xc + 1.0 / 2.0 * kk1;
        stage = 2;
        phase = SRT_PH_STAGE;
      } else if (phase == SRT_PH_STAGE) {
        // ---- raytracer_evalrhs at y (raytracer.f95:282-314)
        double nA[3];
#pragma unroll
        for (int j = 0; j < 3; ++j) nA[j] = srt_div(y[3 + j] * C, wr);  // lanes 0-5: k*C/w; 6,7: k*c/(w+-d)
        const double FA = srt_F(nA, wr, pd, sx);
        SrtPlasmaDerived pc;
        team_bcast_derived(T, pd, 6, pc);
        SrtStix sc;
        srt_stix(pc, w, sc);
        double nB[3], dk_axis;
        team_dfdk_n(y + 3, w, axis, minus, nB, dk_axis);
        const double FB = srt_F(nB, w, pc, sc);
        const double dfdw = (team_get(T, FA, 6) - team_get(T, FA, 7)) / dw / 2.0;
        const int j2 = 2 * (comp % 3);
        const double FAp = team_get(T, FA, j2), FAn = team_get(T, FA, j2 + 1);
        const double FBp = team_get(T, FB, j2), FBn = team_get(T, FB, j2 + 1);
        const double ddx = team_get(T, dx_axis, j2), ddk = team_get(T, dk_axis, j2);
        double rhsc;
        if (comp < 3) rhsc = -srt_div(srt_div(FBp - FBn, ddk) / 2.0, dfdw);
        else rhsc = srt_div(srt_div(FAp - FAn, ddx) / 2.0, dfdw);
        const double kk = dt * rhsc;
        if (adaptive) {
          if (stage == 2) { kk2 = kk; tmpc = xc + (SRT_A31 * kk1 + SRT_A32 * kk2); }
          else if (stage == 3) { kk3 = kk; tmpc = xc + (SRT_A41 * kk1 + SRT_A42 * kk2 + SRT_A43 * kk3); }
          else if (stage == 4) { kk4 = kk; tmpc = xc + (SRT_A51 * kk1 + SRT_A52 * kk2 + SRT_A53 * kk3 + SRT_A54 * kk4); }
          else if (stage == 5) { kk5 = kk; tmpc = xc + (SRT_A61 * kk1 + SRT_A62 * kk2 + SRT_A63 * kk3 + SRT_A64 * kk4 + SRT_A65 * kk5); }
          else {
            kk6 = kk;
            e1c = xc + (SRT_B41 * kk1 + SRT_B42 * kk2 + SRT_B43 * kk3 + SRT_B44 * kk4 + SRT_B45 * kk5 + SRT_B46 * kk6);
            e2c = xc + (SRT_B51 * kk1 + SRT_B52 * kk2 + SRT_B53 * kk3 + SRT_B54 * kk4 + SRT_B55 * kk5 + SRT_B56 * kk6);
            phase = SRT_PH_EST;
          }
        } else {
          if (stage == 2) { kk2 = kk; tmpc = xc + 1.0 / 2.0 * kk2; }
          else if (stage == 3) { kk3 = kk; tmpc = xc + kk3; }
          else {
            kk4 = kk;
            e2c = xc + 1.0 / 6.0 * (kk1 + 2.0 * kk2 + 2.0 * kk3 + kk4);
            e1c = e2c;
            phase = SRT_PH_EST;
          }
        }
        ++stage;
      } else {
#include "srt_team_est.inl"
      }
    }
#if defined(SRT_TEAM_BLOCK_SYNC)
    // ================= shadow groups: after the primary's estimate round every team of the group takes the state of the live one
    if (group > 1) {
      const int fl = __shfl_sync(gmask, did_est | (live_is_S << 1), gbase);
      did_est = 0;
      if (fl & 1) {
        const int src = gbase + 8 * (fl >> 1) + r;
        w = __shfl_sync(gmask, w, src); wl = __shfl_sync(gmask, wl, src); t = __shfl_sync(gmask, t, src); dt = __shfl_sync(gmask, dt, src);
        dfdw_rec = __shfl_sync(gmask, dfdw_rec, src);
        vg[0] = __shfl_sync(gmask, vg[0], src); vg[1] = __shfl_sync(gmask, vg[1], src); vg[2] = __shfl_sync(gmask, vg[2], src);
        xc = __shfl_sync(gmask, xc, src); dfdk_c = __shfl_sync(gmask, dfdk_c, src); rhs1c = __shfl_sync(gmask, rhs1c, src);
        slot = __shfl_sync(gmask, slot, src);
        unsigned f1 = (dropped ? 1u : 0u) | (attempt_start ? 2u : 0u) | (finish_now ? 4u : 0u) | (cone_exit ? 8u : 0u) |
                      ((unsigned)lastrefinedown << 4) | ((unsigned)phase << 8) | ((unsigned)stage << 12) | ((unsigned)stopcond << 16) |
                      ((unsigned)last_fail << 20) | ((unsigned)err_small << 24);
        f1 = __shfl_sync(gmask, f1, src);
        dropped = f1 & 1u; attempt_start = f1 & 2u; finish_now = f1 & 4u; cone_exit = (f1 >> 3) & 1u;
        lastrefinedown = (int)((f1 >> 4) & 0xfu); phase = (int)((f1 >> 8) & 0xfu); stage = (int)((f1 >> 12) & 0xfu);
        stopcond = (int)((f1 >> 16) & 0xfu); last_fail = (int)((f1 >> 20) & 0xfu); err_small = (int)((f1 >> 24) & 1u);
        nstep = __shfl_sync(gmask, nstep, src); nrej = __shfl_sync(gmask, nrej, src); ncone = __shfl_sync(gmask, ncone, src);
        nlog = __shfl_sync(gmask, nlog, src);
        exhausted = true;
        dt_p = dt;
        // The retries the shadows prepare are those after failures of the kind this ray met last.  Error test
        // (raytracer.f95:790-798): dt -> 0.8 dt from the same state, `w` of :821 not yet set by that attempt.  Cone test
        // (:891-899), reached after the growth test of :800-812, taken to go the way this ray's last error estimate would have
        // sent it: dt -> (dt or 1.25 dt) / 2, `w` set; a further refinement halves again (no growth after a refinement).
        // The shadow of depth gq applies gq such failures.
        shadow_kind = last_fail == 1 ? 1 : 2;
        for (int q = 0; q < gq; ++q) {
          if (shadow_kind == 1) {
            dt = 0.8 * dt;
            nrej++;
          } else {
            const bool grow = lastrefinedown == 0 && err_small && dt * 1.25 < P.dtmax;
            dt = grow ? (dt * 1.25) / 2.0 : dt / 2.0;
            ncone++;
            wl = w;
          }
          lastrefinedown = 1;
        }
        shadow_ready = true;
      }
    }
#endif
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    my_accepted += __shfl_down_sync(full, my_accepted, o);
    my_attempts += __shfl_down_sync(full, my_attempts, o);
    my_ff += __shfl_down_sync(full, my_ff, o);
    my_spec += __shfl_down_sync(full, my_spec, o);
  }
  if (lane == 0) {
    if (my_ff) atomicAdd(&B.counters[5], my_ff);
    if (my_spec) atomicAdd(&B.counters[6], my_spec);
    if (threadIdx.x == 0) atomicMax(&B.counters[7], (unsigned long long)sync_iter);
    atomicAdd(&B.counters[1], my_accepted);
    atomicAdd(&B.counters[2], my_attempts);
  }
}
