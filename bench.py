#!/usr/bin/env python
"""bench.py -- accepted ray-steps/s of the batched Haselgrove integrator on N B200s.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...     (the CPU restatement of the reference on the host cores)

One "step" = one pass of the hot path over one batch of synthetic rays.

The line (BASELINE configs[1], "C2"): per GPU 10 000 rays = 100 frequencies 0.2-30 kHz x 100 launch latitudes 10-60 deg,
Ngo density + dipole field, adaptive RK45.  With N > 1 the launch set is ONE lattice of N x 10 000 rays (N launch
longitudes) split over the ranks by ray index through the product's own shard path (stanford_raytracer_b200/shard.py):
fixed work per GPU = weak scaling, no collective on the data path; the host-side gather on rank 0 is inside `e2e`.

value   = accepted ray-steps/s, whole job, inputs resident in HBM, device-timed (CUDA events on the launching stream,
          max over ranks).
e2e     = the same metric through the reference-facing call with HOST buffers: H2D of the rays, the kernels, the
          (ray,row) ordering, D2H of every trajectory row and (N > 1) the gather of all rays on rank 0.
roofline: FP64 (non-tensor) pipe.  achieved = algorithmic flop-equivalents / kernel time; peak = DFMA
          rate measured live by srt_fp64_peak().
extra   = the other BASELINE configs on the same box in the same run (fewer sweeps): C3 (100 k rays, tricubic grid +
          IGRF), a C4-shaped 100 k-ray lattice (strong scaling: sharded over the N ranks) and, at N = 8, the full
          1 M-ray C4; each with value, e2e and roofline.  Also c5 (damping along the config-3 rays), ts05 (the C2 launch set
          with the TS05 external field) and, at N = 1, scattered (modelnum 4: tools/scattered_bench.py in a child process
          with a time limit, its own mismatch counters, no roofline yet -- its kernels had not run on a B200 when it was added).
parity  = rays of the timed sweeps traced by the CPU oracle as well (the cpu_baseline leg) and compared with the
          GPU summaries bit for bit: parity_checked_rays / parity_mismatches (non-zero exits 1).
"""
import argparse
import json
import os
import tempfile
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

import numpy as np  # noqa: E402

METRIC = "accepted_ray_steps_per_sec"
UNIT = "ray-steps/s"
ROW_BYTES = 168         # one log entry

WORKLOADS = {
    # name: (n_freq, n_lat, n_lon, outputper, model, description); n_lon None = one longitude per rank (the line)
    "c2": (100, 100, None, 1, "ngo",
           "C2: 10k-ray sweep per GPU (100 freq 0.2-30 kHz x 100 lat 10-60 deg x one launch longitude per GPU), Ngo density + dipole B, "
           "adaptive RK45 (dt0=1e-3, dtmax=0.1, maxerr=5e-4, tmax=10 s, root=2); N > 1: one lattice sharded by ray index"),
    "c3": (100, 100, 10, 100, "grid",
           "C3: 100k rays (100 freq x 100 lat x 10 lon) through a tricubic 80^3 log-density grid over +-10 R_E (sampled from "
           "the Ngo model by the GPU funcPlasmaParams kernel: GCPM cannot be built here) + IGRF field, adaptive RK45"),
    "c4_100k": (100, 100, 10, 100, "ngo",
                "C4-shaped: 100k-ray launch grid (100 freq x 100 lat x 10 lon), Ngo density + dipole B, adaptive RK45, one lattice sharded over the GPUs"),
    "c4": (100, 1000, 10, 100, "ngo",
           "C4: 1M-ray launch grid (100 freq x 1000 lat x 10 lon), Ngo density + dipole B, adaptive RK45, one lattice sharded over the GPUs"),
}
R_E = 6371.2e3
_GRID_CACHE = os.path.join(tempfile.gettempdir(), "srt_bench_grid80.txt")  # ~50 MB of text: kept out of the repo tree


def work_model():
    """flop-equivalents per accepted step / rejected attempt and where the figures come from"""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "work_tally.json")))
        return t
    except Exception:
        return {"ngo": {"accepted": 4.2e4, "rejected": 4.0e4}, "grid": {"accepted": 6.8e4, "rejected": 6.5e4},
                "source": "SURVEY.md 8(d) hand count (+-20 %)"}


def lattice(name, world):
    import common
    nf, nlat, nlon, outputper, model, desc = WORKLOADS[name]
    freqs = np.logspace(np.log10(200.0), np.log10(30000.0), nf)
    lats = np.linspace(10.0, 60.0, nlat)
    lons = np.arange(nlon if nlon else world) * 36.0
    pos0, dir0, w0 = common.launch_grid(freqs, lats, lons)
    return pos0, dir0, w0, dict(common.C2), common.NGO_CFG


def grid_file(srt, cfg, local_rank):
    from stanford_raytracer_b200 import gridfile
    path = _GRID_CACHE + ".%d" % local_rank
    if not os.path.exists(path):
        ngo = srt.ngosetup(cfg, 2010001, 0, 0, 0)
        ext = (-10 * R_E, 10 * R_E) * 3
        f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 80, 80, 80, ext)
        gridfile.write_grid(path, f, qs, ms, ext)
    return path


def make_model(srt, name, cfg, local_rank):
    """The plasma model of a workload, built through the product's own API (never the oracle)."""
    if WORKLOADS[name][4] == "ngo":
        return srt.ngosetup(cfg, 2010001, 0, 0, 0), cfg
    path = grid_file(srt, cfg, local_rank)
    return srt.interpsetup(path, 2010001, 0, 1, 0), path


class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag, self.max_mhz = index, [], set(), False, None

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for nm, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(nm)
            except Exception:
                pass
            time.sleep(0.2)

    def result(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


# ------------------------------------------------------------------------------------------ CPU side (oracle/: checker + baseline)
def cpu_run(mode, kind, path, use_igrf, pos0, dir0, w0, params, idx, threads):
    """Traces the rays `idx` of a launch set on the host cores, `threads` workers pulling rays from ONE queue (per-ray
    step counts differ by >100x, so static chunks would time the unluckiest thread).
    mode "faithful": oracle/ with the reference's call structure (8 plasma + 14 dispersion_relation calls per RHS);
    "refcost": the same plus RECALC_08 and the dipole-tilt chain on every plasma call, as the Fortran adapters do
    (ngo_dens_model_adapter.f95:169) -- what the reference binary itself would cost; "dedup": the de-duplicated
    evaluation order of the GPU kernels compiled for the host at -O3 (same algorithmic work as the GPU).  All give
    the same bits.
    Returns dict(accepted, seconds, last_row, nrows, stopcond, busy_max/busy_mean of the workers)."""
    import oracle_lib as ol
    n = len(idx)
    last = np.zeros((n, 20))
    nrows = np.zeros(n, dtype=np.int64)
    stop = np.zeros(n, dtype=np.int32)
    nthreads = max(1, min(threads, n))
    if mode in ("faithful", "refcost"):
        models = [ol.OracleModel(kind, path, use_igrf=use_igrf, faithful_cost=(mode == "refcost")) for _ in range(nthreads)]
    else:
        shared = ol.DedupModel(kind, path, use_igrf=use_igrf)
        models = [shared] * nthreads  # read-only during a trace
    lock, head, busy = threading.Lock(), [0], [0.0] * nthreads

    def work(t):
        while True:
            with lock:
                j = head[0]
                head[0] += 1
            if j >= n:
                return
            t0 = time.perf_counter()
            r = int(idx[j])
            s = models[t].trace_summary(pos0[r], dir0[r], w0[r], **params)
            last[j], nrows[j], stop[j] = s["last_row"], s["nrows"], s["stopcond"]
            busy[t] += time.perf_counter() - t0

    t0 = time.perf_counter()
    ts = [threading.Thread(target=work, args=(i,)) for i in range(nthreads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    return dict(accepted=int((nrows - 1).sum()), seconds=dt, last_row=last, nrows=nrows, stopcond=stop, threads=nthreads,
                busy_max=max(busy), busy_mean=float(np.mean(busy)))


def cpu_baselines(kind, path, use_igrf, pos0, dir0, w0, params, n_dedup, n_all, n_one, threads, short):
    """The CPU figures of SURVEY.md 8(d) on bounded samples of one launch set (strided rays): {faithful, dedup} x
    {1 core, all cores}, plus the reference's own cost (RECALC_08 per plasma call) on a small sample.  `short`
    (e.g. {"maxsteps": 300}) shortens the rays of every run but dedup/all-cores, whose full-length rays double as the
    parity sample.  Returns (modes dict, (ray indices, dedup/all-cores run))."""
    idx_all = np.unique(np.linspace(0, len(w0) - 1, n_all).astype(int))
    idx_one = np.unique(np.linspace(0, len(w0) - 1, n_one).astype(int))
    idx_par = np.unique(np.concatenate([idx_all, np.linspace(0, len(w0) - 1, n_dedup).astype(int)]))  # superset of idx_all
    ps = dict(params, **short)
    note = (" first %d steps of each ray" % short["maxsteps"]) if short else ""

    def entry(r, idx, nt):
        e = {"value": r["accepted"] / r["seconds"], "cores": r["threads"], "rays": len(idx), "seconds": r["seconds"]}
        if nt > 1:
            e["slowest_worker_over_mean"] = r["busy_max"] / max(r["busy_mean"], 1e-9)
        return e

    out = {}
    full = cpu_run("dedup", kind, path, use_igrf, pos0, dir0, w0, params, idx_par, threads)
    out["dedup_all_cores"] = entry(full, idx_par, threads)
    out["dedup_1_core"] = dict(entry(cpu_run("dedup", kind, path, use_igrf, pos0, dir0, w0, ps, idx_one, 1), idx_one, 1), note=note)
    tf = min(threads, 8) if kind == "grid" else threads  # every faithful worker parses its own copy of the grid file
    fa = cpu_run("faithful", kind, path, use_igrf, pos0, dir0, w0, ps, idx_all, tf)
    out["faithful_all_cores"] = dict(entry(fa, idx_all, tf), note=note)
    out["faithful_1_core"] = dict(entry(cpu_run("faithful", kind, path, use_igrf, pos0, dir0, w0, ps, idx_one, 1), idx_one, 1), note=note)
    idx_rc = idx_all[:: max(1, len(idx_all) // max(threads, 16))]
    rc = cpu_run("refcost", kind, path, use_igrf, pos0, dir0, w0, ps, idx_rc, tf)
    out["reference_cost_all_cores"] = dict(entry(rc, idx_rc, threads), note="RECALC_08 and the dipole-tilt chain on every plasma call, as the Fortran adapters do;" + note)
    if not short:  # the two statements of the arithmetic must agree bit for bit
        at = np.searchsorted(idx_par, idx_all)
        out["faithful_equals_dedup"] = bool(np.array_equal(fa["last_row"], full["last_row"][at], equal_nan=True) and np.array_equal(fa["nrows"], full["nrows"][at]))
    return out, (idx_par, full)


def parity_check(summary_rows, idx, cpu):
    """GPU summaries ([nrays][24]: last row, nsteps, stopcond, ...) of the rays idx against the CPU traces, bit for bit."""
    g = summary_rows[idx]
    bad = 0
    for j in range(len(idx)):
        ok = np.array_equal(g[j, :20], cpu["last_row"][j], equal_nan=True) and int(g[j, 20]) == int(cpu["nrows"][j]) and int(g[j, 21]) == int(cpu["stopcond"][j])
        bad += 0 if ok else 1
    return len(idx), bad


def translated_reference_run(cfg, pos0, dir0, w0, params, idx, threads):
    """The reference's OWN Fortran of the path -- raytracer.f95, the Ngo adapter and density model, xform_double, Geopack, the vendored
    LAPACK's zgesvd -- as translated to C++ at build time (oracle/_ref/*.so; no Fortran compiler exists), traced on the host cores:
    one dynamic ray queue, every worker thread with its own (thread-local) module state.  Each array expression of the Fortran 90
    is a heap temporary there, so this is slower than a compiled Fortran binary would be: reported beside the port, not instead of it.
    Returns None where oracle/_ref was not built."""
    import ctypes as C
    import oracle_lib as ol
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    names = ["librt_ref.so", "libngo_ref.so", "libxform_ref.so", "libgeopack_ref.so", "libts05_ref.so", "liblapack_ref.so"]
    if not all(os.path.exists(os.path.join(ref_dir, n)) for n in names):
        return None
    R, N, X, G, T, V = (C.CDLL(os.path.join(ref_dir, n)) for n in names)
    vp, d, i, l = C.c_void_p, C.c_double, C.c_int, C.c_long
    N.ngo_ref_readinput.argtypes = [C.c_char_p]
    N.ngo_ref_state.argtypes = [vp]
    R.rt_ref_ngo_adapter.argtypes = [vp] * 6 + [d, i, i, i, i, vp]
    R.rt_ref_use_ngo_adapter.argtypes = [vp]
    R.rt_ref_run.restype = l
    R.rt_ref_run.argtypes = [vp, vp, d, d, d, d, i, d, i, d, i, d, vp, l, vp]
    port = ol.OracleModel("ngo", cfg, libm=True)  # the port on glibc's libm, as the translation is: the rows must be the same bits
    n = len(idx)
    last = np.zeros((n, 20))
    nrows = np.zeros(n, dtype=np.int64)
    nthreads = max(1, min(threads, n))
    lock, head = threading.Lock(), [0]
    fp = lambda f: C.cast(f, vp)  # noqa: E731
    parmod = np.zeros(10)

    def work(t):
        if N.ngo_ref_readinput(cfg.encode()) != 0:
            raise RuntimeError("translated readinput failed")
        st = np.zeros(12)
        N.ngo_ref_state(st.ctypes.data)
        R.rt_ref_ngo_adapter(fp(X.xform_ref_sm_to_gsm), fp(X.xform_ref_gsm_to_sm), fp(G.geopack_ref_recalc), fp(G.geopack_ref_igrf_gsw), fp(T.ts05_ref_t04_s),
                             fp(N.ngo_ref_dens), st[2], 2010001, 0, 0, 0, parmod.ctypes.data)
        R.rt_ref_use_ngo_adapter(fp(V.lapack_ref_csvd_hook))
        rows = np.zeros((params["maxsteps"] + 2, 20))
        while True:
            with lock:
                j = head[0]
                head[0] += 1
            if j >= n:
                return
            r = int(idx[j])
            sc, d0 = C.c_int(), np.array(dir0[r], dtype=np.float64)
            p0 = np.ascontiguousarray(pos0[r], dtype=np.float64)
            k = R.rt_ref_run(p0.ctypes.data, d0.ctypes.data, w0[r], params["dt0"], params["dtmax"], params["maxerr"], params["maxsteps"], params["minalt"],
                             params["root"], params["tmax"], params["fixedstep"], port.del_, rows.ctypes.data, len(rows), C.byref(sc))
            last[j], nrows[j] = rows[k - 1], k

    t0 = time.perf_counter()
    ts = [threading.Thread(target=work, args=(q,)) for q in range(nthreads)]
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    same = all(np.array_equal(port.trace_summary(pos0[int(r)], dir0[int(r)], w0[int(r)], **params)["last_row"], last[j], equal_nan=True) for j, r in enumerate(idx[:8]))
    return {"value": float((nrows - 1).sum() / dt), "unit": UNIT, "cores": nthreads, "rays": int(n), "seconds": dt, "rows_equal_the_port": bool(same),
            "kind": "reference (its Fortran translated to C++ at build time)",
            "sample": "first %d steps of %d rays strided out of the 10k-ray sweep" % (params["maxsteps"], n)}


def run_reference(args):
    """The reference arm: the CPU restatement of the reference on every host core, on the line's workload (C2),
    each step a bounded sample of it.  value = the de-duplicated statement (same algorithmic work as the GPU: the
    conservative denominator); the faithful call structure is reported beside it."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib as ol
    ol.build()
    pos0, dir0, w0, params, cfg = lattice("c2", 1)
    threads = os.cpu_count() or 1
    nsample = max(threads, 256)
    idx = np.unique(np.linspace(0, len(w0) - 1, nsample).astype(int))
    small = np.unique(np.linspace(0, len(w0) - 1, max(threads, 8)).astype(int))
    for _ in range(args.warmup):
        cpu_run("dedup", "ngo", cfg, 0, pos0, dir0, w0, params, small, threads)
    acc_total, t_total = 0, 0.0
    for _ in range(args.steps):
        r = cpu_run("dedup", "ngo", cfg, 0, pos0, dir0, w0, params, idx, threads)
        acc_total += r["accepted"]
        t_total += r["seconds"]
    f = cpu_run("faithful", "ngo", cfg, 0, pos0, dir0, w0, params, idx, threads)
    idx_rc = idx[:: max(1, len(idx) // max(threads, 16))]
    rc = cpu_run("refcost", "ngo", cfg, 0, pos0, dir0, w0, params, idx_rc, threads)
    val = acc_total / t_total
    tr = translated_reference_run(cfg, pos0, dir0, w0, dict(params, maxsteps=1000), idx[:: max(1, len(idx) // (8 * threads))], threads)
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference", "config": config_dict(args.gpus, "c2", 10000, 10000 * args.gpus),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d rays strided out of the 10k-ray sweep per step, one dynamic ray queue; de-duplicated evaluation order (same work as the GPU) "
                                       "of the C++ restatement of the Fortran (no Fortran compiler on the box)" % len(idx),
                             "faithful_call_structure": {"value": f["accepted"] / f["seconds"], "cores": f["threads"], "seconds": f["seconds"], "rays": len(idx)},
                             "reference_cost": {"value": rc["accepted"] / rc["seconds"], "cores": rc["threads"], "seconds": rc["seconds"], "rays": len(idx_rc),
                                                "note": "RECALC_08 and the dipole-tilt chain on every plasma call, as the Fortran adapters do"},
                             "translated_reference": tr},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def config_dict(n_gpus, name, rays_per_gpu, global_rays):
    return {"workload": WORKLOADS[name][5], "rays_per_gpu": int(rays_per_gpu), "global_rays": int(global_rays), "outputper": WORKLOADS[name][3],
            "parallelism": "one launch set sharded by ray index over the ranks, host-side gather on rank 0, no collective",
            "l2": "flushed between timed iterations (256 MiB write)"}


# ------------------------------------------------------------------------------------------ GPU side
class Ctx:
    pass


def measure(C, name, steps, warmup, e2e_steps, with_cpu, cpu_sizes):
    """Times one workload: device-resident sweeps of this rank's shard + the end-to-end sharded call.  Collective."""
    import torch
    import torch.distributed as dist
    srt, shard = C.srt, C.shard
    outputper, kind = WORKLOADS[name][3], WORKLOADS[name][4]
    pos0, dir0, w0, params, cfg = lattice(name, C.world)
    nglobal = len(w0)
    idx = shard.shard_indices(nglobal, C.rank, C.world)
    nrays = len(idx)
    model, model_path = make_model(srt, name, cfg, C.local_rank)
    P = srt.RunParams(params["dt0"], params["dtmax"], params["maxerr"], params["minalt"], params["tmax"], params["maxsteps"], params["root"],
                      params["fixedstep"], outputper)
    dev = C.dev
    rays_h = np.ascontiguousarray(np.concatenate([pos0[idx].T, dir0[idx].T, w0[idx][None, :]], axis=0))  # [7][nrays]
    rays0 = torch.from_numpy(rays_h).to(dev)
    rays = rays0.clone()
    summary = torch.empty((nrays, srt.SUMMARY_WORDS), dtype=torch.float64, device=dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    stream = torch.cuda.current_stream()
    # size the trajectory log from a counting pass (no log attached: nothing recorded, nothing dropped)
    t_count = time.perf_counter()
    srt.trace_batch_device(model, nrays, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    c0 = counters.cpu().numpy()
    if C.rank == 0:
        print("bench[%s]: counting pass %.1f s, %d accepted steps, %d attempts (rank 0 shard: %d rays)" % (name, time.perf_counter() - t_count, int(c0[1]), int(c0[2]), nrays),
              file=sys.stderr, flush=True)
    log_cap = int(c0[1]) // outputper + 2 * nrays + 1024
    log = torch.empty((log_cap, srt.LOG_WORDS), dtype=torch.float64, device=dev)

    def barrier():
        if C.world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(warmup, 0)):
        rays.copy_(rays0)
        srt.trace_batch_device(model, nrays, rays.data_ptr(), P, log.data_ptr(), log_cap, summary.data_ptr(), counters.data_ptr(), stream.cuda_stream)
    torch.cuda.synchronize()
    c = counters.cpu().numpy()
    if c[3] != 0:
        raise SystemExit("trajectory log overflowed in the bench (%d rays dropped)" % c[3])
    accepted, attempts, rows, ff = int(c[1]), int(c[2]), int(c[0]), int(c[5])

    sampler = ClockSampler(C.local_rank)
    sampler.start()
    launches0 = srt.launch_count()
    evs = []
    barrier()
    for _ in range(steps):
        C.flush.zero_()  # L2 flush, outside the per-step event pair
        rays.copy_(rays0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        srt.trace_batch_device(model, nrays, rays.data_ptr(), P, log.data_ptr(), log_cap, summary.data_ptr(), counters.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        evs.append((e0, e1))
    barrier()
    launches = srt.launch_count() - launches0
    kernel_ms = sum(a.elapsed_time(b) for a, b in evs)
    sampler.stop_flag = True
    t_dev = torch.tensor([kernel_ms], dtype=torch.float64, device=dev)
    tot = torch.tensor([accepted * steps, attempts * steps, rows * steps, ff * steps, launches], dtype=torch.float64, device=dev)
    if C.world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    kernel_ms_max = float(t_dev.item())
    total_accepted, total_attempts, total_rows, total_ff, total_launches = [float(v) for v in tot.tolist()]
    secs = kernel_ms_max * 1e-3
    summary_h = summary.cpu().numpy()
    del log

    # ---- end to end: the sharded reference-facing call with host buffers, gather on rank 0 inside the timed region
    e2e = {"value": None, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    if e2e_steps > 0:
        def one():
            res = shard.raytracer_run_sharded(model, pos0, dir0.copy(), w0, C.rank, C.world, C.host_group, outputper=outputper, **params)
            acc = rws = 0
            if C.rank == 0:
                acc, rws = res.accepted_steps, res.total_rows
                touched = 0.0
                for r in range(0, nglobal, max(1, nglobal // 512)):  # the gathered rows are readable on rank 0
                    rr = res.rows(r)
                    if len(rr):
                        touched += float(rr[-1, 0])
            res.close()
            return acc, rws

        try:
            one()  # warm-up (page-locked row buffers are created once and pooled)
            barrier()
            t0 = time.perf_counter()
            acc = rws = 0
            for _ in range(e2e_steps):
                a, r_ = one()
                acc += a
                rws = r_
            torch.cuda.synchronize()
            t_e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
            failed = None
        except Exception as ex:  # (raised on every rank together, shard.run_sharded) e.g. no room for the row buffers in host memory
            failed = "%s: %s" % (type(ex).__name__, ex)
            t_e = torch.tensor([1.0], dtype=torch.float64, device=dev)
            acc = rws = 0
        if C.world > 1:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
        if C.rank == 0 and failed:
            e2e["error"] = failed
        elif C.rank == 0:
            e2e = {"value": acc / float(t_e.item()), "unit": UNIT, "h2d_bytes_per_step": int(nglobal * 56),
                   "d2h_bytes_per_step": int(rws * 160 + nglobal * (srt.SUMMARY_WORDS * 8 + 4 + 56) + 64 * C.world),
                   "gather": "per-ray records (256 B) over gloo, rows mapped from the owners' shared-memory segments" if C.world > 1 else "single process"}

    out = None
    if C.rank == 0:
        wm = work_model()
        fa, fr = wm[kind]["accepted"], wm[kind]["rejected"]
        integrated = accepted - ff  # fast-forwarded NaN steps cost one addition each: not in the work model
        flops_per_launch = integrated * fa + (attempts - accepted) * fr
        avg_launch_s = (kernel_ms / steps) * 1e-3
        achieved = flops_per_launch / avg_launch_s / 1e12
        hbm_ach = (rows * ROW_BYTES + nrays * (56 + srt.SUMMARY_WORDS * 8)) / avg_launch_s / 1e9
        ki = model.kernel_info(nrays)
        roofline = {"bound": "fp64", "achieved": achieved, "peak": C.peak_fma, "unit": "TFLOP/s", "frac": achieved / C.peak_fma,
                    "traffic": traffic_from_profiles(name, ki, accepted),
                    "bound_note": "FP64 (non-tensor) pipe: the path is scalar FP64 with no dense contraction and ~1e-4 of HBM bandwidth, so neither 'hbm' nor 'tensor' applies",
                    "peak_source": "DFMA micro-kernel measured live (srt_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
                    "peak_no_fma": C.peak_nofma, "frac_no_fma": achieved / C.peak_nofma,
                    "work_model": "%.3g flop-eq per integrated accepted RK45 step + %.3g per rejected attempt (%s); fast-forwarded NaN steps count 0" % (fa, fr, wm.get("source", "")),
                    "hbm": {"achieved": hbm_ach, "peak": C.hbm_peak, "unit": "GB/s", "frac": hbm_ach / C.hbm_peak, "of": C.hbm_of}}
        out = {"value": total_accepted / secs, "ms_per_step": kernel_ms_max / steps, "steps": steps, "warmup": warmup,
               "rays_per_sec": float(nglobal) * steps / secs, "attempted_steps_per_sec": total_attempts / secs,
               "value_without_fastforwarded_nan_steps": (total_accepted - total_ff) / secs,
               "fastforwarded_nan_steps_per_sweep": total_ff / steps,
               "accepted_steps_per_sweep": total_accepted / steps, "e2e": e2e, "gpu_launches": int(total_launches), "clocks": sampler.result(),
               "roofline": roofline, "kernel": ki, "config": config_dict(C.world, name, nrays, nglobal)}
        if with_cpu:
            threads = os.cpu_count() or 1
            n_dedup, n_all, n_one, short = cpu_sizes(threads)
            modes, (cidx, crun) = cpu_baselines(kind, model_path, 1 if kind == "grid" else 0, pos0, dir0, w0, params, n_dedup, n_all, n_one, threads, short)
            d = modes["dedup_all_cores"]
            out["cpu_baseline"] = {"value": d["value"], "unit": UNIT, "cores": d["cores"], "kind": "port",
                                   "sample": "%d whole rays strided out of the launch set, %.1f s, one dynamic ray queue; de-duplicated evaluation order (same algorithmic work as "
                                             "the GPU) of the C++ restatement of the reference" % (d["rays"], d["seconds"]),
                                   "modes": modes}
            if kind == "ngo" and name == "c2":  # the reference's own translated source on the same cores, for scale (rank 0, N = 1 only)
                tidx = np.unique(np.linspace(0, len(w0) - 1, 4 * threads).astype(int))
                try:
                    out["cpu_baseline"]["translated_reference"] = translated_reference_run(model_path, pos0, dir0, w0, dict(params, maxsteps=1000), tidx, threads)
                except Exception as e:  # an auxiliary figure: it must not cost the line
                    out["cpu_baseline"]["translated_reference"] = {"unavailable": repr(e)}
            # the same rays on the GPU (this rank's shard holds those with shard_of == 0), bit for bit
            mine = np.searchsorted(idx, cidx)
            ok = (mine < len(idx)) & (idx[np.minimum(mine, len(idx) - 1)] == cidx)
            sel = {k: v[ok] if isinstance(v, np.ndarray) else v for k, v in crun.items()}
            out["parity_checked_rays"], out["parity_mismatches"] = parity_check(summary_h, mine[ok], sel)
    model.close()
    return out


def measure_damping(C, with_cpu):
    """BASELINE config 5: Landau damping along every ray of the C3 run (damping mode 1, Kp = 4, AE = 3, no geometric
    factor: the defaults of damping_main.cpp:54-61).  Rank 0 only.  value = trajectory rows damped per second through
    srt_result_damping (rows H2D + the two kernels + power D2H); the CPU figure is the REFERENCE's own binary
    (oracle/_ref/damping, one core: it is single-threaded) on a strided sample of the same rays, whose powers are also
    compared with the GPU's."""
    if C.rank != 0:
        return None
    import subprocess
    import tempfile as tf
    srt = C.srt
    from stanford_raytracer_b200 import damping
    pos0, dir0, w0, params, cfg = lattice("c3", 1)
    model, _ = make_model(srt, "c3", cfg, C.local_rank)
    res = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=100, **params)
    damping.damp_result(res)  # warm-up
    t0 = time.perf_counter()
    reps = 2
    for _ in range(reps):
        first, power = damping.damp_result(res)
    dt = (time.perf_counter() - t0) / reps
    out = {"metric": "damped_trajectory_rows_per_sec", "unit": "rows/s", "value": res.total_rows / dt, "rows": int(res.total_rows), "rays": int(res.nrays),
           "seconds_per_pass": dt, "h2d_bytes": int(res.total_rows * 160), "d2h_bytes": int(res.total_rows * 8),
           "config": "C5: the C3 run's rows (outputper 100) -> damping mode 1, Kp 4, AE 3, geom_factor 0", "gpu_launches": 2 * reps,
           "work": "per row 50 Gauss-Legendre nodes x 3 resonances x (8 phase-space-density evaluations + 4 Bessel functions)"}
    ref = os.path.join(ROOT, "oracle", "_ref", "damping")
    if with_cpu and os.path.exists(ref):
        idx = np.unique(np.linspace(0, len(w0) - 1, 96).astype(int))
        sub = srt.raytracer_run(model, pos0[idx], dir0[idx].copy(), w0[idx], outputper=100, **params)
        with tf.TemporaryDirectory() as d:
            import shutil
            shutil.copy(os.path.join(ROOT, "oracle", "_ref", "crres_clean.bin"), os.path.join(d, "crres_clean.bin"))
            sub.write_text(os.path.join(d, "in.ray"))
            t0 = time.perf_counter()
            subprocess.run([ref, "-i", "in.ray", "-o", "out.damp"], cwd=d, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            t_ref = time.perf_counter() - t0
            a = np.loadtxt(os.path.join(d, "out.damp")).reshape(-1, 3)
        f2, p2 = damping.damp_result(sub)
        big = np.isfinite(a[:, 2]) & (a[:, 2] > 1e-6)
        rel = np.abs(p2[big] / a[big, 2] - 1.0)
        out["cpu_baseline"] = {"value": sub.total_rows / t_ref, "unit": "rows/s", "cores": 1, "kind": "reference",
                               "sample": "%d rays strided out of the 100k (%d rows), %.1f s incl. reading the ray file; the reference's own damping binary (single-threaded)" % (len(idx), sub.total_rows, t_ref)}
        out["parity"] = {"rows_compared": int(big.sum()), "median_relative_difference": float(np.median(rel)) if rel.size else 0.0,
                         "rows_within_1e-6": float((rel <= 1e-6).mean()) if rel.size else 1.0, "rows_within_1e-3": float((rel <= 1e-3).mean()) if rel.size else 1.0,
                         "max_relative_difference": float(rel.max()) if rel.size else 0.0,
                         "nan_rows_agree": bool(np.array_equal(np.isfinite(a[:, 2]), np.isfinite(p2))),
                         "note": "floating point: the reference's own answer moves by 1e-8 ... 1e-2 when a fitted density parameter moves by 1 ulp (finite differences with "
                                 "relative step 1e-8); the per-row bound is checked in tests/test_gpu_damping.py"}
        sub.close()
    res.close()
    model.close()
    return out


def measure_ts05(C, with_cpu):
    """the C2 launch set with use_tsyganenko = 1 (TS05 external field added to the dipole), the first
    300 steps of every ray -- a bounded sample: the external field is ~100 x the arithmetic of the rest of a plasma
    evaluation.  Rank 0 only.  value = accepted ray-steps/s, rays resident on the device; e2e through raytracer_run with
    host buffers.  CPU: the oracle's adapter with the REFERENCE'S OWN T04_s (tsyganenko/TS05_aka_TS04.for translated at build
    time into oracle/_ref) on a strided sample of the same rays, all host cores; the same rays are compared bit for bit."""
    if C.rank != 0:
        return None
    import torch
    import common
    srt = C.srt
    parmod = [3.0, -45.0, 2.5, -4.0, 0.4, 0.7, 0.3, 0.9, 0.6, 1.1]
    pos0, dir0, w0, params, cfg = lattice("c2", 1)
    params = dict(params, maxsteps=300)
    n = len(w0)
    model = srt.ngosetup(cfg, 2010001, 0, 0, 1).tsyganenko(*parmod[:4], W=parmod[4:])
    P = srt.RunParams(params["dt0"], params["dtmax"], params["maxerr"], params["minalt"], params["tmax"], params["maxsteps"], params["root"], params["fixedstep"], 1)
    rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
    summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
    counters = torch.zeros(8, dtype=torch.int64, device="cuda")
    l0 = srt.lib().srt_launch_count()
    times = []
    for rep in range(3):  # 1 warm + 2 timed
        rays = rays0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        srt.trace_batch_device(model, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1) * 1e-3)
    c = counters.cpu().numpy()
    sec = float(np.mean(times[1:]))
    out = {"metric": METRIC, "unit": UNIT, "value": float(c[1]) / sec, "ms_per_step": sec * 1e3, "steps": 2, "warmup": 1,
           "accepted_steps_per_sweep": int(c[1]), "attempts_per_sweep": int(c[2]), "gpu_launches": int(srt.lib().srt_launch_count() - l0),
           "config": {"workload": "C2 launch set (10k rays), use_tsyganenko = 1: TS05 external field (Pdyn 3 nPa, Dst -45 nT, By 2.5, Bz -4 nT, W1-6 0.3-1.1) on the dipole, "
                                  "first 300 steps of every ray", "rays": n, "maxsteps": 300},
           "kernel": model.kernel_info(n),
           "work": "per plasma evaluation one T04_s call: ~1e5 flop-equivalents (timed: 43 us on one host core), against ~7e2 for Ngo density + dipole"}
    t0 = time.perf_counter()
    res = srt.raytracer_run(model, pos0, dir0, w0, **params)
    dt = time.perf_counter() - t0
    out["e2e"] = {"value": res.accepted_steps / dt, "unit": UNIT, "h2d_bytes_per_step": int(n * 56), "d2h_bytes_per_step": int(res.total_rows * 160 + n * 192)}
    if with_cpu:
        import oracle_lib as ol
        if ol.ts05_ref(True) is None:
            out["cpu_baseline"] = {"unavailable": "oracle/_ref/libts05_ref_srtmath.so not built"}
            return out
        threads = os.cpu_count() or 1
        idx = np.unique(np.linspace(0, n - 1, max(threads, 16)).astype(int))
        models = [ol.OracleModel("ngo", cfg).tsyganenko(parmod) for _ in range(min(threads, len(idx)))]
        last, nrows, lock, head = np.zeros((len(idx), 20)), np.zeros(len(idx), dtype=np.int64), threading.Lock(), [0]

        def work(t):
            while True:
                with lock:
                    j = head[0]
                    head[0] += 1
                if j >= len(idx):
                    return
                r = int(idx[j])
                sm = models[t].trace_summary(pos0[r], dir0[r], w0[r], **params)
                last[j], nrows[j] = sm["last_row"], sm["nrows"]

        t0 = time.perf_counter()
        ts = [threading.Thread(target=work, args=(i,)) for i in range(len(models))]
        [t.start() for t in ts]
        [t.join() for t in ts]
        dt = time.perf_counter() - t0
        g = summary.cpu().numpy()[idx]
        bad = sum(0 if (np.array_equal(g[j, :20], last[j], equal_nan=True) and int(g[j, 20]) == int(nrows[j])) else 1 for j in range(len(idx)))
        out["cpu_baseline"] = {"value": float((nrows - 1).sum()) / dt, "unit": UNIT, "cores": len(models), "kind": "port + reference",
                               "sample": "%d rays strided out of the 10k, first 300 steps, %.1f s: the oracle's adapter and integrator (faithful call structure) around the "
                                         "reference's own T04_s as translated into oracle/_ref" % (len(idx), dt)}
        out["parity_checked_rays"], out["parity_mismatches"] = len(idx), int(bad)
    return out


def measure_scattered(C):
    """SURVEY.md 8(f) row f4, second half: modelnum 4 (scattered samples, moving least squares).  tools/scattered_bench.py in a
    child process with a time limit, rank 0 only: the model's kernels had not run on a B200 when this was written (no GPU
    minutes were left), so nothing they do -- fault, hang, wrong bits -- can reach this process or the line's parity gate;
    its own counters (points_differing, rays_differing) say whether the GPU equals the CPU oracle there."""
    if C.rank != 0:
        return None
    import subprocess
    try:
        r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "scattered_bench.py")], capture_output=True, text=True, timeout=180)
        lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
        if not lines:
            return {"error": "no result (exit code %d): %s" % (r.returncode, r.stderr.strip()[-300:])}
        return json.loads(lines[-1])
    except Exception as e:  # an extra must not take the line down
        return {"error": repr(e)}


def traffic_from_profiles(name, ki, accepted_per_launch):
    """DRAM bytes per launch of the trace kernel on this workload (dram__bytes_read.sum + dram__bytes_write.sum of a
    committed `ncu --set full` capture, profiles/traffic.json).  The captures run the workload's own launch set with
    fewer steps per ray (ncu replays a launch ~40 times), so the figure is bytes per accepted step of the capture x the
    accepted steps of this launch -- and only if the capture measured the kernel this library would launch now (same
    mode, block, registers, local bytes); else null."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        e = t.get(name)
        if e and all(e["kernel"].get(k) == ki.get(k) for k in ("mode", "block", "regs", "local_bytes")):
            return float(e["dram_bytes_per_accepted_step"]) * float(accepted_per_launch)
    except Exception:
        pass
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="c2 is the bench line; the others run as `extra`")
    ap.add_argument("--extras", default="auto", help="auto (c3, c4_100k, c5, ts05, scattered at 1 GPU, and c4 at 8 GPUs), none, or a comma list of workloads")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import stanford_raytracer_b200 as srt
    from stanford_raytracer_b200 import shard

    C = Ctx()
    C.srt, C.shard = srt, shard
    C.world = int(os.environ.get("WORLD_SIZE", "1"))
    C.rank = int(os.environ.get("RANK", "0"))
    C.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libsrt has no CPU path")
    torch.cuda.set_device(C.local_rank)
    C.host_group = None
    if C.world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", C.local_rank))
        C.host_group = shard.host_group(C.world)
    C.dev = torch.device("cuda", C.local_rank)
    srt.init(C.local_rank)
    C.flush = torch.empty(256 << 20, dtype=torch.uint8, device=C.dev)
    C.peak_fma = srt.fp64_peak(True) / 1e3   # TFLOP/s
    C.peak_nofma = srt.fp64_peak(False) / 1e3
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    C.hbm_peak, C.hbm_of = peaks.get("hbm_gbs", 6650.0), ("measured" if "hbm_gbs" in peaks else "fallback")

    t_start = time.perf_counter()
    e2e_steps = 0 if args.no_e2e else max(1, min(args.steps, 3))
    with_cpu = not args.no_cpu_baseline

    def c2_cpu(threads):  # rays of: the de-duplicated all-cores run (= parity sample), the faithful all-cores runs, the 1-core runs
        return 1024, max(threads, 256), max(8, 256 // max(threads, 1)), {}
    # (the de-duplicated all-cores run doubles as the parity sample: cpu_baselines() takes n_all rays for it)

    def c3_cpu(threads):  # the faithful statement of a gridded ray is a dense 64x64 product per species and query: shortened rays there
        return 256, max(threads, 64), 4, {"maxsteps": 300}

    main_out = measure(C, args.workload, args.steps, args.warmup, e2e_steps, with_cpu, c3_cpu if WORKLOADS[args.workload][4] == "grid" else c2_cpu)
    extras = {}
    if args.extras != "none":
        names = (["c3", "c4_100k"] + (["c4"] if C.world >= 8 else [])) if args.extras == "auto" else [s for s in args.extras.split(",") if s]
        for nm in names:
            if nm == args.workload or nm not in WORKLOADS:
                continue
            o = measure(C, nm, 2, 1, 0 if args.no_e2e else 1, with_cpu and nm == "c3", c3_cpu if WORKLOADS[nm][4] == "grid" else c2_cpu)
            if C.rank == 0:
                extras[nm] = o
        if args.extras == "auto" or "c5" in names:
            o = measure_damping(C, with_cpu)
            if C.rank == 0:
                extras["c5"] = o
        if args.extras == "auto" or "ts05" in names:
            try:
                o = measure_ts05(C, with_cpu)
            except Exception as e:  # an extra must not take the line down
                o = {"error": repr(e)}
            if C.rank == 0:
                extras["ts05"] = o
        if (args.extras == "auto" and C.world == 1) or "scattered" in names:
            o = measure_scattered(C)
            if C.rank == 0:
                extras["scattered"] = o
    if C.rank == 0:
        m = main_out
        line = {"metric": METRIC, "value": m["value"], "unit": UNIT, "n_gpus": C.world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": m["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": m["config"], "rays_per_sec": m["rays_per_sec"], "attempted_steps_per_sec": m["attempted_steps_per_sec"],
                "value_without_fastforwarded_nan_steps": m["value_without_fastforwarded_nan_steps"],
                "e2e": m["e2e"], "gpu_launches": m["gpu_launches"], "clocks": m["clocks"], "roofline": m["roofline"],
                "cpu_baseline": m.get("cpu_baseline"), "kernel": m["kernel"],
                "parity_checked_rays": m.get("parity_checked_rays", 0) + sum(e.get("parity_checked_rays", 0) for e in extras.values()),
                "parity_mismatches": m.get("parity_mismatches", 0) + sum(e.get("parity_mismatches", 0) for e in extras.values()),
                "extra": extras, "bench_wall_s": time.perf_counter() - t_start}
        print(json.dumps(line))
        bad = line["parity_mismatches"]
    else:
        bad = 0
    if C.world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if bad:
        raise SystemExit("bench: %d rays of the timed sweep differ from the CPU oracle" % bad)


if __name__ == "__main__":
    main()
