#!/usr/bin/env python
"""bench.py -- accepted ray-steps/s of the batched Haselgrove integrator on N B200s.

    python bench.py --gpus 1 --steps 5 --warmup 3
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...
    python bench.py --impl reference ...     (the CPU restatement of the reference on the host cores)

One "step" = one pass of the hot path over one batch of synthetic rays: BASELINE config 2
(10 000 rays = 100 frequencies 0.2-30 kHz x 100 launch latitudes 10-60 deg, Ngo density +
dipole field, adaptive RK45) per GPU.  With N > 1 every rank traces its own 10 000-ray sweep
(the launch longitude differs per rank), no collective on the data path: weak scaling.

value   = accepted ray-steps/s, whole job, inputs resident in HBM, device-timed (CUDA events on
          the launching stream, max over ranks).
e2e     = the same metric through srt_trace_batch() with HOST buffers: H2D of the rays, the
          kernels, the (ray,row) gather and D2H of every trajectory row inside the timed region.
roofline: FP64 (non-tensor) pipe.  achieved = algorithmic flop-equivalents (SURVEY.md 8d:
          4.2e4 per accepted RK45 step, 4.0e4 per rejected attempt, Ngo+dipole) / kernel time;
          peak = DFMA rate measured live by srt_fp64_peak().
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
    # name: (n_freq, n_lat, n_lon, outputper, model, flop-eq per accepted step, per rejected attempt, description)
    "c2": (100, 100, 1, 1, "ngo", 4.2e4, 4.0e4,
           "C2: 10k-ray sweep (100 freq 0.2-30 kHz x 100 lat 10-60 deg), Ngo density + dipole B, adaptive RK45 "
           "(dt0=1e-3, dtmax=0.1, maxerr=5e-4, tmax=10 s, root=2), per GPU"),
    "c3": (100, 100, 10, 100, "grid", 6.8e4, 6.5e4,
           "C3: 100k rays (100 freq x 100 lat x 10 lon) through a tricubic 80^3 log-density grid over +-10 R_E (sampled from "
           "the Ngo model by the GPU funcPlasmaParams kernel: GCPM cannot be built here) + IGRF field, adaptive RK45, per GPU"),
    "c4": (100, 1000, 10, 100, "ngo", 4.2e4, 4.0e4,
           "C4: 1M-ray launch grid (100 freq x 1000 lat x 10 lon), Ngo density + dipole B, adaptive RK45, per GPU"),
}
R_E = 6371.2e3
_GRID_CACHE = os.path.join(tempfile.gettempdir(), "srt_bench_grid80.txt")  # ~50 MB of text: kept out of the repo tree


def workload(rank, name="c2"):
    import common
    nf, nlat, nlon, outputper, model, fa, fr, desc = WORKLOADS[name]
    freqs = np.logspace(np.log10(200.0), np.log10(30000.0), nf)
    lats = np.linspace(10.0, 60.0, nlat)
    lons = (np.arange(nlon) * 36.0 + 36.0 * rank / max(nlon, 1)) if nlon > 1 else [36.0 * rank]
    pos0, dir0, w0 = common.launch_grid(freqs, lats, lons)
    return pos0, dir0, w0, dict(common.C2), common.NGO_CFG


def make_model(srt, name, cfg):
    """The plasma model of a workload, built through the product's own API (never the oracle)."""
    if WORKLOADS[name][4] == "ngo":
        return srt.ngosetup(cfg, 2010001, 0, 0, 0)
    from stanford_raytracer_b200 import gridfile
    os.makedirs(os.path.dirname(_GRID_CACHE), exist_ok=True)
    path = _GRID_CACHE + ".%d" % int(os.environ.get("LOCAL_RANK", "0"))
    if not os.path.exists(path):
        ngo = srt.ngosetup(cfg, 2010001, 0, 0, 0)
        ext = (-10 * R_E, 10 * R_E) * 3
        f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 80, 80, 80, ext)
        gridfile.write_grid(path, f, qs, ms, ext)
    return srt.interpsetup(path, 2010001, 0, 1, 0)


def config_dict(n_gpus, name="c2", nrays=10000):
    return {"workload": WORKLOADS[name][7], "rays_per_gpu": nrays, "global_rays": nrays * n_gpus, "outputper": WORKLOADS[name][3],
            "parallelism": "rays sharded by rank, no collective", "l2": "flushed between timed iterations (256 MiB write)"}


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


def cpu_reference_run(pos0, dir0, w0, params, cfg, nsample, threads, kind="ngo", use_igrf=0):
    """Times the CPU oracle (the port of the reference) on a strided sample of the workload,
    one model instance per thread.  Returns (accepted steps, seconds)."""
    import oracle_lib as ol
    idx = np.linspace(0, len(w0) - 1, nsample).astype(int)
    chunks = [idx[i::threads] for i in range(threads)]
    chunks = [c for c in chunks if len(c)]
    models = [ol.OracleModel(kind, cfg, use_igrf=use_igrf) for _ in chunks]
    acc = [0] * len(chunks)

    def work(i):
        c = chunks[i]
        acc[i] = models[i].trace_batch_summary(pos0[c], dir0[c], w0[c], **params)["accepted"]

    t0 = time.perf_counter()
    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(chunks))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    return int(sum(acc)), dt


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle_lib as ol
    ol.build()
    pos0, dir0, w0, params, cfg = workload(0, "c2")
    threads = os.cpu_count() or 1
    nsample = max(threads, 256)  # ~12 s of CPU work per step on 16 cores
    for _ in range(args.warmup):
        cpu_reference_run(pos0, dir0, w0, params, cfg, max(threads, 8), threads)
    acc_total, t_total = 0, 0.0
    for _ in range(args.steps):
        a, t = cpu_reference_run(pos0, dir0, w0, params, cfg, nsample, threads)
        acc_total += a
        t_total += t
    val = acc_total / t_total
    line = {"metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * t_total / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "impl": "reference", "config": config_dict(args.gpus),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": "%d rays strided out of the 10k-ray sweep per step (C++ restatement of the Fortran; no Fortran compiler on the box)" % nsample},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg (supplementary workloads whose sweep takes minutes)")
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS), help="c2 is the bench line; c3/c4 are the other BASELINE configs")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    import stanford_raytracer_b200 as srt

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: libsrt has no CPU path")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    srt.init(local_rank)

    wl = args.workload
    outputper = WORKLOADS[wl][3]
    FLOPS_ACCEPTED, FLOPS_REJECTED = WORKLOADS[wl][5], WORKLOADS[wl][6]
    pos0, dir0, w0, params, cfg = workload(rank, wl)
    nrays = len(w0)
    model = make_model(srt, wl, cfg)
    P = srt.RunParams(params["dt0"], params["dtmax"], params["maxerr"], params["minalt"], params["tmax"], params["maxsteps"], params["root"],
                      params["fixedstep"], outputper)
    # device-resident inputs (SoA) and outputs
    rays_h = np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))  # [7][nrays], C order (concatenate keeps F order)
    rays0 = torch.from_numpy(rays_h).to(dev)
    rays = rays0.clone()
    summary = torch.empty((nrays, srt.SUMMARY_WORDS), dtype=torch.float64, device=dev)
    counters = torch.zeros(8, dtype=torch.int64, device=dev)
    # size the trajectory log from a counting pass (no log attached: nothing recorded, nothing dropped)
    t_count = time.perf_counter()
    srt.trace_batch_device(model, nrays, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    if rank == 0:
        print("bench: counting pass %.1f s, %d accepted steps, %d attempts" % (time.perf_counter() - t_count, int(counters[1].item()), int(counters[2].item())),
              file=sys.stderr, flush=True)
    log_cap = int(counters[1].item()) // outputper + 2 * nrays + 1024
    log = torch.empty((log_cap, srt.LOG_WORDS), dtype=torch.float64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream()

    def device_step():
        rays.copy_(rays0)
        srt.trace_batch_device(model, nrays, rays.data_ptr(), P, log.data_ptr(), log_cap, summary.data_ptr(), counters.data_ptr(), stream.cuda_stream)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 0)):
        device_step()
    torch.cuda.synchronize()
    c = counters.cpu().numpy()
    if c[3] != 0:
        raise SystemExit("trajectory log overflowed in the bench (%d rays dropped)" % c[3])
    accepted_per_step, attempts_per_step, rows_per_step = int(c[1]), int(c[2]), int(c[0])

    sampler = ClockSampler(local_rank)
    sampler.start()
    launches0 = srt.launch_count()
    evs = []
    barrier()
    for _ in range(args.steps):
        flush.zero_()  # L2 flush, outside the per-step event pair
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
    if world > 1:
        dist.all_reduce(t_dev, op=dist.ReduceOp.MAX)
    kernel_ms_max = float(t_dev.item())
    total_accepted = accepted_per_step * args.steps * world  # every rank traces a sweep of the same size; counts differ by <1 %
    acc_t = torch.tensor([accepted_per_step * args.steps, attempts_per_step * args.steps, rows_per_step * args.steps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(acc_t)
    total_accepted, total_attempts, total_rows = [float(v) for v in acc_t.tolist()]
    value = total_accepted / (kernel_ms_max * 1e-3)

    # ---- end to end through the C ABI with host buffers
    e2e_value, h2d, d2h = None, 0, 0
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, 3))
        res = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=outputper, **params)  # warm-up (allocations)
        d2h = res.total_rows * 160 + nrays * (srt.SUMMARY_WORDS * 8 + 4 + 56) + 64
        h2d = nrays * 56 + nrays * 8
        res.close()
        barrier()
        t0 = time.perf_counter()
        e2e_acc = 0
        for _ in range(e2e_steps):
            res = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=outputper, **params)
            e2e_acc += res.accepted_steps
            res.close()
        torch.cuda.synchronize()
        t_e2e = time.perf_counter() - t0
        t_e = torch.tensor([t_e2e], dtype=torch.float64, device=dev)
        a_e = torch.tensor([float(e2e_acc)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
            dist.all_reduce(a_e)
        e2e_value = float(a_e.item()) / float(t_e.item())

    # ---- roofline of the dominant kernel (srt_trace_kernel): FP64 pipe
    peak_fma = srt.fp64_peak(True) / 1e3   # TFLOP/s
    peak_nofma = srt.fp64_peak(False) / 1e3
    rejected = attempts_per_step - accepted_per_step
    flops_per_launch = accepted_per_step * FLOPS_ACCEPTED + rejected * FLOPS_REJECTED
    avg_launch_s = (kernel_ms / args.steps) * 1e-3
    achieved = flops_per_launch / avg_launch_s / 1e12
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    hbm_ach = (rows_per_step * ROW_BYTES + nrays * (56 + srt.SUMMARY_WORDS * 8)) / avg_launch_s / 1e9
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "r1j_traffic.json")))
        if tj.get("workload") == wl and model.kernel_info(nrays)["mode"] == "team8":
            traffic = tj["dram_bytes_per_launch"]  # bytes per launch of this kernel on this workload, from the committed ncu pass
    except Exception:
        pass
    roofline = {"bound": "fp64", "achieved": achieved, "peak": peak_fma, "unit": "TFLOP/s", "frac": achieved / peak_fma, "traffic": traffic,
                "bound_note": "FP64 (non-tensor) pipe: the path is scalar FP64 with no dense contraction and ~1e-4 of HBM bandwidth, so neither 'hbm' nor 'tensor' applies",
                "peak_source": "DFMA micro-kernel measured live (srt_fp64_peak); MEASURED_PEAKS.json has no FP64 entry",
                "peak_no_fma": peak_nofma, "frac_no_fma": achieved / peak_nofma,
                "work_model": "%.2g flop-eq per accepted RK45 step + %.2g per rejected attempt (SURVEY.md 8d)" % (FLOPS_ACCEPTED, FLOPS_REJECTED),
                "hbm": {"achieved": hbm_ach, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_ach / hbm_peak,
                        "of": "measured" if "hbm_gbs" in peaks else "fallback"}}

    if rank == 0:
        cpu = None
        if not args.no_cpu_baseline:
            threads = os.cpu_count() or 1
            nsample = max(threads, 256)  # ~12 s of CPU work on 16 cores
            if WORKLOADS[wl][4] == "grid":
                nsample = threads
                a, t = cpu_reference_run(pos0, dir0, w0, dict(params, maxsteps=300), _GRID_CACHE + ".%d" % local_rank, nsample, threads, "grid", 1)
            else:
                a, t = cpu_reference_run(pos0, dir0, w0, params, cfg, nsample, threads)
            cpu = {"value": a / t, "unit": UNIT, "cores": threads, "kind": "port",
                   "sample": "%d rays strided out of the sweep%s, %.1f s, faithful call structure of the reference" % (nsample, " (first 300 steps of each)" if WORKLOADS[wl][4] == "grid" else "", t)}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": kernel_ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": config_dict(world, wl, nrays),
                "rays_per_sec": float(nrays) * world * args.steps / (kernel_ms_max * 1e-3),
                "attempted_steps_per_sec": total_attempts / (kernel_ms_max * 1e-3),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
                "gpu_launches": int(launches), "clocks": sampler.result(), "roofline": roofline, "cpu_baseline": cpu,
                "kernel": model.kernel_info(nrays)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
