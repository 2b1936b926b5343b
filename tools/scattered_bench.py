"""Auxiliary figure for the scattered-sample model (modelnum 4): one JSON line on stdout.

bench.py runs this as a CHILD PROCESS with a time limit (`extra.scattered` of the bench line): the model was written after the
round's GPU minutes were spent, so its first run on a B200 must not be able to cost the bench line -- a fault, a hang or a
wrong result here ends in {"error": ...} or in its own mismatch counters, never in the parent's.

What it does (GPU 0): writes a sample file (40 000 positions uniform in a shell 1.05-5 R_E, log-densities of the Ngo model taken from the
product's own funcPlasmaParams kernel), sets the model up through srt_model_scattered, then
  * times funcPlasmaParams (srt_model_eval: host buffers, copies inside) on 200 000 points;
  * times the trace kernel on the C2 launch set, first 100 steps of every ray, rays resident on the device;
  * compares 400 points and 16 whole rays with the CPU oracle (kd-tree + lsinterp restated; handed the product's neighbour
    order, see tests/test_zy_gpu_scattered.py) bit for bit, and times the oracle on 16 rays (up to 8 host cores).
"""
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
R_E = 6371.2e3
OPTS = dict(window_scale=2.0, order=2, exact=0, local_window_scale=3.0)


def write_samples(srt, path, n, seed=7):
    import common
    rng = np.random.default_rng(seed)
    r = R_E * (1.05 ** 3 + (5.0 ** 3 - 1.05 ** 3) * rng.random(n)) ** (1.0 / 3.0)  # uniform in volume
    u = rng.normal(size=(n, 3))
    u /= np.linalg.norm(u, axis=1)[:, None]
    pos = r[:, None] * u
    ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    raw = ngo.funcPlasmaParams(pos)["raw"]
    logn = np.log(np.maximum(raw[:, 4:8], 1e-30))
    with open(path, "w") as f:
        f.write("4 %.17e %.17e %.17e %.17e %.17e %.17e\n" % tuple(np.ravel([pos.min(0), pos.max(0)], order="F")))
        f.write(" ".join("%.17e" % v for v in raw[0, 0:4]) + "\n")
        f.write(" ".join("%.17e" % v for v in raw[0, 8:12]) + "\n")
        np.savetxt(f, np.hstack([pos, logn]), fmt="%.17e")
    ngo.close()


def main():
    import torch
    import common
    import stanford_raytracer_b200 as srt
    srt.init(0)
    nsamples = 40000  # about 100 neighbours per query at window_scale 2
    out = {"metric": "accepted_ray_steps_per_sec", "unit": "ray-steps/s"}
    tmp = tempfile.mkdtemp(prefix="srt_scat_")
    path = os.path.join(tmp, "samples.txt")
    write_samples(srt, path, nsamples)
    t0 = time.perf_counter()
    model = srt.scatteredinterpsetup(path, OPTS["window_scale"], OPTS["order"], OPTS["exact"], OPTS["local_window_scale"], yearday=2010001)
    out["setup_s"] = time.perf_counter() - t0
    out["model"] = {"samples": model.npoints, "maxnearest_m": model.maxnearest, "radius_m": model.radius, "monomials": model.nmono, "cells": list(model.cells)}
    # ---- funcPlasmaParams
    rng = np.random.default_rng(1)
    u = rng.normal(size=(200000, 3))
    u /= np.linalg.norm(u, axis=1)[:, None]
    x = np.ascontiguousarray(R_E * (1.2 + 3.3 * rng.random(len(u)))[:, None] * u)
    model.funcPlasmaParams(x[:1000])
    t0 = time.perf_counter()
    raw = model.funcPlasmaParams(x)["raw"]
    dt = time.perf_counter() - t0
    out["plasma_evals_per_sec_host_buffers"] = len(x) / dt
    out["plasma_density_not_interpolated"] = int((raw[:, 4] == 1.0).sum())  # status 1 / 2 of lsinterp leaves exp(0)
    # ---- rays: the C2 launch set, first 100 steps
    freqs = np.logspace(np.log10(200.0), np.log10(30000.0), 100)
    pos0, dir0, w0 = common.launch_grid(freqs, np.linspace(10.0, 60.0, 100), (0.0,))
    params = dict(common.C2, maxsteps=100)
    n = len(w0)
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
    out.update({"value": float(c[1]) / sec, "ms_per_step": sec * 1e3, "steps": 2, "warmup": 1, "accepted_steps_per_sweep": int(c[1]),
                "attempts_per_sweep": int(c[2]), "gpu_launches": int(srt.lib().srt_launch_count() - l0), "kernel": model.kernel_info(n),
                "config": {"workload": "C2 launch set (10k rays) through modelnum 4: %d scattered samples of the Ngo model in a shell 1.05-5 R_E, "
                                       "moving least squares of order 2 (10 monomials), dipole field, first 100 steps of every ray" % nsamples,
                           "rays": n, "maxsteps": 100}})
    # ---- parity and CPU rate (own counters: this figure never gates the bench line)
    import oracle_lib as ol
    om = ol.OracleScattered(path, **OPTS).impose_order(srt.scattered_samples(model)[:, :3])
    pts = x[:400]
    out["points_checked"] = len(pts)
    want = om.plasma(pts)
    out["points_differing"] = int((~(np.equal(raw[:400], want) | (np.isnan(raw[:400]) & np.isnan(want))).all(axis=1)).sum())
    threads = min(os.cpu_count() or 1, 8)  # (every worker builds its own kd-tree: about a second each)
    idx = np.unique(np.linspace(0, n - 1, 16).astype(int))
    order = srt.scattered_samples(model)[:, :3]
    oms = [om] + [ol.OracleScattered(path, **OPTS).impose_order(order) for _ in range(threads - 1)]
    last, nrows, lock, head = np.zeros((len(idx), 20)), np.zeros(len(idx), dtype=np.int64), threading.Lock(), [0]

    def work(t):
        while True:
            with lock:
                j = head[0]
                head[0] += 1
            if j >= len(idx):
                return
            r = int(idx[j])
            sm = oms[t].trace_summary(pos0[r], dir0[r], w0[r], **params)
            last[j], nrows[j] = sm["last_row"], sm["nrows"]

    t0 = time.perf_counter()
    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(oms))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    dt = time.perf_counter() - t0
    g = summary.cpu().numpy()[idx]
    out["rays_checked"] = len(idx)
    out["rays_differing"] = int(sum(0 if (np.array_equal(g[j, :20], last[j], equal_nan=True) and int(g[j, 20]) == int(nrows[j])) else 1 for j in range(len(idx))))
    out["cpu_baseline"] = {"value": float((nrows - 1).sum()) / dt, "unit": "ray-steps/s", "cores": len(oms), "kind": "port",
                           "sample": "%d rays strided out of the 10k, first 100 steps, %.1f s: the oracle's kd-tree + lsinterp + integrator" % (len(idx), dt)}
    print(json.dumps(out))


if __name__ == "__main__":
    try:
        main()
    except Exception as e:  # the parent records it
        print(json.dumps({"error": repr(e)}))
