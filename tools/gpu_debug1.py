"""GPU debug: device-path vs host-path counts, kernel timings for both kernel forms."""
import os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
for nf, nl, tmax in ((10, 16, 10.0), (100, 100, 10.0), (100, 1000, 0.5)):
    pos0, dir0, w0 = common.c2_grid(nf, nl)
    n = len(w0)
    p = dict(common.C2, tmax=tmax)
    P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
    rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
    summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
    counters = torch.zeros(8, dtype=torch.int64, device="cuda")
    for mode in ("thread", "team"):
        os.environ["SRT_KERNEL"] = mode
        rays = rays0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
        e1.record(); torch.cuda.synchronize()
        c = counters.cpu().numpy(); s = summary.cpu().numpy()
        print(mode, "n", n, "tmax", tmax, "ms", round(e0.elapsed_time(e1), 1), "counters", c[:5], "sum(nsteps-1)", int(s[:, 20].sum()) - n,
              "steps/s %.3g" % (c[1] / (e0.elapsed_time(e1) * 1e-3)), m.kernel_info(n), flush=True)
    if n <= 10000:
        t0 = time.time()
        res = srt.raytracer_run(m, pos0, dir0.copy(), w0, **p)
        print("host path: accepted", res.accepted_steps, "rows", res.total_rows, "attempts", res.attempts, "s", round(time.time() - t0, 2),
              "nsteps equal to device summary:", np.array_equal(res.nsteps, s[:, 20].astype(np.int64)), flush=True)
