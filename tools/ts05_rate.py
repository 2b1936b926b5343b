"""Cost of the TS05 external field on the GPU (use_tsyganenko = 1): funcPlasmaParams per second through srt_model_eval and
accepted ray-steps per second of both trace kernels, with the term off and on.   python tools/ts05_rate.py [lib.so]"""
import os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
if len(sys.argv) > 1:
    srt._LIB_PATH = os.path.abspath(sys.argv[1])
srt.init(0)
PARMOD = [3.0, -45.0, 2.5, -4.0, 0.4, 0.7, 0.3, 0.9, 0.6, 1.1]
rng = np.random.default_rng(1)
n = 1 << 20
r = 6371.2e3 * rng.uniform(1.1, 7.0, n); th = rng.uniform(0.2, np.pi - 0.2, n); ph = rng.uniform(-np.pi, np.pi, n)
x = np.ascontiguousarray(np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), r * np.cos(th)], axis=1))
pos0, dir0, w0 = common.c2_grid(100, 100)
nr = len(w0)
p = dict(common.C2, maxsteps=300)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((nr, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
for on in (0, 1):
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    if on and hasattr(m, "tsyganenko"):
        try:
            m.tsyganenko(*PARMOD[:4], W=PARMOD[4:])
        except Exception as e:
            print("tsyganenko unavailable in this build:", e); continue
    m.funcPlasmaParams(x[:1000])
    t0 = time.perf_counter(); m.funcPlasmaParams(x); dt = time.perf_counter() - t0
    for mode in ("thread", "team"):
        os.environ["SRT_KERNEL"] = mode
        best = 1e30
        for rep in range(2):
            rays = rays0.clone()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            srt.trace_batch_device(m, nr, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        c = counters.cpu().numpy()
        ki = m.kernel_info(nr)
        print("TS05 %s, %s kernel: %d rays x <=300 steps: %.1f ms, %d accepted steps, %d attempts, %.3g steps/s (regs %d, stack %d B, block %d, %d blocks/SM)"
              % ("on " if on else "off", mode, nr, best, c[1], c[2], c[1] / (best * 1e-3), ki["regs"], ki["local_bytes"], ki["block"], ki["blocks_per_sm"]), flush=True)
    print("TS05 %s: funcPlasmaParams %.3g points/s through srt_model_eval (host buffers, %d points)" % ("on " if on else "off", n / dt, n), flush=True)
