"""BASELINE config 3 shape, bounded: 100k rays through an 80^3 tricubic grid + IGRF, first `maxsteps` steps of every ray.
    python tools/c3_case.py <maxsteps> [thread|team]
Prints set-up times (grid sampling on the GPU, file write, load + coefficient kernel) and accepted ray-steps/s."""
import os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
if len(sys.argv) > 2:
    os.environ["SRT_KERNEL"] = sys.argv[2]
import stanford_raytracer_b200 as srt, common
from stanford_raytracer_b200 import gridfile
maxsteps = int(sys.argv[1])
R_E = 6371.2e3
srt.init(0)
ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
ext = (-10 * R_E, 10 * R_E) * 3
t0 = time.time(); f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 80, 80, 80, ext); t1 = time.time()
path = "/tmp/grid80.txt"
gridfile.write_grid(path, f, qs, ms, ext); t2 = time.time()
m = srt.interpsetup(path, 2010001, 0, 1, 0); t3 = time.time()
print("grid: sample %.1fs (GPU funcPlasmaParams, 512k nodes), write %.1fs, load+FD derivatives+coefficient kernel %.1fs" % (t1 - t0, t2 - t1, t3 - t2), flush=True)
pos0, dir0, w0 = common.c2_grid(100, 100, 10)
n = len(w0)
p = dict(common.C2, maxsteps=maxsteps)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
for rep in range(2):
    rays = rays0.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
    e1.record(); torch.cuda.synchronize()
c = counters.cpu().numpy()
ms_ = e0.elapsed_time(e1)
sc = summary[:, 21].cpu().numpy().astype(int)
print("c3 n", n, "maxsteps", maxsteps, "ms %.1f" % ms_, "accepted", c[1], "attempts", c[2], "steps/s %.3g" % (c[1] / ms_ * 1e3), "stopcond histogram", np.bincount(sc, minlength=7).tolist(), m.kernel_info(n))
