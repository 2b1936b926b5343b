"""One short launch of a trace kernel for ncu: python tools/profile_case.py <thread|team> <nf> <nlat> <maxsteps>"""
import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
os.environ["SRT_KERNEL"] = sys.argv[1]
import stanford_raytracer_b200 as srt, common
nf, nl, maxsteps = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
pos0, dir0, w0 = common.c2_grid(nf, nl)
n = len(w0)
p = dict(common.C2, maxsteps=maxsteps)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
log = torch.empty((n * (maxsteps + 1), 21), dtype=torch.float64, device="cuda")
for rep in range(2):
    rays = rays0.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    srt.trace_batch_device(m, n, rays.data_ptr(), P, log.data_ptr(), log.shape[0], summary.data_ptr(), counters.data_ptr(), 0)
    e1.record(); torch.cuda.synchronize()
c = counters.cpu().numpy()
print(sys.argv[1], "n", n, "ms", e0.elapsed_time(e1), "accepted", c[1], "attempts", c[2], "steps/s %.3g" % (c[1] / e0.elapsed_time(e1) * 1e3), m.kernel_info(n))
