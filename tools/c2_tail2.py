"""Per-round time of the team kernel on subsets of config 2's long-lived rays (latency regime)."""
import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
pos0, dir0, w0 = common.c2_grid(100, 100)
n = len(w0)
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
def run(sel, maxsteps, label):
    p = dict(common.C2, maxsteps=maxsteps)
    P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
    k = len(sel)
    rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0[sel].T, dir0[sel].T, w0[sel][None, :]], axis=0))).cuda()
    summary = torch.zeros((k, 24), dtype=torch.float64, device="cuda")
    counters = torch.zeros(8, dtype=torch.int64, device="cuda")
    best = 1e30
    for rep in range(2):
        rays = rays0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        srt.trace_batch_device(m, k, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    c = counters.cpu().numpy(); ki = m.kernel_info(k)
    print("%-34s rays=%5d block=%4d grid=%3d ms=%8.1f accepted=%9d attempts=%9d steps/s=%.3g repacks=%d rounds avg=%.0f max=%d us/round=%.2f" % (
        label, k, ki["block"], ki["grid"], best, c[1], c[2], c[1] / best * 1e3, c[5], c[6] / ki["grid"], c[7], best * 1e3 / max(c[7], 1)), flush=True)
    return summary.cpu().numpy()
s = run(np.arange(n), 1500, "all rays, 1500 steps")
longr = np.nonzero(s[:, 20] >= 1500)[0]
print("long rays:", len(longr))
rng = np.random.default_rng(0)
for k in (148 * 4, 148 * 8, 148 * 16, len(longr)):
    sel = np.sort(rng.choice(longr, size=min(k, len(longr)), replace=False))
    for blk in (0, 128):
        if blk: os.environ["SRT_TEAM_BLOCK"] = str(blk)
        else: os.environ.pop("SRT_TEAM_BLOCK", None)
        run(sel, 3000, "long subset")
os.environ.pop("SRT_TEAM_BLOCK", None)
short = np.nonzero(s[:, 20] < 1500)[0]
mix = np.sort(np.concatenate([longr, short[: 9472 - len(longr)]]))
run(mix, 3000, "long + short filling the machine")
os.environ["SRT_TEAM_COMPACT"] = "0"
run(mix, 3000, "same, no re-packing")
