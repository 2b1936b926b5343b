"""Config-2 tail analysis: step histogram per ray, rounds per block, effect of queue order / re-packing."""
import os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
nf, nl = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (100, 100)
pos0, dir0, w0 = common.c2_grid(nf, nl)
n = len(w0)
p = dict(common.C2)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
first = True
for order, compact, blk in [(1, 1, 0), (0, 1, 0), (1, 0, 0), (0, 0, 0), (1, 1, 256), (1, 1, 128), (1, 1, 1024)]:
    os.environ["SRT_RAY_ORDER"] = str(order); os.environ["SRT_TEAM_COMPACT"] = str(compact)
    if blk: os.environ["SRT_TEAM_BLOCK"] = str(blk)
    else: os.environ.pop("SRT_TEAM_BLOCK", None)
    best = 1e30
    for rep in range(2 if first else 1):
        rays = rays0.clone()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    c = counters.cpu().numpy(); ki = m.kernel_info(n)
    print("order=%d compact=%d block=%d grid=%d ms=%.1f accepted=%d attempts=%d steps/s=%.3g repacks=%d rounds avg=%.0f max=%d us/round=%.2f" % (
        order, compact, ki["block"], ki["grid"], best, c[1], c[2], c[1] / best * 1e3, c[5], c[6] / ki["grid"], c[7], best * 1e3 / max(c[7], 1)), flush=True)
    if first:
        s = summary.cpu().numpy(); ns = s[:, 20]; att = ns + s[:, 22]
        print("steps per ray: percentiles 10/50/90/99/100 =", np.percentile(ns, [10, 50, 90, 99, 100]), " attempts max", att.max(), " sum", ns.sum())
        h, edges = np.histogram(att, bins=[0, 500, 1000, 2000, 4000, 8000, 16000, 32000, 64000, 1e9]); print("attempt histogram", list(zip(edges[:-1].astype(int), h)))
        print("stopcond", np.bincount(s[:, 21].astype(int), minlength=7))
        idx = np.argsort(-att)[:3000]; print("long rays: index spread min/median/max", idx.min(), np.median(idx), idx.max(), " frac of total attempts in top 3000:", att[idx].sum() / att.sum())
        first = False
