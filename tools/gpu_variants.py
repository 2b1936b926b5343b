"""Times the trace kernels of alternative libsrt builds (tools/_variants/*.so) on one workload."""
import ctypes as C, glob, os, sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
nf, nl, maxsteps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
pos0, dir0, w0 = common.c2_grid(nf, nl)
n = len(w0)
p = dict(common.C2, maxsteps=maxsteps)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
libs = [os.path.join("stanford_raytracer_b200", "libsrt.so")] + sorted(glob.glob("tools/_variants/*.so"))
if os.environ.get("SRT_VARIANT_ONLY"):
    libs = [p for p in libs if os.environ["SRT_VARIANT_ONLY"] in p]
for path in libs:
    srt._lib = None
    srt._LIB_PATH = os.path.abspath(path)
    srt.init(0)
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    for mode in (sys.argv[4].split(",") if len(sys.argv) > 4 else ("thread", "team")):
        os.environ["SRT_KERNEL"] = mode
        best = 1e30
        for rep in range(2):
            rays = rays0.clone()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
            e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        c = counters.cpu().numpy()
        ki = m.kernel_info(n)
        print("%-28s %-6s n=%d ms=%9.1f accepted=%d steps/s=%.3g attempts=%d shadow_attempts=%d rounds_max=%d regs=%d local=%d blocks/SM=%d" % (os.path.basename(path), mode, n, best, c[1], c[1] / (best * 1e-3), c[2], c[6], c[7], ki["regs"], ki["local_bytes"], ki["blocks_per_sm"]), flush=True)
