"""How many rays of a BASELINE workload turn NaN and are then stepped to maxsteps (the reference's controller accepts
every step of a NaN state and no stop condition catches it)?  Host build of the device physics, strided sample.

    python tools/nan_rays_cpu.py c2|c3 [nsample=250] [threads=8]      (c3 needs /tmp/c3cpu/grid80.txt: tools/c3_attempts_cpu.py)
"""
import sys
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, common

which = sys.argv[1] if len(sys.argv) > 1 else "c2"
ns = int(sys.argv[2]) if len(sys.argv) > 2 else 250
threads = int(sys.argv[3]) if len(sys.argv) > 3 else 8
if which == "c3":
    emu = common.HostEmu("grid", "/tmp/c3cpu/grid80.txt", use_igrf=1, precomputed_coef=True)
    pos0, dir0, w0 = common.c2_grid(100, 100, 10)
else:
    emu = common.HostEmu("ngo", common.NGO_CFG)
    pos0, dir0, w0 = common.c2_grid(100, 100)
n = len(w0)
idx = (np.arange(ns) * (n // ns) + (n // ns) // 2 + 7 * np.arange(ns)) % n


def one(i):
    r = emu.trace(pos0[i], [0, 0, 0], w0[i], **common.C2)
    bad = ~np.isfinite(r["rows"]).all(axis=1)
    return len(r["rows"]), int(bad.sum()), r["stopcond"], r["nrejected"], r["ncone"], (int(np.argmax(bad)) if bad.any() else -1)


with ThreadPoolExecutor(threads) as ex:
    res = np.array(list(ex.map(one, idx)))
print(which, "rays", ns, "with a non-finite row:", int((res[:, 1] > 0).sum()), "| non-finite rows", res[:, 1].sum(), "of", res[:, 0].sum())
for sc in range(7):
    m = res[:, 2] == sc
    if m.any():
        print("stop", sc, "rays", int(m.sum()), "mean rows %.0f" % res[m, 0].mean(), "mean rejected %.0f" % res[m, 3].mean(),
              "mean cone retries %.0f" % res[m, 4].mean(), "rays with non-finite rows", int((res[m, 1] > 0).sum()))
b = res[res[:, 1] > 0]
if len(b):
    print("first non-finite row of those rays:", b[:, 5].tolist())
