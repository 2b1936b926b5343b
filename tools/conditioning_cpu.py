#!/usr/bin/env python
"""1-ulp-perturbation conditioning of adaptive rays (parity gates): how far the endpoint of a config-2 ray moves
when its launch position moves by one unit in the last place -- the size of the difference ANY second implementation that
does not share this one's arithmetic bit for bit must expect (the GPU kernels do share it, and agree bit for bit).
    python tools/conditioning_cpu.py [nrays]
Host cores, the de-duplicated CPU build of the kernels' physics (oracle/_build/libsrt_dedup.so)."""
import os, sys, threading
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, common, oracle_lib as ol

n = int(sys.argv[1]) if len(sys.argv) > 1 else 128
m = ol.DedupModel("ngo", common.NGO_CFG)
pos0, dir0, w0 = common.c2_grid(100, 100)
idx = np.unique(np.linspace(0, len(w0) - 1, n).astype(int))
res = {}
lock, head = threading.Lock(), [0]

def work():
    while True:
        with lock:
            j = head[0]; head[0] += 1
        if j >= len(idx): return
        r = int(idx[j])
        a = m.trace_summary(pos0[r], dir0[r], w0[r], **common.C2)
        p = pos0[r].copy(); p[0] = np.nextafter(p[0], np.inf)
        b = m.trace_summary(p, dir0[r], w0[r], **common.C2)
        res[r] = (a, b)
ts = [threading.Thread(target=work) for _ in range(os.cpu_count() or 4)]
[t.start() for t in ts]; [t.join() for t in ts]
same_n = same_stop = 0; dpos, dt, dnn = [], [], []
for r, (a, b) in sorted(res.items()):
    same_n += a["nrows"] == b["nrows"]; same_stop += a["stopcond"] == b["stopcond"]
    la, lb = a["last_row"], b["last_row"]
    if np.isfinite(la).all() and np.isfinite(lb).all():
        dpos.append(np.linalg.norm(la[1:4] - lb[1:4]) / np.linalg.norm(la[1:4]))
        dt.append(abs(la[0] - lb[0]) / max(abs(la[0]), 1e-300))
        dnn.append(np.linalg.norm(la[10:13] - lb[10:13]) / max(np.linalg.norm(la[10:13]), 1e-300))
q = lambda v: "median %.1e, 90 %% %.1e, max %.1e" % (np.median(v), np.percentile(v, 90), np.max(v))
print("config 2, %d rays, launch x moved by 1 ulp (1.2e-16 relative):" % len(res))
print("  same number of rows: %d / %d;  same stop code: %d / %d" % (same_n, len(res), same_stop, len(res)))
print("  relative change of the end position:  " + q(dpos))
print("  relative change of the end time:      " + q(dt))
print("  relative change of the end wave normal:" + q(dnn))
