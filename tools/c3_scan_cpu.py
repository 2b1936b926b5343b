"""Finds, on the host cores, config-3 rays of each kind (NaN state, outside the lattice, every stop code, long lives) for tests/test_gpu_c3_long.py:\n    python tools/c3_scan_cpu.py   (needs /tmp/grid80_cpu.txt: an 80^3 grid file sampled from the Ngo model, see bench.py grid_file)\nUses the de-duplicated CPU build of the kernels' physics (oracle/_build/libsrt_dedup.so); every 97th ray of the 100 000."""
import sys, numpy as np, threading, time
sys.path.insert(0,'tests'); sys.path.insert(0,'.')
import common, oracle_lib as ol
R_E=6371.2e3
m = ol.DedupModel("grid", "/tmp/grid80_cpu.txt", use_igrf=1)
pos0, dir0, w0 = common.c2_grid(100, 100, 10)
idx = np.arange(0, 100000, 97)
P = dict(common.C2)
out = {}
lock = threading.Lock(); head=[0]
def work():
    while True:
        with lock:
            j = head[0]; head[0]+=1
        if j >= len(idx): return
        r = int(idx[j]); s = m.trace_summary(pos0[r], dir0[r], w0[r], **P); out[r] = (s["stopcond"], s["nrows"], s["last_row"][1:4].copy(), s["nrejected"])
t=time.time()
ts=[threading.Thread(target=work) for _ in range(8)]; [x.start() for x in ts]; [x.join() for x in ts]
print("scan", time.time()-t, "s")
nan=[r for r,v in out.items() if np.isnan(v[2]).any()]
outside=[r for r,v in out.items() if not np.isnan(v[2]).any() and (np.abs(v[2])>10*R_E).any()]
far=[r for r,v in out.items() if not np.isnan(v[2]).any() and np.linalg.norm(v[2])>6*R_E]
by={}
for r,v in out.items(): by.setdefault(v[0],[]).append((v[1],r))
print("nan rays", sorted(nan)); print("outside lattice", sorted(outside)); print("far (>6 R_E)", sorted(far)[:10])
for sc in sorted(by): print("stop",sc,len(by[sc]), sorted(by[sc])[:5], sorted(by[sc])[-3:])
np.save("/tmp/c3scan.npy", np.array([(r,v[0],v[1]) for r,v in sorted(out.items())]))
