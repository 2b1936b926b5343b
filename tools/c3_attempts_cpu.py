"""Config 3 on the host cores: attempts per ray over whole lives, for a strided sample of the 100 000 rays.

    python tools/c3_attempts_cpu.py [nsample=128] [threads=8]

Why: the one full-length GPU sweep of config 3 took 120.7 s (profiles/r1z_bench_c3_counting_pass.txt).  Whether that is
a scheduling bound (a few very long rays, as in config 2) or a per-round slowdown can be told from the distribution of
attempts per ray alone: a ray is a serial chain of 6 block rounds per attempt, so the sweep cannot be shorter than
(longest ray's attempts) x 6 x (time of one round), and with the 18 944 team slots of one B200 it cannot be shorter
than (sum of attempts) x 6 / 18 944 rounds either.  Uses the host build of the device physics (tests/csrc/hostemu.cpp,
bit-identical to the kernels); the grid is the bench's 80^3 Ngo-sampled grid, sampled here by the same host build."""
import os, sys, time
from concurrent.futures import ThreadPoolExecutor
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import common
from stanford_raytracer_b200 import gridfile

nsample = int(sys.argv[1]) if len(sys.argv) > 1 else 128
threads = int(sys.argv[2]) if len(sys.argv) > 2 else 8
R_E = 6371.2e3
path = "/tmp/c3cpu/grid80.txt"
t0 = time.time()
if not os.path.exists(path):
    os.makedirs(os.path.dirname(path), exist_ok=True)
    ngo = common.HostEmu("ngo", common.NGO_CFG)
    ext = (-10 * R_E, 10 * R_E) * 3
    f, qs, ms = gridfile.sample_grid(ngo.plasma, 80, 80, 80, ext)
    gridfile.write_grid(path, f, qs, ms, ext)
print("grid file %.1f s" % (time.time() - t0), flush=True)
t0 = time.time()
emu = common.HostEmu("grid", path, use_igrf=1, precomputed_coef=True)  # noqa
print("model set-up %.1f s" % (time.time() - t0), flush=True)
pos0, dir0, w0 = common.c2_grid(100, 100, 10)
n = len(w0)
idx = (np.arange(nsample) * (n // nsample) + (n // nsample) // 2) % n


def one(i):
    r = emu.trace(pos0[i], [0, 0, 0], w0[i], **common.C2)
    acc = len(r["rows"]) - 1
    return acc, acc + r["nrejected"] + r["ncone"], r["stopcond"], float(np.sqrt((r["rows"][:, 1:4] ** 2).sum(axis=1)).max() / R_E)


t0 = time.time()
with ThreadPoolExecutor(threads) as ex:
    res = list(ex.map(one, idx))
dt = time.time() - t0
acc = np.array([r[0] for r in res]); att = np.array([r[1] for r in res]); sc = np.array([r[2] for r in res]); rmax = np.array([r[3] for r in res])
print("c3 sample: %d of %d rays, %d threads, %.1f s, %.0f accepted steps/s on the host" % (nsample, n, threads, dt, acc.sum() / dt))
print("accepted per ray: mean %.0f median %.0f max %d | attempts per ray: mean %.0f median %.0f p90 %.0f p99 %.0f max %d"
      % (acc.mean(), np.median(acc), acc.max(), att.mean(), np.median(att), np.percentile(att, 90), np.percentile(att, 99), att.max()))
print("attempt histogram (edges 0,500,2000,8000,16000,24000,32000,inf):", np.histogram(att, [0, 500, 2000, 8000, 16000, 24000, 32000, 10 ** 9])[0].tolist())
print("stop codes:", np.bincount(sc, minlength=7).tolist(), "| largest radius reached, R_E: median %.2f max %.2f" % (np.median(rmax), rmax.max()))
scale = n / nsample
print("scaled to %d rays: %.0f M attempts; rounds if perfectly packed on 18944 team slots: %.0f k; rounds of the longest sampled ray: %.0f k"
      % (n, att.sum() * scale / 1e6, att.sum() * scale * 6 / 18944 / 1e3, att.max() * 6 / 1e3))
# list scheduling of the whole queue on the team slots (every ray of the sample stands for n/nsample rays, queue in the
# kernel's golden-ratio order = effectively shuffled): the sweep's length in attempts if every round took the same time
import heapq
rng = np.random.default_rng(0)
jobs = np.repeat(att, int(round(scale)))
rng.shuffle(jobs)
for slots in (18944, 9472):
    h = [0] * slots
    heapq.heapify(h)
    for j in jobs:
        heapq.heappush(h, heapq.heappop(h) + int(j) + 1)
    print("list schedule on %d team slots: makespan %d attempts = %.0f k rounds; queue empty after %.0f k rounds"
          % (slots, max(h), max(h) * 6 / 1e3, min(h) * 6 / 1e3))
# where the long rays are launched: mean attempts (thousands) by frequency decile (rows, 0.2 -> 30 kHz) and latitude
# quintile (columns, 10 -> 60 deg); launch order is frequency slowest, then latitude, then longitude (common.launch_grid)
fi, li = idx // 1000, (idx // 10) % 100
tab = np.zeros((10, 5)); cnt = np.zeros((10, 5))
np.add.at(tab, (fi // 10, li // 20), att); np.add.at(cnt, (fi // 10, li // 20), 1)
print("mean attempts/1000 by frequency decile x latitude quintile:")
for row in tab / np.maximum(cnt, 1) / 1e3:
    print("   " + " ".join("%5.1f" % v for v in row))
# the same list schedule with the queue sorted by that table (longest expected first), i.e. what a launch-parameter
# predictor of a ray's length would buy
exp_len = (tab / np.maximum(cnt, 1))[fi // 10, li // 20]
order = np.argsort(-np.repeat(exp_len, int(round(scale))), kind="stable")
h = [0] * 18944
heapq.heapify(h)
for j in np.repeat(att, int(round(scale)))[order]:
    heapq.heappush(h, heapq.heappop(h) + int(j) + 1)
print("list schedule on 18944 slots, queue sorted by the table's expected length: makespan %d attempts = %.0f k rounds" % (max(h), max(h) * 6 / 1e3))
