"""Round-2 experiment for the open item "config 3 at full length" (DESIGN.md): which ingredient of an old config-3 ray
makes a round of the team kernel ten times slower than a round of a young one?

    python tools/c3_age.py --maxsteps 200 [--model grid|ngo] [--igrf 0|1] [--alts 1000,4000,9000,15000,21000] [--nlon 10]

Every run traces 100 freq x 100 lat x nlon rays for at most `maxsteps` accepted steps and prints accepted ray-steps/s, the
time per block round (kernel time / the largest round count of a block, counters[7]) and the stop-condition histogram.
What separates the suspects:
  * --alts spreads the launch points over many grid cells from the first step on (a young ray population with an old
    population's cell footprint): if the rate falls as with old rays, the loss is the coefficient-table working set;
  * --model ngo --igrf 1 keeps IGRF and drops the grid, --model grid --igrf 0 the other way round;
  * SRT_TEAM_PHASE_ALIGN=1 / SRT_TEAM_BLOCK=512 in the environment select the scheduling knobs;
  * --lib tools/_variants/stage.so runs the -DSRT_GRID_STAGE build (team-cooperative fetch of the cell record through
    shared memory, srt_tricubic.h); --save / --compare check that knobs and variants return the same bits;
  * --maxsteps 200 / 1000 / 4000 gives the rate against age.
Product API only (no oracle); needs a GPU."""
import argparse, os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch
import stanford_raytracer_b200 as srt, common
from stanford_raytracer_b200 import gridfile

ap = argparse.ArgumentParser()
ap.add_argument("--maxsteps", type=int, default=200)
ap.add_argument("--model", default="grid", choices=["grid", "ngo"])
ap.add_argument("--igrf", type=int, default=1)
ap.add_argument("--alts", default="1000", help="comma-separated launch altitudes in km, cycled over the rays")
ap.add_argument("--nlon", type=int, default=10)
ap.add_argument("--reps", type=int, default=2)
ap.add_argument("--lib", default="", help="an alternative build of libsrt (tools/build_variant.sh -> tools/_variants/NAME.so)")
ap.add_argument("--save", default="", help="write the per-ray summaries (last row, steps, stop code) to this .npy")
ap.add_argument("--compare", default="", help="require the per-ray summaries to equal, bit for bit, those saved in this .npy")
a = ap.parse_args()

R_E = 6371.2e3
if a.lib:
    srt._LIB_PATH = os.path.abspath(a.lib)
srt.init(0)
ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, a.igrf, 0)
if a.model == "grid":
    path = "/tmp/grid80.txt"
    if not os.path.exists(path):
        ext = (-10 * R_E, 10 * R_E) * 3
        f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 80, 80, 80, ext)
        gridfile.write_grid(path, f, qs, ms, ext)
    m = srt.interpsetup(path, 2010001, 0, a.igrf, 0)
else:
    m = ngo
freqs = np.logspace(np.log10(200.0), np.log10(30000.0), 100)
lats = np.linspace(10.0, 60.0, 100)
lons = np.arange(a.nlon) * 36.0
alts = [float(v) for v in a.alts.split(",")]
parts = [common.launch_grid(freqs, lats, lons, alt_km=alt) for alt in alts]
n = len(parts[0][2])
sel = np.arange(n) % len(alts)  # ray i takes the launch point of altitude i mod len(alts)
pos0 = np.choose(sel[:, None], [p[0] for p in parts])
dir0 = np.zeros_like(pos0)
w0 = parts[0][2]
p = dict(common.C2, maxsteps=a.maxsteps)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays0 = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda")
counters = torch.zeros(8, dtype=torch.int64, device="cuda")
for rep in range(a.reps):
    rays = rays0.clone()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0)
    e1.record(); torch.cuda.synchronize()
c = counters.cpu().numpy()
ms_ = e0.elapsed_time(e1)
sc = summary[:, 21].cpu().numpy().astype(int)
r = np.sqrt((summary[:, 1:4].cpu().numpy() ** 2).sum(axis=1)) / R_E
same = ""
if a.save:
    np.save(a.save, summary.cpu().numpy())
if a.compare:
    same = "| summaries identical to %s: %s" % (a.compare, bool(np.array_equal(np.load(a.compare), summary.cpu().numpy(), equal_nan=True)))
knobs = {k: os.environ[k] for k in ("SRT_TEAM_PHASE_ALIGN", "SRT_TEAM_BLOCK", "SRT_KERNEL", "SRT_TEAM_COMPACT") if k in os.environ}
print("c3_age model", a.model, "igrf", a.igrf, "alts", a.alts, "n", n, "maxsteps", a.maxsteps, "knobs", knobs,
      "| ms %.1f" % ms_, "accepted", c[1], "attempts", c[2], "steps/s %.3g" % (c[1] / ms_ * 1e3),
      "| block rounds max", c[7], "us/round %.1f" % (ms_ * 1e3 / max(int(c[7]), 1)),
      "| final radius R_E min/median/max %.2f/%.2f/%.2f" % (r.min(), np.median(r), r.max()),
      "| stopcond", np.bincount(sc, minlength=7).tolist(), m.kernel_info(n), "lib", a.lib or "default", same, flush=True)
