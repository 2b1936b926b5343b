"""End-to-end timing of the `raytracer` command line on a slice of config 2 (text output included)."""
import os, subprocess, sys, tempfile, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, common
nrays = int(sys.argv[1]) if len(sys.argv) > 1 else 200
pos0, dir0, w0 = common.c2_grid(100, 100)
sel = np.linspace(0, len(w0) - 1, nrays).astype(int)
d = tempfile.mkdtemp()
rays, out = os.path.join(d, "rays.txt"), os.path.join(d, "out.ray")
with open(rays, "w") as f:
    for i in sel:
        f.write(" ".join(repr(float(v)) for v in list(pos0[i]) + [0.0, 0.0, 0.0, w0[i]]) + "\n")
p = common.C2
args = ["stanford_raytracer_b200/raytracer", "--dt0=%r" % p["dt0"], "--dtmax=%r" % p["dtmax"], "--tmax=%r" % p["tmax"], "--root=2", "--fixedstep=0",
        "--maxerr=%r" % p["maxerr"], "--maxsteps=%d" % p["maxsteps"], "--minalt=%r" % p["minalt"], "--inputraysfile=" + rays, "--outputfile=" + out,
        "--modelnum=1", "--ngo_configfile=" + common.NGO_CFG, "--yearday=2010001", "--milliseconds_day=0", "--use_igrf=0", "--use_tsyganenko=0", "--verbose=1"]
t = time.time(); r = subprocess.run(args, capture_output=True, text=True); dt = time.time() - t
print(r.stderr.strip()); print("wall %.2f s, output %.1f MB, rc %d, host cores %d" % (dt, os.path.getsize(out) / 1e6, r.returncode, os.cpu_count()))
os.remove(out)
