#!/usr/bin/env python
"""Generates tests/golden/: frozen outputs of the CPU oracle on small seeded inputs.

    python tools/make_golden.py

The reference is Fortran and no Fortran compiler exists in this environment, so these vectors
come from the oracle itself (PARITY UNPINNED: they pin the oracle -- and through the GPU parity
tests the kernels -- against regressions, not against the Fortran).  Re-run only when the oracle's
arithmetic is changed on purpose; commit the result.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import common  # noqa: E402
import oracle_lib as ol  # noqa: E402
from stanford_raytracer_b200 import gridfile  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
R_E = 6371.2e3


def main():
    os.makedirs(GOLD, exist_ok=True)
    ol.build()
    rng = np.random.default_rng(2010001)
    meta = {"generator": "tools/make_golden.py", "source": "CPU oracle (oracle/), srt_math.h elementary functions", "ngo_rays": [], "grid_rays": []}
    ongo = ol.OracleModel("ngo", common.NGO_CFG)
    x = np.array([ol.launch_point(la, lo, alt) for la, lo, alt in zip(rng.uniform(-80, 80, 256), rng.uniform(0, 360, 256), rng.uniform(200, 30000, 256))])
    x[:32, 1] = 0.0
    out = {"plasma_x": x, "plasma_out": ongo.plasma(x)}
    cases = [(2000.0, 30.0, dict(common.C1, maxsteps=200)), (500.0, 20.0, dict(common.C2, maxsteps=150)), (12000.0, 50.0, dict(common.C2, maxsteps=150))]
    for i, (f, lat, params) in enumerate(cases):
        pos0, _, w0 = common.launch_grid([f], [lat])
        o = ongo.trace(pos0[0], [0, 0, 0], w0[0], **params)
        out["rows_%d" % i] = o["rows"]
        meta["ngo_rays"].append({"pos0": pos0[0].tolist(), "w0": float(w0[0]), "params": params, "stopcond": o["stopcond"], "nrows": len(o["rows"])})
    np.savez_compressed(os.path.join(GOLD, "golden_ngo.npz"), **out)

    # small gridded model: the Ngo model sampled on a 13^3 lattice over +-4 R_E (the grid FILE is the fixture)
    ext = (-4 * R_E, 4 * R_E, -4 * R_E, 4 * R_E, -4 * R_E, 4 * R_E)
    f, qs, ms = gridfile.sample_grid(ongo.plasma, 13, 13, 13, ext)
    gpath = os.path.join(GOLD, "grid13.txt")
    gridfile.write_grid(gpath, f, qs, ms, ext)
    og = ol.OracleModel("grid", gpath, use_igrf=1)
    xg = rng.uniform(-4.4 * R_E, 4.4 * R_E, (512, 3))
    outg = {"plasma_x": xg, "plasma_out": og.plasma(xg)}
    for i, (fr, lat, lon) in enumerate([(1000.0, 25.0, 0.0), (5000.0, 45.0, 72.0)]):
        pos0, _, w0 = common.launch_grid([fr], [lat], [lon])
        params = dict(common.C2, maxsteps=80)
        o = og.trace(pos0[0], [0, 0, 0], w0[0], **params)
        outg["rows_%d" % i] = o["rows"]
        meta["grid_rays"].append({"pos0": pos0[0].tolist(), "w0": float(w0[0]), "params": params, "stopcond": o["stopcond"], "nrows": len(o["rows"])})
    np.savez_compressed(os.path.join(GOLD, "golden_grid.npz"), **outg)
    json.dump(meta, open(os.path.join(GOLD, "golden.json"), "w"), indent=1)
    print("wrote", os.listdir(GOLD))


if __name__ == "__main__":
    main()
