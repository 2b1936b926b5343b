"""Run by tests/test_gpu_zz_phase_align.py in a child process (so that a scheduling bug cannot hang the test session):
the team kernel with SRT_TEAM_PHASE_ALIGN=1 (idle teams refill only on the block's attempt boundary, srt_team.cuh)
must return the same bits as with the knob off, on batches of several machine fills, adaptive and fixed step."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)
import common  # noqa: E402
import stanford_raytracer_b200 as srt  # noqa: E402

os.environ["SRT_KERNEL"] = "team"
srt.init(0)
ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
pos0, dir0, w0 = common.c2_grid(250, 200)  # 50 000 rays: 2.6 fills of 148 x 128 team slots
cases = {"adaptive": dict(common.C2, maxsteps=48, tmax=0.3), "fixed": dict(common.C1, maxsteps=40, tmax=0.02)}
for name, p in cases.items():
    out = {}
    for knob in ("0", "1"):
        os.environ["SRT_TEAM_PHASE_ALIGN"] = knob
        res = srt.raytracer_run(ngo, pos0, np.zeros_like(pos0), w0, outputper=16, **p)
        out[knob] = (res.last_row.copy(), res.nsteps.copy(), res.stopcond.copy(), res.nrejected.copy(), res.accepted_steps,
                     [res[r].rows.copy() for r in (0, 777, 24999, 49999)])
    a, b = out["0"], out["1"]
    assert np.array_equal(a[0], b[0], equal_nan=True), name + ": last rows differ"
    for i in (1, 2, 3):
        assert np.array_equal(a[i], b[i]), name + ": per-ray counters differ"
    assert a[4] == b[4] and a[4] > 0, name
    for x, y in zip(a[5], b[5]):
        assert np.array_equal(x, y, equal_nan=True), name + ": logged rows differ"
    print("phase_align %s: %d rays, %d accepted steps, identical" % (name, len(w0), a[4]), flush=True)
print("PHASE_ALIGN_OK")
