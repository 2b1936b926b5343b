"""Reference behaviour the kernels must reproduce and a planned fast path relies on (DESIGN.md, config-2 / config-3
bounds): a ray whose wave vector turns NaN (here ray 2934 of the config-2 sweep, 868 Hz, at a resonance-cone point where
dt has collapsed to 3e-14 s) is NOT stopped -- raytracer_stopconditions (raytracer.f95:324-353) catches no NaN, the
step controller accepts every step of a NaN state (`err > maxerr` and the cone test `imsum > 0` are both false,
:786-812, :891) -- so it is stepped to `maxsteps` with an unchanged dt, every row being (t + dt, NaN, ..., NaN)."""
import numpy as np

import common
import oracle_lib as ol


def test_nan_ray_is_stepped_to_maxsteps_with_constant_dt(built):
    pos0, dir0, w0 = common.c2_grid(100, 100)
    r = 2934
    p = dict(common.C2, maxsteps=3000)
    o = ol.OracleModel("ngo", common.NGO_CFG).trace(pos0[r], [0, 0, 0], w0[r], **p)
    e = common.HostEmu("ngo", common.NGO_CFG).trace(pos0[r], [0, 0, 0], w0[r], **p)
    assert np.array_equal(o["rows"], e["rows"], equal_nan=True) and o["stopcond"] == e["stopcond"] == 6
    rows = o["rows"]
    assert len(rows) == 3000                                  # maxsteps rows (Q15)
    bad = ~np.isfinite(rows).all(axis=1)
    k = int(np.argmax(bad))
    assert 500 < k < 700 and bad[k:].all() and not bad[:k].any()
    # first the wave vector goes (position still finite), one step later the whole state
    assert np.isfinite(rows[k, 0:4]).all() and np.isnan(rows[k, 4:13]).all()
    assert np.isnan(rows[k + 1:, 1:]).all()
    # from there on: every attempt accepted (no further rejection or cone retry), dt unchanged, t = t + dt
    t = rows[k + 1:, 0]
    assert np.isfinite(t).all()
    dt = t[1] - t[0]
    acc = t[0]
    for j in range(1, len(t)):
        acc = acc + dt
        assert acc == t[j] or np.isclose(acc, t[j], rtol=1e-15, atol=0)
    assert len(np.unique(np.diff(t))) <= 3                    # constant dt up to the rounding of t + dt
    full = ol.OracleModel("ngo", common.NGO_CFG).trace(pos0[r], [0, 0, 0], w0[r], **dict(common.C2, maxsteps=700))
    assert o["nrejected"] == full["nrejected"] and o["ncone"] == full["ncone"]  # nothing rejected after step 700
