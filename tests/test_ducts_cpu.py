"""Field-aligned ducts of the Ngo model (KDUCTS >= 2: fortran/ngo_dens_model.f95:240-337; cards :98-108): Gaussian ducts
(two-sided and one-sided), the sinusoidal perturbation beyond the knee (l0(2) <= 0, with and without further ducts).
Oracle (statement-by-statement restatement) against the host build of the device physics: bit for bit, at points that
reach every branch (inside / above / below the duct ends, both hemispheres, latitude exactly 0) and along rays."""
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

CFGS = ["newray_ducts_gauss.in", "newray_ducts_sin.in", "newray_ducts_sin_gauss.in"]


def duct_points():
    rng = np.random.default_rng(7)
    n = 6000
    lat = rng.uniform(-80, 80, n)
    lat[:200] = 0.0  # latitu == 0 takes both hemispheres' assignments in sequence
    lon = rng.uniform(0, 360, n)
    alt = 10 ** rng.uniform(np.log10(150.0), np.log10(45000.0), n)
    return np.array([ol.launch_point(la, lo, al) for la, lo, al in zip(lat, lon, alt)])


@pytest.mark.parametrize("cfg", CFGS)
def test_duct_density_bitwise_and_effective(built, cfg):
    path = os.path.join(common.DATA, cfg)
    o, h = ol.OracleModel("ngo", path), common.HostEmu("ngo", path)
    base = ol.OracleModel("ngo", common.NGO_CFG)
    x = duct_points()
    a, b = o.plasma(x), h.plasma(x)
    assert np.array_equal(a, b)
    changed = np.abs(a[:, 4] / base.plasma(x)[:, 4] - 1.0) > 1e-6
    assert changed.sum() > 300  # the duct branches did run
    assert np.isfinite(a[:, 4:8]).all() and (a[:, 4] > 0).all()


@pytest.mark.parametrize("cfg", CFGS)
def test_rays_through_ducts_bitwise(built, cfg):
    path = os.path.join(common.DATA, cfg)
    o, h = ol.OracleModel("ngo", path), common.HostEmu("ngo", path)
    P = dict(common.C2, maxsteps=250)
    for f, lat in [(1000.0, 38.0), (3000.0, 50.0), (5000.0, 57.0), (800.0, 62.0)]:
        pos0, dir0, w0 = common.launch_grid([f], [lat])
        a = o.trace(pos0[0], [0, 0, 0], w0[0], **P)
        b = h.trace(pos0[0], [0, 0, 0], w0[0], **P)
        assert a["stopcond"] == b["stopcond"] and a["nrejected"] == b["nrejected"] and a["ncone"] == b["ncone"]
        assert np.array_equal(a["rows"], b["rows"], equal_nan=True)
