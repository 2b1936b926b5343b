"""GPU parity for BASELINE config 3's path: tricubic-interpolated gridded density + IGRF field
(modelnum 3 with use_igrf = 1), through the C ABI, against the CPU oracle: BIT-EXACT."""
import json
import os

import numpy as np
import pytest

import common
import oracle_lib as ol
from stanford_raytracer_b200 import gridfile

pytestmark = pytest.mark.gpu
R_E = 6371.2e3
GOLD = os.path.join(common.ROOT, "tests", "golden")
GRID13 = os.path.join(GOLD, "grid13.txt")


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    s.init(0)
    return s


@pytest.fixture(scope="module")
def grid(srt):
    return srt.interpsetup(GRID13, 2010001, 0, 1, 0)


@pytest.fixture(scope="module")
def ogrid():
    return ol.OracleModel("grid", GRID13, use_igrf=1)


@pytest.fixture(params=["thread", "team"])
def kernel_mode(request, monkeypatch):
    monkeypatch.setenv("SRT_KERNEL", request.param)
    return request.param


def test_grid_plasma_bitwise_inside_edges_outside(srt, grid, ogrid):
    rng = np.random.default_rng(8)
    x = rng.uniform(-1.3 * 4 * R_E, 1.3 * 4 * R_E, (20000, 3))
    nodes = gridfile.node_axes(13, -4 * R_E, 4 * R_E)
    special = np.array([[nodes[0], nodes[3], nodes[5]], [nodes[12], nodes[12], nodes[12]], [nodes[11], 0.0, nodes[12]],
                        [np.nextafter(nodes[12], 0), 1.0, 2.0], [np.nextafter(nodes[0], -1e30), 1.0, 2.0]])
    x = np.concatenate([x, special])
    g = grid.funcPlasmaParams(x)["raw"]
    o = ogrid.plasma(x)
    assert np.array_equal(g, o, equal_nan=True)


def test_precomputed_coefficients_equal_on_the_fly(srt, grid, monkeypatch):
    monkeypatch.setenv("SRT_GRID_COEF", "0")
    slow = srt.interpsetup(GRID13, 2010001, 0, 1, 0)
    x = np.random.default_rng(9).uniform(-4 * R_E, 4 * R_E, (20000, 3))
    assert np.array_equal(grid.funcPlasmaParams(x)["raw"], slow.funcPlasmaParams(x)["raw"], equal_nan=True)
    # the table covers the extended lattice: edge and outside cells, and where a NaN / Inf position locates to
    nan, inf = float("nan"), float("inf")
    wild = np.array([[nan, nan, nan], [nan, 1.0e6, -2.0e7], [3.0e6, nan, nan], [inf, 0.0, 1.0e7], [-inf, -inf, inf], [2.0e8, -2.0e8, 0.0]])
    xo = np.concatenate([np.random.default_rng(10).uniform(-1.4 * 4 * R_E, 1.4 * 4 * R_E, (20000, 3)), wild])
    t, f = grid.funcPlasmaParams(xo)["raw"], slow.funcPlasmaParams(xo)["raw"]
    assert np.array_equal(t, f, equal_nan=True)
    assert np.isfinite(t[-len(wild):, 4:8]).all()  # the clamped corner's density


def test_golden_grid_fixture(srt, grid):
    z = np.load(os.path.join(GOLD, "golden_grid.npz"))
    assert np.array_equal(grid.funcPlasmaParams(z["plasma_x"])["raw"], z["plasma_out"], equal_nan=True)
    meta = json.load(open(os.path.join(GOLD, "golden.json")))
    for i, case in enumerate(meta["grid_rays"]):
        res = srt.raytracer_run(grid, np.array([case["pos0"]]), np.zeros((1, 3)), np.array([case["w0"]]), **case["params"])
        assert res.stopcond[0] == case["stopcond"]
        assert np.array_equal(res[0].rows, z["rows_%d" % i], equal_nan=True)


def test_grid_igrf_rays_bitwise(srt, grid, ogrid, kernel_mode):
    pos0, dir0, w0 = common.launch_grid([500.0, 3000.0, 15000.0], [15.0, 35.0, 55.0], [0.0, 108.0])
    p = dict(common.C2, maxsteps=300)
    res = srt.raytracer_run(grid, pos0, dir0, w0, **p)
    for r in range(len(res)):
        o = ogrid.trace(pos0[r], [0, 0, 0], w0[r], **p)
        assert res.stopcond[r] == o["stopcond"]
        assert res[r].rows.shape == o["rows"].shape
        assert np.array_equal(res[r].rows, o["rows"], equal_nan=True), r
        assert np.array_equal(res.dir0[r], o["dir0"])


def test_ngo_with_igrf_rays_bitwise(srt, kernel_mode):
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 1, 0)
    o = ol.OracleModel("ngo", common.NGO_CFG, use_igrf=1)
    pos0, dir0, w0 = common.launch_grid([2000.0, 8000.0], [25.0, 45.0], [0.0, 200.0])
    p = dict(common.C2, maxsteps=200)
    res = srt.raytracer_run(m, pos0, dir0, w0, **p)
    for r in range(len(res)):
        t = o.trace(pos0[r], [0, 0, 0], w0[r], **p)
        assert res.stopcond[r] == t["stopcond"] and np.array_equal(res[r].rows, t["rows"], equal_nan=True), r


@pytest.mark.parametrize("yearday,msec", [(1997180, 43200000), (2024100, 3600000), (1968300, 0)])
def test_igrf_other_epochs(srt, yearday, msec):
    """IGRF at other epochs and times of day: different brackets of RECALC_08's coefficient interpolation
    (geopack2008.for:890-1012: 1965-70, 1995-2000, extrapolation beyond 2020) and another GEO->GSW rotation / dipole tilt."""
    m = srt.ngosetup(common.NGO_CFG, yearday, msec, 1, 0)
    o = ol.OracleModel("ngo", common.NGO_CFG, yearday=yearday, msec=msec, use_igrf=1)
    rng = np.random.default_rng(yearday)
    x = np.array([ol.launch_point(la, lo, alt) for la, lo, alt in zip(rng.uniform(-85, 85, 1500), rng.uniform(0, 360, 1500), 10 ** rng.uniform(2.2, 4.7, 1500))])
    assert np.array_equal(m.funcPlasmaParams(x)["raw"], o.plasma(x))
    pos0, dir0, w0 = common.launch_grid([3000.0], [35.0], [0.0, 120.0])
    p = dict(common.C2, maxsteps=150)
    res = srt.raytracer_run(m, pos0, dir0, w0, **p)
    for r in range(len(res)):
        t = o.trace(pos0[r], [0, 0, 0], w0[r], **p)
        assert res.stopcond[r] == t["stopcond"] and np.array_equal(res[r].rows, t["rows"], equal_nan=True), r
    m.close()


def test_c3_shape_reduced(srt, tmp_path):
    """Config 3 at reduced size (40^3 grid sampled from the GPU's own Ngo model = the dumpmodel / buildgrid
    role, 1000 rays): size-independent properties + a strided oracle check."""
    ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    ext = (-10 * R_E, 10 * R_E) * 3
    f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 40, 40, 40, ext)
    path = str(tmp_path / "grid40.txt")
    gridfile.write_grid(path, f, qs, ms, ext)
    m = srt.interpsetup(path, 2010001, 0, 1, 0)
    pos0, dir0, w0 = common.c2_grid(10, 10, 10)
    p = dict(common.C2, maxsteps=250)
    res = srt.raytracer_run(m, pos0, dir0, w0, **p)
    assert res.nrays == 1000 and res.total_rows == int(res.nsteps.sum())
    res2 = srt.raytracer_run(m, pos0, np.zeros_like(pos0), w0, **p)
    assert np.array_equal(res.last_row, res2.last_row, equal_nan=True)
    o = ol.OracleModel("grid", path, use_igrf=1)
    for r in range(0, 1000, 125):
        t = o.trace(pos0[r], [0, 0, 0], w0[r], **p)
        assert res.stopcond[r] == t["stopcond"] and np.array_equal(res[r].rows, t["rows"], equal_nan=True), r
