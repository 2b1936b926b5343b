"""CPU checks of the gridded-density + IGRF path (BASELINE config 3): the oracle's restatement of
libtricubic / interp_dens_model_adapter / Geopack-2008 against reference-independent anchors, and the
device physics compiled for the host (tests/csrc/hostemu.cpp) against the oracle, bit for bit."""
import json
import os

import numpy as np
import pytest

import common
import oracle_lib as ol
from stanford_raytracer_b200 import gridfile

R_E = 6371.2e3
GOLD = os.path.join(common.ROOT, "tests", "golden")
GRID13 = os.path.join(GOLD, "grid13.txt")


def _poly_grid(tmp_path, coef, n=(7, 6, 8), ext=(-1.0, 1.5, -1.0, 1.25, 0.0, 2.0), derivs=False):
    """grid of p(x,y,z) = sum c[i][j][k] x^i y^j z^k (degree <= 3 per axis), optionally with exact derivatives"""
    nx, ny, nz = n
    xs, ys, zs = gridfile.node_axes(nx, ext[0], ext[1]), gridfile.node_axes(ny, ext[2], ext[3]), gridfile.node_axes(nz, ext[4], ext[5])
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")

    def ev(dx, dy, dz):
        out = np.zeros_like(X)
        for i in range(4):
            for j in range(4):
                for k in range(4):
                    if i < dx or j < dy or k < dz:
                        continue
                    fac = np.prod([i - t for t in range(dx)]) * np.prod([j - t for t in range(dy)]) * np.prod([k - t for t in range(dz)])
                    out += coef[i, j, k] * fac * X ** (i - dx) * Y ** (j - dy) * Z ** (k - dz)
        return out

    f = np.repeat(ev(0, 0, 0)[..., None], 2, axis=3)
    f[..., 1] *= -0.5
    d = None
    if derivs:
        d = []
        for (a, b, c) in [(1, 0, 0), (0, 1, 0), (0, 0, 1), (1, 1, 0), (1, 0, 1), (0, 1, 1), (1, 1, 1)]:
            g = np.repeat(ev(a, b, c)[..., None], 2, axis=3)
            g[..., 1] *= -0.5
            d.append(g)
    path = str(tmp_path / "poly.txt")
    gridfile.write_grid(path, f, [-1.602e-19, 1.602e-19], [9.1e-31, 1.67e-27], ext, derivs=d)
    return path, ext


def _poly_eval(coef, x):
    out = np.zeros(len(x))
    for i in range(4):
        for j in range(4):
            for k in range(4):
                out += coef[i, j, k] * x[:, 0] ** i * x[:, 1] ** j * x[:, 2] ** k
    return out


def test_tricubic_reproduces_tricubic_polynomials(built, tmp_path):
    """Lekien-Marsden: with exact derivatives at the nodes the interpolant reproduces any polynomial of
    degree <= 3 per axis -- a reference-independent check of Amat, the corner ordering and the evaluation."""
    rng = np.random.default_rng(5)
    coef = rng.normal(size=(4, 4, 4)) * 0.2
    path, ext = _poly_grid(tmp_path, coef, derivs=True)
    og = ol.OracleModel("grid", path)
    x = np.stack([rng.uniform(ext[0], ext[1], 500), rng.uniform(ext[2], ext[3], 500), rng.uniform(ext[4], ext[5], 500)], axis=1)
    got = np.log(og.plasma(x)[:, 4:6])
    want = _poly_eval(coef, x)
    assert np.abs(want).max() < 60
    assert np.allclose(got[:, 0], want, rtol=1e-12, atol=1e-11)
    assert np.allclose(got[:, 1], -0.5 * want, rtol=1e-12, atol=1e-11)
    for pc in (False, True):
        he = common.HostEmu("grid", path, precomputed_coef=pc)
        assert np.array_equal(he.plasma(x), og.plasma(x))


def test_finite_difference_derivatives_exact_for_quadratics(built, tmp_path):
    """compder = 0: centred differences are exact for a polynomial of degree <= 2 per axis at interior
    nodes, so interior cells (two nodes away from the faces) reproduce it."""
    rng = np.random.default_rng(6)
    coef = np.zeros((4, 4, 4))
    coef[:3, :3, :3] = rng.normal(size=(3, 3, 3)) * 0.2
    path, ext = _poly_grid(tmp_path, coef, n=(9, 9, 9), ext=(-2.0, 2.0, -2.0, 2.0, -2.0, 2.0))
    og = ol.OracleModel("grid", path)
    x = rng.uniform(-1.0, 1.0, (400, 3))
    got = np.log(og.plasma(x)[:, 4])
    assert np.allclose(got, _poly_eval(coef, x), rtol=1e-12, atol=1e-11)
    he = common.HostEmu("grid", path, precomputed_coef=True)
    assert np.array_equal(he.plasma(x), og.plasma(x))


def test_edges_and_outside_match_oracle_bitwise(built):
    """Q12: sticky clamp flags at / outside the grid faces (incl. xi < x(1) -> maxloc 0 and xi >= x(n))."""
    og = ol.OracleModel("grid", GRID13, use_igrf=1)
    he = common.HostEmu("grid", GRID13, use_igrf=1, precomputed_coef=True)
    ext = 4 * R_E
    rng = np.random.default_rng(8)
    x = rng.uniform(-1.3 * ext, 1.3 * ext, (3000, 3))
    # exactly on nodes, faces and corners
    nodes = gridfile.node_axes(13, -ext, ext)
    special = np.array([[nodes[0], nodes[3], nodes[5]], [nodes[12], nodes[12], nodes[12]], [nodes[11], 0.0, nodes[12]], [nodes[0], nodes[0], nodes[0]],
                        [np.nextafter(nodes[12], 0), 1.0, 2.0], [np.nextafter(nodes[0], -1e30), 1.0, 2.0]])
    # far outside, and the positions of a ray whose state has become NaN / Inf (about 1 % of the config-3 rays do and
    # are then stepped to maxsteps: raytracer_stopconditions catches no NaN): locate returns 0 or n on those axes
    nan, inf = float("nan"), float("inf")
    wild = np.array([[nan, nan, nan], [nan, 1.0e6, -2.0e7], [3.0e6, nan, nan], [inf, 0.0, 1.0e7], [-inf, -inf, inf], [50 * ext, -50 * ext, 0.0]])
    x = np.concatenate([x, special])
    a, b = og.plasma(x), he.plasma(x)
    assert np.array_equal(a, b, equal_nan=True)
    assert np.isfinite(a[np.linalg.norm(x, axis=1) > R_E]).all()
    # the coefficient table covers the extended lattice: edge / outside / NaN queries read it like interior ones and
    # give the bits of the on-the-fly path (hostemu without a table) and of the oracle
    fly = common.HostEmu("grid", GRID13, use_igrf=1, precomputed_coef=False)
    xa = np.concatenate([x, wild])
    t, f, o = he.plasma(xa), fly.plasma(xa), og.plasma(xa)
    assert np.array_equal(t, f, equal_nan=True)
    assert np.array_equal(t[:, 4:8], o[:, 4:8], equal_nan=True)  # densities; B0 of a NaN position is NaN in all three
    assert np.array_equal(t, o, equal_nan=True)
    assert np.isfinite(t[-len(wild):, 4:8]).all()  # the clamped corner's density: finite even for a NaN position


def test_dense_and_sparse_matmul_agree(built):
    """Q13: skipping the exact zeros of Amat does not change a bit (finite data)."""
    a = ol.OracleModel("grid", GRID13, faithful_cost=True)   # dense 64x64 matmul, RECALC_08 per call
    b = ol.OracleModel("grid", GRID13)
    x = np.random.default_rng(9).uniform(-4 * R_E, 4 * R_E, (300, 3))
    assert np.array_equal(a.plasma(x), b.plasma(x))


def test_igrf_is_dipole_like_and_fp32(built):
    og = ol.OracleModel("ngo", common.NGO_CFG, use_igrf=1)
    od = ol.OracleModel("ngo", common.NGO_CFG, use_igrf=0)
    # surface field magnitude within the known IGRF range (22-67 uT)
    rng = np.random.default_rng(10)
    x = np.array([ol.launch_point(la, lo, 0.0) for la, lo in zip(rng.uniform(-90, 90, 300), rng.uniform(0, 360, 300))])
    b = np.linalg.norm(og.plasma(x)[:, 16:19], axis=1)
    assert 2.0e-5 < b.min() and b.max() < 7.0e-5
    # far field ~ centred dipole within 15 % (dipole moment of 2010 vs the model's 0.312 G, plus quadrupole at 5 R_E)
    x5 = np.array([ol.launch_point(la, lo, 4 * 6371.2) for la, lo in zip(rng.uniform(-60, 60, 100), rng.uniform(0, 360, 100))])
    bi, bd = og.plasma(x5)[:, 16:19], od.plasma(x5)[:, 16:19]
    cosang = (bi * bd).sum(axis=1) / np.linalg.norm(bi, axis=1) / np.linalg.norm(bd, axis=1)
    assert cosang.min() > 0.95
    assert np.all(np.abs(np.linalg.norm(bi, axis=1) / np.linalg.norm(bd, axis=1) - 1) < 0.2)
    # Q3/Q4: every component is a float32 value times 1e-9, rotated: check through |B| granularity instead
    # of bits -- the GSM-frame components are exactly representable in fp32 nT
    he = common.HostEmu("ngo", common.NGO_CFG, use_igrf=1)
    assert np.array_equal(he.plasma(x5), og.plasma(x5))


def test_igrf_epoch_brackets(built):
    """RECALC_08 interpolates linearly inside 5-year brackets and extrapolates after 2020: the dipole
    strength must change smoothly and monotonically (it has been decaying ~0.05 %/yr)."""
    x = np.array([[3 * R_E, 0.0, 0.0], [0.1 * R_E, 0.0, 3 * R_E]])
    mags = []
    for year in (1966, 1975, 1985, 1995, 2005, 2015, 2020, 2024):
        m = ol.OracleModel("ngo", common.NGO_CFG, yearday=year * 1000 + 1, use_igrf=1)
        he = common.HostEmu("ngo", common.NGO_CFG, yearday=year * 1000 + 1, use_igrf=1)
        p = m.plasma(x)
        assert np.array_equal(p, he.plasma(x))
        mags.append(np.linalg.norm(p[1, 16:19]))
    mags = np.array(mags)
    assert np.all(np.diff(mags) < 0) and mags[-1] / mags[0] > 0.93


def test_grid_rays_hostemu_equals_oracle_and_golden(built):
    meta = json.load(open(os.path.join(GOLD, "golden.json")))
    z = np.load(os.path.join(GOLD, "golden_grid.npz"))
    og = ol.OracleModel("grid", GRID13, use_igrf=1)
    he = common.HostEmu("grid", GRID13, use_igrf=1, precomputed_coef=True)
    assert np.array_equal(og.plasma(z["plasma_x"]), z["plasma_out"], equal_nan=True)
    assert np.array_equal(he.plasma(z["plasma_x"]), z["plasma_out"], equal_nan=True)
    for i, case in enumerate(meta["grid_rays"]):
        o = og.trace(np.array(case["pos0"]), [0, 0, 0], case["w0"], **case["params"])
        h = he.trace(np.array(case["pos0"]), [0, 0, 0], case["w0"], **case["params"])
        assert o["stopcond"] == case["stopcond"] == h["stopcond"]
        assert np.array_equal(o["rows"], z["rows_%d" % i], equal_nan=True)
        assert np.array_equal(h["rows"], z["rows_%d" % i], equal_nan=True)


def test_grid_file_roundtrip(tmp_path):
    rng = np.random.default_rng(11)
    f = rng.normal(size=(3, 4, 5, 4))
    ext = (-1.0, 1.0, -2.0, 2.0, -3.0, 3.0)
    p = str(tmp_path / "g.txt")
    gridfile.write_grid(p, f, [1, 2, 3, 4], [5, 6, 7, 8], ext)
    r = gridfile.read_grid(p)
    assert (r["compder"], r["nspec"], r["nx"], r["ny"], r["nz"]) == (0, 4, 5, 4, 3)
    assert np.allclose(r["f"], f, rtol=1e-15, atol=0)
    lines = open(p).read().split("\n")
    assert lines[0] == "%10d%10d%10d%10d%10d" % (0, 4, 5, 4, 3)
    assert len(lines[1]) == 6 * 24 and all(len(ln) == 24 for ln in lines[4:4 + f.size])
    assert "E+000" in lines[1] or "E-00" in lines[1]
