"""TS05 external field (use_tsyganenko = 1), on the CPU.

The yardstick is the REFERENCE'S OWN SOURCE: tsyganenko/TS05_aka_TS04.for translated statement by statement to C++ at
build time (oracle/f77_to_cpp.py -> oracle/_ref/libts05_ref*.so; no Fortran compiler exists here).  The statements the GPU
runs (csrc/srt_ts05.h is host/device code) are compiled for the host and compared with it
  * bit for bit, module by module, when both use the elementary functions of csrc/srt_math.h,
  * within 1e-6 of each module's magnitude when the reference uses glibc's libm (the model differentiates numerically with
    steps of 1e-6 and 1e-4, which amplifies last-bit differences of sin/cos/exp/pow),
and the adapter's field block with the term switched on is compared oracle <-> host build of the device physics."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

NAMES = ["total", "CF", "T1", "T2", "SRC", "PRC", "R11", "R21"]
PARMOD = [3.0, -45.0, 2.5, -4.0, 0.4, 0.7, 0.3, 0.9, 0.6, 1.1]


@pytest.fixture(scope="module")
def probe():
    common.build_hostemu()
    P = C.CDLL(os.path.join(common.BUILD, "ts05_probe.so"))
    d, f = C.c_double, C.c_float
    P.ts05_probe_extern.argtypes = [C.c_void_p, d, d, d, d, C.c_void_p]
    P.ts05_probe_t04.argtypes = [C.c_void_p, f, f, f, f, C.c_void_p]
    return P


def need_ref(srt_math=True):
    R = ol.ts05_ref(srt_math)
    if R is None:
        pytest.skip("oracle/_ref/libts05_ref*.so not built (reference sources absent)")
    return R


def ref_extern(R, par, ps, pt):
    par = np.ascontiguousarray(par, dtype=np.float64)
    o = np.zeros(30)
    R.ts05_ref_extern(par.ctypes.data, ps, pt[0], pt[1], pt[2], o.ctypes.data)
    return o.reshape(10, 3)[[0, 1, 2, 3, 4, 5, 6, 8]]  # the modes R12 and R22 carry no amplitude in EXTERN


def mine_extern(P, par, ps, pt):
    par = np.ascontiguousarray(par, dtype=np.float64)
    o = np.zeros(24)
    P.ts05_probe_extern(par.ctypes.data, ps, pt[0], pt[1], pt[2], o.ctypes.data)
    return o.reshape(8, 3)


def sample(rng, n):
    """parameters over (and beyond) the model's stated range, points from 1.05 to 14 R_E: inside, boundary layer, outside"""
    for _ in range(n):
        par = [rng.uniform(0.5, 8), rng.uniform(-150, 15), rng.uniform(-8, 8), rng.uniform(-10, 8)] + list(rng.uniform(0, 3, 6))
        par[1] = par[1] * 0.8 - 13.0 * np.sqrt(par[0])
        u = rng.uniform(-1, 1, 3)
        yield par, rng.uniform(-0.6, 0.6), u / np.linalg.norm(u) * rng.uniform(1.05, 14.0)


SPECIAL = [(0.0, 0.0, 3.0), (0.0, 0.0, -4.5), (1e-3, 0.0, 5.0), (0.0, 2e-3, -2.0), (3.0, 0.0, 0.0), (-6.0, 0.0, 0.0), (0.0, 7.0, 0.0),
           (10.6, 0.0, 0.0), (10.9, 1.0, 0.5), (11.05, 0.0, 0.2), (14.0, 3.0, 1.0), (-30.0, 5.0, 2.0), (2.0, 2.0, 1e-4)]


def test_every_module_bit_identical_to_the_translated_reference(probe):
    R = need_ref(True)
    rng = np.random.default_rng(7)
    cases = list(sample(rng, 1500)) + [(PARMOD[:1] + [PARMOD[1] * 0.8 - 13 * np.sqrt(PARMOD[0])] + PARMOD[2:], ps, np.array(pt))
                                       for pt in SPECIAL for ps in (0.0, 0.17, -0.45)]
    bad = np.zeros(8, dtype=int)
    branches = set()
    for par, ps, pt in cases:
        a, b = ref_extern(R, par, ps, pt), mine_extern(probe, par, ps, pt)
        for k in range(8):
            bad[k] += not np.array_equal(a[k].view(np.uint64), b[k].view(np.uint64))
        branches.add("inside" if a[1].any() else "outside")
    assert not bad.any(), dict(zip(NAMES, bad))
    assert branches == {"inside", "outside"}


def test_against_the_glibc_build_of_the_reference(probe):
    R = need_ref(False)
    rng = np.random.default_rng(8)
    worst = np.zeros(8)
    for par, ps, pt in sample(rng, 1500):
        a, b = ref_extern(R, par, ps, pt), mine_extern(probe, par, ps, pt)
        for k in range(8):
            worst[k] = max(worst[k], np.abs(a[k] - b[k]).max() / (np.linalg.norm(a[k]) + 1e-300))
    print("TS05 vs reference on glibc libm, worst difference / |B| per module:", dict(zip(NAMES, ["%.1e" % w for w in worst])))
    assert worst.max() < 1e-6


def test_t04_s_single_precision_boundary(probe):
    """T04_s itself: REAL*4 arguments and results around REAL*8 arithmetic, incl. DST* = PARMOD(2)*0.8 - 13*SQRT(PDYN) in mixed precision"""
    R = need_ref(True)
    rng = np.random.default_rng(9)
    pm = np.array(PARMOD, dtype=np.float32)
    for _ in range(500):
        u = rng.uniform(-1, 1, 3)
        x, y, z = (u / np.linalg.norm(u) * rng.uniform(1.1, 13.0)).astype(np.float32)
        ps = np.float32(rng.uniform(-0.5, 0.5))
        a, b = np.zeros(3, dtype=np.float32), np.zeros(3, dtype=np.float32)
        R.ts05_ref_t04_s(pm.ctypes.data, ps, x, y, z, a.ctypes.data)
        probe.ts05_probe_t04(pm.ctypes.data, ps, x, y, z, b.ctypes.data)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), (x, y, z, ps, a, b)


def test_module_fields_are_divergence_free(probe):
    """reference-independent: every module is the curl of a vector potential or a gradient of a harmonic function"""
    par = PARMOD[:1] + [PARMOD[1] * 0.8 - 13 * np.sqrt(PARMOD[0])] + PARMOD[2:]
    h = 1e-4
    for pt in [(-5.0, 0.5, 1.0), (4.0, 1.0, 2.0), (-8.0, 2.0, 1.0), (3.0, 3.0, 3.0), (1.5, -2.0, 0.3)]:
        div = np.zeros(8)
        for i in range(3):
            p, m = np.array(pt), np.array(pt)
            p[i] += h
            m[i] -= h
            div += (mine_extern(probe, par, 0.2, p)[:, i] - mine_extern(probe, par, 0.2, m)[:, i]) / (2 * h)
        f = mine_extern(probe, par, 0.2, np.array(pt))
        scale = np.linalg.norm(f, axis=1)  # nT; gradients are of the order |B| / R_E
        assert (np.abs(div) < 2e-4 * scale).all(), (pt, dict(zip(NAMES, div / scale)))


def points(n, seed):
    rng = np.random.default_rng(seed)
    r = ol.R_E * rng.uniform(1.1, 7.0, n)
    th, ph = rng.uniform(0.2, np.pi - 0.2, n), rng.uniform(-np.pi, np.pi, n)
    return np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), r * np.cos(th)], axis=1)


@pytest.mark.parametrize("use_igrf", [0, 1])
def test_adapter_field_block_with_the_term_on(use_igrf):
    """funcPlasmaParams with use_tsyganenko = 1: oracle (adapter restatement + translated T04_s) == host build of the device physics"""
    need_ref(True)
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 3600000, use_igrf).tsyganenko(PARMOD)
    e = common.HostEmu("ngo", common.NGO_CFG, 2010001, 3600000, use_igrf).tsyganenko(PARMOD)
    off = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 3600000, use_igrf)
    x = points(400, 3)
    a, b, c = o.plasma(x), e.plasma(x), off.plasma(x)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    assert 0.16 < o.L.orc_model_psi(o.h) < 0.2  # `PSI` of the adapters is sin(theta0) of the dipole axis (Q25)
    rel = np.linalg.norm(a[:, 16:19] - c[:, 16:19], axis=1) / np.linalg.norm(c[:, 16:19], axis=1)
    assert rel.max() > 0.05 and np.median(rel) > 1e-3  # the term is there: tens of nT against the main field


def test_rays_through_the_external_field():
    """whole adaptive rays with the term on: oracle == host build of the device physics, every row"""
    need_ref(True)
    o = ol.OracleModel("ngo", common.NGO_CFG).tsyganenko(PARMOD)
    e = common.HostEmu("ngo", common.NGO_CFG).tsyganenko(PARMOD)
    off = ol.OracleModel("ngo", common.NGO_CFG)
    pos0, dir0, w0 = common.c2_grid(3, 3)
    params = dict(common.C2, maxsteps=150)
    moved = 0
    for i in range(len(w0)):
        a = o.trace(pos0[i], dir0[i], w0[i], **params)
        b = e.trace(pos0[i], dir0[i], w0[i], **params)
        c = off.trace(pos0[i], dir0[i], w0[i], **params)
        assert a["stopcond"] == b["stopcond"] and a["rows"].shape == b["rows"].shape
        assert np.array_equal(a["rows"].view(np.uint64), b["rows"].view(np.uint64)), i
        n = min(len(a["rows"]), len(c["rows"]))
        moved += not np.array_equal(a["rows"][:n], c["rows"][:n])
    assert moved == len(w0)
