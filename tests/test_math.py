"""srt_math.h (the elementary-function layer shared by the kernels and the oracle) against mpmath:
every function within 1 ulp on the ranges the hot path uses."""
import ctypes as C
import os

import mpmath as mp
import numpy as np
import pytest

import common

mp.mp.prec = 200


@pytest.fixture(scope="module")
def probe(built):
    return C.CDLL(os.path.join(common.BUILD, "math_probe.so"))


def ulp_err(got, want_mp):
    errs = []
    for g, w in zip(got, want_mp):
        wf = float(w)
        if wf == 0.0 or not np.isfinite(wf):
            continue
        u = np.spacing(abs(wf))
        errs.append(abs(float((mp.mpf(float(g)) - w) / mp.mpf(u))))
    return max(errs)


def run1(probe, name, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.empty_like(x)
    getattr(probe, "probe_" + name)(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), C.c_long(len(x)))
    return y


@pytest.mark.parametrize("name,lo,hi,fn", [
    ("exp", -170.0, 30.0, mp.exp), ("log", 1e-300, 1e300, mp.log), ("sin", -50.0, 50.0, mp.sin), ("cos", -50.0, 50.0, mp.cos),
    ("atan", -1e6, 1e6, mp.atan), ("acos", -1.0, 1.0, mp.acos)])
def test_unary_within_one_ulp(probe, name, lo, hi, fn):
    rng = np.random.default_rng(hash(name) % 1000)
    if name == "log":
        x = 10 ** rng.uniform(-300, 300, 3000)
    elif name == "atan":
        x = np.concatenate([rng.uniform(-4, 4, 2000), rng.uniform(lo, hi, 1000)])
    else:
        x = rng.uniform(lo, hi, 3000)
    y = run1(probe, name, x)
    assert ulp_err(y, [fn(mp.mpf(float(v))) for v in x]) < 1.0


def test_pow_dipole_exponent(probe):
    rng = np.random.default_rng(3)
    x = rng.uniform(0.9, 12.0, 3000)   # R/R_E in bmodel_dipole, exponent -3.0 (Q5); (rconsn/r)**expk in dens
    for e in (-3.0, 3.0, 2.5):
        p = np.full_like(x, e)
        y = np.empty_like(x)
        probe.probe_pow(x.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), C.c_long(len(x)))
        assert ulp_err(y, [mp.power(mp.mpf(float(v)), e) for v in x]) < 1.0


def test_atan2_and_csqrt(probe):
    rng = np.random.default_rng(4)
    a, b = rng.normal(size=3000) * 1e7, rng.normal(size=3000) * 1e7
    y = np.empty_like(a)
    probe.probe_atan2(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p), C.c_long(len(a)))
    assert ulp_err(y, [mp.atan2(mp.mpf(float(u)), mp.mpf(float(v))) for u, v in zip(a, b)]) < 1.0
    re, im = rng.normal(size=2000) * 1e3, np.where(rng.uniform(size=2000) < 0.5, 0.0, rng.normal(size=2000))
    ore, oim = np.empty_like(re), np.empty_like(re)
    probe.probe_csqrt(re.ctypes.data_as(C.c_void_p), im.ctypes.data_as(C.c_void_p), ore.ctypes.data_as(C.c_void_p), oim.ctypes.data_as(C.c_void_p), C.c_long(len(re)))
    for r, i, a_, b_ in zip(re, im, ore, oim):
        w = mp.sqrt(mp.mpc(float(r), float(i)))
        assert abs(a_ - float(w.real)) <= 2 * np.spacing(abs(float(w.real))) + 1e-300
        assert abs(b_ - float(w.imag)) <= 2 * np.spacing(abs(float(w.imag))) + 1e-300
