"""IGRF against the REFERENCE'S OWN SOURCE: tsyganenko/geopack2008.for (RECALC_08, SUN_08, IGRF_GSW_08,
GEOGSW_08) translated statement by statement at build time (oracle/f77_to_cpp.py -> oracle/_ref/libgeopack_ref.so; there is
no Fortran compiler here).  The oracle's restatement (oracle/orc_igrf.h), the product's host set-up (csrc/srt_igrf_setup.h)
and the statement of IGRF_GSW_08 the GPU runs (csrc/srt_igrf.h, compiled for the host) must reproduce it BIT FOR BIT:
the interpolated and Schmidt-normalised coefficients G, H, REC, the GEO->GSW matrix, the words of COMMON /GEOPACK1/ and
the field at random points, for epochs in every interpolation bracket of RECALC_08 (geopack2008.for:890-1012)."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

GRID13 = os.path.join(common.ROOT, "tests", "golden", "grid13.txt")
# (yyyyddd, milliseconds of day): one date per bracket 1965-70 ... 2015-20, the extrapolation after 2020, and both clamps
EPOCHS = [(1967100, 0), (1972200, 3600000), (1977001, 43200000), (1983150, 1000), (1988300, 86399000), (1992060, 7200000), (1997365, 0),
          (2003010, 5000000), (2008180, 0), (2010001, 0), (2013250, 64800000), (2017033, 123000), (2022200, 0), (2024366, 40000000),
          (1960001, 0), (2031001, 0)]
WORDS = ["ST0", "CT0", "SL0", "CL0", "CTCL", "STCL", "CTSL", "STSL", "SFI", "CFI", "SPS", "CPS", "DS3", "CGST", "SGST", "PSI",
         "A11", "A21", "A31", "A12", "A22", "A32", "A13", "A23", "A33", "E11", "E21", "E31", "E12", "E22", "E32", "E13", "E23", "E33"]


@pytest.fixture(scope="module")
def ref():
    path = os.path.join(ol.ORC_DIR, "_ref", "libgeopack_ref.so")
    if not os.path.exists(path):
        ol.ts05_ref(True)  # runs `make ref` where the reference sources exist
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libgeopack_ref.so not built (reference sources absent)")
    R = C.CDLL(path)
    f, i = C.c_float, C.c_int
    R.geopack_ref_recalc.argtypes = [i, i, i, i, i, C.c_void_p, C.c_void_p]
    R.geopack_ref_igrf_gsw.argtypes = [f, f, f, C.c_void_p]
    return R


def ref_state(R, yearday, msec):
    hour = msec // 3600000
    mn = (msec - hour * 3600000) // 60000
    sec = (msec - hour * 3600000 - mn * 60000) // 1000
    g1, g2 = np.zeros(34, dtype=np.float32), np.zeros(315, dtype=np.float32)
    R.geopack_ref_recalc(yearday // 1000, yearday % 1000, hour, mn, sec, g1.ctypes.data, g2.ctypes.data)
    return g1, g2


def points(n, seed):
    rng = np.random.default_rng(seed)
    u = rng.uniform(-1, 1, (n, 3))
    return (u / np.linalg.norm(u, axis=1)[:, None] * rng.uniform(1.02, 9.0, n)[:, None]).astype(np.float32)


@pytest.mark.parametrize("yearday,msec", EPOCHS)
def test_recalc_and_field_equal_the_translated_reference(ref, yearday, msec):
    g1, g2 = ref_state(ref, yearday, msec)
    o = ol.OracleModel("ngo", common.NGO_CFG, yearday, msec, 1)
    L = o.L
    L.orc_model_geopack.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.orc_model_igrf_gsw.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p]
    o1, o2 = np.zeros(34, dtype=np.float32), np.zeros(315, dtype=np.float32)
    L.orc_model_geopack(o.h, o1.ctypes.data, o2.ctypes.data)
    assert np.array_equal(o2.view(np.uint32), g2.view(np.uint32)), "G, H, REC"
    filled = [k for k in range(34) if o1[k] != 0.0 or g1[k] == 0.0]  # the oracle keeps only the words later routines read
    assert {"ST0", "CT0", "STCL", "STSL", "CGST", "SGST", "SPS", "PSI", "A11", "A21", "A31", "A12", "A22", "A32", "A13", "A23", "A33"} <= {WORDS[k] for k in filled}
    bad = [WORDS[k] for k in filled if o1[k:k + 1].view(np.uint32)[0] != g1[k:k + 1].view(np.uint32)[0]]
    assert not bad, bad
    # the product's host set-up and the device statement of IGRF_GSW_08
    e = common.HostEmu("grid", GRID13, yearday, msec, 1)
    E = e.E
    E.emu_model_igrf.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    E.emu_igrf_gsw.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_void_p]
    E.emu_model_ts_ps.restype = C.c_float
    E.emu_model_ts_ps.argtypes = [C.c_void_p]
    a9, p2 = np.zeros(9, dtype=np.float32), np.zeros(315, dtype=np.float32)
    E.emu_model_igrf(e.h, a9.ctypes.data, p2.ctypes.data)
    assert np.array_equal(p2.view(np.uint32), g2.view(np.uint32))
    assert np.array_equal(a9.view(np.uint32), g1[16:25].view(np.uint32))
    e.tsyganenko([2.0, -10.0, 0.0, 0.0, 0, 0, 0, 0, 0, 0])
    # Q25: the adapters' `PSI` names word 1 of COMMON /GEOPACK1/ -- ST0 in the compiled Geopack-2008, not the tilt angle (word 16)
    assert np.float32(E.emu_model_ts_ps(e.h)) == g1[0] and g1[0] != g1[15]
    assert o.L.orc_model_psi(o.h) == float(g1[0])
    br, bo, be = np.zeros(3, dtype=np.float32), np.zeros(3, dtype=np.float32), np.zeros(3, dtype=np.float32)
    for x, y, z in points(300, yearday):
        ref.geopack_ref_igrf_gsw(x, y, z, br.ctypes.data)
        L.orc_model_igrf_gsw(o.h, x, y, z, bo.ctypes.data)
        E.emu_igrf_gsw(e.h, x, y, z, be.ctypes.data)
        assert np.array_equal(br.view(np.uint32), bo.view(np.uint32)) and np.array_equal(br.view(np.uint32), be.view(np.uint32)), (x, y, z, br, bo, be)
    assert 2.0e4 < np.linalg.norm(br) * np.linalg.norm([x, y, z]) ** 3 < 7.0e4  # a dipole-like field of ~3e4 nT R_E^3
