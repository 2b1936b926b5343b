"""The Ngo density model (`readinput` and `dens` incl. the plasmapause and every duct branch) against
the REFERENCE'S OWN SOURCE: fortran/ngo_dens_model.f95 translated statement by statement at build time
(oracle/f77_to_cpp.py -> oracle/_ref/libngo_ref*.so; no Fortran compiler here).  The oracle's restatement must reproduce it
BIT FOR BIT -- the state `readinput` leaves (incl. ane0 normalised through a `dens` call at the set point) and ani(1:4)
of `dens` for the arguments funcPlasmaParams hands it -- on the elementary functions of csrc/srt_math.h (oracle default)
and on glibc's libm (the oracle's libm build); the product then equals the oracle (test_hostemu_cpu.py, test_ducts_cpu.py,
GPU suites)."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

DATA = os.path.join(common.ROOT, "tests", "data")
CONFIGS = [common.NGO_CFG] + [os.path.join(DATA, f) for f in ("newray_ducts_gauss.in", "newray_ducts_sin.in", "newray_ducts_sin_gauss.in")]


def ref_lib(srt_math):
    name = "libngo_ref_srtmath.so" if srt_math else "libngo_ref.so"
    path = os.path.join(ol.ORC_DIR, "_ref", name)
    if not os.path.exists(path):
        ol.ts05_ref(True)  # runs `make ref` where the reference sources exist
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/%s not built (reference sources absent)" % name)
    R = C.CDLL(path)
    d = C.c_double
    R.ngo_ref_readinput.argtypes = [C.c_char_p]
    R.ngo_ref_dens.argtypes = [d, d, d, C.c_void_p]
    R.ngo_ref_state.argtypes = [C.c_void_p]
    return R


@pytest.mark.parametrize("cfg", CONFIGS, ids=[os.path.basename(c) for c in CONFIGS])
@pytest.mark.parametrize("srt_math", [True, False], ids=["srt_math", "glibc"])
def test_readinput_and_dens_equal_the_translated_reference(cfg, srt_math):
    R = ref_lib(srt_math)
    assert R.ngo_ref_readinput(cfg.encode()) == 0
    o = ol.OracleModel("ngo", cfg, libm=not srt_math)
    L = o.L
    d = C.c_double
    L.orc_ngo_dens.argtypes = [C.c_void_p, d, d, d, C.c_void_p]
    L.orc_ngo_state.argtypes = [C.c_void_p, C.c_void_p]
    a, b = np.zeros(12), np.zeros(12)
    R.ngo_ref_state(a.ctypes.data)
    L.orc_ngo_state(o.h, b.ctypes.data)
    names = ["ane0", "rbase", "r0", "num", "kducts", "therm", "rzero", "scbot", "lk", "expk", "ddk", "pi"]
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), {n: (x, y) for n, x, y in zip(names, a, b) if x != y}
    assert a[11] != np.pi and abs(a[11] - np.pi) < 1e-6  # Q14: pi is a default-real literal
    rng = np.random.default_rng(5)
    ra, rb = np.zeros(4), np.zeros(4)
    n, seen_duct = 20000, 0
    for _ in range(n):
        lam = rng.uniform(-75.0, 75.0)                       # latitude, degrees
        lsh = rng.uniform(1.05, 8.0)                         # L shell: through the plasmapause and every duct
        r = 6370.0 * lsh * np.cos(np.deg2rad(lam)) ** 2      # what the adapter hands over as z(1) [km]
        if r < 6470.0:
            continue
        z2 = np.deg2rad(90.0 - lam)
        R.ngo_ref_dens(r, z2, lam, ra.ctypes.data)
        L.orc_ngo_dens(o.h, r, z2, lam, rb.ctypes.data)
        assert np.array_equal(ra.view(np.uint64), rb.view(np.uint64)), (r, z2, lam, ra, rb)
        seen_duct += 1
    assert seen_duct > n // 2 and ra[0] > 0.0
