"""SM <-> GSM of the adapters (xform_double: sm_to_gsm_d / gsm_to_sm_d -> t4_d -> get_q_c_d -> t1_d, t2_d, t0_d, rotations)
against the REFERENCE'S OWN SOURCE, translated statement by statement at build time (oracle/f77_to_cpp.py ->
oracle/_ref/libxform_ref*.so; no Fortran compiler here).  The oracle's restatement (both its cached and its per-call form)
and the product's host set-up (cos/sin of the dipole tilt, csrc/srt_host_setup.h) reproduce it BIT FOR BIT."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

TIMES = [(2010001, 0), (2010001, 3600000), (1999098, 43200000), (2003290, 86399000), (2020172, 12345678), (1985355, 64800000), (2024060, 1000)]


def ref_lib(srt_math):
    name = "libxform_ref_srtmath.so" if srt_math else "libxform_ref.so"
    path = os.path.join(ol.ORC_DIR, "_ref", name)
    if not os.path.exists(path):
        ol.ts05_ref(True)  # runs `make ref` where the reference sources exist
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/%s not built (reference sources absent)" % name)
    R = C.CDLL(path)
    for f in (R.xform_ref_sm_to_gsm, R.xform_ref_gsm_to_sm):
        f.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    return R


@pytest.mark.parametrize("yearday,msec", TIMES)
@pytest.mark.parametrize("srt_math", [True, False], ids=["srt_math", "glibc"])
def test_sm_gsm_equal_the_translated_reference(yearday, msec, srt_math):
    R = ref_lib(srt_math)
    rng = np.random.default_rng(yearday % 97)
    xs = rng.uniform(-5e7, 5e7, (200, 3))
    a, b = np.zeros(3), np.zeros(3)
    for faithful in (False, True):  # cached tilt angle / the whole chain on every call, as the Fortran does
        o = ol.OracleModel("ngo", common.NGO_CFG, yearday, msec, 0, libm=not srt_math, faithful_cost=faithful)
        for fn_r, fn_o in ((R.xform_ref_sm_to_gsm, o.L.orc_sm_to_gsm), (R.xform_ref_gsm_to_sm, o.L.orc_gsm_to_sm)):
            fn_o.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
            for x in xs:
                x = np.ascontiguousarray(x)
                fn_r(yearday, msec, x.ctypes.data, a.ctypes.data)
                fn_o(o.h, x.ctypes.data, b.ctypes.data)
                assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), (x, a, b)
    if srt_math:  # the product's host set-up: cos_mu, sin_mu, used as rotate_y(-+mu) by the device code (srt_rot_y)
        e = common.HostEmu("ngo", common.NGO_CFG, yearday, msec, 0)
        e.E.emu_model_get.argtypes = [C.c_void_p, C.c_void_p]
        g = np.zeros(4)
        e.E.emu_model_get(e.h, g.ctypes.data)
        c, s = g[1], g[2]
        for x in xs[:50]:
            x = np.ascontiguousarray(x)
            R.xform_ref_sm_to_gsm(yearday, msec, x.ctypes.data, a.ctypes.data)
            mine = np.array([x[0] * c + x[2] * (-s), x[1], x[2] * c - x[0] * (-s)])
            assert np.array_equal(a.view(np.uint64), mine.view(np.uint64)), (x, a, mine)
            R.xform_ref_gsm_to_sm(yearday, msec, x.ctypes.data, a.ctypes.data)
            mine = np.array([x[0] * c + x[2] * s, x[1], x[2] * c - x[0] * s])
            assert np.array_equal(a.view(np.uint64), mine.view(np.uint64))


def test_libxformd_entry_points_of_the_damping_code():
    """sm_to_mag_d_, geo_to_mag_d_, mag_to_geo_d_, cart_to_pol_d_, pol_to_cart_d_ (damping/include/coord_transforms.h:31-48): the
    reference's damping binary (oracle/_ref/damping) links the TRANSLATED xform_double (SM_TO_MAG, GEO_TO_MAG, MAG_TO_GEO, Cart_to_pol,
    POL_TO_CART, T0-T5, ...); the hand restatement oracle/ref_stubs/xform_d.cpp (used by tests/csrc/damping_probe.cpp) equals it bit
    for bit"""
    path = os.path.join(ol.ORC_DIR, "_ref", "libxformd_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libxformd_ref.so not built (reference sources absent)")
    ol.build()
    R, S = C.CDLL(path), C.CDLL(os.path.join(ol.ORC_DIR, "_build", "libxform_d_standin.so"))
    rng = np.random.default_rng(71)
    vp = C.c_void_p
    for yearday, msec in TIMES:
        itime = np.array([yearday, msec], dtype=np.int32)
        for x in rng.uniform(-5e7, 5e7, (100, 3)):
            x = np.ascontiguousarray(x)
            for name in ("sm_to_mag_d_", "geo_to_mag_d_", "mag_to_geo_d_"):
                a, b = np.zeros(3), np.zeros(3)
                for lib, out in ((R, a), (S, b)):
                    fn = getattr(lib, name)
                    fn.argtypes = [vp, vp, vp]
                    fn(itime.ctypes.data, x.ctypes.data, out.ctypes.data)
                assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), (name, yearday, msec)
            pa, pb = np.zeros(3), np.zeros(3)
            for lib, out in ((R, pa), (S, pb)):
                lib.cart_to_pol_d_.argtypes = [vp] * 4
                lib.cart_to_pol_d_(x.ctypes.data, out[0:].ctypes.data, out[1:].ctypes.data, out[2:].ctypes.data)
            assert np.array_equal(pa.view(np.uint64), pb.view(np.uint64))
            ca, cb = np.zeros(3), np.zeros(3)
            for lib, out in ((R, ca), (S, cb)):
                lib.pol_to_cart_d_.argtypes = [vp] * 4
                lib.pol_to_cart_d_(pa[0:].ctypes.data, pa[1:].ctypes.data, pa[2:].ctypes.data, out.ctypes.data)
            assert np.array_equal(ca.view(np.uint64), cb.view(np.uint64))
