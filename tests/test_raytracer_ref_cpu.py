"""The integrator itself (dispersion relation, Stix parameters, the three gradients, evalrhs, root
solve, stop conditions, rk4 / rk45, raytracer_run with its step-size control, cone refinement and first-step quirk) against
the REFERENCE'S OWN SOURCE: fortran/raytracer.f95 (+ util.f95, bmodel_dipole.f95, constants.f95) translated at build time
by oracle/f90_to_cpp.py over oracle/ref_stubs/f90rt.h into oracle/_ref/librt_ref.so (no Fortran compiler here).

The translation is syntactic; what a compiler decides about array expressions is stated in f90rt.h (element order of SUM and
DOT_PRODUCT, MAX with a NaN argument, CMPLX without KIND, COMPLEX arithmetic by the Fortran rules).  Two things are hooks:
funcPlasmaParams (the oracle's adapter: its parts are pinned by test_ngo_ref_cpu / test_geopack_ref_cpu / test_xform_ref_cpu /
test_ts05_cpu) and csvd (LAPACK zgesvd in the reference; the oracle's stand-in, see test_hostemu_cpu's handedness tests).
The oracle built on glibc's libm must reproduce the translated reference BIT FOR BIT: every row of every ray."""
import ctypes as C
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

GRID13 = os.path.join(common.ROOT, "tests", "golden", "grid13.txt")


@pytest.fixture(scope="module")
def ref():
    path = os.path.join(ol.ORC_DIR, "_ref", "librt_ref.so")
    if not os.path.exists(path):
        ol.ts05_ref(True)  # runs `make ref` where the reference sources exist
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librt_ref.so not built (reference sources absent)")
    R = C.CDLL(path)
    d, i, l, vp = C.c_double, C.c_int, C.c_long, C.c_void_p
    R.rt_ref_set_hooks.argtypes = [vp, vp]
    R.rt_ref_constant.restype = d
    R.rt_ref_evalrhs.argtypes = [vp, d, vp]
    R.rt_ref_solve.argtypes = [vp, d, vp, vp]
    R.rt_ref_bmodel_cartesian.argtypes = [vp, vp]
    R.rt_ref_cartesian_to_spherical.argtypes = [vp, vp]
    R.rt_ref_run.restype = l
    R.rt_ref_run.argtypes = [vp, vp, d, d, d, d, i, d, i, d, i, d, vp, l, vp]
    return R


def hook(R, o):
    L = o.L
    L.orc_hook_model.argtypes = [C.c_void_p]
    L.orc_hook_model(o.h)
    R.rt_ref_set_hooks(C.cast(L.orc_hook_plasma, C.c_void_p), C.cast(L.orc_hook_csvd, C.c_void_p))


def ref_run(R, o, pos0, dir0, w0, p):
    rows = np.zeros((p["maxsteps"] + 2, 20))
    sc, d0 = C.c_int(), np.array(dir0, dtype=np.float64)
    n = R.rt_ref_run(np.ascontiguousarray(pos0, dtype=np.float64).ctypes.data, d0.ctypes.data, w0, p["dt0"], p["dtmax"], p["maxerr"], p["maxsteps"], p["minalt"],
                     p["root"], p["tmax"], p["fixedstep"], o.del_, rows.ctypes.data, len(rows), C.byref(sc))
    return rows[:n], sc.value, d0


def assert_same(R, o, pos0, dir0, w0, p, what):
    t = o.trace(pos0, dir0, w0, **p)
    rows, sc, d0 = ref_run(R, o, pos0, dir0, w0, p)
    assert len(rows) == len(t["rows"]) and sc == t["stopcond"], (what, len(rows), len(t["rows"]), sc, t["stopcond"])
    assert np.array_equal(rows.view(np.uint64), t["rows"].view(np.uint64)), what
    assert np.array_equal(d0, t["dir0"]), what
    return t


def test_constants_and_pointwise_functions(ref):
    o = ol.OracleModel("ngo", common.NGO_CFG, libm=True)
    hook(ref, o)
    c = [ref.rt_ref_constant(i) for i in range(6)]
    assert c[3] == np.sqrt(1.0 / c[0] / c[2]) and c[4] == 6371.2e3
    rng = np.random.default_rng(2)
    pos0, _, w0 = common.c2_grid(5, 5)
    a, b, k4a, k4b = np.zeros(7), np.zeros(7), np.zeros(4), np.zeros(4)
    for r in range(len(w0)):
        x = np.ascontiguousarray(pos0[r] * rng.uniform(1.0, 3.0))
        kdir = rng.normal(size=3)
        kdir /= np.linalg.norm(kdir)
        o.L.orc_solve(o.h, kdir.ctypes.data, w0[r], x.ctypes.data, k4a.ctypes.data)
        ref.rt_ref_solve(kdir.ctypes.data, w0[r], x.ctypes.data, k4b.ctypes.data)
        assert np.array_equal(k4a.view(np.uint64), k4b.view(np.uint64))
        args = np.array(list(x) + list(kdir * k4a[2]) + [w0[r]])
        o.L.orc_evalrhs(o.h, args.ctypes.data, a.ctypes.data)
        ref.rt_ref_evalrhs(args.ctypes.data, o.del_, b.ctypes.data)
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def test_dipole_field_of_the_adapter(ref):
    """bmodel_cartesian (bmodel_dipole.f95) + util.f95's coordinate helpers: the oracle's field block with use_igrf = 0 at a time
    when the tilt chain is exercised too (adapter: dipole in SM -> GSM -> real*4 nT -> SM)"""
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 0, 0, libm=True)
    rng = np.random.default_rng(3)
    b = np.zeros(3)
    for _ in range(300):
        x = rng.uniform(-4e7, 4e7, 3)
        ref.rt_ref_bmodel_cartesian(x.ctypes.data, b.ctypes.data)
        got = o.plasma(x)[0][16:19]
        # the adapter rounds the field to real*4 nanotesla on the way (Q3): 6e-8 relative
        assert np.abs(got - b).max() <= 2e-7 * np.linalg.norm(b)


@pytest.mark.parametrize("use_igrf", [0, 1])
def test_whole_rays_ngo(ref, use_igrf):
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 3600000, use_igrf, libm=True)
    hook(ref, o)
    pos0, dir0, w0 = common.c2_grid(5, 5)
    stops = set()
    for r in range(len(w0)):
        t = assert_same(ref, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=500), "adaptive ray %d" % r)
        stops.add(t["stopcond"])
        if r % 4 == 0:
            assert_same(ref, o, pos0[r], [0, 0, 0], w0[r], dict(common.C1, maxsteps=150), "fixed-step ray %d" % r)
    assert len(stops) >= 2


def test_cone_refinement_before_the_first_acceptance_and_explicit_directions(ref):
    o = ol.OracleModel("ngo", common.NGO_CFG, libm=True)
    hook(ref, o)
    pos, dirs, ws, dts = common.cone_first_rays(o)
    ncone = 0
    for p0, d0, w, dt0 in zip(pos, dirs, ws, dts):
        t = assert_same(ref, o, p0, d0, w, dict(common.C2, dt0=dt0, maxsteps=120), "cone-first ray")
        ncone += t["ncone"]
        assert_same(ref, o, p0, d0, w, dict(common.C2, dt0=dt0, maxsteps=60, root=1), "root 1")
    assert ncone > 0


def test_whole_rays_gridded_model(ref):
    o = ol.OracleModel("grid", GRID13, 2010001, 0, 1, libm=True)
    hook(ref, o)
    pos0, dir0, w0 = common.launch_grid([1000.0, 6000.0], [25.0, 45.0])
    for r in range(len(w0)):
        assert_same(ref, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=200), "grid ray %d" % r)


def test_rays_with_ducts_and_the_external_field(ref):
    o = ol.OracleModel("ngo", os.path.join(common.ROOT, "tests", "data", "newray_ducts_sin_gauss.in"), libm=True)
    hook(ref, o)
    pos0, dir0, w0 = common.launch_grid([2000.0], [30.0, 50.0])
    for r in range(len(w0)):
        assert_same(ref, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=200), "ducted ray %d" % r)
    if ol.ts05_ref(False) is not None:
        o2 = ol.OracleModel("ngo", common.NGO_CFG, libm=True).tsyganenko([3.0, -45.0, 2.5, -4.0, 0.4, 0.7, 0.3, 0.9, 0.6, 1.1], srt_math=False)
        hook(ref, o2)
        assert_same(ref, o2, pos0[0], [0, 0, 0], w0[0], dict(common.C2, maxsteps=80), "TS05 ray")


# ---------------------------------------------------------------- the whole Ngo path from the reference's translated sources
PARMOD = np.array([3.0, -45.0, 2.5, -4.0, 0.4, 0.7, 0.3, 0.9, 0.6, 1.1])


def full_chain(R, cfg, yearday, msec, use_igrf, use_tsy):
    """funcPlasmaParams = the reference's Ngo adapter (ngo_dens_model_adapter.f95, translated in librt_ref) over the reference's
    ngo_dens_model.f95, xform_double, geopack2008.for and TS05 (each translated into its own library): nothing of the oracle but
    the stand-in for LAPACK's zgesvd is involved."""
    ref_dir = os.path.join(ol.ORC_DIR, "_ref")
    N, X, G, T = (C.CDLL(os.path.join(ref_dir, n)) for n in ("libngo_ref.so", "libxform_ref.so", "libgeopack_ref.so", "libts05_ref.so"))
    vp, d, i = C.c_void_p, C.c_double, C.c_int
    N.ngo_ref_readinput.argtypes = [C.c_char_p]
    N.ngo_ref_state.argtypes = [vp]
    R.rt_ref_ngo_adapter.argtypes = [vp] * 6 + [d, i, i, i, i, vp]
    R.rt_ref_ngo_plasma.argtypes = [C.c_long, vp, vp]
    R.rt_ref_use_ngo_adapter.argtypes = [vp]
    assert N.ngo_ref_readinput(cfg.encode()) == 0
    st = np.zeros(12)
    N.ngo_ref_state(st.ctypes.data)
    fp = lambda f: C.cast(f, vp)  # noqa: E731
    R.rt_ref_ngo_adapter(fp(X.xform_ref_sm_to_gsm), fp(X.xform_ref_gsm_to_sm), fp(G.geopack_ref_recalc), fp(G.geopack_ref_igrf_gsw), fp(T.ts05_ref_t04_s),
                         fp(N.ngo_ref_dens), st[2], yearday, msec, use_igrf, use_tsy, PARMOD.ctypes.data)
    return (N, X, G, T)  # keep the libraries alive


@pytest.mark.parametrize("use_igrf,use_tsy", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("cfg", [common.NGO_CFG, os.path.join(common.ROOT, "tests", "data", "newray_ducts_sin_gauss.in")], ids=["default", "ducts"])
def test_ngo_adapter_from_the_translated_sources(ref, cfg, use_igrf, use_tsy):
    keep = full_chain(ref, cfg, 2010001, 3600000, use_igrf, use_tsy)
    o = ol.OracleModel("ngo", cfg, 2010001, 3600000, use_igrf, libm=True)
    if use_tsy:
        o.tsyganenko(PARMOD, srt_math=False)
    rng = np.random.default_rng(4)
    r = ol.R_E * rng.uniform(1.1, 7.0, 400)
    th, ph = rng.uniform(0.2, np.pi - 0.2, 400), rng.uniform(-np.pi, np.pi, 400)
    x = np.ascontiguousarray(np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), r * np.cos(th)], axis=1))
    out = np.zeros((len(x), 19))
    ref.rt_ref_ngo_plasma(len(x), x.ctypes.data, out.ctypes.data)
    assert np.array_equal(o.plasma(x).view(np.uint64), out.view(np.uint64))
    del keep


@pytest.mark.parametrize("use_igrf,use_tsy", [(0, 0), (1, 0), (0, 1)])
def test_whole_rays_from_the_translated_sources_only(ref, use_igrf, use_tsy):
    """configs 1 and 2 (and the IGRF / TS05 variants), ray for ray: translated driver loop + integrator + adapter + density model +
    coordinate chain + Geopack (+ TS05) against the oracle, every row bit for bit"""
    keep = full_chain(ref, common.NGO_CFG, 2010001, 0, use_igrf, use_tsy)
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 0, use_igrf, libm=True)
    if use_tsy:
        o.tsyganenko(PARMOD, srt_math=False)
    ref.rt_ref_use_ngo_adapter(C.cast(o.L.orc_hook_csvd, C.c_void_p))
    pos0, dir0, w0 = common.c2_grid(3, 4)
    nsteps = 80 if use_tsy else 400
    for r in range(len(w0)):
        assert_same(ref, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=nsteps), "adaptive ray %d" % r)
    assert_same(ref, o, pos0[5], [0, 0, 0], w0[5], dict(common.C1, maxsteps=100), "config-1 ray")
    del keep


# ---------------------------------------------------------------- the gridded path (config 3) from the reference's translated sources
def grid_chain(o, yearday, msec, use_igrf, use_tsy):
    """funcPlasmaParams = the reference's interp adapter + libtricubic.f95 (tricubic_interpolate_at, tricubic_eval, the 64 x 64
    coefficient product), translated in librt_grid_ref, over the translated coordinate chain, Geopack and TS05.  The state arrays
    (f and its seven derivative arrays, filled by the adapter's `setup`: file reader + finite differences, not translated) are
    the oracle's."""
    ref_dir = os.path.join(ol.ORC_DIR, "_ref")
    path = os.path.join(ref_dir, "librt_grid_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/librt_grid_ref.so not built (reference sources absent)")
    R, X, G, T = (C.CDLL(os.path.join(ref_dir, n)) for n in ("librt_grid_ref.so", "libxform_ref.so", "libgeopack_ref.so", "libts05_ref.so"))
    vp, d, i, l = C.c_void_p, C.c_double, C.c_int, C.c_long
    dims, del3, ptrs = (i * 4)(), (d * 3)(), (vp * 13)()
    o.L.orc_grid_state.argtypes = [vp, vp, vp, vp]
    o.L.orc_grid_state(o.h, dims, del3, ptrs)
    arrays = (vp * 8)(*ptrs[5:13])
    fp = lambda f: C.cast(f, vp)  # noqa: E731
    R.rt_ref_grid_adapter.argtypes = [vp] * 5 + [i, i, i, i] + [vp] * 9
    R.rt_ref_grid_adapter(fp(X.xform_ref_sm_to_gsm), fp(X.xform_ref_gsm_to_sm), fp(G.geopack_ref_recalc), fp(G.geopack_ref_igrf_gsw), fp(T.ts05_ref_t04_s),
                          yearday, msec, use_igrf, use_tsy, PARMOD.ctypes.data, dims, del3, ptrs[0], ptrs[1], ptrs[2], ptrs[3], ptrs[4], arrays)
    R.rt_ref_ngo_plasma.argtypes = [l, vp, vp]
    R.rt_ref_use_ngo_adapter.argtypes = [vp]
    R.rt_ref_run.restype = l
    R.rt_ref_run.argtypes = [vp, vp, d, d, d, d, i, d, i, d, i, d, vp, l, vp]
    return R, (X, G, T, arrays)


@pytest.mark.parametrize("use_igrf,use_tsy", [(1, 0), (0, 0), (1, 1)])
def test_interp_adapter_and_libtricubic_from_the_translated_sources(use_igrf, use_tsy):
    o = ol.OracleModel("grid", GRID13, 2010001, 3600000, use_igrf, libm=True, faithful_cost=True)
    if use_tsy:
        o.tsyganenko(PARMOD, srt_math=False)
    R, keep = grid_chain(o, 2010001, 3600000, use_igrf, use_tsy)
    rng = np.random.default_rng(6)
    x = np.ascontiguousarray(rng.uniform(-4.3 * ol.R_E, 4.3 * ol.R_E, (3000, 3)))  # inside the 13^3 lattice, on its edge cells and outside
    x[:50] = np.round(x[:50] / 4247466.666666667) * 4247466.666666667             # on lattice nodes
    x = np.ascontiguousarray(x[np.linalg.norm(x, axis=1) > 0])                     # (the dipole field at the origin is a NaN on both sides)
    out = np.zeros((len(x), 19))
    R.rt_ref_ngo_plasma(len(x), x.ctypes.data, out.ctypes.data)
    a = o.plasma(x)
    assert np.array_equal(a.view(np.uint64), out.view(np.uint64))
    fast = ol.OracleModel("grid", GRID13, 2010001, 3600000, use_igrf, libm=True)  # the oracle's fast form too
    if use_tsy:
        fast.tsyganenko(PARMOD, srt_math=False)
    assert np.array_equal(fast.plasma(x).view(np.uint64), out.view(np.uint64))
    del keep


def test_whole_rays_through_the_translated_gridded_path():
    """config 3's path ray for ray: translated driver loop + integrator + interp adapter + libtricubic + coordinate chain + Geopack"""
    o = ol.OracleModel("grid", GRID13, 2010001, 0, 1, libm=True)
    R, keep = grid_chain(o, 2010001, 0, 1, 0)
    R.rt_ref_use_ngo_adapter(C.cast(o.L.orc_hook_csvd, C.c_void_p))
    pos0, dir0, w0 = common.launch_grid([800.0, 3000.0, 9000.0], [25.0, 40.0, 55.0])
    for r in range(len(w0)):
        assert_same(R, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=250), "grid ray %d" % r)
    del keep


def state_of(get, *head):
    """(dims, del3, 13 arrays) of a grid state: the oracle's (orc_grid_state) or the translated setup's (rt_ref_grid_state)"""
    vp, i, d = C.c_void_p, C.c_int, C.c_double
    dims, del3, ptrs = (i * 4)(), (d * 3)(), (vp * 13)()
    get.argtypes = [vp] * (len(head) + 3)
    get(*head, dims, del3, ptrs)
    ns, nx, ny, nz = dims
    sizes = [ns, ns, nx, ny, nz] + [ns * nx * ny * nz] * 8
    arrs = [np.ctypeslib.as_array(C.cast(ptrs[k], C.POINTER(d)), (sizes[k],)).copy() for k in range(13)]
    return list(dims), np.array(del3), arrs


def translated_setup(path, yearday=2010001, msec=0, use_igrf=1, use_tsy=0):
    ref_dir = os.path.join(common.ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "librt_grid_ref.so")):
        pytest.skip("oracle/_ref/librt_grid_ref.so not built (reference sources absent)")
    R, X, G, T = (C.CDLL(os.path.join(ref_dir, n)) for n in ("librt_grid_ref.so", "libxform_ref.so", "libgeopack_ref.so", "libts05_ref.so"))
    vp, i = C.c_void_p, C.c_int
    fp = lambda f: C.cast(f, vp)  # noqa: E731
    R.rt_ref_grid_setup.argtypes = [C.c_char_p] + [vp] * 5 + [i] * 4 + [vp]
    R.rt_ref_grid_setup(path.encode(), fp(X.xform_ref_sm_to_gsm), fp(X.xform_ref_gsm_to_sm), fp(G.geopack_ref_recalc), fp(G.geopack_ref_igrf_gsw),
                        fp(T.ts05_ref_t04_s), yearday, msec, use_igrf, use_tsy, PARMOD.ctypes.data)
    R.rt_ref_ngo_plasma.argtypes = [C.c_long, vp, vp]
    R.rt_ref_use_ngo_adapter.argtypes = [vp]
    R.rt_ref_run.restype = C.c_long
    d, l = C.c_double, C.c_long
    R.rt_ref_run.argtypes = [vp, vp, d, d, d, d, i, d, i, d, i, d, vp, l, vp]
    return R, (X, G, T)


NAMES13 = ["qs", "ms", "x", "y", "z", "f", "dfdx", "dfdy", "dfdz", "d2fdxdy", "d2fdxdz", "d2fdydz", "d3fdxdydz"]


def test_setup_of_the_interp_adapter_from_the_translated_source():
    """the grid file reader, the node coordinates and tricubic_compute_finite_difference_derivatives (a16's set-up half, row f3's
    derivative build): the oracle's state == the state the translated `setup` leaves, every array bit for bit"""
    o = ol.OracleModel("grid", GRID13, 2010001, 0, 1, libm=True)
    R, keep = translated_setup(GRID13)
    da, ea, A = state_of(o.L.orc_grid_state, o.h)
    db, eb, B = state_of(R.rt_ref_grid_state)
    assert da == db == [4, 13, 13, 13]
    assert np.array_equal(ea.view(np.uint64), eb.view(np.uint64))
    for name, a, b in zip(NAMES13, A, B):
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), name
    assert all(np.abs(a).max() > 0 for a in A[5:])


def test_setup_with_the_derivatives_in_the_file(tmp_path):
    """computederivatives = 1: the seven arrays are read from the file in array-element order, not computed"""
    o = ol.OracleModel("grid", GRID13, 2010001, 0, 1, libm=True)
    dims, del3, A = state_of(o.L.orc_grid_state, o.h)
    rng = np.random.default_rng(12)
    given = [a * rng.uniform(0.5, 1.5, a.shape) for a in A[6:]]  # not what finite differences would give
    path = str(tmp_path / "grid13_with_derivatives.txt")
    with open(path, "w") as f:
        f.write("1 %d %d %d %d\n" % tuple(dims))
        f.write(" ".join("%.17e" % v for v in [A[2][0], A[2][-1], A[3][0], A[3][-1], A[4][0], A[4][-1]]) + "\n")
        f.write(" ".join("%.17e" % v for v in A[0]) + "\n" + " ".join("%.17e" % v for v in A[1]) + "\n")
        for row in A[5].reshape(-1, dims[0]):
            f.write(" ".join("%.17e" % v for v in row) + "\n")
        for g in given:  # one record each (READ of a whole array), here broken over lines of 5 values
            for k in range(0, len(g), 5):
                f.write(" ".join("%.17e" % v for v in g[k:k + 5]) + "\n")
    o2 = ol.OracleModel("grid", path, 2010001, 0, 1, libm=True)
    R, keep = translated_setup(path)
    _, _, A2 = state_of(o2.L.orc_grid_state, o2.h)
    _, _, B2 = state_of(R.rt_ref_grid_state)
    for name, a, b in zip(NAMES13, A2, B2):
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), name
    for g, b in zip(given, B2[6:]):
        assert np.array_equal(g.view(np.uint64), b.view(np.uint64))


def test_whole_rays_from_the_translated_setup_on():
    """grid file -> translated setup -> translated funcPlasmaParams / libtricubic -> translated raytracer_run: nothing of the oracle on
    the reference side but the SVD stand-in; rows equal the oracle's bit for bit"""
    o = ol.OracleModel("grid", GRID13, 2010001, 0, 1, libm=True)
    R, keep = translated_setup(GRID13)
    R.rt_ref_use_ngo_adapter(C.cast(o.L.orc_hook_csvd, C.c_void_p))
    pos0, dir0, w0 = common.launch_grid([800.0, 3000.0, 9000.0], [25.0, 40.0, 55.0])
    for r in range(len(w0)):
        assert_same(R, o, pos0[r], [0, 0, 0], w0[r], dict(common.C2, maxsteps=250), "grid ray %d" % r)
    del keep


# ---- the last stand-in: LAPACK's zgesvd itself (scipy carries it) as the reference's csvd

def lapack_csvd_hook():
    """csvd (blas.f95:180-221) = zgesvd('A', 'A', ...) with the optimal workspace: scipy.linalg.lapack.zgesvd is that LAPACK routine
    (compute_uv = 1, full_matrices = 1 -> JOBU = JOBVT = 'A'), here as the hook the translated is_right_handed calls"""
    lapack = pytest.importorskip("scipy.linalg.lapack")
    calls = [0]

    @C.CFUNCTYPE(None, C.POINTER(C.c_double), C.POINTER(C.c_double))
    def cb(m_re_im, vt_re_im):
        a = np.ctypeslib.as_array(m_re_im, (18,)).copy()
        M = (a[0::2] + 1j * a[1::2]).reshape(3, 3, order="F")
        u, sv, vt, info = lapack.zgesvd(M, compute_uv=1, full_matrices=1)
        assert info == 0
        calls[0] += 1
        flat = np.asarray(vt).reshape(-1, order="F")
        out = np.ctypeslib.as_array(vt_re_im, (18,))
        out[0::2], out[1::2] = flat.real, flat.imag
    return cb, calls


def vendored_lapack():
    """oracle/_ref/liblapack_ref.so: zgesvd and the routines under it from the LAPACK 3.2.1 the reference vendors (lapack-3.2.1/) and
    links, translated by oracle/lapack_to_cpp.py"""
    path = os.path.join(ol.ORC_DIR, "_ref", "liblapack_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/liblapack_ref.so not built (reference sources absent)")
    L = C.CDLL(path)
    L.lapack_ref_csvd3.argtypes = [C.c_void_p] * 4
    L.lapack_ref_csvd3.restype = C.c_int
    return L


def test_translated_lapack_321_against_the_compiled_zgesvd():
    """the translation of the vendored zgesvd itself: singular values equal those of the image's compiled LAPACK (a later version)
    to the last digits, U S VT reconstructs the matrix, and every row of VT equals the compiled routine's up to the sign the two
    versions' Householder / rotation conventions leave open"""
    lapack = pytest.importorskip("scipy.linalg.lapack")
    L = vendored_lapack()
    rng = np.random.default_rng(31)
    signs = set()
    for k in range(1500):
        M = rng.normal(size=(3, 3)) * 10.0 ** rng.uniform(-3, 3) + 1j * rng.normal(size=(3, 3))
        if k % 3 == 0:
            M[:, 2] = M[:, 0] * (1 + 1e-9 * k) + 1e-8 * M[:, 1]  # nearly singular, as a dispersion matrix at its root
        flat = M.reshape(-1, order="F")
        a, s, u, vt = np.zeros(18), np.zeros(3), np.zeros(18), np.zeros(18)
        a[0::2], a[1::2] = flat.real, flat.imag
        assert L.lapack_ref_csvd3(a.ctypes.data, s.ctypes.data, u.ctypes.data, vt.ctypes.data) == 0
        U, VT = ((x[0::2] + 1j * x[1::2]).reshape(3, 3, order="F") for x in (u, vt))
        u2, s2, vt2, info = lapack.zgesvd(M, compute_uv=1, full_matrices=1)
        assert np.abs(s - s2).max() <= 4e-15 * s2[0]
        assert np.abs(U @ np.diag(s) @ VT - M).max() <= 1e-13 * s2[0]
        if s2[2] < 0.999 * s2[1] < 0.998 * s2[0]:  # distinct singular values: the vectors are defined up to a unit factor
            for i in range(3):
                p = np.vdot(vt2[i], VT[i])
                assert abs(abs(p) - 1) < 1e-7 and abs(p.imag) < 1e-7, (k, i, p)
                signs.add(int(np.sign(p.real)))
    assert signs == {1, -1}  # the two LAPACK versions do differ in sign conventions: the handedness decision must not (and does not) care


def test_handedness_with_the_vendored_lapack_under_the_translated_source(ref):
    """a11 end to end in the reference's own code: translated form_dispersion_matrix + is_right_handed (raytracer.f95:355-405)
    calling the translated zgesvd of the vendored LAPACK 3.2.1 == the oracle's handedness, on and off the roots"""
    V = vendored_lapack()
    L = ol.lib()
    ref.rt_ref_set_hooks(None, C.cast(V.lapack_ref_csvd_hook, C.c_void_p))
    d = C.c_double
    ref.rt_ref_is_right_handed.argtypes = [d] * 5
    L.orc_is_right_handed.argtypes = [d] * 5
    rng = np.random.default_rng(22)
    n = 0
    for k in range(20000):
        S, D, P = rng.uniform(-50, 50), rng.uniform(-200, 200), -10 ** rng.uniform(1, 6)
        phi = rng.uniform(0.0, np.pi / 2)
        if k % 3:
            Rr, Ll = S + D, S - D
            A = S * np.sin(phi) ** 2 + P * np.cos(phi) ** 2
            B = Rr * Ll * np.sin(phi) ** 2 + P * S * (1 + np.cos(phi) ** 2)
            disc = B * B - 4 * A * Rr * Ll * P
            if disc < 0 or A == 0:
                continue
            n2 = (B + (1 if k % 2 else -1) * np.sqrt(disc)) / (2 * A)
        else:
            n2 = 10 ** rng.uniform(-1, 4)
        a, b = ref.rt_ref_is_right_handed(n2, phi, S, D, P), L.orc_is_right_handed(n2, phi, S, D, P)
        if a != b:
            sv = np.linalg.svd(_dispersion_matrix(n2, phi, S, D, P), compute_uv=False)
            assert sv[2] > 0.9 * sv[1] or sv[1] == 0, (n2, phi, S, D, P, sv)
        n += 1
    assert n > 15000


def test_whole_rays_in_the_references_own_code_only(ref):
    """Ngo input file -> translated readinput / adapter / density model / coordinate chain / Geopack / integrator / vendored LAPACK:
    every line executed on the reference side is the reference's (translated).  Rows equal the oracle's bit for bit, both roots."""
    V = vendored_lapack()
    keep = full_chain(ref, common.NGO_CFG, 2010001, 0, 1, 0)
    ref.rt_ref_use_ngo_adapter(C.cast(V.lapack_ref_csvd_hook, C.c_void_p))
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 0, 1, libm=True)
    pos0, dir0, w0 = common.c2_grid(3, 3)
    for root in (1, 2):
        for k in range(len(w0)):
            assert_same(ref, o, pos0[k], dir0[k], w0[k], dict(common.C2, maxsteps=80, root=root), "root %d ray %d" % (root, k))
    del keep


def test_handedness_with_lapack_zgesvd_under_the_translated_source(ref):
    """a11: translated form_dispersion_matrix + is_right_handed (raytracer.f95:355-405) calling LAPACK's zgesvd == the oracle's
    stand-in (complex one-sided Jacobi), on the matrices the path meets: n2 a root of the dispersion relation (the null space is
    one-dimensional there) for random Stix parameters and angles, and off-root n2 as well"""
    cb, calls = lapack_csvd_hook()
    L = ol.lib()
    ref.rt_ref_set_hooks(None, C.cast(cb, C.c_void_p))
    d = C.c_double
    ref.rt_ref_is_right_handed.argtypes = [d] * 5
    L.orc_is_right_handed.argtypes = [d] * 5
    rng = np.random.default_rng(21)
    n, skipped = 0, 0
    for k in range(3000):
        S, D, P = rng.uniform(-50, 50), rng.uniform(-200, 200), -10 ** rng.uniform(1, 6)
        phi = rng.uniform(0.0, np.pi / 2) if k % 2 else rng.uniform(0, 90)  # the caller passes radians; larger arguments exercise the same code
        if k % 3:  # a root of A n^4 - B n^2 + R L P = 0 (raytracer.f95:466-478)
            Rr, Ll = S + D, S - D
            A = S * np.sin(phi) ** 2 + P * np.cos(phi) ** 2
            B = Rr * Ll * np.sin(phi) ** 2 + P * S * (1 + np.cos(phi) ** 2)
            disc = B * B - 4 * A * Rr * Ll * P
            if disc < 0 or A == 0:
                continue
            n2 = (B + (1 if k % 2 else -1) * np.sqrt(disc)) / (2 * A)
        else:
            n2 = 10 ** rng.uniform(-1, 4)
        a, b = ref.rt_ref_is_right_handed(n2, phi, S, D, P), L.orc_is_right_handed(n2, phi, S, D, P)
        if a != b:  # only where the smallest singular value is not isolated is the vector (and so the answer) undefined
            sv = np.linalg.svd(_dispersion_matrix(n2, phi, S, D, P), compute_uv=False)
            assert sv[2] > 0.9 * sv[1] or sv[1] == 0, (n2, phi, S, D, P, sv)
            skipped += 1
        n += 1
    print("handedness with LAPACK zgesvd: %d matrices, %d with a degenerate null space" % (n, skipped))
    assert n > 2000 and skipped < 0.01 * n and calls[0] >= n


def _dispersion_matrix(n2, phi, S, D, P):
    s, c = np.sin(phi), np.cos(phi)
    return np.array([[S - n2 * c * c, -1j * D, n2 * c * s], [1j * D, S - n2, 0], [n2 * c * s, 0, P - n2 * s * s]])


def test_whole_rays_with_lapack_zgesvd_and_translated_sources_only(ref):
    """Ngo input file -> translated readinput / adapter / density model / coordinate chain / Geopack / integrator, LAPACK's zgesvd
    as csvd: NOTHING of the oracle on the reference side.  Rows equal the oracle's bit for bit, both roots."""
    cb, calls = lapack_csvd_hook()
    keep = full_chain(ref, common.NGO_CFG, 2010001, 0, 1, 0)
    ref.rt_ref_use_ngo_adapter(C.cast(cb, C.c_void_p))
    o = ol.OracleModel("ngo", common.NGO_CFG, 2010001, 0, 1, libm=True)
    pos0, dir0, w0 = common.c2_grid(3, 3)
    for root in (1, 2):
        for k in range(len(w0)):
            assert_same(ref, o, pos0[k], dir0[k], w0[k], dict(common.C2, maxsteps=80, root=root), "root %d ray %d" % (root, k))
    assert calls[0] >= 2 * len(w0)
    del keep
