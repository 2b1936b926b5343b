"""modelnum 4 (scattered samples, moving least squares) on the CPU: the oracle's restatement of the reference
(oracle/orc_scattered.h: kd-tree, lsinterp, dposv as the Fortran states them) against known answers, and the product's
set-up and device code (csrc/srt_scattered*.h, compiled for the host by tests/csrc/hostemu.cpp) against the oracle.

What can be equal and what cannot: the reference inserts its samples in the order of an unseeded random permutation
(fortran/scattered_interp_dens_model_adapter.f95:141), so the ORDER in which a radius query returns its neighbours is not
defined by the reference's source; the SET is.  Hence:
  * neighbour sets: equal, exactly;
  * lsinterp given the product's neighbour order: equal bit for bit;
  * end to end (each side in its own order): equal to the rounding of sums taken in another order -- the tolerance below is
    measured against the difference between two insertion orders of the oracle itself.
"""
import os

import numpy as np
import pytest

import common
import oracle_lib as ol

R_E = common.R_E
OPTS = dict(window_scale=1.6, order=2, exact=0, local_window_scale=3.0)


@pytest.fixture(scope="module")
def sample_file(tmp_path_factory):
    path = str(tmp_path_factory.mktemp("scat") / "samples.txt")
    pos, logn = common.write_scattered_file(path, 3000, seed=3, duplicates=7, inside=4)
    return path, pos, logn


@pytest.fixture(scope="module")
def models(sample_file):
    path = sample_file[0]
    return common.HostEmuScattered(path, **OPTS), ol.OracleScattered(path, **OPTS)


def query_points(n, seed=11, rmin=1.2, rmax=4.5):
    rng = np.random.default_rng(seed)
    u = rng.normal(size=(n, 3))
    u /= np.linalg.norm(u, axis=1)[:, None]
    return R_E * (rmin + (rmax - rmin) * rng.random(n))[:, None] * u


def test_polynomial_reproduction_of_the_oracle(tmp_path):
    """Moving least squares of degree p reproduces every polynomial of degree <= p exactly (Levin): a known answer that owes
    nothing to the reference's source.  Samples of a quadratic in a box outside the Earth, orders 2 and 3, both windows."""
    rng = np.random.default_rng(5)
    n = 1500
    pos = np.array([3.0, 0.0, 0.0]) * R_E + (rng.random((n, 3)) - 0.5) * 0.6 * R_E
    s = (pos - pos.mean(0)) / (0.3 * R_E)
    f = 1.0 + 0.5 * s[:, 0] - 0.25 * s[:, 1] + 0.1 * s[:, 2] + 0.3 * s[:, 0] * s[:, 1] - 0.2 * s[:, 2] ** 2
    path = str(tmp_path / "quad.txt")
    with open(path, "w") as fh:
        fh.write("1 0 0 0 0 0 0\n-1.602e-19\n9.10938188e-31\n")
        for p, v in zip(pos, f):
            fh.write("%.17e %.17e %.17e %.17e\n" % (p[0], p[1], p[2], v))
    q = np.array([3.0, 0.0, 0.0]) * R_E + (rng.random((40, 3)) - 0.5) * 0.3 * R_E
    sq = (q - pos.mean(0)) / (0.3 * R_E)
    want = 1.0 + 0.5 * sq[:, 0] - 0.25 * sq[:, 1] + 0.1 * sq[:, 2] + 0.3 * sq[:, 0] * sq[:, 1] - 0.2 * sq[:, 2] ** 2
    for order in (2, 3):
        for exact in (0, 1):
            om = ol.OracleScattered(path, window_scale=2.5, order=order, exact=exact, local_window_scale=3.0)
            he = common.HostEmuScattered(path, window_scale=2.5, order=order, exact=exact, local_window_scale=3.0)
            for x, w in zip(q, want):
                st, fi = om.interpolate(x)
                assert st == 0
                assert abs(fi[0] - w) < 2e-7, (order, exact, fi[0], w)
                st2, fi2 = he.lsinterp(x)
                assert st2 == 0 and abs(fi2[0] - w) < 2e-7
            om.close()
    # order 1 does not reproduce a quadratic: the test above is not vacuous
    om = ol.OracleScattered(path, window_scale=2.5, order=1, exact=0, local_window_scale=3.0)
    assert max(abs(om.interpolate(x)[1][0] - w) for x, w in zip(q, want)) > 1e-3
    om.close()


def test_setup_equals_the_oracle(sample_file, models):
    """samples kept (repeated positions dropped), every sample's nearest distance, maxnearest: product's cell lists == kd-tree"""
    path, pos, logn = sample_file
    he, om = models
    assert om.nduplicates == 7 and he.npts == om.nkept == 3000
    assert he.maxnearest == om.maxnearest and he.radius == om.radius
    # brute force over the file's distinct positions
    p = pos[:3000]
    d2 = ((p[:, None, :] - p[None, :, :]) ** 2)
    d2 = (d2[:, :, 0] + d2[:, :, 1]) + d2[:, :, 2]
    np.fill_diagonal(d2, np.inf)
    nearest = np.sqrt(d2.min(1))
    outside = (p[:, 0] * p[:, 0] + p[:, 1] * p[:, 1]) + p[:, 2] * p[:, 2] >= R_E * R_E
    assert outside.sum() == 2996
    assert he.maxnearest == nearest[outside].max()
    rec = he.records
    key = {tuple(v): i for i, v in enumerate(p)}
    idx = np.array([key[tuple(v)] for v in rec[:, :3]])
    assert len(set(idx)) == 3000
    assert np.array_equal(rec[:, 3:7], logn[idx])  # of two samples at one position the first in the file stays
    assert np.array_equal(rec[:, 7], np.where(outside[idx], nearest[idx], 1.0))


def test_root_quirk_of_the_reference_is_restated(sample_file):
    """kdtree_nearest starts from best = root: the root sample's own nearest distance comes out 0 (kdtree_mod.f95:397-418).
    The oracle restates it (root_quirk=True); the product and the comparisons above use the true distance."""
    path = sample_file[0]
    perm = np.r_[10, 0:10, 11:3007]  # (samples 0-3 of the file lie inside the Earth and get no nearest distance at all)
    q = ol.OracleScattered(path, root_quirk=True, perm=perm, **OPTS)
    t = ol.OracleScattered(path, root_quirk=False, perm=perm, **OPTS)
    assert q.root_nearest == 0.0 and t.root_nearest > 1.0e3
    assert q.maxnearest == t.maxnearest  # (unless the root happens to hold the largest distance)
    q.close(); t.close()


def test_neighbour_sets_equal_and_lsinterp_bitwise_in_the_same_order(models):
    he, om = models
    xs = query_points(300)
    nb_counts = []
    for x in xs:
        mine = he.neighbours(x)
        theirs = om.search(x, 4)
        nb_counts.append(len(mine))
        a = he.records[mine]
        # the same set (compared as sorted rows), not the same order
        assert len(a) == len(theirs)
        ka = a[np.lexsort(a[:, :3].T[::-1])]
        kb = theirs[np.lexsort(theirs[:, :3].T[::-1])]
        assert np.array_equal(ka, kb)
        # the oracle's lsinterp fed the product's order == the product's device code, bit for bit
        st_o, fi_o = om.lsinterp_found(x, a[:, :3], a[:, 3:8])
        st_p, fi_p = he.lsinterp(x)
        assert st_o == st_p
        assert np.array_equal(fi_o, fi_p), (fi_o, fi_p)
    assert min(nb_counts) >= 10 and max(nb_counts) > 100  # sparse and dense neighbourhoods both occur


@pytest.mark.parametrize("order,exact,scaled", [(0, 0, 0), (1, 1, 0), (2, 1, 0), (3, 0, 0), (2, 0, 1), (3, 1, 1)])
def test_every_option_bitwise_in_the_same_order(sample_file, order, exact, scaled):
    path = sample_file[0]
    kw = dict(window_scale=1.8, order=order, exact=exact, local_window_scale=2.5, scaled=scaled)
    he, om = common.HostEmuScattered(path, **kw), ol.OracleScattered(path, **kw)
    for x in query_points(60, seed=order * 7 + exact):
        a = he.records[he.neighbours(x)]
        st_o, fi_o = om.lsinterp_found(x, a[:, :3], a[:, 3:8])
        st_p, fi_p = he.lsinterp(x)
        assert st_o == st_p and np.array_equal(fi_o, fi_p)
    om.close()


def test_failures_as_the_reference(sample_file):
    """too few points in the sphere: status 2, fi = 0, density exp(0) = 1 (adapter :247-262); inside the Earth: density 0"""
    path = sample_file[0]
    kw = dict(window_scale=0.05, order=2, exact=0, local_window_scale=3.0)
    he, om = common.HostEmuScattered(path, **kw), ol.OracleScattered(path, **kw)
    x = np.array([[4.4 * R_E, 0.3 * R_E, 0.1 * R_E], [0.5 * R_E, 0.0, 0.0], [np.inf, 0.0, 0.0], [np.nan, 1.0, 1.0]])
    st, fi = he.lsinterp(x[0])
    assert st == 2 and om.interpolate(x[0])[0] == 2 and not fi.any()
    a, b = he.plasma(x), om.plasma(x)
    assert np.array_equal(a[:, 4:8], b[:, 4:8], equal_nan=True)
    assert np.array_equal(a[0, 4:8], np.ones(4)) and not a[1, 4:8].any() and np.array_equal(a[2, 4:8], np.ones(4)) and not a[3, 4:8].any()
    om.close()


def test_end_to_end_plasma_against_the_oracle(sample_file, models):
    """funcPlasmaParams, each side in its own neighbour order.  The yardstick for the tolerance is the reference's own
    indeterminacy: the oracle under two insertion orders."""
    path = sample_file[0]
    he, om = models
    rng = np.random.default_rng(2)
    om2 = ol.OracleScattered(path, perm=np.r_[rng.permutation(3000), 3000:3007], **OPTS)  # (the repeated positions last: the same survivors)
    xs = query_points(400, seed=21)
    a, b, c = he.plasma(xs), om.plasma(xs), om2.plasma(xs)
    assert np.array_equal(a[:, 16:19], b[:, 16:19])  # the field block is the adapters' common one: bit for bit
    assert np.array_equal(a[:, 0:4], b[:, 0:4]) and np.array_equal(a[:, 8:12], b[:, 8:12])
    rel_self = np.abs(c[:, 4:8] / b[:, 4:8] - 1.0).max()
    rel = np.abs(a[:, 4:8] / b[:, 4:8] - 1.0).max()
    print("relative difference of Ns: product vs oracle %.3g; oracle vs oracle under another insertion order %.3g" % (rel, rel_self))
    assert rel < 1e-9 and rel < 50 * max(rel_self, 1e-15)
    # and the interpolant is the Ngo model the samples came from, to the accuracy 3000 samples allow
    ngo = ol.OracleModel("ngo", common.NGO_CFG).plasma(xs)
    err = np.abs(np.log(a[:, 4] / ngo[:, 4]))
    assert np.median(err) < 0.15
    om2.close()


def test_whole_rays_bitwise_given_the_neighbour_order(sample_file, models):
    """raytracer_run over modelnum 4.  With the one thing the reference leaves open -- the order of a query's neighbours --
    set to the product's (OracleScattered.impose_order), the oracle's kd-tree + lsinterp + integrator and the product's cell
    list + device code give the same rows bit for bit: adaptive and fixed step, IGRF and dipole."""
    path = sample_file[0]
    he = models[0]
    pos0, dir0, w0 = common.launch_grid([2000.0, 8000.0], [25.0, 40.0], alt_km=3000.0)
    for use_igrf, fixed in ((0, 0), (1, 0), (0, 1)):
        h = common.HostEmuScattered(path, use_igrf=use_igrf, **OPTS)
        om = ol.OracleScattered(path, use_igrf=use_igrf, **OPTS).impose_order(h.records[:, :3])
        xs = query_points(100, seed=33)
        assert np.array_equal(h.plasma(xs), om.plasma(xs))
        prm = dict(common.C2, maxsteps=60, tmax=0.5, fixedstep=fixed)
        steps = 0
        for r in range(len(w0)):
            o = om.trace(pos0[r], [0, 0, 0], w0[r], **prm)
            g = h.trace(pos0[r], [0, 0, 0], w0[r], **prm)
            assert g["stopcond"] == o["stopcond"] and g["rows"].shape == o["rows"].shape
            assert np.array_equal(g["rows"], o["rows"], equal_nan=True)
            assert g["nrejected"] == o["nrejected"] and g["ncone"] == o["ncone"]
            steps += len(g["rows"])
        assert steps > 100
        om.close()


def test_whole_rays_spread_of_the_reference_itself(sample_file, models):
    """Without that: the normal equations of lsinterp are formed from monomials in metres (entries 1 ... 1e24), a fit keeps
    about 8 digits, and dF/dx differences it over del = 1e-6: the reference's own rays move by per cent between two of its
    insertion orders after a few dozen steps.  The product, in its order, stays within that spread."""
    path = sample_file[0]
    he, om = models
    om2 = ol.OracleScattered(path, perm=np.r_[np.random.default_rng(2).permutation(3000), 3000:3007], **OPTS)
    pos0, dir0, w0 = common.launch_grid([2000.0, 8000.0], [25.0, 40.0], alt_km=3000.0)
    prm = dict(common.C2, maxsteps=12, tmax=0.5, fixedstep=1)
    mine, own = 0.0, 0.0
    for r in range(len(w0)):
        o = om.trace(pos0[r], [0, 0, 0], w0[r], **prm)
        o2 = om2.trace(pos0[r], [0, 0, 0], w0[r], **prm)
        g = he.trace(pos0[r], [0, 0, 0], w0[r], **prm)
        assert g["rows"].shape == o["rows"].shape == o2["rows"].shape
        scale = np.abs(o["rows"][:, 1:4]).max() + 1e-300
        mine = max(mine, np.abs(g["rows"][:, 1:4] - o["rows"][:, 1:4]).max() / scale)
        own = max(own, np.abs(o2["rows"][:, 1:4] - o["rows"][:, 1:4]).max() / scale)
    print("positions after 12 fixed steps: product vs oracle %.3g, oracle vs oracle under another insertion order %.3g" % (mine, own))
    assert mine < 1e-3 and own < 1e-3
    om2.close()


def test_file_reader_follows_list_directed_rules(tmp_path):
    """commas, D exponents, records continuing over lines, and the list ending at the first bad READ (adapter :107-110)"""
    path = str(tmp_path / "odd.txt")
    with open(path, "w") as fh:
        fh.write("2, 0 0 0 0 0 0\n-1.602D-19 1.602d-19\n9.1e-31,\n1.67e-27\n")
        for i in range(30):
            fh.write("%.6e,%.6e\n%.6e %.3f %.3f\n" % (2e7 + 1e5 * i, 1e5 * (i % 5), 2e5 * (i % 3), 20.0 + 0.01 * i, 19.0))
    clean = str(tmp_path / "clean.txt")
    with open(clean, "w") as fh:
        fh.write(open(path).read())
    with open(path, "a") as fh:
        fh.write("oops 1 2 3 4\n3e7 0 0 1 1\n")  # iostat /= 0 ends the list: the record after it is never read
    he = common.HostEmuScattered(path, window_scale=30.0, order=1, exact=0, local_window_scale=3.0)
    om = ol.OracleScattered(clean, window_scale=30.0, order=1, exact=0, local_window_scale=3.0)  # (the oracle's reader knows clean files only)
    assert he.npts == om.nkept == 30 and he.maxnearest == om.maxnearest
    out = he.plasma(np.array([[2.15e7, 2e5, 2e5]]))
    assert out[0, 0] == -1.602e-19 and out[0, 1] == 1.602e-19 and out[0, 9] == 1.67e-27 and out[0, 6] == 0.0
    om.close()


def test_with_the_ts05_external_field(sample_file):
    """use_tsyganenko = 1 on modelnum 4 (adapter :352-360): the field block is the other adapters'; T04_s is the reference's
    own source as translated into oracle/_ref"""
    from test_ts05_cpu import PARMOD
    if ol.ts05_ref(True) is None:
        pytest.skip("oracle/_ref/libts05_ref_srtmath.so not built")
    path = sample_file[0]
    for use_igrf in (0, 1):
        he = common.HostEmuScattered(path, use_igrf=use_igrf, msec=3600000, **OPTS).tsyganenko(PARMOD)
        om = ol.OracleScattered(path, use_igrf=use_igrf, msec=3600000, **OPTS).impose_order(he.records[:, :3]).tsyganenko(PARMOD)
        off = common.HostEmuScattered(path, use_igrf=use_igrf, msec=3600000, **OPTS)
        x = query_points(150, seed=55)
        a, b, c = he.plasma(x), om.plasma(x), off.plasma(x)
        assert np.array_equal(a, b)
        assert np.array_equal(a[:, :16], c[:, :16]) and not np.array_equal(a[:, 16:19], c[:, 16:19])
        om.close()


def _dposv_ref():
    import ctypes as C
    path = os.path.join(common.ROOT, "oracle", "_ref", "libdposv_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref/libdposv_ref.so not built (the reference's lapack-3.2.1 is translated where /root/reference exists)")
    return C.CDLL(path)


def test_blas_and_lapack_calls_against_the_vendored_lapack():
    """lsinterp's three library calls (blas.f95: dposv 'U', dgemm 'T','N', dgemv 'N') as the oracle restates them, against the
    reference's own vendored lapack-3.2.1, translated statement by statement into oracle/_ref/libdposv_ref.so
    (oracle/lapack_to_cpp.py): bit for bit, incl. matrices built like lsinterp's (monomials in metres), a matrix that is not
    positive definite (same INFO, same partial factor) and zeros in the right-hand side (dtrsm / dgemv skip them)."""
    import ctypes as C
    R, L = _dposv_ref(), ol.lib()
    vp, i = C.c_void_p, C.c_int
    for lib_, names in ((R, ("lapack_ref_dposv_upper", "lapack_ref_dgemm_tn", "lapack_ref_dgemv_n")), (L, ("orc_dposv_upper", "orc_mat_trans_mat_mult", "orc_mat_vec_mult"))):
        getattr(lib_, names[0]).argtypes = [i, vp, vp]
        getattr(lib_, names[0]).restype = i
        getattr(lib_, names[1]).argtypes = [i, i, i, vp, vp, vp]
        getattr(lib_, names[2]).argtypes = [i, i, vp, vp, vp]
    rng = np.random.default_rng(8)
    p = lambda a: a.ctypes.data_as(vp)
    cases = 0
    for J in (1, 4, 10, 20):
        for trial in range(12):
            I = J + int(rng.integers(0, 60))
            scale = 10.0 ** rng.uniform(0, 6)
            E = np.asfortranarray(rng.normal(size=(I, J)) * scale ** rng.integers(0, 3, size=J)[None, :])
            if trial == 0:
                E[:, -1] = E[:, 0]  # singular: not positive definite (or barely), INFO > 0 on both sides or the same factor
            A1, A2 = np.zeros((J, J), order="F"), np.zeros((J, J), order="F")
            R.lapack_ref_dgemm_tn(J, J, I, p(E), p(E), p(A1))
            L.orc_mat_trans_mat_mult(J, J, I, p(E), p(E), p(A2))
            assert np.array_equal(A1, A2)
            b1 = np.zeros(J); b1[0] = 1.0
            if trial % 3 == 2:
                b1 = rng.normal(size=J) * (rng.random(J) > 0.3)
            b2 = b1.copy()
            F1, F2 = A1.copy(order="F"), A2.copy(order="F")
            i1, i2 = R.lapack_ref_dposv_upper(J, p(F1), p(b1)), L.orc_dposv_upper(J, p(F2), p(b2))
            assert i1 == i2
            assert np.array_equal(np.triu(F1), np.triu(F2), equal_nan=True)
            if i1 == 0:
                assert np.array_equal(b1, b2, equal_nan=True)
            x = rng.normal(size=J) * (rng.random(J) > 0.2)
            y1, y2 = np.full(I, 7.0), np.full(I, 7.0)
            R.lapack_ref_dgemv_n(I, J, p(E), p(x), p(y1))
            L.orc_mat_vec_mult(I, J, p(E), p(x), p(y2))
            assert np.array_equal(y1, y2)
            cases += 1
    assert cases == 48


def test_second_statement_in_numpy(sample_file, models):
    """lsinterp said a second time, from the MATLAB lines its author left in the Fortran as comments (lsinterp_mod.f95:199-201:
    `etainv = @(r) ((1+eps)./(exp(r.^2./h.^2)-1+eps)).*(.5+.5*cos(r*2*pi/s/2)).*(r<s)`; :388 `a = (Dinv)*E*(A\\c)`) and from
    Levin's definition: the interpolant at x is the value at x of the polynomial of degree <= order that minimises
    sum_i w_i (p(x_i) - f_i)^2 with w_i = 0.5 etainv(r_i).  Solved here by a weighted least-squares fit in coordinates scaled
    by the radius (QR, not the normal equations), on brute-force neighbour sets: agreement to 1e-7 in the log-density."""
    path = sample_file[0]
    he = models[0]
    rec, eps, worst = he.records, 5.0e-16, 0.0
    for exact in (0, 1):
        om = ol.OracleScattered(path, **dict(OPTS, exact=exact))
        for x in query_points(40, seed=90 + exact):
            d = rec[:, :3] - x
            r = np.sqrt((d * d).sum(1))
            nb = r < he.radius
            d, r, vals, near = d[nb], r[nb], rec[nb, 3:7], rec[nb, 7]
            cw = 0.5 + 0.5 * np.cos(r * 2 * np.pi / he.radius / 2)
            h = OPTS["local_window_scale"] * (cw * near).sum() / cw.sum()
            if exact:
                eta = ((1 + eps) / (np.exp(r ** 2 / h ** 2) - 1 + eps)) * cw
            else:
                eta = np.exp(-((r + he.radius * eps) / (h / 4)) ** 1.1) * cw
            keep = eta > 1e-16
            if keep.sum() < 10:
                keep[:] = True
            s = d[keep] / he.radius
            V = np.stack([s[:, 0] ** a * s[:, 1] ** b * s[:, 2] ** c for a in range(3) for b in range(3 - a) for c in range(3 - a - b)], axis=1)
            sw = np.sqrt(0.5 * eta[keep])
            coef = np.linalg.lstsq(V * sw[:, None], vals[keep] * sw[:, None], rcond=None)[0]
            want = coef[0]  # the polynomial's value at the query point: every monomial but the constant vanishes there
            st, fi = om.interpolate(x)
            assert st == 0
            worst = max(worst, np.abs(fi - want).max())
        om.close()
    print("lsinterp against the numpy second statement: largest difference of a log-density %.3g" % worst)
    assert worst < 1e-7


def test_degenerate_geometry_bitwise(tmp_path):
    """Paths the shell of random samples never takes: samples on a plane and on a line (singular normal equations: dposv
    reports INFO > 0 -> status 1, fi = 0, density 1, lsinterp_mod.f95:381-386 -- or a factor that is only rounding), a query that
    coincides with a sample (r = 0: the exact window's weight is (1+eps)/eps), a neighbourhood cut by the weight threshold
    to fewer than J points (mask reset to all, :333-337).  Product (host build of the device code) == oracle bit for bit in
    the product's neighbour order, statuses included."""
    rng = np.random.default_rng(4)
    n = 600
    base = np.array([2.5, 0.3, 0.2]) * R_E
    span = 0.4 * R_E
    clouds = {
        "plane": base + np.c_[rng.random(n) - 0.5, rng.random(n) - 0.5, np.zeros(n)] * span,
        "line": base + np.c_[rng.random(n) - 0.5, np.zeros(n), np.zeros(n)] * span,
        "lattice": base + np.stack(np.meshgrid(*[np.arange(8) - 3.5] * 3, indexing="ij"), -1).reshape(-1, 3) * span / 8,
        "cluster+far": np.vstack([base + (rng.random((40, 3)) - 0.5) * 0.01 * span, base + (rng.random((200, 3)) - 0.5) * span]),
    }
    seen = set()
    for name, pos in clouds.items():
        f = np.sin(pos[:, 0] / span) + 0.3 * np.cos(pos[:, 1] / span) + 20.0
        path = str(tmp_path / (name.replace("+", "_") + ".txt"))
        with open(path, "w") as fh:
            fh.write("2 0 0 0 0 0 0\n-1.602e-19 1.602e-19\n9.10938188e-31 1.6726e-27\n")
            for p, v in zip(pos, f):
                fh.write("%.17e %.17e %.17e %.17e %.17e\n" % (p[0], p[1], p[2], v, v - 1.0))
        for order, exact, lws in ((1, 0, 3.0), (2, 1, 3.0), (3, 0, 0.05), (2, 1, 0.02)):
            kw = dict(window_scale=3.0 if name != "cluster+far" else 1.2, order=order, exact=exact, local_window_scale=lws)
            he = common.HostEmuScattered(path, **kw)
            om = ol.OracleScattered(path, **kw).impose_order(he.records[:, :3])
            assert he.maxnearest == om.maxnearest
            q = np.vstack([base + (rng.random((25, 3)) - 0.5) * span, pos[:5], pos[:3] + 1e-3])
            for x in q:
                a = he.records[he.neighbours(x)]
                st_o, fi_o = om.lsinterp_found(x, a[:, :3], np.c_[a[:, 3:5], a[:, 7]])  # two species, then the nearest distance
                st_p, fi_p = he.lsinterp(x)
                assert st_o == st_p and np.array_equal(fi_o, fi_p, equal_nan=True), (name, kw, x, st_o, st_p, fi_o, fi_p)
                seen.add((name, st_p))
            assert np.array_equal(he.plasma(q), om.plasma(q), equal_nan=True)
            om.close()
    print(sorted(seen))
    assert {s for _, s in seen} >= {0, 1}  # both the solved and the failed factorisation occur


class ReferenceLsinterp:
    """`lsinterp` of fortran/lsinterp_mod.f95 EXECUTED from the reference's own source text by oracle/f90_interp.py (with
    tabular_monomials, etainv, coswindow of the same file and SolveBlasSymmetric / MatTransMatMultBlas / MatVecMultBlas of
    fortran/blas.f95, interpreted too); dposv / dgemm / dgemv underneath are the reference's vendored lapack-3.2.1 as translated
    into oracle/_ref/libdposv_ref.so; kdtree_search is a hook that returns the neighbour arrays the test supplies."""

    def __init__(self):
        import ctypes as C
        import re
        import sys
        ref = "/root/reference/fortran"
        if not os.path.exists(ref + "/lsinterp_mod.f95"):
            pytest.skip("the reference's sources are not present on this machine")
        sys.path.insert(0, os.path.join(common.ROOT, "oracle"))
        import f90_interp as F
        R = _dposv_ref()
        vp, i = C.c_void_p, C.c_int
        R.lapack_ref_dposv_upper.argtypes, R.lapack_ref_dposv_upper.restype = [i, vp, vp], i
        R.lapack_ref_dgemm_tn.argtypes = [i, i, i, vp, vp, vp]
        R.lapack_ref_dgemv_n.argtypes = [i, i, vp, vp, vp]
        units = dict(F.parse_units(ref + "/lsinterp_mod.f95"))
        blas = F.parse_units(ref + "/blas.f95")
        for k in ("solveblassymmetric", "mattransmatmultblas", "matvecmultblas"):
            units[k] = blas[k]
        txt = open(ref + "/constants.f95").read().lower()
        consts = {"pi": float(re.search(r"\bpi\s*=\s*([0-9.]+)_dp", txt).group(1)), "r_e": float(re.search(r"\br_e\s*=\s*([0-9.e+-]+)_dp", txt).group(1))}
        self.found = None

        def ptr(a):
            assert isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.f_contiguous
            return a.ctypes.data

        def dposv(v):
            uplo, n, nrhs, a, lda, b, ldb, info = v
            assert uplo == "u" and nrhs == 1 and lda == n and ldb == n
            return [None] * 7 + [int(R.lapack_ref_dposv_upper(n, ptr(a), ptr(b)))]

        def dgemm(v):
            ta, tb, m, n, k, alpha, a, lda, b, ldb, beta, c, ldc = v
            assert (ta, tb) == ("t", "n") and alpha == 1.0 and beta == 0.0 and lda == k and ldb == k and ldc == m
            R.lapack_ref_dgemm_tn(m, n, k, ptr(a), ptr(b), ptr(c))
            return [None] * 13

        def dgemv(v):
            t, m, n, alpha, a, lda, x, incx, beta, y, incy = v
            assert t == "n" and alpha == 1.0 and beta == 0.0 and lda == m and incx == 1 and incy == 1
            R.lapack_ref_dgemv_n(m, n, ptr(a), ptr(x), ptr(y))
            return [None] * 11

        self.I = F.Interp(units, consts, {"kdtree_search": lambda v: [None, None, None, self.found[0], self.found[1]],
                                          "dposv": dposv, "dgemm": dgemm, "dgemv": dgemv})

    def __call__(self, x, pts, vals, radius, order, exact, scaled, local_window_scale):
        self.found = (np.asfortranarray(pts, dtype=np.float64), np.asfortranarray(vals, dtype=np.float64))
        fi = np.full(vals.shape[1] - 1, np.nan)
        _, fin = self.I.call("lsinterp", [np.array(x, dtype=np.float64), 0, float(radius), int(order), int(exact), int(scaled),
                                          float(local_window_scale), fi, 0])
        return int(fin[8]), np.asarray(fin[7])


def test_lsinterp_against_the_references_own_source_text(sample_file, tmp_path):
    """The pin of the moving-least-squares arithmetic: the oracle's restatement (built on glibc's libm, as a gfortran binary
    here would be) == the reference's lsinterp_mod.f95 + blas.f95 executed statement by statement + its vendored LAPACK, bit
    for bit, statuses included; every (order, exact, scaled), neighbourhoods of 10 to > 100 samples, and the degenerate plane
    (singular normal equations -> status 1; too few neighbours -> status 2)."""
    ref = ReferenceLsinterp()
    path = sample_file[0]
    checked, statuses = 0, set()
    for order, exact, scaled in ((0, 0, 0), (1, 1, 0), (2, 0, 0), (2, 1, 0), (3, 0, 0), (3, 1, 1), (2, 0, 1)):
        kw = dict(window_scale=1.8, order=order, exact=exact, local_window_scale=2.5, scaled=scaled)
        he = common.HostEmuScattered(path, **kw)
        om = ol.OracleScattered(path, libm=True, **kw)
        for x in query_points(12, seed=100 + 10 * order + exact):
            a = he.records[he.neighbours(x)]
            st_r, fi_r = ref(x, a[:, :3], a[:, 3:8], he.radius, order, exact, scaled, 2.5)
            st_o, fi_o = om.lsinterp_found(x, a[:, :3], a[:, 3:8])
            assert st_r == st_o and np.array_equal(fi_r, fi_o[:4]), (order, exact, scaled, st_r, st_o, fi_r, fi_o)
            checked += 1
            statuses.add(st_r)
        om.close()
    rng = np.random.default_rng(4)
    n, base, span = 600, np.array([2.5, 0.3, 0.2]) * R_E, 0.4 * R_E
    pos = base + np.c_[rng.random(n) - 0.5, rng.random(n) - 0.5, np.zeros(n)] * span
    f = np.sin(pos[:, 0] / span) + 0.3 * np.cos(pos[:, 1] / span) + 20.0
    plane = str(tmp_path / "plane.txt")
    with open(plane, "w") as fh:
        fh.write("2 0 0 0 0 0 0\n-1.602e-19 1.602e-19\n9.10938188e-31 1.6726e-27\n")
        for p, v in zip(pos, f):
            fh.write("%.17e %.17e %.17e %.17e %.17e\n" % (p[0], p[1], p[2], v, v - 1.0))
    for order, exact, lws in ((1, 0, 3.0), (2, 1, 3.0), (3, 0, 0.05)):
        kw = dict(window_scale=3.0, order=order, exact=exact, local_window_scale=lws)
        he = common.HostEmuScattered(plane, **kw)
        om = ol.OracleScattered(plane, libm=True, **kw)
        for x in np.vstack([base + (rng.random((8, 3)) - 0.5) * span, pos[:2], base + [0, 0, 5 * span]]):
            a = he.records[he.neighbours(x)]
            vals = np.c_[a[:, 3:5], a[:, 7]].reshape(-1, 3)
            st_r, fi_r = ref(x, a[:, :3].reshape(-1, 3), vals, he.radius, order, exact, 0, lws)
            st_o, fi_o = om.lsinterp_found(x, a[:, :3].reshape(-1, 3), vals)
            assert st_r == st_o and np.array_equal(fi_r, fi_o[:2], equal_nan=True), (order, exact, lws, st_r, st_o, fi_r, fi_o)
            checked += 1
            statuses.add(st_r)
        om.close()
    print("lsinterp: %d queries equal to the executed reference text; statuses met: %s" % (checked, sorted(statuses)))
    assert statuses == {0, 1, 2} and checked == 117


def test_kdtree_and_setup_against_the_references_own_source_text(tmp_path):
    """The pin of the oracle's kd-tree: fortran/kdtree_mod.f95 (kdtree_add, kdtree_nearest, kdtree_find_ptr, kdtree_search and
    its core) EXECUTED from the reference's text by oracle/f90_interp.py (derived type, pointers, recursion, move_alloc), driven
    through the steps of the adapter's setup (scattered_interp_dens_model_adapter.f95:141-191: insert unless the nearest sample
    already in the tree is at distance 0; then every sample's nearest other sample, written into the tree through the pointer
    kdtree_find_ptr returns).  Same insertion order on both sides: the oracle's tree returns the same neighbours IN THE SAME
    ORDER with the same stored distances, keeps and drops the same samples, finds the same maxnearest -- and the root sample's
    distance is 0 in the reference's own text (the quirk the oracle restates and the product does not copy)."""
    import sys
    ref = "/root/reference/fortran/kdtree_mod.f95"
    if not os.path.exists(ref):
        pytest.skip("the reference's sources are not present on this machine")
    sys.path.insert(0, os.path.join(common.ROOT, "oracle"))
    sys.setrecursionlimit(max(sys.getrecursionlimit(), 20000))
    import f90_interp as F
    I = F.Interp(F.parse_units(ref), {}, {})
    path = str(tmp_path / "small.txt")
    pos, logn = common.write_scattered_file(path, 250, seed=12, duplicates=3, inside=2)
    nspec, n = 4, len(pos)
    perm = np.random.default_rng(6).permutation(n)
    tree, kept, dropped = None, 0, 0
    for i in perm:  # adapter :141-158
        p = pos[i].copy()
        _, fin = I.call("kdtree_nearest", [tree, p, 0, None, None])
        best = fin[3]
        if best is None or float(np.dot(best - p, best - p)) > 0.0:
            _, fin = I.call("kdtree_add", [tree, p, np.r_[logn[i], 1.0], 0])
            tree = fin[0]
            kept += 1
        else:
            dropped += 1
    maxnearest = 0.0
    for i in range(n):  # adapter :160-191
        p = pos[i].copy()
        if (p[0] * p[0] + p[1] * p[1]) + p[2] * p[2] >= R_E * R_E:
            _, fin = I.call("kdtree_nearest", [tree, p, 1, None, None])
            best = fin[3]
            _, fin = I.call("kdtree_find_ptr", [tree, p, None, None])
            valptr = fin[3]
            assert valptr is not None
            d = best - p
            dist = float(np.sqrt((d[0] * d[0] + d[1] * d[1]) + d[2] * d[2]))
            valptr[-1] = dist
            maxnearest = max(maxnearest, dist)
    om = ol.OracleScattered(path, perm=perm, root_quirk=True, **OPTS)
    assert (om.nkept, om.nduplicates) == (kept, dropped) == (250, 3)
    assert om.maxnearest == maxnearest
    assert tree.attrs["val"][-1] == 0.0 == om.root_nearest  # kdtree_nearest starts from best = root
    radius = om.radius
    total = 0
    for x in query_points(40, seed=5, rmin=1.1, rmax=4.8):
        _, fin = I.call("kdtree_search", [tree, np.array(x), float(radius), None, None])
        fp, fv = fin[3], fin[4]
        mine = om.search(x, nspec)
        if fp is None or len(fp) == 0:
            assert len(mine) == 0
            continue
        assert np.array_equal(np.c_[fp, fv], mine)  # same samples, same order, same values and stored distances
        total += len(fp)
    assert total > 300
    om.close()
    # ---- and the whole of funcPlasmaParams' density (adapter :214-262) from the reference's text: the adapter's statements,
    # lsinterp, kdtree_search on the tree built above, the BLAS wrappers, the vendored LAPACK.  Field routines are stubbed
    # (that block is the other adapters', pinned through the translated sources).
    ref = ReferenceLsinterp()
    adapter = F.parse_units("/root/reference/fortran/scattered_interp_dens_model_adapter.f95")
    extra = dict(F.parse_units("/root/reference/fortran/kdtree_mod.f95"))
    extra["funcplasmaparams"] = adapter["funcplasmaparams"]
    ref.I.units.update(extra)
    ref.I.trees.update({k: ref.I._structure(u.body) for k, u in extra.items()})
    state = F.Obj()
    state.attrs.update(dict(nspec=nspec, tree=tree, maxnearest=maxnearest, window_scale=OPTS["window_scale"], order=OPTS["order"], exact=OPTS["exact"],
                            scaled=0, local_window_scale=OPTS["local_window_scale"], qs=np.array([-1.0, 1.0, 1.0, 1.0]), ms=np.array([1.0, 2.0, 3.0, 4.0]),
                            itime=np.array([2010001, 0]), use_igrf=0, use_tsyganenko=0, pdyn=0.0, dst=0.0, byimf=0.0, bzimf=0.0,
                            w1=0.0, w2=0.0, w3=0.0, w4=0.0, w5=0.0, w6=0.0))
    datap = F.Obj()
    datap.attrs["p"] = state
    ref.I.hooks.update({"transfer": lambda v: datap, "sm_to_gsm_d": lambda v: [None, None, np.array(v[1])], "gsm_to_sm_d": lambda v: [None, None, np.array(v[1])],
                        "tsy_recalc": lambda v: [None] * 5, "bmodel_cartesian": lambda v: np.zeros(3)})
    om = ol.OracleScattered(path, perm=perm, root_quirk=True, libm=True, **OPTS)
    xs = np.vstack([query_points(30, seed=9, rmin=1.1, rmax=4.8), [[0.3 * R_E, 0.2 * R_E, 0.0]], [[30 * R_E, 0.0, 0.0]]])
    want = om.plasma(xs)[:, 4:8]
    interpolated = 0
    for x, w in zip(xs, want):
        ns = np.full(nspec, np.nan)
        _, fin = ref.I.call("funcplasmaparams", [np.array(x), np.zeros(nspec), ns, np.zeros(nspec), np.zeros(nspec), np.zeros(3), 0])
        assert np.array_equal(fin[2], w), (x, fin[2], w)
        assert np.array_equal(fin[1], state.attrs["qs"]) and np.array_equal(fin[3], state.attrs["ms"]) and not fin[4].any()
        interpolated += int(w[0] != 1.0 and w[0] != 0.0)
    assert interpolated >= 15 and (want[-2] == 0.0).all() and (want[-1] == 1.0).all()  # inside the Earth: 0; nothing in reach: exp(0)
    om.close()


def test_cell_list_finds_the_brute_force_set_for_any_geometry(tmp_path):
    """the walk over the cell list against brute force over the records, for sample counts from 1 to 800, search radii from a
    fraction of the sample spacing to many times the bounding box (one cell), anisotropic boxes, and queries inside, on the
    faces of, and far outside the box: the same samples, each once"""
    rng = np.random.default_rng(17)
    case = 0
    for n in (1, 2, 7, 60, 800):
        for ws in (0.3, 1.0, 6.0, 80.0):
            for shape in ((1.0, 1.0, 1.0), (1.0, 0.02, 0.3)):
                case += 1
                pos = np.array([2.0, 1.0, 0.5]) * R_E + (rng.random((n, 3)) - 0.5) * np.array(shape) * 0.5 * R_E
                path = str(tmp_path / ("c%d.txt" % case))
                with open(path, "w") as fh:
                    fh.write("1 0 0 0 0 0 0\n-1.602e-19\n9.1e-31\n")
                    for p in pos:
                        fh.write("%.17e %.17e %.17e 20.0\n" % tuple(p))
                he = common.HostEmuScattered(path, window_scale=ws, order=1, exact=0, local_window_scale=3.0)
                assert he.npts == n
                rec = he.records
                lo, hi = rec[:, :3].min(0), rec[:, :3].max(0)
                q = np.vstack([lo + (hi - lo) * rng.random((12, 3)), lo, hi, lo - he.radius * 0.999, hi + he.radius * 1.001,
                               lo + (hi - lo) * rng.random((4, 3)) * 40.0 - 20.0 * (hi - lo), rec[:3, :3]])
                for x in q:
                    got = he.neighbours(x)
                    d = rec[:, :3] - x
                    want = np.nonzero((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]) + d[:, 2] * d[:, 2] < he.radius * he.radius)[0]
                    assert len(got) == len(set(got.tolist())) and np.array_equal(np.sort(got), want), (n, ws, shape, x)
                    assert np.array_equal(got, np.sort(got))  # the walk's order is the records' order
    assert case == 40
