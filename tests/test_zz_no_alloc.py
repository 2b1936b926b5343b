"""srt_model_eval_noalloc (include/srt.h): the allocation-free variant of srt_model_eval, whose staging memory is a
static device array.  Called through the C ABI; its output is compared with the ordinary entry point's bit for bit
(same kernel, same inputs) on two calls in a row -- the second with other points and fewer of them, so stale contents
of the shared array would show.  Collected last (file name) so that a failure here hides no other test."""
import ctypes as C

import numpy as np
import pytest

import common

pytestmark = pytest.mark.gpu
R_E = 6371.2e3


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    s.init(0)
    return s


def points(n, seed):
    rng = np.random.default_rng(seed)
    r = R_E * rng.uniform(1.1, 6.0, n)
    th = rng.uniform(0.2, np.pi - 0.2, n)
    ph = rng.uniform(-np.pi, np.pi, n)
    return np.ascontiguousarray(np.stack([r * np.sin(th) * np.cos(ph), r * np.sin(th) * np.sin(ph), r * np.cos(th)], axis=1))


def eval_both(srt, model, x):
    L = srt.lib()
    a = np.full((len(x), 19), np.nan)
    b = np.full((len(x), 19), np.nan)
    assert L.srt_model_eval(model._h, len(x), x.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p)) == 0
    rc = L.srt_model_eval_noalloc(model._h, len(x), x.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p))
    assert rc == 0, L.srt_last_cuda_error().decode()
    return a, b


@pytest.mark.parametrize("use_igrf", [0, 1])
def test_two_calls_in_a_row_equal_the_allocating_entry_point(srt, use_igrf):
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, use_igrf, 0)
    nmax = int(srt.lib().srt_model_eval_noalloc_max_points())
    assert nmax == 16384
    for n, seed in ((nmax, 1), (1000, 2)):
        a, b = eval_both(srt, model, points(n, seed))
        assert np.isfinite(a).all()
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def test_gridded_model_already_on_the_device(srt):
    import os
    grid = srt.interpsetup(os.path.join(common.ROOT, "tests", "golden", "grid13.txt"), 2010001, 0, 1, 0)
    for n, seed in ((5000, 3), (17, 4)):
        a, b = eval_both(srt, grid, points(n, seed))
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64))


def test_more_points_than_the_scratch_holds_is_refused(srt):
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    L = srt.lib()
    n = int(L.srt_model_eval_noalloc_max_points()) + 1
    x = points(n, 5)
    out = np.zeros((n, 19))
    rc = L.srt_model_eval_noalloc(model._h, n, x.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
    assert rc == -7
    assert b"exceed" in L.srt_last_cuda_error() and b"16384" in L.srt_strerror(rc)
    assert not out.any()
