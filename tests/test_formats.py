"""Host-side file formats of the reference driver: byte-level checks."""
import numpy as np

from stanford_raytracer_b200 import rayfile


def test_es24_15e3():
    assert rayfile.es24(7.3712e6) == "  7.371200000000000E+006"
    assert rayfile.es24(-1.602e-19) == " -1.602000000000000E-019"
    assert rayfile.es24(0.0) == "  0.000000000000000E+000"
    assert rayfile.es24(1e-300) == "  1.000000000000000E-300"
    assert all(len(rayfile.es24(v)) == 24 for v in (1.0, -2.5e100, 3e-7, float("nan"), float("inf")))


def test_output_row_layout():
    row = np.arange(20, dtype=np.float64) + 0.5
    qs, ms, nus = [-1.602e-19] + [1.602e-19] * 3, [9.1e-31, 1.67e-27, 6.7e-27, 2.7e-26], [0.0] * 4
    line = rayfile.format_row(3, 1, row, 12566.370614359172, qs, ms, nus)
    # (2i10, 17es24.15e3, i10, 4*nspec es24.15e3) + newline = 822 + 1 characters for nspec = 4 (Appendix B.1)
    assert len(line) == 823 and line.endswith("\n")
    assert line[:20] == "         3         1"
    assert line[20:44] == rayfile.es24(0.5)
    assert line[20 + 16 * 24:20 + 17 * 24] == rayfile.es24(12566.370614359172)
    assert line[20 + 17 * 24:30 + 17 * 24] == "         4"
    body = line[30 + 17 * 24:-1]
    assert len(body) == 16 * 24
    assert body[8 * 24:9 * 24] == rayfile.es24(16.5)  # Ns(1) = row[16]


def test_read_inputrays(tmp_path):
    p = tmp_path / "rays.txt"
    p.write_text("1.0E007 0.0 0.0 0.0 0.0 1.0 6.28e4\n 7.3712d6, 0, 0, 0 0 0 12566.37 trailing-ignored\n\n5 5 5 1 0 0 100\nbad line stops reading\n9 9 9 0 0 1 7\n")
    pos0, dir0, w = rayfile.read_inputrays(str(p))
    assert pos0.shape == (3, 3)
    assert pos0[0, 0] == 1.0e7 and dir0[0, 2] == 1.0 and w[0] == 6.28e4
    assert pos0[1, 0] == 7.3712e6 and w[1] == 12566.37


# ---- the library's formatter / reader (csrc/srt_io.cpp) against the Python statement above

def _awkward_doubles():
    rng = np.random.default_rng(12)
    bits = rng.integers(0, 2 ** 63, 60000, dtype=np.int64).view(np.float64)          # whole exponent range, incl. subnormals / NaNs
    typical = np.exp(rng.uniform(-60, 60, 60000)) * rng.choice([-1.0, 1.0], 60000)
    ties = np.array([0.5, 1.5, 2.5, 9.5, 1.0000000000000005, 9.9999999999999995e99, 9.999999999999999e-100, 1e23, 8.5e-5, 5e-324,
                     2.2250738585072014e-308, 1.7976931348623157e308, 0.0, -0.0, 1.0, -1.0, np.nan, np.inf, -np.inf, 123456789012345675.0])
    return np.concatenate([bits, -bits[:2000], typical, ties])


def test_native_es24_equals_python(built):
    import stanford_raytracer_b200 as srt
    v = _awkward_doubles()
    got = srt.format_es24(v)
    for x, g in zip(v, got):
        assert g == rayfile.es24(x), (x, g)


def test_native_rows_equal_python(built):
    import stanford_raytracer_b200 as srt
    rng = np.random.default_rng(13)
    rows = np.exp(rng.uniform(-40, 40, (37, 20))) * rng.choice([-1.0, 1.0], (37, 20))
    rows[5, 3] = np.nan
    rows[6, 16] = 0.0
    for nspec in (4, 2, 1):
        qs, ms, nus = [-1.602e-19, 1.602e-19, 1.602e-19, 1.602e-19][:nspec], [9.1e-31, 1.67e-27, 6.7e-27, 2.7e-26][:nspec], [0.0] * nspec
        want = "".join(rayfile.format_row(123456, 6, r, 31415.926, qs, ms, nus) for r in rows).encode()
        got = srt.format_rows(123456, 6, rows, 31415.926, qs, ms, nus)
        assert got == want
        assert len(got) == 37 * srt.lib().srt_row_text_bytes(nspec)
    # I10 overflow prints asterisks, as Fortran does
    assert srt.format_rows(12345678901, 0, rows[:1], 1.0, [1.0], [1.0], [0.0])[:10] == b"*" * 10
    assert srt.format_rows(1, 0, rows[:0], 1.0, [1.0], [1.0], [0.0]) == b""


def test_native_reader_equals_python(built, tmp_path):
    import stanford_raytracer_b200 as srt
    p = tmp_path / "rays.txt"
    p.write_text("1.0E007 0.0 0.0 0.0 0.0 1.0 6.28e4\n 7.3712d6, 0, 0, 0 0 0 12566.37 trailing ignored\n\n5 5 5\n 1 0 0 100 200 300\n"
                 "\t1.5+3,2,3,4,5,6,7\r\nbad line stops reading\n9 9 9 0 0 1 7\n")
    a = rayfile.read_inputrays(str(p))
    b = srt.read_inputrays(str(p))
    assert len(b[2]) == 4 and b[0][3, 0] == 1500.0 and b[2][2] == 100.0
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    # a file written at full precision comes back bit for bit
    rng = np.random.default_rng(14)
    rays = rng.normal(size=(500, 7)) * 1e7
    q = tmp_path / "many.txt"
    q.write_text("".join(" ".join(repr(float(v)) for v in r) + "\n" for r in rays))
    pos0, dir0, w0 = srt.read_inputrays(str(q))
    assert np.array_equal(np.concatenate([pos0, dir0, w0[:, None]], axis=1), rays)
    e = tmp_path / "empty.txt"
    e.write_text("")
    assert len(srt.read_inputrays(str(e))[2]) == 0
    import pytest
    with pytest.raises(srt.SrtError):
        srt.read_inputrays(str(tmp_path / "missing.txt"))


def test_cli_usage_and_errors(built, tmp_path):
    """`raytracer`: usage without arguments (driver:75-177), loud failures for missing options, unsupported models and
    -- in this container -- the missing GPU (no CPU path)."""
    import os
    import subprocess
    import common
    exe = os.path.join(common.ROOT, "stanford_raytracer_b200", "raytracer")
    assert os.path.exists(exe)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0 and "--inputraysfile" in r.stdout and "--modelnum" in r.stdout
    r = subprocess.run([exe, "--dt0=1e-3"], capture_output=True, text=True)
    assert r.returncode == 2 and "--dtmax is required" in r.stderr
    rays = tmp_path / "rays.txt"
    rays.write_text("7.3712e6 0 0 0 0 0 12566.37\n")
    common_args = ["--dt0=5e-4", "--dtmax=0.1", "--tmax=0.01", "--root=2", "--fixedstep=1", "--maxerr=5e-4", "--maxsteps=100", "--minalt=6.4712d6",
                   "--inputraysfile=%s" % rays, "--outputfile=%s" % (tmp_path / "out.ray"), "--yearday=2010001", "--milliseconds_day=0"]
    r = subprocess.run([exe] + common_args + ["--modelnum=2"], capture_output=True, text=True)
    import torch
    if not torch.cuda.is_available():
        assert r.returncode == 1 and "CUDA" in r.stderr
    else:
        assert r.returncode == 3 and "not available on the GPU path" in r.stderr


def test_read_output_round_trip(tmp_path):
    """The consumer side of the output format (damping/src/wipp_fileutils.cpp:6-170): what the writer puts is what
    the reader returns, ray by ray, to the 16 digits of the text."""
    from stanford_raytracer_b200 import rayfile
    rng = np.random.default_rng(11)
    qs = np.array([-1.602e-19, 1.602e-19, 1.602e-19, 1.602e-19])
    ms = np.array([9.10938188e-31, 1.6726e-27, 6.6904e-27, 2.67616e-26])
    nus = np.zeros(4)
    p = tmp_path / "out.ray"
    want = {}
    with open(p, "w") as f:
        for raynum, (nrows, sc, w) in enumerate([(3, 1, 2 * np.pi * 2000.0), (1, 2, 2 * np.pi * 200.0), (5, 0, 1.0e5)], start=1):
            rows = rng.standard_normal((nrows, 20)) * 10.0 ** rng.integers(-12, 12, (nrows, 20))
            rows[:, 1] = -np.abs(rows[:, 1])  # negative values fill all 24 columns of their field
            want[raynum] = (rows, sc, w)
            for row in rows:
                f.write(rayfile.format_row(raynum, sc, row, w, qs, ms, nus))
    got = rayfile.read_output(str(p))
    assert list(got) == [1, 2, 3]
    for raynum, (rows, sc, w) in want.items():
        g = got[raynum]
        assert g["stopcond"] == sc and g["nspec"] == 4
        assert np.isclose(g["w"], w, rtol=1e-15, atol=0)
        assert np.array_equal(g["qs"], qs) and np.allclose(g["ms"], ms, rtol=1e-15, atol=0)
        back = np.concatenate([g["time"][:, None], g["pos"], g["vprel"], g["vgrel"], g["n"], g["B0"], g["Ns"]], axis=1)
        assert back.shape == rows.shape
        assert np.allclose(back, rows, rtol=1e-15, atol=0)
        assert np.array_equal(g["nus"], np.zeros((len(rows), 4)))
        # re-writing what was read gives the same bytes
    text = "".join(rayfile.format_row(k, got[k]["stopcond"],
                                      np.concatenate([[got[k]["time"][i]], got[k]["pos"][i], got[k]["vprel"][i], got[k]["vgrel"][i], got[k]["n"][i], got[k]["B0"][i], got[k]["Ns"][i]]),
                                      got[k]["w"], got[k]["qs"], got[k]["ms"], got[k]["nus"][i])
                   for k in got for i in range(len(got[k]["time"])))
    assert text == p.read_text()


def test_output_rows_through_the_references_own_reader(built, tmp_path):
    """f2 / a22: the product's native writer of output.ray, read by the REFERENCE'S OWN consumer of that format --
    read_rayfile of damping/src/wipp_fileutils.cpp, compiled where it lies into oracle/_ref/libwipp_read.so -- gives back every
    field in its place (ray number, stop code, time, pos, vprel, vgrel, n, B0, w, nspec, qs, ms, Ns, nus), to the 16 digits the
    record carries."""
    import ctypes as C
    import os

    import pytest

    import common
    import stanford_raytracer_b200 as srt
    lib = os.path.join(common.ROOT, "oracle", "_ref", "libwipp_read.so")
    if not os.path.exists(lib):
        pytest.skip("oracle/_ref/libwipp_read.so not built (reference sources absent)")
    W = C.CDLL(lib)
    W.wipp_read_ray.restype = C.c_long
    W.wipp_read_ray.argtypes = [C.c_char_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
    rng = np.random.default_rng(15)
    for nspec in (4, 2):
        qs = [-1.602e-19, 1.602e-19, 1.602e-19, 1.602e-19][:nspec]
        ms = [9.10938188e-31, 1.6726e-27, 6.6904e-27, 2.67616e-26][:nspec]
        nus = [0.0, 1.5, 2.5e3, 1e-7][:nspec]
        rays = {3: (1, 41), 17: (6, 7), 1200: (0, 1)}
        path = str(tmp_path / ("out%d.ray" % nspec))
        want = {}
        with open(path, "wb") as f:
            for raynum, (stop, nrows) in rays.items():
                rows = np.exp(rng.uniform(-30, 30, (nrows, 20))) * rng.choice([-1.0, 1.0], (nrows, 20))
                w = float(rng.uniform(1e3, 1e5))
                f.write(srt.format_rows(raynum, stop, rows, w, qs, ms, nus))
                want[raynum] = (stop, rows, w)
        digits16 = lambda a: np.array([float("%.15E" % v) for v in np.ravel(a)]).reshape(np.shape(a))  # noqa: E731
        for raynum, (stop, rows, w) in want.items():
            head, qm, got = np.zeros(5), np.zeros(2 * nspec), np.zeros((len(rows) + 3, 16 + 2 * nspec))
            n = W.wipp_read_ray(path.encode(), raynum, head.ctypes.data, qm.ctypes.data, got.ctypes.data, len(got))
            assert n == len(rows) and head[0] == len(rays) and head[1] == len(rows) and head[2] == nspec and head[3] == stop
            assert head[4] == digits16([w])[0]
            assert np.array_equal(qm, digits16(qs + ms))
            assert np.array_equal(got[:n, :16], digits16(rows[:, :16]))
            assert np.array_equal(got[:n, 16:16 + nspec], digits16(rows[:, 16:16 + nspec]))
            assert np.array_equal(got[:n, 16 + nspec:], np.tile(digits16(nus), (n, 1)))
