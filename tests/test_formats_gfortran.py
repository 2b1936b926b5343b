"""The driver's file formats against gfortran's OWN run-time library.

The reference is built with gfortran: its output.ray is what libgfortran's formatted WRITE produces for i10 / es24.15e3, its ray file
is what libgfortran's list-directed READ accepts (fortran/raytracer_driver.f95:1140-1217).  libgfortran.so.5 travels in the image's
numpy / scipy wheels; oracle/ref_stubs/gfortran_io_shim.c issues the calls gfortran emits for those statements.  The product's
native formatter, writer and reader (csrc, through the C ABI) must agree with it byte for byte / value for value."""
import ctypes as C
import glob
import os
import sys

import numpy as np
import pytest

import common
import oracle_lib as ol
from stanford_raytracer_b200 import rayfile


@pytest.fixture(scope="module")
def gf():
    libs = sorted(p for d in sys.path if d.endswith("site-packages") for p in glob.glob(os.path.join(d, "scipy.libs", "libgfortran*.so*")))
    if not libs:
        pytest.skip("no libgfortran in the image's wheels")
    ol.build()
    G = C.CDLL(os.path.join(ol.ORC_DIR, "_build", "libgfio.so"))
    G.gfio_open_lib.argtypes = [C.c_char_p]
    if G.gfio_open_lib(libs[0].encode()) != 0:
        pytest.skip("libgfortran does not export the GFORTRAN_8 I/O entry points")
    G.gfio_write_es.argtypes = [C.c_double, C.c_char_p]
    G.gfio_write_rayfile.argtypes = [C.c_char_p, C.c_long, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    G.gfio_read_rays.argtypes = [C.c_char_p, C.c_void_p, C.c_long]
    G.gfio_read_rays.restype = C.c_long
    return G


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    return s


def awkward():
    rng = np.random.default_rng(41)
    v = [0.0, -0.0, 1.0, -1.0, 9.9999999999999995e-8, 9.999999999999999e22, 1e23, 5e-324, 2.2250738585072014e-308, 1.7976931348623157e308,
         0.1, 0.5, 1.5, 2.5, 1.0000000000000002, 9.999999999999995, 9.9999999999999995, 1.2345678901234565, 1.2345678901234575,
         float("inf"), float("-inf"), float("nan"), 7.3712e6, -1.602e-19, 8.854187817e-12, 299792458.0, 123456789012345.5, 1e100, 1e-100, 1e-99, 9.5e99]
    v += list(np.ldexp(rng.uniform(0.5, 1.0, 4000), rng.integers(-1070, 1023, 4000)) * rng.choice([-1.0, 1.0], 4000))
    v += list(np.round(rng.uniform(1, 10, 2000), 15) * 10.0 ** rng.integers(-300, 300, 2000))  # decimals with at most 16 digits: ties and near-ties
    v += [float("%d.%015d5e%d" % (rng.integers(1, 10), rng.integers(0, 10 ** 15), rng.integers(-200, 200))) for _ in range(2000)]  # a 17th digit 5
    return np.array(v, dtype=np.float64)


def test_es24_15e3_digit_generation(gf, srt):
    v = awkward()
    mine = srt.format_es24(v)
    buf = C.create_string_buffer(25)
    for x, m in zip(v, mine):
        gf.gfio_write_es(float(x), buf)
        want = buf.raw[:24].decode()
        assert m == want, (repr(float(x)), m, want)
        assert rayfile.es24(float(x)) == want, (repr(float(x)), rayfile.es24(float(x)), want)


def test_output_file_byte_for_byte(gf, srt, tmp_path):
    """the driver's output loop run by libgfortran == the product's writer, whole files (2, 3 and 4 species; ray numbers beyond i10)"""
    rng = np.random.default_rng(42)
    for nspec in (4, 3, 2):
        qs = [-1.602e-19, 1.602e-19, 1.602e-19, 1.602e-19][:nspec]
        ms = [9.10938188e-31, 1.6726e-27, 6.6904e-27, 2.67616e-26][:nspec]
        nus = [0.0, 1.5, 2.5e3, 1e-7][:nspec]
        chunks, head, v17, sp = [], [], [], []
        for raynum, stop, nrows in ((1, 0, 30), (2, 6, 1), (99999, 3, 12), (2147483647, 1, 2), (-5, 2, 1)):
            rows = np.exp(rng.uniform(-60, 60, (nrows, 20))) * rng.choice([-1.0, 1.0], (nrows, 20))
            rows[0, 16:] = [0.0, np.nan, np.inf, -np.inf]
            w = float(rng.uniform(1e3, 1e5))
            chunks.append(srt.format_rows(raynum, stop, rows, w, qs, ms, nus))
            for r in rows:
                head.append([raynum, stop])
                v17.append(list(r[:16]) + [w])
                sp.append(qs + ms + list(r[16:16 + nspec]) + nus)
        head, v17, sp = np.array(head, dtype=np.int32), np.array(v17), np.array(sp)
        path = str(tmp_path / ("gf%d.ray" % nspec))
        gf.gfio_write_rayfile(path.encode(), len(head), head.ctypes.data, v17.ctypes.data, nspec, sp.ctypes.data)
        want = open(path, "rb").read()
        got = b"".join(chunks)
        assert len(want) == len(head) * (20 + 17 * 24 + 10 + 4 * nspec * 24 + 1)
        assert got == want


RAY_FILES = {
    "plain": "1.0E007 0.0 0.0 0.0 0.0 1.0 6.28e4\n7.3712d6 0 0 0 0 0 12566.37\n",
    "commas_tabs_trailing": " 7.3712d6, 0, 0, 0 0 0 12566.37 trailing ignored\n\t1.5+3,2,3,4,5,6,7\r\n1\t2\t3\t4\t5\t6\t7\n",
    "records_continue": "5 5 5\n 1 0 0 100 200 300\n8 8 8\n\n\n 0 0 1\n 42\n",
    "blank_lines_between": "\n\n1 2 3 4 5 6 7\n\n   \n8 9 10 11 12 13 14\n\n",
    "bad_token_stops": "1 2 3 4 5 6 7\nbad line stops reading\n9 9 9 0 0 1 7\n",
    "short_last_record": "1 2 3 4 5 6 7\n8 9 10\n",
    "no_final_newline": "1 2 3 4 5 6 7\n7 6 5 4 3 2 1",
    "more_than_seven": " 1 2 3 4 5 6 7 8 9 10 11 12 13 14\n1,2,3,4,5,6,7,8\n9 9 9 9 9 9 9\n",
    "repeat_counts": "3*1.5 3*0 7\n1 2 3 4 5 6 7\n9*2\n1 2*3 4 5 6 7 8\n",
    "repeat_nulls": "1 1 1 1 1 1 1\n2* 3 4 5 6 7\n2*,8 4 5 6 7 8\n2*1 2* 2*3 4\n",
    "null_values": "1 2 3 4 5 6 7\n8,,10 11 12 13 14\n,2 3 4 5 6 7\n1 2 3 ,, , 6 7 8\n1 2 3,\n,4 5 6 7 8\n1;;3 4 5 6 7\n;2 3 4 5 6 7\n",
    "semicolons": "1;2;3;4;5;6;7\n1 2 3 4 5 6 7;8\n",
    "slash_ends_a_read": "1 2 3 4 5 6 7\n1 2 3 / 9 9 9 9\n5 5 5 5 5 5 5\n/\n1 2 3 4 5 6 7 /\n1 2 3 4 5 6 7/\n1 2 3 4 5 6/7\n8 8 8 8 8 8 8\n",
    "number_forms": "+.5 -.5e-1 1d+3 1.D+3 1.E3 1.e-3 .5\n+1.e3 -.5 1d-3 1.5e+3 2.E-2 -0 +7\n1q2 1.5-3 1+5 5. +5.e1 1e+0003 1E5\n",
    "inf_and_nan": "1 2 3 4 5 6 inf\n-Infinity nan 1 1 1 1 +NaN\n+inf INF NaN nan(abc) -nan 1 1\n",
    "rounding": "12345678901234567890123 0.1000000000000000055511151231257827 1e400 1e-400 4.9e-324 2.4703282292062328e-324 9007199254740993\n",
    "value_then_comment": "1 2 3 4 5 6 7 ! c\n2 2 2 2 2 2 2\n",
    "value_then_split_exponent": "1 2 3 4 5 6 1 e5\n1 2 3 4 5 6 1,5\n",
    "empty": "",
}
# each of these ends the loop at once (iostat > 0): nothing is read
for _k, _t in enumerate(["1e", ".", "+", "1e5.3", "0x10", "1_8", "'a'", "T", "7e+", "1.5E 3", "1.5d0d", "--1", "infinit", "nan(abc", ".e3", "*5", "2**3", "1.0*3",
                         "1-", "1e5e", "(7)", '"7"', "7.5.", "1.5e+3-2", "1.5+3e2", "1e3d2", "infinity5", "in", "na", "0*5", "\x0c4"]):
    RAY_FILES["bad_%02d" % _k] = "1 2 3 4 5 6 %s\n1 1 1 1 1 1 1\n" % _t
RAY_FILES["bad_comment_line"] = "! comment\n1 2 3 4 5 6 7\n"


@pytest.mark.parametrize("name", sorted(RAY_FILES))
def test_ray_file_reader_against_list_directed_read(gf, srt, tmp_path, name):
    """the driver's loop `read(infile,*,iostat=status) pos0, dir0, w` run by libgfortran == the product's reader, ray for ray"""
    path = tmp_path / (name + ".txt")
    path.write_bytes(RAY_FILES[name].encode())
    want = np.zeros((64, 7))
    n = gf.gfio_read_rays(str(path).encode(), want.ctypes.data, 64)
    want = want[:n]
    for reader in (srt.read_inputrays, rayfile.read_inputrays):
        pos0, dir0, w0 = reader(str(path))
        got = np.concatenate([pos0, dir0, w0[:, None]], axis=1) if len(w0) else np.zeros((0, 7))
        assert got.shape == want.shape, (name, reader.__module__, got, want)
        assert np.array_equal(np.isnan(got), np.isnan(want)), (name, reader.__module__, got, want)
        a, b = np.where(np.isnan(got), 0.0, got), np.where(np.isnan(want), 0.0, want)
        assert np.array_equal(a.view(np.uint64), b.view(np.uint64)), (name, reader.__module__, got, want)  # incl. the sign of zero
    if name == "bad_comment_line" or name[4:6].isdigit():
        assert n == 0, name


def test_q13_the_runtime_matmul_is_machine_dependent(gf):
    """Q13 (libtricubic.f95:719, `a = matmul(Amat, x)`): gfortran calls libgfortran's matmul_r8, which is built with -ffast-math and
    picks an AVX / AVX2+FMA / AVX-512 variant at run time: its sums are reassociated and fused per machine.  The oracle and the
    kernels fix the order (ascending column, unfused).  Recorded here: the library's result on THIS host agrees with that order to
    a few units of the last place of the sum of magnitudes and, where the CPU has FMA, not bit for bit -- so no fixed order can equal
    the reference on every machine, and the difference is rounding noise of the coefficient product."""
    gf.gfio_matmul_r8.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int]
    rng = np.random.default_rng(51)
    differ, total, worst = 0, 0, 0.0
    for _ in range(100):
        A = np.asfortranarray(np.where(rng.random((64, 64)) < 0.2, rng.integers(-27, 28, (64, 64)), 0).astype(np.float64))  # as sparse and small-integer as Amat
        x = rng.uniform(-20, 20, 64)
        c = np.zeros(64)
        gf.gfio_matmul_r8(A.ctypes.data, x.ctypes.data, c.ctypes.data, 64, 64)
        ref = np.zeros(64)
        for col in range(64):  # ascending column order, one rounding per product and per sum
            ref = ref + A[:, col] * x[col]
        differ += int((c.view(np.uint64) != ref.view(np.uint64)).sum())
        total += 64
        worst = max(worst, float((np.abs(c - ref) / np.maximum(np.abs(A) @ np.abs(x), 1e-300)).max()))
    print("libgfortran matmul_r8 on this host: %d of %d results differ in bits from the fixed order; worst |difference| / sum|terms| = %.1e" % (differ, total, worst))
    assert worst < 2e-15


def test_q24_integer_powers_against_libgcc_and_libgfortran(gf):
    """Q24 (libtricubic.f95:677-679, `x**(i-derx)`): gfortran turns REAL(8)**INTEGER(4) into __builtin_powi = libgcc's __powidf2; other
    kinds go to libgfortran's pow_r8_i8 and the like.  The oracle's restatement equals both, bit for bit, for every exponent the path can meet."""
    import ctypes.util
    L = ol.lib()
    L.orc_pow_r8_i4.restype = C.c_double
    L.orc_pow_r8_i4.argtypes = [C.c_double, C.c_int]
    gcc = C.CDLL(ctypes.util.find_library("gcc_s") or "libgcc_s.so.1")
    gcc.__powidf2.restype = C.c_double
    gcc.__powidf2.argtypes = [C.c_double, C.c_int]
    libs = sorted(p for d in sys.path if d.endswith("site-packages") for p in glob.glob(os.path.join(d, "scipy.libs", "libgfortran*.so*")))
    gfl = C.CDLL(libs[0])
    gfl._gfortran_pow_r8_i8.restype = C.c_double  # (REAL(8)**INTEGER(4) has no library routine: it is the builtin; this is the INTEGER(8) one)
    gfl._gfortran_pow_r8_i8.argtypes = [C.c_double, C.c_int64]
    rng = np.random.default_rng(52)
    xs = list(rng.uniform(-2, 2, 300)) + list(np.exp(rng.uniform(-200, 200, 100))) + [0.0, -0.0, 1.0, 5e-324, 1e308]
    for x in xs:
        for n in range(0, 12):
            a, b = L.orc_pow_r8_i4(x, n), gcc.__powidf2(x, n)
            assert np.float64(a).tobytes() == np.float64(b).tobytes(), (x, n, a, b)
            assert np.float64(a).tobytes() == np.float64(gfl._gfortran_pow_r8_i8(x, n)).tobytes(), (x, n)


def test_dump_file_byte_for_byte(gf, built, tmp_path):
    """the end of dumpmodel.f95 (:1284-1289) run by libgfortran == dumpmodel.write_dump, whole file"""
    from stanford_raytracer_b200 import dumpmodel as dm
    gf.gfio_write_dump.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_long]
    rng = np.random.default_rng(61)
    for nspec in (4, 3):
        m = 4 * nspec + 3
        f = rng.standard_normal((3, 4, 5, m)) * 10.0 ** rng.integers(-300, 300, (3, 4, 5, m))
        f[0, 0, 0, :4] = [-1.602e-19, 0.0, -0.0, np.inf]
        ext = np.array([-6371200.0, 6371200.0, 0.0, 0.0, -1.27424e7, 1.27424e7])
        mine, theirs = str(tmp_path / ("mine%d.txt" % nspec)), str(tmp_path / ("gf%d.txt" % nspec))
        dm.write_dump(mine, f, tuple(ext))
        head = np.array([nspec, 5, 4, 3], dtype=np.int32)
        flat = np.ascontiguousarray(f).ravel()
        gf.gfio_write_dump(theirs.encode(), head.ctypes.data, ext.ctypes.data, flat.ctypes.data, len(flat))
        assert open(mine, "rb").read() == open(theirs, "rb").read()


def _grid_text(style, compder, rng):
    """a small grid file in several writing styles the list-directed READs of `setup` accept"""
    ns, nx, ny, nz = 3, 3, 2, 4
    nodes = nx * ny * nz
    vals = rng.uniform(5, 25, (8 if compder else 1, nodes, ns))
    fmt = {"one_per_line": lambda v: "".join("%24.15E\n" % x for x in v),
           "node_per_line": lambda v: "".join(" ".join("%.17g" % x for x in row) + "\n" for row in v.reshape(-1, ns)),
           "commas_and_d": lambda v: "".join(",".join(("%.16e" % x).replace("e", "D") for x in row) + "\n" for row in v.reshape(-1, ns)),
           "trailing_junk": lambda v: "".join(" ".join("%.17g" % x for x in row) + " 999 ignored\n" for row in v.reshape(-1, ns))}[style]
    junk = "  77 88 trailing\n" if style == "trailing_junk" else "\n"
    text = "%d %d %d %d %d%s" % (compder, ns, nx, ny, nz, junk)
    text += "-2.5e7 2.5e7 -1e7,1e7 -3e7 3.0D+7%s" % junk
    text += "-1.602e-19 1.602e-19 1.602e-19%s" % junk
    text += "9.1e-31\n1.67e-27\n6.69e-27%s" % junk  # a READ continuing over records
    text += fmt(vals[0].ravel())
    for q in range(1, len(vals)):  # whole-array READs: the values may sit on lines of any length; the next array starts on a new record
        flat = vals[q].ravel()
        text += "".join(" ".join("%.17g" % x for x in flat[k:k + 5]) + "\n" for k in range(0, len(flat), 5)) if style != "one_per_line" else fmt(flat)
    return text, (ns, nx, ny, nz)


@pytest.mark.parametrize("compder", [0, 1])
@pytest.mark.parametrize("style", ["one_per_line", "node_per_line", "commas_and_d", "trailing_junk"])
def test_grid_file_reader_against_the_runtime(gf, built, tmp_path, style, compder):
    """the READ statements of the interp adapter's setup run by libgfortran == the product's loader (csrc/srt_grid_setup.h, here its host
    build) == the oracle's, every array"""
    text, (ns, nx, ny, nz) = _grid_text(style, compder, np.random.default_rng(62))
    path = str(tmp_path / "grid.txt")
    open(path, "w").write(text)
    nodes = nx * ny * nz
    gf.gfio_read_grid.argtypes = [C.c_char_p] + [C.c_void_p] * 6 + [C.c_long]
    head, lims, qs, ms = np.zeros(5, dtype=np.int32), np.zeros(6), np.zeros(16), np.zeros(16)
    f, deriv = np.zeros(nodes * ns), np.zeros(7 * nodes * ns)
    assert gf.gfio_read_grid(path.encode(), head.ctypes.data, lims.ctypes.data, qs.ctypes.data, ms.ctypes.data, f.ctypes.data, deriv.ctypes.data, len(f)) == 0
    assert list(head) == [compder, ns, nx, ny, nz]
    e = common.HostEmu("grid", path, 2010001, 0, 0)
    E = e.E
    E.emu_grid_nodes.restype = C.c_long
    E.emu_grid_nodes.argtypes = [C.c_void_p] * 4 + [C.c_long]
    dims, qm, out = np.zeros(4, dtype=np.int32), np.zeros(8), np.zeros(8 * nodes * ns)
    assert E.emu_grid_nodes(e.h, dims.ctypes.data, qm.ctypes.data, out.ctypes.data, len(out)) == nodes
    assert list(dims) == [ns, nx, ny, nz]
    assert np.array_equal(qm[:ns], qs[:ns]) and np.array_equal(qm[ns:2 * ns], ms[:ns])
    assert np.array_equal(out[:nodes * ns].view(np.uint64), f.view(np.uint64))
    if compder:
        assert np.array_equal(out[nodes * ns:].view(np.uint64), deriv.view(np.uint64))
    o = ol.OracleModel("grid", path, 2010001, 0, 0, libm=True)
    import test_raytracer_ref_cpu as T
    _, _, A = T.state_of(o.L.orc_grid_state, o.h)
    assert np.array_equal(A[5].view(np.uint64), f.view(np.uint64))
    assert np.array_equal(A[2], np.arange(nx) * ((lims[1] - lims[0]) / (nx - 1.0)) + lims[0])
    if compder:
        assert np.array_equal(np.concatenate(A[6:]).view(np.uint64), deriv.view(np.uint64))


CARD_FILES = {
    "newray_default": None,   # the files of the test suite themselves
    "newray_ducts": None,
    "cards_in_odd_styles": "0,1 0 200.\n -1.0 , 50\n3 0;1 0 2*0 1.d3 4.0 0. 3e3\n0.0 1600.\n4* /\n7370. 3000 .08 .02, ,\n"
                           "6470 140 6370 1 30 / the rest is not read\n 4.0 .13 .07 1.e9 50\n1 2 3 4 5 6 7 8 9 10 11 12 13\n0 0 0 0 0 0 0 0\n",
    "truncated": "0 1 0 200.\n-1 50\n3 0 1 0 0\n",
    "bad_number": "0 1 0 200.\n-1 5O\n3 0 1 0 0 0 1 2 3 4\n",
}
CARD_COUNTS = [4, 2, 10, 5, 5, 5, 5, 12, 8]


@pytest.mark.parametrize("name", sorted(CARD_FILES))
def test_config_cards_against_the_runtime(gf, built, tmp_path, name):
    """the READ statements of readinput (ngo_dens_model.f95:61-115: 4, 2, 10, 5, 5, 5, 5, 12, 8 items) run by libgfortran == the
    product's list-directed reader (csrc/srt_listread.h), item for item, with the same READ failing the same way"""
    if name == "newray_default":
        path = common.NGO_CFG
    elif name == "newray_ducts":
        path = os.path.join(common.ROOT, "tests", "data", "newray_ducts_sin_gauss.in")
    else:
        path = str(tmp_path / (name + ".in"))
        open(path, "w").write(CARD_FILES[name])
    counts = np.array(CARD_COUNTS, dtype=np.int32)
    common.build_hostemu()
    E = C.CDLL(os.path.join(common.BUILD, "libsrt_hostemu.so"))
    res = []
    for fn in (gf.gfio_read_counts, E.emu_read_counts):
        fn.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        out, status = np.full(int(counts.sum()), -777.0), np.zeros(len(counts), dtype=np.int32)
        done = fn(path.encode(), counts.ctypes.data, len(counts), out.ctypes.data, status.ctypes.data)
        res.append((done, out, status))
    (da, oa, sa), (db, ob, sb) = res
    assert da == db, (name, da, db, sa, sb)
    assert np.array_equal(np.sign(sa), np.where(sb == -2, 1, sb)), (name, sa, sb)  # 0 / end of file (< 0 in Fortran, -1 here) / bad value (> 0, -2 here)
    k = int(counts[:da].sum())
    assert np.array_equal(oa[:k].view(np.uint64), ob[:k].view(np.uint64)), (name, oa[:k], ob[:k])
    if name == "truncated":
        assert da == 2 and sa[2] < 0
    if name == "bad_number":
        assert da == 1 and sa[1] > 0
    if name.startswith("newray"):
        assert da >= 7


def test_scattered_sample_file_against_the_runtime(gf, built, tmp_path):
    """the READ statements of the scattered adapter's setup (scattered_interp_dens_model_adapter.f95:88-110: 7 items, nspec,
    nspec, then `read(infile,*,iostat=reason) pos, dat%Ns` until a READ fails) run by libgfortran == what the product's set-up
    (csrc/srt_scattered_setup.h) takes from the same file: the same number of samples, the same values, the list ending at the
    same record (a word that is not a number; the record after it is never read)."""
    path = str(tmp_path / "odd.txt")
    with open(path, "w") as fh:
        fh.write("2, 0 0 0 0 0 0\n-1.602D-19 1.602d-19\n9.1e-31,\n1.67e-27\n")
        for i in range(30):
            fh.write("%.6e,%.6e\n%.6e %.3f %.3f trailing words are dropped with the record\n" % (2e7 + 1e5 * i, 1e5 * (i % 5), 2e5 * (i % 3), 20.0 + 0.01 * i, 19.0))
        fh.write("oops 1 2 3 4\n3e7 0 0 1 1\n")
    counts = np.array([7, 2, 2] + [5] * 40, dtype=np.int32)
    gf.gfio_read_counts.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    out, status = np.full(int(counts.sum()), -777.0), np.zeros(len(counts), dtype=np.int32)
    done = gf.gfio_read_counts(path.encode(), counts.ctypes.data, len(counts), out.ctypes.data, status.ctypes.data)
    assert done == 33 and status[33] > 0  # header, qs, ms, 30 samples; the 31st READ meets `oops`
    samples = out[11:11 + 5 * 30].reshape(30, 5)
    he = common.HostEmuScattered(path, window_scale=30.0, order=1, exact=0, local_window_scale=3.0)
    assert he.npts == 30
    rec = he.records[np.lexsort(he.records[:, :3].T[::-1])]
    ref = samples[np.lexsort(samples[:, :3].T[::-1])]
    assert np.array_equal(rec[:, :5].view(np.uint64), ref.view(np.uint64)) and not rec[:, 5:7].any()
    o = he.plasma(np.array([[2.15e7, 2e5, 2e5]]))
    assert np.array_equal(o[0, 0:2], out[7:9]) and np.array_equal(o[0, 8:10], out[9:11])  # qs, ms as libgfortran read them
