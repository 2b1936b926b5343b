"""Damping along the ray (config 5), CPU side.
  * oracle/_ref/damping is the REFERENCE's own post-processor (damping/src/*.cpp compiled where they lie, oracle/Makefile
    `ref`; MATLAB's libmat replaced by oracle/ref_stubs/, the Fortran transforms by the reference's own xform_double translated).  Its outputs on a committed ray file
    are the golden vectors tests/golden/damping_ref_*.damp; here the binary must reproduce them byte for byte (the same
    binary is the checker of the GPU path in tests/test_gpu_damping.py).
  * the host set-up of the GPU kernels (Gauss-Legendre table, SM->MAG matrix, Greenwich longitude) against numpy and
    against the stand-ins the reference build links."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import common

REF = os.path.join(common.ROOT, "oracle", "_ref", "damping")
GOLD = os.path.join(common.ROOT, "tests", "golden")
# (the reference's getopt string "i:o:m:a:k:t:u:" lacks v, so `-v 1` is ignored by damping_main.cpp:80: the long option works)
CASES = [("damping_ref_AE3_Kp4_geom0.damp", ["-a", "3", "-k", "4", "--geom_factor", "0", "-u", "0"]),
         ("damping_ref_AE1_Kp2.5_geom1.damp", ["-a", "1", "-k", "2.5", "--geom_factor", "1", "-u", "43200000"]),
         ("damping_ref_AE2_Kp5.5_geom0.damp", ["-a", "2", "-k", "5.5", "--geom_factor", "0", "-u", "21600000"])]


def case_options(args):
    kw = dict(zip(args[0::2], args[1::2]))
    return dict(AE_level=float(kw["-a"]), Kp=float(kw["-k"]), geom_factor=kw["--geom_factor"] == "1", msec=int(kw["-u"]), yearday=2010001)


def run_reference_damping(inp, out, args, cwd):
    """runs oracle/_ref/damping in `cwd` (it opens "crres_clean.mat" there: the stand-in reads crres_clean.bin)"""
    table = os.path.join(cwd, "crres_clean.bin")
    if not os.path.exists(table):
        import shutil
        shutil.copy(os.path.join(common.ROOT, "oracle", "_ref", "crres_clean.bin"), table)
    subprocess.run([REF, "-i", inp, "-o", out] + args, cwd=cwd, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)


@pytest.fixture(scope="module")
def ref_binary():
    if not os.path.exists(REF):
        subprocess.check_call(["make", "-s", "-C", os.path.join(common.ROOT, "oracle"), "ref"])
    assert os.path.exists(REF), "oracle/_ref/damping is built from /root/reference by `make -C oracle ref`"
    return REF


@pytest.mark.parametrize("name,args", CASES)
def test_reference_binary_reproduces_golden(ref_binary, tmp_path, name, args):
    out = str(tmp_path / "out.damp")
    run_reference_damping(os.path.join(GOLD, "damping_in.ray"), out, args, str(tmp_path))
    assert open(out).read() == open(os.path.join(GOLD, name)).read()


def test_crres_table_is_the_packaged_one():
    a = open(os.path.join(common.ROOT, "oracle", "_ref", "crres_clean.bin"), "rb").read()
    b = open(os.path.join(common.ROOT, "stanford_raytracer_b200", "data", "crres_clean.bin"), "rb").read()
    assert a == b and len(a) == 16 + 12 * 8 * (2 + 70 + 24 + 70 * 24)


@pytest.fixture(scope="module")
def probe():
    os.makedirs(common.BUILD, exist_ok=True)
    out = os.path.join(common.BUILD, "damping_probe.so")
    srcs = [os.path.join(common.ROOT, "tests", "csrc", "damping_probe.cpp"), os.path.join(common.ROOT, "oracle", "ref_stubs", "xform_d.cpp"),
            os.path.join(common.ROOT, "stanford_raytracer_b200", "csrc", "srt_damping_setup.h")]
    if common._newer(out, srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-o", out, srcs[0], srcs[1]])
    L = C.CDLL(out)
    L.probe_greenwich.restype = C.c_double
    L.probe_set_perturbation.argtypes = [C.c_double] * 3
    L.probe_set_noise.argtypes = [C.c_double, C.c_uint64]
    L.probe_damping.argtypes = [C.c_long, C.c_void_p, C.c_long, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int,
                                C.c_int, C.c_char_p, C.c_void_p]
    return L


def load_rows(path):
    """rows of an output file as the damping entry points take them: (first[nrays], rows[n][20], w[nrays], qs, ms, raynums, times)"""
    from stanford_raytracer_b200 import rayfile
    rays = rayfile.read_output(path)
    nums = sorted(rays)
    rows = []
    for k in nums:
        r = rays[k]
        ns = np.zeros((len(r["time"]), 4))
        ns[:, :r["nspec"]] = r["Ns"]
        rows.append(np.concatenate([r["time"][:, None], r["pos"], r["vprel"], r["vgrel"], r["n"], r["B0"], ns], axis=1))
    n = [len(r) for r in rows]
    first = np.concatenate([[0], np.cumsum(n)[:-1]]).astype(np.int64)
    return first, np.ascontiguousarray(np.concatenate(rows)), np.array([rays[k]["w"] for k in nums]), np.ascontiguousarray(rays[nums[0]]["qs"]), np.ascontiguousarray(rays[nums[0]]["ms"])


def host_damping(probe, path, AE_level, Kp, geom_factor, msec, yearday, perturb=(0.0, 0.0, 0.0), noise=(0.0, 0)):
    """the GPU path's arithmetic (csrc/srt_damping_row.h) run on the host with glibc (tests/csrc/damping_probe.cpp)"""
    first, rows, w, qs, ms = load_rows(path)
    power = np.zeros(len(rows))
    table = os.path.join(common.ROOT, "stanford_raytracer_b200", "data", "crres_clean.bin")
    probe.probe_set_perturbation(*perturb)
    probe.probe_set_noise(*noise)
    rc = probe.probe_damping(len(first), first.ctypes.data, len(rows), rows.ctypes.data, w.ctypes.data, len(qs), qs.ctypes.data, ms.ctypes.data, Kp, AE_level,
                             yearday, msec, int(geom_factor), table.encode(), power.ctypes.data)
    probe.probe_set_perturbation(0.0, 0.0, 0.0)
    probe.probe_set_noise(0.0, 0)
    assert rc == 0
    return power


def conditioning(probe, path, **opts):
    """Per row: how well the reference's own answer is determined.  The largest relative change of the power when (a) ONE
    fitted density parameter (power-law exponent, its amplitude, the plasmapause L) is moved by one unit in the last place,
    (b) every pow / log / exp of the density model gets a +-1 ulp error of its own (what separates two math libraries).
    The integrand differentiates the density with a relative step of 1e-8 (integrand.cpp:99-160), the suprathermal
    density a/v^4 - b/v^5 + c/v^6 cancels around v = 1.7e7 m/s, and the blended density is exp(mean of logs): the result
    is determined to ~1e-8 for benign rows, ~1e-5 for rays whose resonant velocities sit in the cancellation, and up to
    1e-2 for rays that cross the plasmapause region far out (measured)."""
    base = host_damping(probe, path, **opts)
    cond = np.zeros(len(base))
    eps = 2.220446049250313e-16
    runs = [dict(perturb=p) for p in ((eps, 0, 0), (0, eps, 0), (0, 0, eps), (-eps, 0, 0))] + [dict(noise=(eps, seed)) for seed in (1, 2, 3)]
    for kw in runs:
        p = host_damping(probe, path, **kw, **opts)
        with np.errstate(divide="ignore", invalid="ignore"):
            c = np.abs(p / base - 1.0)
        cond = np.maximum(cond, np.where(np.isfinite(c), c, 0.0))
    return base, cond


def assert_power_agrees(got, ref, cond, what, slack=20.0, floor=1e-6):
    """|got / ref - 1| <= floor + slack * (1-ulp conditioning of the row; slack = 20: CUDA documents up to 2 ulp for pow) wherever the reference's power is above its own
    cut-off 1e-6 (WAVE_PWR_THRESH, consts.h); NaN rows agree."""
    assert np.array_equal(np.isfinite(got), np.isfinite(ref)), what
    big = np.isfinite(ref) & (ref > 1e-6)
    rel = np.abs(got[big] / ref[big] - 1.0)
    bound = floor + slack * cond[big]
    bad = rel > bound
    assert not bad.any(), (what, int(bad.sum()), float(rel[bad].max()), float(bound[bad].min()))
    return rel


@pytest.mark.parametrize("name,args", CASES)
def test_gpu_arithmetic_on_the_host_against_the_reference(ref_binary, probe, name, args):
    """same inputs, same libm, two statements of the algorithm (the reference's; the kernels' row arithmetic run sequentially):
    they agree to the rows' own conditioning, and to 1e-8 in the median"""
    path = os.path.join(GOLD, "damping_in.ray")
    ref = np.loadtxt(os.path.join(GOLD, name)).reshape(-1, 3)[:, 2]
    base, cond = conditioning(probe, path, **case_options(args))
    rel = assert_power_agrees(base, ref, cond, name)
    assert np.median(rel) < 1e-5 and ref.min() < 0.9  # one ill-conditioned row moves every later row of its ray (running product)


def test_gauss_legendre_table(probe):
    x, w = np.zeros(25), np.zeros(25)
    probe.probe_gl(50, x.ctypes.data, w.ctypes.data)
    xr, wr = np.polynomial.legendre.leggauss(50)
    # the reference stops Newton at 1e-10 and takes the weight with the derivative of the previous iterate: abscissae are
    # exact to rounding, weights to ~1e-10 -- reproduced, not improved
    assert np.allclose(x, xr[25:], rtol=0, atol=1e-14) and np.allclose(w, wr[25:], rtol=1e-9, atol=0)
    assert abs(2 * w.sum() - 2.0) < 1e-9


@pytest.mark.parametrize("yearday,msec", [(2010001, 0), (2010001, 43200000), (1999200, 7200000)])
def test_sm_to_mag_matrix_equals_the_transform_chain(probe, yearday, msec):
    M = np.zeros(9)
    probe.probe_sm2mag(yearday, msec, M.ctypes.data)
    M = M.reshape(3, 3)
    assert np.allclose(M @ M.T, np.eye(3), atol=1e-14)
    rng = np.random.default_rng(3)
    for _ in range(20):
        x = rng.uniform(-5e7, 5e7, 3)
        o = np.zeros(3)
        probe.probe_ref_sm2mag(yearday, msec, x.ctypes.data, o.ctypes.data)
        assert np.allclose(M @ x, o, rtol=0, atol=1e-7)  # metres
    # Greenwich (51.48 N, 0 E) in MAG
    lat = np.deg2rad(51.48)
    g = np.array([np.cos(lat), 0.0, np.sin(lat)])
    o = np.zeros(3)
    probe.probe_ref_geo2mag(yearday, msec, g.ctypes.data, o.ctypes.data)
    assert abs(probe.probe_greenwich(yearday, msec) - np.rad2deg(np.arctan2(o[1], o[0]))) < 1e-12
