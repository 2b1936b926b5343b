"""Damping along the ray (config 5), CPU side.
  * oracle/_ref/damping is the REFERENCE's own post-processor (damping/src/*.cpp compiled where they lie, oracle/Makefile
    `ref`; MATLAB's libmat and the Fortran transforms replaced by oracle/ref_stubs/).  Its outputs on a committed ray file
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
CASES = [("damping_ref_AE3_Kp4_geom0.damp", ["-a", "3", "-k", "4", "-v", "0", "-u", "0"]),
         ("damping_ref_AE1_Kp2.5_geom1.damp", ["-a", "1", "-k", "2.5", "-v", "1", "-u", "43200000"])]


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
    return L


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
