"""TS05 external field on the GPU (use_tsyganenko = 1), through the C ABI, against the CPU
oracle whose T04_s is the reference's own source translated into oracle/_ref: BIT-EXACT (both sides on csrc/srt_math.h)."""
import os

import numpy as np
import pytest

import common
import oracle_lib as ol
from test_ts05_cpu import PARMOD, points

pytestmark = pytest.mark.gpu
GRID13 = os.path.join(common.ROOT, "tests", "golden", "grid13.txt")


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    s.init(0)
    if ol.ts05_ref(True) is None:
        pytest.skip("oracle/_ref/libts05_ref_srtmath.so not built")
    return s


@pytest.mark.parametrize("kind,use_igrf", [("ngo", 0), ("ngo", 1), ("grid", 1)])
def test_plasma_parameters_with_the_external_field(srt, kind, use_igrf):
    path = common.NGO_CFG if kind == "ngo" else GRID13
    setup = srt.ngosetup if kind == "ngo" else srt.interpsetup
    m = setup(path, 2010001, 3600000, use_igrf, 1).tsyganenko(*PARMOD[:4], W=PARMOD[4:])
    o = ol.OracleModel(kind, path, 2010001, 3600000, use_igrf).tsyganenko(PARMOD)
    off = setup(path, 2010001, 3600000, use_igrf, 0)
    x = points(3000, 11)
    a, b = m.funcPlasmaParams(x)["raw"], o.plasma(x)
    assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
    c = off.funcPlasmaParams(x)["raw"]
    assert not np.array_equal(a[:, 16:19], c[:, 16:19]) and np.array_equal(a[:, :16].view(np.uint64), c[:, :16].view(np.uint64))


def test_switching_the_term_on_after_setup_and_parameter_change(srt):
    """srt_model_tsyganenko on a model set up without the term, then with other parameters: the device copy follows"""
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    x = points(200, 12)
    base = m.funcPlasmaParams(x)["raw"]
    m.tsyganenko(*PARMOD[:4], W=PARMOD[4:])
    o = ol.OracleModel("ngo", common.NGO_CFG).tsyganenko(PARMOD)
    assert np.array_equal(m.funcPlasmaParams(x)["raw"].view(np.uint64), o.plasma(x).view(np.uint64))
    other = [1.5, -5.0, 0.0, 3.0, 0, 0, 0, 0, 0, 0]
    m.tsyganenko(*other[:4], W=other[4:])
    o2 = ol.OracleModel("ngo", common.NGO_CFG).tsyganenko(other)
    got = m.funcPlasmaParams(x)["raw"]
    assert np.array_equal(got.view(np.uint64), o2.plasma(x).view(np.uint64)) and not np.array_equal(got, base)
    assert m.kernel_info(1000)["block"] <= 512  # the kernels with the external field exist in the 128-register budget


@pytest.mark.parametrize("mode", ["thread", "team"])
def test_rays_through_the_external_field(srt, mode, monkeypatch):
    monkeypatch.setenv("SRT_KERNEL", mode)
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 1).tsyganenko(*PARMOD[:4], W=PARMOD[4:])
    o = ol.OracleModel("ngo", common.NGO_CFG).tsyganenko(PARMOD)
    pos0, dir0, w0 = common.c2_grid(4, 4)
    params = dict(common.C2, maxsteps=400)
    res = srt.raytracer_run(m, pos0, dir0, w0, **params)
    for r in range(len(res)):
        t = o.trace(pos0[r], [0, 0, 0], w0[r], **params)
        rows = res[r].rows
        assert rows.shape == t["rows"].shape and res.stopcond[r] == t["stopcond"], r
        assert np.array_equal(rows.view(np.uint64), t["rows"].view(np.uint64)), r
        assert res.nrejected[r] == t["nrejected"] + t["ncone"]


def test_command_line_with_the_drivers_tsyganenko_options(srt, tmp_path):
    """raytracer --use_tsyganenko=1 --tsyganenko_Pdyn=... (raytracer_driver.f95:262-340): output file byte-identical to the oracle's rows"""
    import subprocess
    import test_gpu_cli as cli
    om = ol.OracleModel("ngo", common.NGO_CFG).tsyganenko(PARMOD)
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    pos0, dir0, w0 = common.launch_grid([800.0, 5000.0], [25.0, 50.0])
    rays, outp = tmp_path / "rays.txt", tmp_path / "out.ray"
    cli._write_rays(rays, pos0, dir0, w0)
    params = dict(common.C2, maxsteps=120)
    names = ("Pdyn", "Dst", "ByIMF", "BzIMF", "W1", "W2", "W3", "W4", "W5", "W6")
    args = [cli.EXE, "--dt0=%r" % params["dt0"], "--dtmax=%r" % params["dtmax"], "--tmax=%r" % params["tmax"], "--root=%d" % params["root"],
            "--fixedstep=%d" % params["fixedstep"], "--maxerr=%r" % params["maxerr"], "--maxsteps=%d" % params["maxsteps"],
            "--minalt=%r" % params["minalt"], "--inputraysfile=%s" % rays, "--outputfile=%s" % outp, "--yearday=2010001", "--milliseconds_day=0",
            "--modelnum=1", "--ngo_configfile=%s" % common.NGO_CFG, "--use_igrf=0", "--use_tsyganenko=1"] + ["--tsyganenko_%s=%r" % (k, v) for k, v in zip(names, PARMOD)]
    r = subprocess.run(args, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr
    want = cli._oracle_text(om, pos0, w0, params, 1, m.qs, m.ms, m.nus)
    assert open(outp, "rb").read() == want
