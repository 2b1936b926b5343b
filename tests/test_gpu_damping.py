"""Damping along the ray on the GPU (srt_damping_foust, csrc/srt_damping.cu) against the REFERENCE's own post-processor
(oracle/_ref/damping = damping/src/*.cpp compiled unchanged, see tests/test_damping_cpu.py) -- the one row of the scope
table whose checker is the reference itself.

Tolerance (floating point): the integrand differentiates the phase-space density with a relative step of 1e-8
(integrand.cpp:99-160), so last-bit differences between CUDA's and glibc's pow/log/exp/jn -- or between two algebraically
equal statements on one libm -- are amplified ~1e8-fold in each gradient.  How well the reference's own answer is
determined is measured per row (test_damping_cpu.conditioning: the change of the power when one fitted parameter moves by
one unit in the last place: 1e-8 for benign rows, up to 1e-2 for rays crossing the plasmapause region far out).  Bar, on
every row whose reference power is above the reference's own cut-off 1e-6 (WAVE_PWR_THRESH):
    |power_gpu / power_ref - 1| <= 1e-6 + 20 x (1-ulp conditioning of that row);
NaN rows agree; ray numbers and times are copied, so equal."""
import os

import numpy as np
import pytest

import common
from test_damping_cpu import CASES, GOLD, REF, run_reference_damping, case_options, conditioning, assert_power_agrees, probe  # noqa: F401

pytestmark = pytest.mark.gpu


def read_damp(path):
    a = np.loadtxt(path).reshape(-1, 3)
    return a[:, 0].astype(int), a[:, 1], a[:, 2]


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    s.init(0)
    return s


@pytest.mark.parametrize("name,args", CASES)
def test_golden_ray_file(srt, probe, tmp_path, name, args):
    from stanford_raytracer_b200 import damping
    out = str(tmp_path / "gpu.damp")
    inp = os.path.join(GOLD, "damping_in.ray")
    damping.damp_rayfile(inp, out, **case_options(args))
    rg, tg, pg = read_damp(out)
    rr, tr, pr = read_damp(os.path.join(GOLD, name))
    assert np.array_equal(rg, rr) and np.array_equal(tg, tr)
    _, cond = conditioning(probe, inp, **case_options(args))
    rel = assert_power_agrees(pg, pr, cond, name)
    assert pr.min() < 0.9  # the rays are damped: the comparison is not of ones
    print("damping %s: %d rows, relative difference median %.1e, worst %.1e (worst 1-ulp conditioning %.1e)" % (name, len(pr), np.median(rel), rel.max(), cond.max()))


def test_rays_traced_on_the_gpu_through_the_file_formats(srt, probe, tmp_path):
    """tracer -> output file (library writer) -> damping on the GPU (module CLI) and by the reference binary"""
    from stanford_raytracer_b200 import damping
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref/damping not built")
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    pos0, dir0, w0 = common.c2_grid(6, 5)
    res = srt.raytracer_run(model, pos0, dir0, w0, outputper=20, **dict(common.C2, maxsteps=3000))
    ray = str(tmp_path / "out.ray")
    res.write_text(ray)
    damping.main(["--inp_file=" + ray, "--out_file=" + str(tmp_path / "gpu.damp"), "--Kp=5.5", "--AE=2", "--msec=21600000"])
    run_reference_damping(ray, str(tmp_path / "ref.damp"), ["-k", "5.5", "-a", "2", "-u", "21600000"], str(tmp_path))
    rg, tg, pg = read_damp(str(tmp_path / "gpu.damp"))
    rr, tr, pr = read_damp(str(tmp_path / "ref.damp"))
    assert len(pr) == res.total_rows and np.array_equal(rg, rr) and np.array_equal(tg, tr)
    assert np.isfinite(pr).sum() > 0.9 * len(pr)
    _, cond = conditioning(probe, ray, AE_level=2.0, Kp=5.5, geom_factor=False, msec=21600000, yearday=2010001)
    rel = assert_power_agrees(pg, pr, cond, "traced rays")
    print("damping of %d traced rows: relative difference median %.1e, worst %.1e" % (len(pr), np.median(rel), rel.max()))


def test_direct_call_on_a_batch_result(srt):
    """damping_foust on the rows of a BatchResult (no files): first rows are 1, power is finite and not negative"""
    from stanford_raytracer_b200 import damping
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    pos0, dir0, w0 = common.c2_grid(3, 3)
    res = srt.raytracer_run(model, pos0, dir0, w0, outputper=50, **dict(common.C2, maxsteps=2000))
    rows = [res.rows(r) for r in range(len(res))]
    power = damping.damping_foust(rows, w0, model.qs, model.ms)
    first, flat = damping.damp_result(res)
    assert np.array_equal(np.concatenate(power), flat) and np.array_equal(first, np.concatenate([[0], np.cumsum([len(r) for r in rows])[:-1]]))
    for p, r in zip(power, rows):
        assert len(p) == len(r) and p[0] == 1.0
        assert not np.any(p < 0)  # (a ray may overflow to inf or turn NaN, as in the reference: compared in the tests above)
