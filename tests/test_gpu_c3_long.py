"""Config 3 at full length on the real lattice (80^3 tricubic grid over +-10 R_E sampled from the Ngo model, IGRF field):
whole rays to maxsteps = 20 000 against the oracle, bit for bit -- the population that broke config 3 in round 1:
rays that live 10^4 steps, leave the lattice, turn NaN and are stepped on to maxsteps (NaN fast-forward), end by the
dt < 1e-14 stop (code 5) after a refinement collapse, by altitude, by group velocity, by tmax.  Chosen on the host by
tools/c3_scan_cpu.py (every 97th ray of the 100 000).  Both kernel forms; the coefficient table, and (fewer, shorter
rays) the on-the-fly tricubic path SRT_GRID_COEF=0."""
import os
import tempfile
import threading

import numpy as np
import pytest

import common
import oracle_lib as ol
from stanford_raytracer_b200 import gridfile

pytestmark = pytest.mark.gpu
R_E = common.R_E

NAN_RAYS = [35502, 38509, 44620, 54417]          # state turns NaN, stepped to maxsteps (stop code 6)
OUTSIDE = [21825, 31719, 55678, 56648]            # end outside the lattice (|x| > 10 R_E on an axis) / beyond 6 R_E
CODE5 = [48500, 68288, 46463, 64505, 94381]       # dt < 1e-14 after 700 ... 16 700 rows
CODE6 = [37442, 38412, 39479, 99716]              # 20 000 rows, finite state
CODE0 = [16296, 13289, 32398, 3686]               # tmax after 1 800 ... 17 600 rows
CODE1 = [41322, 81965, 77600, 97097]              # below minalt after 50 ... 5 200 rows
CODE3 = [6402, 5529]                              # |vgrel| > 1.01
CODE2 = [12319, 79346]                            # evanescent start: one row
OTHERS = [97, 10088, 50052]
RAYS = NAN_RAYS + OUTSIDE + CODE5 + CODE6 + CODE0 + CODE1 + CODE3 + CODE2 + OTHERS


@pytest.fixture(scope="module")
def srt(built):
    import stanford_raytracer_b200 as s
    s.init(0)
    return s


@pytest.fixture(scope="module")
def grid80(srt):
    path = os.path.join(tempfile.gettempdir(), "srt_test_grid80.txt")
    if not os.path.exists(path):
        ngo = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
        ext = (-10 * R_E, 10 * R_E) * 3
        f, qs, ms = gridfile.sample_grid(lambda x: ngo.funcPlasmaParams(x)["raw"], 80, 80, 80, ext)
        gridfile.write_grid(path, f, qs, ms, ext)
    return path


def oracle_rays(path, pos0, w0, params, nthreads=0):
    """the oracle's rows of every ray (a pool of workers, one oracle model each)"""
    out = [None] * len(w0)
    nthreads = nthreads or max(4, min(16, os.cpu_count() or 4))
    models = [ol.OracleModel("grid", path, use_igrf=1) for _ in range(min(nthreads, len(w0)))]
    lock, head = threading.Lock(), [0]

    def work(t):
        while True:
            with lock:
                j = head[0]
                head[0] += 1
            if j >= len(w0):
                return
            out[j] = models[t].trace(pos0[j], [0, 0, 0], w0[j], **params)  # ~1.6 ms per step: a 20 000-row ray takes half a minute

    ts = [threading.Thread(target=work, args=(i,)) for i in range(len(models))]
    [t.start() for t in ts]
    [t.join() for t in ts]
    return out


@pytest.fixture(scope="module")
def launch():
    pos0, dir0, w0 = common.c2_grid(100, 100, 10)
    return pos0[RAYS], w0[RAYS]


@pytest.fixture(scope="module")
def oracle_full(grid80, launch):
    return oracle_rays(grid80, launch[0], launch[1], dict(common.C2))


@pytest.mark.parametrize("mode", ["team", "thread"])
def test_long_rays_on_the_real_lattice(srt, grid80, launch, oracle_full, mode, monkeypatch):
    monkeypatch.setenv("SRT_KERNEL", mode)
    pos0, w0 = launch
    m = srt.interpsetup(grid80, 2010001, 0, 1, 0)
    res = srt.raytracer_run(m, pos0, np.zeros_like(pos0), w0, **common.C2)
    codes = []
    for j, r in enumerate(RAYS):
        o = oracle_full[j]
        assert res.stopcond[j] == o["stopcond"], (r, res.stopcond[j], o["stopcond"])
        assert res.nsteps[j] == len(o["rows"]), (r, res.nsteps[j], len(o["rows"]))
        assert res.nrejected[j] == o["nrejected"] + o["ncone"], r
        assert np.array_equal(res.rows(j), o["rows"], equal_nan=True), r
        codes.append(o["stopcond"])
    # the population is what the header says
    nan = [j for j in range(len(RAYS)) if np.isnan(oracle_full[j]["rows"][-1][1:4]).any()]
    assert len(nan) >= 2 and all(len(oracle_full[j]["rows"]) == 20000 for j in nan)
    outside = [j for j in range(len(RAYS)) if j not in nan and (np.abs(oracle_full[j]["rows"][:, 1:4]) > 10 * R_E).any()]
    assert len(outside) >= 2
    assert codes.count(5) >= 2 and {0, 1, 2, 3, 5, 6} <= set(codes)
    assert max(len(o["rows"]) for j, o in enumerate(oracle_full) if j not in nan) == 20000
    assert res.fastforwarded_steps > 10000  # the NaN rays' tails were written by the fast-forward
    m.close()


def test_on_the_fly_tricubic_path_on_old_rays(srt, grid80, launch, monkeypatch):
    """SRT_GRID_COEF=0 (no coefficient table: every query builds the cell's coefficients, edge clamps included) on rays that
    leave the lattice and turn NaN, first 2 500 rows, against the oracle."""
    monkeypatch.setenv("SRT_GRID_COEF", "0")
    pick = [RAYS.index(r) for r in (35502, 21825, 31719, 48500, 16296, 97)]
    pos0, w0 = launch[0][pick], launch[1][pick]
    p = dict(common.C2, maxsteps=2500)
    m = srt.interpsetup(grid80, 2010001, 0, 1, 0)
    res = srt.raytracer_run(m, pos0, np.zeros_like(pos0), w0, **p)
    ref = oracle_rays(grid80, pos0, w0, p)
    for j in range(len(w0)):
        assert res.stopcond[j] == ref[j]["stopcond"] and np.array_equal(res.rows(j), ref[j]["rows"], equal_nan=True), pick[j]
    m.close()
