#!/usr/bin/env python
"""Exact dynamic call counts per accepted step / rejected attempt for the bench workloads -> profiles/work_tally.json
("hand counts ... replaced by exact tallies").

    python tools/work_tally.py [grid80.txt]

Runs strided samples of the launch sets through oracle/_build/libsrt_tally.so: the de-duplicated evaluation order of the
kernels (csrc/srt_physics.h) compiled for the host with call counters (-DSRT_TALLY).  Steps of rays whose state has turned
NaN are left out (the kernels fast-forward them).  What is tallied exactly: evaluations of the plasma model, of the
plasma-derived quantities, of the Stix set, of the dispersion-function tail, root solves, handedness tests, IGRF and
tricubic evaluations, and calls of every elementary function.  What stays a source count: the
IEEE operations inside one unit.  Convention of SURVEY.md 8(d): IEEE add/sub/mul/div/sqrt = 1 flop, a transcendental
call = 40 flops, FMA never assumed; fp32 IGRF operations count half."""
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import common  # noqa: E402
import oracle_lib as ol  # noqa: E402

NAMES = ["div", "sqrt", "exp", "pow", "sincos", "atan2", "acos", "csqrt", "plasma", "derive", "stix", "ftail", "root", "hand", "igrf", "tricubic"]
# IEEE FP64 operations inside one unit, outside the elementary functions counted separately
UNIT_FLOPS = {
    "plasma_ngo": 118.0,     # Ngo adapter geometry ~20 + dens live code ~49 + dipole, two rotations, fp32 round trip ~45 (+ its transcendentals, tallied)
    "plasma_grid": 1100.0,   # 4 species x 64 terms x (3 mul + 1 add) + cell search and local offsets (+ 4 exp, tallied)
    "igrf_fp32": 0.5 * (105 * 23 + 60),  # <= 105 harmonic terms x ~23 fp32 operations + rotations, at half weight
    "derive": 35.0,          # wps2, wcs (12 divisions), |B0|, free-space threshold
    "stix": 53.0,            # 12 divisions + sums
    "ftail": 33.0,           # dispersion_relation after the Stix set, incl. the k C / w scaling
    "root": 120.0,           # Booker quartic in complex arithmetic (+ 3 complex sqrt, acos: tallied)
    "hand": 130.0,           # 3x3 symmetric eigen-solve of the float-rounded wave matrix
    "attempt": 360.0,        # RK45 stage algebra + error norms
}
T_FLOPS = 40.0


def tally(model, pos0, dir0, w0, idx, params):
    E = model.E
    n = E.emu_tally_slots()
    assert n == len(NAMES)
    E.emu_trace_tally.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_double] * 4 + [C.c_int, C.c_double, C.c_int, C.c_double, C.c_int, C.c_void_p]
    tot = np.zeros((2, n + 1), dtype=np.int64)
    for r in idx:
        out = np.zeros((2, n + 1), dtype=np.int64)
        d0 = np.array(dir0[r], dtype=np.float64)
        E.emu_trace_tally(model.h, pos0[r].ctypes.data, d0.ctypes.data, w0[r], params["dt0"], params["dtmax"], params["maxerr"], params["maxsteps"],
                          params["minalt"], params["root"], params["tmax"], params["fixedstep"], out.ctypes.data)
        tot += out
    return tot


def summarise(tot, kind):
    res = {}
    for k, label in enumerate(("accepted", "rejected")):
        nat = max(int(tot[k, -1]), 1)
        per = {nm: tot[k, i] / nat for i, nm in enumerate(NAMES)}
        # one srt_sincos call stands for the reference's separate sin() and cos() of that angle: 2 calls where both values are
        # used, 1 where only the cosine is (Ngo adapter: sin+cos of the polar angle, cos of the latitude, sin+cos in dens,
        # sin+cos of the azimuth = 7 calls from 4; is_right_handed: 2 from 1)
        T = per["exp"] + per["pow"] + per["atan2"] + per["acos"] + per["csqrt"] + (1.75 if kind == "ngo" else 2.0) * per["sincos"]
        flops = (per["plasma"] * UNIT_FLOPS["plasma_" + kind] + per["igrf"] * UNIT_FLOPS["igrf_fp32"] + per["derive"] * UNIT_FLOPS["derive"] +
                 per["stix"] * UNIT_FLOPS["stix"] + per["ftail"] * UNIT_FLOPS["ftail"] + per["root"] * UNIT_FLOPS["root"] + per["hand"] * UNIT_FLOPS["hand"] +
                 UNIT_FLOPS["attempt"])
        res[label] = {"attempts_tallied": int(tot[k, -1]), "calls_per_attempt": {k2: round(v, 3) for k2, v in per.items()}, "transcendental_calls": round(T, 2),
                      "ieee_flops": round(flops, 1), "flop_equivalents": round(flops + T_FLOPS * T, 1)}
    return res


class TallyModel(ol.DedupModel):
    def __init__(self, kind, path, use_igrf=0):
        p = os.path.join(ol.ORC_DIR, "_build", "libsrt_tally.so")
        if not os.path.exists(p):
            ol.build()
        E = C.CDLL(p)
        for f in (E.emu_model_ngo, E.emu_model_grid):
            f.restype = C.c_void_p
            f.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int]
        self.E = E
        fn = E.emu_model_ngo if kind == "ngo" else E.emu_model_grid
        self.h = fn(path.encode(), 2010001, 0, use_igrf | (2 if kind != "ngo" else 0))
        assert self.h


def main():
    grid = sys.argv[1] if len(sys.argv) > 1 else "/tmp/grid80_cpu.txt"
    out = {"source": "exact call counts per attempt from tools/work_tally.py (host build of the kernels' physics with counters) x per-unit IEEE operation counts of "
                     "SURVEY.md Appendix C; 40 flops per transcendental call; steps of NaN states excluded", "unit_flops": UNIT_FLOPS, "workloads": {}}
    ngo = TallyModel("ngo", common.NGO_CFG)
    pos0, dir0, w0 = common.c2_grid(100, 100, 1)
    t = summarise(tally(ngo, pos0, dir0, w0, np.linspace(0, len(w0) - 1, 64).astype(int), common.C2), "ngo")
    out["workloads"]["c2"] = t
    pos0, dir0, w0 = common.c2_grid(100, 100, 10)
    out["workloads"]["c4_100k"] = summarise(tally(ngo, pos0, dir0, w0, np.linspace(0, len(w0) - 1, 64).astype(int), common.C2), "ngo")
    if os.path.exists(grid):
        gm = TallyModel("grid", grid, use_igrf=1)
        out["workloads"]["c3"] = summarise(tally(gm, pos0, dir0, w0, np.linspace(0, len(w0) - 1, 32).astype(int), common.C2), "grid")
    # what bench.py reads: flop-equivalents per accepted step / rejected attempt by model kind
    out["ngo"] = {"accepted": out["workloads"]["c2"]["accepted"]["flop_equivalents"], "rejected": out["workloads"]["c2"]["rejected"]["flop_equivalents"]}
    if "c3" in out["workloads"]:
        out["grid"] = {"accepted": out["workloads"]["c3"]["accepted"]["flop_equivalents"], "rejected": out["workloads"]["c3"]["rejected"]["flop_equivalents"]}
    else:
        out["grid"] = {"accepted": 6.8e4, "rejected": 6.5e4}
    path = os.path.join(ROOT, "profiles", "work_tally.json")
    json.dump(out, open(path, "w"), indent=1)
    for k, v in out["workloads"].items():
        print(k, "accepted: %.0f flop-eq (%.0f IEEE + %.0f T calls), rejected: %.0f" % (v["accepted"]["flop_equivalents"], v["accepted"]["ieee_flops"],
                                                                                      v["accepted"]["transcendental_calls"], v["rejected"]["flop_equivalents"]))
        print("   calls per accepted step:", v["accepted"]["calls_per_attempt"])


if __name__ == "__main__":
    main()
