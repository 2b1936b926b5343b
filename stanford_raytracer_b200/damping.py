"""Landau damping along the ray on the GPU: the reference's `damping` post-processor (damping/src/damping_main.cpp) in its
default mode 1 (damping_foust), for every ray of an output file or of a BatchResult.

    python -m stanford_raytracer_b200.damping --inp_file=out.ray --out_file=out.damp [--AE=3 --Kp=4 --yearday=2010001 --msec=0 --geom_factor=0]

Option names and defaults are the reference's (damping_main.cpp:44-98); `--mode` must be 1 (the legacy 2-D mode 0 is
not on the path).  The output file is what write_damping emits (wipp_fileutils.cpp:234-252): `ray  time  power` as
`%d  %.15e  %.15e` per recorded row, rays in ascending ray number."""
import ctypes as C
import os
import sys

import numpy as np

from . import lib, _check, _ptr, rayfile, SrtError

CRRES_TABLE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data", "crres_clean.bin")


def damping_foust(rows_per_ray, w, qs, ms, Kp=4.0, AE_level=3.0, yearday=2010001, msec=0, geom_factor=False, crres_table=CRRES_TABLE):
    """rows_per_ray: list of [n_i][20] row arrays (t, pos, vprel, vgrel, n, B0, Ns); w[nrays]; qs, ms[nspec].
    Returns the list of power arrays [n_i]."""
    L = lib()
    L.srt_damping_foust.argtypes = [C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double,
                                    C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_void_p]
    L.srt_damping_last_error.restype = C.c_char_p
    n = [len(r) for r in rows_per_ray]
    first = np.concatenate([[0], np.cumsum(n)[:-1]]).astype(np.int64) if n else np.zeros(0, dtype=np.int64)
    rows = np.ascontiguousarray(np.concatenate([np.asarray(r, dtype=np.float64).reshape(-1, 20) for r in rows_per_ray], axis=0)) if n else np.zeros((0, 20))
    w = np.ascontiguousarray(w, dtype=np.float64)
    qs, ms = np.ascontiguousarray(qs, dtype=np.float64), np.ascontiguousarray(ms, dtype=np.float64)
    power = np.empty(len(rows))
    rc = L.srt_damping_foust(len(n), _ptr(first), len(rows), _ptr(rows), _ptr(w), len(qs), _ptr(qs), _ptr(ms), float(Kp), float(AE_level), int(yearday),
                             int(msec), int(bool(geom_factor)), os.fsencode(crres_table), _ptr(power))
    if rc:
        raise SrtError("srt_damping_foust failed: %s [%s]" % (L.srt_strerror(rc).decode(), L.srt_damping_last_error().decode()))
    return [power[f:f + k] for f, k in zip(first, n)]


def damp_result(res, Kp=4.0, AE_level=3.0, yearday=2010001, msec=0, geom_factor=False, crres_table=CRRES_TABLE):
    """Damping for every ray of a BatchResult (rows stay in the library).  Returns (first_row[nrays], power[total_rows])."""
    L = lib()
    L.srt_result_damping.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int, C.c_char_p, C.c_void_p, C.c_void_p]
    L.srt_damping_last_error.restype = C.c_char_p
    first = np.zeros(res.nrays, dtype=np.int64)
    power = np.empty(res.total_rows)
    w0 = np.ascontiguousarray(res.w0, dtype=np.float64)
    rc = L.srt_result_damping(res._h, res.model._h, _ptr(w0), float(Kp), float(AE_level), int(yearday), int(msec), int(bool(geom_factor)),
                              os.fsencode(crres_table), _ptr(first), _ptr(power))
    if rc:
        raise SrtError("srt_result_damping failed: %s [%s]" % (L.srt_strerror(rc).decode(), L.srt_damping_last_error().decode()))
    return first, power


def damp_rayfile(inp_file, out_file, **opts):
    """The post-processor on an output file of the raytracer; returns {raynum: (time, power)}."""
    rays = rayfile.read_output(inp_file)
    nums = sorted(rays)  # the reference keeps the rays in a std::map: ascending ray number
    if not nums:
        open(out_file, "w").close()
        return {}
    first = rays[nums[0]]
    rows = []
    for k in nums:
        r = rays[k]
        ns = np.zeros((len(r["time"]), 4))
        ns[:, :r["nspec"]] = r["Ns"]
        rows.append(np.concatenate([r["time"][:, None], r["pos"], r["vprel"], r["vgrel"], r["n"], r["B0"], ns], axis=1))
    power = damping_foust(rows, [rays[k]["w"] for k in nums], first["qs"], first["ms"], **opts)
    write_damping(out_file, nums, [rays[k]["time"] for k in nums], power)
    return {k: (rays[k]["time"], p) for k, p in zip(nums, power)}


def write_damping(path, raynums, times, powers):
    """wipp_fileutils.cpp:234-252"""
    with open(path, "w") as f:
        for k, t, p in zip(raynums, times, powers):
            for a, b in zip(t, p):
                f.write("%d  %.15e  %.15e\n" % (k, a, b))


def main(argv=None):
    opts = dict(inp_file="input.ray", out_file="output.damp", mode="1", AE="3", Kp="4", yearday="2010001", msec="0", geom_factor="0")
    for a in (sys.argv[1:] if argv is None else argv):
        if a.startswith("--") and "=" in a:
            k, v = a[2:].split("=", 1)
            if k in opts:
                opts[k] = v
    if int(opts["mode"]) != 1:
        raise SystemExit("only --mode=1 (damping_foust) is implemented on the GPU path")
    damp_rayfile(opts["inp_file"], opts["out_file"], Kp=float(opts["Kp"]), AE_level=float(opts["AE"]), yearday=int(opts["yearday"]), msec=int(opts["msec"]),
                 geom_factor=int(opts["geom_factor"]) == 1)


if __name__ == "__main__":
    main()
