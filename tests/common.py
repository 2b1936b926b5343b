"""Shared helpers of the test-suite: builds, synthetic launch sets, ctypes
wrapper of the test-only host emulation of the device physics."""
import ctypes as C
import os
import subprocess

import numpy as np

import oracle_lib as ol

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "tests", "data")
NGO_CFG = os.path.join(DATA, "newray_default.in")
BUILD = os.path.join(ROOT, "tests", "_build")
R_E = 6371.2e3
MINALT = R_E + 100e3

C1 = dict(dt0=5e-4, dtmax=0.1, maxerr=5e-4, maxsteps=2000, minalt=MINALT, root=2, tmax=0.5, fixedstep=1)
C2 = dict(dt0=1e-3, dtmax=0.1, maxerr=5e-4, maxsteps=20000, minalt=MINALT, root=2, tmax=10.0, fixedstep=0)


def _newer(target, srcs):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in srcs)


def build_hostemu():
    os.makedirs(BUILD, exist_ok=True)
    csrc = os.path.join(ROOT, "stanford_raytracer_b200", "csrc")
    srcs = [os.path.join(csrc, f) for f in os.listdir(csrc)] + [os.path.join(ROOT, "tests", "csrc", "hostemu.cpp")]
    out = os.path.join(BUILD, "libsrt_hostemu.so")
    if _newer(out, srcs):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-DEMU_HAVE_GRID", "-Wno-attributes", "-shared", "-o", out,
                               os.path.join(ROOT, "tests", "csrc", "hostemu.cpp")])
    out3 = os.path.join(BUILD, "ts05_probe.so")
    src3 = os.path.join(ROOT, "tests", "csrc", "ts05_probe.cpp")
    if _newer(out3, [src3] + [os.path.join(csrc, f) for f in ("srt_ts05.h", "srt_ts05_tables.h", "srt_math.h")]):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-fno-fast-math", "-Wno-attributes", "-I" + csrc, "-shared", "-o", out3, src3])
    out2 = os.path.join(BUILD, "math_probe.so")
    src2 = os.path.join(ROOT, "tests", "csrc", "math_probe.cpp")
    if _newer(out2, [src2, os.path.join(csrc, "srt_math.h")]):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-ffp-contract=off", "-shared", "-o", out2, src2])
    return out


def build_all():
    ol.build()
    build_hostemu()
    from stanford_raytracer_b200 import build as b
    b.build()


def launch_grid(freqs_hz, lats_deg, lons_deg=(0.0,), alt_km=1000.0):
    """Lattice of launch points: frequency slowest, then latitude, then longitude.
    dir0 = 0 (field-aligned convention)."""
    pos, w = [], []
    for f in freqs_hz:
        for la in lats_deg:
            for lo in lons_deg:
                pos.append(ol.launch_point(la, lo, alt_km))
                w.append(2 * np.pi * f)
    pos = np.array(pos)
    return pos, np.zeros_like(pos), np.array(w)


def c2_grid(nf=100, nlat=100, nlon=1):
    freqs = np.logspace(np.log10(200.0), np.log10(30000.0), nf)
    lats = np.linspace(10.0, 60.0, nlat)
    lons = np.arange(nlon) * 36.0
    return launch_grid(freqs, lats, lons)


class HostEmu:
    """tests/csrc/hostemu.cpp: the device physics compiled for the CPU (test-only)."""

    def __init__(self, kind, path, yearday=2010001, msec=0, use_igrf=0, precomputed_coef=False):
        build_hostemu()
        E = C.CDLL(os.path.join(BUILD, "libsrt_hostemu.so"))
        d, i, vp, l = C.c_double, C.c_int, C.c_void_p, C.c_long
        E.emu_model_ngo.restype = vp
        E.emu_model_ngo.argtypes = [C.c_char_p, i, i, i]
        E.emu_model_grid.restype = vp
        E.emu_model_grid.argtypes = [C.c_char_p, i, i, i]
        E.emu_plasma.argtypes = [vp, l, vp, vp]
        E.emu_model_tsyganenko.argtypes = [vp, vp, i, i]
        self.yearday, self.msec = yearday, msec
        E.emu_is_right_handed.restype = i
        E.emu_is_right_handed.argtypes = [d] * 5
        E.emu_trace.restype = l
        E.emu_trace.argtypes = [vp, vp, vp, d, d, d, d, i, d, i, d, i, vp, l, vp, vp, vp, vp]
        self.E = E
        fn = E.emu_model_ngo if kind == "ngo" else E.emu_model_grid
        self.h = fn(path.encode(), yearday, msec, use_igrf | (2 if (precomputed_coef and kind != "ngo") else 0))
        if not self.h:
            raise RuntimeError("hostemu model setup failed")

    def tsyganenko(self, parmod):
        pm = np.ascontiguousarray(parmod, dtype=np.float64)
        self.E.emu_model_tsyganenko(self.h, pm.ctypes.data, self.yearday, self.msec)
        return self

    def plasma(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 3)
        out = np.empty((len(x), 19))
        self.E.emu_plasma(self.h, len(x), x.ctypes.data, out.ctypes.data)
        return out

    def is_right_handed(self, n2, phi, S, D, P):
        return bool(self.E.emu_is_right_handed(n2, phi, S, D, P))

    def trace(self, pos0, dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep):
        pos0 = np.ascontiguousarray(pos0, dtype=np.float64)
        dir0 = np.array(dir0, dtype=np.float64)
        rows = np.empty((maxsteps + 1, 20))
        sc, nr, nc, ce = C.c_int(), C.c_long(), C.c_long(), C.c_int()
        n = self.E.emu_trace(self.h, pos0.ctypes.data, dir0.ctypes.data, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep,
                             rows.ctypes.data, len(rows), C.byref(sc), C.byref(nr), C.byref(nc), C.byref(ce))
        return dict(rows=rows[:n].copy(), stopcond=sc.value, nrejected=nr.value, ncone=nc.value, fixed_cone_exit=ce.value, dir0=dir0)


class HostEmuScattered(HostEmu):
    """modelnum 4 through the product's own set-up and device code, compiled for the CPU (test-only)."""

    def __init__(self, path, window_scale, order, exact, local_window_scale, scaled=0, yearday=2010001, msec=0, use_igrf=0):
        build_hostemu()
        E = C.CDLL(os.path.join(BUILD, "libsrt_hostemu.so"))
        d, i, vp, l = C.c_double, C.c_int, C.c_void_p, C.c_long
        E.emu_model_scattered.restype = vp
        E.emu_model_scattered.argtypes = [C.c_char_p, i, i, i, d, i, i, i, d]
        E.emu_scattered_state.restype = l
        E.emu_scattered_state.argtypes = [vp, vp, l, vp, vp]
        E.emu_scattered_neighbours.restype = l
        E.emu_scattered_neighbours.argtypes = [vp, vp, vp, l]
        E.emu_scattered_lsinterp.restype = i
        E.emu_scattered_lsinterp.argtypes = [vp, vp, vp]
        E.emu_plasma.argtypes = [vp, l, vp, vp]
        E.emu_model_tsyganenko.argtypes = [vp, vp, i, i]
        E.emu_trace.restype = l
        E.emu_trace.argtypes = [vp, vp, vp, d, d, d, d, i, d, i, d, i, vp, l, vp, vp, vp, vp]
        self.E, self.yearday, self.msec = E, yearday, msec
        self.h = E.emu_model_scattered(path.encode(), yearday, msec, use_igrf, window_scale, order, exact, scaled, local_window_scale)
        if not self.h:
            raise RuntimeError("hostemu scattered model setup failed")
        info = np.zeros(5, dtype=np.int32)
        dinfo = np.zeros(2)
        E.emu_scattered_state(self.h, None, 0, info.ctypes.data, dinfo.ctypes.data)
        self.npts, self.nmono, self.cells = int(info[0]), int(info[1]), tuple(int(v) for v in info[2:5])
        self.maxnearest, self.radius = dinfo[0], dinfo[1]
        self.records = np.empty((self.npts, 8))
        E.emu_scattered_state(self.h, self.records.ctypes.data, self.records.size, info.ctypes.data, dinfo.ctypes.data)

    def neighbours(self, x):
        """indices into self.records, in the order the device code walks them"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = self.E.emu_scattered_neighbours(self.h, x.ctypes.data, None, 0)
        out = np.empty(n, dtype=np.int32)
        self.E.emu_scattered_neighbours(self.h, x.ctypes.data, out.ctypes.data, n)
        return out

    def lsinterp(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        fi = np.zeros(4)
        st = self.E.emu_scattered_lsinterp(self.h, x.ctypes.data, fi.ctypes.data)
        return st, fi


def write_scattered_file(path, n, seed=1, rmin=1.05, rmax=5.0, nspec=4, duplicates=0, inside=0, digits=17):
    """A sample file for modelnum 4 in the format of gcpm_dens_model_buildgrid_random (header, qs, ms, then `pos, log(Ns)` per
    sample): positions drawn with a density falling with radius inside a spherical shell (SM, metres), log-densities of the
    Ngo model there (the product's own srt_model_eval is NOT used: values come from the CPU oracle).  duplicates: that many
    samples repeated verbatim with another value; inside: that many samples inside the Earth."""
    rng = np.random.default_rng(seed)
    r = R_E * (rmin + (rmax - rmin) * rng.random(n) ** 1.7)
    u = rng.normal(size=(n, 3))
    u /= np.linalg.norm(u, axis=1)[:, None]
    pos = r[:, None] * u
    if inside:
        pos[:inside] *= (0.5 * R_E / r[:inside])[:, None]
    om = ol.OracleModel("ngo", NGO_CFG)
    pl = om.plasma(pos)
    om.close()
    logn = np.log(np.maximum(pl[:, 4:4 + nspec], 1e-30))
    if inside:
        logn[:inside] = 0.0
    if duplicates:
        pos = np.vstack([pos, pos[5:5 + duplicates]])
        logn = np.vstack([logn, logn[5:5 + duplicates] + 1.0])
    with open(path, "w") as f:
        f.write("%d %.17e %.17e %.17e %.17e %.17e %.17e\n" % ((nspec,) + tuple(np.ravel([pos.min(0), pos.max(0)], order="F"))))
        f.write(" ".join("%.17e" % v for v in pl[0, 0:nspec]) + "\n")
        f.write(" ".join("%.17e" % v for v in pl[0, 8:8 + nspec]) + "\n")
        fmt = "%%.%de" % digits
        for p, v in zip(pos, logn):
            f.write(" ".join(fmt % t for t in list(p) + list(v)) + "\n")
    return pos, logn


# Rays whose FIRST attempt that passes the error test leaves the resonance cone (launched 1-5 degrees inside the cone with a
# large dt0): raytracer.f95:821 sets the local `w` before the cone test at :891, so the retry already carries the dF/dk
# error term (with `w` set only on acceptance, 57 of 95 such rays differed from the oracle).
# (launch latitude deg, frequency Hz, wave-normal angle from B0 deg, dt0)
CONE_FIRST = [(15.0, 30000.0, 94.05, 0.05), (20.0, 8000.0, 91.3, 0.02), (20.0, 20000.0, 92.925, 0.1), (30.0, 10000.0, 85.7, 0.05),
              (30.0, 10000.0, 86.1, 0.1), (30.0, 15000.0, 91.6, 0.05), (20.0, 30000.0, 86.95, 0.05), (45.0, 30000.0, 87.525, 0.05)]


def cone_first_rays(oracle_model):
    """pos0, dir0, w0, params of the CONE_FIRST rays (direction built from the oracle's field at the launch point)."""
    pos, dirs, ws, dts = [], [], [], []
    for lat, f, ang, dt0 in CONE_FIRST:
        p0 = ol.launch_point(lat, 0.0, 1000.0)
        B = oracle_model.plasma(p0)[0][16:19]
        b = B / np.linalg.norm(B)
        e = np.cross(b, [0.0, 1.0, 0.0])
        e /= np.linalg.norm(e)
        a = np.deg2rad(ang)
        pos.append(p0)
        dirs.append(np.cos(a) * b + np.sin(a) * e)
        ws.append(2 * np.pi * f)
        dts.append(dt0)
    return np.array(pos), np.array(dirs), np.array(ws), dts


def stop_code_rays(oracle_model):
    """Rays that end by each stop condition of raytracer_stopconditions (raytracer.f95:324-353):
    [(expected code, pos0, dir0, w0, params)].  Codes 2, 3, 5 are rays of the config-2 sweep (found with the CPU oracle);
    code 1 is a downgoing ray with minalt raised to R_E + 500 km; 6 and 0 by maxsteps / tmax."""
    pos0, dir0, w0 = c2_grid(100, 100)
    P = dict(C2, maxsteps=4000)
    out = []
    for f, lat in [(5000.0, 40.0), (2000.0, 30.0)]:
        p = ol.launch_point(lat, 0.0, 1000.0)
        out.append((1, p, -p / np.linalg.norm(p), 2 * np.pi * f, dict(P, minalt=R_E + 500e3)))
    out.append((2, pos0[7033], np.zeros(3), w0[7033], P))
    for r in (2067, 2236):
        out.append((3, pos0[r], np.zeros(3), w0[r], P))
    for r in (8931, 7332):
        out.append((5, pos0[r], np.zeros(3), w0[r], P))
    out.append((6, pos0[1001], np.zeros(3), w0[1001], dict(P, maxsteps=150)))
    out.append((0, pos0[1755], np.zeros(3), w0[1755], dict(P, tmax=0.4)))
    return out
