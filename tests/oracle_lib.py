"""ctypes wrapper over the CPU oracle (oracle/_build/libsrt_oracle*.so).

Test infrastructure only: imported by tests/, bench.py's cpu_baseline / --impl reference
legs and __graft_entry__.smoke().  Never imported by the product package.
"""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORC_DIR = os.path.join(ROOT, "oracle")
ROW = 20
R_E = 6371.2e3


def build():
    subprocess.check_call(["make", "-s", "-C", ORC_DIR])


_libs = {}


def lib(libm=False):
    name = "libsrt_oracle_libm.so" if libm else "libsrt_oracle.so"
    if name in _libs:
        return _libs[name]
    path = os.path.join(ORC_DIR, "_build", name)
    if not os.path.exists(path):
        build()
    L = C.CDLL(path)
    vp, d, i, l = C.c_void_p, C.c_double, C.c_int, C.c_long
    L.orc_model_ngo.restype = vp
    L.orc_model_ngo.argtypes = [C.c_char_p, i, i, i, i]
    if hasattr(L, "orc_model_grid"):
        L.orc_model_grid.restype = vp
        L.orc_model_grid.argtypes = [C.c_char_p, i, i, i, i]
    L.orc_model_free.argtypes = [vp]
    L.orc_model_tsyganenko.argtypes = [vp, vp, vp]
    L.orc_model_psi.restype = d
    L.orc_model_psi.argtypes = [vp]
    L.orc_model_del.restype = d
    L.orc_model_del.argtypes = [vp]
    L.orc_plasma.argtypes = [vp, l, vp, vp]
    L.orc_dispersion_relation.restype = d
    L.orc_dispersion_relation.argtypes = [vp, vp, d, vp]
    L.orc_stix.argtypes = [vp, d, vp, vp]
    L.orc_evalrhs.argtypes = [vp, vp, vp]
    L.orc_solve.argtypes = [vp, vp, d, vp, vp]
    L.orc_is_right_handed.restype = i
    L.orc_is_right_handed.argtypes = [d, d, d, d, d]
    L.orc_form_dispersion_matrix.argtypes = [d, d, d, d, d, vp]
    L.orc_trace.restype = vp
    L.orc_trace.argtypes = [vp, vp, vp, d, d, d, d, i, d, i, d, i]
    for f in ("nrows", "nrejected", "ncone", "plasma_calls"):
        getattr(L, "orc_trace_" + f).restype = l
        getattr(L, "orc_trace_" + f).argtypes = [vp]
    L.orc_trace_stopcond.restype = i
    L.orc_trace_stopcond.argtypes = [vp]
    L.orc_trace_fixed_cone_exit.restype = i
    L.orc_trace_fixed_cone_exit.argtypes = [vp]
    L.orc_trace_dir0.argtypes = [vp, vp]
    L.orc_trace_rows.argtypes = [vp, vp]
    L.orc_trace_species.argtypes = [vp, vp]
    L.orc_trace_free.argtypes = [vp]
    L.orc_trace_batch_summary.restype = l
    L.orc_trace_batch_summary.argtypes = [vp, l, vp, vp, vp, d, d, d, i, d, i, d, i, vp, vp, vp, vp]
    _libs[name] = L
    return L


def ts05_ref(srt_math=True):
    """oracle/_ref/libts05_ref*.so: the reference's TS05 source, translated at build time (oracle/f77_to_cpp.py).
    srt_math: the build on the elementary functions of csrc/srt_math.h (bit-for-bit yardstick); else glibc's."""
    name = "libts05_ref_srtmath.so" if srt_math else "libts05_ref.so"
    if name not in _libs:
        path = os.path.join(ORC_DIR, "_ref", name)
        if not os.path.exists(path):
            subprocess.check_call(["make", "-s", "-C", ORC_DIR, "ref"])
        if not os.path.exists(path):
            return None
        R = C.CDLL(path)
        d, f = C.c_double, C.c_float
        R.ts05_ref_extern.argtypes = [C.c_void_p, d, d, d, d, C.c_void_p]
        R.ts05_ref_t04_s.argtypes = [C.c_void_p, f, f, f, f, C.c_void_p]
        _libs[name] = R
    return _libs[name]


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class OracleModel:
    def __init__(self, kind, path, yearday=2010001, msec=0, use_igrf=0, libm=False, faithful_cost=False):
        self.L = lib(libm)
        fn = self.L.orc_model_ngo if kind == "ngo" else self.L.orc_model_grid
        self.h = fn(path.encode(), yearday, msec, use_igrf, int(faithful_cost))
        if not self.h:
            raise RuntimeError("oracle model setup failed: %s" % path)
        self.del_ = self.L.orc_model_del(self.h)

    def tsyganenko(self, parmod, srt_math=True):
        """use_tsyganenko = 1: T04_s is the reference's own source as translated into oracle/_ref (ts05_ref())."""
        R = ts05_ref(srt_math)
        pm = np.ascontiguousarray(parmod, dtype=np.float64)
        assert pm.shape == (10,)
        self.L.orc_model_tsyganenko(self.h, _p(pm), C.cast(R.ts05_ref_t04_s, C.c_void_p))
        return self

    def close(self):
        if self.h:
            self.L.orc_model_free(self.h)
            self.h = None

    def plasma(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 3)
        out = np.empty((len(x), 19))
        self.L.orc_plasma(self.h, len(x), _p(x), _p(out))
        return out

    def dispersion_relation(self, n, w, x):
        n = np.ascontiguousarray(n, dtype=np.float64)
        x = np.ascontiguousarray(x, dtype=np.float64)
        return self.L.orc_dispersion_relation(self.h, _p(n), w, _p(x))

    def stix(self, w, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        o = np.empty(5)
        self.L.orc_stix(self.h, w, _p(x), _p(o))
        return dict(zip("SDPRL", o))

    def evalrhs(self, args7):
        a = np.ascontiguousarray(args7, dtype=np.float64)
        o = np.empty(7)
        self.L.orc_evalrhs(self.h, _p(a), _p(o))
        return o

    def solve(self, k, w, x):
        k = np.ascontiguousarray(k, dtype=np.float64)
        x = np.ascontiguousarray(x, dtype=np.float64)
        o = np.empty(4)
        self.L.orc_solve(self.h, _p(k), w, _p(x), _p(o))
        return complex(o[0], o[1]), complex(o[2], o[3])

    def trace(self, pos0, dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep):
        pos0 = np.ascontiguousarray(pos0, dtype=np.float64)
        dir0 = np.ascontiguousarray(dir0, dtype=np.float64)
        t = self.L.orc_trace(self.h, _p(pos0), _p(dir0), w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep)
        n = self.L.orc_trace_nrows(t)
        rows = np.empty((n, ROW))
        self.L.orc_trace_rows(t, _p(rows))
        d = np.empty(3)
        self.L.orc_trace_dir0(t, _p(d))
        sp = np.empty(12)
        self.L.orc_trace_species(t, _p(sp))
        res = dict(rows=rows, stopcond=self.L.orc_trace_stopcond(t), nrejected=self.L.orc_trace_nrejected(t),
                   ncone=self.L.orc_trace_ncone(t), plasma_calls=self.L.orc_trace_plasma_calls(t),
                   fixed_cone_exit=self.L.orc_trace_fixed_cone_exit(t), dir0=d, qs=sp[0:4], ms=sp[4:8], nus=sp[8:12])
        self.L.orc_trace_free(t)
        return res

    def trace_summary(self, pos0, dir0, w0, **params):
        """endpoint of one ray (last row, rows, stop code)"""
        s = self.trace_batch_summary(np.asarray(pos0).reshape(1, 3), np.asarray(dir0).reshape(1, 3), np.array([w0]), **params)
        return dict(last_row=s["last_row"][0], nrows=int(s["nrows"][0]), stopcond=int(s["stopcond"][0]), nrejected=int(s["nrejected"][0]))

    def trace_batch_summary(self, pos0, dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep):
        pos0 = np.ascontiguousarray(pos0, dtype=np.float64).reshape(-1, 3)
        dir0 = np.ascontiguousarray(dir0, dtype=np.float64).reshape(-1, 3)
        w0 = np.ascontiguousarray(w0, dtype=np.float64)
        n = len(w0)
        last = np.empty((n, ROW))
        nrows = np.empty(n, dtype=np.int64)
        stop = np.empty(n, dtype=np.int32)
        nrej = np.empty(n, dtype=np.int64)
        acc = self.L.orc_trace_batch_summary(self.h, n, _p(pos0), _p(dir0), _p(w0), dt0, dtmax, maxerr, maxsteps, minalt, root,
                                             tmax, fixedstep, _p(last), _p(nrows), _p(stop), _p(nrej))
        return dict(accepted=acc, last_row=last, nrows=nrows, stopcond=stop, nrejected=nrej)


class OracleScattered(OracleModel):
    """modelnum 4 (oracle/orc_scattered.h): kd-tree + lsinterp as the reference states them.  perm = the insertion order the
    reference leaves to an unseeded random_number (None: file order); root_quirk: the root sample's nearest distance 0."""

    def __init__(self, path, window_scale, order, exact, local_window_scale, scaled=0, yearday=2010001, msec=0, use_igrf=0, perm=None,
                 root_quirk=False, libm=False):
        self.L = lib(libm)
        L, d, i, l, vp = self.L, C.c_double, C.c_int, C.c_long, C.c_void_p
        L.orc_model_scattered.restype = vp
        L.orc_model_scattered.argtypes = [C.c_char_p, i, i, i, d, i, i, i, d, vp, l, i]
        L.orc_scattered_info.argtypes = [vp, vp]
        L.orc_scattered_search.restype = l
        L.orc_scattered_search.argtypes = [vp, vp, vp, l]
        L.orc_scattered_impose_order.restype = l
        L.orc_scattered_impose_order.argtypes = [vp, l, vp]
        L.orc_scattered_interpolate.restype = i
        L.orc_scattered_interpolate.argtypes = [vp, vp, vp]
        L.orc_lsinterp_found.restype = i
        L.orc_lsinterp_found.argtypes = [vp, l, vp, vp, i, d, i, i, i, d, vp]
        pm = None if perm is None else np.ascontiguousarray(perm, dtype=np.int64)
        self.h = L.orc_model_scattered(path.encode(), yearday, msec, use_igrf, window_scale, order, exact, scaled, local_window_scale, _p(pm),
                                       0 if pm is None else len(pm), int(root_quirk))
        if not self.h:
            raise RuntimeError("oracle scattered model setup failed: %s" % path)
        self.del_ = L.orc_model_del(self.h)
        self.opts = (order, exact, scaled, local_window_scale)
        info = np.empty(4)
        L.orc_scattered_info(self.h, _p(info))
        self.maxnearest, self.nkept, self.nduplicates, self.root_nearest = info[0], int(info[1]), int(info[2]), info[3]
        self.radius = self.maxnearest * window_scale

    def search(self, x, nspec):
        """neighbours of x in the tree's traversal order: [I][3 + nspec + 1]"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        n = self.L.orc_scattered_search(self.h, _p(x), None, 0)
        out = np.empty((n, 3 + nspec + 1))
        self.L.orc_scattered_search(self.h, _p(x), _p(out), n)
        return out

    def impose_order(self, pts):
        """from now on a query returns its neighbours in the order of pts [n][3] (the one thing the reference's source leaves
        open); every sample must be among them"""
        pts = np.ascontiguousarray(pts, dtype=np.float64)
        missing = self.L.orc_scattered_impose_order(self.h, len(pts), _p(pts))
        assert missing == 0 and len(pts) == self.nkept
        return self

    def interpolate(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        fi = np.zeros(4)
        st = self.L.orc_scattered_interpolate(self.h, _p(x), _p(fi))
        return st, fi

    def lsinterp_found(self, x, pts, vals):
        """lsinterp on neighbour arrays in the caller's order (vals: [I][nspec + 1], last = nearest distance)"""
        x = np.ascontiguousarray(x, dtype=np.float64)
        pts = np.ascontiguousarray(pts, dtype=np.float64)
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        fi = np.zeros(4)
        order, exact, scaled, lws = self.opts
        st = self.L.orc_lsinterp_found(_p(x), len(pts), _p(pts), _p(vals), vals.shape[1], self.radius, order, exact, scaled, lws, _p(fi))
        return st, fi


class DedupModel:
    """oracle/_build/libsrt_dedup.so: the GPU kernels' evaluation order on the host at -O3 (bench.py cpu_baseline,
    "deduplicated" mode).  Read-only during a trace: one instance serves every worker thread."""

    def __init__(self, kind, path, yearday=2010001, msec=0, use_igrf=0):
        p = os.path.join(ORC_DIR, "_build", "libsrt_dedup.so")
        if not os.path.exists(p):
            build()
        E = C.CDLL(p)
        d, i, vp, l = C.c_double, C.c_int, C.c_void_p, C.c_long
        for f in (E.emu_model_ngo, E.emu_model_grid):
            f.restype = vp
            f.argtypes = [C.c_char_p, i, i, i]
        E.emu_trace_summary.restype = l
        E.emu_trace_summary.argtypes = [vp, vp, vp, d, d, d, d, i, d, i, d, i, vp, vp, vp]
        self.E = E
        fn = E.emu_model_ngo if kind == "ngo" else E.emu_model_grid
        self.h = fn(path.encode(), yearday, msec, use_igrf | (2 if kind != "ngo" else 0))  # grid: per-cell coefficient table, as on the GPU
        if not self.h:
            raise RuntimeError("dedup model setup failed: %s" % path)

    def trace_summary(self, pos0, dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep):
        pos0 = np.ascontiguousarray(pos0, dtype=np.float64)
        d0 = np.array(dir0, dtype=np.float64)
        last = np.empty(ROW)
        sc, nr = C.c_int(), C.c_long()
        n = self.E.emu_trace_summary(self.h, _p(pos0), _p(d0), w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep, _p(last),
                                     C.byref(sc), C.byref(nr))
        return dict(last_row=last, nrows=int(n), stopcond=sc.value, nrejected=nr.value)


def launch_point(lat_deg, lon_deg=0.0, alt_km=1000.0):
    """SM position of a launch point (radius R_E + alt)."""
    r = R_E + alt_km * 1e3
    la, lo = np.deg2rad(lat_deg), np.deg2rad(lon_deg)
    return np.array([r * np.cos(la) * np.cos(lo), r * np.cos(la) * np.sin(lo), r * np.sin(la)])
