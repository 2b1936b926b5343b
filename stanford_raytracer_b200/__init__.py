"""stanford_raytracer_b200 -- host-side mirror of the reference's raytracer interface over
libsrt.so (hand-written sm_100a kernels behind the C ABI of include/srt.h).

The names follow the reference: `raytracer_run` takes the arguments of
fortran/raytracer.f95:609-642 (for a whole batch of rays instead of one) and returns, per
ray, the arrays it allocates (pos, time, vprel, vgrel, n, B0, qs, ms, Ns, nus, stopcond);
`ngosetup` / `interpsetup` mirror the model set-up of fortran/raytracer_driver.f95:256-369 /
:497-615; `write_output` / `read_inputrays` are the driver's file formats (:1146,
:1197-1217).  There is no CPU path: if libsrt.so is missing or no GPU is present the
calls raise.
"""
import ctypes as C
import os

import numpy as np

from . import rayfile

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.environ.get("SRT_LIB") or os.path.join(_HERE, "libsrt.so")  # SRT_LIB: tuning hook, an alternative build (tools/build_variant.sh)
ROW_DOUBLES = 20
LOG_WORDS = 21
SUMMARY_WORDS = 24
R_E = 6371.2e3

_lib = None


class SrtError(RuntimeError):
    pass


class RunParams(C.Structure):
    _fields_ = [("dt0", C.c_double), ("dtmax", C.c_double), ("maxerr", C.c_double), ("minalt", C.c_double), ("tmax", C.c_double),
                ("maxsteps", C.c_int32), ("root", C.c_int32), ("fixedstep", C.c_int32), ("outputper", C.c_int32)]


def lib():
    """Loads libsrt.so; raises if the CUDA extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(_LIB_PATH):
        raise SrtError("libsrt.so is not built (run `python -m stanford_raytracer_b200.build`); there is no CPU fallback")
    L = C.CDLL(_LIB_PATH)
    vp, i, i64, d = C.c_void_p, C.c_int, C.c_int64, C.c_double
    L.srt_init.argtypes = [i]
    L.srt_strerror.restype = C.c_char_p
    L.srt_strerror.argtypes = [i]
    L.srt_last_cuda_error.restype = C.c_char_p
    L.srt_model_ngo.argtypes = [C.POINTER(vp), C.c_char_p, i, i, i, i]
    L.srt_model_grid.argtypes = [C.POINTER(vp), C.c_char_p, i, i, i, i]
    L.srt_model_scattered.argtypes = [C.POINTER(vp), C.c_char_p, i, i, i, i, d, i, i, i, d]
    L.srt_model_scattered_info.argtypes = [vp, C.POINTER(i64), C.POINTER(d), C.POINTER(d), C.POINTER(i), vp]
    L.srt_model_scattered_samples.restype = i64
    L.srt_model_scattered_samples.argtypes = [vp, vp, i64]
    L.srt_model_free.argtypes = [vp]
    L.srt_model_free.restype = None
    L.srt_model_species.argtypes = [vp, C.POINTER(i), vp, vp, vp]
    L.srt_model_eval.argtypes = [vp, i64, vp, vp]
    L.srt_model_tsyganenko.argtypes = [vp, vp]
    L.srt_model_eval_noalloc.argtypes = [vp, i64, vp, vp]
    L.srt_model_eval_noalloc_max_points.restype = i64
    L.srt_model_eval_noalloc_max_points.argtypes = []
    L.srt_trace_batch.argtypes = [vp, i64, vp, vp, vp, C.POINTER(RunParams), C.POINTER(vp)]
    L.srt_init_devices.argtypes = [vp, i]
    L.srt_set_rows_shm.argtypes = [C.c_char_p]
    L.srt_result_chunk_count.argtypes = [vp]
    L.srt_result_chunk_info.argtypes = [vp, i, C.c_char_p, i, C.POINTER(i64), C.POINTER(vp)]
    L.srt_result_ray_location.argtypes = [vp, i64, vp, vp, vp]
    L.srt_result_devices.argtypes = [vp]
    for name in ("nrays", "total_rows", "accepted_steps", "attempts", "fastforwarded_steps"):
        f = getattr(L, "srt_result_" + name)
        f.restype = i64
        f.argtypes = [vp]
    for name in ("nrows", "nsteps"):
        f = getattr(L, "srt_result_" + name)
        f.restype = i64
        f.argtypes = [vp, i64]
    L.srt_result_stopcond.argtypes = [vp, i64]
    L.srt_result_flags.argtypes = [vp, i64]
    L.srt_result_copy_ray.argtypes = [vp, i64, vp]
    L.srt_result_ray_view.argtypes = [vp, i64, C.POINTER(vp), C.POINTER(i64)]
    L.srt_result_summary.argtypes = [vp, vp, vp, vp, vp]
    L.srt_result_free.argtypes = [vp]
    L.srt_result_free.restype = None
    L.srt_trace_batch_device.argtypes = [vp, i64, vp, C.POINTER(RunParams), vp, i64, vp, vp, vp]
    L.srt_launch_count.restype = i64
    L.srt_kernel_info.argtypes = [vp, i64] + [C.POINTER(i)] * 6
    L.srt_fp64_peak.argtypes = [i, i, C.POINTER(d)]
    L.srt_format_es24.argtypes = [vp, i64, vp]
    L.srt_row_text_bytes.restype = i64
    L.srt_row_text_bytes.argtypes = [i]
    L.srt_format_rows.restype = i64
    L.srt_format_rows.argtypes = [i64, i, vp, i64, d, i, vp, vp, vp, vp]
    L.srt_result_write_text.argtypes = [vp, vp, vp, C.c_char_p, i64, i, i]
    L.srt_read_inputrays.argtypes = [C.c_char_p, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), C.POINTER(i64)]
    L.srt_free.argtypes = [vp]
    L.srt_free.restype = None
    _lib = L
    return L


def _check(rc, what):
    if rc != 0:
        L = lib()
        msg = L.srt_strerror(rc).decode()
        cuda = L.srt_last_cuda_error().decode()
        raise SrtError("%s failed: %s%s" % (what, msg, (" [" + cuda + "]") if cuda and rc in (-5, -1) else ""))


def init(device=-1):
    """Selects the CUDA device of this process (device < 0: the current one)."""
    _check(lib().srt_init(device), "srt_init")


def init_devices(devices=None):
    """Selects the devices raytracer_run spreads a batch over (None: every visible device)."""
    if devices is None:
        _check(lib().srt_init_devices(None, 0), "srt_init_devices")
    else:
        a = np.ascontiguousarray(devices, dtype=np.int32)
        _check(lib().srt_init_devices(_ptr(a), len(a)), "srt_init_devices")
    return lib().srt_device_count()


def set_rows_shm(prefix):
    """Row buffers of the following batches in POSIX shared memory `/<prefix>.<pid>.<n>` (None: pinned memory)."""
    _check(lib().srt_set_rows_shm(prefix.encode() if prefix else None), "srt_set_rows_shm")


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Model:
    """A plasma model living on the device (the reference's funcPlasmaParams + state data)."""

    def __init__(self, handle, kind):
        self._h = handle
        self.kind = kind
        n = C.c_int()
        qs, ms, nus = np.zeros(4), np.zeros(4), np.zeros(4)
        _check(lib().srt_model_species(self._h, C.byref(n), _ptr(qs), _ptr(ms), _ptr(nus)), "srt_model_species")
        self.nspec = n.value
        self.qs, self.ms, self.nus = qs[: n.value], ms[: n.value], nus[: n.value]

    def close(self):
        if self._h:
            lib().srt_model_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def funcPlasmaParams(self, x):
        """qs, Ns, ms, nus, B0 at positions x[n][3] (SM, metres), evaluated on the GPU."""
        x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1, 3)
        out = np.empty((len(x), 19))
        _check(lib().srt_model_eval(self._h, len(x), _ptr(x), _ptr(out)), "srt_model_eval")
        return dict(qs=out[:, 0:4], Ns=out[:, 4:8], ms=out[:, 8:12], nus=out[:, 12:16], B0=out[:, 16:19], raw=out)

    def tsyganenko(self, Pdyn, Dst, ByIMF, BzIMF, W=(0, 0, 0, 0, 0, 0)):
        """use_tsyganenko = 1 with the driver's --tsyganenko_* parameters (TS05 external field, T04_s)."""
        parmod = np.array([Pdyn, Dst, ByIMF, BzIMF] + list(W), dtype=np.float64)
        assert parmod.shape == (10,)
        _check(lib().srt_model_tsyganenko(self._h, _ptr(parmod)), "srt_model_tsyganenko")
        return self

    def kernel_info(self, nrays):
        v = [C.c_int() for _ in range(6)]
        _check(lib().srt_kernel_info(self._h, int(nrays), *[C.byref(t) for t in v]), "srt_kernel_info")
        d = dict(zip(("mode", "regs", "local_bytes", "blocks_per_sm", "grid", "block"), [t.value for t in v]))
        d["mode"] = "team8" if d["mode"] else "thread"
        return d


def ngosetup(ngo_configfile, yearday=0, milliseconds_day=0, use_igrf=0, use_tsyganenko=0):
    """modelnum 1 (driver:256-369)."""
    h = C.c_void_p()
    _check(lib().srt_model_ngo(C.byref(h), os.fsencode(ngo_configfile), int(yearday), int(milliseconds_day), int(use_igrf), int(use_tsyganenko)),
           "srt_model_ngo")
    return Model(h, "ngo")


def interpsetup(interp_interpfile, yearday=0, milliseconds_day=0, use_igrf=0, use_tsyganenko=0):
    """modelnum 3 (driver:497-615)."""
    h = C.c_void_p()
    _check(lib().srt_model_grid(C.byref(h), os.fsencode(interp_interpfile), int(yearday), int(milliseconds_day), int(use_igrf), int(use_tsyganenko)),
           "srt_model_grid")
    return Model(h, "grid")


def scatteredinterpsetup(interp_interpfile, window_scale, order, exact, local_window_scale, scaled=0, yearday=0, milliseconds_day=0, use_igrf=0,
                         use_tsyganenko=0):
    """modelnum 4 (driver:617-770): density by moving least squares over scattered samples; the driver's
    --scattered_interp_window_scale, _order, _exact, _local_window_scale (scaled is 0 in the driver)."""
    h = C.c_void_p()
    _check(lib().srt_model_scattered(C.byref(h), os.fsencode(interp_interpfile), int(yearday), int(milliseconds_day), int(use_igrf),
                                     int(use_tsyganenko), float(window_scale), int(order), int(exact), int(scaled), float(local_window_scale)),
           "srt_model_scattered")
    m = Model(h, "scattered")
    n, mx, rad, nm = C.c_int64(), C.c_double(), C.c_double(), C.c_int()
    cells = (C.c_int * 3)()
    _check(lib().srt_model_scattered_info(h, C.byref(n), C.byref(mx), C.byref(rad), C.byref(nm), cells), "srt_model_scattered_info")
    m.npoints, m.maxnearest, m.radius, m.nmono, m.cells = n.value, mx.value, rad.value, nm.value, tuple(cells)
    return m


def scattered_samples(model):
    """The samples set-up kept, in the order a query walks them: [npoints][8] = x, y, z, log-density of species 1..4,
    distance to the nearest other sample."""
    out = np.empty((model.npoints, 8))
    got = lib().srt_model_scattered_samples(model._h, _ptr(out), out.size)
    if got != model.npoints:
        raise SrtError("srt_model_scattered_samples: %d" % got)
    return out


class RayOutput:
    """What raytracer_run returns for one ray (column-major (3,n) arrays as in the reference)."""

    __slots__ = ("rows", "stopcond", "nsteps", "flags", "qs", "ms", "nus", "w", "outputper")

    def __init__(self, rows, stopcond, nsteps, flags, model, w, outputper):
        self.rows, self.stopcond, self.nsteps, self.flags = rows, stopcond, nsteps, flags
        self.qs, self.ms, self.nus, self.w, self.outputper = model.qs, model.ms, model.nus, w, outputper

    time = property(lambda s: s.rows[:, 0])
    pos = property(lambda s: s.rows[:, 1:4].T)
    vprel = property(lambda s: s.rows[:, 4:7].T)
    vgrel = property(lambda s: s.rows[:, 7:10].T)
    n = property(lambda s: s.rows[:, 10:13].T)
    B0 = property(lambda s: s.rows[:, 13:16].T)
    Ns = property(lambda s: s.rows[:, 16:16 + len(s.qs)].T)


class BatchResult:
    def __init__(self, handle, model, w0, outputper):
        self._h, self.model, self.w0, self.outputper = handle, model, w0, outputper
        L = lib()
        self.nrays = L.srt_result_nrays(handle)
        self.total_rows = L.srt_result_total_rows(handle)
        self.accepted_steps = L.srt_result_accepted_steps(handle)
        self.attempts = L.srt_result_attempts(handle)
        self.fastforwarded_steps = L.srt_result_fastforwarded_steps(handle)
        self.ndevices = L.srt_result_devices(handle)
        self.last_row = np.empty((self.nrays, ROW_DOUBLES))
        self.nsteps = np.empty(self.nrays, dtype=np.int64)
        self.nrejected = np.empty(self.nrays, dtype=np.int64)
        self.stopcond = np.empty(self.nrays, dtype=np.int32)
        _check(L.srt_result_summary(handle, _ptr(self.last_row), _ptr(self.nsteps), _ptr(self.nrejected), _ptr(self.stopcond)), "srt_result_summary")

    def __len__(self):
        return self.nrays

    def rows(self, ray):
        L = lib()
        n = L.srt_result_nrows(self._h, ray)
        if n < 0:
            raise IndexError(ray)
        out = np.empty((n, ROW_DOUBLES))
        _check(L.srt_result_copy_ray(self._h, ray, _ptr(out)), "srt_result_copy_ray")
        return out

    def chunks(self):
        """[(shm_name or "", nrows, address)] of the row buffers, and per ray (chunk, first_row, nrows)."""
        L = lib()
        out = []
        for c in range(L.srt_result_chunk_count(self._h)):
            name = C.create_string_buffer(96)
            n, p = C.c_int64(), C.c_void_p()
            _check(L.srt_result_chunk_info(self._h, c, name, 96, C.byref(n), C.byref(p)), "srt_result_chunk_info")
            out.append((name.value.decode(), n.value, p.value))
        chunk = np.empty(self.nrays, dtype=np.int32)
        first = np.empty(self.nrays, dtype=np.int64)
        nrows = np.empty(self.nrays, dtype=np.int64)
        _check(L.srt_result_ray_location(self._h, self.nrays, _ptr(chunk), _ptr(first), _ptr(nrows)), "srt_result_ray_location")
        return out, chunk, first, nrows

    def __getitem__(self, ray):
        L = lib()
        return RayOutput(self.rows(ray), int(self.stopcond[ray]), int(self.nsteps[ray]), L.srt_result_flags(self._h, ray), self.model,
                         float(self.w0[ray]), self.outputper)

    def write_text(self, path, first_raynum=1, append=False, nthreads=0):
        """The driver's output file for this batch (fortran/raytracer_driver.f95:1197-1217), written by the
        library's multi-threaded formatter."""
        w0 = np.ascontiguousarray(self.w0, dtype=np.float64)
        _check(lib().srt_result_write_text(self._h, self.model._h, _ptr(w0), os.fsencode(path), int(first_raynum), int(bool(append)),
                                           int(nthreads)), "srt_result_write_text")

    def close(self):
        if self._h:
            lib().srt_result_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def format_es24(values):
    """ES24.15E3 text of an array of doubles (24 characters each) by the library's formatter."""
    v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
    out = C.create_string_buffer(24 * len(v))
    _check(lib().srt_format_es24(_ptr(v), len(v), out), "srt_format_es24")
    raw = out.raw
    return [raw[24 * k:24 * k + 24].decode() for k in range(len(v))]


def format_rows(raynum, stopcond, rows, w, qs, ms, nus):
    """The output lines of one ray as bytes (driver:1197-1217)."""
    rows = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, ROW_DOUBLES)
    qs, ms, nus = (np.ascontiguousarray(a, dtype=np.float64) for a in (qs, ms, nus))
    L = lib()
    out = C.create_string_buffer(int(L.srt_row_text_bytes(len(qs))) * len(rows) + 1)
    n = L.srt_format_rows(int(raynum), int(stopcond), _ptr(rows), len(rows), float(w), len(qs), _ptr(qs), _ptr(ms), _ptr(nus), out)
    if n < 0:
        _check(int(n), "srt_format_rows")
    return out.raw[:n]


def read_inputrays(path):
    """inputraysfile -> pos0[n][3], dir0[n][3], w[n] by the library's reader (driver:1146)."""
    L = lib()
    p, dd, w, n = C.c_void_p(), C.c_void_p(), C.c_void_p(), C.c_int64()
    _check(L.srt_read_inputrays(os.fsencode(path), C.byref(p), C.byref(dd), C.byref(w), C.byref(n)), "srt_read_inputrays")
    try:
        k = n.value
        pos0 = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_double)), shape=(max(k, 1) * 3,))[:3 * k].reshape(k, 3).copy()
        dir0 = np.ctypeslib.as_array(C.cast(dd, C.POINTER(C.c_double)), shape=(max(k, 1) * 3,))[:3 * k].reshape(k, 3).copy()
        w0 = np.ctypeslib.as_array(C.cast(w, C.POINTER(C.c_double)), shape=(max(k, 1),))[:k].copy()
    finally:
        L.srt_free(p); L.srt_free(dd); L.srt_free(w)
    return pos0, dir0, w0


def raytracer_run(model, pos0, dir0, w0, dt0, dtmax, maxerr, maxsteps, minalt, root, tmax, fixedstep, outputper=1):
    """raytracer_run (fortran/raytracer.f95:609) for a batch: pos0[n][3], dir0[n][3] (updated in
    place when all-zero, as the reference's intent(inout) dir0), w0[n]."""
    pos0 = np.ascontiguousarray(pos0, dtype=np.float64).reshape(-1, 3)
    if not (isinstance(dir0, np.ndarray) and dir0.dtype == np.float64 and dir0.flags.c_contiguous and dir0.shape == pos0.shape):
        dir0 = np.ascontiguousarray(dir0, dtype=np.float64).reshape(-1, 3).copy()
    w0 = np.ascontiguousarray(w0, dtype=np.float64).reshape(-1)
    if not (len(pos0) == len(dir0) == len(w0)):
        raise ValueError("pos0, dir0, w0 disagree in length")
    p = RunParams(dt0, dtmax, maxerr, minalt, tmax, int(maxsteps), int(root), int(fixedstep), int(outputper))
    h = C.c_void_p()
    _check(lib().srt_trace_batch(model._h, len(w0), _ptr(pos0), _ptr(dir0), _ptr(w0), C.byref(p), C.byref(h)), "srt_trace_batch")
    res = BatchResult(h, model, w0, max(1, int(outputper)))
    res.dir0 = dir0
    return res


def trace_batch_device(model, nrays, d_rays_ptr, params, d_log_ptr, log_capacity, d_summary_ptr, d_counters_ptr, stream_ptr=0):
    """srt_trace_batch_device with raw device pointers (bench.py passes torch data_ptr()s)."""
    _check(lib().srt_trace_batch_device(model._h, int(nrays), C.c_void_p(d_rays_ptr), C.byref(params), C.c_void_p(d_log_ptr), int(log_capacity),
                                        C.c_void_p(d_summary_ptr), C.c_void_p(d_counters_ptr), C.c_void_p(stream_ptr)), "srt_trace_batch_device")


def fp64_peak(fma=True, iters=20000):
    g = C.c_double()
    _check(lib().srt_fp64_peak(1 if fma else 0, iters, C.byref(g)), "srt_fp64_peak")
    return g.value


def launch_count():
    return lib().srt_launch_count()


write_output = rayfile.write_output  # pure-Python writer (tests compare it with BatchResult.write_text)
