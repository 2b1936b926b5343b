"""`dumpmodel` (fortran/dumpmodel.f95): the plasma parameters of a model on a regular lattice, written in the
reference's dump format.  The lattice is evaluated by the product's funcPlasmaParams kernel (`srt_model_eval`,
one launch for all nodes) instead of the reference's triple loop over the adapter callback
(dumpmodel.f95:374-378 Ngo, :649-653 interpolated).

    python -m stanford_raytracer_b200.dumpmodel --modelnum 1 --ngo_configfile newray.in --yearday 2010001 \\
        --milliseconds_day 0 --use_igrf 0 --use_tsyganenko 0 --minx .. --maxx .. --miny .. --maxy .. --minz .. --maxz .. \\
        --nx 64 --ny 1 --nz 64 --filename dump.txt

Format (dumpmodel.f95:1284-1289):
    (5i10)          nspec nx ny nz
    (6es24.15e3)    minx maxx miny maxy minz maxz
    (es24.15e3)     one value per line of f(4*nspec+3, nx, ny, nz), first index fastest:
                    qs(nspec) Ns(nspec) ms(nspec) nus(nspec) B0(3) per node, x fastest, then y, then z.
Coordinates are SM metres (models 1 and 3).  No CPU path: the evaluation needs the CUDA library and a GPU.
"""
import argparse
import sys

import numpy as np

NSPEC = 4


def lattice_axis(n, lo, hi):
    """`(/ (ind, ind=0,n-1) /)*del + min` with del = (max-min)/(n-1.0); a single node sits at min
    (dumpmodel.f95:219-238)."""
    n = int(n)
    if n == 1:
        return np.array([float(lo)])
    d = (float(hi) - float(lo)) / (n - 1.0)
    return np.arange(n) * d + float(lo)


def lattice_points(nx, ny, nz, extent):
    """Node positions in the dump's order (x fastest, z slowest): [nz*ny*nx][3]."""
    xs, ys, zs = lattice_axis(nx, extent[0], extent[1]), lattice_axis(ny, extent[2], extent[3]), lattice_axis(nz, extent[4], extent[5])
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    return np.ascontiguousarray(np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1))


def dump(model, nx, ny, nz, extent, chunk=1 << 22):
    """f[nz][ny][nx][19] = (qs4, Ns4, ms4, nus4, B03) of `model` (a stanford_raytracer_b200 model handle) at the
    lattice nodes."""
    P = lattice_points(nx, ny, nz, extent)
    out = np.empty((len(P), 4 * NSPEC + 3))
    for i in range(0, len(P), chunk):
        out[i:i + chunk] = model.funcPlasmaParams(P[i:i + chunk])["raw"]
    n = int(getattr(model, "nspec", NSPEC))
    if n < NSPEC:
        # the reference allocates f(nspec*4+3, nx, ny, nz) with the model's own nspec (dumpmodel.f95:1284-1289): a gridded
        # file with fewer than four species dumps nspec columns per quantity, not four with phantom zeros
        cols = [c + q * NSPEC for q in range(4) for c in range(n)] + [4 * NSPEC, 4 * NSPEC + 1, 4 * NSPEC + 2]
        out = out[:, cols]
    return out.reshape(int(nz), int(ny), int(nx), 4 * n + 3)


def _es24_block(values):
    """One ES24.15E3 value per line, by the library's exact formatter (csrc/srt_io.cpp)."""
    import ctypes as C
    from . import lib
    v = np.ascontiguousarray(values, dtype=np.float64).reshape(-1)
    buf = C.create_string_buffer(24 * len(v))
    rc = lib().srt_format_es24(v.ctypes.data_as(C.c_void_p), len(v), buf)
    if rc != 0:
        raise RuntimeError("srt_format_es24 failed: %d" % rc)
    a = np.frombuffer(buf.raw, dtype=np.uint8).reshape(len(v), 24)
    lines = np.empty((len(v), 25), dtype=np.uint8)
    lines[:, :24] = a
    lines[:, 24] = 10
    return lines.tobytes()


def write_dump(path, f, extent):
    """f[nz][ny][nx][4*nspec+3] -> the reference's dump file."""
    nz, ny, nx, m = f.shape
    nspec = (m - 3) // 4
    from .rayfile import es24
    with open(path, "wb") as fh:
        fh.write(("%10d%10d%10d%10d\n" % (nspec, nx, ny, nz)).encode())
        fh.write(("".join(es24(v) for v in extent) + "\n").encode())
        flat = np.ascontiguousarray(f, dtype=np.float64).ravel()
        for i in range(0, len(flat), 1 << 20):
            fh.write(_es24_block(flat[i:i + (1 << 20)]))


def read_dump(path):
    """Returns dict(nspec, nx, ny, nz, extent, f[nz][ny][nx][4*nspec+3]) (the reader of
    matlab/readdump.m's role)."""
    with open(path) as fh:
        toks = fh.read().replace("D", "E").replace("d", "e").split()
    nspec, nx, ny, nz = (int(t) for t in toks[:4])
    v = np.array(toks[4:], dtype=np.float64)
    m = 4 * nspec + 3
    if len(v) != 6 + m * nx * ny * nz:
        raise ValueError("dump file %s: %d values, expected %d" % (path, len(v), 6 + m * nx * ny * nz))
    return dict(nspec=nspec, nx=nx, ny=ny, nz=nz, extent=v[:6], f=v[6:].reshape(nz, ny, nx, m))


def main(argv=None):
    ap = argparse.ArgumentParser(prog="dumpmodel", description="plasma parameters of a model on a lattice (reference: fortran/dumpmodel.f95)")
    for k in ("minx", "maxx", "miny", "maxy", "minz", "maxz"):
        ap.add_argument("--" + k, type=float, default=0.0)
    for k in ("nx", "ny", "nz"):
        ap.add_argument("--" + k, type=int, default=0)
    ap.add_argument("--filename", default="")
    ap.add_argument("--modelnum", type=int, default=0)
    ap.add_argument("--ngo_configfile", default="")
    ap.add_argument("--interp_interpfile", default="")
    ap.add_argument("--scattered_interp_window_scale", type=float, default=0.0)  # modelnum 4 (dumpmodel.f95:152-160)
    ap.add_argument("--scattered_interp_order", type=float, default=0.0)
    ap.add_argument("--scattered_interp_exact", type=float, default=0.0)
    ap.add_argument("--scattered_interp_local_window_scale", type=float, default=0.0)
    for k in ("yearday", "milliseconds_day", "use_tsyganenko", "use_igrf"):
        ap.add_argument("--" + k, type=int, default=0)  # unset options are 0 in the reference build (-finit-local-zero)
    TSY = ("Pdyn", "Dst", "ByIMF", "BzIMF", "W1", "W2", "W3", "W4", "W5", "W6")
    for k in TSY:
        ap.add_argument("--tsyganenko_" + k, type=float, default=0.0)
    a, _ = ap.parse_known_args(argv)  # the GCPM options are accepted and unused (out of scope)
    if min(a.nx, a.ny, a.nz) < 1 or not a.filename:
        ap.error("--nx --ny --nz (>= 1) and --filename are required")
    import stanford_raytracer_b200 as srt
    srt.init(0)
    if a.modelnum == 1:
        model = srt.ngosetup(a.ngo_configfile, a.yearday, a.milliseconds_day, a.use_igrf, a.use_tsyganenko)
    elif a.modelnum == 3:
        model = srt.interpsetup(a.interp_interpfile, a.yearday, a.milliseconds_day, a.use_igrf, a.use_tsyganenko)
    elif a.modelnum == 4:
        import math
        model = srt.scatteredinterpsetup(a.interp_interpfile, a.scattered_interp_window_scale, math.floor(a.scattered_interp_order),
                                         math.floor(a.scattered_interp_exact), a.scattered_interp_local_window_scale, 0, a.yearday,
                                         a.milliseconds_day, a.use_igrf, a.use_tsyganenko)
    else:
        print("dumpmodel: modelnum %d is not supported (1 = Ngo, 3 = gridded interpolation, 4 = scattered samples)" % a.modelnum, file=sys.stderr)
        return 2
    if a.use_tsyganenko == 1:
        pm = [getattr(a, "tsyganenko_" + k) for k in TSY]
        model.tsyganenko(*pm[:4], W=pm[4:])
    extent = (a.minx, a.maxx, a.miny, a.maxy, a.minz, a.maxz)
    write_dump(a.filename, dump(model, a.nx, a.ny, a.nz, extent), extent)
    return 0


if __name__ == "__main__":
    sys.exit(main())
