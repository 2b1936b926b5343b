"""The gridded-density file of modelnum 3 (host side): the format written by
fortran/gcpm_dens_model_buildgrid.f95:302-327 and read by
fortran/interp_dens_model_adapter.f95:58-116.

    line 1  (5i10)          compder nspec nx ny nz
    line 2  (6es24.15e3)    minx maxx miny maxy minz maxz     [m, SM]
    line 3  (nspec es24)    qs
    line 4  (nspec es24)    ms
    then one (es24.15e3) value per line: f(s,i,j,k) = ln(Ns [m^-3]), s fastest, then x, y, z;
    iff compder == 1 the seven derivative arrays follow in the same order.

`sample_grid` fills such a grid from any plasma-parameter function (on the GPU box: the
product's own funcPlasmaParams kernel, the reference's gcpm_dens_model_buildgrid role with the
Ngo model standing in for GCPM, which cannot be built here).
"""
import numpy as np

R_E = 6371.2e3


def _es24_lines(values):
    """ES24.15E3 of every value, one per line."""
    from .rayfile import es24 as f
    return "\n".join(f(v) for v in values)


def es24(v):
    from .rayfile import es24 as f
    return f(v)


def node_axes(n, lo, hi):
    """Node coordinates as the reader regenerates them (interp_dens_model_adapter.f95:90-95)."""
    d = (hi - lo) / (n - 1.0)
    return np.arange(n) * d + lo


def sample_grid(plasma_fn, nx, ny, nz, extent, min_radius=R_E):
    """ln(Ns) at the nodes of an nx x ny x nz lattice over extent = (minx, maxx, miny, maxy, minz, maxz).
    plasma_fn(x[n][3]) -> raw[n][19] (qs4, Ns4, ms4, nus4, B03).  Nodes inside `min_radius` (inside the
    Earth, where the analytic models are meaningless) take the value of the point radially above them.
    Returns f[nz][ny][nx][4], qs[4], ms[4]."""
    xs, ys, zs = node_axes(nx, extent[0], extent[1]), node_axes(ny, extent[2], extent[3]), node_axes(nz, extent[4], extent[5])
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    P = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    r = np.sqrt((P * P).sum(axis=1))
    inside = r < min_radius
    Pq = P.copy()
    zero = r == 0.0
    Pq[zero] = [min_radius, 0.0, 0.0]
    sel = inside & ~zero
    Pq[sel] = P[sel] * (min_radius / r[sel])[:, None]
    # the analytic (Ngo) model is singular on the dipole axis (L = r / sin^2(colat)): nudge axis nodes off it
    axis = (Pq[:, 0] == 0.0) & (Pq[:, 1] == 0.0)
    Pq[axis, 0] = 1e-3 * np.abs(Pq[axis, 2])
    raw = plasma_fn(np.ascontiguousarray(Pq))
    Ns = raw[:, 4:8]
    f = np.log(Ns).reshape(nz, ny, nx, 4)
    if not np.isfinite(f).all():
        raise ValueError("sample_grid: %d non-finite log-densities" % int((~np.isfinite(f)).sum()))
    return f, raw[0, 0:4].copy(), raw[0, 8:12].copy()


def write_grid(path, f, qs, ms, extent, derivs=None):
    """f[nz][ny][nx][nspec]; derivs = None (compder 0) or a list of seven arrays of the same shape."""
    nz, ny, nx, nspec = f.shape
    with open(path, "w") as fh:
        fh.write("%10d%10d%10d%10d%10d\n" % (1 if derivs is not None else 0, nspec, nx, ny, nz))
        fh.write("".join(es24(v) for v in extent) + "\n")
        fh.write("".join(es24(v) for v in qs[:nspec]) + "\n")
        fh.write("".join(es24(v) for v in ms[:nspec]) + "\n")
        for arr in [f] + (list(derivs) if derivs is not None else []):
            flat = np.ascontiguousarray(arr, dtype=np.float64).ravel()
            for i in range(0, len(flat), 1 << 18):
                fh.write(_es24_lines(flat[i:i + (1 << 18)]) + "\n")


def read_grid(path):
    """Returns dict(compder, nspec, nx, ny, nz, extent, qs, ms, f[, derivs])."""
    with open(path) as fh:
        toks = fh.read().replace("D", "E").replace("d", "e").split()
    compder, nspec, nx, ny, nz = (int(float(t)) for t in toks[:5])
    v = np.array(toks[5:], dtype=np.float64)
    extent = v[:6]
    qs, ms = v[6:6 + nspec], v[6 + nspec:6 + 2 * nspec]
    n = nspec * nx * ny * nz
    off = 6 + 2 * nspec
    out = dict(compder=compder, nspec=nspec, nx=nx, ny=ny, nz=nz, extent=extent, qs=qs, ms=ms, f=v[off:off + n].reshape(nz, ny, nx, nspec))
    if compder == 1:
        out["derivs"] = [v[off + (q + 1) * n:off + (q + 2) * n].reshape(nz, ny, nx, nspec) for q in range(7)]
    return out
