"""The reference driver's two file formats (host side, no GPU involved).

* inputraysfile (fortran/raytracer_driver.f95:1146): list-directed `pos0(3) dir0(3) w`,
  one ray per READ; reading stops at the first record that cannot be parsed.
* output file (fortran/raytracer_driver.f95:1197-1217): per recorded row
  `(i10, i10, 17es24.15e3, i10)` = raynum, stopcond, time, pos, vprel, vgrel, n, B0, w,
  nspec, then qs, ms, Ns, nus each as nspec * `es24.15e3`, then a newline.
"""
import math
import re

import numpy as np

_SEP = re.compile(r"[,\s]+")


def _fortran_float(tok):
    t = tok.strip().lower().replace("d", "e")
    # Fortran accepts 1.0+5 as 1.0e+5
    m = re.match(r"^([+-]?(\d+\.?\d*|\.\d+))([+-]\d+)$", t)
    if m:
        t = m.group(1) + "e" + m.group(3)
    return float(t)


def read_inputrays(path):
    """Returns pos0[n][3], dir0[n][3], w[n].  A READ takes seven items, continuing over records if it has to, and
    drops what is left of the record it ends in; the first item that is not a number ends the file (iostat /= 0).
    Pure-Python statement of the library's srt_read_inputrays (csrc/srt_io.cpp)."""
    vals = []
    toks = []
    stop = False
    with open(path) as f:
        for line in f:
            for t in _SEP.split(line.strip()):
                if not t:
                    continue
                toks.append(t)
                if len(toks) == 7:
                    try:
                        vals.append([_fortran_float(x) for x in toks])
                    except ValueError:
                        stop = True
                    toks = []
                    break  # rest of the record is dropped
            if stop:
                break
    a = np.array(vals, dtype=np.float64).reshape(-1, 7)
    return a[:, 0:3].copy(), a[:, 3:6].copy(), a[:, 6].copy()


def es24(x):
    """Fortran ES24.15E3 edit descriptor."""
    x = float(x)
    if math.isnan(x):
        return "NaN".rjust(24)
    if math.isinf(x):
        return ("Infinity" if x > 0 else "-Infinity").rjust(24)
    mant, exp = ("%.15E" % x).split("E")
    e = int(exp)
    if abs(e) > 999:
        return "*" * 24
    return ("%sE%s%03d" % (mant, "-" if e < 0 else "+", abs(e))).rjust(24)


def format_row(raynum, stopcond, row, w, qs, ms, nus):
    """One output line for a 20-double row (t, pos, vprel, vgrel, n, B0, Ns)."""
    nspec = len(qs)
    parts = ["%10d%10d" % (raynum, stopcond)]
    parts.extend(es24(v) for v in row[0:16])
    parts.append(es24(w))
    parts.append("%10d" % nspec)
    parts.extend(es24(v) for v in qs)
    parts.extend(es24(v) for v in ms)
    parts.extend(es24(v) for v in row[16:16 + nspec])
    parts.extend(es24(v) for v in nus)
    return "".join(parts) + "\n"


def write_output(path, batch, w=None):
    """Writes a BatchResult exactly as the driver's loop :1197-1217 would (rows are already
    decimated by outputper on the device).  raynum is 1-based."""
    with open(path, "w") as f:
        for r in range(len(batch)):
            rows = batch.rows(r)
            wr = float(batch.w0[r]) if w is None else float(w[r])
            sc = int(batch.stopcond[r])
            for row in rows:
                f.write(format_row(r + 1, sc, row, wr, batch.model.qs, batch.model.ms, batch.model.nus))


def read_output(path):
    """Reads an output file back (the consumer side of the driver's format: damping/src/wipp_fileutils.cpp:6-170,
    damping/matlab/readrayoutput.m:24-47).  Returns {raynum: dict} in file order, each with w, stopcond, nspec (taken
    from the ray's first line), qs[nspec], ms[nspec] (first line) and per recorded row time[n], pos/vprel/vgrel/n/B0
    [n][3], Ns/nus[n][nspec].  The fields are fixed-width (i10, es24.15e3), so the line is cut by column and a value
    that touches its neighbour (a four-digit exponent cannot occur, a 24-character negative number can) still parses."""
    rays = {}
    acc = {}
    with open(path) as f:
        for line in f:
            line = line.rstrip("\n")
            if not line.strip():
                continue
            raynum, stopcond = int(line[0:10]), int(line[10:20])
            v = [_fortran_float(line[20 + 24 * i:44 + 24 * i]) for i in range(17)]
            nspec = int(line[428:438])
            tail = [_fortran_float(line[438 + 24 * i:462 + 24 * i]) for i in range(4 * nspec)]
            if raynum not in rays:
                rays[raynum] = dict(w=v[16], stopcond=stopcond, nspec=nspec, qs=np.array(tail[0:nspec]), ms=np.array(tail[nspec:2 * nspec]))
                acc[raynum] = []
            acc[raynum].append(v[0:16] + tail[2 * nspec:4 * nspec])
    for raynum, rows in acc.items():
        a = np.array(rows, dtype=np.float64)
        ns = rays[raynum]["nspec"]
        rays[raynum].update(time=a[:, 0], pos=a[:, 1:4], vprel=a[:, 4:7], vgrel=a[:, 7:10], n=a[:, 10:13], B0=a[:, 13:16],
                            Ns=a[:, 16:16 + ns], nus=a[:, 16 + ns:16 + 2 * ns])
    return rays
