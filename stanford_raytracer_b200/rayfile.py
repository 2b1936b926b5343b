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
