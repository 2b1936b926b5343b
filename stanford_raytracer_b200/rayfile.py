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

_NUMBER = re.compile(r"^([+-]?)(\d+\.?\d*|\.\d+)(?:[EDQedq]([+-]?\d+)|([+-]\d+))?$")
_INFNAN = re.compile(r"^([+-]?)(inf|infinity|nan|nan\([^)]*\))$", re.IGNORECASE)
_BLANK = " \t\r\n"
_END = _BLANK + ",;/"


def _fortran_float(tok):
    """One value of a list-directed READ into a REAL item, as gfortran's run-time library accepts it: digits with an optional
    point, an exponent E|D|Q [sign] digits or just sign digits (1.5+3), or INF / INFINITY / NAN / NAN(...) in any case."""
    tok = tok.strip(" ")  # (fixed-width fields of read_output come with their padding)
    m = _NUMBER.match(tok)
    if m:
        exp = m.group(3) or m.group(4)
        return float(m.group(1) + m.group(2) + ("e" + exp if exp else ""))
    m = _INFNAN.match(tok)
    if not m:
        raise ValueError(tok)
    if m.group(2).lower().startswith("nan"):
        return float("nan")
    return float("-inf") if m.group(1) == "-" else float("inf")


def read_inputrays(path):
    """Returns pos0[n][3], dir0[n][3], w[n]: the driver's loop `read(infile,*,iostat=status) pos0, dir0, w` under gfortran's
    list-directed rules.  Values are separated by blanks (tab, CR, end of record), one comma or semicolon, or a slash, which ends
    the READ; nothing between two separators is a null value and leaves the item as the previous READ left it (zeros at first);
    r*c is r copies, r* is r nulls; a READ takes records until its seven items are served and drops the rest of the record it ends
    in; the first token that is not a number, or the end of the file inside a READ, ends the loop.
    Pure-Python statement of the library's srt_read_inputrays (csrc/srt_io.cpp)."""
    with open(path, "rb") as f:
        text = f.read().decode("latin-1")
    vals, cur, p, n = [], [0.0] * 7, 0, 0
    n = len(text)
    stop = False
    while not stop:
        idx, expecting, slash = 0, True, False
        while idx < 7 and not stop and not slash:
            while p < n and text[p] in _BLANK:
                p += 1
            if p >= n:
                stop = True
                break
            c = text[p]
            if c in ",;":
                p += 1
                if expecting:
                    idx += 1
                else:
                    expecting = True
                continue
            if c == "/":
                p += 1
                slash = True
                break
            q = p
            while q < n and text[q] not in _END:
                q += 1
            tok, p, rep = text[p:q], q, 1
            m = re.match(r"^(\d+)\*(.*)$", tok, re.DOTALL)
            if m:
                rep, tok = int(m.group(1)), m.group(2)
                if rep == 0:
                    stop = True
                    break
                if tok == "":
                    idx += min(rep, 7)
                    expecting = False
                    continue
            try:
                v = _fortran_float(tok)
            except ValueError:
                stop = True
                break
            while rep > 0 and idx < 7:
                cur[idx] = v
                idx += 1
                rep -= 1
            expecting = False
        if stop:
            break
        vals.append(list(cur))
        while p < n and text[p] != "\n":
            p += 1
        p += 1
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
