#!/usr/bin/env python
"""crres_clean.mat (the CRRES suprathermal electron flux survey the damping code fits its phase-space density to,
damping/data/crres_clean.mat, read by damping/src/psd_model.cpp:14-33,113-256) -> a flat little-endian table:

    stanford_raytracer_b200/data/crres_clean.bin
      int32  magic 0x53525443 ("CTRS"), n_entries (12), nL (70), nMLT (24)
      per entry:  float64 E [keV], float64 AE (1, 2 or 3), float64 L[nL], float64 MLT[nMLT],
                  float64 J_int[nL * nMLT]  column-major: J(l, m) at l + nL * m  (MATLAB order)

The .mat stores L and MLT as full nL x nMLT arrays; the code only ever looks at L(:, 1) and MLT(1, :), and the arrays are
exactly those vectors repeated (checked here), so the table keeps the vectors.  `t_b` is unused by the damping code.
Run in the container where /root/reference exists; the table is published data, committed like the IGRF coefficients."""
import os
import struct
import sys

import numpy as np
import scipy.io

src = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/damping/data/crres_clean.mat"
dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "stanford_raytracer_b200", "data", "crres_clean.bin")
c = scipy.io.loadmat(src)["crres"]
n = c.shape[1]
out = bytearray()
nL, nM = c[0, 0]["L"].shape
out += struct.pack("<4i", 0x53525443, n, nL, nM)
for k in range(n):
    e = c[0, k]
    L, M, J = e["L"].astype(np.float64), e["MLT"].astype(np.float64), e["J_int"].astype(np.float64)
    assert L.shape == M.shape == J.shape == (nL, nM)
    assert np.array_equal(L, np.repeat(L[:, :1], nM, axis=1)) and np.array_equal(M, np.repeat(M[:1, :], nL, axis=0))
    out += struct.pack("<2d", float(e["E"].ravel()[0]), float(e["AE"].ravel()[0]))
    out += L[:, 0].tobytes() + M[0, :].tobytes() + np.asfortranarray(J).tobytes(order="F")
open(dst, "wb").write(bytes(out))
print(dst, len(out), "bytes;", n, "entries", nL, "x", nM)
