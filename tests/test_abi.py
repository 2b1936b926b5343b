"""The C-ABI library: loads, exports every symbol include/srt.h declares, and FAILS LOUDLY without a
GPU (no CPU fallback).  No compute call succeeds here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import common


def declared_symbols():
    txt = open(os.path.join(common.ROOT, "include", "srt.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(srt_[a-z0-9_]+)\s*\(", txt)))


def test_header_symbols_exported(built):
    L = C.CDLL(os.path.join(common.ROOT, "stanford_raytracer_b200", "libsrt.so"))
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(L, s), "libsrt.so does not export %s" % s


def test_no_torch_types_in_header():
    txt = open(os.path.join(common.ROOT, "include", "srt.h")).read()
    assert "torch" not in txt.lower() and "at::" not in txt and "std::" not in txt


def test_compute_fails_loudly_without_gpu(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import stanford_raytracer_b200 as srt
    L = srt.lib()
    assert L.srt_init(-1) == -1  # SRT_E_NODEVICE
    assert b"no CPU path" in L.srt_strerror(-1) or b"CUDA" in L.srt_strerror(-1)
    # host-only set-up works (it is file parsing), compute does not
    m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    pos0, dir0, w0 = common.launch_grid([2000.0], [30.0])
    with pytest.raises(srt.SrtError):
        srt.raytracer_run(m, pos0, dir0, w0, **common.C1)
    with pytest.raises(srt.SrtError):
        m.funcPlasmaParams(pos0)
    with pytest.raises(srt.SrtError):
        srt.interpsetup(os.path.join(common.ROOT, "tests", "golden", "grid13.txt"), 2010001, 0, 1, 0)
    with pytest.raises(srt.SrtError):  # modelnum 4 places its samples on the devices at set-up
        srt.scatteredinterpsetup(os.path.join(common.ROOT, "tests", "golden", "grid13.txt"), 1.5, 2, 0, 3.0)
    assert L.srt_model_scattered_samples(None, None, 0) == -3 and L.srt_model_scattered_info(None, None, None, None, None, None) == -3


def test_argument_errors(built):
    import stanford_raytracer_b200 as srt
    with pytest.raises(srt.SrtError):
        srt.ngosetup("/nonexistent/newray.in")
    with pytest.raises(srt.SrtError):
        srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 2)  # use_tsyganenko is 0 or 1
    L = srt.lib()
    assert L.srt_trace_batch(None, 1, None, None, None, None, None) == -3
    assert L.srt_result_nrays(None) == 0


def test_product_never_touches_oracle():
    """the product package must not import, link or execute anything under oracle/"""
    pkg = os.path.join(common.ROOT, "stanford_raytracer_b200")
    for root, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".h", ".cu", ".cuh", ".inl", ".cpp")):
                txt = open(os.path.join(root, f), errors="replace").read()
                assert "oracle_lib" not in txt and "orc_" not in txt and "/oracle/" not in txt, f
