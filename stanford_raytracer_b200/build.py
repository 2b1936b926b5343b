"""Builds libsrt.so (the C-ABI library with the sm_100a kernels) in-tree with nvcc.

    python -m stanford_raytracer_b200.build

Flags that matter for parity: -fmad=false (no FMA contraction: the reference is built
without it, SURVEY.md Q19, and the CPU oracle is compiled -ffp-contract=off), IEEE
division and square root (nvcc defaults), no fast-math.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libsrt.so")
CLI = os.path.join(HERE, "raytracer")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-fmad=false", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-fast-math",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def sources():
    out = [os.path.join(HERE, "..", "include", "srt.h")]
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".h", ".cu", ".cuh", ".inl", ".cpp")):
            out.append(os.path.join(CSRC, f))
    return out


def build(force=False, verbose=False):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    srcs = sources()
    if force or _newer(LIB, srcs):
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-o", LIB, os.path.join(CSRC, "srt_api.cu")]
        subprocess.check_call(cmd)
    cli_src = os.path.join(CSRC, "srt_cli.cpp")
    if os.path.exists(cli_src) and (force or _newer(CLI, srcs + [LIB])):
        gxx = shutil.which("g++") or "g++"
        subprocess.check_call([gxx, "-O2", "-std=c++17", "-ffp-contract=off", "-o", CLI, cli_src, "-I", os.path.join(HERE, "..", "include"),
                               "-L", HERE, "-lsrt", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
