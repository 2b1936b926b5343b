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
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-fmad=false", "-lineinfo", "-w",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fno-fast-math",
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def sources():
    out = [os.path.join(HERE, "..", "include", "srt.h"), os.path.abspath(__file__)]  # a flag change here rebuilds too
    for f in sorted(os.listdir(CSRC)):
        if f.endswith((".h", ".cu", ".cuh", ".inl", ".cpp")):
            out.append(os.path.join(CSRC, f))
    return out


def build(force=False, verbose=False):
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    srcs = sources()
    if force or _newer(LIB, srcs):
        extra = (["-Xptxas", "-v"] if verbose else [])
        # two translation units, compiled concurrently: the thread-per-ray kernel + C ABI and the
        # team-of-8 kernel in two register budgets; elementary functions inlined in both; see srt_team_tu.cu
        objs = [os.path.join(HERE, "srt_api.o"), os.path.join(HERE, "srt_team_tu.o"), os.path.join(HERE, "srt_damping.o")]
        procs = [subprocess.Popen([nvcc] + NVCC_FLAGS + extra + ["-DSRT_INLINE_MATH", "-c", "-o", objs[0], os.path.join(CSRC, "srt_api.cu")]),
                 subprocess.Popen([nvcc] + NVCC_FLAGS + extra + ["-DSRT_INLINE_MATH", "-DSRT_INLINE_STIX", "-DSRT_INLINE_PLASMA", "-DSRT_INLINE_ACCEPT", "-c", "-o", objs[1], os.path.join(CSRC, "srt_team_tu.cu")])]
        # damping along the ray (config 5): its own unit
        procs.append(subprocess.Popen([nvcc] + NVCC_FLAGS + extra + ["-c", "-o", objs[2], os.path.join(CSRC, "srt_damping.cu")]))
        # host-only unit (text formats of the driver), plain g++
        gxx = shutil.which("g++") or "g++"
        objs.append(os.path.join(HERE, "srt_io.o"))
        procs.append(subprocess.Popen([gxx, "-O2", "-std=c++17", "-fPIC", "-pthread", "-c", "-o", objs[3], os.path.join(CSRC, "srt_io.cpp")]))
        for pr in procs:
            if pr.wait() != 0:
                raise subprocess.CalledProcessError(pr.returncode, pr.args)
        subprocess.check_call([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs + ["-lpthread"])
        for o in objs:
            os.remove(o)
    cli_src = os.path.join(CSRC, "srt_cli.cpp")
    if os.path.exists(cli_src) and (force or _newer(CLI, [cli_src, srcs[0], LIB])):
        gxx = shutil.which("g++") or "g++"
        subprocess.check_call([gxx, "-O2", "-std=c++17", "-ffp-contract=off", "-o", CLI, cli_src, "-I", os.path.join(HERE, "..", "include"),
                               "-L", HERE, "-lsrt", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
