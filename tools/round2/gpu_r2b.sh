#!/bin/bash
# Round 2, second GPU call: the new default build (phase alignment, cell staging, NaN fast-forward always on):
# whole GPU test-suite, FP64 latency probe, timings of configs 2/3 and the throughput regime, ncu of three geometries.
set -u
mkdir -p gpurun_out
out=gpurun_out/r2b.txt
: > $out
run() { echo "## $*" >> $out; timeout 900 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests -m gpu -x -q
run SRT_X=0 tools/_bin/fp64_latency
run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 100 100 20000 team
run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 200 200 200 team
run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 50 40 300 team
run SRT_X=0 python tools/c3_age.py --maxsteps 4000
run SRT_X=0 python bench.py --workload c3 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline
# ncu (one tool per call): plain run first, then the capture of the second launch
N="ncu --set full --clock-control none --import-source on -k regex:srt_trace_team -s 1 -c 1"
python tools/profile_case.py team 50 40 300 > gpurun_out/r2b_plain1.log 2>&1 && $N -o gpurun_out/r2b_team_small python tools/profile_case.py team 50 40 300 >> $out 2>&1
python tools/profile_case.py team 200 200 200 > gpurun_out/r2b_plain2.log 2>&1 && $N -o gpurun_out/r2b_team_40k python tools/profile_case.py team 200 200 200 >> $out 2>&1
python tools/c3_age.py --maxsteps 1000 --nlon 2 --reps 2 > gpurun_out/r2b_plain3.log 2>&1 && $N -o gpurun_out/r2b_team_c3 python tools/c3_age.py --maxsteps 1000 --nlon 2 --reps 2 >> $out 2>&1
tail -40 $out
