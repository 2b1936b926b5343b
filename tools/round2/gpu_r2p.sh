#!/bin/bash
# Round 2: shadow teams (srt_team.cuh): parity suites first (every run under its own timeout: a phase error between a primary and
# its shadow would hang a warp), then A/B against SRT_TEAM_SHADOW=0 on the same box
set -u
mkdir -p gpurun_out
out=gpurun_out/r2p.txt
: > $out
run() { echo "## $*" >> $out; timeout 600 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_parity.py tests/test_gpu_c3_long.py tests/test_gpu_grid.py tests/test_gpu_ts05.py -q -x
run SRT_TEAM_SHADOW=0 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_SHADOW=1 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_SHADOW=0 python tools/gpu_variants.py 200 200 200 team
run SRT_TEAM_SHADOW=1 python tools/gpu_variants.py 200 200 200 team
run SRT_TEAM_SHADOW=0 python tools/gpu_variants.py 50 40 300 team
run SRT_TEAM_SHADOW=1 python tools/gpu_variants.py 50 40 300 team
run SRT_TEAM_SHADOW=0 python bench.py --workload c4_100k --steps 1 --warmup 1 --extras c3 --no-cpu-baseline
run SRT_TEAM_SHADOW=1 python bench.py --workload c4_100k --steps 1 --warmup 1 --extras c3 --no-cpu-baseline
tail -c 2500 $out
