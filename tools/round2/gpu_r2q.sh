#!/bin/bash
# Round 2: shadow teams with the guess following the ray's last failure kind: parity, then A/B on the same box
set -u
mkdir -p gpurun_out
out=gpurun_out/r2q.txt
: > $out
run() { echo "## $*" >> $out; timeout 600 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_parity.py tests/test_gpu_c3_long.py -q -x
run SRT_TEAM_SHADOW=1 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_SHADOW=0 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_SHADOW=1 python tools/gpu_variants.py 200 200 200 team
run SRT_TEAM_SHADOW=0 python tools/gpu_variants.py 200 200 200 team
run SRT_TEAM_SHADOW=1 python bench.py --workload c3 --steps 1 --warmup 1 --extras c4_100k --no-cpu-baseline --no-e2e
tail -c 2500 $out
