#!/bin/bash
# Round 2: block size of the team kernel on config 2 (10 000 rays = 68 per SM: barely more than one wave of a 512-thread block)
set -u
mkdir -p gpurun_out
out=gpurun_out/r2y.txt
: > $out
run() { echo "## $*" >> $out; timeout 100 env "$@" 2>&1 | grep -v Warning >> $out || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_BLOCK=256 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_BLOCK=384 python tools/gpu_variants.py 100 100 20000 team
run SRT_TEAM_BLOCK=128 python tools/gpu_variants.py 100 100 20000 team
cat $out
