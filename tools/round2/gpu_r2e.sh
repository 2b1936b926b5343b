#!/bin/bash
# Round 2, fifth GPU call: damping tests against the reference binary, bench with the c5 extra (short)
set -u
mkdir -p gpurun_out
out=gpurun_out/r2e.txt
: > $out
run() { echo "## $*" >> $out; timeout 1500 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_damping.py tests/test_damping_cpu.py -q -s
run SRT_X=0 python bench.py --steps 2 --warmup 1 --extras c5
tail -c 2500 $out
