#!/bin/bash
# Round 2, fourth GPU call: whole GPU suite with the new parity tests, then the default bench line with extras
set -u
mkdir -p gpurun_out
out=gpurun_out/r2d.txt
: > $out
run() { echo "## $*" >> $out; timeout 1500 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests -m gpu -q --durations=8
run SRT_X=0 python bench.py --steps 5 --warmup 3
tail -c 2500 $out
