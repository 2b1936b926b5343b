#!/bin/bash
# Round 2, last GPU seconds: shadow groups of four (SRT_TEAM_SHADOW=4) -- bit identity with 0 and 1, and config 2 timed with 4 and 1 on one box
set -u
mkdir -p gpurun_out
out=gpurun_out/r2zb.txt
: > $out
(timeout 45 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k shadow_teams 2>&1 | tail -6) >> $out
(SRT_TEAM_SHADOW=4 timeout 40 python tools/gpu_variants.py 100 100 20000 team 2>&1 | grep -v Warn | tail -2) >> $out
(SRT_TEAM_SHADOW=1 timeout 40 python tools/gpu_variants.py 100 100 20000 team 2>&1 | grep -v Warn | tail -2) >> $out
cat $out
