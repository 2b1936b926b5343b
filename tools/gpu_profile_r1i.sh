#!/bin/bash
# r1i: register budget of the scarce-ray team kernel on full config 2
mkdir -p gpurun_out
L=gpurun_out/r1i_variants.log; : > $L
SRT_VARIANT_ONLY=libsrt timeout 100 python tools/gpu_variants.py 100 100 20000 team >> $L 2>&1
SRT_VARIANT_ONLY=maxt256 SRT_TEAM_BLOCK=256 timeout 100 python tools/gpu_variants.py 100 100 20000 team >> $L 2>&1
SRT_VARIANT_ONLY=maxt320 SRT_TEAM_BLOCK=320 timeout 100 python tools/gpu_variants.py 100 100 20000 team >> $L 2>&1
SRT_VARIANT_ONLY=maxt320 SRT_TEAM_BLOCK=128 timeout 100 python tools/gpu_variants.py 50 40 700 team >> $L 2>&1
SRT_VARIANT_ONLY=maxt320 SRT_TEAM_BLOCK=320 timeout 100 python tools/gpu_variants.py 100 150 300 team >> $L 2>&1
SRT_VARIANT_ONLY=libsrt timeout 100 python tools/gpu_variants.py 100 150 300 team >> $L 2>&1
cat $L
