#!/bin/bash
# r1l: stix / F / derive inlined into the team kernel's rounds
mkdir -p gpurun_out
L=gpurun_out/r1l_variants.log; : > $L
timeout 150 python tools/gpu_variants.py 100 100 20000 team >> $L 2>&1
timeout 100 python tools/gpu_variants.py 50 40 700 team >> $L 2>&1
timeout 100 python tools/gpu_variants.py 200 200 200 team >> $L 2>&1
cat $L
