#!/bin/bash
# Round 2, sixth GPU call: interleaved division (srt_divn): self-test, parity, A/B against the build without it; damping tests
set -u
mkdir -p gpurun_out
out=gpurun_out/r2f.txt
: > $out
run() { echo "## $*" >> $out; timeout 1500 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_parity.py -q -k "division or plasma_eval or c1_fixed or first_attempt or stop_code or ducts or evanescent or nan_state"
run SRT_X=0 python -m pytest tests/test_gpu_damping.py tests/test_gpu_grid.py -q -s
run SRT_X=0 python tools/gpu_variants.py 100 100 20000 team
run SRT_X=0 python tools/gpu_variants.py 200 200 200 team
run SRT_X=0 python tools/gpu_variants.py 50 40 300 team
run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --save /tmp/c3_1000.npy
run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --lib tools/_variants/nodivn.so --compare /tmp/c3_1000.npy
tail -c 3000 $out
