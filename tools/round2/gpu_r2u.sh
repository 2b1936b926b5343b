#!/bin/bash
# Round 2, last GPU call: the final tree (oracle/_ref rebuilt from scratch, tests added since r2s) -- smoke, the quick GPU tests,
# the allocation-free entry point alone, and a short bench line
set -u
mkdir -p gpurun_out
out=gpurun_out/r2u.txt
: > $out
run() { echo "## $*" >> $out; timeout 400 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -c "import __graft_entry__ as g; g.smoke()"
run SRT_X=0 python -m pytest tests/test_zz_no_alloc.py -m gpu -q
run SRT_X=0 python -m pytest tests/test_gpu_parity.py tests/test_gpu_ts05.py tests/test_gpu_grid.py -m gpu -q -x --durations=3 -k "not thread"
tail -c 1200 $out
