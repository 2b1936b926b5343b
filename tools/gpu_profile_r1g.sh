#!/bin/bash
# r1g: zero-numerator division shortcut.  parity tests, bench, mode timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "not full_c2" > gpurun_out/r1g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1g_pytest.log
timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/r1g_bench.json 2> gpurun_out/r1g_bench.err; echo "bench rc=$?" >> gpurun_out/r1g_bench.err
timeout 200 python tools/gpu_variants.py 200 200 200 team > gpurun_out/r1g_modes.log 2>&1
timeout 200 python tools/gpu_variants.py 50 40 700 team >> gpurun_out/r1g_modes.log 2>&1
tail -3 gpurun_out/r1g_pytest.log; tail -1 gpurun_out/r1g_bench.json; cat gpurun_out/r1g_modes.log

timeout 300 python tools/c2_tail2.py > gpurun_out/r1g_c2_tail2.log 2>&1; cat gpurun_out/r1g_c2_tail2.log | cut -c1-250
exit 0
