#!/bin/bash
# r1e: live-team compaction.  parity tests, bench, mode timings
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "not full_c2" > gpurun_out/r1e_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1e_pytest.log
timeout 400 python bench.py --steps 3 --warmup 3 > gpurun_out/r1e_bench.json 2> gpurun_out/r1e_bench.err; echo "bench rc=$?" >> gpurun_out/r1e_bench.err
timeout 200 python tools/gpu_variants.py 200 200 400 team > gpurun_out/r1e_modes.log 2>&1
timeout 200 python tools/gpu_variants.py 50 40 700 team >> gpurun_out/r1e_modes.log 2>&1
tail -3 gpurun_out/r1e_pytest.log; tail -1 gpurun_out/r1e_bench.json; cat gpurun_out/r1e_modes.log
exit 0
