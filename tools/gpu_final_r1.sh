#!/bin/bash
# round-1 final check: the driver's own sequence (gpu tests, smoke, bench, reference arm)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r1z_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1z_pytest.log
tail -3 gpurun_out/r1z_pytest.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1z_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r1z_smoke.log
tail -2 gpurun_out/r1z_smoke.log
timeout 600 python bench.py > gpurun_out/r1z_bench.json 2> gpurun_out/r1z_bench.err; echo "bench rc=$?"
tail -1 gpurun_out/r1z_bench.json
timeout 400 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1z_bench_ref.json 2> gpurun_out/r1z_bench_ref.err; echo "ref rc=$?"
tail -1 gpurun_out/r1z_bench_ref.json
exit 0
