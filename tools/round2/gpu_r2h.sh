#!/bin/bash
# Round 2, eighth GPU call: damping tests with the final bound; ncu captures at the bench geometries (traffic.json,
# summaries, per-function breakdowns); launch list of the bench command
set -u
mkdir -p gpurun_out
out=gpurun_out/r2h.txt
: > $out
run() { echo "## $*" >> $out; timeout 1200 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_damping.py -q -s
N="ncu --set full --clock-control none --import-source on -k regex:srt_trace_team -s 1 -c 1"
python tools/profile_case.py team 100 100 2000 > gpurun_out/r2h_c2.plain.log 2>&1 && $N -o gpurun_out/r2h_c2 python tools/profile_case.py team 100 100 2000 >> $out 2>&1
python tools/profile_case.py team 100 1000 300 > gpurun_out/r2h_c4.plain.log 2>&1 && $N -o gpurun_out/r2h_c4 python tools/profile_case.py team 100 1000 300 >> $out 2>&1
python tools/c3_age.py --maxsteps 300 --reps 2 > gpurun_out/r2h_c3.plain.log 2>&1 && $N -o gpurun_out/r2h_c3 python tools/c3_age.py --maxsteps 300 --reps 2 >> $out 2>&1
python bench.py --steps 2 --warmup 1 --extras none --no-cpu-baseline > gpurun_out/r2h_bench.plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2h_bench_launches.csv python bench.py --steps 2 --warmup 1 --extras none --no-cpu-baseline > /dev/null 2>&1
tail -c 1500 $out
