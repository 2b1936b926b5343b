#!/bin/bash
# NOT YET RUN (the round's GPU minutes were spent when modelnum 4 was written).  The first GPU call for the scattered-sample
# model: its parity tests, its bench figure, then -- each only after the plain command has exited 0 -- the launch list and one
# `ncu --set full` capture of its evaluation kernel and of its team kernel (tools/ncu_summary.py turns the reports into
# profiles/ summaries).
#   gpurun --timeout 900 -- bash tools/round2/gpu_next_scattered.sh
set -u
mkdir -p gpurun_out
out=gpurun_out/scattered.txt
: > $out
(timeout 300 python -m pytest tests/test_zy_gpu_scattered.py tests/test_zz_no_alloc.py -m gpu -q -x 2>&1 | tail -15) >> $out
(timeout 200 python tools/scattered_bench.py 2>&1 | tail -1) >> $out
if grep -q '"value"' $out; then
  (timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/scattered_launches.csv \
     python tools/scattered_bench.py > gpurun_out/scattered_ncu_list.log 2>&1)
  (timeout 400 ncu --set full --clock-control none --import-source on -k regex:srt_trace_team_kernel -c 1 -o gpurun_out/scattered_team \
     python tools/scattered_bench.py > gpurun_out/scattered_ncu_team.log 2>&1)
  (timeout 400 ncu --set full --clock-control none --import-source on -k regex:srt_eval_kernel -s 2 -c 1 -o gpurun_out/scattered_eval \
     python tools/scattered_bench.py > gpurun_out/scattered_ncu_eval.log 2>&1)
fi
cat $out
