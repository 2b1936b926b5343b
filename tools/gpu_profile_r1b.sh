#!/bin/bash
# round-1 (second pass, after the block-synchronous kernels): launch list + DRAM bytes of the bench, one full
# ncu capture per trace-kernel form.  Run under gpurun.
mkdir -p gpurun_out
B="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
$B > gpurun_out/r1b_plain_bench.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r1b_launches.csv $B > gpurun_out/r1b_ncu_bench.log 2>&1
T="python tools/profile_case.py team 100 96 200"
$T > gpurun_out/r1b_plain_team.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1b_team $T > gpurun_out/r1b_ncu_team.log 2>&1
T="python tools/profile_case.py thread 200 200 150"
$T > gpurun_out/r1b_plain_thread.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1b_thread $T > gpurun_out/r1b_ncu_thread.log 2>&1
cat gpurun_out/r1b_plain_team.log gpurun_out/r1b_plain_thread.log
tail -2 gpurun_out/r1b_ncu_team.log gpurun_out/r1b_ncu_thread.log
