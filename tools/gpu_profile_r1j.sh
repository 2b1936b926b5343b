#!/bin/bash
# r1j: ncu captures of the final round-1 team kernel (latency regime, full set) and the bench launch list with DRAM bytes
mkdir -p gpurun_out
T="python tools/profile_case.py team 50 40 300"
$T > gpurun_out/r1j_plain_team_small.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1j_team_small $T > gpurun_out/r1j_ncu_team_small.log 2>&1
cat gpurun_out/r1j_plain_team_small.log
BN="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $BN > gpurun_out/r1j_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1j_launches.csv $BN > gpurun_out/r1j_ncu_bench.log 2>&1
tail -1 gpurun_out/r1j_plain_bench.log | cut -c1-300
ls -la gpurun_out
exit 0
