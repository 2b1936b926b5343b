#!/bin/bash
# r1h: 255-register variant timing, final ncu captures, full-size config 3 and config 4 lines
mkdir -p gpurun_out
( SRT_TEAM_BLOCK=256 timeout 120 python tools/gpu_variants.py 50 40 700 team; SRT_TEAM_BLOCK=256 timeout 200 python tools/gpu_variants.py 100 100 3000 team; timeout 100 python tools/gpu_variants.py 50 40 700 team ) > gpurun_out/r1h_variants.log 2>&1
cat gpurun_out/r1h_variants.log
T="python tools/profile_case.py team 50 40 300"
$T > gpurun_out/r1h_plain_team_small.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1h_team_small $T > gpurun_out/r1h_ncu_team_small.log 2>&1
T="python tools/profile_case.py team 200 200 150"
$T > gpurun_out/r1h_plain_team.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1h_team $T > gpurun_out/r1h_ncu_team.log 2>&1
cat gpurun_out/r1h_plain_team_small.log gpurun_out/r1h_plain_team.log
BN="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 300 $BN > gpurun_out/r1h_plain_bench.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1h_launches.csv $BN > gpurun_out/r1h_ncu_bench.log 2>&1
tail -1 gpurun_out/r1h_plain_bench.log | cut -c1-300
timeout 900 python bench.py --workload c3 --steps 1 --warmup 3 > gpurun_out/r1h_bench_c3.json 2> gpurun_out/r1h_bench_c3.err; echo "c3 rc=$?"
tail -1 gpurun_out/r1h_bench_c3.json | cut -c1-1500
timeout 1100 python bench.py --workload c4 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/r1h_bench_c4.json 2> gpurun_out/r1h_bench_c4.err; echo "c4 rc=$?"
tail -1 gpurun_out/r1h_bench_c4.json | cut -c1-1500
exit 0
