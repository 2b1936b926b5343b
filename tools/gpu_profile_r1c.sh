#!/bin/bash
# round-1c: full ncu capture of the team kernel (throughput regime and latency regime)
mkdir -p gpurun_out
T="python tools/profile_case.py team 100 96 200"
$T > gpurun_out/r1c_plain_team.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1c_team $T > gpurun_out/r1c_ncu_team.log 2>&1
T="python tools/profile_case.py team 50 40 300"
$T > gpurun_out/r1c_plain_team_small.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -f -o gpurun_out/r1c_team_small $T > gpurun_out/r1c_ncu_team_small.log 2>&1
cat gpurun_out/r1c_plain_team.log gpurun_out/r1c_plain_team_small.log
