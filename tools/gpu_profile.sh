#!/bin/bash
# tools/gpu_profile.sh TAG -- CMD...      (run under gpurun; ONE ncu tool per call)
# The profiling recipe of /opt/skills/guides/B200_PROFILING.md for one case: the plain run first (must exit 0), then the
# launch list and one `--set full` capture of the second launch of the trace kernel, into gpurun_out/TAG.*; read here with
#   python tools/ncu_summary.py gpurun_out/TAG.ncu-rep profiles/TAG_ncu_summary.txt
#   python tools/ncu_by_function.py gpurun_out/TAG.ncu-rep profiles/TAG_by_function.txt
# Cases used in round 2 (the exact calls are kept under tools/round2/):
#   python tools/profile_case.py team 50 40 300       latency regime (2 000 rays, 128-thread blocks)
#   python tools/profile_case.py team 200 200 200     throughput regime (40 000 rays, 1024-thread blocks)
#   python tools/c3_age.py --maxsteps 1000 --nlon 2 --reps 2      config-3 slice (grid + IGRF, 20 000 rays)
set -u
tag=$1; shift; [ "$1" = "--" ] && shift
mkdir -p gpurun_out
"$@" > gpurun_out/$tag.plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/$tag.plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 50 --csv --log-file gpurun_out/$tag.launches.csv "$@" > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:srt_trace -s 1 -c 1 -o gpurun_out/$tag "$@" > gpurun_out/$tag.ncu.log 2>&1
tail -2 gpurun_out/$tag.ncu.log
