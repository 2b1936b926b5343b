#!/bin/bash
# Round 2, eighth GPU call (second try: the first one brought back more than gpurun's 64 MiB and was lost): damping tests
# with the final bound; 768-thread / 80-register A/B; ncu captures at the bench geometries, summarised ON THE BOX
# (ncu summary, per-function shares, traffic.json) so that only text comes back; launch list of the bench command
set -u
mkdir -p gpurun_out
out=gpurun_out/r2h.txt
: > $out
run() { echo "## $*" >> $out; timeout 1200 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_damping.py -q -s
run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 200 200 200 team
run SRT_VARIANT_ONLY=mid768 SRT_TEAM_BLOCK=768 python tools/gpu_variants.py 200 200 200 team
run SRT_VARIANT_ONLY=libsrt python tools/gpu_variants.py 100 1000 300 team
run SRT_VARIANT_ONLY=mid768 SRT_TEAM_BLOCK=768 python tools/gpu_variants.py 100 1000 300 team
run SRT_X=0 python tools/c3_age.py --maxsteps 1000 --save /tmp/c3_1000.npy
run SRT_TEAM_BLOCK=768 python tools/c3_age.py --maxsteps 1000 --lib tools/_variants/mid768.so --compare /tmp/c3_1000.npy
N="ncu --set full --clock-control none --import-source on -k regex:srt_trace_team -s 1 -c 1"
cap() {  # tag workload cmd...
  tag=$1; wl=$2; shift 2
  "$@" > gpurun_out/$tag.plain.log 2>&1 && $N -o /tmp/$tag "$@" > /tmp/$tag.ncu.log 2>&1
  python tools/ncu_summary.py /tmp/$tag.ncu-rep gpurun_out/${tag}_ncu_summary.txt > /dev/null 2>&1
  python tools/ncu_by_function.py /tmp/$tag.ncu-rep gpurun_out/${tag}_by_function.txt > /dev/null 2>&1
  python tools/traffic_from_ncu.py $wl /tmp/$tag.ncu-rep gpurun_out/$tag.plain.log >> $out 2>&1
  rm -f /tmp/$tag.ncu-rep
}
cap r2h_c2 c2 python tools/profile_case.py team 100 100 1000
cap r2h_c4 c4_100k python tools/profile_case.py team 100 1000 150
cap r2h_c3 c3 python tools/c3_age.py --maxsteps 150 --reps 2
cp profiles/traffic.json gpurun_out/traffic.json
python bench.py --steps 2 --warmup 1 --extras none --no-cpu-baseline > gpurun_out/r2h_bench.plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2h_bench_launches.csv python bench.py --steps 2 --warmup 1 --extras none --no-cpu-baseline > /dev/null 2>&1
du -sh gpurun_out >> $out
tail -c 1500 $out
