#!/bin/bash
# Round 2, ninth GPU call: the driver's own sequence on the final code (GPU suite, smoke, bench with the driver's arguments,
# reference arm), and an ncu capture of the damping kernels summarised on the box
set -u
mkdir -p gpurun_out
out=gpurun_out/r2i.txt
: > $out
run() { echo "## $*" >> $out; timeout 1500 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests -m gpu -q --durations=5
run SRT_X=0 python -c "import __graft_entry__ as g; g.smoke()"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2i_bench_line.json 2> gpurun_out/r2i_bench.err; echo "bench rc=$?" >> $out
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2i_bench_reference.json 2> gpurun_out/r2i_bench_reference.err; echo "reference rc=$?" >> $out
cat > /tmp/dampcase.py <<'PY'
import sys; sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, time, common, stanford_raytracer_b200 as srt
from stanford_raytracer_b200 import damping
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
pos0, dir0, w0 = common.c2_grid(40, 50)
res = srt.raytracer_run(m, pos0, dir0, w0, outputper=50, **common.C2)
for _ in range(2):
    t = time.perf_counter(); f, p = damping.damp_result(res); dt = time.perf_counter() - t
print("damping case: rays", res.nrays, "rows", res.total_rows, "s %.3f" % dt, "rows/s %.3g" % (res.total_rows / dt))
PY
python /tmp/dampcase.py > gpurun_out/r2i_damp.plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:srt_damp_row -s 1 -c 1 -o /tmp/r2i_damp python /tmp/dampcase.py > /tmp/r2i_damp.ncu.log 2>&1
python tools/ncu_summary.py /tmp/r2i_damp.ncu-rep gpurun_out/r2i_damp_ncu_summary.txt > /dev/null 2>&1
cat gpurun_out/r2i_damp.plain.log >> $out
du -sh gpurun_out >> $out
tail -c 1500 $out
