#!/bin/bash
# Round 2, final GPU call: the driver's own sequence on the final code (GPU suite, smoke, bench with the driver's arguments,
# reference arm), and an ncu capture of the team kernel with the TS05 external field summarised on the box
set -u
mkdir -p gpurun_out
out=gpurun_out/r2s.txt
: > $out
run() { echo "## $*" >> $out; timeout 1500 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests -m gpu -q --durations=5
run SRT_X=0 python -c "import __graft_entry__ as g; g.smoke()"
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2s_bench_line.json 2> gpurun_out/r2s_bench.err; echo "bench rc=$?" >> $out
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2s_bench_reference.json 2> gpurun_out/r2s_bench_reference.err; echo "reference rc=$?" >> $out
# ncu of the team kernel with the TS05 external field (use_tsyganenko = 1), summarised on the box
cat > /tmp/ts05case.py <<'PY'
import sys; sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch, common, stanford_raytracer_b200 as srt
srt.init(0)
m = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 1).tsyganenko(3.0, -45.0, 2.5, -4.0, W=[0.4, 0.7, 0.3, 0.9, 0.6, 1.1])
pos0, dir0, w0 = common.c2_grid(100, 100)
n = len(w0)
p = dict(common.C2, maxsteps=40)
P = srt.RunParams(p["dt0"], p["dtmax"], p["maxerr"], p["minalt"], p["tmax"], p["maxsteps"], p["root"], p["fixedstep"], 1)
rays = torch.from_numpy(np.ascontiguousarray(np.concatenate([pos0.T, dir0.T, w0[None, :]], axis=0))).cuda()
summary = torch.zeros((n, 24), dtype=torch.float64, device="cuda"); counters = torch.zeros(8, dtype=torch.int64, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); srt.trace_batch_device(m, n, rays.data_ptr(), P, 0, 0, summary.data_ptr(), counters.data_ptr(), 0); e1.record(); torch.cuda.synchronize()
c = counters.cpu().numpy()
print("ts05 case: rays", n, "accepted", c[1], "attempts", c[2], "ms %.1f" % e0.elapsed_time(e1), "steps/s %.3g" % (c[1] / e0.elapsed_time(e1) * 1e3))
PY
python /tmp/ts05case.py > gpurun_out/r2s_ts05.plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:srt_trace_team -c 1 -o /tmp/r2s_ts05 python /tmp/ts05case.py > /tmp/r2s_ts05.ncu.log 2>&1
python tools/ncu_summary.py /tmp/r2s_ts05.ncu-rep gpurun_out/r2s_ts05_ncu_summary.txt > /dev/null 2>&1
python tools/ncu_by_function.py /tmp/r2s_ts05.ncu-rep gpurun_out/r2s_ts05_by_function.txt > /dev/null 2>&1
cat gpurun_out/r2s_ts05.plain.log >> $out
du -sh gpurun_out >> $out
tail -c 1500 $out
