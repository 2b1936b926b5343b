#!/bin/bash
# Round 2, third GPU call (2 GPUs): multi-device library fan-out, sharded two-rank path, device FD kernel (grid tests), bench at N=2
set -u
mkdir -p gpurun_out
out=gpurun_out/r2c.txt
: > $out
run() { echo "## $*" >> $out; timeout 1200 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_multi.py tests/test_gpu_grid.py tests/test_dumpmodel.py tests/test_gpu_cli.py -x -q -m gpu
run SRT_X=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 1
tail -30 $out
