#!/bin/bash
# Round 2, seventh GPU call (8 GPUs): the sharded bench at N = 8 with the C4-shaped and full C4 (1 M rays) extras; damping tests
set -u
mkdir -p gpurun_out
out=gpurun_out/r2g.txt
: > $out
run() { echo "## $*" >> $out; timeout 1300 env "$@" >> $out 2>&1 || echo "FAILED rc=$?" >> $out; }
run SRT_X=0 python -m pytest tests/test_gpu_damping.py tests/test_gpu_multi.py -q -s
run SRT_X=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 2 --warmup 1 --extras c4_100k,c4
df -h /dev/shm >> $out 2>&1; free -g >> $out 2>&1
tail -c 3000 $out
