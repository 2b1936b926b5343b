"""Multi-GPU paths of the product against its own single-GPU result, bit for bit:
  * one process per GPU (shard.raytracer_run_sharded: ONE launch set split by ray index, per-ray records gathered on
    rank 0 over gloo, rows mapped from the owners' shared-memory segments) -- world_size 2; on a one-GPU box both
    ranks share device 0, which exercises the same host path;
  * one process, several devices (srt_init_devices + srt_trace_batch's internal fan-out) -- needs two GPUs.
The single-GPU result itself is compared with the oracle in tests/test_gpu_parity.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as tmp

import common

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, common.ROOT)
    sys.path.insert(0, os.path.join(common.ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import stanford_raytracer_b200 as srt
    from stanford_raytracer_b200 import shard
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    srt.init(dev)
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    pos0, dir0, w0 = common.c2_grid(6, 7)
    p = dict(common.C2, maxsteps=400)
    res = shard.raytracer_run_sharded(model, pos0, dir0.copy(), w0, rank, world, outputper=3, **p)
    if rank == 0:
        srt.set_rows_shm(None)
        one = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=3, **p)
        ok = (np.array_equal(res.last_row, one.last_row, equal_nan=True) and np.array_equal(res.nsteps, one.nsteps) and
              np.array_equal(res.stopcond, one.stopcond) and np.array_equal(res.nrejected, one.nrejected) and np.array_equal(res.dir0, one.dir0) and
              res.accepted_steps == one.accepted_steps and res.total_rows == one.total_rows and res.attempts == one.attempts)
        for r in range(len(w0)):
            ok = ok and np.array_equal(np.array(res.rows(r)), one.rows(r), equal_nan=True)
        open(out_path, "w").write("ok" if ok else "MISMATCH")
        one.close()
    res.close()
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_two_ranks_equal_single_gpu(built, tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "verdict.txt")
    tmp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert open(out).read() == "ok"
    assert not [f for f in os.listdir("/dev/shm") if f.startswith("srt_rows.")]


def test_library_fanout_equals_single_device(built):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (run with gpurun --gpus 2)")
    import stanford_raytracer_b200 as srt
    pos0, dir0, w0 = common.c2_grid(8, 9)
    p = dict(common.C2, maxsteps=500)
    srt.init(0)
    model = srt.ngosetup(common.NGO_CFG, 2010001, 0, 0, 0)
    one = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=2, **p)
    assert one.ndevices == 1
    n = srt.init_devices(None)
    assert n >= 2
    try:
        many = srt.raytracer_run(model, pos0, dir0.copy(), w0, outputper=2, **p)
        assert many.ndevices == min(n, len(w0))
        assert np.array_equal(many.last_row, one.last_row, equal_nan=True) and np.array_equal(many.nsteps, one.nsteps)
        assert np.array_equal(many.stopcond, one.stopcond) and np.array_equal(many.dir0, one.dir0)
        assert many.accepted_steps == one.accepted_steps and many.total_rows == one.total_rows
        for r in range(len(w0)):
            assert np.array_equal(many.rows(r), one.rows(r), equal_nan=True)
    finally:
        srt.init(0)
