"""The N > 1 host logic (rays shard by index, host-side gather on rank 0, no data-path collective) on
CPU: world_size 2 over gloo.  The per-shard tracer is injected -- here the CPU oracle stands in for the
GPU call, which cannot run in this container; the product's own tracer is raytracer_run_sharded."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as tmp

import common
from stanford_raytracer_b200 import shard


def test_shard_indices_partition():
    for n in (0, 1, 7, 10000):
        for world in (1, 2, 3, 8):
            got = np.sort(np.concatenate([shard.shard_indices(n, r, world) for r in range(world)]))
            assert np.array_equal(got, np.arange(n))
            sizes = [len(shard.shard_indices(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_merge_detects_loss_and_duplication():
    with pytest.raises(ValueError):
        shard.merge_shards(4, [(np.array([0, 2]), ["a", "b"])])
    with pytest.raises(ValueError):
        shard.merge_shards(2, [(np.array([0, 1]), ["a", "b"]), (np.array([1]), ["c"])])
    assert shard.merge_shards(3, [(np.array([0, 2]), ["a", "c"]), (np.array([1]), ["b"])]) == ["a", "b", "c"]


def _worker(rank, world, port, out_path):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, common.ROOT)
    sys.path.insert(0, os.path.join(common.ROOT, "tests"))
    import torch.distributed as dist
    import oracle_lib as ol
    dist.init_process_group("gloo", rank=rank, world_size=world)
    model = ol.OracleModel("ngo", common.NGO_CFG)
    P = dict(common.C2, maxsteps=40)

    def trace(p, d, w):
        return [(lambda o: (o["rows"], o["stopcond"]))(model.trace(p[i], d[i], w[i], **P)) for i in range(len(w))]

    pos0, dir0, w0 = common.c2_grid(3, 3)
    res = shard.run_sharded(trace, pos0, dir0, w0, rank, world)
    if rank == 0:
        np.savez(out_path, n=len(res), **{"rows_%d" % i: r[0] for i, r in enumerate(res)}, stop=np.array([r[1] for r in res]))
    else:
        assert res is None
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_equals_single_process(built, tmp_path):
    import oracle_lib as ol
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    out = str(tmp_path / "gathered.npz")
    tmp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    z = np.load(out)
    model = ol.OracleModel("ngo", common.NGO_CFG)
    pos0, dir0, w0 = common.c2_grid(3, 3)
    assert int(z["n"]) == len(w0)
    for i in range(len(w0)):
        o = model.trace(pos0[i], [0, 0, 0], w0[i], **dict(common.C2, maxsteps=40))
        assert np.array_equal(z["rows_%d" % i], o["rows"], equal_nan=True) and z["stop"][i] == o["stopcond"]
