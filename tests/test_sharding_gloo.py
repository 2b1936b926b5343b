"""The N > 1 host logic (ONE launch set sharded by ray index, fixed-size per-ray records gathered on rank 0 as
contiguous arrays, trajectory rows mapped from the owning rank's shared-memory segment, no data-path collective) on
CPU: world_size 2 over gloo.  The per-shard tracer is injected -- here the CPU oracle stands in for the GPU call, which
cannot run in this container, and puts its rows into a POSIX shared-memory segment the way libsrt does
(srt_set_rows_shm); the product's own tracer is shard.raytracer_run_sharded (GPU test: tests/test_gpu_multi.py)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as tmp

import common
from stanford_raytracer_b200 import shard


def test_shard_indices_partition():
    for n in (0, 1, 7, 64, 10000):
        for world in (1, 2, 3, 8):
            got = np.sort(np.concatenate([shard.shard_indices(n, r, world) for r in range(world)]))
            assert np.array_equal(got, np.arange(n))
            sizes = [len(shard.shard_indices(n, r, world)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1


def test_shards_mix_the_fastest_lattice_index():
    """8 longitudes (fastest index) on 8 ranks: every rank still sees every longitude."""
    world, nlon = 8, 8
    i = np.arange(100 * nlon)
    for r in range(world):
        mine = shard.shard_indices(len(i), r, world)
        assert len(set(mine % nlon)) == nlon


class _FakeKeep:
    def __init__(self, shm):
        self.shm = shm

    def close(self):
        if self.shm is not None:
            self.shm.close()
            self.shm.unlink()
            self.shm = None


def _oracle_tracer(model, P, use_shm, tag):
    from multiprocessing import shared_memory

    def trace(p, d, w):
        outs = [model.trace(p[i], d[i], w[i], **P) for i in range(len(w))]
        nrows = np.array([len(o["rows"]) for o in outs], dtype=np.int64)
        first = np.concatenate([[0], np.cumsum(nrows)[:-1]]).astype(np.int64) if len(outs) else np.zeros(0, dtype=np.int64)
        total = int(nrows.sum())
        shm, name = None, ""
        if use_shm:
            name = "srt_test_%s_%d" % (tag, os.getpid())
            shm = shared_memory.SharedMemory(name=name, create=True, size=max(total, 1) * 160)
            buf = np.ndarray((total, 20), dtype=np.float64, buffer=shm.buf)
        else:
            buf = np.empty((total, 20))
        for o, f in zip(outs, first):
            buf[f:f + len(o["rows"])] = o["rows"]
        meta = shard.pack_meta(np.array([o["rows"][-1] for o in outs]).reshape(-1, 20), nrows, [o["stopcond"] for o in outs],
                               [o["nrejected"] + o["ncone"] for o in outs], np.array([o["dir0"] for o in outs]).reshape(-1, 3),
                               np.zeros(len(outs)), first, nrows)
        return dict(meta=meta, segments=[(name, total, None if use_shm else buf)], totals=(int((nrows - 1).sum()), 0, total, 0), keep=_FakeKeep(shm))

    return trace


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
    pos0, dir0, w0 = common.c2_grid(3, 3)
    res = shard.run_sharded(_oracle_tracer(model, P, True, "r%d" % rank), pos0, dir0, w0, rank, world)
    if rank == 0:
        assert isinstance(res, shard.ShardedResult) and len(res) == len(w0)
        assert set(res.owner.tolist()) == {0, 1}
        np.savez(out_path, n=len(res), **{"rows_%d" % i: np.array(res.rows(i)) for i in range(len(res))}, stop=res.stopcond, last=res.last_row,
                 dir0=res.dir0, accepted=res.accepted_steps)
    else:
        assert not isinstance(res, shard.ShardedResult)
    res.close()  # collective: the owners unlink their segments only after rank 0 dropped its maps
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
    acc = 0
    for i in range(len(w0)):
        o = model.trace(pos0[i], [0, 0, 0], w0[i], **dict(common.C2, maxsteps=40))
        assert np.array_equal(z["rows_%d" % i], o["rows"], equal_nan=True) and z["stop"][i] == o["stopcond"]
        assert np.array_equal(z["last"][i], o["rows"][-1], equal_nan=True) and np.array_equal(z["dir0"][i], o["dir0"])
        acc += len(o["rows"]) - 1
    assert int(z["accepted"]) == acc
    assert not [f for f in os.listdir("/dev/shm") if f.startswith("srt_test_")]  # segments released


def test_single_rank_path_needs_no_shared_memory(built):
    import oracle_lib as ol
    model = ol.OracleModel("ngo", common.NGO_CFG)
    P = dict(common.C2, maxsteps=30)
    pos0, dir0, w0 = common.c2_grid(2, 3)
    res = shard.run_sharded(_oracle_tracer(model, P, False, "solo"), pos0, dir0, w0, 0, 1)
    for i in range(len(w0)):
        o = model.trace(pos0[i], [0, 0, 0], w0[i], **P)
        assert np.array_equal(res.rows(i), o["rows"], equal_nan=True)
    res.close()
