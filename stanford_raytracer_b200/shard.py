"""Multi-GPU: rays are independent, so the batch shards by ray index across the ranks of one box
(one process per GPU) and the only cross-rank step is a host-side gather of the
per-ray results on rank 0, which then emits them in ray order like the reference's driver loop
(fortran/raytracer_driver.f95:1143,1230).  No collective on the data path, no NCCL traffic:
`gather_object` below moves host (numpy) buffers through whatever process group the launcher
created (NCCL ranks use a side gloo group; CPU tests run it with world_size 2 over gloo).
"""
import numpy as np


def shard_indices(nrays, rank, world):
    """Ray i belongs to rank i mod world (interleaved: neighbouring launch points, which have similar
    step counts, spread over all GPUs)."""
    return np.arange(rank, nrays, world, dtype=np.int64)


def merge_shards(nrays, shards):
    """shards: list over ranks of (indices, per_ray_list).  Returns the per-ray list in ray order."""
    out = [None] * nrays
    seen = 0
    for idx, items in shards:
        if len(idx) != len(items):
            raise ValueError("shard size mismatch")
        for i, it in zip(idx, items):
            if out[int(i)] is not None:
                raise ValueError("ray %d traced twice" % int(i))
            out[int(i)] = it
            seen += 1
    if seen != nrays:
        raise ValueError("gather incomplete: %d of %d rays" % (seen, nrays))
    return out


def run_sharded(trace_fn, pos0, dir0, w0, rank=0, world=1, group=None):
    """trace_fn(pos0_shard, dir0_shard, w0_shard) -> list of per-ray results (any picklable object, e.g.
    (rows, stopcond, dir0)).  Every rank traces its shard; rank 0 returns the merged list in ray
    order, the other ranks return None."""
    pos0 = np.asarray(pos0, dtype=np.float64).reshape(-1, 3)
    dir0 = np.asarray(dir0, dtype=np.float64).reshape(-1, 3)
    w0 = np.asarray(w0, dtype=np.float64).reshape(-1)
    idx = shard_indices(len(w0), rank, world)
    mine = trace_fn(pos0[idx], dir0[idx].copy(), w0[idx])
    if world == 1:
        return merge_shards(len(w0), [(idx, mine)])
    import torch.distributed as dist
    gathered = [None] * world if rank == 0 else None
    dist.gather_object((idx, mine), gathered, dst=0, group=group)
    if rank != 0:
        return None
    return merge_shards(len(w0), gathered)


def raytracer_run_sharded(model, pos0, dir0, w0, rank=0, world=1, group=None, **run_args):
    """raytracer_run for a batch sharded over `world` ranks (each rank owns one GPU and its own copy of
    the model).  Rank 0 gets [(rows[n][20], stopcond, nsteps, dir0[3])] in ray order."""
    from . import raytracer_run

    def trace(p, d, w):
        res = raytracer_run(model, p, d, w, **run_args)
        out = [(res.rows(r), int(res.stopcond[r]), int(res.nsteps[r]), res.dir0[r].copy()) for r in range(len(res))]
        res.close()
        return out

    return run_sharded(trace, pos0, dir0, w0, rank, world, group)
