"""Multi-GPU, one process per GPU: rays are independent, so ONE launch set shards by ray index across the ranks of a box
 and the only cross-rank step is a host-side gather on rank 0, which then has every ray in input order
like the reference's driver loop (fortran/raytracer_driver.f95:1143,1230).  No collective on the data path, no NCCL
traffic.

What moves at the gather:
  * per ray, a fixed-size record (last row, nsteps, stopcond, rejected attempts, dir0, location of its rows): one
    contiguous float64 array per rank through `torch.distributed.gather` on a gloo group -- 256 bytes per ray;
  * the trajectory rows themselves do NOT move: every rank lets libsrt place its row buffers in POSIX shared memory
    (srt_set_rows_shm: the D2H copy lands there directly) and rank 0 maps the segments.  `ShardedResult.rows(ray)` is a
    zero-copy view into the owning rank's segment; `close()` is collective and releases the segments.

(In one process the library itself spreads a batch over the devices of the box: srt_init_devices + raytracer_run.)
"""
import os

import numpy as np

META_WORDS = 32  # last row (20), nsteps, stopcond, rejected, dir0 (3), chunk, first row, rows held, 3 spare
_SHM_DIR = "/dev/shm"


def shard_of(i, world):
    """Rank of ray i: (i + i // world) mod world.  Every run of `world` consecutive rays is dealt one to each rank,
    and the deal rotates from run to run, so a lattice whose fastest index has `world` values (e.g. 8 launch
    longitudes on 8 GPUs) is still mixed over the ranks instead of giving every rank one longitude."""
    i = np.asarray(i, dtype=np.int64)
    return (i + i // world) % world


def shard_indices(nrays, rank, world):
    """The rays of `rank`, ascending.  Shard sizes differ by at most one."""
    i = np.arange(nrays, dtype=np.int64)
    return i[shard_of(i, world) == rank]


def pack_meta(last_row, nsteps, stopcond, nrejected, dir0, chunk, first, nrows):
    n = len(nsteps)
    m = np.zeros((n, META_WORDS))
    m[:, 0:20] = last_row
    m[:, 20], m[:, 21], m[:, 22] = nsteps, stopcond, nrejected
    m[:, 23:26] = dir0
    m[:, 26], m[:, 27], m[:, 28] = chunk, first, nrows
    return m


class ShardedResult:
    """Rank 0's view of a sharded batch, rays in input order."""

    def __init__(self, nrays, world):
        self.nrays, self.world = nrays, world
        self.last_row = np.zeros((nrays, 20))
        self.nsteps = np.zeros(nrays, dtype=np.int64)
        self.stopcond = np.zeros(nrays, dtype=np.int32)
        self.nrejected = np.zeros(nrays, dtype=np.int64)
        self.dir0 = np.zeros((nrays, 3))
        self.owner = np.zeros(nrays, dtype=np.int32)
        self._chunk = np.zeros(nrays, dtype=np.int32)
        self._first = np.zeros(nrays, dtype=np.int64)
        self._nrows = np.zeros(nrays, dtype=np.int64)
        self._segments = {}  # (rank, chunk) -> ndarray [rows][20]
        self._names = {}     # (rank, chunk) -> (shm name, rows)
        self.accepted_steps = self.attempts = self.total_rows = self.fastforwarded_steps = 0
        self._closer = None

    def __len__(self):
        return self.nrays

    def _segment(self, rank, chunk):
        key = (rank, chunk)
        if key not in self._segments:
            name, rows = self._names[key]
            self._segments[key] = np.memmap(os.path.join(_SHM_DIR, name.lstrip("/")), dtype=np.float64, mode="r", shape=(rows, 20))
        return self._segments[key]

    def rows(self, ray):
        """Trajectory rows of `ray`, [n][20], a view into the owning rank's row buffer."""
        n = int(self._nrows[ray])
        if n == 0:
            return np.empty((0, 20))
        seg = self._segment(int(self.owner[ray]), int(self._chunk[ray]))
        f = int(self._first[ray])
        return seg[f:f + n]

    def close(self):
        self._segments.clear()
        if self._closer:
            self._closer()
            self._closer = None


class _RankHandle:
    """What the ranks other than 0 hold: their rows stay alive until the collective close()."""

    def __init__(self, closer):
        self._closer = closer

    def close(self):
        if self._closer:
            self._closer()
            self._closer = None


_gloo_groups = {}


def host_group(world):
    """A gloo group over all ranks for the host-side gather (NCCL groups cannot move host buffers)."""
    import torch.distributed as dist
    if dist.get_backend() == "gloo":
        return None
    if world not in _gloo_groups:
        _gloo_groups[world] = dist.new_group(backend="gloo")
    return _gloo_groups[world]


def run_sharded(trace_fn, pos0, dir0, w0, rank=0, world=1, group=None):
    """trace_fn(pos0_shard, dir0_shard, w0_shard) -> dict(meta=[n][META_WORDS] (pack_meta), segments=[(name, rows,
    array-or-None)], totals=(accepted, attempts, total_rows, fastforwarded), keep=object that owns the rows, with
    .close()).  Every rank traces its shard; rank 0 returns a ShardedResult, the others a handle; close() on all."""
    pos0 = np.asarray(pos0, dtype=np.float64).reshape(-1, 3)
    dir0 = np.asarray(dir0, dtype=np.float64).reshape(-1, 3)
    w0 = np.asarray(w0, dtype=np.float64).reshape(-1)
    nrays = len(w0)
    idx = shard_indices(nrays, rank, world)
    err, mine = None, None
    try:
        mine = trace_fn(pos0[idx], dir0[idx].copy(), w0[idx])
    except Exception as e:  # the gather below is collective: a rank that failed must not leave the others waiting
        err = e
    if world > 1:
        import torch
        import torch.distributed as dist
        okay = torch.tensor([0 if err else 1], dtype=torch.int32)
        dist.all_reduce(okay, op=dist.ReduceOp.MIN, group=group)
        if int(okay.item()) == 0:
            if mine is not None:
                mine["keep"].close()
            raise err if err else RuntimeError("the sharded trace failed on another rank")
    elif err:
        raise err
    if world == 1:
        metas, names, totals = [mine["meta"]], [[(s[0], s[1]) for s in mine["segments"]]], [mine["totals"]]
    else:
        import torch
        import torch.distributed as dist
        nmax = (nrays + world - 1) // world
        send = torch.zeros((nmax, META_WORDS), dtype=torch.float64)
        send[:len(idx)] = torch.from_numpy(np.ascontiguousarray(mine["meta"]))
        recv = [torch.zeros((nmax, META_WORDS), dtype=torch.float64) for _ in range(world)] if rank == 0 else None
        dist.gather(send, recv, dst=0, group=group)
        small = [None] * world if rank == 0 else None
        dist.gather_object(([(s[0], s[1]) for s in mine["segments"]], tuple(int(v) for v in mine["totals"])), small, dst=0, group=group)
        if rank == 0:
            metas = [r.numpy() for r in recv]
            names = [s[0] for s in small]
            totals = [s[1] for s in small]

    def closer():
        if world > 1:
            import torch.distributed as dist
            dist.barrier(group=group)  # rank 0 has dropped its maps before any owner frees a segment
        mine["keep"].close()

    if rank != 0:
        return _RankHandle(closer)
    out = ShardedResult(nrays, world)
    seen = np.zeros(nrays, dtype=bool)
    for r in range(world):
        ridx = shard_indices(nrays, r, world)
        m = metas[r][:len(ridx)]
        if seen[ridx].any():
            raise ValueError("a ray was traced twice")
        seen[ridx] = True
        out.last_row[ridx] = m[:, 0:20]
        out.nsteps[ridx] = m[:, 20].astype(np.int64)
        out.stopcond[ridx] = m[:, 21].astype(np.int32)
        out.nrejected[ridx] = m[:, 22].astype(np.int64)
        out.dir0[ridx] = m[:, 23:26]
        out.owner[ridx] = r
        out._chunk[ridx] = m[:, 26].astype(np.int32)
        out._first[ridx] = m[:, 27].astype(np.int64)
        out._nrows[ridx] = m[:, 28].astype(np.int64)
        for c, (name, rows) in enumerate(names[r]):
            out._names[(r, c)] = (name, rows)
        out.accepted_steps += totals[r][0]
        out.attempts += totals[r][1]
        out.total_rows += totals[r][2]
        out.fastforwarded_steps += totals[r][3]
    if not seen.all():
        raise ValueError("gather incomplete: %d of %d rays" % (int(seen.sum()), nrays))
    if world == 1:  # no shared memory needed: the views come from the local row buffers
        for c, s in enumerate(mine["segments"]):
            out._segments[(0, c)] = s[2]
    out._closer = closer
    return out


def raytracer_run_sharded(model, pos0, dir0, w0, rank=0, world=1, group=None, **run_args):
    """raytracer_run for ONE launch set sharded over `world` ranks (each rank owns one GPU and its own copy of the
    model).  Collective.  Rank 0 gets a ShardedResult (rays in input order), the other ranks a handle; every rank
    calls .close() when rank 0 is done with the rows."""
    from . import raytracer_run, set_rows_shm, lib
    import ctypes as C

    if world > 1:
        set_rows_shm("srt_rows")

    def trace(p, d, w):
        res = raytracer_run(model, p, d, w, **run_args)
        chunks, chunk, first, nrows = res.chunks()
        segs = []
        for name, n, addr in chunks:
            arr = None
            if world == 1:
                arr = np.ctypeslib.as_array(C.cast(addr, C.POINTER(C.c_double)), shape=(max(n, 1) * 20,))[:n * 20].reshape(n, 20)
            segs.append((name, n, arr))
        meta = pack_meta(res.last_row, res.nsteps, res.stopcond, res.nrejected, res.dir0, chunk, first, nrows)
        return dict(meta=meta, segments=segs, totals=(res.accepted_steps, res.attempts, res.total_rows, res.fastforwarded_steps), keep=res)

    return run_sharded(trace, pos0, dir0, w0, rank, world, group)
