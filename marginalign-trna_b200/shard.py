"""Sharding of guide alignments across the GPUs of one box.

The reference's only parallelism is a process farm over reads (jobTree workers running one cPecanRealign per read,
src/margin/utils.py:127-180; one per alignment chunk per EM iteration, submodules/cPecan/cPecanEm.py:166-171) whose
"reduction" is Python summing `%f` text files (cPecanEm.py:184-188).  Here every rank (one process per GPU,
torch.distributed) owns a shard of the alignments:

  * posterior mode has NO collective: alignments are independent work items; results are gathered to the caller's
    rank in input order (the reference consumes cigars positionally, marginAlignLib.py:334);
  * expectation mode has exactly one: an all-reduce(sum) of the S*S + S*16 + 1 doubles per EM iteration.

This module only moves indices and small vectors; the arithmetic is PairHmmEngine's (CUDA, no CPU fallback).
"""
import numpy as np

from .pairhmm import AlignmentParameters, PosteriorPairs


def estimate_cells(batch, params):
    """Per-alignment DP cell estimate used for load balancing: diagonals x (expansion + band slack).  The exact
    count comes from the plan; balancing only needs proportionality."""
    lx = batch.off_x[1:] - batch.off_x[:-1]
    ly = batch.off_y[1:] - batch.off_y[:-1]
    return (lx + ly + 1).astype(np.int64) * (int(params.diagonal_expansion) + 3)


def partition(weights, world):
    """Deterministic balanced partition of items into `world` shards.

    Items are sorted by weight (descending, stable) and dealt in boustrophedon order (0..w-1, w-1..0, ...), which keeps
    the shard sums within one max-weight of each other for sorted input and is O(n log n) vectorised -- 1 M reads
    partition in well under a second.  Each shard's indices are returned ascending so a shard preserves input order.
    """
    weights = np.asarray(weights)
    n = len(weights)
    if world < 1:
        raise ValueError("world must be >= 1")
    order = np.argsort(-weights, kind="stable")
    pos = np.arange(n)
    lap, col = pos // world, pos % world
    owner_sorted = np.where(lap % 2 == 0, col, world - 1 - col)
    owner = np.empty(n, dtype=np.int64)
    owner[order] = owner_sorted
    return [np.flatnonzero(owner == r) for r in range(world)]


def _dist():
    import torch.distributed as dist
    return dist


def world_info(group=None):
    dist = _dist()
    if not (dist.is_available() and dist.is_initialized()):
        return 0, 1
    return dist.get_rank(group), dist.get_world_size(group)


def allreduce_expectations(vec, device=None, group=None):
    """Sum of the expectation vector over ranks (the path's only collective).  `vec` is a numpy array or a torch
    tensor; a tensor is reduced in place on its own device (NCCL for CUDA tensors, gloo for CPU tensors)."""
    import torch
    dist = _dist()
    rank, world = world_info(group)
    if world == 1:
        return vec
    if isinstance(vec, np.ndarray):
        t = torch.from_numpy(np.ascontiguousarray(vec, dtype=np.float64))
        if device is not None:
            t = t.to(device)
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
        return t.cpu().numpy()
    dist.all_reduce(vec, op=dist.ReduceOp.SUM, group=group)
    return vec


def _gather_varlen(arr, dst, device, group):
    """Gathers 1-D numpy arrays of different lengths to rank `dst` (returns list of arrays there, None elsewhere)."""
    import torch
    dist = _dist()
    rank, world = world_info(group)
    n = torch.tensor([len(arr)], dtype=torch.int64, device=device)
    sizes = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    m = max(sizes) if sizes else 0
    pad = torch.zeros(max(m, 1), dtype=torch.from_numpy(arr[:0]).dtype, device=device)
    if len(arr):
        pad[:len(arr)] = torch.from_numpy(np.ascontiguousarray(arr)).to(device)
    out = [torch.empty_like(pad) for _ in range(world)] if rank == dst else None
    dist.gather(pad, out, dst=dst, group=group)
    if rank != dst:
        return None
    return [o[:s].cpu().numpy() for o, s in zip(out, sizes)]


def merge_pairs(parts, shards, n_total):
    """Reassembles per-shard PosteriorPairs into one CSR result in input order.  parts[r] = (offsets, x, y, p) of
    shard r, whose alignments are shards[r] (ascending input indices)."""
    counts = np.zeros(n_total, dtype=np.int64)
    for (off, _, _, _), idx in zip(parts, shards):
        counts[idx] = np.diff(off)
    offsets = np.zeros(n_total + 1, dtype=np.int64)
    np.cumsum(counts, out=offsets[1:])
    total = int(offsets[-1])
    x = np.empty(total, dtype=np.int32)
    y = np.empty(total, dtype=np.int32)
    p = np.empty(total, dtype=np.int32)
    for (off, px, py, pp), idx in zip(parts, shards):
        if len(idx) == 0:
            continue
        lens = np.diff(off)
        # destination position of every pair of this shard: start of its alignment in the merged CSR + rank inside
        dst = np.repeat(offsets[idx] - off[:-1], lens) + np.arange(off[-1])
        x[dst], y[dst], p[dst] = px, py, pp
    return PosteriorPairs(offsets, x, y, p)


class ShardedPairHmm:
    """Runs the two hot-path operations over all ranks of the default (or given) process group.

    `engine` is this rank's PairHmmEngine (one GPU per process).  With no process group initialised it degrades to
    the single-GPU engine.  `comm_device` is where the small collective buffers live: the engine's CUDA device under
    NCCL, "cpu" under gloo.
    """

    def __init__(self, engine, group=None, comm_device=None):
        self.engine = engine
        self.group = group
        self.rank, self.world = world_info(group)
        self.comm_device = comm_device if comm_device is not None else getattr(engine, "device", "cpu")
        self.last_shards = None
        self.last_phases = None

    def shards(self, batch, params):
        return partition(estimate_cells(batch, params), self.world)

    def get_aligned_pairs_using_anchors(self, model, batch, params=None, dst=0):
        """Every rank must call this with the same batch; rank `dst` receives the merged result (others None)."""
        params = params or AlignmentParameters()
        if self.world == 1:
            return self.engine.get_aligned_pairs_using_anchors(model, batch, params)
        import time
        t0 = time.perf_counter()
        shards = self.shards(batch, params)
        self.last_shards = shards
        mine = shards[self.rank]
        sub = batch.subset(mine)
        t1 = time.perf_counter()
        res = self.engine.get_aligned_pairs_using_anchors(model, sub, params)
        t2 = time.perf_counter()
        cols = []
        for a in (res.offsets, res.x, res.y, res.p):
            cols.append(_gather_varlen(np.asarray(a), dst, self.comm_device, self.group))
        t3 = time.perf_counter()
        self.last_phases = {"partition_s": t1 - t0, "compute_s": t2 - t1, "gather_s": t3 - t2, "merge_s": 0.0}
        if self.rank != dst:
            return None
        parts = [(cols[0][r], cols[1][r], cols[2][r], cols[3][r]) for r in range(self.world)]
        out = merge_pairs(parts, shards, batch.n)
        self.last_phases["merge_s"] = time.perf_counter() - t3
        return out

    def get_expectations_using_anchors(self, model, batch, params=None, pseudo_count=1e-12):
        """All ranks receive the same summed (transitions, emissions, likelihood).  The pseudo-count is added once,
        as hmm_constructEmpty(1e-12) does for one cPecanRealign process (cPecanRealign.c:531)."""
        params = params or AlignmentParameters()
        if self.world == 1:
            return self.engine.get_expectations_using_anchors(model, batch, params, pseudo_count=pseudo_count)
        shards = self.shards(batch, params)
        self.last_shards = shards
        mine = shards[self.rank]
        T, E, lik = self.engine.get_expectations_using_anchors(model, batch.subset(mine), params, pseudo_count=0.0)
        vec = allreduce_expectations(np.concatenate([T, E, [lik]]), self.comm_device, self.group)
        nT, nE = len(T), len(E)
        return vec[:nT] + pseudo_count, vec[nT:nT + nE] + pseudo_count, float(vec[-1])
