"""Sharding of guide alignments across the GPUs of one box.

The reference's only parallelism is a process farm over reads (jobTree workers running one cPecanRealign per read,
src/margin/utils.py:127-180; one per alignment chunk per EM iteration, submodules/cPecan/cPecanEm.py:166-171) whose
"reduction" is Python summing `%f` text files (cPecanEm.py:184-188).  Here every rank (one process per GPU,
torch.distributed) owns a shard of the alignments:

  * posterior mode has NO collective: alignments are independent work items; shards are contiguous slices balanced by
    estimated cells, results reach the caller's rank in input order (the reference consumes cigars positionally,
    marginAlignLib.py:334) through shared memory on one node (a padded gather over the backend otherwise);
  * expectation mode has exactly one: an all-reduce(sum) of the S*S + S*16 + 1 doubles per EM iteration.

This module only moves indices and small vectors; the arithmetic is PairHmmEngine's (CUDA, no CPU fallback).
"""
import inspect

import numpy as np

from .pairhmm import AlignmentParameters, PosteriorPairs


def estimate_cells(batch, params):
    """Per-alignment DP cell estimate used for load balancing: diagonals x (expansion + band slack).  The exact
    count comes from the plan; balancing only needs proportionality."""
    lx = batch.off_x[1:] - batch.off_x[:-1]
    ly = batch.off_y[1:] - batch.off_y[:-1]
    return (lx + ly + 1).astype(np.int64) * (int(params.diagonal_expansion) + 3)


def partition_bounds(weights, world):
    """Contiguous balanced partition: world + 1 non-decreasing boundaries; shard r owns items bounds[r]:bounds[r+1].

    Item i goes to the shard that holds the START of its weight interval in the running sum, so every shard's total is
    within one item weight of total / world, a shard is a slice of the input (no copy of the sequence data) and the
    results of the shards concatenate back in input order -- which is the order the reference consumes them in
    (marginAlignLib.py:334)."""
    weights = np.asarray(weights, dtype=np.float64)
    n = len(weights)
    if world < 1:
        raise ValueError("world must be >= 1")
    if n == 0:
        return np.zeros(world + 1, dtype=np.int64)
    start = np.concatenate([[0.0], np.cumsum(weights)[:-1]])
    total = float(start[-1] + weights[-1])
    if total <= 0:
        owner = (np.arange(n) * world) // n
    else:
        owner = np.minimum((start * world / total).astype(np.int64), world - 1)
    return np.searchsorted(owner, np.arange(world + 1), side="left").astype(np.int64)


def partition(weights, world):
    """The shards of partition_bounds as index arrays (ascending, contiguous)."""
    b = partition_bounds(weights, world)
    return [np.arange(b[r], b[r + 1], dtype=np.int64) for r in range(world)]


def _dist():
    import torch.distributed as dist
    return dist


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def bind_rank_to_local_cores(device_index, local_rank, local_world):
    """Pins the calling process (one rank per GPU) to its share of the host cores next to its GPU: the cores of the
    GPU's NUMA node (sysfs local_cpulist of its PCI device) split among the ranks whose GPUs sit on that node, or --
    without that information -- an even slice of the allowed cores.  Must run before the engine allocates pinned host
    memory, so that first-touch places the staging and result buffers on the GPU's own node.  Returns a description
    (for the bench line); never fails."""
    import os
    try:
        allowed = sorted(os.sched_getaffinity(0))
    except (AttributeError, OSError):
        return None
    info = {"allowed": len(allowed), "numa": None}
    mine = None
    try:
        import torch
        groups = {}  # local cpu list -> ranks (device indices) that share it
        for dev in range(local_world):
            bus = torch.cuda.get_device_properties(dev).pci_bus_id if hasattr(
                torch.cuda.get_device_properties(dev), "pci_bus_id") else None
            if bus is None:
                raise OSError("no pci bus id")
            path = "/sys/bus/pci/devices/%04x:%02x:%02x.0/local_cpulist" % (
                torch.cuda.get_device_properties(dev).pci_domain_id, bus, torch.cuda.get_device_properties(dev).pci_device_id)
            with open(path) as fh:
                cpus = tuple(sorted(_parse_cpulist(fh.read()) & set(allowed)))
            groups.setdefault(cpus, []).append(dev)
        for cpus, devs in groups.items():
            if device_index in devs and cpus:
                per = max(1, len(cpus) // len(devs))
                k = devs.index(device_index)
                mine = list(cpus[k * per:(k + 1) * per]) or list(cpus)
                info["numa"] = "%d local cores shared by %d GPUs" % (len(cpus), len(devs))
    except (OSError, AttributeError, RuntimeError, ValueError):
        mine = None
    if not mine:
        per = max(1, len(allowed) // max(1, local_world))
        mine = allowed[local_rank * per:(local_rank + 1) * per] or allowed
    try:
        os.sched_setaffinity(0, mine)
    except OSError:
        return info
    info["cores"] = len(mine)
    return info


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
    shard r, whose alignments are shards[r] (ascending input indices; contiguous shards concatenate without a scatter,
    each shard staying a block of the result)."""
    counts = np.zeros(n_total, dtype=np.int64)
    for (off, _, _, _), idx in zip(parts, shards):
        counts[idx] = np.diff(off)
    offsets = np.zeros(n_total + 1, dtype=np.int64)
    np.cumsum(counts, out=offsets[1:])
    contiguous = all(len(idx) == 0 or (idx[-1] - idx[0] + 1 == len(idx)) for idx in shards)
    if contiguous:
        blocks = [(int(idx[0]), int(idx[-1]) + 1, int(offsets[idx[0]]), px, py, pp)
                  for (off, px, py, pp), idx in zip(parts, shards) if len(idx)]
        blocks.sort(key=lambda b: b[0])
        if not blocks:
            blocks = [(0, 0, 0, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32))]
        return PosteriorPairs(offsets, blocks=blocks)
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


_SHM_DIR = "/dev/shm"
_shm_counter = [0]


def _gather_shared_memory(arrays, dst, group):
    """Single-node transport of a rank's result arrays to rank `dst` through a tmpfs file: every other rank writes its
    arrays into one file under /dev/shm, the paths travel in a tiny gather_object, `dst` maps the files (np.memmap) and
    returns views; the files are unlinked at once and their pages live as long as the returned arrays.  Returns None
    when /dev/shm cannot be used (the caller falls back to the collective)."""
    import os
    dist = _dist()
    rank, world = world_info(group)
    meta, path, ok = None, None, 1
    if rank != dst:
        try:
            _shm_counter[0] += 1
            path = os.path.join(_SHM_DIR, "mab_%d_%d.bin" % (os.getpid(), _shm_counter[0]))
            at, layout = 0, []
            with open(path, "wb") as fh:
                for a in arrays:
                    a = np.ascontiguousarray(a)
                    fh.write(memoryview(a).cast("B"))
                    layout.append((str(a.dtype), int(a.size), at))
                    at += int(a.nbytes)
            meta = (path, layout)
        except OSError:
            ok = 0
    flags = [None] * world
    dist.all_gather_object(flags, ok, group=group)
    if not all(flags):
        if path is not None and os.path.exists(path):
            os.remove(path)
        return None
    metas = [None] * world if rank == dst else None
    dist.gather_object(meta, metas, dst=dst, group=group)
    out = ()  # other ranks: done (an empty tuple, not None: None means "shared memory unusable, fall back")
    if rank == dst:
        out = []
        for r in range(world):
            if r == dst:
                out.append([np.asarray(a) for a in arrays])
                continue
            name, layout = metas[r]
            out.append([np.memmap(name, dtype=np.dtype(dt), mode="r", offset=at, shape=(n,)) if n else
                        np.zeros(0, dtype=np.dtype(dt)) for dt, n, at in layout])
        out = (out, [])
    dist.barrier(group=group)  # dst has mapped every file: the owners can drop the names
    if path is not None:
        try:
            os.remove(path)
        except FileNotFoundError:
            pass
    return out


class ShardedPairHmm:
    """Runs the two hot-path operations over all ranks of the default (or given) process group.

    `engine` is this rank's PairHmmEngine (one GPU per process).  With no process group initialised it degrades to
    the single-GPU engine.  `comm_device` is where the small collective buffers live: the engine's CUDA device under
    NCCL, "cpu" under gloo.
    """

    def __init__(self, engine, group=None, comm_device=None, shared_memory=True):
        self.engine = engine
        self.shared_memory = shared_memory  # single-node result transport through /dev/shm (falls back if unusable)
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
        sub = batch.slice(mine[0], mine[-1] + 1) if len(mine) else batch.slice(0, 0)  # shards are contiguous: views
        t1 = time.perf_counter()
        res = self.engine.get_aligned_pairs_using_anchors(model, sub, params)
        t2 = time.perf_counter()
        arrays = (res.offsets, res.x, res.y, res.p)
        got = _gather_shared_memory(arrays, dst, self.group) if self.shared_memory else None
        transport = "shared memory"
        if got is None:  # no usable /dev/shm (or more than one node): padded gather over the collective backend
            transport = "collective gather"
            cols = [_gather_varlen(np.asarray(a), dst, self.comm_device, self.group) for a in arrays]
            got = ([[cols[k][r] for k in range(4)] for r in range(self.world)], []) if self.rank == dst else None
        t3 = time.perf_counter()
        self.last_phases = {"partition_s": t1 - t0, "compute_s": t2 - t1, "gather_s": t3 - t2, "merge_s": 0.0,
                            "transport": transport}
        if self.rank != dst:
            return None
        parts, keep = got
        out = merge_pairs([tuple(p) for p in parts], shards, batch.n)
        out._keep = keep  # shared-memory segments behind the blocks of the other ranks
        self.last_phases["merge_s"] = time.perf_counter() - t3
        return out

    def realign(self, model, batch, params=None, gap_gamma=0.5, match_gamma=0.0, dst=0):
        """PairHmmEngine.realign over all ranks: every rank realigns its contiguous shard on its GPU (posteriors, AMAP
        reweight, MEA chain, cigar), only the cigars travel to rank `dst` -- a few bytes per read instead of the pairs --
        and are concatenated in input order, the order marginAlignLib.py:334 consumes.  Others return None."""
        from .pairhmm import RealignedCigars
        params = params or AlignmentParameters()
        if self.world == 1:
            return self.engine.realign(model, batch, params, gap_gamma=gap_gamma, match_gamma=match_gamma)
        import time
        t0 = time.perf_counter()
        shards = self.shards(batch, params)
        self.last_shards = shards
        mine = shards[self.rank]
        sub = batch.slice(mine[0], mine[-1] + 1) if len(mine) else batch.slice(0, 0)
        t1 = time.perf_counter()
        res = self.engine.realign(model, sub, params, gap_gamma=gap_gamma, match_gamma=match_gamma)
        t2 = time.perf_counter()
        arrays = (np.asarray(res.off_ops, dtype=np.int64), np.asarray(res.ops, dtype=np.uint8),
                  np.asarray(res.lens, dtype=np.int32))
        got = _gather_shared_memory(arrays, dst, self.group) if self.shared_memory else None
        transport = "shared memory"
        if got is None:
            transport = "collective gather"
            cols = [_gather_varlen(a, dst, self.comm_device, self.group) for a in arrays]
            got = ([[cols[k][r] for k in range(3)] for r in range(self.world)], []) if self.rank == dst else None
        t3 = time.perf_counter()
        self.last_phases = {"partition_s": t1 - t0, "compute_s": t2 - t1, "gather_s": t3 - t2, "merge_s": 0.0,
                            "transport": transport}
        if self.rank != dst:
            return None
        parts, _keep = got
        off, base = [np.zeros(1, dtype=np.int64)], 0
        for p in parts:  # shards are contiguous and in rank order: concatenation restores the input order
            off.append(np.asarray(p[0][1:]) + base)
            base += int(p[0][-1]) if len(p[0]) else 0
        out = RealignedCigars(np.concatenate(off), np.concatenate([np.asarray(p[1]) for p in parts]),
                              np.concatenate([np.asarray(p[2]) for p in parts]))
        self.last_phases["merge_s"] = time.perf_counter() - t3
        return out

    def get_expectations_using_anchors(self, model, batch, params=None, pseudo_count=1e-12, plan_cache=None):
        """All ranks receive the same summed (transitions, emissions, likelihood).  The pseudo-count is added once,
        as hmm_constructEmpty(1e-12) does for one cPecanRealign process (cPecanRealign.c:531).  plan_cache: a
        pairhmm.PlanCache; every rank keeps the device plans of its own shard between calls."""
        params = params or AlignmentParameters()
        extra = {}
        if plan_cache is not None and "plan_cache" in inspect.signature(
                self.engine.get_expectations_using_anchors).parameters:
            extra["plan_cache"] = plan_cache
        if self.world == 1:
            return self.engine.get_expectations_using_anchors(model, batch, params, pseudo_count=pseudo_count, **extra)
        sub = None
        if plan_cache is not None:  # the shard object must be the same one for the cache to recognise it
            if getattr(plan_cache, "_shard_of", None) is batch:
                sub = plan_cache._shard
        if sub is None:
            shards = self.shards(batch, params)
            self.last_shards = shards
            mine = shards[self.rank]
            sub = batch.slice(mine[0], mine[-1] + 1) if len(mine) else batch.slice(0, 0)
            if plan_cache is not None:
                plan_cache._shard_of, plan_cache._shard = batch, sub
        T, E, lik = self.engine.get_expectations_using_anchors(model, sub, params, pseudo_count=0.0, **extra)
        vec = allreduce_expectations(np.concatenate([T, E, [lik]]), self.comm_device, self.group)
        nT, nE = len(T), len(E)
        return vec[:nT] + pseudo_count, vec[nT:nT + nE] + pseudo_count, float(vec[-1])
