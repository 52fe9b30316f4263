"""Host side of the B200 pair-HMM path: models, parameters, batch planning and the two public operations.

These are the in-process replacements for what the reference does by spawning `cPecanRealign` once per read:
  get_aligned_pairs_using_anchors   <- getAlignedPairsUsingAnchors   (cPecan/impl/pairwiseAligner.c:1172)
  get_expectations_using_anchors    <- getExpectationsUsingAnchors   (cPecan/impl/pairwiseAligner.c:1201)
applied to a whole batch of guide alignments at once.  PyTorch is used only for device/pinned buffers and streams;
all arithmetic happens in libmargin_b200.so (CUDA, sm_100a).  There is no CPU fallback: without a CUDA device the
operations raise RuntimeError.
"""
import ctypes as C
import os
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from dataclasses import dataclass

import numpy as np

from . import capi
from .hmm import Hmm
from .synth import AlignmentBatch


@dataclass
class AlignmentParameters:
    """PairwiseAlignmentParameters (cPecan/inc/pairwiseAligner.h:28-39) as set by cPecanRealign's flags.

    split_matrix_bigger_than is the *flag value* (--splitMatrixBiggerThanThis); it is squared on the way in exactly
    as cPecanRealign.c:459 does.  Defaults are the realign call site's (src/margin/marginAlignLib.py:315-318).
    """
    threshold: float = 0.01
    min_diags_between_traceback: int = 1000
    traceback_diagonals: int = 40
    diagonal_expansion: int = 10
    split_matrix_bigger_than: int = 3000
    ragged_left: bool = True
    ragged_right: bool = True

    def to_c(self):
        return capi.MabParams(self.threshold, self.min_diags_between_traceback, self.traceback_diagonals,
                              self.diagonal_expansion, int(self.split_matrix_bigger_than) ** 2,
                              int(self.ragged_left), int(self.ragged_right))


class PairHmmModel:
    """A loaded state machine (replaces hmm_loadFromFile + hmm_getStateMachine / stateMachine5_construct)."""

    def __init__(self, handle):
        self._h = handle
        L = capi.load()
        self.state_number = L.mab_model_state_number(handle)
        self.type = L.mab_model_type(handle)

    @classmethod
    def from_hmm(cls, hmm: Hmm):
        return cls.from_arrays(hmm._modelTypeInt(), hmm.transitions, hmm.emissions)

    @classmethod
    def from_arrays(cls, hmm_type, transitions, emissions):
        L = capi.load()
        S = 5 if hmm_type <= 1 else 3
        T = np.ascontiguousarray(transitions, dtype=np.float64)
        E = np.ascontiguousarray(emissions, dtype=np.float64)
        if T.size != S * S or E.size != S * 16:
            raise RuntimeError("hmm type %d needs %d transitions and %d emissions" % (hmm_type, S * S, S * 16))
        h = capi.vp()
        capi.check(L.mab_model_from_hmm(int(hmm_type), T.ctypes.data_as(capi.dp), E.ctypes.data_as(capi.dp),
                                        C.byref(h)))
        m = cls(h)
        m.src_transitions, m.src_emissions = T.copy(), E.copy()  # the `.hmm` numbers this machine was built from
        return m

    @classmethod
    def from_file(cls, path):
        return cls.from_hmm(Hmm.loadHmm(path))

    @classmethod
    def default(cls, hmm_type=0):
        """The built-in machine cPecanRealign uses without --loadHmm (cPecanRealign.c:526): fiveState."""
        L = capi.load()
        h = capi.vp()
        capi.check(L.mab_model_default(int(hmm_type), C.byref(h)))
        return cls(h)

    def transitions(self):
        L = capi.load()
        out = np.zeros((13, 4))
        n = L.mab_model_transitions(self._h, out.ctypes.data_as(capi.dp), 13)
        return out[:n]

    def boundary(self, which):
        out = np.zeros(self.state_number)
        capi.check(capi.load().mab_model_boundary(self._h, which, out.ctypes.data_as(capi.dp)))
        return out

    def emissions(self):
        m, gx, gy = np.zeros(25), np.zeros(5), np.zeros(5)
        capi.check(capi.load().mab_model_emissions(self._h, m.ctypes.data_as(capi.dp), gx.ctypes.data_as(capi.dp),
                                                   gy.ctypes.data_as(capi.dp)))
        return m, gx, gy

    @property
    def n_expectations(self):
        S = self.state_number
        return S * S + S * 16 + 1

    def __del__(self):
        try:
            if self._h:
                capi.load().mab_model_destroy(self._h)
                self._h = None
        except Exception:
            pass


def _ptr(a, typ):
    return a.ctypes.data_as(typ)


class Plan:
    """Host work plan of one batch (regions, bands, traceback schedule); see csrc/plan.cpp."""

    def __init__(self, handle):
        self._h = handle
        self.info = capi.MabPlanInfo()
        capi.check(capi.load().mab_plan_get_info(handle, C.byref(self.info)))

    @classmethod
    def from_batch(cls, batch: AlignmentBatch, params: AlignmentParameters, n_threads=0):
        L = capi.load()
        p = params.to_c()
        h = capi.vp()
        sx = np.ascontiguousarray(batch.seq_x, dtype=np.uint8)
        sy = np.ascontiguousarray(batch.seq_y, dtype=np.uint8)
        ox = np.ascontiguousarray(batch.off_x, dtype=np.int64)
        oy = np.ascontiguousarray(batch.off_y, dtype=np.int64)
        ops = np.ascontiguousarray(batch.ops, dtype=np.uint8)
        lens = np.ascontiguousarray(batch.lens, dtype=np.int32)
        oo = np.ascontiguousarray(batch.off_ops, dtype=np.int64)
        capi.check(L.mab_plan_from_cigars(C.byref(p), batch.n, _ptr(sx, capi.u8p), _ptr(ox, capi.i64p),
                                          _ptr(sy, capi.u8p), _ptr(oy, capi.i64p), _ptr(ops, capi.u8p),
                                          _ptr(lens, capi.i32p), _ptr(oo, capi.i64p), n_threads, C.byref(h)))
        return cls(h)

    @classmethod
    def from_anchors(cls, xs, ys, anchors, params: AlignmentParameters, n_threads=0):
        """xs, ys: list of bytes; anchors: list of (k,2) int arrays (explicit anchors instead of a cigar)."""
        L = capi.load()
        b = AlignmentBatch.from_lists(xs, ys, [([], [])] * len(xs))
        anc = [np.asarray(a, dtype=np.int64).reshape(-1, 2) for a in anchors]
        off = np.zeros(len(anc) + 1, dtype=np.int64)
        np.cumsum([len(a) for a in anc], out=off[1:])
        flat = np.ascontiguousarray(np.concatenate(anc) if anc else np.zeros((0, 2), np.int64))
        p = params.to_c()
        h = capi.vp()
        capi.check(L.mab_plan_from_anchors(C.byref(p), b.n, _ptr(b.seq_x, capi.u8p), _ptr(b.off_x, capi.i64p),
                                           _ptr(b.seq_y, capi.u8p), _ptr(b.off_y, capi.i64p), _ptr(flat, capi.i64p),
                                           _ptr(off, capi.i64p), n_threads, C.byref(h)))
        return cls(h)

    # ---- introspection (tests)
    def anchors(self, aln):
        L = capi.load()
        n = L.mab_plan_anchors(self._h, aln, None, 0)
        out = np.zeros((max(n, 1), 2), dtype=np.int64)
        L.mab_plan_anchors(self._h, aln, _ptr(out, capi.i64p), n)
        return out[:n]

    def regions(self):
        out = np.zeros((max(self.info.n_regions, 1), 5), dtype=np.int64)
        n = capi.load().mab_plan_regions(self._h, _ptr(out, capi.i64p), len(out))
        return out[:n]

    def band(self, region):
        L = capi.load()
        n = L.mab_plan_band(self._h, region, None, 0)
        out = np.zeros((n, 2), dtype=np.int64)
        L.mab_plan_band(self._h, region, _ptr(out, capi.i64p), n)
        return out[:, 0].copy(), out[:, 1].copy()

    def tracebacks(self):
        out = np.zeros((max(self.info.n_tracebacks, 1), 5), dtype=np.int64)
        n = capi.load().mab_plan_tracebacks(self._h, _ptr(out, capi.i64p), len(out))
        return out[:n]

    def write_blob(self, dst_ptr, nbytes):
        capi.check(capi.load().mab_plan_write_blob(self._h, dst_ptr, nbytes))

    def __del__(self):
        try:
            if self._h:
                capi.load().mab_plan_destroy(self._h)
                self._h = None
        except Exception:
            pass


class PosteriorPairs:
    """CSR result of posterior mode: alignment i owns rows offsets[i]:offsets[i+1] of (x, y, p_int), in the order
    the reference's getAlignedPairsUsingAnchors returns them; p_int = floor(min(p,1) * 1e7).

    The engine returns one block per chunk (pinned host memory filled by D2H copies); .x / .y / .p concatenate the
    blocks on first use, .pairs(i) reads the owning block directly.  x and y are int16 arrays when every sub-sequence of
    the batch is shorter than 32768 bases (PairHmmEngine(compact_pairs=True), the default), else int32; p is int32."""

    def __init__(self, offsets, x=None, y=None, p=None, blocks=None):
        self.offsets = np.asarray(offsets, dtype=np.int64)
        if blocks is None:
            blocks = [(0, int(len(self.offsets) - 1), 0, np.asarray(x), np.asarray(y), np.asarray(p))]
        self._blocks = blocks  # (first alignment, end alignment, first pair, x, y, p)
        self._flat = None
        self._starts = np.array([b[0] for b in blocks], dtype=np.int64)

    def _cat(self):
        if self._flat is None:
            if len(self._blocks) == 1:
                self._flat = self._blocks[0][3:6]
            else:
                self._flat = tuple(np.concatenate([b[k] for b in self._blocks]) for k in (3, 4, 5))
        return self._flat

    @property
    def x(self):
        return self._cat()[0]

    @property
    def y(self):
        return self._cat()[1]

    @property
    def p(self):
        return self._cat()[2]

    @property
    def n_pairs(self):
        return int(self.offsets[-1])

    def pairs(self, i):
        b = self._blocks[int(np.searchsorted(self._starts, i, side="right")) - 1]
        s = slice(self.offsets[i] - b[2], self.offsets[i + 1] - b[2])
        return np.stack([b[5][s].astype(np.int64), b[3][s].astype(np.int64), b[4][s].astype(np.int64)], axis=1)


def _torch():
    import torch
    return torch


def require_cuda():
    torch = _torch()
    if not torch.cuda.is_available() or capi.load().mab_device_count() == 0:
        raise RuntimeError("marginalign_trna_b200: no CUDA device; the pair-HMM path has no CPU fallback")
    return torch


class DevicePlan:
    """A plan whose blob is resident in HBM, plus its scratch workspace.

    Two ways to get one: from a host-built Plan (the blob is copied over PCIe), or DevicePlan.build_on_device, which
    uploads only the raw inputs and runs the planner on the GPU (mab_plan_build_device)."""

    def __init__(self, plan: Plan, device, expectations=False, stream=None, _resident=None, _alloc_ws=True):
        torch = require_cuda()
        self.plan = plan
        self.device = device
        if _resident is None:
            nb = plan.info.blob_bytes
            self.h_blob = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
            plan.write_blob(self.h_blob.data_ptr(), nb)
            self.d_blob = torch.empty(nb, dtype=torch.uint8, device=device)
            with torch.cuda.stream(stream) if stream is not None else _null():
                self.d_blob.copy_(self.h_blob, non_blocking=True)
            self.blob_ptr = self.d_blob.data_ptr()
            self.h2d_bytes = nb
            self._inputs = None
        else:
            self.d_blob, self.blob_ptr, self.h2d_bytes, self._inputs = _resident
        ws = plan.info.workspace_bytes_expectations if expectations else plan.info.workspace_bytes_posteriors
        self.ws_bytes = int(ws)
        self.d_ws = None
        if _alloc_ws:
            self.alloc_workspace()

    def alloc_workspace(self):
        """Scratch for the DP kernels, allocated on the caller's current stream (256-byte aligned view)."""
        if self.d_ws is None:
            self.d_ws = _torch().empty(self.ws_bytes + 256, dtype=_torch().uint8, device=self.device)
            self.ws_ptr = (self.d_ws.data_ptr() + 255) & ~255

    def drop_workspace(self):
        self.d_ws, self.ws_ptr = None, None

    def use_workspace(self, buf):
        """Scratch owned by somebody else (PairHmmEngine keeps one grow-only buffer for all its chunks)."""
        assert buf.numel() >= self.ws_bytes + 256
        self.d_ws = buf
        self.ws_ptr = (buf.data_ptr() + 255) & ~255

    def adopt(self, stream, workspace=None):
        """Hands a plan built on another stream (already synchronised) over to `stream`: the caching allocator must
        not recycle the blob or the inputs while kernels of `stream` still read them.  workspace: a function
        bytes -> buffer that provides the scratch (default: an allocation of its own)."""
        for t in (self.d_blob,) + tuple(x for x in (self._inputs or ()) if hasattr(x, "record_stream")):
            if t is not None and t.is_cuda:
                t.record_stream(stream)
        if workspace is not None:
            self.use_workspace(workspace(self.ws_bytes + 256))
        else:
            self.alloc_workspace()

    @staticmethod
    def pack_inputs(batch, staging=None):
        """Packs bases + cigars of `batch` into one pinned host buffer (host-only work; safe to run on a helper
        thread while the GPU is busy).  Returns the tuple build_on_device consumes."""
        torch = _torch()
        n = batch.n
        parts = [np.ascontiguousarray(batch.off_x, dtype=np.int64), np.ascontiguousarray(batch.off_y, dtype=np.int64),
                 np.ascontiguousarray(batch.off_ops, dtype=np.int64), np.ascontiguousarray(batch.lens, dtype=np.int32),
                 np.ascontiguousarray(batch.seq_x, dtype=np.uint8), np.ascontiguousarray(batch.seq_y, dtype=np.uint8),
                 np.ascontiguousarray(batch.ops, dtype=np.uint8)]
        offs, total = [], 0
        for a in parts:
            offs.append(total)
            total = (total + a.nbytes + 255) & ~255
        total = max(total, 256)
        h_in = staging.pinned(total) if staging is not None else torch.empty(total, dtype=torch.uint8, pin_memory=True)
        hv = h_in.numpy()
        for a, o in zip(parts, offs):
            if a.nbytes:
                hv[o:o + a.nbytes] = a.view(np.uint8).reshape(-1)
        len_x, len_y = (int(batch.off_x[-1]), int(batch.off_y[-1])) if n else (0, 0)
        return h_in, offs, total, n, len_x, len_y

    @staticmethod
    def upload_inputs(packed, device):
        """H2D copy of a packed input buffer on the current stream; returns the device tensor."""
        torch = _torch()
        h_in, _offs, total = packed[0], packed[1], packed[2]
        d_in = torch.empty(total + 256, dtype=torch.uint8, device=device)
        base = (d_in.data_ptr() + 255) & ~255
        d_in[base - d_in.data_ptr():base - d_in.data_ptr() + total].copy_(h_in[:total], non_blocking=True)
        return d_in

    @classmethod
    def build_on_device(cls, batch, params, device, expectations=False, staging=None, packed=None,
                        alloc_ws=True, d_in=None):
        """Uploads bases + cigars (one pinned staging buffer, one H2D copy; skipped when the caller already did it and
        passes d_in) and plans on the GPU, on the current stream."""
        torch = require_cuda()
        L = capi.load()
        packed = packed if packed is not None else cls.pack_inputs(batch, staging)
        h_in, offs, total, n, len_x, len_y = packed
        if d_in is None:
            d_in = cls.upload_inputs(packed, device)
        base = (d_in.data_ptr() + 255) & ~255
        ptr = [base + o for o in offs]
        p = params.to_c()
        need = int(L.mab_plan_device_arena_hint(C.byref(p), n, len_x, len_y))
        stream = C.c_void_p(torch.cuda.current_stream(device).cuda_stream)
        for _ in range(3):
            arena = torch.empty(need + 256, dtype=torch.uint8, device=device)
            aptr = (arena.data_ptr() + 255) & ~255
            h, needed = capi.vp(), C.c_int64(0)
            rc = L.mab_plan_build_device(C.byref(p), n, ptr[4], ptr[0], len_x, ptr[5], ptr[1], len_y, ptr[6], ptr[3],
                                         ptr[2], None, None, aptr, need, stream, C.byref(h), C.byref(needed))
            if rc == -4 and needed.value > need:  # MAB_E_CAPACITY: retry with the exact requirement
                need = int(needed.value)
                del arena
                continue
            capi.check(rc)
            break
        else:
            raise RuntimeError("mab_plan_build_device: arena sizing did not converge")
        plan = Plan(h)
        out = cls(plan, device, expectations=expectations, _resident=(arena, aptr, total, (d_in, h_in)),
                  _alloc_ws=alloc_ws)
        out.input_offsets = (ptr[0], ptr[1], len_x, len_y)  # device addresses of off_x, off_y (int64 [n+1]) in d_in
        return out


class _Staging:
    """Grow-only pinned host buffer reused between chunks (page-locking memory is slow; do it once)."""

    def __init__(self):
        self.buf = None

    def pinned(self, nbytes):
        torch = _torch()
        if self.buf is None or self.buf.numel() < nbytes:
            self.buf = torch.empty(int(nbytes * 1.25) + 4096, dtype=torch.uint8, pin_memory=True)
        return self.buf[:nbytes]


def torch_props(torch, device):
    return torch.cuda.get_device_properties(device)


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


class PlanCache:
    """Device plans of one batch kept in HBM between calls.  The band geometry of an alignment (anchors, split regions,
    bands, checkpoint schedule: cPecanRealign recomputes all of it in every EM iteration, cPecanEm.py:139-165 ->
    cPecanRealign.c:567) does not depend on the HMM, so the EM loop plans its alignments once and every later
    iteration starts at the forward sweep: no packing, no upload, no planner kernels.  Pass the same object to
    get_expectations_using_anchors for every call on the same (batch, parameters); a cache that was filled for another
    batch, other banding parameters or the other mode is refilled.  Holds the plan blobs only (~5.5 kB per tRNA read;
    workspaces are scratch and allocated per call); `max_bytes` bounds what it keeps (default: a quarter of the HBM
    free when it is filled) -- a batch whose plans do not fit is simply planned again every call."""

    def __init__(self, max_bytes=None):
        self.max_bytes = max_bytes
        self.batch, self.key, self.plans = None, None, None
        self.hits = 0

    @staticmethod
    def _key(params, expectations, device):
        return tuple(getattr(params, f) for f in sorted(vars(params)) if not f.startswith("_")) + (
            bool(expectations), str(device))

    def lookup(self, batch, params, expectations, device):
        """The kept plans if they were made for this very batch object (the cache holds a reference to it, so the
        identity test cannot be fooled by a recycled id), these parameters, this mode and this GPU."""
        if self.plans is not None and self.batch is batch and self.key == self._key(params, expectations, device):
            self.hits += 1
            return self.plans
        return None

    def store(self, batch, params, expectations, device, plans):
        self.batch, self.key, self.plans = batch, self._key(params, expectations, device), plans

    def clear(self):
        self.batch, self.key, self.plans = None, None, None


class PairHmmEngine:
    """Runs batches on one GPU.  `max_cells` bounds the number of DP cells resident at once (28 B of HBM each in
    posterior mode, 80 B in expectation mode).  plan_on = "device" (default): raw inputs are uploaded and the plan is
    built on the GPU; "host": the C++ host planner builds the blob and it is copied over PCIe."""

    def __init__(self, device=None, max_cells=None, plan_threads=0, plan_on="device", plan_ahead=None,
                 compact_pairs=True, chunk_ramp=None):
        torch = require_cuda()
        self.torch = torch
        self.device = torch.device(device if device is not None else "cuda:%d" % torch.cuda.current_device())
        self.max_cells = int(max_cells) if max_cells else None  # None: chosen per batch (_auto_max_cells)
        self.plan_threads = plan_threads
        if plan_on not in ("device", "host"):
            raise ValueError("plan_on must be 'device' or 'host'")
        self.plan_on = plan_on
        # get_aligned_pairs_using_anchors returns x / y as int16 arrays when every sub-sequence is shorter than 32768
        # bases (same values; a third less data leaves the GPU); False keeps int32 always
        self.compact_pairs = bool(compact_pairs)
        if chunk_ramp is not None:  # divisors of the chunk size for the first chunks (default CHUNK_RAMP)
            self.CHUNK_RAMP = tuple(int(d) for d in chunk_ramp)
        if plan_ahead is not None:  # chunks planned concurrently ahead of the one computing (default PLAN_AHEAD)
            self.PLAN_AHEAD = max(1, int(plan_ahead))
        self.last_stats = {}
        self._t = {"wait_plan_s": 0.0, "plan_s": 0.0, "pack_s": 0.0, "compute_s": 0.0, "host_alloc_s": 0.0, "drain_s": 0.0}
        self._ws_buf = None
        self._pool, self._slot_free = {}, [None, None]
        self._stagings = (_Staging(), _Staging())
        self._copy_stream = torch.cuda.Stream(device=self.device)
        self._plan_streams = [torch.cuda.Stream(device=self.device)]
        self._t_lock = threading.Lock()

    def _workspace(self, nbytes):
        """The DP scratch of the chunk being computed: ONE grow-only buffer per engine.  Chunks run one after the other
        on the caller's stream, so they can share it; allocating it per chunk let the caching allocator carve smaller
        tensors (pair outputs, plan arenas) out of a freed 46 GB block and then open another 46 GB segment for the next
        chunk -- 20 000 reads of 10 kb at expansion 100 ran out of memory with 139 GB reserved and 1.4 GB in use.
        release_workspace() gives it back."""
        if self._ws_buf is None or self._ws_buf.numel() < nbytes:
            self._ws_buf = None  # never hold two: the old block is recycled in stream order
            self._ws_buf = self.torch.empty(int(nbytes), dtype=self.torch.uint8, device=self.device)
        return self._ws_buf

    def release_workspace(self):
        self._ws_buf = None
        self._pool.clear()
        self._slot_free = [None, None]

    def _pooled(self, slot, name, numel, dtype):
        """A view of `numel` elements of a grow-only device buffer owned by the engine (one per pipeline slot and
        name).  The realign path sizes its per-chunk arrays by bounds (pairs: a quarter of the band cells, tail scratch
        36 B per pair): allocated per chunk, 14 chunks of 10 kb reads cycled ~18 GB each through the caching allocator
        (different sizes every time, cudaMalloc / cudaFree with device syncs) and e2e ran at 4.7 instead of 12 GCUPS."""
        key = (slot, name)
        buf = self._pool.get(key)
        if buf is None or buf.numel() < numel or buf.dtype != dtype:
            self._pool[key] = None
            buf = self._pool[key] = self.torch.empty(int(numel * 1.1) + 64, dtype=dtype, device=self.device)
        return buf[:numel]

    # ---------------------------------------------------------------- chunking
    CHUNK_RAMP = (4,)         # the first chunk is a quarter of the others, see _chunks
    PLAN_AHEAD = 2            # chunks planned ahead of the one computing
    # chunk size for short reads (tRNA: ~285 k alignments, 22 GB of workspace).  Measured on 1 M tRNA reads (r2bi / r2bj):
    # 2.5e8 -> e2e 15.5 GCUPS, 4e8 -> 16.4, 8e8 -> 17.0, 1.2e9 -> 16.7, 1.6e9 -> 16.6 (larger chunks: fewer planner
    # round trips and kernel tails; smaller ones: more of the first upload and the last download overlapped)
    BASE_CELLS = 800_000_000

    def _auto_max_cells(self, mean_est, params, expectations):
        """Chunk size when the caller gave none.  The forward sweep of one region is sequential over its diagonals, so
        a chunk must hold enough regions to fill the GPU (~6000 resident warps / the warps one region gets, see
        wave_launch in csrc/kernels.cu): a few hundred 10 kb reads per chunk would leave the forward kernel
        latency-bound.  Bounded by free HBM (posterior mode ~28 B of workspace per cell + outputs, expectation mode
        160 B)."""
        # HBM this call can use: what the driver reports free, what the caching allocator holds without using it, and
        # the engine's own grow-only buffers, which the chunks of this call will re-use.  (Counting only the first
        # made the second call on 20 000 reads of 10 kb cut the batch into 236 chunks of 85 reads: the buffers of
        # the first call looked like somebody else's memory.)
        free, _total = self.torch.cuda.mem_get_info(self.device)
        free += self.torch.cuda.memory_reserved(self.device) - self.torch.cuda.memory_allocated(self.device)
        free += self._ws_buf.numel() if self._ws_buf is not None else 0
        free += sum(b.numel() * b.element_size() for b in self._pool.values() if b is not None)
        mem_cap = min(int(0.55 * free / (170 if expectations else 40)), 2_000_000_000)  # 32-bit side offsets
        # regions resident at once in the forward kernel: the register kernels hold 5 blocks of 128 threads per SM, a
        # half-warp per region while the band span fits 16 lanes (expansion <= 13), a warp up to expansion 29
        # (mab_band_class); a chunk slightly larger than that would run the sweep twice, the second time for a handful
        # of regions.  (The forward sweep of 10 kb reads stays bound by its 20 000 sequential diagonals, ~1.2 us each:
        # 8000 reads at expansion 10 take 70 ms forward against 36 ms backward whatever the chunking.)
        sms = torch_props(self.torch, self.device).multi_processor_count
        if params.diagonal_expansion <= 13:
            want_alignments = sms * 5 * 8
        elif params.diagonal_expansion <= 29:
            want_alignments = sms * 5 * 4
        else:
            want_alignments = 6144 // min(8, max(1, -(-(params.diagonal_expansion + 2) // 16)))
        target = int(mean_est * want_alignments)
        return max(min(max(self.BASE_CELLS, target), mem_cap), 1), want_alignments

    def _chunks(self, batch, params, expectations=False):
        """Contiguous alignment ranges of at most ~max_cells estimated DP cells (cells ~ diagonals x band width; the
        exact count comes from the plan)."""
        if batch.n == 0:
            return [(0, 0)]
        # estimate per alignment: (lX + lY + 1) (expansion + 4); its running sum telescopes over the offsets, so it is
        # built without the per-alignment array (this runs before any GPU work can start: 1 M reads, 14 -> 5 ms)
        cum = np.asarray(batch.off_x[1:], dtype=np.int64) + batch.off_y[1:]
        cum += np.arange(1, batch.n + 1, dtype=np.int64)
        cum -= int(batch.off_x[0]) + int(batch.off_y[0])
        cum *= params.diagonal_expansion + 4
        mean_est = float(cum[-1]) / batch.n
        if self.max_cells:
            max_cells, want_alignments = self.max_cells, 0
        else:
            max_cells, want_alignments = self._auto_max_cells(mean_est, params, expectations)
        bounds, start, base = [], 0, 0
        # the first chunk is a quarter of the others: nothing overlaps its packing, upload and planning, so the pipeline
        # should fill quickly -- unless such a chunk would hold too few alignments to fill the GPU.  (CHUNK_RAMP can
        # ask for a longer ramp, e.g. 8,4,2; measured equal on 1 M tRNA reads: 190 vs 186 ms.)
        ramp = [max(max_cells // d, 1) for d in self.CHUNK_RAMP]
        ramp = [r for r in ramp if not (want_alignments and r < mean_est * want_alignments)]
        quarter_ok = any(r <= max(max_cells // 4, 1) for r in ramp)
        while start < batch.n:
            limit = ramp.pop(0) if ramp else max_cells
            end = int(np.searchsorted(cum, base + limit, side="right"))
            end = max(end, start + 1)
            bounds.append((start, min(end, batch.n)))
            start = bounds[-1][1]
            base = int(cum[start - 1])
        # ... and so is the last one where results go back to the host (posterior mode): nothing overlaps the D2H copy
        # of its pairs
        if quarter_ok and not expectations and len(bounds) >= 3:
            a, b = bounds[-1]
            first_cell = int(cum[a - 1])
            if int(cum[b - 1]) - first_cell > max_cells // 2:
                mid = int(np.searchsorted(cum, int(cum[b - 1]) - max_cells // 4, side="right"))
                if a < mid < b:
                    bounds[-1:] = [(a, mid), (mid, b)]
        return bounds

    def _stream_ptr(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def _device_plans(self, batch, params, expectations, cache=None):
        """Yields one DevicePlan per chunk.  Host planning of chunk k+1 overlaps the GPU work on chunk k.  With a
        PlanCache the plans of a batch are kept and re-used by the next call (see PlanCache)."""
        if cache is None:
            yield from self._build_device_plans(batch, params, expectations)
            return
        kept = cache.lookup(batch, params, expectations, self.device)
        if kept is not None:
            for dplan in kept:
                dplan.use_workspace(self._workspace(dplan.ws_bytes + 256))
                yield dplan
                dplan.drop_workspace()  # scratch only: freed in stream order, the next chunk's takes its place
            return
        cache.clear()
        budget = cache.max_bytes
        if budget is None:
            budget = self.torch.cuda.mem_get_info(self.device)[0] // 4
        keep, used = [], 0
        for dplan in self._build_device_plans(batch, params, expectations):
            yield dplan
            if keep is not None:
                used += int(dplan.d_blob.numel())
                if used > budget:
                    keep = None  # does not fit: this batch is planned again next time
                else:
                    dplan.drop_workspace()
                    keep.append(dplan)
        if keep is not None:
            cache.store(batch, params, expectations, self.device, keep)

    def _build_device_plans(self, batch, params, expectations):
        chunks = self._chunks(batch, params, expectations)
        torch = self.torch
        if self.plan_on == "device":
            # chunks k+1 .. k+PLAN_AHEAD are packed into pinned memory, uploaded and planned by helper threads, each on
            # its own stream, while the DP kernels of chunk k run on the caller's stream.  Packing (host memcpy, GIL
            # released) of one chunk overlaps the planner kernels and their host round trips of another.  Staging
            # buffers rotate over PLAN_AHEAD + 1 slots: the H2D copy of a buffer has completed before its helper hands
            # the plan over, and the slot is not packed again before that plan has been consumed.
            ahead = max(1, min(self.PLAN_AHEAD, len(chunks)))
            while len(self._stagings) < ahead + 1:
                self._stagings += (_Staging(),)
            while len(self._plan_streams) < ahead:
                self._plan_streams.append(torch.cuda.Stream(device=self.device))

            def make(i):
                t0 = time.perf_counter()
                stream = self._plan_streams[i % ahead]
                with torch.cuda.device(self.device), torch.cuda.stream(stream):
                    packed = DevicePlan.pack_inputs(batch.slice(*chunks[i]), self._stagings[i % (ahead + 1)])
                    t1 = time.perf_counter()
                    dplan = DevicePlan.build_on_device(None, params, self.device, expectations, packed=packed,
                                                       alloc_ws=False)
                    stream.synchronize()
                t2 = time.perf_counter()
                with self._t_lock:
                    self._t["pack_s"] += t1 - t0
                    self._t["plan_s"] += t2 - t0
                return dplan

            with ThreadPoolExecutor(ahead) as pool:
                futs = [pool.submit(make, i) for i in range(ahead)]
                for ci in range(len(chunks)):
                    t0 = time.perf_counter()
                    dplan = futs[ci].result()
                    futs[ci] = None
                    self._t["wait_plan_s"] += time.perf_counter() - t0
                    if ci + ahead < len(chunks):
                        futs.append(pool.submit(make, ci + ahead))
                    dplan.adopt(torch.cuda.current_stream(self.device), self._workspace)
                    yield dplan
            return
        with ThreadPoolExecutor(1) as pool:
            make = lambda b: Plan.from_batch(batch.slice(*b), params, self.plan_threads)
            fut = pool.submit(make, chunks[0])
            for ci in range(len(chunks)):
                plan = fut.result()
                if ci + 1 < len(chunks):
                    fut = pool.submit(make, chunks[ci + 1])
                yield DevicePlan(plan, self.device, expectations)

    # ---------------------------------------------------------------- posteriors
    def posteriors_device(self, model, dplan):
        """Device-resident half of posterior mode: returns (offsets, x, y, p) as torch device tensors."""
        torch, L = self.torch, capi.load()
        n = dplan.plan.info.n_alignments
        d_off = torch.empty(n + 1, dtype=torch.int64, device=self.device)
        d_np = torch.empty(1, dtype=torch.int64, device=self.device)
        capi.check(L.mab_posteriors_compute(model._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr,
                                            dplan.ws_bytes, d_off.data_ptr(), d_np.data_ptr(), self._stream_ptr()))
        total = int(d_np.item())  # one sync: the output size is data dependent
        d_x = torch.empty(total, dtype=torch.int32, device=self.device)
        d_y = torch.empty(total, dtype=torch.int32, device=self.device)
        d_p = torch.empty(total, dtype=torch.int32, device=self.device)
        capi.check(L.mab_posteriors_emit(dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes,
                                         d_x.data_ptr(), d_y.data_ptr(), d_p.data_ptr(), total, self._stream_ptr()))
        return d_off, d_x, d_y, d_p

    def _get_aligned_pairs_sync(self, model, batch, params=None):
        """Posterior match probabilities of every alignment of `batch` (host in, host out).

        One host sync per chunk (the pair count, which sizes the outputs).  The host-side work for the results of
        chunk k (pinned blocks, queueing the D2H copies on the copy stream) is done after the DP kernels of chunk k+1
        have been queued, so the GPU does not wait for it."""
        params = params or AlignmentParameters()
        torch, L = self.torch, capi.load()
        offsets, blocks = np.zeros(batch.n + 1, dtype=np.int64), []
        stats = dict(cells=0, regions=0, diagonals=0, h2d_bytes=0, d2h_bytes=0, chunks=0)
        pending = []  # D2H copies in flight: (first alignment, n alignments, pinned tensors, event, device tensors)

        def start_copy(a0, na, dev, done):
            t1 = time.perf_counter()
            # pinned host blocks (torch's pinned allocator recycles them between calls); the D2H copies run on a
            # second stream so they overlap the kernels of the next chunks
            host = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in dev]
            with torch.cuda.stream(self._copy_stream):
                self._copy_stream.wait_event(done)
                for h, t in zip(host, dev):
                    h.copy_(t, non_blocking=True)
                    t.record_stream(self._copy_stream)
                ev = torch.cuda.Event()
                ev.record(self._copy_stream)
            pending.append((a0, na, host, ev, None))  # record_stream above keeps the device arrays until the copies ran
            self._t["host_alloc_s"] += time.perf_counter() - t1

        # coordinates are relative to the sub-sequences: 16 bits are enough unless one of them reaches 32768 bases, and
        # the pairs are what this call moves (8 instead of 12 bytes each over PCIe and into host memory)
        longest = int(max(np.diff(batch.off_x).max(initial=0), np.diff(batch.off_y).max(initial=0))) if batch.n else 0
        narrow = longest < 32768 and self.compact_pairs
        xy_dtype, emit = (torch.int16, L.mab_posteriors_emit_i16) if narrow else (torch.int32, L.mab_posteriors_emit)
        first, prev = 0, None
        with torch.cuda.device(self.device):
            for k in self._t:
                self._t[k] = 0.0
            for dplan in self._device_plans(batch, params, False):
                info = dplan.plan.info
                n = info.n_alignments
                t0 = time.perf_counter()
                d_off = torch.empty(n + 1, dtype=torch.int64, device=self.device)
                d_np = torch.empty(1, dtype=torch.int64, device=self.device)
                capi.check(L.mab_posteriors_compute(model._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr,
                                                    dplan.ws_bytes, d_off.data_ptr(), d_np.data_ptr(),
                                                    self._stream_ptr()))
                if prev is not None:
                    start_copy(*prev)
                t1 = time.perf_counter()
                total = int(d_np.item())  # the sync: the output size is data dependent
                d_x = torch.empty(total, dtype=xy_dtype, device=self.device)
                d_y = torch.empty(total, dtype=xy_dtype, device=self.device)
                d_p = torch.empty(total, dtype=torch.int32, device=self.device)
                capi.check(emit(dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes,
                                d_x.data_ptr(), d_y.data_ptr(), d_p.data_ptr(), total, self._stream_ptr()))
                done = torch.cuda.Event()
                done.record(torch.cuda.current_stream(self.device))
                self._t["compute_s"] += time.perf_counter() - t1
                prev = (first, n, (d_off, d_x, d_y, d_p), done)
                first += n
                stats["cells"] += info.n_cells
                stats["regions"] += info.n_regions
                stats["diagonals"] += info.n_diagonals
                stats["h2d_bytes"] += dplan.h2d_bytes
                stats["chunks"] += 1
                del dplan
            if prev is not None:
                start_copy(*prev)
            t_drain = time.perf_counter()
            for a0, na, host, ev, _keep in pending:
                ev.synchronize()
                off = host[0].numpy()
                base = int(offsets[a0])
                np.add(off[1:], base, out=offsets[a0 + 1:a0 + na + 1])
                blocks.append((a0, a0 + na, base, host[1].numpy(), host[2].numpy(), host[3].numpy()))
                stats["d2h_bytes"] += sum(h.numel() * h.element_size() for h in host)
            self._t["drain_s"] = time.perf_counter() - t_drain
        stats["host_wall"] = dict(self._t)
        if not blocks:
            blocks = [(0, 0, 0, np.zeros(0, np.int32), np.zeros(0, np.int32), np.zeros(0, np.int32))]
        self.last_stats = stats
        return PosteriorPairs(offsets, blocks=blocks)

    PAIRS_PER_CELL_CAP = 0.25  # output capacity of the sync-free path (measured: 0.07 pairs per cell at threshold 0.01)

    def posteriors_async(self, model, dplan, slot=None):
        """posteriors_device without the host sync: outputs are sized by a bound (PAIRS_PER_CELL_CAP x cells; the emit
        kernel clips at the capacity), the pair count follows asynchronously into pinned memory.  Returns a state
        dict; `posteriors_settle` must be called before the pairs are trusted."""
        torch, L = self.torch, capi.load()
        info = dplan.plan.info
        n = info.n_alignments
        cap = int(info.n_cells * self.PAIRS_PER_CELL_CAP) + 1024
        if slot is None:
            new = lambda name, numel, dtype: torch.empty(numel, dtype=dtype, device=self.device)
        else:  # the engine's buffers of this pipeline slot: the copies of the chunk that used them last must be done
            new = lambda name, numel, dtype: self._pooled(slot, name, numel, dtype)
            if self._slot_free[slot] is not None:
                torch.cuda.current_stream(self.device).wait_event(self._slot_free[slot])
        st = dict(dplan=dplan, model=model, n=n, cap=cap, slot=slot, new=new,
                  d_off=new("d_off", n + 1, torch.int64),
                  d_np=torch.empty(1, dtype=torch.int64, device=self.device),
                  d_x=new("d_x", cap, torch.int32),
                  d_y=new("d_y", cap, torch.int32),
                  d_p=new("d_p", cap, torch.int32),
                  h_np=torch.empty(1, dtype=torch.int64, pin_memory=True))
        sp = self._stream_ptr()
        capi.check(L.mab_posteriors_compute(model._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes,
                                            st["d_off"].data_ptr(), st["d_np"].data_ptr(), sp))
        capi.check(L.mab_posteriors_emit(dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes,
                                         st["d_x"].data_ptr(), st["d_y"].data_ptr(), st["d_p"].data_ptr(), cap, sp))
        st["h_np"].copy_(st["d_np"], non_blocking=True)
        st["counted"] = torch.cuda.Event()
        st["counted"].record(torch.cuda.current_stream(self.device))
        return st

    def posteriors_settle(self, st):
        """Waits for the pair count of an async chunk; if the bound was too small (more than PAIRS_PER_CELL_CAP pairs
        per band cell: a threshold near 0) the chunk is computed again and its pairs are emitted into exactly sized
        buffers.  Computed again, not just emitted again: the chunks of a call share the engine's workspace, and the
        next chunk's kernels -- queued before this one is settled -- have overwritten it.  Returns the number of
        pairs."""
        torch, L = self.torch, capi.load()
        st["counted"].synchronize()
        total = int(st["h_np"][0])
        if total > st["cap"]:
            dplan, sp = st["dplan"], self._stream_ptr()
            dplan.use_workspace(self._workspace(dplan.ws_bytes + 256))
            for k in ("d_x", "d_y", "d_p"):
                st[k] = torch.empty(total, dtype=torch.int32, device=self.device)
            capi.check(L.mab_posteriors_compute(st["model"]._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr,
                                                dplan.ws_bytes, st["d_off"].data_ptr(), st["d_np"].data_ptr(), sp))
            capi.check(L.mab_posteriors_emit(dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr, dplan.ws_bytes,
                                             st["d_x"].data_ptr(), st["d_y"].data_ptr(), st["d_p"].data_ptr(), total,
                                             sp))
            st["cap"] = total
            st["overflowed"] = True
        st["total"] = total
        return total

    def get_aligned_pairs_using_anchors(self, model, batch, params=None):
        """Posterior match probabilities of every alignment of `batch` (host in, host out): one host sync per chunk
        (the pair count).  (A sync-free variant with bound-sized outputs -- what PairHmmEngine.realign builds on --
        measured slower for plain posteriors, profiles/r1_ncu_summary.md r1g, and was removed.)"""
        return self._get_aligned_pairs_sync(model, batch, params)

    # ---------------------------------------------------------------- expectations
    def expectations_device(self, model, dplan, d_out):
        capi.check(capi.load().mab_expectations(model._h, dplan.plan._h, dplan.blob_ptr, dplan.ws_ptr,
                                                dplan.ws_bytes, d_out.data_ptr(), self._stream_ptr()))

    def get_expectations_using_anchors(self, model, batch, params=None, pseudo_count=1e-12, as_hmm=False,
                                       plan_cache=None):
        """EM expectations summed over the batch: (transitions[S*S], emissions[S*16], likelihood).

        Starts from hmm_constructEmpty(1e-12) like cPecanRealign.c:531 (pseudo-count on transitions and emissions,
        likelihood 0).  plan_cache: a PlanCache that keeps the batch's device plans for the next call (the EM loop
        passes one: only the model changes between iterations)."""
        params = params or AlignmentParameters()
        torch = self.torch
        S = model.state_number
        init = np.full(model.n_expectations, pseudo_count)
        init[-1] = 0.0
        stats = dict(cells=0, regions=0, diagonals=0, h2d_bytes=0, d2h_bytes=0, chunks=0)
        with torch.cuda.device(self.device):
            d_out = torch.from_numpy(init).to(self.device)
            done = []  # one event per chunk: at most two chunks' workspaces are alive, and the host never waits for
            for dplan in self._device_plans(batch, params, True, plan_cache):  # the chunk it has just queued
                info = dplan.plan.info
                if len(done) >= 2:
                    done[-2].synchronize()
                self.expectations_device(model, dplan, d_out)
                ev = torch.cuda.Event()
                ev.record(torch.cuda.current_stream(self.device))
                done.append(ev)
                stats["cells"] += info.n_cells
                stats["regions"] += info.n_regions
                stats["diagonals"] += info.n_diagonals
                stats["h2d_bytes"] += dplan.h2d_bytes
                stats["chunks"] += 1
                del dplan
            out = d_out.cpu().numpy()
            stats["d2h_bytes"] = out.nbytes
        self.last_stats = stats
        T, E, lik = out[:S * S].copy(), out[S * S:S * S + S * 16].copy(), float(out[-1])
        if as_hmm:
            h = Hmm(model.type)
            h.transitions, h.emissions, h.likelihood = T, E, lik
            return h
        return T, E, lik


# ------------------------------------------------------------------------------------------------ realign tail
class RealignedCigars:
    """Result of PairHmmEngine.realign: the new cigar of every alignment (op codes 0 M, 1 D = reference only,
    2 I = read only, relative to the alignment's sub-sequences, like the guide cigars of an AlignmentBatch), and the
    chain of aligned pairs when it was asked for."""

    def __init__(self, off_ops, ops, lens, chain_off=None, chain_x=None, chain_y=None):
        # ops, lens, chain_x, chain_y: arrays, or lists of per-chunk arrays (pinned host blocks) that are joined the
        # first time they are read -- the engine does not pay for a copy the caller may never need
        self.off_ops, self.chain_off = off_ops, chain_off
        self._v = dict(ops=ops, lens=lens, chain_x=chain_x, chain_y=chain_y)

    def _get(self, k):
        v = self._v[k]
        if isinstance(v, list):
            v = self._v[k] = np.concatenate(v) if len(v) > 1 else v[0]
        return v

    ops = property(lambda self: self._get("ops"))
    lens = property(lambda self: self._get("lens"))
    chain_x = property(lambda self: self._get("chain_x"))
    chain_y = property(lambda self: self._get("chain_y"))

    @property
    def n(self):
        return len(self.off_ops) - 1

    def cigar(self, i):
        s = slice(int(self.off_ops[i]), int(self.off_ops[i + 1]))
        return self.ops[s], self.lens[s]

    def chain(self, i):
        s = slice(int(self.chain_off[i]), int(self.chain_off[i + 1]))
        return np.stack([self.chain_x[s], self.chain_y[s]], axis=1)


_JITTER = {}


def jitter_table(n_draws):
    """The first n_draws values of glibc rand() after srand(1) (every cPecanRealign process starts there): the stream
    behind st_random() in makeAlignmentWeight (multipleAligner.c:140-147)."""
    have = _JITTER.get("t")
    if have is None or len(have) < n_draws:
        n = max(1 << 16, 1 << int(np.ceil(np.log2(max(n_draws, 2)))))
        have = np.empty(n, dtype=np.int32)
        capi.check(capi.load().mab_rand_sequence(1, n, have.ctypes.data))
        _JITTER["t"] = have
    return have


def _realign_launch(engine, model, dplan, sub, expansion, gap_gamma, match_gamma, want_chain, slot=None):
    """Queues everything one chunk needs -- posteriors, realign chain, cigar emission -- without a host sync: the
    outputs are sized by bounds (at most 2 (lX + lY) + 1 operations and min(lX, lY) chain pairs per alignment)."""
    torch, L = engine.torch, capi.load()
    dev = engine.device
    st = engine.posteriors_async(model, dplan, slot)
    n, new = st["n"], st["new"]
    off_x, off_y = sub.off_x, sub.off_y
    if getattr(dplan, "input_offsets", None) is not None:  # planned on the device: the offsets are already in HBM
        ox_ptr, oy_ptr, len_x, len_y = dplan.input_offsets
        h_offs = d_offs = None
    else:
        len_x, len_y = (int(off_x[-1]), int(off_y[-1])) if n else (0, 0)
        h_offs = torch.from_numpy(np.stack([np.ascontiguousarray(off_x, dtype=np.int64),
                                            np.ascontiguousarray(off_y, dtype=np.int64)])).pin_memory()
        d_offs = h_offs.to(dev, non_blocking=True)
        ox_ptr, oy_ptr = d_offs[0].data_ptr(), d_offs[1].data_ptr()
    # jitter stream: two draws per pair.  An alignment has at most one pair per band cell, i.e. at most (diagonals x the
    # widest diagonal of the batch) pairs: a bound, not an estimate (the kernels refuse to wrap the stream, and
    # _engine_realign raises if an alignment was refused because the 2^25-draw cap was hit)
    est = int(((off_x[1:] - off_x[:-1]) + (off_y[1:] - off_y[:-1]) + 1).max() * max(int(dplan.plan.info.max_width), 1)) if n else 0
    table = jitter_table(min(2 * est + 2, 1 << 25))
    # grow-only on the device: a longer prefix of the same stream gives the same draws
    if engine._jitter_dev is None or engine._jitter_dev[0] != str(dev) or engine._jitter_dev[1].numel() < len(table):
        engine._jitter_dev = (str(dev), torch.from_numpy(table).to(dev))
    d_rand = engine._jitter_dev[1]
    st["n_rand"] = int(d_rand.numel())
    st["span"] = np.asarray((off_x[1:] - off_x[:-1]) + (off_y[1:] - off_y[:-1]))
    cap = st["cap"]
    ws_bytes = int(L.mab_realign_workspace_bytes(n, cap, len_x, len_y))
    d_ws = new("rl_ws", ws_bytes + 256, torch.uint8)
    ws_ptr = (d_ws.data_ptr() + 255) & ~255
    cap_ops = 2 * (len_x + len_y) + n + 1
    cap_chain = min(len_x, len_y) + 1
    st.update(len_x=len_x, len_y=len_y, d_offs=d_offs, h_offs=h_offs, ox_ptr=ox_ptr, oy_ptr=oy_ptr, d_ws=d_ws, ws_ptr=ws_ptr, ws_bytes=ws_bytes,
              d_op_off=new("d_op_off", n + 1, torch.int64),
              d_ch_off=new("d_ch_off", n + 1, torch.int64),
              d_ops=new("d_ops", cap_ops, torch.uint8),
              d_lens=new("d_lens", cap_ops, torch.int32),
              d_cx=new("d_cx", cap_chain, torch.int32) if want_chain else None,
              d_cy=new("d_cy", cap_chain, torch.int32) if want_chain else None,
              h_tot=torch.empty(2, dtype=torch.int64, pin_memory=True), d_rand=d_rand,
              gammas=(gap_gamma, match_gamma), want_chain=want_chain)
    _realign_kernels(engine, st)
    return st


def _realign_kernels(engine, st):
    torch, L = engine.torch, capi.load()
    n, sp = st["n"], engine._stream_ptr()
    ox, oy = st["ox_ptr"], st["oy_ptr"]
    capi.check(L.mab_realign_chain(n, ox, oy, st["len_x"], st["len_y"], st["d_off"].data_ptr(),
                                   st["d_x"].data_ptr(), st["d_y"].data_ptr(), st["d_p"].data_ptr(), st["cap"],
                                   st["gammas"][0], st["gammas"][1], st["d_rand"].data_ptr(), st["d_rand"].numel(),
                                   st["ws_ptr"], st["ws_bytes"], st["d_op_off"].data_ptr(),
                                   st["d_ch_off"].data_ptr(), sp))
    wc = st["want_chain"]
    capi.check(L.mab_realign_emit(n, ox, oy, st["len_x"], st["len_y"], st["cap"], st["ws_ptr"],
                                  st["ws_bytes"], st["d_op_off"].data_ptr(), st["d_ch_off"].data_ptr(),
                                  st["d_ops"].data_ptr(), st["d_lens"].data_ptr(),
                                  st["d_cx"].data_ptr() if wc else None, st["d_cy"].data_ptr() if wc else None, sp))
    st["h_tot"].copy_(torch.stack([st["d_op_off"][n], st["d_ch_off"][n]]), non_blocking=True)
    st["done"] = torch.cuda.Event()
    st["done"].record(torch.cuda.current_stream(engine.device))


def _realign_finish(engine, st):
    """Waits for a launched chunk and brings its cigars to the host (copy stream: overlaps the next chunk's kernels)."""
    torch, L = engine.torch, capi.load()
    before = st["cap"]
    engine.posteriors_settle(st)
    if st["cap"] != before:  # the pair bound was too small (never seen at threshold 0.01): redo the tail, exact sizes
        ws_bytes = int(L.mab_realign_workspace_bytes(st["n"], st["cap"], st["len_x"], st["len_y"]))
        st["d_ws"] = st["new"]("rl_ws", ws_bytes + 256, torch.uint8)
        st["ws_ptr"], st["ws_bytes"] = (st["d_ws"].data_ptr() + 255) & ~255, ws_bytes
        _realign_kernels(engine, st)
    st["done"].synchronize()
    n_ops, n_chain = int(st["h_tot"][0]), int(st["h_tot"][1])
    dev = [st["d_op_off"], st["d_ops"][:n_ops], st["d_lens"][:n_ops]]
    if st["want_chain"]:
        dev += [st["d_ch_off"], st["d_cx"][:n_chain], st["d_cy"][:n_chain]]
    host = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in dev]
    with torch.cuda.stream(engine._copy_stream):
        for h, t in zip(host, dev):
            h.copy_(t, non_blocking=True)
            t.record_stream(engine._copy_stream)
        ev = torch.cuda.Event()
        ev.record(engine._copy_stream)
    # the chunk's device buffers (pair arrays sized by a bound, tail scratch: ~50 B per band cell) are done with once the
    # copies above are queued; the caller keeps `st` until the end of the batch, so only what it reads later stays
    # (20 000 reads of 10 kb ran out of HBM with 14 chunks' buffers alive)
    if st.get("slot") is not None:
        engine._slot_free[st["slot"]] = ev  # the slot's buffers may be written again once these copies have run
    keep = {k: st[k] for k in ("span", "n_rand", "total")}
    st.clear()
    st.update(keep)
    return host, ev, st


def _engine_realign(self, model, batch, params=None, gap_gamma=0.5, match_gamma=0.0, want_chain=False):
    """Realigns every guide alignment of `batch`: posterior pairs -> AMAP reweight -> maximum-expected-accuracy chain
    -> cigar, all on the GPU (cPecanRealign's default mode as marginAlignLib.py:315-318 drives it: --gapGamma,
    --matchGamma, --diagonalExpansion 10 --splitMatrixBiggerThanThis 3000 --loadHmm).  Only the cigars come back.
    Software-pipelined by one chunk like get_aligned_pairs_using_anchors."""
    params = params or AlignmentParameters()
    torch = self.torch
    copies = []
    stats = dict(cells=0, chunks=0, h2d_bytes=0, d2h_bytes=0, pairs=0)
    first, prev = 0, None
    with torch.cuda.device(self.device):
        for k in self._t:
            self._t[k] = 0.0
        for dplan in self._device_plans(batch, params, False):
            info = dplan.plan.info
            n = info.n_alignments
            t0 = time.perf_counter()
            st = _realign_launch(self, model, dplan, batch.slice(first, first + n), params.diagonal_expansion, gap_gamma,
                                 match_gamma, want_chain, slot=stats["chunks"] & 1)
            t1 = time.perf_counter()
            self._t["host_alloc_s"] += t1 - t0  # queueing a chunk: buffers + launches, no sync
            if prev is not None:
                copies.append(_realign_finish(self, prev))
            self._t["compute_s"] += time.perf_counter() - t1  # waiting for the previous chunk's kernels
            prev = st
            first += n
            stats["cells"] += info.n_cells
            stats["chunks"] += 1
            stats["h2d_bytes"] += dplan.h2d_bytes + 16 * (n + 1)
            del dplan, st
        if prev is not None:
            t1 = time.perf_counter()
            copies.append(_realign_finish(self, prev))
            self._t["compute_s"] += time.perf_counter() - t1
            prev = None
        parts = []
        t_drain = time.perf_counter()
        for host, ev, st in copies:
            ev.synchronize()
            arrs = [h.numpy() for h in host]
            # a non-empty alignment without a single cigar operation was refused by the tail kernels (k_rl_chain): it
            # has more posterior pairs than the jitter stream holds, and wrapping the stream would change the cigar
            if ((np.diff(arrs[0]) == 0) & (st["span"] > 0)).any():
                raise RuntimeError("realign: an alignment has more posterior pairs than the jitter stream holds "
                                   "(%d draws); split the batch or lower diagonalExpansion" % st["n_rand"])
            parts.append(arrs + [None] * (6 - len(arrs)))
            stats["pairs"] += st["total"]
            stats["d2h_bytes"] += sum(a.nbytes for a in arrs)
        self._t["drain_s"] = time.perf_counter() - t_drain
    stats["host_wall"] = dict(self._t)
    self.last_stats = stats

    def cat_off(idx):
        out, base = [np.zeros(1, dtype=np.int64)], 0
        for p in parts:
            out.append(p[idx][1:] + base)
            base += int(p[idx][-1])
        return np.concatenate(out)

    def cat(idx, dtype):  # joined lazily by RealignedCigars
        return [p[idx] for p in parts] if parts else np.zeros(0, dtype=dtype)

    if want_chain:
        return RealignedCigars(cat_off(0), cat(1, np.uint8), cat(2, np.int32), cat_off(3), cat(4, np.int32),
                               cat(5, np.int32))
    return RealignedCigars(cat_off(0), cat(1, np.uint8), cat(2, np.int32))


PairHmmEngine.realign = _engine_realign
PairHmmEngine._jitter_dev = None
