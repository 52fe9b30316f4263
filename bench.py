#!/usr/bin/env python
"""bench.py -- throughput of the pair-HMM hot path on B200 (BASELINE.json metric: pair-HMM GCUPS).

A *step* is one pass of the hot path (forward, backward, totals, posteriors, ordered pair emission) over one batch of
synthetic guide alignments.  Default workload = BASELINE.json configs[2]: synthetic direct-tRNA reads (70-150 nt, ~15 %
error) against 400 tRNA references, posterior mode, 5-state asymmetric model, expansion 10.

  value  : whole-job cell updates per second with the plan (bands, schedule, symbols) already resident in HBM
  e2e    : the same metric through the public host API (PairHmmEngine.get_aligned_pairs_using_anchors): host buffers
           in, host arrays out; planning, H2D of the plan and D2H of the pairs are inside the timed region
  --impl reference : the reference's own C code (oracle/_ref, unmodified cPecan) on all host cores, bounded sample

Launch: python bench.py --gpus N --steps K --warmup W   (N>1: torchrun, one rank per GPU; weak scaling: every rank
processes its own --reads alignments; posterior mode has no collective, EM mode all-reduces 106 doubles).
"""
import argparse
import ctypes as C
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOPS_PER_CELL = {  # algorithmic flops per cell (add/sub/mul = 1, no FMA possible), 5-state / 3-state
    ("posterior", 5): 280, ("em", 5): 353, ("posterior", 3): 194, ("em", 3): 243,
}
KERNEL_FLOPS_5 = {"k_forward": 130, "k_backward": 130}  # 13 transitions x 10 flops each


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="trna", choices=["trna", "long"])
    ap.add_argument("--mode", default="posterior", choices=["posterior", "em", "realign"],
                    help="posterior: getAlignedPairsUsingAnchors; em: getExpectationsUsingAnchors; realign: posterior "
                         "+ AMAP reweight + MEA chain + cigar (cPecanRealign's default mode)")
    ap.add_argument("--states", type=int, default=5, choices=[5, 3],
                    help="5: fiveStateAsymmetric with the tests/input.hmm values (default); 3: the built-in threeState "
                         "model (stateMachine3_construct), our arm only")
    ap.add_argument("--reads", type=int, default=0, help="alignments per GPU per step (default: 1M tRNA / 2000 long)")
    ap.add_argument("--expansion", type=int, default=0, help="diagonalExpansion (default 10 tRNA / 50 long)")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--ref-sample", type=int, default=0, help="alignments in the CPU reference sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="skip the end-to-end leg (profiling runs)")
    ap.add_argument("--max-cells", type=int, default=0,
                    help="DP cells resident per chunk (0: the engine sizes chunks from the band width and free HBM)")
    return ap.parse_args()


def make_workload(args, rank):
    import marginalign_trna_b200 as mab
    if args.workload == "trna":
        n = args.reads or 1_000_000
        exp = args.expansion or 10
        batch = mab.synth.make_trna_batch(n, n_refs=400, seed=1234 + rank)
        params = mab.AlignmentParameters(diagonal_expansion=exp, split_matrix_bigger_than=3000)
        name = "synthetic direct-tRNA reads (70-150 nt, 5/5/5 % sub/ins/del) vs 400 tRNA refs, posterior mode"
    else:
        n = args.reads or 2000
        exp = args.expansion or 50
        batch = mab.synth.make_long_batch(n, read_len=10000, genome_len=5_000_000, seed=4321 + rank)
        params = mab.AlignmentParameters(diagonal_expansion=exp, split_matrix_bigger_than=300)
        name = "synthetic 10 kb nanopore reads vs 5 Mb genome"
    return batch, params, name


# ------------------------------------------------------------------------------------------------ clocks

class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                smax = float(f[2])
            except ValueError:
                continue
            for k, nme in enumerate(names):
                if f[4 + k].lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU reference arm

_W = {}


def _ref_init(model_path, params_tuple):
    from oracle import pyref
    import marginalign_trna_b200 as mab
    h = mab.Hmm.loadHmm(model_path)
    _W["model"] = pyref.RefModel.from_hmm(h._modelTypeInt(), h.transitions, h.emissions)
    _W["pyref"] = pyref
    _W["params"] = params_tuple


def _port_init(model_path, params_tuple):
    from oracle import pyoracle
    import marginalign_trna_b200 as mab
    h = mab.Hmm.loadHmm(model_path)
    _W["model"] = pyoracle.model_from_hmm(h._modelTypeInt(), h.transitions, h.emissions)
    _W["po"] = pyoracle
    _W["params"] = params_tuple


def _ref_work(item):
    """One alignment through the reference's getAlignedPairsUsingAnchors / getExpectationsUsingAnchors."""
    x, y, anchors, mode = item
    exp, split = _W["params"]
    if "pyref" in _W:
        pr = _W["pyref"]
        if mode == "posterior":
            return len(pr.aligned_pairs(_W["model"], x, y, anchors, expansion=exp, split=split))
        if mode == "realign":  # cPecanRealign.c:584-611 through the reference library
            pairs = pr.aligned_pairs(_W["model"], x, y, anchors, expansion=exp, split=split)
            return len(pr.realign_tail(pairs, x, y, 0.5, 0.0, seed=1))
        pr.expectations(_W["model"], x, y, anchors, expansion=exp, split=split)
        return 0
    po = _W["po"]
    P = po.make_params(expansion=exp, split=split)
    if mode == "posterior":
        return len(po.aligned_pairs(_W["model"], x, y, anchors, P))
    if mode == "realign":
        pairs = po.aligned_pairs(_W["model"], x, y, anchors, P)
        return len(po.realign_chain(pairs, len(x), len(y), 0.5, 0.0, seed=1)[0])
    po.expectations(_W["model"], x, y, anchors, P)
    return 0


def cpu_reference_throughput(batch, params, mode, n_sample, steps=1, warmup=0):
    """Times the reference C path (oracle/_ref if built, else our C port) over all host cores on the first n_sample
    alignments; returns dict(value GCUPS, cores, kind, sample, ms_per_step)."""
    import marginalign_trna_b200 as mab
    from oracle import pyoracle as po
    from oracle import pyref
    kind = "reference" if pyref.available() else "port"
    cores = os.cpu_count() or 1
    n_sample = min(n_sample, batch.n)
    sub = batch.subset(np.arange(n_sample))
    plan = mab.Plan.from_batch(sub, params)
    cells = plan.info.n_cells
    items = []
    for i in range(sub.n):
        items.append((sub.x(i), sub.y(i), plan.anchors(i), mode))
    model_path = os.path.join(ROOT, "tests", "golden", "input_model.hmm")
    ptuple = (params.diagonal_expansion, params.split_matrix_bigger_than ** 2)
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_ref_init if kind == "reference" else _port_init,
                  initargs=(model_path, ptuple)) as pool:
        chunk = max(1, len(items) // (cores * 8))
        for _ in range(warmup):
            pool.map(_ref_work, items, chunksize=chunk)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_ref_work, items, chunksize=chunk)
        dt = (time.perf_counter() - t0) / max(1, steps)
    return dict(value=cells / dt / 1e9, unit="GCUPS", cores=cores, kind=kind, ms_per_step=dt * 1e3,
                reads_per_s=n_sample / dt, cells=int(cells),
                sample="first %d alignments of the workload (%d cells), %s mode, one %s call per alignment in a "
                       "%d-process pool" % (n_sample, cells, mode, {"posterior": "getAlignedPairsUsingAnchors",
                                             "realign": "getAlignedPairsUsingAnchors + reweightAlignedPairs2 + "
                                                        "filterPairwiseAlignmentToMakePairsOrdered",
                                             "em": "getExpectationsUsingAnchors"}[mode], cores))


def run_reference(args, rank, world):
    if rank != 0:
        return
    batch, params, name = make_workload(args, 0)
    n_sample = args.ref_sample or (20000 if args.workload == "trna" else 16)
    r = cpu_reference_throughput(batch, params, args.mode, n_sample, steps=args.steps, warmup=args.warmup)
    line = {
        "impl": "reference", "metric": "pair-HMM cell updates per second (%s mode)" % args.mode,
        "value": r["value"], "unit": "GCUPS", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "mode": args.mode, "diagonal_expansion": params.diagonal_expansion,
                   "model": "fiveStateAsymmetric (tests/input.hmm values)", "sample_alignments": n_sample},
        "reads_per_s": r["reads_per_s"],
        "cpu_baseline": {"value": r["value"], "unit": "GCUPS", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"]},
        "e2e": {"value": r["value"], "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ our arm

def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def run_ours(args, rank, world, local_rank):
    import torch
    import marginalign_trna_b200 as mab
    from marginalign_trna_b200 import capi
    L = capi.load()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback on this path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    batch, params, name = make_workload(args, rank)
    hmm = mab.Hmm.loadHmm(os.path.join(ROOT, "tests", "golden", "input_model.hmm"))
    model = mab.PairHmmModel.from_hmm(hmm) if args.states == 5 else mab.PairHmmModel.default(2)
    S = model.state_number
    eng = mab.PairHmmEngine(device=dev, max_cells=args.max_cells or None)
    em = args.mode == "em"

    # ---- resident setup: plan every chunk and copy its blob into HBM; one workspace sized for the largest chunk
    t_plan0 = time.perf_counter()
    chunks = eng._chunks(batch, params, em)
    plans = [mab.Plan.from_batch(batch if b == (0, batch.n) else batch.subset(np.arange(b[0], b[1])), params)
             for b in chunks]
    t_plan = time.perf_counter() - t_plan0
    cells = sum(p.info.n_cells for p in plans)
    n_reads = batch.n
    ws_need = max((p.info.workspace_bytes_expectations if em else p.info.workspace_bytes_posteriors) for p in plans)
    d_ws = torch.empty(ws_need + 256, dtype=torch.uint8, device=dev)
    ws_ptr = (d_ws.data_ptr() + 255) & ~255
    d_blobs = []
    for p in plans:
        hb = torch.empty(p.info.blob_bytes, dtype=torch.uint8, pin_memory=True)
        p.write_blob(hb.data_ptr(), hb.numel())
        d_blobs.append(hb.to(dev, non_blocking=True))
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream(dev)
    sptr = C.c_void_p(stream.cuda_stream)
    max_n = max(p.info.n_alignments for p in plans)
    d_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)
    d_np = torch.zeros(1, dtype=torch.int64, device=dev)
    h_np = torch.zeros(1, dtype=torch.int64, pin_memory=True)
    # output capacity: generous fixed bound so the resident loop has no host sync (pairs/cell ~ 0.07)
    cap = int(max(p.info.n_cells for p in plans) * 0.25) + 1024
    d_x = torch.empty(cap, dtype=torch.int32, device=dev)
    d_y = torch.empty(cap, dtype=torch.int32, device=dev)
    d_p = torch.empty(cap, dtype=torch.int32, device=dev)
    init = np.full(model.n_expectations, 1e-12)
    init[-1] = 0.0
    d_exp = torch.from_numpy(init).to(dev)
    realign = args.mode == "realign"
    rl = []
    if realign:  # per chunk: sequence offsets on the device; shared: jitter stream, scratch, cigar outputs
        d_rand = torch.from_numpy(mab.jitter_table(1 << 22)).to(dev)
        for b in chunks:
            sub = batch.slice(*b)
            rl.append((sub.n, torch.from_numpy(np.ascontiguousarray(sub.off_x, dtype=np.int64)).to(dev),
                       torch.from_numpy(np.ascontiguousarray(sub.off_y, dtype=np.int64)).to(dev),
                       int(sub.off_x[-1]), int(sub.off_y[-1])))
        rl_ws = max(int(L.mab_realign_workspace_bytes(n, cap, lx, ly)) for n, _, _, lx, ly in rl)
        d_rl_ws = torch.empty(rl_ws + 256, dtype=torch.uint8, device=dev)
        rl_ptr = (d_rl_ws.data_ptr() + 255) & ~255
        cap_ops = max(2 * (lx + ly) + n for n, _, _, lx, ly in rl)
        d_ops = torch.empty(cap_ops, dtype=torch.uint8, device=dev)
        d_lens = torch.empty(cap_ops, dtype=torch.int32, device=dev)
        d_op_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)
        d_ch_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)

    def resident_step():
        for p, db in zip(plans, d_blobs):
            if em:
                capi.check(L.mab_expectations(model._h, p._h, db.data_ptr(), ws_ptr, ws_need, d_exp.data_ptr(), sptr))
            else:
                capi.check(L.mab_posteriors_compute(model._h, p._h, db.data_ptr(), ws_ptr, ws_need, d_off.data_ptr(),
                                                    d_np.data_ptr(), sptr))
                capi.check(L.mab_posteriors_emit(p._h, db.data_ptr(), ws_ptr, ws_need, d_x.data_ptr(), d_y.data_ptr(),
                                                 d_p.data_ptr(), cap, sptr))
                if realign:
                    n, ox, oy, lx, ly = rl[plans.index(p)]
                    capi.check(L.mab_realign_chain(n, ox.data_ptr(), oy.data_ptr(), lx, ly, d_off.data_ptr(),
                                                   d_x.data_ptr(), d_y.data_ptr(), d_p.data_ptr(), cap, 0.5, 0.0,
                                                   d_rand.data_ptr(), d_rand.numel(), rl_ptr, rl_ws,
                                                   d_op_off.data_ptr(), d_ch_off.data_ptr(), sptr))
                    capi.check(L.mab_realign_emit(n, ox.data_ptr(), oy.data_ptr(), lx, ly, cap, rl_ptr, rl_ws,
                                                  d_op_off.data_ptr(), d_ch_off.data_ptr(), d_ops.data_ptr(),
                                                  d_lens.data_ptr(), None, None, sptr))
        if em and dist is not None:
            dist.all_reduce(d_exp)  # the path's only collective: 106 (58) doubles per EM iteration

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(args.warmup):
        resident_step()
    barrier()
    L.mab_launch_count(1)
    L.mab_kernel_timing(1)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        resident_step()
    ev1.record(stream)
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    launches = int(L.mab_launch_count(0))
    nk = L.mab_kernel_count()
    kms = (C.c_double * nk)()
    kn = (C.c_int64 * nk)()
    capi.check(L.mab_kernel_times(kms, kn, 1))
    L.mab_kernel_timing(0)
    kernel_ms = {L.mab_kernel_name(k).decode(): (kms[k], int(kn[k])) for k in range(nk) if kn[k]}
    # pairs/cell sanity: the emit capacity must not have been exceeded
    h_np.copy_(d_np)
    torch.cuda.synchronize()
    if not em and int(h_np.item()) > cap:
        raise RuntimeError("pair capacity exceeded in the resident loop: %d > %d" % (int(h_np.item()), cap))

    t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_step = float(t.item()) / args.steps
    total_cells = cells * world  # weak scaling: every rank holds `cells` (same generator, different seed: ~equal)
    if dist is not None:
        c = torch.tensor([cells], dtype=torch.float64, device=dev)
        dist.all_reduce(c)
        total_cells = float(c.item())
    value = total_cells / (ms_step * 1e-3) / 1e9

    # ---- e2e through the public API (host in, host out)
    e2e = None
    e2e_steps = max(1, min(args.e2e_steps, args.steps))
    run_e2e = (lambda: eng.get_expectations_using_anchors(model, batch, params)) if em else \
              (lambda: eng.realign(model, batch, params)) if realign else \
              (lambda: eng.get_aligned_pairs_using_anchors(model, batch, params))
    if realign:
        del d_rl_ws, d_ops, d_lens
    del d_x, d_y, d_p, d_ws, d_blobs
    torch.cuda.empty_cache()
    if not args.no_e2e:
        for _ in range(max(1, min(args.warmup, 3))):
            run_e2e()  # warm-up: the device and page-locked allocator pools reach their steady state in 2-3 calls
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            run_e2e()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        st = dict(eng.last_stats)
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        e2e = {"value": total_cells / dt / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": int(st["h2d_bytes"]),
               "d2h_bytes_per_step": int(st["d2h_bytes"]), "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "reads_per_s": n_reads * world / dt, "host_wall_s": st.get("host_wall"),
               "note": "packing of bases+cigars into pinned memory, H2D, planning on the GPU, DP kernels, %s, per step "
                       "(rank 0 breakdown in host_wall_s)" % (
                           "realign tail on the GPU, D2H of the cigars" if realign else
                           "D2H of the expectation vector" if em else "D2H of the pairs into pinned host blocks")}

    if rank != 0:
        return
    peaks, peaks_src = load_peaks()
    # roofline of the dominant kernel: algorithmic flops / device time from CUDA events on the launch stream
    dom = max(("k_forward", "k_backward"), key=lambda k: kernel_ms.get(k, (0, 0))[0])
    dom_ms, dom_n = kernel_ms.get(dom, (0.0, 0))
    flops_cell_kernel = (130 if S == 5 else 90)
    roof = None
    if dom_n:
        per_launch_ms = dom_ms / dom_n
        cells_per_launch = cells * args.steps / dom_n
        achieved = cells_per_launch * flops_cell_kernel / (per_launch_ms * 1e-3) / 1e12
        # FP64 non-tensor, non-FMA issue peak: SMs x 64 FP64 lanes x clock (one add OR mul per lane per clock); the
        # bit-exact recurrence cannot use FMA.  No driver-measured FP64 figure exists in MEASURED_PEAKS.json, so the
        # denominator is nominal (148 x 64 x sm_max_mhz) and says so.
        smax = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        peak = 148 * 64 * smax * 1e6 / 1e12
        # algorithmic HBM bytes per cell of one DP kernel: posterior mode writes the match plane (8 B) and
        # all S states on the diagonals that carry a total (1 in 10); expectation mode writes all S states
        bytes_cell = S * 8 if em else 8 + S * 8 / 10.0
        roof = {"bound": "fp64", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                "frac": achieved / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the ncu --set full capture
                # (profiles/r1_ncu_summary.md r1h: k_forward<5,0,16>, 0.62 GB read + 3.76 GB written for a chunk of
                # 3.13e8 cells = 14.0 B per cell against 12 B algorithmic), scaled to this run's cells per launch
                "traffic": (cells_per_launch * 14.0 if (S == 5 and not em and args.workload == "trna") else None),
                "peak_source": "nominal non-FMA FP64 issue rate 148 SM x 64 lanes x %.0f MHz (no measured FP64 peak "
                               "in MEASURED_PEAKS.json)" % smax,
                "flops_per_cell": flops_cell_kernel, "ms_per_launch": per_launch_ms,
                "hbm": {"achieved": cells_per_launch * bytes_cell / (per_launch_ms * 1e-3) / 1e9,
                        "peak": peaks["hbm_gbs"], "unit": "GB/s", "peak_source": peaks_src + " MEASURED_PEAKS.json",
                        "bytes_per_cell": bytes_cell}}
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        n_sample = args.ref_sample or (20000 if args.workload == "trna" else 16)
        r = cpu_reference_throughput(batch, params, args.mode, n_sample)
        cpu = {"value": r["value"], "unit": "GCUPS", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
               "reads_per_s": r["reads_per_s"]}
    line = {
        "metric": "pair-HMM cell updates per second (%s mode)" % args.mode, "value": value, "unit": "GCUPS",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "mode": args.mode, "alignments_per_gpu": n_reads, "cells_per_gpu": int(cells),
                   "diagonal_expansion": params.diagonal_expansion, "chunks": len(plans),
                   "model": "fiveStateAsymmetric (tests/input.hmm values)" if S == 5 else "threeState (built-in)",
                   "l2": "inputs larger than L2: %.1f GB of DP state per step" % (cells * 80 / 1e9),
                   "parallelism": "reads sharded, dp%d" % world},
        "reads_per_s": n_reads * world / (ms_step * 1e-3),
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches,
        "kernel_ms_per_step": {k: v[0] / args.steps for k, v in kernel_ms.items()},
        "roofline": roof, "cpu_baseline": cpu, "host_plan_s": t_plan,
    }
    print(json.dumps(line), flush=True)


def main():
    if os.environ.get("MAB_SWITCH_INTERVAL"):  # experiment knob: GIL hand-over latency between the engine's threads
        sys.setswitchinterval(float(os.environ["MAB_SWITCH_INTERVAL"]))
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
