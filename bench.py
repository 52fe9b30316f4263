#!/usr/bin/env python
"""bench.py -- throughput of the pair-HMM hot path on B200 (BASELINE.json metric: pair-HMM GCUPS).

A *step* is one pass of the hot path (forward, backward, totals, posteriors, ordered pair emission) over one batch of
synthetic guide alignments.  Default workload = BASELINE.json configs[2]: synthetic direct-tRNA reads (70-150 nt, ~15 %
error) against 400 tRNA references, posterior mode, 5-state asymmetric model, expansion 10.

  value  : whole-job cell updates per second with the INPUTS (bases + guide cigars) resident in HBM: every step plans
           on the GPU (anchors, split regions, bands, checkpoint schedule: SURVEY 8 a3-a6) and runs the DP
  e2e    : the same metric through the public host API (PairHmmEngine.get_aligned_pairs_using_anchors): host buffers
           in, host arrays out; packing, H2D, planning and D2H of the pairs are inside the timed region
  em     : seconds per EM iteration through the public API, the NCCL all-reduce of the expectation vector and the
           M-step inside the timed region (weak scaling: every rank its own reads)
  strong : ONE batch split over the ranks by the product's sharder (ShardedPairHmm), results merged in input order
  --impl reference : the reference's own C code (oracle/_ref, unmodified cPecan) on all host cores, bounded sample;
           cpu_baseline.pattern = the reference as a user runs it, one cPecanRealign process per read

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
RR_TRAFFIC_B_PER_CELL = 13.1  # ncu r2b: k_backward_rr<5,0,16>, 0.154 GB read + 0.896 GB written per 8.0e7 cells


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
    ap.add_argument("--no-em-leg", action="store_true", help="skip the seconds-per-EM-iteration leg")
    ap.add_argument("--no-realign-leg", action="store_true", help="skip the realign-mode end-to-end leg")
    ap.add_argument("--no-strong-leg", action="store_true", help="skip the strong-scaling leg (ShardedPairHmm)")
    ap.add_argument("--em-reads", type=int, default=300000, help="alignments per GPU in the EM leg")
    ap.add_argument("--em-iterations", type=int, default=2)
    ap.add_argument("--int32-pairs", action="store_true", help="e2e leg: 32-bit pair coordinates even where 16 bits do")
    ap.add_argument("--chunk-ramp", default="",
                    help="divisors of the chunk size for the first chunks of a call, e.g. 8,2 (default: the engine's)")
    ap.add_argument("--no-pipelined-leg", action="store_true",
                    help="skip the resident loop with planning overlapped on a second stream")
    ap.add_argument("--plan-ahead", type=int, default=0, help="chunks planned ahead of the one computing (0: engine default)")
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
MODEL_PATH = os.path.join(ROOT, "tests", "golden", "input_model.hmm")


def read_hmm_file(path):
    """(type, transitions, emissions) of a `.hmm` file (cPecan/impl/stateMachine.c:145-202): the CPU arm parses it
    itself so that it imports nothing from the product."""
    with open(path) as fh:
        l1 = fh.readline().split()
        l2 = fh.readline().split()
    return int(l1[0]), np.array([float(v) for v in l1[1:-1]]), np.array([float(v) for v in l2])


def _ref_init(model_path, params_tuple):
    from oracle import pyref
    _W["model"] = pyref.RefModel.from_hmm(*read_hmm_file(model_path))
    _W["pyref"] = pyref
    _W["params"] = params_tuple


def _port_init(model_path, params_tuple):
    from oracle import pyoracle
    _W["model"] = pyoracle.model_from_hmm(*read_hmm_file(model_path))
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


def _pattern_work(item):
    """The reference as a user runs it (src/margin/marginAlignLib.py:315-318, marginCallerLib.py:54-57,
    cPecanEm.py:178-180): ONE cPecanRealign process per read, FASTA files and the exonerate cigar written first."""
    i, x, y, ops, lens, mode, exp, split, tmp, binary = item
    d = os.path.join(tmp, "w%d" % (i % 4096))
    os.makedirs(d, exist_ok=True)
    ref, read = os.path.join(d, "ref%d.fa" % i), os.path.join(d, "read%d.fa" % i)
    with open(ref, "w") as fh:
        fh.write(">ref\n%s\n" % x.decode())
    with open(read, "w") as fh:
        fh.write(">read\n%s\n" % y.decode())
    line = "cigar: read 0 %d + ref 0 %d + 1.0 %s\n" % (len(y), len(x), " ".join("%s %d" % ("MDI"[o], n)
                                                                                  for o, n in zip(ops, lens)))
    cmd = [binary, "--diagonalExpansion=%d" % exp, "--splitMatrixBiggerThanThis=%d" % split, "--loadHmm=" + MODEL_PATH]
    out = os.path.join(d, "out%d.txt" % i)
    if mode == "posterior":
        cmd.append("--outputAllPosteriorProbs=" + out)
    elif mode == "em":
        cmd.append("--outputExpectations=" + out)
    else:
        cmd += ["--gapGamma=0.5", "--matchGamma=0.0"]
    r = subprocess.run(cmd + [ref, read], input=line, capture_output=True, text=True)
    for f in (ref, read, out):
        if os.path.exists(f):
            os.remove(f)
    return r.returncode


def cpu_reference_pattern(batch, params, mode, n_sample, cells_per_read):
    """Process-per-read throughput of the unmodified CLI over all host cores; None when oracle/_ref is not built."""
    import tempfile
    from oracle import pyref
    binary = getattr(pyref, "REALIGN_BIN", os.path.join(ROOT, "oracle", "_ref", "cPecanRealign"))
    if not os.path.exists(binary):
        return None
    cores = os.cpu_count() or 1
    n_sample = min(n_sample, batch.n)
    with tempfile.TemporaryDirectory() as tmp:
        items = [(i, bytes(batch.x(i)), bytes(batch.y(i)), [int(o) for o in batch.cigar(i)[0]],
                  [int(n) for n in batch.cigar(i)[1]], mode, params.diagonal_expansion,
                  params.split_matrix_bigger_than, tmp, binary) for i in range(n_sample)]
        ctx = mp.get_context("fork")
        with ctx.Pool(cores) as pool:
            pool.map(_pattern_work, items[:cores], chunksize=1)  # page the binary in
            t0 = time.perf_counter()
            rcs = pool.map(_pattern_work, items, chunksize=max(1, n_sample // (cores * 8)))
            dt = time.perf_counter() - t0
    if any(rcs):
        return {"error": "cPecanRealign exited non-zero on %d of %d reads" % (sum(1 for r in rcs if r), n_sample)}
    return {"value": cells_per_read * n_sample / dt / 1e9, "unit": "GCUPS", "reads_per_s": n_sample / dt, "cores": cores,
            "sample": "first %d alignments, one oracle/_ref/cPecanRealign process per read (FASTA + cigar files "
                      "written per read), %d at a time" % (n_sample, cores)}


def cpu_reference_throughput(batch, params, mode, n_sample, steps=1, warmup=0, pattern_sample=0):
    """Times the reference C path (oracle/_ref if built, else our C port) over all host cores on the first n_sample
    alignments; returns dict(value GCUPS, cores, kind, sample, ms_per_step).  Nothing of the product is used: anchors
    come from the oracle's restatement of the cigar -> anchor rule, the cell count from the oracle's band walk."""
    from oracle import pyoracle as po
    from oracle import pyref
    kind = "reference" if pyref.available() else "port"
    cores = os.cpu_count() or 1
    n_sample = min(n_sample, batch.n)
    oP = po.make_params(expansion=params.diagonal_expansion, split=params.split_matrix_bigger_than ** 2)
    items, cells = [], 0
    for i in range(n_sample):
        x, y = batch.x(i), batch.y(i)
        anc = po.anchors_from_cigar(*batch.cigar(i), x, y)
        cells += int(po.count_cells(len(x), len(y), anc, oP))
        items.append((x, y, anc, mode))
    ptuple = (params.diagonal_expansion, params.split_matrix_bigger_than ** 2)
    ctx = mp.get_context("fork")
    with ctx.Pool(cores, initializer=_ref_init if kind == "reference" else _port_init,
                  initargs=(MODEL_PATH, ptuple)) as pool:
        chunk = max(1, len(items) // (cores * 8))
        for _ in range(warmup):
            pool.map(_ref_work, items, chunksize=chunk)
        t0 = time.perf_counter()
        for _ in range(steps):
            pool.map(_ref_work, items, chunksize=chunk)
        dt = (time.perf_counter() - t0) / max(1, steps)
    out = dict(value=cells / dt / 1e9, unit="GCUPS", cores=cores, kind=kind, ms_per_step=dt * 1e3,
               reads_per_s=n_sample / dt, cells=int(cells),
               sample="first %d alignments of the workload (%d cells), %s mode, one %s call per alignment in a "
                      "%d-process pool" % (n_sample, cells, mode, {"posterior": "getAlignedPairsUsingAnchors",
                                            "realign": "getAlignedPairsUsingAnchors + reweightAlignedPairs2 + "
                                                       "filterPairwiseAlignmentToMakePairsOrdered",
                                            "em": "getExpectationsUsingAnchors"}[mode], cores))
    if pattern_sample:
        out["pattern"] = cpu_reference_pattern(batch, params, mode, pattern_sample, cells / max(1, n_sample))
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return
    batch, params, name = make_workload(args, 0)
    n_sample = args.ref_sample or (20000 if args.workload == "trna" else 16)
    r = cpu_reference_throughput(batch, params, args.mode, n_sample, steps=args.steps, warmup=args.warmup,
                                 pattern_sample=(2000 if args.workload == "trna" else 16))
    line = {
        "impl": "reference", "metric": "pair-HMM cell updates per second (%s mode)" % args.mode,
        "value": r["value"], "unit": "GCUPS", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "mode": args.mode, "diagonal_expansion": params.diagonal_expansion,
                   "model": "fiveStateAsymmetric (tests/input.hmm values)", "sample_alignments": n_sample,
                   "same_config": "a %d-alignment sample of the same workload per step (GCUPS and reads/s are "
                                  "rates, so the sample size cancels)" % n_sample},
        "reads_per_s": r["reads_per_s"],
        "cpu_baseline": {"value": r["value"], "unit": "GCUPS", "cores": r["cores"], "kind": r["kind"],
                         "sample": r["sample"], "pattern": r.get("pattern")},
        "e2e": {"value": r["value"], "unit": "GCUPS", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ our arm

def load_peaks():
    """HBM peak from the driver-written MEASURED_PEAKS.json; FP64 non-FMA issue peak from the pipe microbenchmark of
    this repo (tools/pipe_peaks.cu, run on a B200 of this pool; newest profiles/*pipe_peaks.json)."""
    import glob
    peaks, src = {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback (B200_PROFILING.md)"
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            peaks, src = json.load(fh), "measured (MEASURED_PEAKS.json)"
    fp64 = None
    for f in sorted(glob.glob(os.path.join(ROOT, "profiles", "*pipe_peaks.json"))):
        with open(f) as fh:
            d = json.load(fh)
        if "dadd_tops" in d:
            fp64 = (float(d["dadd_tops"]), os.path.relpath(f, ROOT))
    return peaks, src, fp64


def run_ours(args, rank, world, local_rank):
    import torch
    import marginalign_trna_b200 as mab
    from marginalign_trna_b200 import capi
    from marginalign_trna_b200.pairhmm import DevicePlan
    L = capi.load()
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device (no CPU fallback on this path)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    cores_mine = mab.shard.bind_rank_to_local_cores(local_rank, local_rank, world) if world > 1 else None
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)

    batch, params, name = make_workload(args, rank)
    hmm = mab.Hmm.loadHmm(os.path.join(ROOT, "tests", "golden", "input_model.hmm"))
    model = mab.PairHmmModel.from_hmm(hmm) if args.states == 5 else mab.PairHmmModel.default(2)
    S = model.state_number
    eng = mab.PairHmmEngine(device=dev, max_cells=args.max_cells or None, plan_ahead=args.plan_ahead or None,
                            compact_pairs=not args.int32_pairs,
                            chunk_ramp=[int(v) for v in args.chunk_ramp.split(",")] if args.chunk_ramp else None)
    em = args.mode == "em"
    realign = args.mode == "realign"
    stream = torch.cuda.current_stream(dev)
    sptr = C.c_void_p(stream.cuda_stream)

    # ---- resident setup: the raw inputs of every chunk (bases, offsets, guide cigars) are copied into HBM once; a
    # first planning pass sizes the shared workspace and the output buffers.  Every timed step then plans each chunk
    # on the GPU (mab_plan_build_device) and runs the DP on it.
    chunks = eng._chunks(batch, params, em)
    packed, d_ins = [], []
    for b in chunks:
        pk = DevicePlan.pack_inputs(batch.slice(*b))
        packed.append(pk)
        d_ins.append(DevicePlan.upload_inputs(pk, dev))
    torch.cuda.synchronize()

    def plan_chunk(k):
        return DevicePlan.build_on_device(None, params, dev, em, packed=packed[k], alloc_ws=False, d_in=d_ins[k])

    cells, ws_need, max_n, max_cells_chunk, rl = 0, 0, 0, 0, []
    for k, b in enumerate(chunks):
        dp = plan_chunk(k)
        info = dp.plan.info
        cells += info.n_cells
        ws_need = max(ws_need, info.workspace_bytes_expectations if em else info.workspace_bytes_posteriors)
        max_n = max(max_n, info.n_alignments)
        max_cells_chunk = max(max_cells_chunk, info.n_cells)
        if realign:
            rl.append((info.n_alignments, dp.input_offsets))
        del dp
    n_reads = batch.n
    d_ws = torch.empty(ws_need + 256, dtype=torch.uint8, device=dev)
    ws_ptr = (d_ws.data_ptr() + 255) & ~255
    d_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)
    d_np = torch.zeros(1, dtype=torch.int64, device=dev)
    h_np = torch.zeros(1, dtype=torch.int64, pin_memory=True)
    # output capacity: generous fixed bound so the loop has no data-dependent host sync (pairs/cell ~ 0.07)
    cap = int(max_cells_chunk * 0.25) + 1024
    d_x = torch.empty(cap, dtype=torch.int32, device=dev)
    d_y = torch.empty(cap, dtype=torch.int32, device=dev)
    d_p = torch.empty(cap, dtype=torch.int32, device=dev)
    init = np.full(model.n_expectations, 1e-12)
    init[-1] = 0.0
    d_exp = torch.from_numpy(init).to(dev)
    if realign:  # shared: jitter stream, scratch, cigar outputs
        d_rand = torch.from_numpy(mab.jitter_table(1 << 22)).to(dev)
        rl_ws = max(int(L.mab_realign_workspace_bytes(n, cap, o[2], o[3])) for n, o in rl)
        d_rl_ws = torch.empty(rl_ws + 256, dtype=torch.uint8, device=dev)
        rl_ptr = (d_rl_ws.data_ptr() + 255) & ~255
        cap_ops = max(2 * (o[2] + o[3]) + n for n, o in rl)
        d_ops = torch.empty(cap_ops, dtype=torch.uint8, device=dev)
        d_lens = torch.empty(cap_ops, dtype=torch.int32, device=dev)
        d_op_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)
        d_ch_off = torch.empty(max_n + 1, dtype=torch.int64, device=dev)

    # The timed region runs planner and DP kernels of every chunk back to back on one stream, so that the per-kernel
    # CUDA-event durations behind `roofline` are those of kernels that have the GPU to themselves.  pipelined = True
    # (the `pipelined` leg) queues the planning of chunk k + 1 on a second stream while the DP kernels of chunk k run --
    # what PairHmmEngine does for host callers; the planner's two host round trips then block only that stream.  A
    # plan's arena is freed only after the DP that reads it has finished.
    side_stream = torch.cuda.Stream(device=dev)

    def resident_step(pipelined=False):
        plan_stream = side_stream if pipelined else stream

        def plan_on_side(k):
            with torch.cuda.stream(plan_stream):
                dp = plan_chunk(k)
                ev = torch.cuda.Event()
                ev.record(plan_stream)
            return dp, ev

        nxt = plan_on_side(0)
        inflight = []
        for k in range(len(chunks)):
            dp, planned = nxt
            stream.wait_event(planned)
            ph, blob = dp.plan._h, dp.blob_ptr
            if em:
                capi.check(L.mab_expectations(model._h, ph, blob, ws_ptr, ws_need, d_exp.data_ptr(), sptr))
            else:
                capi.check(L.mab_posteriors_compute(model._h, ph, blob, ws_ptr, ws_need, d_off.data_ptr(),
                                                    d_np.data_ptr(), sptr))
                capi.check(L.mab_posteriors_emit(ph, blob, ws_ptr, ws_need, d_x.data_ptr(), d_y.data_ptr(),
                                                 d_p.data_ptr(), cap, sptr))
                if realign:
                    n, (ox, oy, lx, ly) = rl[k]
                    capi.check(L.mab_realign_chain(n, ox, oy, lx, ly, d_off.data_ptr(), d_x.data_ptr(), d_y.data_ptr(),
                                                   d_p.data_ptr(), cap, 0.5, 0.0, d_rand.data_ptr(), d_rand.numel(),
                                                   rl_ptr, rl_ws, d_op_off.data_ptr(), d_ch_off.data_ptr(), sptr))
                    capi.check(L.mab_realign_emit(n, ox, oy, lx, ly, cap, rl_ptr, rl_ws, d_op_off.data_ptr(),
                                                  d_ch_off.data_ptr(), d_ops.data_ptr(), d_lens.data_ptr(), None, None,
                                                  sptr))
            done = torch.cuda.Event()
            done.record(stream)
            inflight.append((dp, done))
            del dp
            if k + 1 < len(chunks):
                nxt = plan_on_side(k + 1)
            while len(inflight) > 1:
                inflight.pop(0)[1].synchronize()
        inflight.pop(0)[1].synchronize()
        if em and dist is not None:
            dist.all_reduce(d_exp)  # the path's only collective: 106 (58) doubles per EM iteration

    def barrier():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def max_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if dist is not None:
            dist.all_reduce(t)
        return float(t.item())

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

    ms_step = max_over_ranks(ms_total) / args.steps
    total_cells = sum_over_ranks(cells) if dist is not None else float(cells)
    value = total_cells / (ms_step * 1e-3) / 1e9

    # ---- the same resident loop with planning overlapped (second stream), no per-kernel events
    pipelined_leg = None
    if not args.no_pipelined_leg:
        resident_step(True)
        barrier()
        ev0.record(stream)
        for _ in range(args.steps):
            resident_step(True)
        ev1.record(stream)
        barrier()
        ms_pipe = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
        pipelined_leg = {"value": total_cells / (ms_pipe * 1e-3) / 1e9, "unit": "GCUPS", "ms_per_step": ms_pipe,
                         "steps": args.steps,
                         "what": "the timed loop again with the planning of chunk k+1 queued on a second stream while "
                                 "chunk k computes (PairHmmEngine's pipeline); same work, same buffers"}

    # ---- e2e through the public API (host in, host out)
    e2e = None
    e2e_steps = max(1, min(args.e2e_steps, args.steps))
    run_e2e = (lambda: eng.get_expectations_using_anchors(model, batch, params)) if em else \
              (lambda: eng.realign(model, batch, params)) if realign else \
              (lambda: eng.get_aligned_pairs_using_anchors(model, batch, params))
    if realign:
        del d_rl_ws, d_ops, d_lens
    del d_x, d_y, d_p, d_ws, d_ins, packed
    torch.cuda.empty_cache()
    if not args.no_e2e:
        for _ in range(max(1, min(args.warmup, 3))):
            run_e2e()  # warm-up: the device and page-locked allocator pools reach their steady state in 2-3 calls
        torch.cuda.synchronize()
        barrier()
        ms0 = torch.cuda.memory_stats(dev)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            run_e2e()
        torch.cuda.synchronize()
        dt = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
        ms1 = torch.cuda.memory_stats(dev)
        st = dict(eng.last_stats)
        e2e = {"value": total_cells / dt / 1e9, "unit": "GCUPS", "h2d_bytes_per_step": int(st["h2d_bytes"]),
               # device allocator traffic inside the timed calls (cudaMalloc / cudaFree synchronise the device)
               "allocator": {k: int(ms1.get(k, 0) - ms0.get(k, 0)) for k in
                             ("num_device_alloc", "num_device_free", "num_alloc_retries")},
               "chunks": int(st.get("chunks", 0)),
               "d2h_bytes_per_step": int(st["d2h_bytes"]), "ms_per_step": dt * 1e3, "steps": e2e_steps,
               "reads_per_s": n_reads * world / dt, "host_wall_s": st.get("host_wall"),
               "note": "packing of bases+cigars into pinned memory, H2D, planning on the GPU, DP kernels, %s, per step "
                       "(rank 0 breakdown in host_wall_s)" % (
                           "realign tail on the GPU, D2H of the cigars" if realign else
                           "D2H of the expectation vector" if em else "D2H of the pairs into pinned host blocks")}

    # ---- realign mode end to end (what marginAlign does with every read: posteriors + AMAP reweight + MEA chain + cigar,
    # all on the GPU; only cigars come back, ~0.1 GB per 1 M reads instead of 2.2 GB of pairs)
    realign_leg = None
    if not args.no_realign_leg and args.workload == "trna" and not em and not realign:
        torch.cuda.empty_cache()
        for _ in range(2):
            eng.realign(model, batch, params)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            eng.realign(model, batch, params)
        torch.cuda.synchronize()
        dt = max_over_ranks((time.perf_counter() - t0) / e2e_steps)
        st = dict(eng.last_stats)
        realign_leg = {"value": total_cells / dt / 1e9, "unit": "GCUPS", "reads_per_s": n_reads * world / dt,
                       "ms_per_step": dt * 1e3, "steps": e2e_steps, "h2d_bytes_per_step": int(st["h2d_bytes"]),
                       "d2h_bytes_per_step": int(st["d2h_bytes"]),
                       "note": "PairHmmEngine.realign through the public API, host buffers in, cigars out"}

    # ---- strong scaling through the product's sharder: ONE batch (rank 0's), split by ShardedPairHmm.shards()
    strong = None
    if not args.no_strong_leg and args.workload == "trna" and not em and not realign:
        sb = batch if rank == 0 or world == 1 else make_workload(args, 0)[0]
        sh = mab.shard.ShardedPairHmm(eng)
        torch.cuda.empty_cache()
        res = sh.get_aligned_pairs_using_anchors(model, sb, params)  # warm-up
        del res  # its pinned blocks go back to the allocator: the timed call must not page-lock fresh memory
        barrier()
        t0 = time.perf_counter()
        res = sh.get_aligned_pairs_using_anchors(model, sb, params)
        torch.cuda.synchronize()
        dt = max_over_ranks(time.perf_counter() - t0)
        sc = sum_over_ranks(eng.last_stats["cells"]) if dist is not None else float(eng.last_stats["cells"])
        phases = getattr(sh, "last_phases", None)
        # the same batch in realign mode: only cigars travel to rank 0
        cig = sh.realign(model, sb, params)
        del cig
        barrier()
        t0 = time.perf_counter()
        cig = sh.realign(model, sb, params)
        torch.cuda.synchronize()
        dt_rl = max_over_ranks(time.perf_counter() - t0)
        strong = {"alignments": sb.n, "cells": sc, "s": dt, "value": sc / dt / 1e9, "unit": "GCUPS",
                  "reads_per_s": sb.n / dt, "phases_s": phases,
                  "realign": {"s": dt_rl, "value": sc / dt_rl / 1e9, "unit": "GCUPS", "reads_per_s": sb.n / dt_rl,
                              "phases_s": getattr(sh, "last_phases", None)},
                  "pairs_on_rank0": int(res.offsets[-1]) if res is not None else None,
                  "note": "ShardedPairHmm.get_aligned_pairs_using_anchors: every rank runs its shard through the public "
                          "API, results gathered to rank 0 and merged in input order; wall clock, max over ranks"}

    torch.cuda.empty_cache()
    # ---- seconds per EM iteration (BASELINE.json metric), the all-reduce and the M-step inside the timed region
    em_leg = None
    if not args.no_em_leg and args.workload == "trna":
        n_em = min(batch.n, args.em_reads)
        em_batch = batch if n_em == batch.n else batch.slice(0, n_em)
        em_model, cur = model, mab.Hmm(hmm.modelType if S == 5 else "threeState")
        em_cells = 0

        # the band plans do not depend on the model: like em.expectation_maximisation, the loop keeps them in HBM
        # (PlanCache); the first iteration, which plans, is timed separately
        em_cache = mab.PlanCache()

        def em_iteration():
            nonlocal em_model, em_cells
            T, E, lik = eng.get_expectations_using_anchors(em_model, em_batch, params, pseudo_count=0.0,
                                                           plan_cache=em_cache)
            em_cells = eng.last_stats["cells"]
            vec = torch.from_numpy(np.concatenate([T, E, [lik]])).to(dev)
            if dist is not None:
                dist.all_reduce(vec)  # NCCL: S*S + 16 S + 1 doubles
            v = vec.cpu().numpy()
            new = mab.Hmm(cur.modelType)
            new.addExpectations(v[:S * S] + 1e-12, v[S * S:S * S + 16 * S] + 1e-12, v[-1])
            new.normalise()  # the M-step, identical on every rank (cPecanEm.py:182-202)
            em_model = mab.PairHmmModel.from_hmm(new)

        em_iteration()  # warm-up (allocator, pinned buffers); its plans are dropped again
        em_cache.clear()
        barrier()
        t0 = time.perf_counter()
        em_iteration()  # iteration 0 of a run: packs, uploads and plans the batch
        torch.cuda.synchronize()
        dt_first = max_over_ranks(time.perf_counter() - t0)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.em_iterations):
            em_iteration()
        torch.cuda.synchronize()
        dt = max_over_ranks((time.perf_counter() - t0) / args.em_iterations)
        tot = sum_over_ranks(em_cells) if dist is not None else float(em_cells)
        em_leg = {"s_per_iteration": dt, "first_iteration_s": dt_first, "plan_cache_hits": em_cache.hits,
                  "iterations": args.em_iterations, "alignments_per_gpu": n_em,
                  "cells_per_iteration": tot, "value": tot / dt / 1e9, "unit": "GCUPS", "scaling": "weak",
                  "collective": "NCCL all-reduce of %d doubles per iteration" % (S * S + 16 * S + 1) if dist is not None
                  else "none (one rank)"}

    if rank != 0:
        return
    peaks, peaks_src, fp64 = load_peaks()
    # roofline of the dominant kernel: algorithmic flops / device time from CUDA events on the launch stream
    dom = max(("k_forward", "k_backward"), key=lambda k: kernel_ms.get(k, (0, 0))[0])
    dom_ms, dom_n = kernel_ms.get(dom, (0.0, 0))
    flops_cell_kernel = (130 if S == 5 else 90)
    roof = None
    if dom_n:
        per_step_ms = dom_ms / args.steps          # all launches of the kernel in one step cover every cell once
        achieved = cells * flops_cell_kernel / (per_step_ms * 1e-3) / 1e12
        smax = (clocks or {}).get("sm_max_mhz") or peaks.get("sm_max_mhz") or 1965.0
        if fp64:
            peak, peak_src = fp64[0], "measured non-FMA FP64 issue rate (DADD, tools/pipe_peaks.cu -> %s)" % fp64[1]
        else:
            peak = 148 * 64 * smax * 1e6 / 1e12
            peak_src = "nominal non-FMA FP64 issue rate 148 SM x 64 lanes x %.0f MHz" % smax
        # algorithmic HBM bytes per cell of one DP kernel: posterior mode writes the match plane (8 B) and
        # all S states on the diagonals that carry a total (1 in 10); expectation mode writes all S states
        bytes_cell = S * 8 if em else 8 + S * 8 / 10.0
        roof = {"bound": "fp64 (non-FMA issue; the bit-exact recurrence cannot contract into FMA)", "kernel": dom,
                "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                # dram__bytes_read.sum + dram__bytes_write.sum of this kernel from the ncu --set full capture, per
                # cell (profiles/: r2 capture of k_backward_rr), scaled to this run's cells per step
                "traffic": (cells * RR_TRAFFIC_B_PER_CELL if (S == 5 and not em and args.workload == "trna") else None),
                "peak_source": peak_src, "flops_per_cell": flops_cell_kernel, "ms_per_step": per_step_ms,
                "launches_per_step": dom_n / args.steps,
                "hbm": {"achieved": cells * bytes_cell / (per_step_ms * 1e-3) / 1e9,
                        "peak": peaks["hbm_gbs"], "unit": "GB/s", "peak_source": peaks_src,
                        "bytes_per_cell": bytes_cell}}
    cpu = None
    if not args.no_cpu_baseline and world == 1:
        n_sample = args.ref_sample or (20000 if args.workload == "trna" else 16)
        r = cpu_reference_throughput(batch, params, args.mode, n_sample,
                                     pattern_sample=(2000 if args.workload == "trna" else 8))
        cpu = {"value": r["value"], "unit": "GCUPS", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
               "reads_per_s": r["reads_per_s"], "pattern": r.get("pattern")}
    line = {
        "metric": "pair-HMM cell updates per second (%s mode)" % args.mode, "value": value, "unit": "GCUPS",
        "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "mode": args.mode, "alignments_per_gpu": n_reads, "cells_per_gpu": int(cells),
                   "diagonal_expansion": params.diagonal_expansion, "chunks": len(chunks),
                   "model": "fiveStateAsymmetric (tests/input.hmm values)" if S == 5 else "threeState (built-in)",
                   "value_includes": "GPU planning of every chunk (anchors, split regions, bands, schedule) + DP + "
                                     "ordered pair emission, back to back on one stream; inputs (bases, cigars) "
                                     "resident in HBM",
                   "l2": "inputs larger than L2: %.1f GB of DP state per step" % (cells * (80 if em else 28) / 1e9),
                   "parallelism": "reads sharded, dp%d" % world, "host_cores_per_rank": cores_mine},
        "reads_per_s": n_reads * world / (ms_step * 1e-3),
        "clocks": clocks, "e2e": e2e, "e2e_realign": realign_leg, "em": em_leg, "strong": strong, "pipelined": pipelined_leg,
        "gpu_launches": launches,
        "kernel_ms_per_step": {k: v[0] / args.steps for k, v in kernel_ms.items()},
        "roofline": roof, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)


def main():
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
