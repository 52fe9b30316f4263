"""Host planner (csrc/plan.cpp) and model builder (csrc/model.cpp) against the oracle; C-ABI surface checks.
No compute calls: this file runs without a GPU."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import ROOT, oracle_anchors, oracle_model, oracle_params
from marginalign_trna_b200 import capi
from oracle import pyoracle as po


def _params(p):
    return mab.AlignmentParameters(threshold=p["threshold"], min_diags_between_traceback=p["min_diags"],
                                   traceback_diagonals=p["traceback_diagonals"], diagonal_expansion=p["expansion"],
                                   split_matrix_bigger_than=p["split"], ragged_left=bool(p["ragged_left"]),
                                   ragged_right=bool(p["ragged_right"]))


def plan_for_case(c):
    P = _params(c["params"])
    if "anchors" in c:
        return mab.Plan.from_anchors([c["x"]], [c["y"]], [c["anchors"]], P)
    return mab.Plan.from_batch(mab.AlignmentBatch.from_lists([c["x"]], [c["y"]], [(c["ops"], c["lens"])]), P)


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "margin_b200.h")).read()
    declared = set(re.findall(r"\b(mab_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 25
    L = capi.load()
    for name in sorted(declared):
        assert hasattr(L, name), "libmargin_b200.so does not export %s" % name
    assert declared == set(capi.SIGNATURES), declared ^ set(capi.SIGNATURES)
    assert b"margin_b200" in L.mab_version()


def test_plan_matches_oracle_geometry(golden):
    for c in golden.cases:
        plan = plan_for_case(c)
        anc = oracle_anchors(po, c)
        assert np.array_equal(plan.anchors(0), anc), c["name"]
        p = c["params"]
        sp = po.split_points(anc, len(c["x"]), len(c["y"]), p["split"] ** 2, p["ragged_left"], p["ragged_right"])
        sp = sp[(sp[:, 2] - sp[:, 0]) + (sp[:, 3] - sp[:, 1]) > 0] if len(sp) else sp  # empty regions are dropped
        reg = plan.regions()
        assert np.array_equal(reg[:, 1:], sp.reshape(-1, 4)), c["name"]
        cells = 0
        j = 0
        for r, (x1, y1, x2, y2) in enumerate(sp.tolist()):
            sub = []
            while j < len(anc) and anc[j, 0] + anc[j, 1] < x2 + y2:
                sub.append([anc[j, 0] - x1, anc[j, 1] - y1])
                j += 1
            L, R = po.band(sub, x2 - x1, y2 - y1, p["expansion"])
            bl, br = plan.band(r)
            assert np.array_equal(bl, L) and np.array_equal(br, R), (c["name"], r)
            cells += int(((R - L) // 2 + 1).sum())
        assert plan.info.n_cells == cells
        assert cells == po.count_cells(len(c["x"]), len(c["y"]), anc, oracle_params(po, p), p["ragged_left"],
                                       p["ragged_right"])


def test_traceback_schedule():
    # 2000 diagonals, contiguous anchors, expansion 10: checkpoints at 1000 and 1959 (= 959 + 1000), end at 2000
    x = "ACGT" * 250
    plan = mab.Plan.from_anchors([x], [x], [[[i, i] for i in range(1000)]], mab.AlignmentParameters())
    tb = plan.tracebacks()
    assert tb.tolist() == [[0, 0, 1000, 959, 0], [0, 959, 1959, 1918, 0], [0, 1918, 2000, 2000, 1]]
    # posterior diagonals partition 1..D; one total per 10 of them
    assert plan.info.n_totals == sum(-(-(f - lo) // 10) for _, lo, _, f, _ in tb.tolist())


def test_plan_is_thread_count_independent():
    b = mab.synth.make_trna_batch(3000, seed=3)
    P = mab.AlignmentParameters()
    blobs = []
    for nt in (1, 3, 8):
        plan = mab.Plan.from_batch(b, P, n_threads=nt)
        buf = np.zeros(plan.info.blob_bytes, dtype=np.uint8)
        plan.write_blob(buf.ctypes.data, buf.nbytes)
        blobs.append(buf)
    assert np.array_equal(blobs[0], blobs[1]) and np.array_equal(blobs[0], blobs[2])


def test_plan_rejects_bad_input():
    P = mab.AlignmentParameters()
    with pytest.raises(RuntimeError, match="does not span"):
        mab.Plan.from_batch(mab.AlignmentBatch.from_lists(["ACGT"], ["ACGT"], [([0], [3])]), P)
    with pytest.raises(RuntimeError, match="runs past"):
        mab.Plan.from_batch(mab.AlignmentBatch.from_lists(["ACGT"], ["ACGT"], [([0], [9])]), P)
    with pytest.raises(RuntimeError, match="anchor"):
        mab.Plan.from_anchors(["ACGT"], ["ACGT"], [[[2, 2], [1, 3]]], P)
    with pytest.raises(RuntimeError, match="invalid parameters"):
        mab.Plan.from_anchors(["ACGT"], ["ACGT"], [[]], mab.AlignmentParameters(diagonal_expansion=3))
    with pytest.raises(RuntimeError, match="invalid parameters"):
        mab.Plan.from_anchors(["ACGT"], ["ACGT"], [[]], mab.AlignmentParameters(min_diags_between_traceback=41))


def test_empty_batch_and_empty_sequences():
    P = mab.AlignmentParameters()
    plan = mab.Plan.from_batch(mab.AlignmentBatch.from_lists([], [], []), P)
    assert plan.info.n_alignments == 0 and plan.info.n_regions == 0 and plan.info.n_cells == 0
    plan = mab.Plan.from_anchors(["", "ACG"], ["", ""], [[], []], P)
    assert plan.info.n_alignments == 2 and plan.info.n_regions == 1  # the 0x0 alignment has no diagonals


def test_model_block_matches_oracle(golden):
    for key in golden.models:
        spec = golden.model_arrays(key)
        om = oracle_model(po, spec)
        gm = mab.PairHmmModel.default(0) if spec is None else mab.PairHmmModel.from_arrays(*spec)
        assert gm.state_number == om.S
        want = np.array(po.model_transitions(om))
        assert np.array_equal(gm.transitions(), want), key
        for which, name in enumerate(["start", "ragged_start", "end", "ragged_end"]):
            assert np.array_equal(gm.boundary(which), np.array(getattr(om, name)[:om.S])), (key, name)
        m, gx, gy = gm.emissions()
        assert np.array_equal(m, np.array(om.em_match[:]))
        assert np.array_equal(gx, np.array(om.em_gap_x[:])) and np.array_equal(gy, np.array(om.em_gap_y[:]))


def test_model_rejects_bad_type():
    with pytest.raises(RuntimeError):
        mab.PairHmmModel.from_arrays(1, np.ones(9), np.ones(48))
    L = capi.load()
    h = capi.vp()
    assert L.mab_model_default(7, C.byref(h)) == -1
    assert b"bad hmm type" in L.mab_last_error()


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mab.PairHmmEngine()
    # the C ABI itself refuses too
    L = capi.load()
    m = mab.PairHmmModel.default(0)
    plan = mab.Plan.from_anchors(["ACGT"], ["ACGT"], [[]], mab.AlignmentParameters())
    rc = L.mab_posteriors_compute(m._h, plan._h, 1, 256, 1 << 30, 8, 8, None)
    assert rc == -5 and b"no CPU fallback" in L.mab_last_error()


def test_chunker_tiles_the_batch_in_order():
    """PairHmmEngine._chunks is host logic: ranges tile [0, n) in order, the first and (posterior mode) the last chunk
    are short so the pipeline fills and drains quickly, and no chunk exceeds the cell budget by more than one read."""
    from marginalign_trna_b200 import synth
    from marginalign_trna_b200.pairhmm import PairHmmEngine
    batch = synth.trna_batch(3000, seed=5) if hasattr(synth, "trna_batch") else None
    if batch is None:
        rng = np.random.default_rng(5)
        xs = ["".join(rng.choice(list("ACGT"), int(n))) for n in rng.integers(70, 150, 3000)]
        batch = mab.AlignmentBatch.from_lists(xs, xs, [([0], [len(x)]) for x in xs])
    eng = object.__new__(PairHmmEngine)  # no GPU: only the chunker is exercised
    P = mab.AlignmentParameters(diagonal_expansion=10)
    est = ((batch.off_x[1:] - batch.off_x[:-1]) + (batch.off_y[1:] - batch.off_y[:-1]) + 1) * 14
    for budget in (50_000, 400_000, 10 ** 9):
        eng.max_cells = budget
        for expectations in (False, True):
            bounds = eng._chunks(batch, P, expectations)
            assert bounds[0][0] == 0 and bounds[-1][1] == batch.n
            assert all(a < b for a, b in bounds) and all(bounds[i][1] == bounds[i + 1][0] for i in range(len(bounds) - 1))
            sizes = [int(est[a:b].sum()) for a, b in bounds]
            assert all(s <= budget + int(est.max()) for s in sizes)
            if len(bounds) >= 4:
                assert sizes[0] <= budget // 4 + int(est.max())
                if not expectations:
                    assert sizes[-1] <= budget // 2 + int(est.max())
    eng.max_cells = 1000
    assert eng._chunks(mab.AlignmentBatch.from_lists([], [], []), P) == [(0, 0)]


def _blob(plan):
    raw = np.zeros(plan.info.blob_bytes, dtype=np.uint8)
    plan.write_blob(raw.ctypes.data, raw.nbytes)
    return raw, raw[:256].view(np.int64)


def test_packed_symbols_match_the_symbol_codes():
    """2-bit codes + N mask (BlobHeader::off_pack_x; symbol_convertStringToSymbols pairwiseAligner.c:281-289 as bits)
    decode to the byte codes the other kernels read, N and lower-case bases included."""
    b = mab.AlignmentBatch.from_lists(["ACGTNacgtnRYACGT" * 5, "TTGACCA"], ["ACGTTNNNacgtACGT" * 5, "TTGNCCA"],
                                      [([0], [80]), ([0], [7])])
    raw, h = _blob(mab.Plan.from_batch(b, mab.AlignmentParameters()))
    for off_sym, off_pack, n in ((int(h[15]), int(h[30]), int(h[9])), (int(h[16]), int(h[31]), int(h[10]))):
        sym = raw[off_sym:off_sym + n]
        cw, mw = (n + 512 + 15) // 16, (n + 512 + 31) // 32
        words = raw[off_pack:off_pack + 4 * (cw + mw)].view(np.uint32)
        pos = np.arange(n) + 256
        code = (words[pos >> 4] >> (2 * (pos & 15)).astype(np.uint32)) & 3
        mask = (words[cw + (pos >> 5)] >> (pos & 31).astype(np.uint32)) & 1
        assert np.array_equal(np.where(mask == 1, 4, code), sym)
        # nothing outside the sequence is set
        total = np.unpackbits(words.view(np.uint8)).sum()
        assert total == np.unpackbits(code.astype(np.uint8)).sum() + mask.sum()


def test_band_edges_move_by_at_most_one_base_per_diagonal():
    """What the register kernels' symbol rings rely on (csrc/wavefront_reg.cuh): along a region the lowest / highest x
    and y of the band never decrease and grow by at most one per anti-diagonal (band_construct,
    pairwiseAligner.c:128-179: the rectangles of successive anchor pairs are nested in both coordinates)."""
    batches = [(mab.synth.make_trna_batch(300, seed=3), mab.AlignmentParameters()),
               (mab.synth.make_long_batch(3, read_len=3000, genome_len=40_000, seed=4),
                mab.AlignmentParameters(split_matrix_bigger_than=300)),
               (mab.synth.make_long_batch(2, read_len=2000, genome_len=40_000, seed=5, p_ins=0.12, p_del=0.12),
                mab.AlignmentParameters(diagonal_expansion=20, split_matrix_bigger_than=30))]
    for b, P in batches:
        plan = mab.Plan.from_batch(b, P)
        for r in range(plan.info.n_regions):
            lo, hi = plan.band(r)
            d = np.arange(len(lo))
            x_lo, x_hi = (d + lo) // 2, (d + hi) // 2
            y_lo, y_hi = (d - hi) // 2, (d - lo) // 2
            for e in (x_lo, x_hi, y_lo, y_hi):
                step = np.diff(e)
                assert step.min() >= 0 and step.max() <= 1, (r, step.min(), step.max())
