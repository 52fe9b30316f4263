"""Parity of the CUDA path (through the C ABI) against the reference.

Bars (BASELINE.json north_star): posteriors and expectations within 1e-5 relative; aligned-pair sets identical
except for pairs whose posterior lies within tolerance of the threshold.  What we actually assert is tighter:
  * forward / backward / total-probability values: bit-identical to the oracle by construction, so
  * p_int differs from the reference by at most 1 unit of 1e-7 (CUDA exp vs glibc exp, last-ulp differences), and a
    pair may appear/disappear only when its probability is within 1e-7 of the threshold;
  * expectations: relative error <= 1e-9 on every entry (summation order differs), the spec allows 1e-5.
"""
import json
import os

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN, oracle_anchors, oracle_model, oracle_params
from oracle import pyoracle as po
from test_host_plan import _params

pytestmark = pytest.mark.gpu

P_INT_TOL = 1          # units of 1e-7
EXP_RTOL = 1e-9        # spec: 1e-5


@pytest.fixture(scope="module")
def engine():
    return mab.PairHmmEngine()


def gpu_model(spec):
    return mab.PairHmmModel.default(0) if spec is None else mab.PairHmmModel.from_arrays(*spec)


def assert_pairs_match(got, want, threshold, ctx):
    """got/want rows (p_int, x, y) in list order."""
    if len(got) == len(want) and np.array_equal(got[:, 1:], want[:, 1:]):
        assert np.abs(got[:, 0] - want[:, 0]).max(initial=0) <= P_INT_TOL, ctx
        return
    # tolerate pairs sitting on the threshold (within one p_int unit)
    thr = int(np.floor(threshold * 1e7))
    gd = {(int(x), int(y)): int(p) for p, x, y in got}
    wd = {(int(x), int(y)): int(p) for p, x, y in want}
    for k in set(gd) ^ set(wd):
        p = gd.get(k, wd.get(k))
        assert abs(p - thr) <= P_INT_TOL, (ctx, k, p)
    for k in set(gd) & set(wd):
        assert abs(gd[k] - wd[k]) <= P_INT_TOL, (ctx, k)
    # order of the common pairs is the reference's
    common = set(gd) & set(wd)
    assert [tuple(r[1:]) for r in got.tolist() if tuple(r[1:]) in common] == \
           [tuple(r[1:]) for r in want.tolist() if tuple(r[1:]) in common], ctx


def assert_exp_match(got, want, ctx, rtol=EXP_RTOL):
    got, want = np.asarray(got), np.asarray(want)
    assert got.shape == want.shape, ctx
    scale = np.maximum(np.abs(want), 1e-300)
    err = np.abs(got - want) / scale
    # entries that only hold the 1e-12 pseudo-count are exact
    assert err.max() <= rtol, (ctx, int(err.argmax()), got[err.argmax()], want[err.argmax()])


def batch_for_case(c):
    return mab.AlignmentBatch.from_lists([c["x"]], [c["y"]], [(c["ops"], c["lens"])])


def run_case_gpu(engine, c, model):
    P = _params(c["params"])
    if "anchors" in c:
        # explicit anchors go through mab_plan_from_anchors
        plan = mab.Plan.from_anchors([c["x"]], [c["y"]], [c["anchors"]], P)
        dplan = mab.DevicePlan(plan, engine.device)
        d_off, d_x, d_y, d_p = engine.posteriors_device(model, dplan)
        pairs = np.stack([d_p.cpu().numpy(), d_x.cpu().numpy(), d_y.cpu().numpy()], axis=1).astype(np.int64)
        import torch
        init = np.full(model.n_expectations, 1e-12)
        init[-1] = 0
        d_out = torch.from_numpy(init).to(engine.device)
        dplan2 = mab.DevicePlan(plan, engine.device, expectations=True)
        engine.expectations_device(model, dplan2, d_out)
        return pairs, d_out.cpu().numpy()
    b = batch_for_case(c)
    res = engine.get_aligned_pairs_using_anchors(model, b, P)
    T, E, lik = engine.get_expectations_using_anchors(model, b, P)
    return res.pairs(0), np.concatenate([T, E, [lik]])


def test_golden_cases_against_reference_outputs(engine, golden):
    """Every golden case (fixture reads, KAT example, all four model types, N/lower-case symbols, ragged and
    non-ragged ends, multi-traceback, sparse/no anchors, splits, empty inputs) vs the reference C code's outputs."""
    for c in golden.cases:
        model = gpu_model(golden.model_arrays(c["model"]))
        pairs, exp = run_case_gpu(engine, c, model)
        assert_pairs_match(pairs, golden.pairs(c["name"]), c["params"]["threshold"], c["name"])
        assert_exp_match(exp, golden.exp(c["name"]), c["name"])


def test_dp_matrices_bit_identical_to_reference_totals(engine, golden):
    """The likelihood entry is a plain sum of the per-diagonal total probabilities, which depend on every forward and
    backward cell: with bit-identical DP values it matches the reference to the last few ulps of the sum."""
    for name in ("fixture0_realign", "multi_tb", "trna_rand5sym_0", "trna_three_asym_1"):
        c = next(k for k in golden.cases if k["name"] == name)
        model = gpu_model(golden.model_arrays(c["model"]))
        _, exp = run_case_gpu(engine, c, model)
        want = golden.exp(name)[-1]
        assert abs(exp[-1] - want) <= 1e-13 * abs(want), (name, exp[-1], want)


def test_batched_trna_reads_against_oracle(engine, golden):
    """A batch of 300 synthetic tRNA-length reads in ONE launch vs the oracle read by read (5- and 3-state)."""
    b = mab.synth.make_trna_batch(300, seed=99)
    P = mab.AlignmentParameters()
    oP = po.make_params(expansion=10, split=3000 ** 2)
    for key in ("fixture5asym", "three_asym"):
        spec = golden.model_arrays(key)
        model, om = gpu_model(spec), oracle_model(po, spec)
        res = engine.get_aligned_pairs_using_anchors(model, b, P)
        T, E, lik = engine.get_expectations_using_anchors(model, b, P)
        S = om.S
        oT, oE, olik = np.full(S * S, 1e-12), np.full(S * 16, 1e-12), 0.0
        for i in range(b.n):
            anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
            want = po.aligned_pairs(om, b.x(i), b.y(i), anc, oP)
            assert_pairs_match(res.pairs(i), want, 0.01, (key, i))
            oT, oE, olik = po.expectations(om, b.x(i), b.y(i), anc, oP, T=oT, E=oE, lik=olik)
        assert_exp_match(np.concatenate([T, E, [lik]]), np.concatenate([oT, oE, [olik]]), key)


def test_bigmutations_reads_against_reference_digests(engine):
    """All 30 reads of the reference's inputBigMutations fixture (294-7214 nt), marginCaller parameters
    (expansion 10, split 100) for posteriors and EM parameters (split 300) for expectations."""
    with open(os.path.join(GOLDEN, "bigmutations_digest.json")) as fh:
        dig = json.load(fh)
    with open(os.path.join(GOLDEN, "bigmutations_alignments.json")) as fh:
        alns = json.load(fh)
    model = mab.PairHmmModel.from_file(os.path.join(GOLDEN, "input_model.hmm"))
    om = po.model_from_hmm(1, mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm")).transitions,
                           mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm")).emissions)
    b = mab.AlignmentBatch.from_lists([a["x"] for a in alns], [a["y"] for a in alns],
                                      [(a["ops"], a["lens"]) for a in alns])
    res = engine.get_aligned_pairs_using_anchors(model, b, mab.AlignmentParameters(split_matrix_bigger_than=100))
    for i, d in enumerate(dig):
        got = res.pairs(i)
        assert abs(len(got) - d["n_pairs"]) <= 2, (i, len(got), d["n_pairs"])
        assert abs(int(got[:, 0].sum()) - d["p_sum"]) <= len(got) + 2 * 100000, i
        if i < 6:  # full comparison against the oracle for a few reads
            anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
            want = po.aligned_pairs(om, b.x(i), b.y(i), anc, po.make_params(expansion=10, split=100 ** 2))
            assert_pairs_match(got, want, 0.01, i)
    # expectations: one read at a time against the stored reference vectors, then the whole batch against their sum
    Pem = mab.AlignmentParameters(split_matrix_bigger_than=300)
    total = np.zeros(106)
    for i, d in enumerate(dig):
        want = np.array(d["exp"])
        total += want - np.concatenate([np.full(105, 1e-12), [0.0]])
        if i % 5 == 0:
            T, E, lik = engine.get_expectations_using_anchors(model, b.subset([i]), Pem)
            assert_exp_match(np.concatenate([T, E, [lik]]), want, ("exp", i))
    T, E, lik = engine.get_expectations_using_anchors(model, b, Pem)
    assert_exp_match(np.concatenate([T, E, [lik]]), total + np.concatenate([np.full(105, 1e-12), [0.0]]), "batch",
                     rtol=1e-8)


def test_chunking_does_not_change_results(golden):
    b = mab.synth.make_trna_batch(200, seed=5)
    model = gpu_model(golden.model_arrays("fixture5asym"))
    one = mab.PairHmmEngine(max_cells=10 ** 9)
    many = mab.PairHmmEngine(max_cells=50_000)
    r1 = one.get_aligned_pairs_using_anchors(model, b)
    r2 = many.get_aligned_pairs_using_anchors(model, b)
    assert many.last_stats["chunks"] > 3 and one.last_stats["chunks"] == 1
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(r1, a), getattr(r2, a)), a
    e1 = one.get_expectations_using_anchors(model, b)
    e2 = many.get_expectations_using_anchors(model, b)
    assert_exp_match(np.concatenate([e1[0], e1[1], [e1[2]]]), np.concatenate([e2[0], e2[1], [e2[2]]]), "chunks",
                     rtol=1e-11)


def test_wide_band_and_long_read(engine, golden):
    """A 3 kb read with expansion 100 (diagonals up to ~100 cells wide, several chunks of 32 lanes per diagonal and
    multiple checkpoints)."""
    b = mab.synth.make_long_batch(2, read_len=3000, genome_len=50_000, seed=8)
    spec = golden.model_arrays("fixture5asym")
    model, om = gpu_model(spec), oracle_model(po, spec)
    P = mab.AlignmentParameters(diagonal_expansion=100, split_matrix_bigger_than=300)
    oP = po.make_params(expansion=100, split=300 ** 2)
    res = engine.get_aligned_pairs_using_anchors(model, b, P)
    T, E, lik = engine.get_expectations_using_anchors(model, b, P)
    oT, oE, olik = np.full(25, 1e-12), np.full(80, 1e-12), 0.0
    for i in range(b.n):
        anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
        assert_pairs_match(res.pairs(i), po.aligned_pairs(om, b.x(i), b.y(i), anc, oP), 0.01, i)
        oT, oE, olik = po.expectations(om, b.x(i), b.y(i), anc, oP, T=oT, E=oE, lik=olik)
    assert_exp_match(np.concatenate([T, E, [lik]]), np.concatenate([oT, oE, [olik]]), "wide")


def test_warp_group_size_does_not_change_results(engine, golden):
    """The 5-state wavefront kernels may give one work item to 1, 2, 4 or 8 warps (wide bands); every group size
    must produce bit-identical pairs and expectations, and the wide-band ones must match the oracle."""
    L = mab.capi.load()
    spec = golden.model_arrays("fixture5asym")
    model, om = gpu_model(spec), oracle_model(po, spec)
    cases = [
        (mab.synth.make_trna_batch(300, seed=11), mab.AlignmentParameters(), None),
        (mab.synth.make_long_batch(3, read_len=2500, genome_len=40_000, seed=12),
         mab.AlignmentParameters(diagonal_expansion=50, split_matrix_bigger_than=300),
         po.make_params(expansion=50, split=300 ** 2)),
    ]
    try:
        for b, P, oP in cases:
            base = None
            for g in (1, 2, 4, 8, 0):
                mab.capi.check(L.mab_debug_set_warp_group(g))
                res = engine.get_aligned_pairs_using_anchors(model, b, P)
                T, E, lik = engine.get_expectations_using_anchors(model, b, P)
                got = (res.offsets.copy(), res.x.copy(), res.y.copy(), res.p.copy(), np.concatenate([T, E, [lik]]))
                if base is None:
                    base = got
                    if oP is not None:
                        for i in range(b.n):
                            anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
                            assert_pairs_match(res.pairs(i), po.aligned_pairs(om, b.x(i), b.y(i), anc, oP), 0.01, i)
                else:
                    for k in range(4):
                        assert np.array_equal(base[k], got[k]), (g, k)
                    assert_exp_match(got[4], base[4], ("group", g), rtol=1e-11)
    finally:
        L.mab_debug_set_warp_group(0)
    assert L.mab_debug_set_warp_group(3) != 0


def test_full_size_properties(engine, golden):
    """Size-independent properties on a batch too large for the oracle (50k tRNA reads, ~140 M cells):
    (1) every pair is in range and unique per alignment, 0 < p_int <= 1e7  (the reference's
        test_getAlignedPairsWithBanding assertions, pairwiseAlignerTest.c:369-404);
    (2) permuting the batch permutes the result (alignments are independent work items);
    (3) posterior mass per read position never exceeds 1 by more than the logAdd approximation error;
    (4) expectations are additive over a partition of the batch."""
    n = 50_000
    b = mab.synth.make_trna_batch(n, seed=2024)
    model = gpu_model(golden.model_arrays("fixture5asym"))
    res = engine.get_aligned_pairs_using_anchors(model, b)
    assert engine.last_stats["cells"] > 100_000_000
    lens_x = (b.off_x[1:] - b.off_x[:-1])
    lens_y = (b.off_y[1:] - b.off_y[:-1])
    owner = np.repeat(np.arange(n), np.diff(res.offsets))
    assert (res.p > 0).all() and (res.p <= 10_000_000).all()
    assert (res.x >= 0).all() and (res.x < lens_x[owner]).all()
    assert (res.y >= 0).all() and (res.y < lens_y[owner]).all()
    key = (owner.astype(np.int64) << 20) | (res.x.astype(np.int64) << 10) | res.y.astype(np.int64)
    assert len(np.unique(key)) == len(key)
    mass = np.bincount((owner.astype(np.int64) << 8) | res.y.astype(np.int64), weights=res.p / 1e7)
    assert mass.max() < 1.01
    perm = np.random.default_rng(0).permutation(n)[:5000]
    sub = engine.get_aligned_pairs_using_anchors(model, b.subset(perm))
    for k in (0, 17, 4999):
        assert np.array_equal(sub.pairs(k), res.pairs(perm[k]))
    half = n // 2
    ea = engine.get_expectations_using_anchors(model, b.subset(np.arange(0, half)), pseudo_count=0.0)
    eb = engine.get_expectations_using_anchors(model, b.subset(np.arange(half, n)), pseudo_count=0.0)
    ec = engine.get_expectations_using_anchors(model, b, pseudo_count=0.0)
    tot = np.concatenate([ea[0] + eb[0], ea[1] + eb[1], [ea[2] + eb[2]]])
    assert_exp_match(np.concatenate([ec[0], ec[1], [ec[2]]]), tot, "additive", rtol=1e-9)
    # transitions into match consume one (x,y) pair each: their expected count is bounded by min(lX, lY) summed
    assert ec[0].reshape(5, 5)[:, 0].sum() <= np.minimum(lens_x, lens_y).sum()


def _blob_arrays(raw):
    """Splits a plan blob (host bytes) into its arrays using the header offsets (csrc/mab_internal.h BlobHeader)."""
    h = raw[:256].view(np.int64)
    names = ["magic", "n_alignments", "n_regions", "n_diagonals", "n_cells", "n_tracebacks", "n_totals", "n_anchors",
             "max_width", "len_sym_x", "len_sym_y", "off_regions", "off_diags", "off_tracebacks", "off_totals",
             "off_sym_x", "off_sym_y", "off_order_fwd", "off_order_bwd", "off_aln_region", "blob_bytes", "n_edge_cells",
             "n_full_cells", "off_diag_region", "n_class_regions0", "n_class_regions1", "n_class_regions2",
             "n_class_tbs0", "n_class_tbs1", "n_class_tbs2", "off_pack_x", "off_pack_y"]
    H = {k: int(h[i]) for i, k in enumerate(names)}
    words = lambda n: (n + 512 + 15) // 16 + (n + 512 + 31) // 32  # mab_pack_words (csrc/mab_internal.h)
    cut = lambda off, nbytes: raw[off:off + nbytes]
    return H, {
        "regions": cut(H["off_regions"], H["n_regions"] * 64),
        "diags": cut(H["off_diags"], (H["n_diagonals"] + 1) * 16),
        "tracebacks": cut(H["off_tracebacks"], H["n_tracebacks"] * 32),
        "totals": cut(H["off_totals"], H["n_totals"] * 16),
        "sym_x": cut(H["off_sym_x"] - 16, H["len_sym_x"] + 32),
        "sym_y": cut(H["off_sym_y"] - 16, H["len_sym_y"] + 32),
        "order_fwd": cut(H["off_order_fwd"], H["n_regions"] * 4),
        "order_bwd": cut(H["off_order_bwd"], H["n_tracebacks"] * 4),
        "aln_region": cut(H["off_aln_region"], (H["n_alignments"] + 1) * 8),
        "diag_region": cut(H["off_diag_region"], (H["n_diagonals"] + 1) * 4),
        "pack_x": cut(H["off_pack_x"], words(H["len_sym_x"]) * 4),
        "pack_y": cut(H["off_pack_y"], words(H["len_sym_y"]) * 4),
    }


def _assert_device_plan_equals_host_plan(batch, P, device):
    import torch
    hp = mab.Plan.from_batch(batch, P)
    hb = np.zeros(hp.info.blob_bytes, dtype=np.uint8)
    hp.write_blob(hb.ctypes.data, hb.nbytes)
    dp = mab.DevicePlan.build_on_device(batch, P, device)
    torch.cuda.synchronize()
    off = dp.blob_ptr - dp.d_blob.data_ptr()
    db = dp.d_blob[off:off + dp.plan.info.blob_bytes].cpu().numpy()
    Hh, Ah = _blob_arrays(hb)
    Hd, Ad = _blob_arrays(db)
    for k in ("n_alignments", "n_regions", "n_diagonals", "n_cells", "n_tracebacks", "n_totals", "max_width",
              "len_sym_x", "len_sym_y", "n_edge_cells", "n_full_cells", "n_class_regions0", "n_class_regions1",
              "n_class_regions2", "n_class_tbs0", "n_class_tbs1", "n_class_tbs2"):
        assert Hh[k] == Hd[k], (k, Hh[k], Hd[k])
    for k in Ah:
        assert np.array_equal(Ah[k], Ad[k]), k
    for f in ("n_cells", "n_regions", "workspace_bytes_posteriors", "workspace_bytes_expectations"):
        assert getattr(hp.info, f) == getattr(dp.plan.info, f), f


def test_device_planner_matches_host_planner(engine, golden):
    """mab_plan_build_device must produce the host planner's blob, array by array, byte for byte: tRNA batch, long reads
    with several checkpoints, small split limit (many regions per alignment), expansion 0, N / lower-case bases, and
    every cigar-driven golden case."""
    dev = engine.device
    _assert_device_plan_equals_host_plan(mab.synth.make_trna_batch(5000, seed=77), mab.AlignmentParameters(), dev)
    long_b = mab.synth.make_long_batch(6, read_len=4000, genome_len=60_000, seed=12)
    for P in (mab.AlignmentParameters(split_matrix_bigger_than=300),
              mab.AlignmentParameters(split_matrix_bigger_than=3),
              mab.AlignmentParameters(diagonal_expansion=0, split_matrix_bigger_than=1),
              mab.AlignmentParameters(diagonal_expansion=60, split_matrix_bigger_than=100, ragged_left=False,
                                      ragged_right=False)):
        _assert_device_plan_equals_host_plan(long_b, P, dev)
    cases = [c for c in golden.cases if "anchors" not in c]
    b = mab.AlignmentBatch.from_lists([c["x"] for c in cases], [c["y"] for c in cases],
                                      [(c["ops"], c["lens"]) for c in cases])
    for P in (mab.AlignmentParameters(), mab.AlignmentParameters(split_matrix_bigger_than=10)):
        _assert_device_plan_equals_host_plan(b, P, dev)
    _assert_device_plan_equals_host_plan(mab.AlignmentBatch.from_lists([], [], []), mab.AlignmentParameters(), dev)
    with pytest.raises(RuntimeError, match="does not span"):
        mab.DevicePlan.build_on_device(mab.AlignmentBatch.from_lists(["ACGT"], ["ACGT"], [([0], [3])]),
                                       mab.AlignmentParameters(), dev)


def test_host_and_device_planned_runs_agree(golden):
    b = mab.synth.make_trna_batch(400, seed=31)
    model = gpu_model(golden.model_arrays("fixture5asym"))
    r1 = mab.PairHmmEngine(plan_on="host").get_aligned_pairs_using_anchors(model, b)
    r2 = mab.PairHmmEngine(plan_on="device", max_cells=300_000).get_aligned_pairs_using_anchors(model, b)
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(r1, a), getattr(r2, a)), a


# ---------------------------------------------------------------------------------------------- round-2 additions

def _oracle_check(engine, model, om, b, P, oP, ctx, expectations=True):
    res = engine.get_aligned_pairs_using_anchors(model, b, P)
    S = om.S
    oT, oE, olik = np.full(S * S, 1e-12), np.full(S * 16, 1e-12), 0.0
    for i in range(b.n):
        anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
        assert_pairs_match(res.pairs(i), po.aligned_pairs(om, b.x(i), b.y(i), anc, oP), P.threshold, (ctx, i))
        if expectations:
            oT, oE, olik = po.expectations(om, b.x(i), b.y(i), anc, oP, T=oT, E=oE, lik=olik)
    if expectations:
        T, E, lik = engine.get_expectations_using_anchors(model, b, P)
        assert_exp_match(np.concatenate([T, E, [lik]]), np.concatenate([oT, oE, [olik]]), ctx)
    return res


@pytest.mark.parametrize("expansion", [200, 500])
def test_band_sweep_wide_expansions(engine, golden, expansion):
    """BASELINE config 4's band sweep beyond expansion 100: two 3 kb reads at expansion 200 and 500, posteriors and
    expectations against the oracle, with the launch shape the library picks by itself (band_construct
    pairwiseAligner.c:128-179, driver :638-759)."""
    b = mab.synth.make_long_batch(2, read_len=3000, genome_len=50_000, seed=80 + expansion)
    spec = golden.model_arrays("fixture5asym")
    model, om = gpu_model(spec), oracle_model(po, spec)
    P = mab.AlignmentParameters(diagonal_expansion=expansion, split_matrix_bigger_than=300)
    _oracle_check(engine, model, om, b, P, po.make_params(expansion=expansion, split=300 ** 2), ("exp", expansion))


def _with_gap_box(b, gap):
    """Appends one tRNA-like alignment whose guide cigar holds a `gap`-base deletion followed by a `gap`-base insertion:
    an unanchored gap x gap box inside an otherwise narrow batch (one wide diagonal in a narrow batch)."""
    rng = np.random.default_rng(gap)
    acgt = np.frombuffer(b"ACGT", dtype=np.uint8)
    left = acgt[rng.integers(0, 4, 40)].tobytes().decode()
    right = acgt[rng.integers(0, 4, 40)].tobytes().decode()
    gx = acgt[rng.integers(0, 4, gap)].tobytes().decode()
    gy = acgt[rng.integers(0, 4, gap)].tobytes().decode()
    xs = [bytes(b.x(i)).decode() for i in range(b.n)] + [left + gx + right]
    ys = [bytes(b.y(i)).decode() for i in range(b.n)] + [left + gy + right]
    cig = [b.cigar(i) for i in range(b.n)] + [([0, 1, 2, 0], [40, gap, gap, 40])]
    return mab.AlignmentBatch.from_lists(xs, ys, [(list(o), list(l)) for o, l in cig])


@pytest.mark.parametrize("key", ["fixture5asym", "three_asym"])
def test_wide_box_inside_a_narrow_batch(engine, golden, key):
    """Band classes mixed in one batch: 60 tRNA reads (half-warp items), plus alignments with a 12 x 12 (a warp per
    item) and a 300 x 300 unanchored box (shared-memory-ring kernels), for both state counts."""
    b = _with_gap_box(_with_gap_box(mab.synth.make_trna_batch(60, seed=41), 12), 300)
    spec = golden.model_arrays(key)
    model, om = gpu_model(spec), oracle_model(po, spec)
    P = mab.AlignmentParameters()
    plan = mab.Plan.from_batch(b, P)
    raw = np.zeros(plan.info.blob_bytes, dtype=np.uint8)
    plan.write_blob(raw.ctypes.data, raw.nbytes)
    classes = raw[:256].view(np.int64)[24:27]
    assert classes.sum() == plan.info.n_regions and (classes > 0).all(), classes
    _oracle_check(engine, model, om, b, P, po.make_params(expansion=10, split=3000 ** 2), key)


def test_band_class_dispatch_does_not_change_results(engine, golden):
    """Register-resident kernels (half-warp / warp per item) and shared-memory-ring kernels are bit-identical: the same
    batch forced through each family (mab_debug_set_min_band_class), 5- and 3-state, posteriors and DP matrices."""
    L = mab.capi.load()
    b = _with_gap_box(mab.synth.make_trna_batch(500, seed=43), 25)
    try:
        for key in ("fixture5asym", "three_asym"):
            model = gpu_model(golden.model_arrays(key))
            base = None
            for cls in (0, 1, 2):
                mab.capi.check(L.mab_debug_set_min_band_class(cls))
                res = engine.get_aligned_pairs_using_anchors(model, b)
                T, E, lik = engine.get_expectations_using_anchors(model, b)
                got = (res.offsets.copy(), res.x.copy(), res.y.copy(), res.p.copy(), np.concatenate([T, E, [lik]]))
                if base is None:
                    base = got
                else:
                    for k in range(4):
                        assert np.array_equal(base[k], got[k]), (key, cls, k)
                    assert_exp_match(got[4], base[4], (key, cls), rtol=1e-11)
    finally:
        L.mab_debug_set_min_band_class(0)
    assert L.mab_debug_set_min_band_class(3) != 0


def test_bigmutations_all_reads_exact_against_oracle(engine):
    """All 30 reads of inputBigMutations (marginCaller parameters: expansion 10, split 100): the pair set, its order and
    every p_int (within one unit of 1e-7) against the oracle."""
    with open(os.path.join(GOLDEN, "bigmutations_alignments.json")) as fh:
        alns = json.load(fh)
    hmm = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    model = mab.PairHmmModel.from_hmm(hmm)
    om = po.model_from_hmm(1, hmm.transitions, hmm.emissions)
    b = mab.AlignmentBatch.from_lists([a["x"] for a in alns], [a["y"] for a in alns],
                                      [(a["ops"], a["lens"]) for a in alns])
    P = mab.AlignmentParameters(split_matrix_bigger_than=100)
    _oracle_check(engine, model, om, b, P, po.make_params(expansion=10, split=100 ** 2), "bigmut", expectations=False)


def test_fused_backward_posterior_equals_the_separate_kernels(engine, golden):
    """N2: posterior mode runs backward sweep, total probabilities and posteriors in one kernel for band classes 0 / 1
    (pairwiseAligner.c:695-733).  Same pairs, bit for bit, as the separate backward / k_totals / k_posterior_diag
    kernels, on a batch that mixes all three classes, several tracebacks per region (long reads) and both state counts."""
    L = mab.capi.load()
    cases = [(_with_gap_box(_with_gap_box(mab.synth.make_trna_batch(400, seed=51), 12), 300), mab.AlignmentParameters()),
             (mab.synth.make_long_batch(4, read_len=3000, genome_len=50_000, seed=52),
              mab.AlignmentParameters(split_matrix_bigger_than=300))]
    try:
        for key in ("fixture5asym", "three_asym"):
            model = gpu_model(golden.model_arrays(key))
            for b, P in cases:
                got = []
                for fused in (1, 0):
                    mab.capi.check(L.mab_debug_set_fused_posterior(fused))
                    res = engine.get_aligned_pairs_using_anchors(model, b, P)
                    got.append((res.offsets.copy(), res.x.copy(), res.y.copy(), res.p.copy()))
                for k in range(4):
                    assert np.array_equal(got[0][k], got[1][k]), (key, k)
                assert len(got[0][1]) > 0
    finally:
        L.mab_debug_set_fused_posterior(0)


def test_compact_pair_coordinates_are_the_same_pairs(golden):
    """16-bit coordinates (mab_posteriors_emit_i16, the default when every sub-sequence is shorter than 32768 bases)
    carry the same pairs as the 32-bit ones; a batch with a longer sequence falls back to 32 bits by itself."""
    model = gpu_model(golden.model_arrays("fixture5asym"))
    wide, compact = mab.PairHmmEngine(compact_pairs=False), mab.PairHmmEngine()
    b = mab.synth.make_trna_batch(500, seed=61)
    r32, r16 = wide.get_aligned_pairs_using_anchors(model, b), compact.get_aligned_pairs_using_anchors(model, b)
    assert r32.x.dtype == np.int32 and r16.x.dtype == np.int16 and r16.p.dtype == np.int32
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(r32, a), getattr(r16, a)), a
    assert np.array_equal(r16.pairs(7), r32.pairs(7))
    long_b = mab.synth.make_long_batch(1, read_len=33_000, genome_len=100_000, seed=62)
    P = mab.AlignmentParameters(split_matrix_bigger_than=300)
    l16 = compact.get_aligned_pairs_using_anchors(model, long_b, P)
    assert l16.x.dtype == np.int32 and int(l16.x.max()) >= 32768
    l32 = wide.get_aligned_pairs_using_anchors(model, long_b, P)
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(l32, a), getattr(l16, a)), a


def test_over_wide_gap_box_runs_on_the_global_ring(engine, golden):
    """A 2200 x 2200 unanchored gap box (marginAlign's realign split limit allows up to 3000 x 3000): its diagonals
    are wider than any shared-memory ring (~1850 cells for five states), which failed the whole batch in round 1.  The
    class-2 items of such a batch run with the ring in global memory; pairs and expectations match the oracle, and the
    narrow alignments in the same batch are untouched."""
    b = _with_gap_box(mab.synth.make_trna_batch(20, seed=71), 2200)
    spec = golden.model_arrays("fixture5asym")
    model, om = gpu_model(spec), oracle_model(po, spec)
    P = mab.AlignmentParameters()
    assert mab.Plan.from_batch(b, P).info.max_width > 1900
    _oracle_check(engine, model, om, b, P, po.make_params(expansion=10, split=3000 ** 2), "over-wide")
