"""Realign tail on the GPU (mab_realign_chain / mab_realign_emit through PairHmmEngine.realign) against the reference.

Bars: cigars and chains identical to the reference CLI's (tests/golden/realign_cases.json) and to the oracle on the
same inputs.  The only tolerated difference upstream is a posterior within 1 unit of 1e-7 (CUDA exp vs glibc exp);
it can move a reweighted weight by one unit and so, in principle, flip an exact tie; the tests would show it."""
import json
import os

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    return mab.PairHmmEngine()


@pytest.fixture(scope="module")
def models():
    hmm = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    return mab.PairHmmModel.from_hmm(hmm), po.model_from_hmm(hmm._modelTypeInt(), hmm.transitions, hmm.emissions)


def oracle_realign(om, b, i, oP, gg, mg):
    anc = po.anchors_from_cigar(*b.cigar(i), b.x(i), b.y(i))
    pairs = po.aligned_pairs(om, b.x(i), b.y(i), anc, oP)
    chain, _ = po.realign_chain(pairs, len(b.x(i)), len(b.y(i)), gg, mg, seed=1)
    return po.cigar_from_pairs(chain, len(b.x(i)), len(b.y(i))), chain


def assert_same(res, i, ops, chain, ctx):
    g_ops, g_lens = res.cigar(i)
    assert g_ops.tolist() == [int(o) for o in ops[:, 0]] and g_lens.tolist() == [int(n) for n in ops[:, 1]], ctx
    if res.chain_off is not None:
        assert np.array_equal(res.chain(i), chain), ctx


def test_golden_cigars_of_the_reference_cli(engine, models):
    with open(os.path.join(GOLDEN, "realign_cases.json")) as fh:
        cases = json.load(fh)
    by_gamma = {}
    for c in cases:
        by_gamma.setdefault((c["gap_gamma"], c["match_gamma"]), []).append(c)
    for (gg, mg), cs in by_gamma.items():
        b = mab.AlignmentBatch.from_lists([c["x"] for c in cs], [c["y"] for c in cs],
                                          [(c["ops"], c["lens"]) for c in cs])
        res = engine.realign(models[0], b, gap_gamma=gg, match_gamma=mg, want_chain=True)
        assert res.n == len(cs)
        for i, c in enumerate(cs):
            ops = np.stack([c["out_ops"], c["out_lens"]], axis=1)
            assert_same(res, i, ops, np.asarray(c["chain"]).reshape(-1, 2), c["name"])


def test_batched_trna_reads_against_oracle(engine, models):
    b = mab.synth.make_trna_batch(400, seed=31)
    oP = po.make_params(expansion=10, split=3000 ** 2)
    for gg, mg in ((0.5, 0.0), (0.3, 0.4)):
        res = engine.realign(models[0], b, gap_gamma=gg, match_gamma=mg, want_chain=True)
        for i in range(b.n):
            ops, chain = oracle_realign(models[1], b, i, oP, gg, mg)
            assert_same(res, i, ops, chain, (gg, mg, i))


def test_long_reads_and_chunking(models):
    """3 kb reads (several checkpoints, thousands of pairs per alignment), one chunk per read vs a single chunk."""
    b = mab.synth.make_long_batch(3, read_len=3000, genome_len=50_000, seed=33)
    oP = po.make_params(expansion=10, split=3000 ** 2)
    one = mab.PairHmmEngine(max_cells=10 ** 9)
    many = mab.PairHmmEngine(max_cells=60_000)
    r1 = one.realign(models[0], b, want_chain=True)
    r2 = many.realign(models[0], b, want_chain=True)
    assert many.last_stats["chunks"] >= 3 and one.last_stats["chunks"] == 1
    for a in ("off_ops", "ops", "lens", "chain_off", "chain_x", "chain_y"):
        assert np.array_equal(getattr(r1, a), getattr(r2, a)), a
    for i in range(b.n):
        ops, chain = oracle_realign(models[1], b, i, oP, 0.5, 0.0)
        assert_same(r1, i, ops, chain, i)


def test_repeated_calls_reuse_the_engine_buffers(models):
    """The engine keeps its DP workspace and the bound-sized arrays of the realign path between chunks and calls
    (pairhmm.py: _workspace, _pooled): chunks of different sizes and repeated calls on one engine -- several chunks, so
    that both pipeline slots are written again while copies of an earlier chunk may still be queued -- give the same
    cigars as a fresh engine, and the chunking of a call does not depend on what earlier calls left allocated."""
    big = mab.synth.make_trna_batch(6000, seed=35)
    small = mab.synth.make_trna_batch(700, seed=36)
    eng = mab.PairHmmEngine(max_cells=2_000_000)
    want_big = mab.PairHmmEngine(max_cells=2_000_000).realign(models[0], big, want_chain=True)
    want_small = mab.PairHmmEngine(max_cells=2_000_000).realign(models[0], small, want_chain=True)
    auto = mab.PairHmmEngine()
    chunks = []
    for b, want in ((small, want_small), (big, want_big), (small, want_small), (big, want_big)):
        got = eng.realign(models[0], b, want_chain=True)
        assert eng.last_stats["chunks"] >= (3 if b is big else 1)
        for a in ("off_ops", "ops", "lens", "chain_off", "chain_x", "chain_y"):
            assert np.array_equal(getattr(got, a), getattr(want, a)), a
        auto.realign(models[0], b)
        chunks.append(auto.last_stats["chunks"])
    assert chunks[0] == chunks[2] and chunks[1] == chunks[3]
    pairs = [eng.get_aligned_pairs_using_anchors(models[0], small) for _ in range(2)]
    for a in ("offsets", "x", "y", "p"):
        assert np.array_equal(getattr(pairs[0], a), getattr(pairs[1], a)), a


def test_pair_bound_overflow_recomputes_the_chunk(models):
    """At threshold 0 nearly every band cell yields a pair, four times the bound the sync-free path sizes its pair
    arrays by (PAIRS_PER_CELL_CAP): the chunk is then computed again with exact sizes -- after the NEXT chunk's kernels
    have already run in the shared workspace.  Several chunks with the overflow == an engine whose bound is never
    exceeded == the oracle."""
    b = mab.synth.make_trna_batch(900, seed=37)
    P = mab.AlignmentParameters(threshold=0.0)
    tight = mab.PairHmmEngine(max_cells=400_000)
    roomy = mab.PairHmmEngine(max_cells=400_000)
    roomy.PAIRS_PER_CELL_CAP = 1.25
    got = tight.realign(models[0], b, P, want_chain=True)
    assert tight.last_stats["chunks"] >= 4 and tight.last_stats["pairs"] > 0.5 * tight.last_stats["cells"]
    want = roomy.realign(models[0], b, P, want_chain=True)
    for a in ("off_ops", "ops", "lens", "chain_off", "chain_x", "chain_y"):
        assert np.array_equal(getattr(got, a), getattr(want, a)), a
    oP = po.make_params(threshold=0.0, expansion=10, split=3000 ** 2)
    for i in (0, 449, 899):
        ops, chain = oracle_realign(models[1], b, i, oP, 0.5, 0.0)
        assert_same(got, i, ops, chain, i)


def test_full_size_properties(engine, models):
    """100 k reads: every cigar spans its two sequences, match runs equal the chain, chains strictly increase."""
    b = mab.synth.make_trna_batch(100_000, seed=35)
    res = engine.realign(models[0], b, want_chain=True)
    lx = (b.off_x[1:] - b.off_x[:-1]).astype(np.int64)
    ly = (b.off_y[1:] - b.off_y[:-1]).astype(np.int64)
    aln = np.repeat(np.arange(b.n), np.diff(res.off_ops))
    lens = res.lens.astype(np.int64)
    span_x = np.bincount(aln, weights=lens * (res.ops != 2), minlength=b.n).astype(np.int64)
    span_y = np.bincount(aln, weights=lens * (res.ops != 1), minlength=b.n).astype(np.int64)
    assert np.array_equal(span_x, lx) and np.array_equal(span_y, ly)
    matched = np.bincount(aln, weights=lens * (res.ops == 0), minlength=b.n).astype(np.int64)
    assert np.array_equal(matched, np.diff(res.chain_off))
    caln = np.repeat(np.arange(b.n), np.diff(res.chain_off))
    same = caln[1:] == caln[:-1]
    assert np.all(np.diff(res.chain_x.astype(np.int64))[same] > 0) and np.all(np.diff(res.chain_y.astype(np.int64))[same] > 0)
    # no two adjacent ops of the same kind
    same_op = (aln[1:] == aln[:-1]) & (res.ops[1:] == res.ops[:-1])
    assert not same_op.any()
    # spot-check 50 reads against the oracle
    oP = po.make_params(expansion=10, split=3000 ** 2)
    for i in np.random.default_rng(1).choice(b.n, 50, replace=False):
        ops, chain = oracle_realign(models[1], b, int(i), oP, 0.5, 0.0)
        assert_same(res, int(i), ops, chain, int(i))


def test_empty_and_degenerate(engine, models):
    b = mab.AlignmentBatch.from_lists(["ACGTACGTAC", "A"], ["ACGTTCGTAC", "C"], [([0], [10]), ([0], [1])])
    res = engine.realign(models[0], b, want_chain=True)
    oP = po.make_params(expansion=10, split=3000 ** 2)
    for i in range(b.n):
        ops, chain = oracle_realign(models[1], b, i, oP, 0.5, 0.0)
        assert_same(res, i, ops, chain, i)
