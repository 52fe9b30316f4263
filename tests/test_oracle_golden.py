"""Pins the CPU oracle (oracle/pairhmm_oracle.c) to the reference: bit-for-bit against fixtures generated from the
reference's own C code (tests/golden/, make_golden.py) and against the known answers in the reference's unit tests
(submodules/cPecan/tests/pairwiseAlignerTest.c)."""
import hashlib
import json
import os

import numpy as np
import pytest

from conftest import GOLDEN, oracle_anchors, oracle_model, oracle_params
from oracle import pyoracle as po


def test_golden_cases_bit_exact(golden):
    assert len(golden.cases) >= 40
    for c in golden.cases:
        m = oracle_model(po, golden.model_arrays(c["model"]))
        anc = oracle_anchors(po, c)
        P = oracle_params(po, c["params"])
        rl, rr = c["params"]["ragged_left"], c["params"]["ragged_right"]
        pairs = po.aligned_pairs(m, c["x"], c["y"], anc, P, rl, rr)
        assert np.array_equal(pairs, golden.pairs(c["name"])), c["name"]
        T, E, lik = po.expectations(m, c["x"], c["y"], anc, P, rl, rr)
        got = np.concatenate([T, E, [lik]])
        want = golden.exp(c["name"])
        assert got.shape == want.shape, c["name"]
        # identical operation order and libm: the doubles match exactly
        assert np.array_equal(got, want), (c["name"], np.abs(got - want).max())


def test_bigmutations_digests():
    with open(os.path.join(GOLDEN, "bigmutations_digest.json")) as fh:
        dig = json.load(fh)
    with open(os.path.join(GOLDEN, "bigmutations_alignments.json")) as fh:
        alns = json.load(fh)
    h = [l.split() for l in open(os.path.join(GOLDEN, "input_model.hmm"))]
    T = np.array([float(v) for v in h[0][1:26]])
    E = np.array([float(v) for v in h[1]])
    m = po.model_from_hmm(int(h[0][0]), T, E)
    assert len(dig) == len(alns) == 30
    for d, a in zip(dig[:12], alns[:12]):  # 12 reads keep the CPU suite short; the GPU suite covers all 30
        anc = po.anchors_from_cigar(a["ops"], a["lens"], a["x"], a["y"])
        pairs = po.aligned_pairs(m, a["x"], a["y"], anc, po.make_params(expansion=10, split=100 ** 2))
        assert len(pairs) == d["n_pairs"]
        assert hashlib.sha256(np.ascontiguousarray(pairs).tobytes()).hexdigest() == d["sha256"]
        Tx, Ex, lik = po.expectations(m, a["x"], a["y"], anc, po.make_params(expansion=10, split=300 ** 2))
        assert np.array_equal(np.concatenate([Tx, Ex, [lik]]), np.array(d["exp"]))


def test_logadd_samples():
    s = np.load(os.path.join(GOLDEN, "logadd_samples.npy"))
    for a, b, want in s:
        got = po.logadd(a, b)
        assert got == want or (np.isnan(got) and np.isnan(want)), (a, b, got, want)
    # property of the reference's test_logAdd (pairwiseAlignerTest.c:133-143): within 1e-3 of the exact value
    rng = np.random.default_rng(1)
    for _ in range(2000):
        a, b = rng.random(), rng.random()
        assert abs(po.logadd(np.log(a), np.log(b)) - np.log(a + b)) < 1e-3


def test_symbol_kat():
    # test_symbol, pairwiseAlignerTest.c:145-152
    assert [po.symbol(c) for c in "AcGtNnXa"] == [0, 1, 2, 3, 4, 4, 4, 0]


def test_bands_kat():
    # test_bands, pairwiseAlignerTest.c:68-131: anchors (1,0),(2,1),(3,3), lX=6, lY=5, expansion 2
    L, R = po.band([[1, 0], [2, 1], [3, 3]], 6, 5, 2)
    want = [(0, 0), (-1, 1), (-2, 2), (-1, 3), (-2, 4), (-1, 3), (-2, 4), (-3, 3), (-2, 2), (-1, 3), (0, 2), (1, 1)]
    assert list(zip(L.tolist(), R.tolist())) == want


def test_split_points_kat():
    # test_getSplitPoints, pairwiseAlignerTest.c:536-605
    ms = 2000 * 2000
    assert po.split_points([], 3000, 1000, ms, 0, 0).tolist() == [[0, 0, 3000, 1000]]
    lX, lY = 20000, 25000
    assert po.split_points([], lX, lY, ms, 1, 1).tolist() == []
    assert po.split_points([], lX, lY, ms, 1, 0).tolist() == [[18000, 23000, lX, lY]]
    assert po.split_points([], lX, lY, ms, 0, 1).tolist() == [[0, 0, 2000, 2000]]
    assert po.split_points([], lX, lY, ms, 0, 0).tolist() == [[0, 0, 2000, 2000], [18000, 23000, lX, lY]]
    anchors = [[2000, 2000], [4002, 4001], [5000, 5000], [8000, 6000], [9000, 9000], [10000, 14000], [15000, 15000],
               [16000, 16000]]
    assert po.split_points(anchors, lX, lY, ms, 0, 0).tolist() == [
        [0, 0, 3001, 3001], [3002, 3001, 9500, 11001], [9501, 12000, 12001, 14500], [13000, 14501, 18000, 18001],
        [18001, 23000, 20000, 25000]]


def test_diagonal_dp_kat(golden):
    # test_diagonalDPCalculations, pairwiseAlignerTest.c:241-323: AGCG vs AGTTCG, built-in 5-state model,
    # threshold 0.2 -> exactly the pairs (0,0) (1,1) (2,4) (3,5)
    m = po.model_default(0)
    P = po.make_params(threshold=0.2, expansion=2)
    pairs = po.aligned_pairs(m, "AGCG", "AGTTCG", [], P, False, False)
    assert sorted(map(tuple, pairs[:, 1:].tolist())) == [(0, 0), (1, 1), (2, 4), (3, 5)]
    assert np.all(pairs[:, 0] > 0) and np.all(pairs[:, 0] <= 10000000)


def test_ragged_ends_property():
    # test_getAlignedPairsWithRaggedEnds, pairwiseAlignerTest.c:634-669 (fewer repetitions): a 100-nt core shared by
    # two 300-nt sequences is recovered on its diagonal when both ends are ragged
    rng = np.random.default_rng(5)
    m = po.model_default(0)
    for _ in range(20):
        core = "".join(rng.choice(list("ACGT"), 100))
        x = "".join(rng.choice(list("ACGT"), 100)) + core + "".join(rng.choice(list("ACGT"), 100))
        y = "".join(rng.choice(list("ACGT"), 100)) + core + "".join(rng.choice(list("ACGT"), 100))
        anc = [[100 + i, 100 + i] for i in range(100)]
        pairs = po.aligned_pairs(m, x, y, anc, po.make_params(expansion=10), True, True)
        strong = pairs[pairs[:, 0] >= 5000000]
        on_diag = strong[(strong[:, 1] == strong[:, 2]) & (strong[:, 1] >= 100) & (strong[:, 1] < 200)]
        assert len(on_diag) >= 95


def test_model_parameter_block(golden):
    # the 3-/5-state transition lists are in cellCalculate order (stateMachine.c:399-429, 636-661)
    t5 = po.model_transitions(po.model_default(0))
    assert [(g, f, t) for g, f, t, _ in t5] == [(0, 0, 1), (0, 1, 1), (0, 0, 3), (0, 3, 3), (1, 0, 0), (1, 1, 0),
                                                (1, 2, 0), (1, 3, 0), (1, 4, 0), (2, 0, 2), (2, 2, 2), (2, 0, 4),
                                                (2, 4, 4)]
    t3 = po.model_transitions(po.model_default(2))
    assert [(g, f, t) for g, f, t, _ in t3] == [(0, 0, 1), (0, 1, 1), (0, 2, 1), (1, 0, 0), (1, 1, 0), (1, 2, 0),
                                                (2, 0, 2), (2, 2, 2), (2, 1, 2)]


def test_cell_count_matches_band():
    rng = np.random.default_rng(3)
    x = "".join(rng.choice(list("ACGT"), 90))
    anc = [[i, i] for i in range(0, 90, 3)]
    P = po.make_params(expansion=10)
    L, R = po.band(anc, 90, 90, 10)
    assert po.count_cells(90, 90, anc, P) == int(((R - L) // 2 + 1).sum())
