"""BASELINE config 2 on the GPU: EM training of the 5-state asymmetric model, 10 iterations, on the reference's own
fixture alignments (tests/inputBigMutations.sam reads vs tests/referencesMutated.fa, stored in
tests/golden/bigmutations_alignments.json), marginAlign's EM parameters (expansion 10, split 300, trainEmissions;
src/margin/marginAlign.py:41-52, cPecanEm.py:107-202).  The same loop is run with the CPU oracle computing the
expectations; the per-iteration likelihoods and the final normalised model must agree within 1e-5 relative (spec),
what we assert is 1e-7."""
import json
import os

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu

ITERATIONS = 10
RTOL = 1e-7  # spec: 1e-5


def reference_m_step(modelType, T, E, lik):
    """calculateMaximisation (cPecanEm.py:182-202, trainEmissions, no tying) = Hmm.normalise (:61-73), written out with
    plain lists and left-to-right sums like the reference -- independent of mab.Hmm / mab.em.maximisation, which
    tests/test_py_reference.py pins to the reference's own code."""
    S = 5 if modelType.startswith("five") else 3
    T, E = list(T), list(E)
    for s in range(S):
        j = sum(T[s * S:(s + 1) * S])
        T[s * S:(s + 1) * S] = [v / j for v in T[s * S:(s + 1) * S]]
        j = sum(E[s * 16:(s + 1) * 16])
        E[s * 16:(s + 1) * 16] = [v / j for v in E[s * 16:(s + 1) * 16]]
    hmm = mab.Hmm(modelType)
    hmm.transitions, hmm.emissions, hmm.likelihood = np.array(T), np.array(E), float(lik)
    return hmm


def oracle_em(batch, options):
    """cPecanEm.py:107-202 with the oracle as the E-step and reference_m_step as the M-step."""
    hmm = mab.em.initial_hmm(options)
    oP = po.make_params(expansion=options.diagonalExpansion, split=options.splitMatrixBiggerThanThis ** 2)
    running = []
    for _ in range(options.iterations):
        om = po.model_from_hmm(hmm._modelTypeInt(), hmm.transitions, hmm.emissions)
        S = om.S
        T, E, lik = np.full(S * S, 1e-12), np.full(S * 16, 1e-12), 0.0
        for i in range(batch.n):
            anc = po.anchors_from_cigar(*batch.cigar(i), batch.x(i), batch.y(i))
            T, E, lik = po.expectations(om, batch.x(i), batch.y(i), anc, oP, T=T, E=E, lik=lik)
        hmm = reference_m_step(hmm.modelType, T.tolist(), E.tolist(), lik)
        running.append(hmm.likelihood)
    return hmm, running


@pytest.mark.parametrize("model_type,start", [("fiveStateAsymmetric", "fixture"), ("threeState", "equal")])
def test_em_ten_iterations_match_the_oracle(tmp_path, model_type, start):
    with open(os.path.join(GOLDEN, "bigmutations_alignments.json")) as fh:
        alns = json.load(fh)[:5]
    b = mab.AlignmentBatch.from_lists([a["x"] for a in alns], [a["y"] for a in alns],
                                      [(a["ops"], a["lens"]) for a in alns])
    kw = dict(iterations=ITERATIONS, trials=1, modelType=model_type)
    if start == "fixture":
        kw["inputModel"] = os.path.join(GOLDEN, "input_model.hmm")
    else:
        kw["randomStart"] = False
    out = str(tmp_path / "trained.hmm")
    eng = mab.PairHmmEngine()
    got = mab.em.expectation_maximisation_trials(eng, b, out, mab.em.EmOptions(**kw))
    want, running = oracle_em(b, mab.em.EmOptions(**kw))
    assert len(got.runningLikelihoods) == ITERATIONS
    np.testing.assert_allclose(got.runningLikelihoods, running, rtol=RTOL)
    np.testing.assert_allclose(got.transitions, want.transitions, rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(got.emissions, want.emissions, rtol=RTOL, atol=1e-300)
    # EM gains likelihood (the reference's test_em_* check, pairwiseAlignerTest.c:769-832).  What the reference calls
    # likelihood is the sum of the (stale, approximate-logAdd) totals over all diagonals, so near convergence it may
    # wobble by ~1e-5 relative on these reads: only the overall gain is asserted.
    assert running[-1] > running[0] and got.runningLikelihoods[-1] > got.runningLikelihoods[0]
    # the written model loads back to the same parameters (the .hmm file stays drop-in)
    back = mab.Hmm.loadHmm(out)
    np.testing.assert_allclose(back.transitions, got.transitions, rtol=1e-6)


def test_config2_on_the_two_alignments_of_input_sam(tmp_path, golden):
    """BASELINE config 2 as worded: fiveStateAsymmetric, 10 iterations from the shipped model, on the two alignments of
    the reference's tests/input.sam (golden cases fixture0 / fixture1: 1751- and 6222-nt reads, ~190 k cells per
    iteration), expansion 10, split 300, trainEmissions."""
    cs = [next(c for c in golden.cases if c["name"] == n) for n in ("fixture0_realign", "fixture1_realign")]
    b = mab.AlignmentBatch.from_lists([c["x"] for c in cs], [c["y"] for c in cs], [(c["ops"], c["lens"]) for c in cs])
    kw = dict(iterations=ITERATIONS, trials=1, modelType="fiveStateAsymmetric",
              inputModel=os.path.join(GOLDEN, "input_model.hmm"))
    got = mab.em.expectation_maximisation_trials(mab.PairHmmEngine(), b, str(tmp_path / "m.hmm"), mab.em.EmOptions(**kw))
    want, running = oracle_em(b, mab.em.EmOptions(**kw))
    np.testing.assert_allclose(got.runningLikelihoods, running, rtol=RTOL)
    np.testing.assert_allclose(got.transitions, want.transitions, rtol=RTOL, atol=1e-300)
    np.testing.assert_allclose(got.emissions, want.emissions, rtol=RTOL, atol=1e-300)
    assert engine_cells(b) > 150_000


def engine_cells(b):
    plan = mab.Plan.from_batch(b, mab.AlignmentParameters(split_matrix_bigger_than=300))
    return plan.info.n_cells


def test_plan_cache_reuses_the_device_plans_between_iterations(golden):
    """The EM loop plans its batch once (pairhmm.PlanCache: the band geometry does not depend on the model,
    cPecanRealign.c:567 recomputes it every iteration).  Several chunks, two models, three calls: cached calls give what
    uncached calls give (the same kernels on the same plan; sums differ only by the order of the atomic additions),
    another batch or other banding parameters refill the cache."""
    eng = mab.PairHmmEngine(max_cells=150_000)
    b = mab.synth.make_trna_batch(400, seed=91)
    P = mab.AlignmentParameters()
    assert len(eng._chunks(b, P, True)) >= 3
    models = [mab.PairHmmModel.from_arrays(*golden.model_arrays("fixture5asym")), mab.PairHmmModel.default(0)]
    cache = mab.PlanCache()
    for it, m in enumerate([models[0], models[1], models[0]]):
        got = np.concatenate([np.ravel(v) for v in eng.get_expectations_using_anchors(m, b, P, plan_cache=cache)])
        want = np.concatenate([np.ravel(v) for v in eng.get_expectations_using_anchors(m, b, P)])
        np.testing.assert_allclose(got, want, rtol=1e-11, atol=0)
        assert cache.hits == it and len(cache.plans) >= 3
    b2 = mab.synth.make_trna_batch(400, seed=92)
    eng.get_expectations_using_anchors(models[0], b2, P, plan_cache=cache)
    assert cache.hits == 2 and cache.batch is b2
    P2 = mab.AlignmentParameters(diagonal_expansion=12)
    got = np.concatenate([np.ravel(v) for v in eng.get_expectations_using_anchors(models[0], b2, P2, plan_cache=cache)])
    want = np.concatenate([np.ravel(v) for v in eng.get_expectations_using_anchors(models[0], b2, P2)])
    np.testing.assert_allclose(got, want, rtol=1e-11, atol=0)
    assert cache.hits == 2
    tiny = mab.PlanCache(max_bytes=1)  # nothing fits: every call plans again, results unchanged
    for _ in range(2):
        got = np.concatenate([np.ravel(v) for v in eng.get_expectations_using_anchors(models[0], b, P, plan_cache=tiny)])
    assert tiny.hits == 0 and tiny.plans is None
