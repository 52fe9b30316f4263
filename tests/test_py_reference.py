"""Host-side rows pinned to the reference's own PYTHON code.

tests/golden/py_reference_vectors.json was produced by tests/golden/make_py_golden.py, which executes the reference's
source (cPecanEm.py Hmm class, marginCallerLib.py column functions, marginAlignLib.py model post-processing) read
from /root/reference with three mechanical Python-2 -> 3 shims.  These tests hold the product's Python 3 / numpy
restatements -- and the test-only caller oracle -- to those outputs."""
import json
import os
import random

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN
from marginalign_trna_b200 import hmm as hmm_mod
from oracle import caller_oracle as co


@pytest.fixture(scope="module")
def vec():
    with open(os.path.join(GOLDEN, "py_reference_vectors.json")) as fh:
        return json.load(fh)


def test_hmm_methods_match_the_reference_class(vec):
    """randomise (same Mersenne-Twister draws in the same order), normalise (the EM M-step, cPecanEm.py:184-188),
    tieEmissions, equalise + setEmissionsToJukesCantor: bit for bit."""
    for rec in vec["hmm"]:
        rng = random.Random(rec["seed"])
        h = mab.Hmm(rec["modelType"])
        h.randomise(rng)
        assert h.transitions.tolist() == rec["randomise"][0] and h.emissions.tolist() == rec["randomise"][1], rec["modelType"]
        h = mab.Hmm(rec["modelType"])
        h.transitions, h.emissions = np.array(rec["raw"][0]), np.array(rec["raw"][1])
        h.normalise()
        assert h.transitions.tolist() == rec["normalise"][0]
        assert h.emissions.tolist() == rec["normalise"][1]
        h.tieEmissions()
        np.testing.assert_allclose(h.emissions, rec["tieEmissions"], rtol=2e-16, atol=0)
        h = mab.Hmm(rec["modelType"])
        h.equalise()
        h.setEmissionsToJukesCantor(0.2)
        assert h.transitions.tolist() == rec["equalise_jc"][0] and h.emissions.tolist() == rec["equalise_jc"][1]


def test_maximisation_is_the_reference_m_step(vec):
    """calculateMaximisation (cPecanEm.py:182-202) = add the expectations, normalise, tie or keep emissions."""
    rec = next(r for r in vec["hmm"] if r["modelType"] == "fiveStateAsymmetric")
    prev = mab.Hmm(rec["modelType"])
    prev.equalise()
    for train, tie, want_e in ((True, False, rec["normalise"][1]), (True, True, rec["tieEmissions"]),
                               (False, False, prev.emissions.tolist())):
        new = mab.Hmm(rec["modelType"])
        new.addExpectations(rec["raw"][0], rec["raw"][1], -12.5)
        out = mab.em.maximisation(new, prev, mab.em.EmOptions(trainEmissions=train, tieEmissions=tie))
        assert out.transitions.tolist() == rec["normalise"][0]
        np.testing.assert_allclose(out.emissions, want_e, rtol=2e-16, atol=0)
        assert out.likelihood == -12.5


def test_hmm_file_loads_like_the_reference(vec):
    h = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    assert h.modelType == vec["load"]["modelType"] and h.likelihood == vec["load"]["likelihood"]
    assert h.transitions.tolist() == vec["load"]["transitions"] and h.emissions.tolist() == vec["load"]["emissions"]


def test_model_postprocessing_matches_the_reference(vec):
    """learnModelFromSamFileTargetFn2's two steps and the variation-rate helper (marginAlignLib.py:227-263)."""
    for rec in vec["postprocess"]:
        h = mab.Hmm(rec["modelType"])
        h.transitions, h.emissions = np.array(rec["before"][0]), np.array(rec["before"][1])
        hmm_mod.setHmmIndelEmissionsToBeFlat(h)
        assert h.emissions.tolist() == rec["flat"]
        hmm_mod.normaliseHmmByReferenceGCContent(h, 0.5)
        np.testing.assert_allclose(h.emissions, rec["flat_gc50"], rtol=2e-16, atol=0)
        h.emissions = np.array(rec["before"][1])
        hmm_mod.normaliseHmmByReferenceGCContent(h, 0.35)
        np.testing.assert_allclose(h.emissions, rec["gc35"], rtol=2e-16, atol=0)
        h.emissions = np.array(rec["before"][1])
        hmm_mod.modifyHmmEmissionsByExpectedVariationRate(h, 0.1)
        np.testing.assert_allclose(h.emissions, rec["variation_0.1"], rtol=1e-15, atol=0)


def test_learn_model_postprocess_writes_the_realign_model(vec, tmp_path):
    rec = vec["postprocess"][0]
    h = mab.Hmm(rec["modelType"])
    h.transitions, h.emissions = np.array(rec["before"][0]), np.array(rec["before"][1])
    src, dst = str(tmp_path / "unnormalised.hmm"), str(tmp_path / "out.hmm")
    h.write(src)
    hmm_mod.learn_model_postprocess(src, dst)
    back = mab.Hmm.loadHmm(dst)
    np.testing.assert_allclose(back.emissions, rec["flat_gc50"], rtol=2e-16, atol=0)
    assert back.transitions.tolist() == rec["before"][0]


def test_caller_oracle_is_pinned_to_the_reference_functions(vec):
    """oracle/caller_oracle.py (the checker of the GPU column kernels) against marginCallerLib.py's own
    loadHmmSubstitutionMatrix and calcBasePosteriorProbs: identical doubles."""
    h = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    err = co.substitution_matrix(h.emissions)
    assert [err[k] for k in sorted(err, key=lambda k: ("ACGT".index(k[0]), "ACGT".index(k[1])))] == vec["substitution_matrix"]
    m = mab.caller.load_hmm_substitution_matrix(h)
    assert m.reshape(-1).tolist() == vec["substitution_matrix"]
    for rec in vec["caller"]:
        evo = co.null_matrix() if rec["evo"] == "null" else err
        post = co.calc_base_posterior_probs(dict(zip("ACGT", rec["obs"])), rec["ref"], evo, err)
        assert [post[b] for b in "ACGT"] == rec["post"], rec


def test_caller_collation_and_snv_loop_are_pinned(vec):
    """The loop bodies of posteriorProbabilityCalculationTargetFn (marginCallerLib.py:59-69: `%f` round trip, float sums,
    N / lower-case read bases) and variantCallSamFileTargetFn (:189-209), cut out of the reference and run on posteriors
    of the reference C library: oracle/caller_oracle.py gives the same expectations and the same calls, to the bit."""
    with open(os.path.join(GOLDEN, "caller_case.json")) as fh:
        case = json.load(fh)
    c = vec["collate"]
    where = [case["where"][r["aln"]] for r in c["reads"]]
    exp = co.collate(c["pairs"], [r["y"] for r in c["reads"]], [w[0] for w in where], [w[1] for w in where])
    want = {(e[0], e[1]): e[2:] for e in c["expectations"]}
    assert set(exp) == set(want) and len(want) > 1000
    for k, v in want.items():
        assert [exp[k][b] for b in "ACGT"] == v, k
    h = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    calls, _post = co.variant_calls(exp, case["references"], co.substitution_matrix(h.emissions), c["threshold"])
    assert sorted(map(tuple, calls)) == sorted(tuple(x) for x in c["calls"]) and len(calls) > 0
