"""`.hmm` container (hmm.py) -- mirrors the reference's test_hmm_* round trips (pairwiseAlignerTest.c:675-767) and
the Python Hmm class semantics (cPecanEm.py:19-105)."""
import os
import random

import numpy as np

import marginalign_trna_b200 as mab
from conftest import GOLDEN


def test_load_fixture_model():
    h = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    assert h.modelType == "fiveStateAsymmetric" and h.stateNumber == 5
    assert len(h.transitions) == 25 and len(h.emissions) == 80
    assert abs(h.transitions[0] - 0.861018861829) < 1e-15
    assert abs(h.likelihood - (-6.14898969935e+11)) < 1


def test_write_load_round_trip_all_types(tmp_path):
    for name in mab.hmm.MODEL_TYPES:
        h = mab.Hmm(name)
        h.randomise(random.Random(3))
        h.likelihood = -1234.5
        h.runningLikelihoods = [-3.0, -2.0]
        p = str(tmp_path / (name + ".hmm"))
        h.write(p)
        g = mab.Hmm.loadHmm(p)
        assert g.modelType == name
        assert np.array_equal(g.transitions, h.transitions) and np.array_equal(g.emissions, h.emissions)
        assert g.likelihood == h.likelihood and g.runningLikelihoods == h.runningLikelihoods
        h.write(p, py2_compat=True)  # 12 significant digits, what the Python-2 reference writes
        g = mab.Hmm.loadHmm(p)
        assert np.allclose(g.transitions, h.transitions, rtol=1e-11, atol=0)


def test_normalise_and_expectation_sums():
    h = mab.Hmm("threeState")
    h.addExpectations(np.arange(1, 10, dtype=float), np.ones(48), -5.0)
    h.addExpectations(np.arange(1, 10, dtype=float), np.ones(48), -7.0)
    assert h.likelihood == -12.0 and h.transitions[8] == 18.0
    h.normalise()
    assert np.allclose(h.transitions.reshape(3, 3).sum(axis=1), 1.0)
    assert np.allclose(h.emissions.reshape(3, 16).sum(axis=1), 1.0)


def test_equalise_jukes_cantor_tie():
    h = mab.Hmm("fiveState")
    h.equalise()
    assert np.allclose(h.transitions, 0.2) and np.allclose(h.emissions, 1 / 16)
    h.setEmissionsToJukesCantor(0.2)
    E = h.emissions.reshape(5, 4, 4)
    assert np.allclose(E.sum(axis=(1, 2)), 1.0) and E[0, 0, 0] > E[0, 0, 1]
    h.randomise(random.Random(1))
    ident = np.trace(h.emissions.reshape(5, 4, 4)[2])
    h.tieEmissions()
    E = h.emissions.reshape(5, 4, 4)
    assert np.allclose(np.trace(E[2]), ident) and np.allclose(E[2].sum(), 1.0)
    assert np.allclose(E[2][0, 1], E[2][3, 2])
