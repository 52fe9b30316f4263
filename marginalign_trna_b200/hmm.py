"""`.hmm` model files: the Python-side `Hmm` container of the EM driver (submodules/cPecan/cPecanEm.py:19-105),
re-stated for Python 3 / numpy, same method names and file layout.

File layout (shared with hmm_write / hmm_loadFromFile, cPecan/impl/stateMachine.c:133-202):
  line 1: <type int> <S*S transitions, row-major from*S+to> <likelihood>
  line 2: <S*16 emissions, state*16 + x*4 + y>
  line 3 (optional): running likelihoods of the EM iterations (cPecanEm.py:49, 173-176; the C loader ignores it)
"""
import math
import random

import numpy as np

SYMBOL_NUMBER = 4
MODEL_TYPES = {"fiveState": 0, "fiveStateAsymmetric": 1, "threeState": 2, "threeStateAsymmetric": 3}
MODEL_NAMES = {v: k for k, v in MODEL_TYPES.items()}


def _fmt(v, py2_compat):
    # Python 2's str(float) printed 12 significant digits (what the reference writes); repr round-trips exactly
    return ("%.12g" % v) if py2_compat else repr(float(v))


class Hmm:
    def __init__(self, modelType):
        if isinstance(modelType, int):
            modelType = MODEL_NAMES[modelType]
        self.modelType = modelType
        self.stateNumber = 5 if modelType.startswith("five") else 3
        self.transitions = np.zeros(self.stateNumber ** 2)
        self.emissions = np.zeros(SYMBOL_NUMBER ** 2 * self.stateNumber)
        self.likelihood = 0.0
        self.runningLikelihoods = []

    def _modelTypeInt(self):
        return MODEL_TYPES[self.modelType]

    def write(self, file, py2_compat=False):
        with open(file, "w") as f:
            f.write("%s " % self._modelTypeInt() + " ".join(_fmt(v, py2_compat) for v in self.transitions)
                    + " %s\n" % _fmt(self.likelihood, py2_compat))
            f.write(" ".join(_fmt(v, py2_compat) for v in self.emissions) + "\n")
            if self.runningLikelihoods:
                f.write(" ".join(_fmt(v, py2_compat) for v in self.runningLikelihoods) + "\n")

    def addExpectationsFile(self, file):
        with open(file) as fh:
            l = [float(x) for x in fh.readline().split()]
            if len(l) != len(self.transitions) + 2 or int(l[0]) != self._modelTypeInt():
                raise RuntimeError("malformed hmm file %s" % file)
            self.likelihood += l[-1]
            self.transitions = self.transitions + np.array(l[1:-1])
            e = [float(x) for x in fh.readline().split()]
            if len(e) != len(self.emissions):
                raise RuntimeError("malformed hmm emissions in %s" % file)
            self.emissions = self.emissions + np.array(e)
            self.runningLikelihoods = [float(x) for x in fh.readline().split()]
        return self

    def addExpectations(self, transitions, emissions, likelihood):
        """In-memory equivalent of addExpectationsFile (what the GPU E-step returns)."""
        self.transitions = self.transitions + np.asarray(transitions, dtype=np.float64)
        self.emissions = self.emissions + np.asarray(emissions, dtype=np.float64)
        self.likelihood += float(likelihood)
        return self

    @staticmethod
    def loadHmm(file):
        with open(file) as fh:
            l = fh.readline().split()
        if not l:
            raise RuntimeError("empty hmm file %s" % file)
        return Hmm(MODEL_NAMES[int(l[0])]).addExpectationsFile(file)

    def normalise(self):
        """cPecanEm.py:61-73.  The row sums are Python's left-to-right `sum` (numpy's pairwise sum of 16 values rounds
        differently in the last place; tests/test_py_reference.py pins this against the reference's own code)."""
        S = self.stateNumber
        T = self.transitions.reshape(S, S)
        self.transitions = np.array([v / sum(row) for row in T.tolist() for v in row])
        E = self.emissions.reshape(S, SYMBOL_NUMBER ** 2)
        self.emissions = np.array([v / sum(row) for row in E.tolist() for v in row])

    def randomise(self, rng=random):
        self.transitions = np.array([rng.random() for _ in range(self.stateNumber ** 2)])
        self.emissions = np.array([rng.random() for _ in range(self.stateNumber * SYMBOL_NUMBER ** 2)])
        self.normalise()

    def equalise(self):
        self.transitions = np.full(self.stateNumber ** 2, 1.0 / self.stateNumber)
        self.emissions = np.full(self.stateNumber * SYMBOL_NUMBER ** 2, 1.0 / SYMBOL_NUMBER ** 2)

    def setEmissionsToJukesCantor(self, divergence):
        i = (0.25 + 0.75 * math.exp(-4.0 * divergence / 3.0)) / 4.0
        j = (0.25 - 0.25 * math.exp(-4.0 * divergence / 3.0)) / 4.0
        block = np.full((SYMBOL_NUMBER, SYMBOL_NUMBER), j)
        np.fill_diagonal(block, i)
        self.emissions = np.tile(block.reshape(-1), self.stateNumber)

    def tieEmissions(self):
        n = SYMBOL_NUMBER
        E = self.emissions.reshape(self.stateNumber, n, n).copy()
        for s in range(self.stateNumber):
            # cPecanEm.py:97-101: the identity mass is summed over all 16 entries in order, off-diagonals adding 0.0
            ident = sum(float(E[s, i, i]) for i in range(n))
            E[s] = (1.0 - ident) / (n * n - n)
            np.fill_diagonal(E[s], ident / n)
        self.emissions = E.reshape(-1)


# ---- model post-processing after EM (src/margin/marginAlignLib.py:227-263) -----------------------------------------

def setHmmIndelEmissionsToBeFlat(hmm):
    """marginAlignLib.py:251-256: every gap state emits uniformly."""
    n2 = SYMBOL_NUMBER ** 2
    hmm.emissions = np.asarray(hmm.emissions, dtype=np.float64).copy()
    hmm.emissions[n2:] = 1.0 / n2


def normaliseHmmByReferenceGCContent(hmm, gcContent):
    """marginAlignLib.py:238-249: rows (reference base) of the match and gap-X emission blocks are rescaled so that the
    reference bases have GC content `gcContent`; the insert states 2 and 4 (no reference base) are left alone."""
    n = SYMBOL_NUMBER
    E = np.asarray(hmm.emissions, dtype=np.float64).reshape(hmm.stateNumber, n, n).copy()
    for state in range(hmm.stateNumber):
        if state in (2, 4):
            continue
        for i in range(n):
            row_sum = sum(E[state, i].tolist())
            w = gcContent / 2.0 if i in (1, 2) else (1.0 - gcContent) / 2.0
            E[state, i] = [(v / row_sum) * w for v in E[state, i].tolist()]
    hmm.emissions = E.reshape(-1)


def modifyHmmEmissionsByExpectedVariationRate(hmm, substitutionRate):
    """marginAlignLib.py:258-263: match emissions multiplied by a uniform substitution matrix."""
    n = SYMBOL_NUMBER
    m = np.full((n, n), substitutionRate / (n - 1))
    np.fill_diagonal(m, 1.0 - substitutionRate)
    hmm.emissions = np.asarray(hmm.emissions, dtype=np.float64).copy()
    hmm.emissions[:n * n] = np.dot(hmm.emissions[:n * n].reshape(n, n), m).reshape(-1)


def learn_model_postprocess(unnormalised_model_path, output_model_path, py2_compat=False):
    """learnModelFromSamFileTargetFn2 (marginAlignLib.py:227-232): what turns the model EM wrote into the one
    marginAlign realigns with."""
    hmm = Hmm.loadHmm(unnormalised_model_path)
    setHmmIndelEmissionsToBeFlat(hmm)
    normaliseHmmByReferenceGCContent(hmm, 0.5)
    hmm.write(output_model_path, py2_compat=py2_compat)
    return hmm
