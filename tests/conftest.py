import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


class GoldenCases:
    """tests/golden/cases.{json,npz}: inputs + outputs of the reference C code (made by make_golden.py)."""

    def __init__(self):
        with open(os.path.join(GOLDEN, "cases.json")) as fh:
            d = json.load(fh)
        self.models = d["models"]
        self.cases = d["cases"]
        self.arrays = np.load(os.path.join(GOLDEN, "cases.npz"))

    def pairs(self, name):
        return self.arrays[name + "/pairs"]

    def exp(self, name):
        return self.arrays[name + "/exp"]

    def model_arrays(self, key):
        m = self.models[key]
        if m is None:
            return None
        return m["type"], np.array(m["T"]), np.array(m["E"])


@pytest.fixture(scope="session")
def golden():
    return GoldenCases()


def oracle_model(po, spec):
    return po.model_default(0) if spec is None else po.model_from_hmm(*spec)


def oracle_anchors(po, case):
    if "anchors" in case:
        return np.asarray(case["anchors"], dtype=np.int64).reshape(-1, 2)
    return po.anchors_from_cigar(case["ops"], case["lens"], case["x"], case["y"])


def oracle_params(po, p):
    return po.make_params(threshold=p["threshold"], min_diags=p["min_diags"],
                          traceback_diagonals=p["traceback_diagonals"], expansion=p["expansion"],
                          split=p["split"] ** 2)
