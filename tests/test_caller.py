"""marginCaller columns (BASELINE config 5): tests/inputBigMutations.sam vs tests/referencesMutated.fa.

CPU: the oracle chain (oracle posteriors -> oracle/caller_oracle.py, the restated marginCallerLib.py) against the
golden calls made from the UNMODIFIED reference C posteriors (tests/golden/caller_case.json).
GPU: ColumnCaller (mab_columns_accumulate / mab_columns_call) against the oracle chain: identical call set, expected
base counts within the 1e-6 steps a last-place difference in p_int can cause, column posteriors within 1e-9 of the
oracle's calcBasePosteriorProbs evaluated on the SAME counts and within 1e-5 (the spec) of the all-oracle chain.
The oracle's column functions are pinned to the reference's own Python in tests/test_py_reference.py."""
import json
import os

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN
from oracle import caller_oracle as co
from oracle import pyoracle as po


@pytest.fixture(scope="module")
def case():
    with open(os.path.join(GOLDEN, "caller_case.json")) as fh:
        c = json.load(fh)
    with open(os.path.join(GOLDEN, "bigmutations_alignments.json")) as fh:
        c["alns"] = json.load(fh)
    c["hmm"] = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    return c


@pytest.fixture(scope="module")
def oracle_run(case):
    om = po.model_from_hmm(case["hmm"]._modelTypeInt(), case["hmm"].transitions, case["hmm"].emissions)
    oP = po.make_params(expansion=10, split=100 ** 2)
    pairs = []
    for a in case["alns"]:
        anc = po.anchors_from_cigar(a["ops"], a["lens"], a["x"], a["y"])
        pairs.append(po.aligned_pairs(om, a["x"], a["y"], anc, oP).tolist())
    exp = co.collate(pairs, [a["y"] for a in case["alns"]], [w[0] for w in case["where"]],
                     [w[1] for w in case["where"]])
    calls, post = co.variant_calls(exp, case["references"], co.substitution_matrix(case["hmm"].emissions),
                                   case["threshold"])
    return exp, calls, post


def test_oracle_calls_match_reference_posteriors(case, oracle_run):
    exp, calls, _post = oracle_run
    assert len(exp) == case["n_columns_observed"]
    assert abs(sum(sum(v.values()) for v in exp.values()) - case["expectation_checksum"]) < 1e-6
    want = {(c[0], c[1], c[2]): c[3] for c in case["calls"]}
    got = {(c[0], c[1], c[2]): c[3] for c in calls}
    assert set(got) == set(want)
    for k in want:
        assert abs(got[k] - want[k]) <= 1e-9 * max(1.0, want[k]), k
    # the reference's own acceptance check (tests/tests.py:62-78): precision / recall against tests/mutations.txt
    truth = {(t[0], t[1], t[2]) for t in case["truth"]}
    hit = len(set(got) & truth)
    assert hit / len(got) > 0.85 and hit / len(truth) > 0.65


def test_substitution_matrix_and_vcf_round_trip(case, tmp_path):
    m = mab.caller.load_hmm_substitution_matrix(case["hmm"])
    d = co.substitution_matrix(case["hmm"].emissions)
    for i, a in enumerate("ACGT"):
        for j, b in enumerate("ACGT"):
            assert m[i, j] == d[(a, b)]
    calls = [(c[0], c[1], c[2], c[3]) for c in case["calls"]]
    path = str(tmp_path / "out.vcf")
    mab.caller.write_vcf(path, "referencesMutated.fa", case["references"], calls)
    assert mab.caller.read_vcf_calls(path) == {(c[0], c[1] + 1, c[2]) for c in calls}
    lines = [l for l in open(path) if not l.startswith("#")]
    assert len(lines) == len({(c[0], c[1]) for c in calls})
    f = lines[0].rstrip("\n").split("\t")
    assert len(f) == 8 and f[2] == "." and f[5] == "." and f[6] == "PASS"


def test_columns_have_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    L = mab.capi.load()
    z = np.zeros(8, dtype=np.int64)
    rc = L.mab_columns_accumulate(1, z.ctypes.data, z.ctypes.data, z.ctypes.data, z.ctypes.data, z.ctypes.data,
                                  z.ctypes.data, z.ctypes.data, z.ctypes.data, 1, None)
    assert rc == -5 and b"no CPU fallback" in L.mab_last_error()


@pytest.mark.gpu
def test_gpu_columns_match_the_oracle(case, oracle_run):
    exp, calls, post = oracle_run
    eng = mab.PairHmmEngine()
    model = mab.PairHmmModel.from_hmm(case["hmm"])
    alns = case["alns"]
    b = mab.AlignmentBatch.from_lists([a["x"] for a in alns], [a["y"] for a in alns],
                                      [(a["ops"], a["lens"]) for a in alns])
    cc = mab.caller.ColumnCaller(eng, case["references"])
    # two batches into the same table, like the reference's per-job pickles merged at the end (:176-186)
    half = len(alns) // 2
    for lo, hi in ((0, half), (half, len(alns))):
        cc.add(model, b.subset(np.arange(lo, hi)), [w[0] for w in case["where"][lo:hi]],
               [w[1] for w in case["where"][lo:hi]])
    table = cc.expectations()
    n_obs = 0
    for name in cc.names:
        for pos in range(len(case["references"][name])):
            row = table[cc.offset[name] + pos]
            e = exp.get((name, pos))
            if e is None:
                assert not row.any()
                continue
            n_obs += 1
            # a posterior may differ by one unit of 1e-7 from the oracle's (CUDA exp): at most one 1e-6 step per pair
            np.testing.assert_allclose(row, [e[c] for c in "ACGT"], rtol=0, atol=3e-6)
    assert n_obs == len(exp)
    err = mab.caller.load_hmm_substitution_matrix(case["hmm"])
    gp = cc.posteriors(err)
    err_d = co.substitution_matrix(case["hmm"].emissions)
    for (name, pos), probs in post.items():
        row = gp[cc.offset[name] + pos]
        # (1) the column kernel itself: the oracle's calcBasePosteriorProbs on the counts the GPU accumulated
        counts = table[cc.offset[name] + pos]
        same = co.calc_base_posterior_probs(dict(zip("ACGT", (counts / counts.sum()).tolist())),
                                            case["references"][name][pos], co.null_matrix(), err_d)
        np.testing.assert_allclose(row, [same[c] for c in "ACGT"], rtol=1e-9, atol=1e-300)
        # (2) the whole chain at the spec's bar.  A count may sit one 1e-6 step away from the oracle's (see above);
        # log P(base) is linear in the normalised counts with slopes log(err) of a few units, so one step on a column
        # of total mass >= 1 moves a posterior by ~1e-5 relative at most: columns observed less than that are skipped
        if counts.sum() >= 3.0:
            np.testing.assert_allclose(row, [probs[c] for c in "ACGT"], rtol=1e-5, atol=1e-9)
    got = {(c[0], c[1], c[2]) for c in cc.calls(err, case["threshold"])}
    want = {(c[0], c[1], c[2]) for c in calls}
    near = {(c[0], c[1], c[2]) for c in calls if abs(c[3] - case["threshold"]) < 1e-4}
    assert got - near == want - near
