"""Generates tests/golden/caller_case.json (run in the build container only): BASELINE config 5,
tests/inputBigMutations.sam (30 reads) vs tests/referencesMutated.fa with the known mutations of tests/mutations.txt
(the inputs of the reference's own marginCaller test, tests/tests.py:62-78, 150-170).  Stores where every fixture
alignment of bigmutations_alignments.json lies on its reference, the reference sequences and the truth set; the
expected calls come from the UNMODIFIED reference C code for the posteriors (oracle/_ref, getAlignedPairsUsingAnchors
with marginCaller's parameters: expansion 10, split 100) pushed through the restated Python of marginCallerLib.py
(oracle/caller_oracle.py; the reference Python itself is Python 2 + pysam and cannot run here)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/tests"

import marginalign_trna_b200 as mab  # noqa: E402
from marginalign_trna_b200 import seqio  # noqa: E402
from oracle import caller_oracle as co  # noqa: E402
from oracle import pyoracle as po  # noqa: E402
from oracle import pyref  # noqa: E402


def main():
    _, recs = seqio.sam_read(os.path.join(REF, "inputBigMutations.sam"))
    refs = seqio.fasta_dict(os.path.join(REF, "referencesMutated.fa"))
    alns = json.load(open(os.path.join(HERE, "bigmutations_alignments.json")))
    where = []
    k = 0
    for r in recs:
        if r.is_unmapped:
            continue
        g = seqio.guide_from_sam(r)
        a = alns[k]
        assert a["name"] == r.qname and a["x"] == refs[g.ref_name][g.ref_start:g.ref_end], r.qname
        where.append([g.ref_name, g.ref_start])
        k += 1
    assert k == len(alns)
    truth = [[f[0], int(f[1]), f[2]] for f in (l.split() for l in open(os.path.join(REF, "mutations.txt"))) if f]
    model = pyref.RefModel.load(os.path.join(HERE, "input_model.hmm"))
    pairs = []
    for a in alns:
        anc = po.anchors_from_cigar(a["ops"], a["lens"], a["x"], a["y"])
        pairs.append(pyref.aligned_pairs(model, a["x"], a["y"], anc, expansion=10, split=100 * 100).tolist())
    exp = co.collate(pairs, [a["y"] for a in alns], [w[0] for w in where], [w[1] for w in where])
    hmm = mab.Hmm.loadHmm(os.path.join(HERE, "input_model.hmm"))
    calls, post = co.variant_calls(exp, refs, co.substitution_matrix(hmm.emissions), 0.3)
    out = dict(references=refs, where=where, truth=truth, threshold=0.3,
               calls=sorted([c[0], c[1], c[2], c[3]] for c in calls),
               n_columns_observed=len(exp),
               expectation_checksum=sum(sum(v.values()) for v in exp.values()))
    with open(os.path.join(HERE, "caller_case.json"), "w") as fh:
        json.dump(out, fh)
    got = {(c[0], c[1], c[2]) for c in calls}
    want = {(t[0], t[1], t[2]) for t in truth}
    print("calls %d, truth %d, precision %.3f recall %.3f" % (len(got), len(want), len(got & want) / max(1, len(got)),
                                                               len(got & want) / max(1, len(want))))


if __name__ == "__main__":
    main()
