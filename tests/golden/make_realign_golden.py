"""Generates tests/golden/realign_cases.json from the UNMODIFIED reference (run in the build container only).

For every case the reference CLI `oracle/_ref/cPecanRealign` (compiled from /root/reference by oracle/Makefile) is run
the way realignCigarTargetFn does (src/margin/marginAlignLib.py:315-318): one process per read,
  echo <cigar> | cPecanRealign --diagonalExpansion=10 --splitMatrixBiggerThanThis=3000 --gapGamma=G --matchGamma=M
                 --loadHmm=tests/golden/input_model.hmm ref.fa read.fa
and the cigar it prints is stored.  The chain of aligned pairs comes from the reference library entry points
(getAlignedPairsUsingAnchors -> reweightAlignedPairs2 -> filterPairwiseAlignmentToMakePairsOrdered through
oracle/ref_shim.c) after srand(1), which is the state a fresh cPecanRealign process starts in.
"""
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import pyref  # noqa: E402
import marginalign_trna_b200 as mab  # noqa: E402

LETTERS = "MDI"
HMM = os.path.join(HERE, "input_model.hmm")


def cli_realign(x, y, ops, lens, gap_gamma, match_gamma, expansion=10, split=3000):
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "ref.fa"), "w") as fh:
            fh.write(">ref\n%s\n" % x)
        with open(os.path.join(d, "read.fa"), "w") as fh:
            fh.write(">read\n%s\n" % y)
        body = " ".join("%s %d" % (LETTERS[o], n) for o, n in zip(ops, lens))
        line = "cigar: read 0 %d + ref 0 %d + 1.0 %s\n" % (len(y), len(x), body)
        out = subprocess.run([pyref.REALIGN_BIN, "--diagonalExpansion=%d" % expansion,
                              "--splitMatrixBiggerThanThis=%d" % split, "--gapGamma=%s" % gap_gamma,
                              "--matchGamma=%s" % match_gamma, "--loadHmm=" + HMM,
                              os.path.join(d, "ref.fa"), os.path.join(d, "read.fa")],
                             input=line, capture_output=True, text=True, check=True).stdout
    f = out.split()
    assert f[0] == "cigar:" and f[1] == "read" and f[5] == "ref", out
    assert (int(f[2]), int(f[3]), int(f[6]), int(f[7])) == (0, len(y), 0, len(x)), out
    return [LETTERS.index(c) for c in f[10::2]], [int(v) for v in f[11::2]]


def lib_chain(model, x, y, ops, lens, gap_gamma, match_gamma, expansion=10, split=3000):
    from oracle import pyoracle as po
    anchors = po.anchors_from_cigar(ops, lens, x, y)
    pairs = pyref.aligned_pairs(model, x, y, anchors, expansion=expansion, split=split * split)
    chain = pyref.realign_tail(pairs, x, y, gap_gamma, match_gamma, seed=1)
    return len(pairs), chain


def main():
    pyref.build()
    hmm = mab.Hmm.loadHmm(HMM)
    model = pyref.RefModel.load(HMM)
    cases = []

    def add(name, x, y, ops, lens, gg=0.5, mg=0.0):
        x, y = (s.decode() if isinstance(s, bytes) else s for s in (x, y))
        ops, lens = [int(o) for o in ops], [int(n) for n in lens]
        c_ops, c_lens = cli_realign(x, y, ops, lens, gg, mg)
        n_pairs, chain = lib_chain(model, x, y, ops, lens, gg, mg)
        cases.append(dict(name=name, x=x, y=y, ops=ops, lens=lens, gap_gamma=gg, match_gamma=mg, n_pairs=n_pairs,
                          out_ops=c_ops, out_lens=c_lens, chain=np.asarray(chain).reshape(-1).tolist()))
        print(name, len(x), len(y), n_pairs, len(chain), len(c_ops))

    # the reference's own fixture alignments (tests/inputBigMutations.sam vs tests/referencesMutated.fa): first 6
    fixture = json.load(open(os.path.join(HERE, "bigmutations_alignments.json")))
    for a in fixture[:6]:
        add("fixture:" + a["name"], a["x"], a["y"], a["ops"], a["lens"])
    add("fixture-g0.2-m0.3:" + fixture[1]["name"], fixture[1]["x"], fixture[1]["y"], fixture[1]["ops"],
        fixture[1]["lens"], 0.2, 0.3)
    add("fixture-g0:" + fixture[2]["name"], fixture[2]["x"], fixture[2]["y"], fixture[2]["ops"], fixture[2]["lens"],
        0.0, 0.0)
    # synthetic tRNA-length alignments, several gamma settings
    b = mab.synth.make_trna_batch(24, seed=21)
    gammas = [(0.5, 0.0), (0.5, 0.0), (0.5, 0.0), (0.2, 0.3), (0.9, 0.5), (0.0, 0.0)]
    for i in range(b.n):
        gg, mg = gammas[i % len(gammas)]
        add("trna%d" % i, b.x(i), b.y(i), *b.cigar(i), gg, mg)
    with open(os.path.join(HERE, "realign_cases.json"), "w") as fh:
        json.dump(cases, fh)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
