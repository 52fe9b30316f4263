#!/usr/bin/env python
"""Generates the golden fixtures under tests/golden/ from the UNMODIFIED reference C code.

Runs only in the build container (needs /root/reference and oracle/_ref built by `make -C oracle ref`).  The outputs
are small and committed; nothing under tests/ reads /root/reference at test time.

  input_model.hmm          copy of the reference's 5-state asymmetric fixture model (tests/input.hmm), re-written
                           through our Hmm class so the numbers are identical
  fixture_alignments.json  the two guide alignments of tests/input.sam as (X, Y, cigar ops/lens)
  cases.npz                for every case: pairs (p_int,x,y) from getAlignedPairsUsingAnchors and the
                           expectation vector (T,E,lik) from getExpectationsUsingAnchors, full precision
  cases.json               the inputs of every case (sequences, anchors or cigar, model, parameters)
  bigmutations_digest.json per-read digests (count, sha256, expectation vector) for the 30 reads of
                           tests/inputBigMutations.sam vs tests/referencesMutated.fa
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import marginalign_trna_b200 as mab  # noqa: E402
from marginalign_trna_b200 import seqio, synth  # noqa: E402
from oracle import pyoracle as po  # noqa: E402
from oracle import pyref as pr  # noqa: E402

REF = "/root/reference"


def load_hmm_arrays(path):
    h = mab.Hmm.loadHmm(path)
    return h._modelTypeInt(), h.transitions.copy(), h.emissions.copy()


def alignments_from_sam(sam_path, fasta_path):
    _, recs = seqio.sam_read(sam_path)
    refs = seqio.fasta_dict(fasta_path)
    out = []
    for r in recs:
        if r.is_unmapped:
            continue
        g = seqio.guide_from_sam(r)
        out.append(dict(name=r.qname, x=refs[g.ref_name][g.ref_start:g.ref_end], y=r.seq[g.read_start:g.read_end],
                        ops=list(map(int, g.ops)), lens=list(map(int, g.lens))))
    return out


def random_seq(rng, n, alphabet="ACGT"):
    return "".join(rng.choice(list(alphabet), size=n)) if n else ""


def mutate(rng, s, rate=0.15):
    """Returns (read, ops, lens): the read and its generating alignment."""
    out, ops = [], []
    for i, ch in enumerate(s):
        u = rng.random()
        edge = i == 0 or i == len(s) - 1
        if u < rate / 3 and not edge:
            ops.append(1)
        else:
            if u < 2 * rate / 3:
                ch = rng.choice([c for c in "ACGT" if c != ch.upper()])
            out.append(ch)
            ops.append(0)
        if rng.random() < rate / 3 and i != len(s) - 1:
            out.append(rng.choice(list("ACGT")))
            ops.append(2)
    rl_ops, rl_lens = [], []
    for o in ops:
        if rl_ops and rl_ops[-1] == o:
            rl_lens[-1] += 1
        else:
            rl_ops.append(o)
            rl_lens.append(1)
    return "".join(out), rl_ops, rl_lens


def build_cases():
    rng = np.random.default_rng(20240611)
    typ, T5, E5 = load_hmm_arrays(os.path.join(REF, "tests", "input.hmm"))
    T3, E3 = synth.three_state_from_five(T5, E5)
    models = {
        "fixture5asym": (1, T5, E5),
        "three_asym": (3, T3, E3),
        "three_sym": (2, T3, E3),
        "rand5sym": (0,) + synth.random_hmm(0, 11),
        "rand5asym": (1,) + synth.random_hmm(1, 12),
        "rand3sym": (2,) + synth.random_hmm(2, 13),
        "rand3asym": (3,) + synth.random_hmm(3, 14),
        "default5": None,  # stateMachine5_construct(fiveState)
    }
    base = dict(threshold=0.01, min_diags=1000, traceback_diagonals=40, expansion=10, split=3000, ragged_left=1,
                ragged_right=1)
    cases = []

    def add(name, x, y, model, cigar=None, anchors=None, **kw):
        p = dict(base)
        p.update(kw)
        c = dict(name=name, x=x, y=y, model=model, params=p)
        if cigar is not None:
            c["ops"], c["lens"] = list(map(int, cigar[0])), list(map(int, cigar[1]))
        else:
            c["anchors"] = [list(map(int, a)) for a in anchors]
        cases.append(c)

    # (1) the reference's own fixture reads
    for i, a in enumerate(alignments_from_sam(os.path.join(REF, "tests", "input.sam"),
                                              os.path.join(REF, "tests", "references.fa"))):
        add("fixture%d_realign" % i, a["x"], a["y"], "fixture5asym", cigar=(a["ops"], a["lens"]))
        if i == 0:
            add("fixture0_caller_split100", a["x"], a["y"], "fixture5asym", cigar=(a["ops"], a["lens"]), split=100)
            add("fixture0_em_split300", a["x"], a["y"], "fixture5asym", cigar=(a["ops"], a["lens"]), split=300)
            add("fixture0_three_asym", a["x"], a["y"], "three_asym", cigar=(a["ops"], a["lens"]))
            add("fixture0_three_sym", a["x"], a["y"], "three_sym", cigar=(a["ops"], a["lens"]))
            add("fixture0_default5", a["x"], a["y"], "default5", cigar=(a["ops"], a["lens"]))
            add("fixture0_exp0", a["x"], a["y"], "fixture5asym", cigar=(a["ops"], a["lens"]), expansion=0, split=1)
            add("fixture0_exp40", a["x"], a["y"], "fixture5asym", cigar=(a["ops"], a["lens"]), expansion=40)
    # (2) the reference's unit-test example (test_diagonalDPCalculations): AGCG vs AGTTCG, no anchors
    add("kat_agcg", "AGCG", "AGTTCG", "default5", anchors=[], expansion=2, threshold=0.2, ragged_left=0, ragged_right=0)
    add("kat_agcg_ragged", "AGCG", "AGTTCG", "default5", anchors=[], expansion=2, threshold=0.2)
    # (3) tRNA-length synthetic pairs across all model types, with N and lower case mixed in
    for k, model in enumerate(["fixture5asym", "three_asym", "three_sym", "rand5sym", "rand5asym", "rand3sym",
                               "rand3asym", "default5"]):
        for rep in range(3):
            x = random_seq(rng, int(rng.integers(70, 150)))
            y, ops, lens = mutate(rng, x)
            if rep == 1:  # sprinkle N / lower case / IUPAC
                xl, yl = list(x), list(y)
                for _ in range(4):
                    xl[int(rng.integers(len(xl)))] = rng.choice(list("NnacgtR"))
                    yl[int(rng.integers(len(yl)))] = rng.choice(list("NnacgtY"))
                x, y = "".join(xl), "".join(yl)
            add("trna_%s_%d" % (model, rep), x, y, model, cigar=(ops, lens),
                ragged_left=int(rep != 2), ragged_right=int(rep != 2))
    # (4) short checkpoint distance -> many tracebacks on a short pair; explicit sparse anchors; splits
    x = random_seq(rng, 400)
    y, ops, lens = mutate(rng, x)
    add("multi_tb", x, y, "fixture5asym", cigar=(ops, lens), min_diags=100, traceback_diagonals=20)
    add("multi_tb_three", x, y, "rand3asym", cigar=(ops, lens), min_diags=60, traceback_diagonals=10, expansion=4)
    anc = po.anchors_from_cigar(ops, lens, x, y)
    add("sparse_anchors", x, y, "fixture5asym", anchors=anc[::25], expansion=20)
    add("sparse_anchors_split", x, y, "fixture5asym", anchors=anc[::40], expansion=6, split=20)
    add("no_anchors_small", x[:60], y[:50], "rand5asym", anchors=[], expansion=10)
    add("no_anchors_split", x[:200], y[:180], "fixture5asym", anchors=[], expansion=10, split=50)
    add("split_not_ragged", x, y, "fixture5asym", anchors=anc[::40], expansion=6, split=20, ragged_left=0,
        ragged_right=0)
    # (5) degenerate shapes
    add("empty_x", "", "ACGTAC", "fixture5asym", anchors=[])
    add("empty_y", "ACGTAC", "", "fixture5asym", anchors=[])
    add("empty_both", "", "", "fixture5asym", anchors=[])
    add("single", "A", "A", "fixture5asym", anchors=[[0, 0]])
    add("all_n", "NNNNNNNN", "NNNNNNN", "fixture5asym", anchors=[])
    return models, cases


def run_case(models, c):
    spec = models[c["model"]]
    rm = pr.RefModel.default5() if spec is None else pr.RefModel.from_hmm(*spec)
    if "anchors" in c:
        anc = np.asarray(c["anchors"], dtype=np.int64).reshape(-1, 2)
    else:
        anc = po.anchors_from_cigar(c["ops"], c["lens"], c["x"], c["y"])
    p = c["params"]
    kw = dict(threshold=p["threshold"], min_diags=p["min_diags"], traceback_diagonals=p["traceback_diagonals"],
              expansion=p["expansion"], split=p["split"] ** 2, ragged_left=p["ragged_left"],
              ragged_right=p["ragged_right"])
    pairs = pr.aligned_pairs(rm, c["x"], c["y"], anc, **kw)
    T, E, lik = pr.expectations(rm, c["x"], c["y"], anc, **kw)
    return pairs, np.concatenate([T, E, [lik]])


def main():
    assert pr.available() or os.path.exists(REF), "reference oracle not available"
    pr.build()
    # model fixture
    h = mab.Hmm.loadHmm(os.path.join(REF, "tests", "input.hmm"))
    h.write(os.path.join(HERE, "input_model.hmm"))
    h2 = mab.Hmm.loadHmm(os.path.join(HERE, "input_model.hmm"))
    assert np.array_equal(h.transitions, h2.transitions) and np.array_equal(h.emissions, h2.emissions)
    models, cases = build_cases()
    arrays = {}
    for c in cases:
        pairs, exp = run_case(models, c)
        arrays[c["name"] + "/pairs"] = pairs
        arrays[c["name"] + "/exp"] = exp
        print("%-28s lX %5d lY %5d pairs %6d lik %.6f" % (c["name"], len(c["x"]), len(c["y"]), len(pairs), exp[-1]))
    np.savez_compressed(os.path.join(HERE, "cases.npz"), **arrays)
    jm = {k: (None if v is None else dict(type=int(v[0]), T=[float(t) for t in v[1]], E=[float(e) for e in v[2]]))
          for k, v in models.items()}
    with open(os.path.join(HERE, "cases.json"), "w") as fh:
        json.dump(dict(models=jm, cases=cases), fh)
    # reference known answers for the structural pieces (bands / split points / logAdd samples)
    rng = np.random.default_rng(7)
    la = []
    for _ in range(2000):
        a, b = float(rng.uniform(-30, 0)), float(rng.uniform(-30, 0))
        la.append([a, b, pr.logadd(a, b)])
    for a, b in [(-np.inf, -1.0), (-1.0, -np.inf), (-np.inf, -np.inf), (-2.0, -2.0), (-1.0, -8.5), (-1.0, -8.5000001),
                 (0.0, -1.0), (0.0, -2.5), (0.0, -4.5), (0.0, -7.5)]:
        la.append([a, b, pr.logadd(a, b)])
    np.save(os.path.join(HERE, "logadd_samples.npy"), np.array(la))
    # big-mutations digests
    typ, T5, E5 = load_hmm_arrays(os.path.join(REF, "tests", "input.hmm"))
    rm = pr.RefModel.from_hmm(typ, T5, E5)
    dig = []
    alns = alignments_from_sam(os.path.join(REF, "tests", "inputBigMutations.sam"),
                               os.path.join(REF, "tests", "referencesMutated.fa"))
    for a in alns:
        anc = po.anchors_from_cigar(a["ops"], a["lens"], a["x"], a["y"])
        pairs = pr.aligned_pairs(rm, a["x"], a["y"], anc, expansion=10, split=100 ** 2)
        T, E, lik = pr.expectations(rm, a["x"], a["y"], anc, expansion=10, split=300 ** 2)
        dig.append(dict(name=a["name"], lX=len(a["x"]), lY=len(a["y"]), n_pairs=int(len(pairs)),
                        sha256=hashlib.sha256(np.ascontiguousarray(pairs).tobytes()).hexdigest(),
                        p_sum=int(pairs[:, 0].sum()), exp=[float(v) for v in np.concatenate([T, E, [lik]])]))
    with open(os.path.join(HERE, "bigmutations_digest.json"), "w") as fh:
        json.dump(dig, fh)
    with open(os.path.join(HERE, "bigmutations_alignments.json"), "w") as fh:
        json.dump(alns, fh)
    print("wrote", len(cases), "cases and", len(dig), "digests")


if __name__ == "__main__":
    main()
