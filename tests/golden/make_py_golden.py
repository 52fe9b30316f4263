"""Golden vectors from the reference's own PYTHON code for the host-side rows.

The reference is Python 2 (xrange, list-returning map, integer `/`) and imports pysam / jobTree at module level, so its
modules cannot be imported here.  This script reads the source of the pure functions it needs straight from
/root/reference, applies three mechanical Python-2 -> 3 shims to the TEXT IN MEMORY (nothing is copied into the repo),
executes it and stores inputs + outputs in tests/golden/py_reference_vectors.json:

  submodules/cPecan/cPecanEm.py:19-105     class Hmm: normalise, randomise, equalise, setEmissionsToJukesCantor,
                                           tieEmissions, loadHmm / addExpectationsFile
  src/margin/marginCallerLib.py:76-106     getProb, calcBasePosteriorProbs, loadHmmSubstitutionMatrix,
                                           getNullSubstitutionMatrix
  src/margin/marginAlignLib.py:207-256     the two model post-processing steps of learnModelFromSamFileTargetFn2
                                           (normaliseHmmByReferenceGCContent, setHmmIndelEmissionsToBeFlat)

Shims: xrange -> range; map -> list(map) (Python 2 map returns a list); `/` between integers -> `//` at the three
places the reference relies on integer division (tieEmissions' `i / SYMBOL_NUMBER`, loadHmmSubstitutionMatrix's `i/4`).
Run in the build container only:  python tests/golden/make_py_golden.py
"""
import builtins
import functools
import json
import math
import os
import random
import re
from itertools import product

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def grab(path, start_pat, stop_pat):
    """Source text from the line matching start_pat up to (not including) the next line matching stop_pat."""
    lines = open(os.path.join(REF, path)).read().split("\n")
    i = next(k for k, l in enumerate(lines) if re.match(start_pat, l))
    j = next(k for k in range(i + 1, len(lines)) if re.match(stop_pat, lines[k]))
    return "\n".join(lines[i:j]) + "\n"


def py3(src):
    src = src.replace("xrange", "range")
    src = src.replace("(i / SYMBOL_NUMBER)", "(i // SYMBOL_NUMBER)")        # cPecanEm.py:99-100
    src = src.replace("m[4*(i/4):4*(1 + i/4)]", "m[4*(i//4):4*(1 + i//4)]")  # marginCallerLib.py:99
    return src


def main():
    env = {"SYMBOL_NUMBER": 4, "random": random, "math": math, "product": product, "reduce": functools.reduce,
           "map": lambda f, *a: list(builtins.map(f, *a)), "BASES": "ACGT"}
    exec(py3(grab("submodules/cPecan/cPecanEm.py", r"^class Hmm", r"^def ")), env)
    exec(py3(grab("src/margin/marginCallerLib.py", r"^def getProb", r"^def vcfRead")), env)
    Hmm = env["Hmm"]
    out = {"hmm": [], "caller": [], "postprocess": []}

    # ---- Hmm methods
    for model_type in ("fiveState", "fiveStateAsymmetric", "threeState", "threeStateAsymmetric"):
        for seed in (1, 7):
            random.seed(seed)
            h = Hmm(model_type)
            h.randomise()
            rec = {"modelType": model_type, "seed": seed, "randomise": [list(h.transitions), list(h.emissions)]}
            # un-normalised expectations -> normalise (the M-step, cPecanEm.py:184-188)
            rng = random.Random(100 + seed)
            raw_t = [rng.random() * 50 + 1e-12 for _ in h.transitions]
            raw_e = [rng.random() * 900 + 1e-12 for _ in h.emissions]
            h2 = Hmm(model_type)
            h2.transitions, h2.emissions = list(raw_t), list(raw_e)
            h2.normalise()
            rec["raw"] = [raw_t, raw_e]
            rec["normalise"] = [list(h2.transitions), list(h2.emissions)]
            h2.tieEmissions()
            rec["tieEmissions"] = list(h2.emissions)
            h3 = Hmm(model_type)
            h3.equalise()
            h3.setEmissionsToJukesCantor(0.2)
            rec["equalise_jc"] = [list(h3.transitions), list(h3.emissions)]
            out["hmm"].append(rec)
    model_path = os.path.join(HERE, "input_model.hmm")
    h = Hmm.loadHmm(model_path)
    out["load"] = {"modelType": h.modelType, "transitions": list(h.transitions), "emissions": list(h.emissions),
                   "likelihood": h.likelihood}

    # ---- marginCaller column posteriors
    err = env["loadHmmSubstitutionMatrix"](model_path)
    null = env["getNullSubstitutionMatrix"]()
    out["substitution_matrix"] = [err[k] for k in product("ACGT", "ACGT")]
    rng = random.Random(5)
    for k in range(40):
        obs = {b: (rng.random() * (30 if rng.random() < 0.7 else 0.5)) for b in "ACGT"}
        tot = sum(obs.values())
        obs = {b: v / tot for b, v in obs.items()}  # variantCallSamFileTargetFn normalises before the call (:196-199)
        ref_base = "ACGTacgt"[k % 8]
        for name, evo in (("null", null), ("err", err)):
            post = env["calcBasePosteriorProbs"](obs, ref_base, evo, err)
            out["caller"].append({"obs": [obs[b] for b in "ACGT"], "ref": ref_base, "evo": name,
                                  "post": [post[b] for b in "ACGT"]})

    # ---- model post-processing after EM (learnModelFromSamFileTargetFn2, marginAlignLib.py:227-232: load,
    # setHmmIndelEmissionsToBeFlat, normaliseHmmByReferenceGCContent(hmm, 0.5), write) + the variation-rate helper
    env["np"] = __import__("numpy")
    exec(py3(grab("src/margin/marginAlignLib.py", r"^toMatrix = ", r"^def realignSamFileTargetFn")
             .replace("i / SYMBOL_NUMBER else", "i // SYMBOL_NUMBER else")), env)   # marginAlignLib.py:260-261
    for model_type in ("fiveStateAsymmetric", "threeState"):
        random.seed(11)
        h = Hmm(model_type)
        h.randomise()
        before = [list(h.transitions), list(h.emissions)]
        env["setHmmIndelEmissionsToBeFlat"](h)
        flat = list(h.emissions)
        env["normaliseHmmByReferenceGCContent"](h, 0.5)
        gc50 = list(h.emissions)
        h2 = Hmm(model_type)
        h2.transitions, h2.emissions = list(before[0]), list(before[1])
        env["normaliseHmmByReferenceGCContent"](h2, 0.35)
        gc35 = list(h2.emissions)
        h3 = Hmm(model_type)
        h3.transitions, h3.emissions = list(before[0]), list(before[1])
        env["modifyHmmEmissionsByExpectedVariationRate"](h3, 0.1)
        out["postprocess"].append({"modelType": model_type, "before": before, "flat": flat, "flat_gc50": gc50,
                                   "gc35": gc35, "variation_0.1": [float(v) for v in h3.emissions]})
    # ---- marginCaller collation + SNV loop (marginCallerLib.py:59-69 and :189-209): the loop bodies are cut out of the
    # two job functions (whose surroundings are jobTree / pysam plumbing) and run on a posterior file written the way
    # cPecanRealign's writePosteriorProbs does ("%d\t%d\t%f", cPecanRealign.c:346-351) from posteriors of the reference C
    # library (oracle/_ref) for three short fixture reads, one of them with an N and a lower-case base planted
    import sys
    import tempfile
    import textwrap
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    from oracle import pyoracle as po
    from oracle import pyref
    lib_src = open(os.path.join(REF, "src/margin/marginCallerLib.py")).read().split("\n")
    i0 = next(k for k, l in enumerate(lib_src) if "#Now collate the reference position expectations" in l)
    i1 = next(k for k, l in enumerate(lib_src) if "#Pickle the posterior probs" in l)
    collate_src = textwrap.dedent("\n".join(lib_src[i0:i1]))
    j0 = next(k for k, l in enumerate(lib_src) if "#Now do the SNV calculations" in l)
    j1 = next(k for k, l in enumerate(lib_src) if "# Write to VCF" in l)
    snv_src = textwrap.dedent("\n".join(lib_src[j0:j1]))
    case = json.load(open(os.path.join(HERE, "caller_case.json")))
    alns = json.load(open(os.path.join(HERE, "bigmutations_alignments.json")))
    model = pyref.RefModel.load(model_path)
    picked, reads_in, pairs_in = [13, 19, 27], [], []

    class Opt:
        noMargin = False
        threshold = case["threshold"]

    expectations = {}
    for n, k in enumerate(picked):
        a = alns[k]
        y = a["y"]
        if n == 0:
            y = y[:40] + "N" + y[41:60] + y[60:80].lower() + y[80:]
        anc = po.anchors_from_cigar(a["ops"], a["lens"], a["x"], y)
        pairs = pyref.aligned_pairs(model, a["x"], y, anc, expansion=10, split=100 * 100)
        with tempfile.NamedTemporaryFile("w", suffix=".txt", delete=False) as fh:
            for p_int, x, yy in pairs.tolist():
                fh.write("%d\t%d\t%f\n" % (case["where"][k][1] + x, yy, float(p_int) / 10000000))
            path = fh.name
        env2 = dict(env)
        env2.update(tempPosteriorProbsFile=path, referenceSequenceName=case["where"][k][0], querySequence=y,
                    options=Opt, expectationsOfBasesAtEachPosition=expectations)
        exec(collate_src, env2)
        os.remove(path)
        reads_in.append({"aln": k, "y": y})
        pairs_in.append(pairs.tolist())
    env3 = dict(env)
    env3.update(expectationsOfBasesAtEachPosition=expectations, refSequences=case["references"], options=Opt,
                variantCalls=[], errorSubstitutionMatrix=err, evolutionarySubstitutionMatrix=null)
    exec(snv_src, env3)
    out["collate"] = {"reads": reads_in, "pairs": pairs_in,
                      "expectations": [[k[0], k[1]] + [v[b] for b in "ACGT"] for k, v in expectations.items()],
                      "calls": [list(c) for c in env3["variantCalls"]], "threshold": case["threshold"]}
    with open(os.path.join(HERE, "py_reference_vectors.json"), "w") as fh:
        json.dump(out, fh)
    print({k: len(v) for k, v in out.items()})


if __name__ == "__main__":
    main()
