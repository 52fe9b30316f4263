"""Generates tests/golden/realign_sam_golden.json from the UNMODIFIED reference CLI (build container only).

Input: the reference's own fixture tests/inputBigMutations.sam (30 aligned reads, hard / soft clips, both strands)
against tests/referencesMutated.fa.  For every aligned record the guide cigar is written exactly as
getExonerateCigarFormatString does (src/margin/utils.py:35-55) and oracle/_ref/cPecanRealign is run once per read with
realignCigarTargetFn's flags (src/margin/marginAlignLib.py:315-318) and the shipped model tests/input.hmm.  The printed
cigars are then spliced into the records the way realignSamFile3TargetFn does (marginAlignLib.py:340-363; written out
here with plain string handling because pysam is not available) and the resulting SAM text is stored next to the input.
"""
import json
import os
import re
import subprocess
import sys
import tempfile
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_TESTS = "/root/reference/tests"
BIN = os.path.join(ROOT, "oracle", "_ref", "cPecanRealign")
HMM = os.path.join(HERE, "input_model.hmm")


def fasta(path):
    out, name = {}, None
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            name = line[1:].split()[0]
            out[name] = []
        elif name:
            out[name].append(line)
    return {k: "".join(v) for k, v in out.items()}


def cli(ref_seq, read_seq, cigar_line):
    with tempfile.TemporaryDirectory() as d:
        with open(os.path.join(d, "ref.fa"), "w") as fh:
            fh.write(">ref\n%s\n" % ref_seq)
        with open(os.path.join(d, "read.fa"), "w") as fh:
            fh.write(">read\n%s\n" % read_seq)
        return subprocess.run([BIN, os.path.join(d, "ref.fa"), os.path.join(d, "read.fa"), "--diagonalExpansion=10",
                               "--splitMatrixBiggerThanThis=3000", "--loadHmm=" + HMM, "--gapGamma=0.5",
                               "--matchGamma=0.0"], input=cigar_line + "\n", capture_output=True, text=True,
                              check=True).stdout


def main():
    refs = fasta(os.path.join(REF_TESTS, "referencesMutated.fa"))
    lines = [l.rstrip("\n") for l in open(os.path.join(REF_TESTS, "inputBigMutations.sam")) if l.strip()]
    header = [l for l in lines if l.startswith("@")]
    recs = [l.split("\t") for l in lines if not l.startswith("@")]
    jobs = []
    for f in recs:
        if f[2] == "*" or int(f[1]) & 4:
            continue
        cig = [(int(n), op) for n, op in re.findall(r"(\d+)([MIDNSHP=X])", f[5])]
        assert all(op in "MIDSH" for _, op in cig)
        seq = f[9]
        qs = sum(n for n, op in cig[:2] if op == "S" and (cig.index((n, op)) == 0 or cig[0][1] == "H"))
        # query_alignment_start: leading soft clip (after an optional hard clip)
        k = 1 if cig and cig[0][1] == "H" else 0
        qs = cig[k][0] if len(cig) > k and cig[k][1] == "S" else 0
        qe = qs + sum(n for n, op in cig if op in "MI")
        rs = int(f[3]) - 1
        re_ = rs + sum(n for n, op in cig if op in "MD")
        body = " ".join("%s %d" % (op, n) for n, op in cig if op in "MID")
        # read and reference are given as the whole SEQ / whole reference, coordinates select the aligned parts
        line = "cigar: read %d %d + ref %d %d + 1 %s" % (qs, qe, rs, re_, body)
        jobs.append((f, cig, qs, qe, rs, re_, line))
    with ThreadPoolExecutor(8) as pool:
        outs = list(pool.map(lambda j: cli(refs[j[0][2]], j[0][9], j[6]), jobs))
    out_lines = []
    for (f, cig, qs, qe, rs, re_, _line), out in zip(jobs, outs):
        t = out.split()
        assert t[0] == "cigar:" and (int(t[2]), int(t[3]), int(t[6]), int(t[7])) == (qs, qe, rs, re_), out[:200]
        new = [(int(n), op) for op, n in zip(t[10::2], t[11::2])]
        ops = []
        if cig and cig[0][1] == "H":
            ops.append(cig[0])
        if qs > 0:
            ops.append((qs, "S"))
        ops += new
        if qe < len(f[9]):
            ops.append((len(f[9]) - qe, "S"))
        if len(cig) > 1 and cig[-1][1] == "H":
            ops.append(cig[-1])
        assert sum(n for n, op in ops if op in "MIS") == sum(n for n, op in cig if op in "MIS")
        assert sum(n for n, op in ops if op in "MD") == re_ - rs
        g = list(f)
        g[5] = "".join("%d%s" % (n, op) for n, op in ops)
        out_lines.append("\t".join(g))
    with open(os.path.join(HERE, "realign_sam_golden.json"), "w") as fh:
        json.dump({"references": refs, "sam_in": lines, "sam_out": header + out_lines,
                   "n_changed": sum(1 for (f, *_), o in zip(jobs, out_lines) if f[5] != o.split("\t")[5])}, fh)
    print(len(jobs), "records;", sum(1 for (f, *_), o in zip(jobs, out_lines) if f[5] != o.split("\t")[5]), "cigars changed")


if __name__ == "__main__":
    sys.exit(main())
