"""Golden vectors for the chaining stage from the reference's own Python (build container only).

The reference's chainFn / mergeChainedAlignedSegments (src/margin/marginAlignLib.py:9-157) and the position helpers
(src/margin/utils.py:12-33) are read from /root/reference, shimmed for Python 3 in memory (xrange -> range) and executed
on stand-in alignment objects that implement the handful of pysam.AlignedSegment attributes those functions touch
(pysam itself is not installed).  Inputs: the reference's tests/inputBigMutations.sam + tests/reads.fq, every alignment
cut into two or three local pieces (hard- or soft-clipped, a few operations dropped in between so that the chain has to
bridge a gap) plus decoys that must NOT be chained (overlapping in the read, wrong strand, too far away).
Output: tests/golden/chain_golden.json (input SAM text, reads, references, expected chained records).
"""
import json
import os
import random
import re

REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
COMP = str.maketrans("ACGTNacgtn", "TGCANtgcan")


class Seg:
    """Stand-in for pysam.AlignedSegment: only what chainFn / mergeChainedAlignedSegments read or write."""

    def __init__(self):
        self.query_name = None
        self.query_sequence = None
        self.reference_id = -1
        self.reference_start = 0
        self.is_reverse = False
        self.next_reference_id = -1
        self.cigar = ()

    @property
    def reference_end(self):
        return self.reference_start + sum(n for op, n in self.cigar if op in (0, 2, 3, 7, 8))

    @property
    def query_alignment_start(self):
        n = 0
        for op, length in self.cigar:
            if op == 5:
                continue
            if op == 4:
                n += length
            else:
                break
        return n

    @property
    def query_alignment_end(self):
        return self.query_alignment_start + sum(n for op, n in self.cigar if op in (0, 1, 7, 8))


def grab(path, start_pat, stop_pat):
    lines = open(os.path.join(REF, path)).read().split("\n")
    i = next(k for k, l in enumerate(lines) if re.match(start_pat, l))
    j = next(k for k in range(i + 1, len(lines)) if re.match(stop_pat, lines[k]))
    return "\n".join(lines[i:j]) + "\n"


def load_reference_functions():
    class FakePysam:
        AlignedSegment = Seg

    env = {"pysam": FakePysam, "reverseComplement": lambda s: s.translate(COMP)[::-1]}
    exec(grab("src/margin/utils.py", r"^def getFirstNonClippedPositionInRead", r"^def getExonerateCigarFormatString"), env)
    src = grab("src/margin/marginAlignLib.py", r"^def mergeChainedAlignedSegments", r"^def chainSamFile")
    exec(src.replace("xrange", "range"), env)
    return env


def parse_cigar(text):
    return [("MIDNSHP=X".index(op), int(n)) for n, op in re.findall(r"(\d+)([MIDNSHP=X])", text)]


def cigar_text(ops):
    return "".join("%d%s" % (n, "MIDNSHP=X"[op]) for op, n in ops)


def trim_to_matches(ops):
    while ops and ops[0][0] != 0:
        ops = ops[1:]
    while ops and ops[-1][0] != 0:
        ops = ops[:-1]
    return ops


def pieces_of(fields, rng, wide_gap):
    """Cuts one SAM record into local alignments; returns SAM field lists."""
    cig = parse_cigar(fields[5])
    seq = fields[9]
    lead = cig[0][1] if cig[0][0] == 4 else 0
    core = [c for c in cig if c[0] in (0, 1, 2)]
    n = len(core)
    cuts = sorted({int(n * 0.4), int(n * 0.7)})
    spans, a = [], 0
    for c in cuts + [n]:
        spans.append((a, c))
        a = c + (rng.randint(0, 3) if not wide_gap else 60)  # operations dropped between the pieces
    out = []
    for k, (i0, i1) in enumerate(spans):
        ops = core[i0:i1]
        lo = 0
        while lo < len(ops) and ops[lo][0] != 0:
            lo += 1
        ops_t = trim_to_matches(ops)
        if not ops_t:
            continue
        first = i0 + lo
        ref0 = int(fields[3]) - 1 + sum(m for op, m in core[:first] if op in (0, 2))
        q0 = lead + sum(m for op, m in core[:first] if op in (0, 1))
        q1 = q0 + sum(m for op, m in ops_t if op in (0, 1))
        hard = (k + int(fields[3])) % 2 == 0
        clip = 5 if hard else 4
        c = ([(clip, q0)] if q0 else []) + ops_t + ([(clip, len(seq) - q1)] if q1 < len(seq) else [])
        f = list(fields[:11])
        f[3], f[5], f[9], f[10] = str(ref0 + 1), cigar_text(c), (seq[q0:q1] if hard else seq), "*"
        out.append(f)
    # a decoy: the start of the first piece again, on the other strand -- shorter, and never chainable with the rest
    d = list(out[0])
    dc = [c for c in parse_cigar(d[5]) if c[0] in (0, 1, 2)][:3]
    dc = trim_to_matches(dc)
    d[1] = str(int(d[1]) ^ 16)
    q = sum(m for op, m in dc if op in (0, 1))
    full = len(seq)
    d[5], d[9] = cigar_text(dc + ([(4, full - q)] if q < full else [])), seq
    out.insert(1, d)
    return out


def main():
    env = load_reference_functions()
    tests = os.path.join(REF, "tests")
    lines = [l.rstrip("\n") for l in open(os.path.join(tests, "inputBigMutations.sam")) if l.strip()]
    header = [l for l in lines if l.startswith("@")]
    ref_names = [re.search(r"SN:(\S+)", h).group(1) for h in header if h.startswith("@SQ")]
    refs, name = {}, None
    for l in open(os.path.join(tests, "referencesMutated.fa")):
        l = l.strip()
        if l.startswith(">"):
            name = l[1:].split()[0]
            refs[name] = ""
        elif name:
            refs[name] += l
    fq = open(os.path.join(tests, "reads.fq")).read().split("\n")
    reads = [(fq[i][1:], fq[i + 1], fq[i + 3]) for i in range(0, len(fq) - 3, 4)]
    rng = random.Random(17)
    sam_in, want = list(header), []
    records = [l.split("\t") for l in lines if not l.startswith("@")][:12]  # 12 reads keep the fixture small
    by_name = {}
    for k, f in enumerate(records):
        for p in pieces_of(f, rng, wide_gap=(k % 7 == 3)):
            sam_in.append("\t".join(p))
            s = Seg()
            s.query_name, s.query_sequence = p[0], p[9]
            s.reference_id, s.reference_start = ref_names.index(p[2]), int(p[3]) - 1
            s.is_reverse, s.cigar = bool(int(p[1]) & 16), tuple(parse_cigar(p[5]))
            by_name.setdefault(p[0], []).append(s)
    chained = []
    for rname, seq, _q in reads:
        rname = rname.split()[0]
        if rname in by_name:
            segs = by_name.pop(rname)
            ref = refs[ref_names[segs[0].reference_id]]
            chain = env["chainFn"](segs, ref, seq)
            chained.append((len(chain), env["mergeChainedAlignedSegments"](chain, ref, seq)))
    assert not by_name
    chained.sort(key=lambda c: (ref_names[c[1].reference_id], c[1].reference_start, c[1].reference_end))
    for n_chain, c in chained:
        want.append({"qname": c.query_name, "reverse": bool(c.is_reverse), "rname": ref_names[c.reference_id],
                     "reference_start": c.reference_start, "cigar": cigar_text(c.cigar), "seq": c.query_sequence,
                     "n_chained": n_chain})
    with open(os.path.join(HERE, "chain_golden.json"), "w") as fh:
        json.dump({"sam_in": sam_in, "reads": reads, "references": refs, "chained": want}, fh)
    print(len(sam_in) - len(header), "local alignments ->", len(want), "chained;", "members per chain:",
          sorted({w["n_chained"] for w in want}))


if __name__ == "__main__":
    main()
