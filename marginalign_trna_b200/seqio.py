"""Pure-Python FASTA / FASTQ / SAM / exonerate-cigar I/O for the realign and EM call sites.

Replaces what the reference gets from pysam and sonLib.bioio (neither exists for Python 3 here):
  fastaRead / fastqRead / fastaWrite      submodules/sonLib/bioio.py
  getExonerateCigarFormatString           src/margin/utils.py:35-55
  cigarRead / cigarWrite (exonerate text) submodules/sonLib/C/impl/pairwiseAlignment.c:135-240
Op codes follow sonLib: 0 = M (PAIRWISE_MATCH), 1 = D (PAIRWISE_INDEL_X, consumes reference only),
2 = I (PAIRWISE_INDEL_Y, consumes read only).
"""
import re
from dataclasses import dataclass, field
from typing import Iterator, List, Tuple

OP_MATCH, OP_INDEL_X, OP_INDEL_Y = 0, 1, 2
_SAM_CIGAR_RE = re.compile(r"(\d+)([MIDNSHP=X])")


def fasta_read(path) -> Iterator[Tuple[str, str]]:
    name, chunks = None, []
    with open(path) as fh:
        for line in fh:
            line = line.rstrip("\n")
            if line.startswith(">"):
                if name is not None:
                    yield name, "".join(chunks)
                name, chunks = line[1:], []
            elif line and name is not None:
                chunks.append(line.strip())
    if name is not None:
        yield name, "".join(chunks)


def fasta_dict(path):
    """First word of each header -> sequence (utils.py getFastaDictionary)."""
    d = {}
    for name, seq in fasta_read(path):
        key = name.split()[0]
        if key in d:
            raise RuntimeError("duplicate fasta sequence name: %s" % key)
        d[key] = seq
    return d


def fasta_write(fh, name, seq, width=100):
    fh.write(">%s\n" % name)
    for i in range(0, len(seq), width):
        fh.write(seq[i:i + width] + "\n")


def fastq_read(path) -> Iterator[Tuple[str, str, str]]:
    with open(path) as fh:
        while True:
            head = fh.readline()
            if not head:
                return
            head = head.rstrip("\n")
            if not head:
                continue
            if not head.startswith("@"):
                raise RuntimeError("malformed fastq header: %r" % head)
            seq = fh.readline().rstrip("\n")
            plus = fh.readline()
            if not plus.startswith("+"):
                raise RuntimeError("malformed fastq record for %s" % head)
            qual = fh.readline().rstrip("\n")
            yield head[1:], seq, qual


@dataclass
class SamRecord:
    qname: str
    flag: int
    rname: str
    pos: int          # 1-based, as in the file
    mapq: int
    cigar: List[Tuple[str, int]]
    rnext: str
    pnext: int
    tlen: int
    seq: str
    qual: str
    tags: List[str] = field(default_factory=list)

    @property
    def is_reverse(self):
        return bool(self.flag & 16)

    @property
    def is_unmapped(self):
        return bool(self.flag & 4) or self.rname == "*"

    @property
    def reference_start(self):
        return self.pos - 1

    @property
    def reference_end(self):
        return self.reference_start + sum(n for op, n in self.cigar if op in "MDN=X")

    @property
    def query_alignment_start(self):
        """Offset of the first aligned base in SEQ (soft clips are in SEQ, hard clips are not)."""
        n = 0
        for op, length in self.cigar:
            if op == "H":
                continue
            if op == "S":
                n += length
            else:
                break
        return n

    @property
    def query_alignment_end(self):
        return self.query_alignment_start + sum(n for op, n in self.cigar if op in "MI=X")

    def cigar_string(self):
        return "".join("%d%s" % (n, op) for op, n in self.cigar) or "*"

    def to_line(self):
        cols = [self.qname, str(self.flag), self.rname, str(self.pos), str(self.mapq), self.cigar_string(), self.rnext,
                str(self.pnext), str(self.tlen), self.seq, self.qual] + self.tags
        return "\t".join(cols)


def parse_sam_cigar(text):
    if text == "*":
        return []
    out = [(op, int(n)) for n, op in _SAM_CIGAR_RE.findall(text)]
    if "".join("%d%s" % (n, op) for op, n in out) != text:
        raise RuntimeError("malformed SAM cigar: %r" % text)
    return out


def sam_read(path):
    """Returns (header_lines, [SamRecord])."""
    header, records = [], []
    with open(path) as fh:
        for line in fh:
            line = line.rstrip("\n")
            if not line:
                continue
            if line.startswith("@"):
                header.append(line)
                continue
            f = line.split("\t")
            if len(f) < 11:
                raise RuntimeError("malformed SAM line: %r" % line[:80])
            records.append(SamRecord(f[0], int(f[1]), f[2], int(f[3]), int(f[4]), parse_sam_cigar(f[5]), f[6],
                                     int(f[7]), int(f[8]), f[9], f[10], f[11:]))
    return header, records


def sam_write(path, header, records):
    with open(path, "w") as fh:
        for h in header:
            fh.write(h + "\n")
        for r in records:
            fh.write(r.to_line() + "\n")


@dataclass
class GuideAlignment:
    """One exonerate-style cigar: read (contig2 / Y) against reference (contig1 / X), both on '+'."""
    read_name: str
    read_start: int
    read_end: int
    ref_name: str
    ref_start: int
    ref_end: int
    score: float
    ops: List[int]
    lens: List[int]

    def to_line(self):
        letters = "MDI"
        body = " ".join("%s %d" % (letters[o], n) for o, n in zip(self.ops, self.lens))
        head = "cigar: %s %d %d + %s %d %d + %f" % (self.read_name, self.read_start, self.read_end, self.ref_name,
                                                   self.ref_start, self.ref_end, self.score)
        return head + (" " + body if body else "")


def guide_from_sam(rec: SamRecord) -> GuideAlignment:
    """getExonerateCigarFormatString (utils.py:35-55): M/I/D only, coordinates on SEQ and on the reference."""
    ops, lens = [], []
    for op, n in rec.cigar:
        if op not in "MIDSH":
            raise RuntimeError("unsupported SAM cigar op %r in %s" % (op, rec.qname))
        if op == "M":
            ops.append(OP_MATCH)
        elif op == "D":
            ops.append(OP_INDEL_X)
        elif op == "I":
            ops.append(OP_INDEL_Y)
        else:
            continue
        lens.append(n)
    return GuideAlignment(rec.qname, rec.query_alignment_start, rec.query_alignment_end, rec.rname,
                          rec.reference_start, rec.reference_end, 1.0, ops, lens)


def parse_cigar_line(line) -> GuideAlignment:
    """cigarRead: 'cigar: <read> <qs> <qe> <+|-> <ref> <rs> <re> <+|-> <score> (M|I|D <len>)*'."""
    f = line.split()
    if len(f) < 10 or f[0] != "cigar:":
        raise RuntimeError("malformed cigar line: %r" % line[:80])
    if f[4] != "+" or f[8] != "+":
        raise RuntimeError("only forward-strand cigars are produced by the margin call sites")
    code = {"M": OP_MATCH, "D": OP_INDEL_X, "I": OP_INDEL_Y}
    ops = [code[c] for c in f[10::2]]
    lens = [int(n) for n in f[11::2]]
    return GuideAlignment(f[1], int(f[2]), int(f[3]), f[5], int(f[6]), int(f[7]), float(f[9]), ops, lens)
