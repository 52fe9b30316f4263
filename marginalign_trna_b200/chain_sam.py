"""Chaining of local alignments into one alignment per read: what marginAlign does to the mapper's SAM
before it realigns (src/margin/marginAlignLib.py:9-199, position helpers src/margin/utils.py:12-33), re-stated for
Python 3 on seqio.SamRecord instead of pysam.

  get_first_non_clipped_position   getFirstNonClippedPositionInRead   utils.py:12-27
  chain_fn                         chainFn                            marginAlignLib.py:112-157  (quadratic best chain)
  merge_chained_aligned_segments   mergeChainedAlignedSegments        marginAlignLib.py:9-110
  chain_sam_file                   chainSamFile                       marginAlignLib.py:159-199

Host-only text processing; nothing here touches the GPU.  Pinned to the reference's own functions, executed on
stand-in alignment objects, by tests/golden/make_chain_golden.py -> tests/test_chain_sam.py.
"""
from typing import Dict, List, Sequence

from . import seqio

_COMPLEMENT = str.maketrans("ACGTNacgtn", "TGCANtgcan")
_OP_INT = {"M": 0, "I": 1, "D": 2, "S": 4, "H": 5}
_OP_CHR = {v: k for k, v in _OP_INT.items()}


def reverse_complement(seq: str) -> str:
    """bioio.reverseComplement: characters outside ACGTN are kept as they are."""
    return seq.translate(_COMPLEMENT)[::-1]


def get_first_non_clipped_position(rec: seqio.SamRecord, read_seq: str) -> int:
    """utils.py:12-27.  Coordinate of the first aligned read base relative to the complete read (hard clips included);
    negative on the reverse strand (the 2nd base of the read is -1 there)."""
    offset = rec.cigar[0][1] if rec.cigar and rec.cigar[0][0] == "H" else 0
    if rec.is_reverse:
        offset = -(len(read_seq) - 1 - offset)
    return offset + rec.query_alignment_start


def get_last_non_clipped_position(rec: seqio.SamRecord, read_seq: str) -> int:
    """utils.py:29-33."""
    return get_first_non_clipped_position(rec, read_seq) + rec.query_alignment_end - rec.query_alignment_start - 1


def chain_fn(records: Sequence[seqio.SamRecord], ref_seq: str, read_seq: str, score_fn=None, max_gap=200):
    """marginAlignLib.py:112-157: the highest-scoring chain of alignments on one strand.  Score of an alignment = its
    number of aligned (M) positions; B may follow A if it starts after A ends in both sequences, on the same strand,
    with at most max_gap unaligned positions in between.  Same quadratic scan, same tie rules (stable sort by reference
    start, strict improvement, first maximum)."""
    if score_fn is None:
        score_fn = lambda r, ref, read: sum(n for op, n in r.cigar if op == "M")
    score = [score_fn(r, ref_seq, read_seq) for r in records]
    coord = [(r.reference_start, get_first_non_clipped_position(r, read_seq), r.reference_end - 1,
              get_last_non_clipped_position(r, read_seq)) for r in records]
    order = sorted(range(len(records)), key=lambda k: coord[k][0])
    best = {k: score[k] for k in order}
    pointer = {}
    for a, i in enumerate(order):
        r_start, q_start, _r_end, _q_end = coord[i]
        own = score[i]
        for j in order[:a]:
            _s2, _q2, r_end2, q_end2 = coord[j]
            if (r_start > r_end2 and q_start > q_end2 and records[i].is_reverse == records[j].is_reverse and
                    r_start - r_end2 + q_start - q_end2 <= max_gap and own + best[j] > best[i]):
                best[i] = own + best[j]
                pointer[i] = j
    k = max(order, key=lambda i: best[i])
    chain = [k]
    while k in pointer:
        k = pointer[k]
        chain.append(k)
    chain.reverse()
    return [records[i] for i in chain]


def merge_chained_aligned_segments(chain: Sequence[seqio.SamRecord], ref_seq: str, read_seq: str) -> seqio.SamRecord:
    """marginAlignLib.py:9-110: one alignment for the chain; unaligned reference stretches between its members become
    deletions, unaligned read stretches insertions, the read's unaligned ends soft clips; SEQ is the complete read
    (reverse-complemented on the reverse strand), no hard clips remain."""
    first = chain[0]
    reverse = first.is_reverse
    ops: List[tuple] = []
    p_pos = first.reference_start
    p_qpos = -(len(read_seq) - 1) if reverse else 0
    for k, rec in enumerate(chain):
        if rec.is_reverse != reverse or rec.reference_start < p_pos:
            raise RuntimeError("chain of %s is not co-linear" % first.qname)
        if rec.reference_start > p_pos:
            ops.append(("D", rec.reference_start - p_pos))
            p_pos = rec.reference_start
        q_pos = get_first_non_clipped_position(rec, read_seq)
        if q_pos < p_qpos:
            raise RuntimeError("chain of %s overlaps in the read" % first.qname)
        if q_pos > p_qpos:
            ops.append(("S" if k == 0 else "I", q_pos - p_qpos))
            p_qpos = q_pos
        for op, n in rec.cigar:
            if op not in "MIDSH":
                raise RuntimeError("unsupported SAM cigar op %r in %s" % (op, rec.qname))
            if op in "MID":
                ops.append((op, n))
            if op in "MD":
                p_pos += n
            if op in "MI":
                p_qpos += n
    if p_pos > len(ref_seq):
        raise RuntimeError("chain of %s runs past the reference" % first.qname)
    if reverse:
        if p_qpos > 1:
            raise RuntimeError("chain of %s runs past the read" % first.qname)
        if p_qpos < 1:
            ops.append(("S", -p_qpos + 1))
    else:
        if p_qpos > len(read_seq):
            raise RuntimeError("chain of %s runs past the read" % first.qname)
        if p_qpos < len(read_seq):
            ops.append(("S", len(read_seq) - p_qpos))
    out = seqio.SamRecord(first.qname, 16 if reverse else 0, first.rname, first.reference_start + 1, 0, ops, "*", 0, 0,
                          reverse_complement(read_seq) if reverse else read_seq, "*", [])
    # the reference's closing assertions (marginAlignLib.py:97-108)
    if not (0 <= out.reference_start < len(ref_seq) and out.reference_end <= len(ref_seq) and
            0 <= out.query_alignment_start < len(read_seq) and out.query_alignment_end <= len(read_seq)):
        raise RuntimeError("chained alignment of %s is out of range" % first.qname)
    return out


def chain_records(records: Sequence[seqio.SamRecord], reads: Sequence[tuple], references: Dict[str, str],
                  chain=chain_fn) -> List[seqio.SamRecord]:
    """chainSamFile's body: aligned records bucketed by (read, reference), one chained alignment per bucket in the order
    of the reads file, then sorted by (reference name, start, end)."""
    buckets: Dict[str, Dict[str, list]] = {}
    for r in records:
        if not r.is_unmapped:
            buckets.setdefault(r.qname, {}).setdefault(r.rname, []).append(r)
    out = []
    for name, seq, _qual in reads:
        name = name.split()[0]
        if name in buckets:
            for rname, recs in buckets.pop(name).items():
                ref = references[rname]
                out.append(merge_chained_aligned_segments(chain(recs, ref, seq), ref, seq))
    if buckets:  # marginAlignLib.py:188: every read of the SAM must be in the reads file
        raise RuntimeError("reads missing from the fastq file: %s" % ", ".join(sorted(buckets)[:5]))
    out.sort(key=lambda r: (r.rname, r.reference_start, r.reference_end))
    return out


def chain_sam_file(sam_in, sam_out, read_fastq, reference_fasta, chain=chain_fn):
    header, records = seqio.sam_read(sam_in)
    out = chain_records(records, list(seqio.fastq_read(read_fastq)), seqio.fasta_dict(reference_fasta), chain)
    seqio.sam_write(sam_out, header, out)
    return len(out)
