"""SAM in -> realigned SAM out: the host stages either side of PairHmmEngine.realign.

Re-statement for Python 3 of what marginAlign does between reading the (chained) SAM and writing the realigned one:
  paralleliseSamProcessingTargetFn   src/margin/utils.py:127-180        grouping of the alignments into jobs
  getExonerateCigarFormatString      src/margin/utils.py:35-55          SAM record -> guide cigar on SEQ[qstart:qend]
  realignCigarTargetFn               src/margin/marginAlignLib.py:296-318   one cPecanRealign per read  (-> ONE batched
                                                                             GPU call here, no temp files)
  realignSamFile3TargetFn            src/margin/marginAlignLib.py:320-370   hard / soft clips re-attached, length
                                                                             assertions, SAM rebuilt positionally
  chainSamFile (optional first step)  src/margin/marginAlignLib.py:159-199   chain_sam.py
No pysam: records are parsed by seqio.sam_read.  Unaligned records are dropped, as samIterator does (utils.py:108-114).
"""
from typing import Dict, List, Sequence

import numpy as np

from . import seqio
from .pairhmm import AlignmentParameters
from .synth import AlignmentBatch

# parameters cPecanRealign gets from realignCigarTargetFn (marginAlignLib.py:315-318)
REALIGN_EXPANSION = 10
REALIGN_SPLIT = 3000
SAM_OP = {0: "M", 1: "D", 2: "I"}  # library op codes (0 M, 1 reference only, 2 read only) as SAM letters


def sam_jobs(records: Sequence[seqio.SamRecord], max_alignment_length_per_job=700000) -> List[List[int]]:
    """Job grouping of paralleliseSamProcessingTargetFn (utils.py:157-176): a new job starts whenever the summed SEQ
    length of the current one exceeds maxAlignmentLengthPerJob or the reference name changes.  Returns lists of record
    indices (aligned records only, in file order)."""
    jobs, total, ref_name = [], None, None
    for i, r in enumerate(records):
        if r.is_unmapped:
            continue
        if total is None or total > max_alignment_length_per_job or ref_name != r.rname:
            jobs.append([])
            total = 0
        jobs[-1].append(i)
        total += len(r.seq)
        ref_name = r.rname
    return jobs


def batch_from_records(records: Sequence[seqio.SamRecord], references: Dict[str, str]) -> AlignmentBatch:
    """The sub-sequences and guide cigars cPecanRealign would be given for these records: X = reference
    [reference_start, reference_end), Y = SEQ[query_alignment_start, query_alignment_end) (SEQ is already reverse-
    complemented for flag-16 reads, so every cigar is forward-strand: SURVEY App. A)."""
    xs, ys, cigs = [], [], []
    for r in records:
        g = seqio.guide_from_sam(r)
        if r.rname not in references:
            raise RuntimeError("SAM record %s refers to unknown reference %s" % (r.qname, r.rname))
        ref = references[r.rname]
        if g.ref_end > len(ref) or g.read_end > len(r.seq):
            raise RuntimeError("SAM record %s runs past its sequences" % r.qname)
        xs.append(ref[g.ref_start:g.ref_end])
        ys.append(r.seq[g.read_start:g.read_end])
        cigs.append((g.ops, g.lens))
    return AlignmentBatch.from_lists(xs, ys, cigs)


def rebuild_record(rec: seqio.SamRecord, ops, lens) -> seqio.SamRecord:
    """realignSamFile3TargetFn's loop body (marginAlignLib.py:340-363): hard-clipped prefix, soft clip of the unaligned
    prefix, the realigned operations, soft clip of the unaligned suffix, hard-clipped suffix; then the two length
    checks (read and reference spans unchanged)."""
    old = rec.cigar
    new = []
    if old and old[0][0] == "H":
        new.append(old[0])
    if rec.query_alignment_start > 0:
        new.append(("S", rec.query_alignment_start))
    new += [(SAM_OP[int(o)], int(n)) for o, n in zip(ops, lens)]
    if rec.query_alignment_end < len(rec.seq):
        new.append(("S", len(rec.seq) - rec.query_alignment_end))
    if len(old) > 1 and old[-1][0] == "H":
        new.append(old[-1])
    on_read = lambda c: sum(n for op, n in c if op in "MIS")
    on_ref = lambda c: sum(n for op, n in c if op in "MD")
    if on_read(new) != on_read(old) or on_ref(new) != rec.reference_end - rec.reference_start:
        raise RuntimeError("realigned cigar of %s does not span the original alignment" % rec.qname)
    out = seqio.SamRecord(rec.qname, rec.flag, rec.rname, rec.pos, rec.mapq, new, rec.rnext, rec.pnext, rec.tlen,
                          rec.seq, rec.qual, list(rec.tags))
    return out


def realign_records(engine, model, records, references, gap_gamma=0.5, match_gamma=0.0,
                    max_alignment_length_per_job=700000):
    """Realigns the aligned records of a SAM (one batched GPU call) and returns the rebuilt records in file order."""
    order = [i for job in sam_jobs(records, max_alignment_length_per_job) for i in job]
    todo = [records[i] for i in order]
    if not todo:
        return []
    batch = batch_from_records(todo, references)
    params = AlignmentParameters(diagonal_expansion=REALIGN_EXPANSION, split_matrix_bigger_than=REALIGN_SPLIT)
    res = engine.realign(model, batch, params, gap_gamma=gap_gamma, match_gamma=match_gamma)
    if res.n != len(todo):  # the reference's cigarIterator yields None to fail here (marginAlignLib.py:333-335)
        raise RuntimeError("got %d realigned cigars for %d alignments" % (res.n, len(todo)))
    out = []
    for k, rec in enumerate(todo):
        ops, lens = res.cigar(k)
        out.append(rebuild_record(rec, ops, lens))
    return out


def realign_sam(sam_in, reference_fasta, sam_out, model, engine, gap_gamma=0.5, match_gamma=0.0, read_fastq=None):
    """realignSamFileTargetFn with a given model (marginAlignLib.py:265-296, no EM): sam_in -> sam_out with every
    aligned record's cigar replaced by its pair-HMM realignment.  With `read_fastq` the local alignments are first
    chained into one alignment per read (chainSamFile, chain_sam.py); without it the SAM is taken as it is
    (--noChain)."""
    header, records = seqio.sam_read(sam_in)
    references = seqio.fasta_dict(reference_fasta)
    if read_fastq is not None:
        from . import chain_sam
        records = chain_sam.chain_records(records, list(seqio.fastq_read(read_fastq)), references)
    out = realign_records(engine, model, records, references, gap_gamma, match_gamma)
    seqio.sam_write(sam_out, header, out)
    return len(out)
