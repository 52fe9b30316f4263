"""Chaining of local alignments (chainSamFile, marginAlignLib.py:9-199).  Golden: the reference's own
chainFn / mergeChainedAlignedSegments executed on stand-in alignment objects (tests/golden/make_chain_golden.py) over
pieces cut from the reference's fixture alignments, with decoys that must not be chained."""
import json
import os

import pytest

from conftest import GOLDEN
from marginalign_trna_b200 import chain_sam, seqio


@pytest.fixture(scope="module")
def case(tmp_path_factory):
    with open(os.path.join(GOLDEN, "chain_golden.json")) as fh:
        c = json.load(fh)
    d = tmp_path_factory.mktemp("chain")
    c["sam"], c["fq"], c["fa"] = str(d / "in.sam"), str(d / "reads.fq"), str(d / "ref.fa")
    with open(c["sam"], "w") as fh:
        fh.write("\n".join(c["sam_in"]) + "\n")
    with open(c["fq"], "w") as fh:
        for name, seq, qual in c["reads"]:
            fh.write("@%s\n%s\n+\n%s\n" % (name, seq, qual))
    with open(c["fa"], "w") as fh:
        for name, seq in c["references"].items():
            seqio.fasta_write(fh, name, seq)
    return c


def test_position_helpers():
    """getFirst/LastNonClippedPositionInRead (utils.py:12-33): hard clips count, soft clips are inside SEQ, reverse-strand
    coordinates are negative."""
    read = "A" * 50
    fwd = seqio.SamRecord("q", 0, "r", 1, 0, seqio.parse_sam_cigar("5H3S10M2I8M22S"), "*", 0, 0, "A" * 45, "*")
    assert chain_sam.get_first_non_clipped_position(fwd, read) == 8
    assert chain_sam.get_last_non_clipped_position(fwd, read) == 27
    rev = seqio.SamRecord("q", 16, "r", 1, 0, seqio.parse_sam_cigar("5H3S10M32S"), "*", 0, 0, "A" * 45, "*")
    assert chain_sam.get_first_non_clipped_position(rev, read) == -(50 - 1 - 5) + 3
    assert chain_sam.reverse_complement("ACGTNacgtX") == "XacgtNACGT"


def test_chained_sam_matches_the_reference_functions(case, tmp_path):
    out = str(tmp_path / "chained.sam")
    n = chain_sam.chain_sam_file(case["sam"], out, case["fq"], case["fa"])
    assert n == len(case["chained"])
    _header, got = seqio.sam_read(out)
    assert {w["n_chained"] for w in case["chained"]} >= {1, 3}  # the fixture holds broken and complete chains
    for g, w in zip(got, case["chained"]):
        assert (g.qname, g.is_reverse, g.rname, g.reference_start, g.cigar_string(), g.seq) == \
               (w["qname"], w["reverse"], w["rname"], w["reference_start"], w["cigar"], w["seq"]), w["qname"]
        assert "H" not in g.cigar_string() and g.query_alignment_end <= len(g.seq)


def test_missing_read_is_an_error(case):
    _h, recs = seqio.sam_read(case["sam"])
    with pytest.raises(RuntimeError, match="missing from the fastq"):
        chain_sam.chain_records(recs, case["reads"][:1], case["references"])
