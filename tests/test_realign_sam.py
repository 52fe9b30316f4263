"""SAM in -> realigned SAM out.  Golden: the reference's tests/inputBigMutations.sam realigned by the
UNMODIFIED cPecanRealign CLI, one process per read, and spliced as realignSamFile3TargetFn does
(tests/golden/make_realign_sam_golden.py).  CPU: the host logic (job grouping, guide cigars, clip re-attachment, SAM
text) around a stub engine that answers with the CPU oracle.  GPU: the same function around PairHmmEngine.realign;
the SAM text must be identical, byte for byte."""
import json
import os

import numpy as np
import pytest

import marginalign_trna_b200 as mab
from conftest import GOLDEN
from marginalign_trna_b200 import realign_sam as rs
from marginalign_trna_b200 import seqio
from oracle import pyoracle as po


@pytest.fixture(scope="module")
def case(tmp_path_factory):
    with open(os.path.join(GOLDEN, "realign_sam_golden.json")) as fh:
        c = json.load(fh)
    d = tmp_path_factory.mktemp("sam")
    c["sam_in_path"], c["fa_path"] = str(d / "in.sam"), str(d / "ref.fa")
    with open(c["sam_in_path"], "w") as fh:
        fh.write("\n".join(c["sam_in"]) + "\n")
    with open(c["fa_path"], "w") as fh:
        for name, seq in c["references"].items():
            seqio.fasta_write(fh, name, seq)
    return c


class OracleEngine:
    """Stands in for PairHmmEngine.realign on a machine without a GPU: test infrastructure, one alignment at a time
    through the oracle's restatement of cPecanRealign's default mode."""

    def __init__(self, hmm):
        self.om = po.model_from_hmm(hmm._modelTypeInt(), hmm.transitions, hmm.emissions)

    def realign(self, model, batch, params, gap_gamma=0.5, match_gamma=0.0):
        oP = po.make_params(expansion=params.diagonal_expansion, split=params.split_matrix_bigger_than ** 2)
        off, ops, lens = [0], [], []
        for i in range(batch.n):
            anc = po.anchors_from_cigar(*batch.cigar(i), batch.x(i), batch.y(i))
            rows, _chain = po.realign(self.om, batch.x(i), batch.y(i), anc, oP, gap_gamma, match_gamma)
            ops += rows[:, 0].tolist()
            lens += rows[:, 1].tolist()
            off.append(len(ops))
        return mab.RealignedCigars(np.array(off), np.array(ops, dtype=np.uint8), np.array(lens, dtype=np.int32))


def test_job_grouping_follows_the_reference():
    """paralleliseSamProcessingTargetFn (utils.py:157-176): a new job when the reference changes or after the running
    SEQ length has exceeded the limit; unaligned records are skipped (samIterator)."""
    R = lambda name, ref, n, flag=0: seqio.SamRecord(name, flag, ref, 1, 0, [("M", n)], "*", 0, 0, "A" * n, "*")
    recs = [R("a", "r1", 40), R("b", "r1", 70), R("u", "*", 5, 4), R("c", "r1", 10), R("d", "r2", 10), R("e", "r1", 10)]
    assert rs.sam_jobs(recs, max_alignment_length_per_job=100) == [[0, 1], [3], [4], [5]]
    assert rs.sam_jobs(recs) == [[0, 1, 3], [4], [5]]
    assert rs.sam_jobs([]) == []


def test_clip_reattachment_and_length_checks():
    rec = seqio.SamRecord("q", 16, "r", 5, 9, seqio.parse_sam_cigar("3H2S4M1I2M1D3M4S5H"), "*", 0, 0, "A" * 16, "I" * 16,
                          ["NM:i:1"])
    out = rs.rebuild_record(rec, [0, 2, 0, 1, 0], [3, 2, 3, 2, 2])  # 3M 2I 3M 2D 2M: read 10, reference 10
    assert out.cigar_string() == "3H2S3M2I3M2D2M4S5H" and out.tags == ["NM:i:1"] and out.flag == 16
    with pytest.raises(RuntimeError, match="does not span"):
        rs.rebuild_record(rec, [0], [9])


def test_realigned_sam_matches_the_reference_cli(case, tmp_path):
    hmm = mab.Hmm.loadHmm(os.path.join(GOLDEN, "input_model.hmm"))
    out = str(tmp_path / "out.sam")
    n = rs.realign_sam(case["sam_in_path"], case["fa_path"], out, None, OracleEngine(hmm))
    assert n == 30 and case["n_changed"] == 30
    assert open(out).read().split("\n")[:-1] == case["sam_out"]


@pytest.mark.gpu
def test_gpu_realigned_sam_is_bit_identical(case, tmp_path):
    model = mab.PairHmmModel.from_file(os.path.join(GOLDEN, "input_model.hmm"))
    out = str(tmp_path / "out.sam")
    assert rs.realign_sam(case["sam_in_path"], case["fa_path"], out, model, mab.PairHmmEngine()) == 30
    assert open(out).read().split("\n")[:-1] == case["sam_out"]
