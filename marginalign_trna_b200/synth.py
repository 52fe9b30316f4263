"""Synthetic workloads and models for tests and bench.py.

Nothing here is on the product path; it only manufactures inputs of the shapes BASELINE.json names:
  * direct-tRNA-like reads (70-150 nt, ~15 % error: 5 % substitution / 5 % insertion / 5 % deletion) against a few
    hundred short references with fixed 5'/3' adapters, guide cigar = the generating alignment;
  * long nanopore-like reads (default 10 kb) against windows of a random genome;
  * `.hmm`-style models: the reference's shipped 5-state asymmetric model values, the documented reduction to a
    3-state model (same reduction as cPecanEm.py:305-308), and seeded random models of all 4 types.
All generators are vectorised numpy so that 1 M reads take seconds.
"""
from dataclasses import dataclass

import numpy as np

ADAPTER_5 = "CCTAAGAGCAAGAAGAAGCCTGG"  # fixed synthetic adapters (not from the reference)
ADAPTER_3 = "GGCTTCTTCTTGCTCTTAGG"
_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)


@dataclass
class AlignmentBatch:
    """Flat, concatenated batch of guide alignments (what the C-ABI consumes).

    seq_x / seq_y are the aligned sub-sequences (reference window X, read Y) as ASCII bytes; cigar ops use sonLib
    codes (0 M, 1 D = reference only, 2 I = read only) and are relative to those sub-sequences.
    """
    seq_x: np.ndarray   # uint8
    off_x: np.ndarray   # int64 [n+1]
    seq_y: np.ndarray
    off_y: np.ndarray
    ops: np.ndarray     # uint8
    lens: np.ndarray    # int32
    off_ops: np.ndarray  # int64 [n+1]

    @property
    def n(self):
        return len(self.off_x) - 1

    def x(self, i):
        return self.seq_x[self.off_x[i]:self.off_x[i + 1]].tobytes()

    def y(self, i):
        return self.seq_y[self.off_y[i]:self.off_y[i + 1]].tobytes()

    def cigar(self, i):
        s = slice(self.off_ops[i], self.off_ops[i + 1])
        return self.ops[s], self.lens[s]

    def subset(self, idx):
        idx = np.asarray(idx, dtype=np.int64)

        def take(data, off):
            lens = off[idx + 1] - off[idx]
            new_off = np.zeros(len(idx) + 1, dtype=np.int64)
            np.cumsum(lens, out=new_off[1:])
            pos = np.repeat(off[idx] - new_off[:-1], lens) + np.arange(new_off[-1])
            return data[pos], new_off

        sx, ox = take(self.seq_x, self.off_x)
        sy, oy = take(self.seq_y, self.off_y)
        ops, oo = take(self.ops, self.off_ops)
        ln, _ = take(self.lens, self.off_ops)
        return AlignmentBatch(sx, ox, sy, oy, ops, ln, oo)

    def slice(self, a, b):
        """Alignments a..b-1 as a batch of views (no copy of the sequence data, offsets rebased)."""
        a, b = int(a), int(b)
        if a == 0 and b == self.n:
            return self
        ox, oy, oo = self.off_x[a:b + 1], self.off_y[a:b + 1], self.off_ops[a:b + 1]
        return AlignmentBatch(self.seq_x[ox[0]:ox[-1]], ox - ox[0], self.seq_y[oy[0]:oy[-1]], oy - oy[0],
                              self.ops[oo[0]:oo[-1]], self.lens[oo[0]:oo[-1]], oo - oo[0])

    @staticmethod
    def from_lists(xs, ys, cigars):
        """xs, ys: bytes/str per alignment; cigars: (ops, lens) per alignment."""
        def cat(seqs):
            bs = [s.encode() if isinstance(s, str) else bytes(s) for s in seqs]
            off = np.zeros(len(bs) + 1, dtype=np.int64)
            np.cumsum([len(b) for b in bs], out=off[1:])
            return np.frombuffer(b"".join(bs), dtype=np.uint8).copy(), off
        sx, ox = cat(xs)
        sy, oy = cat(ys)
        oo = np.zeros(len(cigars) + 1, dtype=np.int64)
        np.cumsum([len(c[0]) for c in cigars], out=oo[1:])
        ops = np.concatenate([np.asarray(c[0], dtype=np.uint8) for c in cigars]) if cigars else np.zeros(0, np.uint8)
        lens = np.concatenate([np.asarray(c[1], dtype=np.int32) for c in cigars]) if cigars else np.zeros(0, np.int32)
        return AlignmentBatch(sx, ox, sy, oy, ops.astype(np.uint8), lens.astype(np.int32), oo)


def _mutate(rng, ref_codes, ref_len, p_sub, p_ins, p_del):
    """Vectorised read generator.

    ref_codes: int8 [n, L] (0..3, padded), ref_len: [n].  Every reference column emits M (maybe substituted) or D,
    optionally followed by a 1-base insertion.  Returns the AlignmentBatch pieces for Y and the cigar.
    """
    n, L = ref_codes.shape
    col = np.arange(L)[None, :]
    valid = col < ref_len[:, None]
    u = rng.random((n, L))
    is_del = valid & (u < p_del)
    is_match = valid & ~is_del
    sub = is_match & (rng.random((n, L)) < p_sub / max(1e-12, 1.0 - p_del))
    ins = valid & (rng.random((n, L)) < p_ins)
    # keep both ends aligned (a guide alignment never starts or ends in an indel after clipping)
    last = (ref_len - 1)[:, None]
    edge = (col == 0) | (col == last)
    is_del &= ~edge
    is_match = valid & ~is_del
    ins &= ~(col == last)
    read_base = np.where(sub, (ref_codes + rng.integers(1, 4, size=(n, L))) % 4, ref_codes)
    ins_base = rng.integers(0, 4, size=(n, L))
    # elementary op stream: slot 2c = M/D of column c, slot 2c+1 = I after column c
    ops = np.full((n, 2 * L), 255, dtype=np.uint8)
    ops[:, 0::2] = np.where(is_match, 0, np.where(is_del, 1, 255))
    ops[:, 1::2] = np.where(ins, 2, 255)
    emit = np.full((n, 2 * L), 255, dtype=np.uint8)
    emit[:, 0::2] = np.where(is_match, read_base, 255)
    emit[:, 1::2] = np.where(ins, ins_base, 255)
    # read sequences
    ymask = emit != 255
    y_len = ymask.sum(axis=1)
    seq_y = _ACGT[emit[ymask]]
    off_y = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(y_len, out=off_y[1:])
    # run-length encode the op stream per row
    omask = ops != 255
    flat_ops = ops[omask]
    row = np.repeat(np.arange(n), omask.sum(axis=1))
    change = np.ones(len(flat_ops), dtype=bool)
    change[1:] = (flat_ops[1:] != flat_ops[:-1]) | (row[1:] != row[:-1])
    starts = np.flatnonzero(change)
    run_ops = flat_ops[starts]
    run_lens = np.diff(np.append(starts, len(flat_ops))).astype(np.int32)
    run_row = row[starts]
    off_ops = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(run_row, minlength=n), out=off_ops[1:])
    return seq_y, off_y, run_ops, run_lens, off_ops


def make_trna_batch(n_reads, n_refs=400, seed=1234, p_sub=0.05, p_ins=0.05, p_del=0.05, body=(70, 95)):
    """Config 3: tRNA-length references (random body + fixed adapters, <= 150 nt) and reads drawn from them."""
    rng = np.random.default_rng(seed)
    a5 = np.frombuffer(ADAPTER_5.encode(), dtype=np.uint8)
    a3 = np.frombuffer(ADAPTER_3.encode(), dtype=np.uint8)
    lut = np.zeros(256, dtype=np.int8)
    lut[_ACGT] = np.arange(4)
    body_len = rng.integers(body[0], body[1] + 1, size=n_refs)
    ref_len = body_len + len(a5) + len(a3)
    L = int(ref_len.max())
    refs = np.zeros((n_refs, L), dtype=np.int8)
    for r in range(n_refs):  # 400 iterations
        b = rng.integers(0, 4, size=body_len[r]).astype(np.int8)
        refs[r, :ref_len[r]] = np.concatenate([lut[a5], b, lut[a3]])
    pick = rng.integers(0, n_refs, size=n_reads)
    ref_codes = refs[pick]
    rl = ref_len[pick]
    seq_y, off_y, ops, lens, off_ops = _mutate(rng, ref_codes, rl, p_sub, p_ins, p_del)
    col = np.arange(L)[None, :]
    seq_x = _ACGT[ref_codes[col < rl[:, None]]]
    off_x = np.zeros(n_reads + 1, dtype=np.int64)
    np.cumsum(rl, out=off_x[1:])
    return AlignmentBatch(seq_x, off_x, seq_y, off_y, ops, lens, off_ops)


def make_long_batch(n_reads, read_len=10000, genome_len=5_000_000, seed=4321, p_sub=0.05, p_ins=0.05, p_del=0.05):
    """Config 4: windows of a random genome, mutated into long reads; guide cigar = the generating alignment."""
    rng = np.random.default_rng(seed)
    genome = rng.integers(0, 4, size=genome_len).astype(np.int8)
    start = rng.integers(0, genome_len - read_len, size=n_reads)
    ref_codes = genome[start[:, None] + np.arange(read_len)[None, :]]
    rl = np.full(n_reads, read_len, dtype=np.int64)
    seq_y, off_y, ops, lens, off_ops = _mutate(rng, ref_codes, rl, p_sub, p_ins, p_del)
    seq_x = _ACGT[ref_codes.reshape(-1)]
    off_x = np.arange(n_reads + 1, dtype=np.int64) * read_len
    return AlignmentBatch(seq_x, off_x, seq_y, off_y, ops, lens, off_ops)


# ---------------------------------------------------------------------------------------------------- models

# Values of a trained 5-state asymmetric nanopore model in the `.hmm` layout (type, 25 transitions row-major
# from*5+to, then 80 emissions state*16+x*4+y).  Same numbers as tests/golden/input_model.hmm, which
# tests/golden/make_golden.py writes from the reference's fixture so tests do not depend on /root/reference.
def default_hmm_path():
    import os
    return os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden",
                        "input_model.hmm")


def three_state_from_five(T5, E5, floor=1e-3):
    """SURVEY App. A recipe: top-left 3x3 of the 5x5 transitions floored at `floor`, rows renormalised; first 48
    emissions."""
    T5 = np.asarray(T5, dtype=np.float64).reshape(5, 5)
    T3 = np.maximum(T5[:3, :3], floor)
    T3 = T3 / T3.sum(axis=1, keepdims=True)
    return T3.reshape(-1), np.asarray(E5, dtype=np.float64)[:48].copy()


def random_hmm(hmm_type, seed):
    """hmm_randomise-like (stateMachine.c:114-131): uniform random then row-normalised; returns (T, E)."""
    rng = np.random.default_rng(seed)
    S = 5 if hmm_type <= 1 else 3
    T = rng.random((S, S)) + 1e-3
    T /= T.sum(axis=1, keepdims=True)
    E = rng.random((S, 16)) + 1e-3
    E /= E.sum(axis=1, keepdims=True)
    return T.reshape(-1), E.reshape(-1)
