"""marginCaller's use of the posteriors, on the GPU.

Mirrors src/margin/marginCallerLib.py: posteriorProbabilityCalculationTargetFn (:22-74, the marginalised branch:
--diagonalExpansion=10 --splitMatrixBiggerThanThis=100 --outputAllPosteriorProbs) and variantCallSamFileTargetFn
(:171-217).  The posterior pairs stay in HBM; `ColumnCaller.add` accumulates them into one [column][A,C,G,T] table
(mab_columns_accumulate), `posteriors` evaluates calcBasePosteriorProbs per column (mab_columns_call), `calls` applies
the threshold and `write_vcf` prints the reference's minimal VCF (vcfWrite, :112-166).  The noMargin branch
(--rescoreOriginalAlignment, weight 1.0 per aligned pair) involves no pair-HMM and stays out of scope.
"""
import ctypes as C
import datetime

import numpy as np

from . import capi
from .hmm import Hmm
from .pairhmm import AlignmentParameters, require_cuda

BASES = "ACGT"


def load_hmm_substitution_matrix(hmm_or_path):
    """loadHmmSubstitutionMatrix (marginCallerLib.py:94-100): the match emissions, each row normalised.
    Returns a 4x4 array [true base][observed base]."""
    hmm = hmm_or_path if isinstance(hmm_or_path, Hmm) else Hmm.loadHmm(hmm_or_path)
    m = np.asarray(hmm.emissions[:16], dtype=np.float64).reshape(4, 4)
    return m / m.sum(axis=1, keepdims=True)


def null_substitution_matrix():
    """getNullSubstitutionMatrix (marginCallerLib.py:102-105)."""
    return np.ones((4, 4))


class ColumnCaller:
    """Column table over a set of reference sequences: columns of all references are concatenated in the order of
    `references` (a dict name -> sequence, like getFastaDictionary)."""

    def __init__(self, engine, references):
        torch = require_cuda()
        self.torch, self.engine = torch, engine
        self.names = list(references)
        self.seqs = {k: references[k] for k in self.names}
        self.offset, total = {}, 0
        for k in self.names:
            self.offset[k] = total
            total += len(references[k])
        self.n_columns = total
        ref = np.frombuffer("".join(references[k] for k in self.names).encode(), dtype=np.uint8)
        self.d_ref = torch.from_numpy(ref.copy()).to(engine.device)
        self.d_table = torch.zeros(max(total, 1) * 4, dtype=torch.int64, device=engine.device)

    def add(self, model, batch, ref_names, ref_starts, params=None):
        """Posteriors of every guide alignment of `batch` (alignment i lies on reference ref_names[i], its window
        starting at ref_starts[i]) accumulated into the column table."""
        params = params or AlignmentParameters(diagonal_expansion=10, split_matrix_bigger_than=100)
        torch, eng, L = self.torch, self.engine, capi.load()
        base_all = np.array([self.offset[r] + int(s) for r, s in zip(ref_names, ref_starts)], dtype=np.int64)
        first = 0
        with torch.cuda.device(eng.device):
            for dplan in eng._device_plans(batch, params, False):
                n = dplan.plan.info.n_alignments
                sub = batch.slice(first, first + n)
                d_off, d_x, d_y, d_p = eng.posteriors_device(model, dplan)
                d_seq_y = torch.from_numpy(np.ascontiguousarray(sub.seq_y, dtype=np.uint8)).to(eng.device)
                d_off_y = torch.from_numpy(np.ascontiguousarray(sub.off_y, dtype=np.int64)).to(eng.device)
                d_base = torch.from_numpy(base_all[first:first + n].copy()).to(eng.device)
                capi.check(L.mab_columns_accumulate(n, d_off.data_ptr(), d_x.data_ptr(), d_y.data_ptr(), d_p.data_ptr(),
                                                    d_seq_y.data_ptr(), d_off_y.data_ptr(), d_base.data_ptr(),
                                                    self.d_table.data_ptr(), self.n_columns, eng._stream_ptr()))
                torch.cuda.current_stream(eng.device).synchronize()
                first += n
                del dplan

    def expectations(self):
        """[n_columns][4] expected base counts (sums of the 6-decimal probabilities the reference re-reads)."""
        return self.d_table.cpu().numpy().reshape(-1, 4)[:self.n_columns] * 1e-6

    def posteriors(self, error_matrix, evolutionary_matrix=None):
        """calcBasePosteriorProbs for every column: [n_columns][4]; zeros where nothing was observed."""
        torch, eng = self.torch, self.engine
        evo = null_substitution_matrix() if evolutionary_matrix is None else np.asarray(evolutionary_matrix)
        with np.errstate(divide="ignore"):
            d_err = torch.from_numpy(np.log(np.asarray(error_matrix, dtype=np.float64)).reshape(-1).copy()).to(eng.device)
            d_evo = torch.from_numpy(np.log(evo.astype(np.float64)).reshape(-1).copy()).to(eng.device)
        d_post = torch.empty(max(self.n_columns, 1) * 4, dtype=torch.float64, device=eng.device)
        with torch.cuda.device(eng.device):
            capi.check(capi.load().mab_columns_call(self.d_table.data_ptr(), self.n_columns, self.d_ref.data_ptr(),
                                                    d_err.data_ptr(), d_evo.data_ptr(), d_post.data_ptr(),
                                                    eng._stream_ptr()))
        return d_post.cpu().numpy().reshape(-1, 4)[:self.n_columns]

    def calls(self, error_matrix, threshold=0.3, evolutionary_matrix=None):
        """Variant calls (reference name, 0-based position, alt base, posterior), as variantCallSamFileTargetFn
        collects them (:205-209): non-reference bases whose posterior reaches the threshold."""
        return calls_from_posteriors(self.names, self.seqs, self.offset, self.posteriors(error_matrix,
                                                                                         evolutionary_matrix), threshold)


def calls_from_posteriors(names, seqs, offset, post, threshold):
    out = []
    for name in names:
        seq = seqs[name].upper()
        block = post[offset[name]:offset[name] + len(seq)]
        hit_c, hit_b = np.nonzero(block >= threshold)
        for c, b in zip(hit_c.tolist(), hit_b.tolist()):
            if BASES[b] != seq[c] and block[c].sum() > 0.0:
                out.append((name, c, BASES[b], float(block[c, b])))
    return out


def write_vcf(path, reference_fasta_name, references, variant_calls, today=None):
    """vcfWrite (marginCallerLib.py:112-166): one line per reference position that has calls."""
    by_pos = {}
    for name, pos, base, prob in variant_calls:
        by_pos.setdefault((name, pos), []).append((base, prob))
    today = today or datetime.datetime.now().date()
    with open(path, "w") as fh:
        fh.write("##fileformat=VCFv4.2\n")
        fh.write("##fileDate=" + str(today).replace("-", "") + "\n")
        fh.write("##source=marginCaller\n")
        fh.write("##reference=" + reference_fasta_name + "\n")
        fh.write("#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\n")
        for name in references:
            for pos, ref_base in enumerate(references[name]):
                if (name, pos) in by_pos:
                    alts = by_pos[(name, pos)]
                    fh.write("\t".join([name, str(pos + 1), ".", ref_base, ",".join(a for a, _ in alts), ".", "PASS",
                                        ",".join(repr(p) for _, p in alts)]) + "\n")


def read_vcf_calls(path):
    """vcfRead (marginCallerLib.py:107-112) without PyVCF: {(chrom, 1-based pos, ALT)}."""
    calls = set()
    with open(path) as fh:
        for line in fh:
            if line.startswith("#"):
                continue
            f = line.rstrip("\n").split("\t")
            for alt in f[4].split(","):
                calls.add((f[0], int(f[1]), alt.upper()))
    return calls
