"""ctypes binding of oracle/libpairhmm_oracle.so (TEST INFRASTRUCTURE ONLY; see oracle/pairhmm_oracle.h)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libpairhmm_oracle.so")


class PhModel(C.Structure):
    _fields_ = [
        ("type", C.c_int32), ("S", C.c_int32), ("n_trans", C.c_int32 * 3),
        ("frm", (C.c_int32 * 5) * 3), ("to", (C.c_int32 * 5) * 3), ("tp", (C.c_double * 5) * 3),
        ("start", C.c_double * 5), ("ragged_start", C.c_double * 5), ("end", C.c_double * 5),
        ("ragged_end", C.c_double * 5),
        ("em_match", C.c_double * 25), ("em_gap_x", C.c_double * 5), ("em_gap_y", C.c_double * 5),
    ]


class PhParams(C.Structure):
    _fields_ = [
        ("threshold", C.c_double), ("min_diags_between_traceback", C.c_int64), ("traceback_diagonals", C.c_int64),
        ("diagonal_expansion", C.c_int64), ("split_matrix_bigger_than", C.c_int64),
    ]


def build(force=False):
    """Compile the C restatement in place (gcc only; no reference sources involved)."""
    src = os.path.join(_HERE, "pairhmm_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        i64p, dp = C.POINTER(C.c_int64), C.POINTER(C.c_double)
        L.ph_model_from_hmm.argtypes = [C.c_int, dp, dp, C.POINTER(PhModel)]
        L.ph_model_default.argtypes = [C.c_int, C.POINTER(PhModel)]
        L.ph_logadd.argtypes = [C.c_double, C.c_double]
        L.ph_logadd.restype = C.c_double
        L.ph_symbol.argtypes = [C.c_char]
        L.ph_band.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, i64p, i64p]
        L.ph_band.restype = C.c_int64
        L.ph_split_points.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, i64p,
                                      C.c_int64]
        L.ph_split_points.restype = C.c_int64
        L.ph_aligned_pairs.argtypes = [C.POINTER(PhModel), C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, i64p,
                                       C.c_int64, C.POINTER(PhParams), C.c_int, C.c_int, i64p, C.c_int64]
        L.ph_aligned_pairs.restype = C.c_int64
        L.ph_expectations.argtypes = [C.POINTER(PhModel), C.c_char_p, C.c_int64, C.c_char_p, C.c_int64, i64p,
                                      C.c_int64, C.POINTER(PhParams), C.c_int, C.c_int, dp, dp, dp]
        L.ph_expectations.restype = None
        L.ph_count_cells.argtypes = [C.c_int64, C.c_int64, i64p, C.c_int64, C.POINTER(PhParams), C.c_int, C.c_int]
        L.ph_count_cells.restype = C.c_int64
        L.ph_anchors_from_cigar.argtypes = [C.POINTER(C.c_int32), i64p, C.c_int64, C.c_char_p, C.c_char_p, i64p,
                                            C.c_int64]
        L.ph_anchors_from_cigar.restype = C.c_int64
        L.ph_realign_chain.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, C.c_double, C.c_double, C.c_int, i64p,
                                       i64p]
        L.ph_realign_chain.restype = C.c_int64
        L.ph_cigar_from_pairs.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, i64p, C.c_int64]
        L.ph_cigar_from_pairs.restype = C.c_int64
        _lib = L
    return _lib


def _i64(a):
    a = np.ascontiguousarray(a, dtype=np.int64)
    return a, a.ctypes.data_as(C.POINTER(C.c_int64))


def _f64(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(C.POINTER(C.c_double))


def _b(s):
    return s if isinstance(s, bytes) else s.encode()


def make_params(threshold=0.01, min_diags=1000, traceback_diagonals=40, expansion=10, split=3000 * 3000):
    return PhParams(threshold, min_diags, traceback_diagonals, expansion, split)


def model_from_hmm(hmm_type, T, E):
    m = PhModel()
    T, tp = _f64(T)
    E, ep = _f64(E)
    if lib().ph_model_from_hmm(int(hmm_type), tp, ep, C.byref(m)) != 0:
        raise ValueError("bad hmm type %r" % (hmm_type,))
    return m


def model_default(hmm_type=0):
    m = PhModel()
    if lib().ph_model_default(int(hmm_type), C.byref(m)) != 0:
        raise ValueError("bad hmm type %r" % (hmm_type,))
    return m


def model_transitions(m):
    """[(group, from, to, tP)] in cellCalculate order."""
    out = []
    for g in range(3):
        for k in range(m.n_trans[g]):
            out.append((g, m.frm[g][k], m.to[g][k], m.tp[g][k]))
    return out


def logadd(x, y):
    return lib().ph_logadd(x, y)


def symbol(ch):
    return lib().ph_symbol(_b(ch))


def band(anchors, lX, lY, expansion):
    a, ap = _i64(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    L = np.zeros(lX + lY + 1, dtype=np.int64)
    R = np.zeros(lX + lY + 1, dtype=np.int64)
    lib().ph_band(ap, len(a), lX, lY, expansion, L.ctypes.data_as(C.POINTER(C.c_int64)),
                  R.ctypes.data_as(C.POINTER(C.c_int64)))
    return L, R


def split_points(anchors, lX, lY, split, ragged_left, ragged_right):
    a, ap = _i64(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    n = lib().ph_split_points(ap, len(a), lX, lY, split, int(ragged_left), int(ragged_right), None, 0)
    out = np.zeros((max(n, 1), 4), dtype=np.int64)
    lib().ph_split_points(ap, len(a), lX, lY, split, int(ragged_left), int(ragged_right),
                          out.ctypes.data_as(C.POINTER(C.c_int64)), n)
    return out[:n]


def aligned_pairs(m, sX, sY, anchors, params, ragged_left=True, ragged_right=True):
    """rows (p_int, x, y) in the reference's list order."""
    sX, sY = _b(sX), _b(sY)
    a, ap = _i64(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    cap = 1 << 16
    while True:
        out = np.zeros((cap, 3), dtype=np.int64)
        n = lib().ph_aligned_pairs(C.byref(m), sX, len(sX), sY, len(sY), ap, len(a), C.byref(params),
                                   int(ragged_left), int(ragged_right), out.ctypes.data_as(C.POINTER(C.c_int64)), cap)
        if n <= cap:
            return out[:n]
        cap = n


def expectations(m, sX, sY, anchors, params, ragged_left=True, ragged_right=True, T=None, E=None, lik=0.0,
                 pseudo=1e-12):
    """Returns (T[S*S], E[S*16], lik) after accumulating one alignment into the given (or fresh) accumulators."""
    S = m.S
    T = np.full(S * S, pseudo) if T is None else np.array(T, dtype=np.float64)
    E = np.full(S * 16, pseudo) if E is None else np.array(E, dtype=np.float64)
    likc = C.c_double(lik)
    sX, sY = _b(sX), _b(sY)
    a, ap = _i64(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    lib().ph_expectations(C.byref(m), sX, len(sX), sY, len(sY), ap, len(a), C.byref(params), int(ragged_left),
                          int(ragged_right), T.ctypes.data_as(C.POINTER(C.c_double)),
                          E.ctypes.data_as(C.POINTER(C.c_double)), C.byref(likc))
    return T, E, likc.value


def count_cells(lX, lY, anchors, params, ragged_left=True, ragged_right=True):
    a, ap = _i64(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    return lib().ph_count_cells(lX, lY, ap, len(a), C.byref(params), int(ragged_left), int(ragged_right))


def anchors_from_cigar(ops, lens, sX, sY):
    ops = np.ascontiguousarray(ops, dtype=np.int32)
    lens = np.ascontiguousarray(lens, dtype=np.int64)
    cap = int(lens.sum()) + 1
    out = np.zeros((cap, 2), dtype=np.int64)
    n = lib().ph_anchors_from_cigar(ops.ctypes.data_as(C.POINTER(C.c_int32)),
                                    lens.ctypes.data_as(C.POINTER(C.c_int64)), len(ops), _b(sX), _b(sY),
                                    out.ctypes.data_as(C.POINTER(C.c_int64)), cap)
    return out[:n]


def realign_chain(pairs, lX, lY, gap_gamma=0.5, match_gamma=0.0, seed=1):
    """AMAP reweight + MEA chain of one alignment.  pairs: rows (p_int, x, y) in list order.  gap_gamma/match_gamma
    are floats in the reference (pairwiseAligner.h:38, cPecanRealign.c:390).  Returns (chain xy rows, weights)."""
    p, pp = _i64(np.asarray(pairs, dtype=np.int64).reshape(-1, 3))
    out = np.zeros((len(p) + 1, 2), dtype=np.int64)
    w = np.zeros(len(p) + 1, dtype=np.int64)
    m = lib().ph_realign_chain(pp, len(p), lX, lY, float(np.float32(gap_gamma)), float(np.float32(match_gamma)), seed,
                               out.ctypes.data_as(C.POINTER(C.c_int64)), w.ctypes.data_as(C.POINTER(C.c_int64)))
    return out[:m], w[:len(p)]


def cigar_from_pairs(xy, lX, lY):
    """Rows (op, length), op 0 M / 1 D (reference only) / 2 I (read only)."""
    a, ap = _i64(np.asarray(xy, dtype=np.int64).reshape(-1, 2))
    cap = 3 * len(a) + 4
    out = np.zeros((cap, 2), dtype=np.int64)
    n = lib().ph_cigar_from_pairs(ap, len(a), lX, lY, out.ctypes.data_as(C.POINTER(C.c_int64)), cap)
    return out[:n]


def realign(m, sX, sY, anchors, params, gap_gamma=0.5, match_gamma=0.0, seed=1):
    """Posterior pairs -> chain -> cigar ops of one alignment (cPecanRealign.c:584-631)."""
    pairs = aligned_pairs(m, sX, sY, anchors, params)
    xy, _ = realign_chain(pairs, len(sX), len(sY), gap_gamma, match_gamma, seed)
    return cigar_from_pairs(xy, len(sX), len(sY)), xy
