"""ctypes binding of oracle/_ref/libcpecan_ref.so: the UNMODIFIED reference C code (TEST INFRASTRUCTURE ONLY).

Available only where oracle/_ref was built (this container: `make -C oracle ref`; it travels to the GPU box as a
prebuilt, git-ignored artefact).  ``available()`` says whether it can be used.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libcpecan_ref.so")
REALIGN_BIN = os.path.join(_HERE, "_ref", "cPecanRealign")
_REF_SRC = "/root/reference/submodules/cPecan/impl/pairwiseAligner.c"


def build():
    if os.path.exists(_REF_SRC):
        subprocess.check_call(["make", "-C", _HERE, "ref", "-j8"], stdout=subprocess.DEVNULL)


def available():
    return os.path.exists(_LIB_PATH)


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not available():
            build()
        L = C.CDLL(_LIB_PATH)
        i64p, dp, vp = C.POINTER(C.c_int64), C.POINTER(C.c_double), C.c_void_p
        L.ref_model_create.argtypes = [C.c_int, dp, dp]
        L.ref_model_create.restype = vp
        L.ref_model_load.argtypes = [C.c_char_p]
        L.ref_model_load.restype = vp
        L.ref_model_default5.restype = vp
        L.ref_model_destroy.argtypes = [vp]
        L.ref_model_state_number.argtypes = [vp]
        L.ref_model_state_number.restype = C.c_int64
        L.ref_model_type.argtypes = [vp]
        L.ref_model_boundary.argtypes = [vp, C.c_int, dp]
        L.ref_model_transitions.argtypes = [vp, C.c_int, C.c_int, dp, C.c_int64]
        L.ref_model_transitions.restype = C.c_int64
        L.ref_logadd.argtypes = [C.c_double, C.c_double]
        L.ref_logadd.restype = C.c_double
        L.ref_symbol.argtypes = [C.c_char]
        L.ref_band.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, i64p]
        L.ref_band.restype = C.c_int64
        L.ref_split_points.argtypes = [i64p, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_int, C.c_int, i64p,
                                       C.c_int64]
        L.ref_split_points.restype = C.c_int64
        common = [vp, C.c_char_p, C.c_char_p, i64p, C.c_int64, C.c_double, C.c_int64, C.c_int64, C.c_int64, C.c_int64,
                  C.c_int, C.c_int]
        L.ref_aligned_pairs.argtypes = common + [i64p, C.c_int64]
        L.ref_aligned_pairs.restype = C.c_int64
        L.ref_expectations.argtypes = common + [dp, dp, dp]
        L.ref_expectations.restype = None
        L.ref_realign_tail.argtypes = [i64p, C.c_int64, C.c_char_p, C.c_char_p, C.c_double, C.c_double, C.c_int, i64p,
                                       C.c_int64]
        L.ref_realign_tail.restype = C.c_int64
        _lib = L
    return _lib


def _b(s):
    return s if isinstance(s, bytes) else s.encode()


def _anch(anchors):
    a = np.ascontiguousarray(np.asarray(anchors, dtype=np.int64).reshape(-1, 2))
    return a, a.ctypes.data_as(C.POINTER(C.c_int64))


class RefModel:
    def __init__(self, handle):
        self.h = handle
        self.S = lib().ref_model_state_number(handle)
        self.type = lib().ref_model_type(handle)

    @classmethod
    def from_hmm(cls, hmm_type, T, E):
        T = np.ascontiguousarray(T, dtype=np.float64)
        E = np.ascontiguousarray(E, dtype=np.float64)
        return cls(lib().ref_model_create(int(hmm_type), T.ctypes.data_as(C.POINTER(C.c_double)),
                                          E.ctypes.data_as(C.POINTER(C.c_double))))

    @classmethod
    def load(cls, path):
        return cls(lib().ref_model_load(_b(path)))

    @classmethod
    def default5(cls):
        return cls(lib().ref_model_default5())

    def boundary(self, which):
        out = np.zeros(self.S)
        lib().ref_model_boundary(self.h, which, out.ctypes.data_as(C.POINTER(C.c_double)))
        return out

    def transitions(self, cX, cY):
        """rows (group, from, to, eP, tP) of cellCalculate for symbols cX, cY."""
        out = np.zeros((16, 5))
        n = lib().ref_model_transitions(self.h, cX, cY, out.ctypes.data_as(C.POINTER(C.c_double)), 16)
        return out[:n]


def logadd(x, y):
    return lib().ref_logadd(x, y)


def symbol(ch):
    return lib().ref_symbol(_b(ch))


def band(anchors, lX, lY, expansion):
    a, ap = _anch(anchors)
    out = np.zeros((lX + lY + 1, 3), dtype=np.int64)
    lib().ref_band(ap, len(a), lX, lY, expansion, out.ctypes.data_as(C.POINTER(C.c_int64)))
    return out[:, 1].copy(), out[:, 2].copy()


def split_points(anchors, lX, lY, split, ragged_left, ragged_right):
    a, ap = _anch(anchors)
    cap = len(a) + 2
    out = np.zeros((cap, 4), dtype=np.int64)
    n = lib().ref_split_points(ap, len(a), lX, lY, split, int(ragged_left), int(ragged_right),
                               out.ctypes.data_as(C.POINTER(C.c_int64)), cap)
    return out[:n]


def aligned_pairs(model, sX, sY, anchors, threshold=0.01, min_diags=1000, traceback_diagonals=40, expansion=10,
                  split=3000 * 3000, ragged_left=True, ragged_right=True):
    a, ap = _anch(anchors)
    cap = 1 << 16
    while True:
        out = np.zeros((cap, 3), dtype=np.int64)
        n = lib().ref_aligned_pairs(model.h, _b(sX), _b(sY), ap, len(a), threshold, min_diags, traceback_diagonals,
                                    expansion, split, int(ragged_left), int(ragged_right),
                                    out.ctypes.data_as(C.POINTER(C.c_int64)), cap)
        if n <= cap:
            return out[:n]
        cap = n


def expectations(model, sX, sY, anchors, threshold=0.01, min_diags=1000, traceback_diagonals=40, expansion=10,
                 split=3000 * 3000, ragged_left=True, ragged_right=True, T=None, E=None, lik=0.0, pseudo=1e-12):
    S = model.S
    T = np.full(S * S, pseudo) if T is None else np.array(T, dtype=np.float64)
    E = np.full(S * 16, pseudo) if E is None else np.array(E, dtype=np.float64)
    likc = C.c_double(lik)
    a, ap = _anch(anchors)
    lib().ref_expectations(model.h, _b(sX), _b(sY), ap, len(a), threshold, min_diags, traceback_diagonals, expansion,
                           split, int(ragged_left), int(ragged_right), T.ctypes.data_as(C.POINTER(C.c_double)),
                           E.ctypes.data_as(C.POINTER(C.c_double)), C.byref(likc))
    return T, E, likc.value


def realign_tail(pairs, sX, sY, gap_gamma, match_gamma, seed=1):
    p = np.ascontiguousarray(pairs, dtype=np.int64).reshape(-1, 3)
    out = np.zeros((len(p) + 1, 2), dtype=np.int64)
    n = lib().ref_realign_tail(p.ctypes.data_as(C.POINTER(C.c_int64)), len(p), _b(sX), _b(sY), gap_gamma, match_gamma,
                               seed, out.ctypes.data_as(C.POINTER(C.c_int64)), len(out))
    return out[:n]
