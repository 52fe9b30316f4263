"""ctypes binding of libmargin_b200.so (the C ABI declared in include/margin_b200.h).

The library is built in-tree by build.py; if it is missing or a symbol is absent this module raises -- there is no
Python or CPU fallback for the pair-HMM path.
"""
import ctypes as C
import os

from . import build as _build

i64p = C.POINTER(C.c_int64)
i32p = C.POINTER(C.c_int32)
u8p = C.POINTER(C.c_uint8)
dp = C.POINTER(C.c_double)
vp = C.c_void_p


class MabParams(C.Structure):
    _fields_ = [
        ("threshold", C.c_double),
        ("min_diags_between_traceback", C.c_int64),
        ("traceback_diagonals", C.c_int64),
        ("diagonal_expansion", C.c_int64),
        ("split_matrix_bigger_than", C.c_int64),
        ("ragged_left", C.c_int32),
        ("ragged_right", C.c_int32),
    ]


class MabPlanInfo(C.Structure):
    _fields_ = [(n, C.c_int64) for n in (
        "n_alignments", "n_regions", "n_diagonals", "n_cells", "n_tracebacks", "n_totals", "n_anchors", "max_width",
        "blob_bytes", "workspace_bytes_posteriors", "workspace_bytes_expectations")]


# name -> (restype, argtypes); every function include/margin_b200.h declares
SIGNATURES = {
    "mab_last_error": (C.c_char_p, []),
    "mab_version": (C.c_char_p, []),
    "mab_device_count": (C.c_int, []),
    "mab_model_from_hmm": (C.c_int, [C.c_int, dp, dp, C.POINTER(vp)]),
    "mab_model_default": (C.c_int, [C.c_int, C.POINTER(vp)]),
    "mab_model_destroy": (None, [vp]),
    "mab_model_state_number": (C.c_int, [vp]),
    "mab_model_type": (C.c_int, [vp]),
    "mab_model_transitions": (C.c_int, [vp, dp, C.c_int]),
    "mab_model_boundary": (C.c_int, [vp, C.c_int, dp]),
    "mab_model_emissions": (C.c_int, [vp, dp, dp, dp]),
    "mab_params_default": (None, [C.POINTER(MabParams)]),
    "mab_plan_from_cigars": (C.c_int, [C.POINTER(MabParams), C.c_int64, u8p, i64p, u8p, i64p, u8p, i32p, i64p, C.c_int,
                                       C.POINTER(vp)]),
    "mab_plan_from_anchors": (C.c_int, [C.POINTER(MabParams), C.c_int64, u8p, i64p, u8p, i64p, i64p, i64p, C.c_int,
                                        C.POINTER(vp)]),
    "mab_plan_destroy": (None, [vp]),
    "mab_plan_get_info": (C.c_int, [vp, C.POINTER(MabPlanInfo)]),
    "mab_plan_write_blob": (C.c_int, [vp, vp, C.c_int64]),
    "mab_plan_anchors": (C.c_int64, [vp, C.c_int64, i64p, C.c_int64]),
    "mab_plan_regions": (C.c_int64, [vp, i64p, C.c_int64]),
    "mab_plan_band": (C.c_int64, [vp, C.c_int64, i64p, C.c_int64]),
    "mab_plan_tracebacks": (C.c_int64, [vp, i64p, C.c_int64]),
    "mab_plan_device_arena_hint": (C.c_int64, [C.POINTER(MabParams), C.c_int64, C.c_int64, C.c_int64]),
    "mab_plan_build_device": (C.c_int, [C.POINTER(MabParams), C.c_int64, vp, vp, C.c_int64, vp, vp, C.c_int64, vp, vp,
                                        vp, vp, vp, vp, C.c_int64, vp, C.POINTER(vp), i64p]),
    "mab_plan_device_blob": (vp, [vp]),
    "mab_posteriors_compute": (C.c_int, [vp, vp, vp, vp, C.c_int64, vp, vp, vp]),
    "mab_posteriors_compute_noalloc": (C.c_int, [vp, vp, vp, vp, C.c_int64, vp, vp, vp]),
    "mab_posteriors_emit": (C.c_int, [vp, vp, vp, C.c_int64, vp, vp, vp, C.c_int64, vp]),
    "mab_posteriors_emit_i16": (C.c_int, [vp, vp, vp, C.c_int64, vp, vp, vp, C.c_int64, vp]),
    "mab_expectations": (C.c_int, [vp, vp, vp, vp, C.c_int64, vp, vp]),
    "mab_debug_copy_matrices": (C.c_int, [vp, vp, dp, dp, dp, vp]),
    "mab_realign_workspace_bytes": (C.c_int64, [C.c_int64, C.c_int64, C.c_int64, C.c_int64]),
    "mab_realign_chain": (C.c_int, [C.c_int64, vp, vp, C.c_int64, C.c_int64, vp, vp, vp, vp, C.c_int64, C.c_double,
                                    C.c_double, vp, C.c_int64, vp, C.c_int64, vp, vp, vp]),
    "mab_realign_emit": (C.c_int, [C.c_int64, vp, vp, C.c_int64, C.c_int64, C.c_int64, vp, C.c_int64, vp, vp, vp, vp,
                                   vp, vp, vp]),
    "mab_rand_sequence": (C.c_int, [C.c_uint32, C.c_int64, vp]),
    "mab_columns_accumulate": (C.c_int, [C.c_int64, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int64, vp]),
    "mab_columns_call": (C.c_int, [vp, C.c_int64, vp, vp, vp, vp, vp]),
    "mab_debug_set_warp_group": (C.c_int, [C.c_int]),
    "mab_debug_set_min_band_class": (C.c_int, [C.c_int]),
    "mab_debug_set_fused_posterior": (C.c_int, [C.c_int]),
    "mab_launch_count": (C.c_int64, [C.c_int]),
    "mab_kernel_timing": (C.c_int, [C.c_int]),
    "mab_kernel_count": (C.c_int, []),
    "mab_kernel_name": (C.c_char_p, [C.c_int]),
    "mab_kernel_times": (C.c_int, [dp, i64p, C.c_int]),
}

_lib = None


def lib_path():
    return _build.LIB_PATH


def load(build_if_missing=True):
    """Loads the shared library and binds every declared symbol; raises if anything is missing."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path):
        if not build_if_missing:
            raise RuntimeError("libmargin_b200.so is not built (run __graft_entry__.build())")
        _build.build_library()
    L = C.CDLL(path)
    for name, (restype, argtypes) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the symbol is not exported
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise RuntimeError("libmargin_b200: %s (code %d)" % (load().mab_last_error().decode(), rc))
