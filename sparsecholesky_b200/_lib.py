"""ctypes loader for libspchol.so (the C-ABI declared in include/spchol.h)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libspchol.so")

c_int_p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
c_i64_p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
c_dbl_p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")

# every symbol include/spchol.h declares: (restype, argtypes)
_VP = C.c_void_p
SIGNATURES = {
    "spchol_last_error": (C.c_char_p, []),
    "spchol_version": (C.c_char_p, []),
    "spchol_device_count": (C.c_int, []),
    "spchol_triplet_to_csc": (C.c_int, [C.c_int64, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, C.POINTER(C.c_int64)]),
    "spchol_triplet_to_csc_device": (C.c_int, [C.c_int64, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, C.POINTER(C.c_int64), _VP]),
    "spchol_triplet_to_csc_device_noalloc": (C.c_int, [C.c_int64, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, C.POINTER(C.c_int64), _VP]),
    "spchol_triplet_to_csc_gpu": (C.c_int, [C.c_int64, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, C.POINTER(C.c_int64), C.POINTER(C.c_double)]),
    "spchol_mtx_parse_device": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP, _VP, _VP, C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_int64), _VP]),
    "spchol_mtx_read_gpu": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int64), _VP, _VP, _VP, C.POINTER(C.c_double)]),
    "spchol_mtx_read": (C.c_int, [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int64), _VP, _VP, _VP]),
    "spchol_nd_perm_grid": (C.c_int, [C.c_int, C.c_int, C.c_int, _VP]),
    "spchol_order_amd": (C.c_int, [C.c_int, _VP, _VP, _VP]),
    "spchol_permute_upper": (C.c_int, [C.c_int, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "spchol_etree": (C.c_int, [C.c_int, _VP, _VP, _VP]),
    "spchol_post_order": (C.c_int, [C.c_int, _VP, _VP]),
    "spchol_col_count": (C.c_int, [C.c_int, _VP, _VP, _VP, _VP, _VP]),
    "spchol_ereach": (C.c_int, [C.c_int, _VP, _VP, _VP, C.c_int, _VP, _VP, _VP, _VP, C.c_int]),
    "spchol_compute_levels": (C.c_int, [C.c_int, _VP, _VP, _VP]),
    "spchol_compute_supernodes": (C.c_int, [C.c_int, _VP, _VP, _VP, _VP]),
    "spchol_atree": (C.c_int, [C.c_int, _VP, _VP, C.c_int, _VP, _VP]),
    "spchol_symbolic_create": (C.c_int, [C.c_int, _VP, _VP, C.POINTER(_VP)]),
    "spchol_symbolic_create_ex": (C.c_int, [C.c_int, _VP, _VP, C.c_int, C.POINTER(_VP)]),
    "spchol_symbolic_destroy": (None, [_VP]),
    "spchol_symbolic_n": (C.c_int, [_VP]),
    "spchol_symbolic_nnzL": (C.c_int64, [_VP]),
    "spchol_symbolic_nnzA": (C.c_int64, [_VP]),
    "spchol_symbolic_gather_csc64": (C.c_int, [_VP, _VP, _VP, _VP, _VP]),
    "spchol_factor_resident": (C.c_int, [_VP, _VP, C.POINTER(C.c_int)]),
    "spchol_symbolic_flops": (C.c_double, [_VP]),
    "spchol_symbolic_parent": (C.c_int, [_VP, _VP]),
    "spchol_symbolic_colcount": (C.c_int, [_VP, _VP]),
    "spchol_symbolic_Lp64": (C.c_int, [_VP, _VP]),
    "spchol_symbolic_Lp": (C.c_int, [_VP, _VP]),
    "spchol_symbolic_Li": (C.c_int, [_VP, _VP]),
    "spchol_symbolic_supernodes": (C.c_int, [_VP, C.c_int, C.POINTER(C.c_int), _VP, _VP]),
    "spchol_symbolic_stats": (C.c_int, [_VP, _VP]),
    "spchol_plan_create": (C.c_int, [_VP, C.c_int, C.POINTER(_VP)]),
    "spchol_plan_destroy": (None, [_VP]),
    "spchol_plan_stats": (C.c_int, [_VP, _VP]),
    "spchol_plan_dryrun": (C.c_int, [_VP, _VP, _VP]),
    "spchol_plan_dryrun_dist": (C.c_int, [_VP, C.c_int, C.c_int, _VP, _VP, _VP, _VP]),
    "spchol_comm_unique_id": (C.c_int, [_VP]),
    "spchol_comm_create": (C.c_int, [C.c_int, C.c_int, _VP, C.c_int, C.POINTER(_VP)]),
    "spchol_comm_destroy": (None, [_VP]),
    "spchol_comm_rank": (C.c_int, [_VP]),
    "spchol_comm_world": (C.c_int, [_VP]),
    "spchol_plan_create_dist": (C.c_int, [_VP, _VP, _VP, C.POINTER(_VP)]),
    "spchol_plan_dist_stats": (C.c_int, [_VP, _VP, _VP]),
    "spchol_map_fronts": (C.c_int, [_VP, C.c_int, C.c_int, _VP, _VP, _VP]),
    "spchol_factor": (C.c_int, [_VP, _VP, _VP, C.POINTER(C.c_int)]),
    "spchol_factor_device": (C.c_int, [_VP, _VP]),
    "spchol_factor_batched": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP]),
    "spchol_factor_batched_device": (C.c_int, [_VP, C.c_int, _VP, _VP, _VP]),
    "spchol_export_csc_device": (C.c_int, [_VP, _VP]),
    "spchol_sync": (C.c_int, [_VP, C.POINTER(C.c_int)]),
    "spchol_factor_lbuf": (C.c_int, [_VP, _VP, _VP, C.POINTER(C.c_int)]),
    "spchol_host_alloc": (C.c_int, [C.c_int64, C.POINTER(_VP)]),
    "spchol_host_free": (C.c_int, [_VP]),
    "spchol_fp64_peak": (C.c_int, [C.c_int, C.POINTER(C.c_double)]),
    "spchol_lbuf": (C.c_int, [_VP, C.POINTER(_VP), C.POINTER(C.c_int64)]),
    "spchol_lbuf_to_host": (C.c_int, [_VP, _VP]),
    "spchol_last_timing": (C.c_int, [_VP, _VP]),
    "spchol_set_profile": (C.c_int, [_VP, C.c_int]),
    "spchol_get_profile": (C.c_int, [_VP, _VP, _VP, _VP]),
    "spchol_get_gemm_classes": (C.c_int, [_VP, _VP]),
    "spchol_plan_gemm_bytes": (C.c_int, [_VP, C.POINTER(C.c_double), C.POINTER(C.c_int64)]),
    "spchol_stream": (_VP, [_VP]),
    "spchol_last_launches": (C.c_int64, [_VP]),
    "spchol_symbolic_panel_layout": (C.c_int, [_VP, _VP, _VP, _VP, _VP, C.POINTER(C.c_int64)]),
    "spchol_symbolic_front_rows": (C.c_int, [_VP, C.c_int, _VP]),
    "spchol_residual": (C.c_int, [_VP, _VP, C.POINTER(C.c_double)]),
    "spchol_solve": (C.c_int, [_VP, _VP, _VP, C.c_int]),
}

_lib = None


def lib():
    """Load libspchol.so; fails loudly if the CUDA extension has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -m sparsecholesky_b200.build` "
                "(there is no CPU fallback for the numeric path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def ptr(a):
    """raw pointer of a C-contiguous numpy array (or None)"""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


class SpcholError(RuntimeError):
    pass


def check(rc, what="spchol"):
    if rc < 0:
        msg = lib().spchol_last_error().decode()
        raise SpcholError(f"{what} failed ({rc}): {msg}")
    return rc
