"""ctypes binding of ``libb200km.so`` (declared in ``include/b200km.h``).

There is no CPU fallback: if the shared library is missing or cannot be loaded every use of the
package raises ``B200LibraryError`` (build it with ``python -c "import __graft_entry__ as g;
g.build()"`` or ``make -C dask_ml_b200/csrc``).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_double, c_float, c_int, c_int64, c_size_t, c_uint32, c_uint64, c_void_p

__all__ = ["load", "check", "B200LibraryError", "B200KernelError", "LIB_PATH", "PROTOTYPES", "OPT_K1_VARIANT",
           "OPT_K2_VARIANT", "K1_AUTO", "K1_FFMA", "K1_TCGEN05", "K1_TCGEN05_PAIR"]

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200km.so")


class B200LibraryError(RuntimeError):
    """libb200km.so is missing / not loadable -- the CUDA path is the only path."""


class B200KernelError(RuntimeError):
    """An entry point returned a CUDA error (positive code)."""


# option keys / values of b200km_set_option / b200km_get_option (include/b200km.h)
OPT_K1_VARIANT = 1      # B200KM_OPT_K1_VARIANT
OPT_K2_VARIANT = 2      # B200KM_OPT_K2_VARIANT
OPT_FUSED_VARIANT = 3   # B200KM_OPT_FUSED_VARIANT (1 = the sorted-tile fused kernel always)
OPT_K4_VARIANT = 4      # B200KM_OPT_K4_VARIANT (1 = the register-path K4 kernel always)
K1_AUTO, K1_FFMA, K1_TCGEN05, K1_TCGEN05_PAIR = 0, 1, 2, 3
CDIST_SQUARED, CDIST_DIST, CDIST_RBF = 0, 1, 2   # b200km_cdist_pack_f32 modes

_P = c_void_p  # device pointers and the stream are passed as plain addresses

# name -> (restype, argtypes).  Kept in the order of include/b200km.h; tests/test_abi.py checks
# that this table, the header and the exported symbols agree.
PROTOTYPES = {
    "b200km_set_option": (c_int, [c_int, c_int]),
    "b200km_get_option": (c_int, [c_int]),
    "b200km_version": (c_int, []),
    "b200km_error_string": (c_char_p, [c_int]),
    "b200km_pack_bytes": (c_size_t, [c_int, c_int]),
    "b200km_pack_centers": (c_int, [_P, c_int, c_int, _P, _P]),
    "b200km_assign_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, c_int, _P]),
    "b200km_assign_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P]),
    "b200km_assign_lloyd_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, c_int, _P]),
    "b200km_refine_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, _P]),
    "b200km_center_convert_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P]),
    "b200km_assign_centred_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, c_int, _P]),
    "b200km_refine_x64_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, _P]),
    "b200km_rowdist_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P]),
    "b200km_refine_pruned_x64_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, _P, c_int, _P, _P, _P, _P, _P]),
    "b200km_refine_pruned_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, _P]),
    "b200km_refine_pruned_len_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int64, _P, _P, _P, _P]),
    "b200km_refine_pruned_x64_len_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, _P, c_int, _P, c_int64, _P, _P, _P, _P]),
    "b200km_assign_static_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P]),
    "b200km_label_counts": (c_int, [_P, c_int64, c_int, _P, _P]),
    "b200km_acc_len": (c_size_t, [c_int, c_int]),
    "b200km_update_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, c_int, _P, c_int, _P]),
    "b200km_update_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, c_int, _P, c_int, _P]),
    "b200km_update_inertia_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, c_int, _P, _P]),
    "b200km_update_inertia_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, _P, c_int, _P, _P]),
    "b200km_assign_update_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P]),
    "b200km_assign_update_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P]),
    "b200km_lloyd_small_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, _P, _P, c_double, _P]),
    "b200km_lloyd_small_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, _P, _P, _P, c_double, _P]),
    "b200km_reduce_sum_f32": (c_int, [_P, c_int64, c_int, _P, _P]),
    "b200km_reduce_sum_f64": (c_int, [_P, c_int64, c_int, _P, _P]),
    "b200km_finalize": (c_int, [_P, c_int, c_int, _P, _P, _P, _P]),
    "b200km_mindist_update_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, _P, _P]),
    "b200km_mindist_update_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, _P, _P]),
    "b200km_mindist_tc_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, _P, _P]),
    "b200km_mindist_merge_f32": (c_int, [_P, c_int64, _P, c_int, _P, _P]),
    "b200km_mindist_merge_f64": (c_int, [_P, c_int64, _P, c_int, _P, _P]),
    "b200km_mindist_pots_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P]),
    "b200km_mindist_pots_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P]),
    "b200km_bernoulli_pick_f32": (c_int, [_P, c_int64, c_double, c_uint64, c_uint32, c_int64, _P, _P, _P, c_int, _P]),
    "b200km_bernoulli_pick_f64": (c_int, [_P, c_int64, c_double, c_uint64, c_uint32, c_int64, _P, _P, _P, c_int, _P]),
    "b200km_gather_rows_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P]),
    "b200km_gather_rows_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P]),
    "b200km_cdist_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, c_int, _P]),
    "b200km_cdist_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, _P, _P, c_int, _P]),
    "b200km_cdist_pack_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int64, c_int, c_double, _P]),
    "b200km_pairwise_kernel_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, c_double, c_double, c_int, _P]),
    "b200km_pairwise_kernel_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P, c_int, c_double, c_double, c_int, _P]),
    "b200km_row_norms_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P]),
    "b200km_row_norms_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, _P]),
    "b200km_count_nonfinite_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, _P]),
    "b200km_count_nonfinite_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, _P]),
    "b200km_make_blobs_f32": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, c_float, c_uint64, c_int64, _P, _P]),
    "b200km_make_blobs_f64": (c_int, [_P, c_int64, c_int, c_int64, _P, c_int, c_double, c_uint64, c_int64, _P, _P]),
    "b200km_probe_ffma": (c_int, [_P, c_int, c_int, c_int, _P]),
    "b200km_probe_dfma": (c_int, [_P, c_int, c_int, c_int, _P]),
}

_lib = None


def load():
    """Load libb200km.so once and attach the prototypes.  Raises B200LibraryError loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200LibraryError(
            f"{LIB_PATH} not found. dask_ml_b200 has no CPU fallback: build the CUDA library "
            "(`make -C dask_ml_b200/csrc` or `__graft_entry__.build()`)."
        )
    try:
        lib = ctypes.CDLL(LIB_PATH)
    except OSError as e:  # pragma: no cover
        raise B200LibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (restype, argtypes) in PROTOTYPES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise B200LibraryError(f"{LIB_PATH} does not export {name}; rebuild it") from e
        fn.restype = restype
        fn.argtypes = argtypes
    _lib = lib
    return lib


def check(code: int, what: str = "") -> None:
    """Map a return code to the exception the reference API would raise:
    argument errors -> ValueError, CUDA errors -> B200KernelError."""
    if code == 0:
        return
    msg = load().b200km_error_string(code).decode()
    if code < 0:
        raise ValueError(f"{what}: {msg}" if what else msg)
    raise B200KernelError(f"{what}: CUDA error {code}: {msg}" if what else f"CUDA error {code}: {msg}")
