"""ctypes binding of libpcc_b200.so (the C-ABI declared in include/pcc_b200.h).

The product path has no CPU fallback: if the library is missing or a call fails, this module
raises.  Build with ``python __graft_entry__.py build`` (nvcc, sm_100a).
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# PCC_B200_LIB: another build of the same library (kernel A/B experiments); default = the in-tree build
LIB_PATH = os.environ.get("PCC_B200_LIB") or os.path.join(_HERE, "libpcc_b200.so")

P = ctypes.c_void_p
I64 = ctypes.c_int64
I32 = ctypes.c_int32
F32 = ctypes.c_float
SZ = ctypes.c_size_t


class PccError(RuntimeError):
    pass


class ConvArgs(ctypes.Structure):
    """struct pcc_conv_args (include/pcc_b200.h)."""
    _fields_ = [
        ("in_", P), ("n_in", I64), ("cin", I32), ("cout", I32),
        ("weight", P), ("bias", P), ("nbr", P), ("nbr_stride", I64), ("tapmask", P),
        ("K", I32), ("act", I32), ("slope", F32), ("film_c", I32),
        ("film", P), ("film_rows", P), ("residual", P), ("res_rows", P),
        ("out", P), ("n_out", I64), ("out_stride", I64), ("mode", I32), ("flags", I32),
        ("weight_t", P), ("order", P), ("n_groups", I32), ("reserved", I32), ("tile_order", P),
    ]


class WgradArgs(ctypes.Structure):
    """struct pcc_wgrad_args (include/pcc_b200.h)."""
    _fields_ = [
        ("in_", P), ("gout", P), ("cin", I32), ("cout", I32), ("nbr", P), ("nbr_stride", I64),
        ("tapmask", P), ("order", P), ("n_out", I64), ("K", I32), ("n_groups", I32),
        ("dweight", P), ("ws", P), ("ws_bytes", SZ), ("mode", I32), ("reserved", I32),
    ]


# name -> (restype, argtypes); every symbol include/pcc_b200.h declares
SIGNATURES = {
    "pcc_version": (I32, []),
    "pcc_last_error": (ctypes.c_char_p, []),
    "pcc_device_info": (I32, [P, P]),
    "pcc_coords_pack": (I32, [P, I64, P, P, P]),
    "pcc_coords_unpack": (I32, [P, I64, P, P]),
    "pcc_sort_ws_bytes": (SZ, [I64]),
    "pcc_sort_keys": (I32, [P, I64, P, P, P, SZ, P]),
    "pcc_hash_capacity": (I64, [I64]),
    "pcc_hash_build": (I32, [P, I64, P, P, I64, P]),
    "pcc_hash_probe": (I32, [P, P, I64, P, I64, P, P]),
    "pcc_stride_ws_bytes": (SZ, [I64]),
    "pcc_coords_stride": (I32, [P, I64, I32, P, P, P, SZ, P]),
    "pcc_expand_ws_bytes": (SZ, [I64, I32]),
    "pcc_coords_expand": (I32, [P, P, I64, P, I64, I32, I32, P, P, P, SZ, P]),
    "pcc_kernel_map_ws_bytes": (SZ, [I64]),
    "pcc_kernel_map": (I32, [P, P, I64, P, I64, I32, I32, I32, P, I64, P, P, P, I64, I32, P, P, SZ, P]),
    "pcc_conv_fwd": (I32, [P, P]),
    "pcc_weight_transpose": (I32, [P, I32, I32, I32, P, P]),
    "pcc_conv_head0_ws_bytes": (SZ, [I64]),
    "pcc_conv_head0": (I32, [P, I64, I32, P, I32, P, P, I64, P, I64, I32, P, P, SZ, P]),
    "pcc_gather_rows": (I32, [P, I32, P, I64, P, P, P]),
    "pcc_topk_ws_bytes": (SZ, [I64]),
    "pcc_topk_select": (I32, [P, I64, I64, I64, P, P, SZ, P]),
    "pcc_take_keys": (I32, [P, P, I64, P, P]),
    "pcc_gc_quantize_index": (I32, [P, P, P, I64, I32, P, I32, F32, P, P, P]),
    "pcc_gc_build_indexes": (I32, [P, P, I64, I32, P, I32, F32, P, P]),
    "pcc_gc_dequantize": (I32, [P, P, P, I64, I32, P, P]),
    "pcc_gc_likelihood": (I32, [P, P, P, I64, I32, F32, F32, P, P, P]),
    "pcc_eb_quantize": (I32, [P, P, I64, I32, P, P, P]),
    "pcc_eb_dequantize": (I32, [P, P, I64, I32, P, P]),
    "pcc_eb_likelihood": (I32, [P, P, P, I64, I32, F32, P, P, P]),
    "pcc_color_finalize": (I32, [P, P, I64, I32, P, P]),
    "pcc_conv_wgrad_ws_bytes": (SZ, [I32, I32, I32]),
    "pcc_conv_wgrad": (I32, [P, P]),
    "pcc_scatter_rows_add": (I32, [P, I32, P, I64, P, P]),
    "pcc_col_sum_ws_bytes": (SZ, [I32]),
    "pcc_col_sum": (I32, [P, I64, I32, P, P, SZ, P]),
    "pcc_col_sum_static": (I32, [P, I64, I32, P, P]),
    "pcc_gc_likelihood_train_fwd": (I32, [P, P, P, I64, I32, F32, F32, P, P]),
    "pcc_gc_likelihood_train_bwd": (I32, [P, P, P, I64, I32, F32, F32, P, P, P, P]),
    "pcc_octree_ws_bytes": (SZ, [I64]),
    "pcc_octree_encode": (I32, [P, I64, I32, I32, P, I64, P, P, SZ, P]),
    "pcc_octree_decode": (I32, [P, I64, I32, I32, P, I64, P, P, SZ, P]),
    "pcc_rans_encode_with_indexes": (I64, [P, P, I64, P, I32, P, P, I32, P, I64]),
    "pcc_rans_decode_with_indexes": (I32, [P, I64, P, I64, P, I32, P, P, I32, P]),
    "pcc_pmf_to_quantized_cdf": (I32, [P, I32, I32, P]),
    "pcc_gaussian_cdf_tables": (I32, [P, I32, ctypes.c_double, P, P, P, P]),
    "pcc_factorized_cdf_tables": (I32, [P, P, P, P, I32, P, P, P, P]),
}

_lib = None


def load():
    """Load the shared library and bind every declared symbol.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PccError(
            f"{LIB_PATH} not found: the CUDA extension is not built "
            "(run `python __graft_entry__.py build`); there is no CPU fallback")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export it
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what=""):
    """Raise PccError for a negative status."""
    if rc is not None and rc < 0:
        msg = load().pcc_last_error().decode("utf-8", "replace")
        raise PccError(f"{what} failed ({rc}): {msg}")
    return rc
