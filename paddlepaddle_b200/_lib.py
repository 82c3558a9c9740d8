"""ctypes binding of libpdb200.so (the C-ABI declared in include/pdb200.h).

The library is the product: if it is missing or does not load, importing the ops fails
loudly.  There is no Python, PyTorch or CPU fallback for any compute call.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpdb200.so")

# dtype codes, include/pdb200.h
BOOL, UINT8, INT8, INT16, INT32, INT64, FLOAT16, BFLOAT16, FLOAT32, FLOAT64 = range(10)
OK, INVALID_ARGUMENT, UNSUPPORTED, CUDA_ERROR, WORKSPACE_TOO_SMALL = range(5)
REDUCE_SUM, REDUCE_MEAN, REDUCE_MAX, REDUCE_MIN, REDUCE_PROD = range(5)
LESS_THAN, LESS_EQUAL, GREATER_THAN, GREATER_EQUAL, EQUAL, NOT_EQUAL = range(6)
LOGICAL_AND, LOGICAL_OR, LOGICAL_XOR = range(3)
BITWISE_AND, BITWISE_OR, BITWISE_XOR = range(3)
GRAD_ADD, GRAD_MULTIPLY = range(2)
MAX_RANK = 9

_c_i64p = ctypes.POINTER(ctypes.c_int64)
_vp = ctypes.c_void_p
_int = ctypes.c_int
_sz = ctypes.c_size_t

# name -> (restype, argtypes); every symbol include/pdb200.h declares
SIGNATURES = {
    "pdb200_last_error": (ctypes.c_char_p, []),
    "pdb200_version": (ctypes.c_char_p, []),
    "pdb200_launch_count": (ctypes.c_uint64, []),
    "pdb200_broadcast_shape": (_int, [_c_i64p, _int, _c_i64p, _int, _c_i64p, ctypes.POINTER(_int)]),
    "pdb200_add": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_multiply": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_subtract": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_divide": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_maximum": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_minimum": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_compare": (_int, [_int, _vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_logical": (_int, [_int, _vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_bitwise": (_int, [_int, _vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_bitwise_shift": (_int, [_int, _int, _vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp]),
    "pdb200_logical_not": (_int, [_vp, ctypes.c_int64, _int, _vp, _vp]),
    "pdb200_bitwise_not": (_int, [_vp, ctypes.c_int64, _int, _vp, _vp]),
    "pdb200_scale": (_int, [_vp, ctypes.c_int64, _int, ctypes.c_double, ctypes.c_double, _int, _vp, _vp]),
    "pdb200_cast": (_int, [_vp, ctypes.c_int64, _int, _int, _vp, _vp]),
    "pdb200_where": (_int, [_vp, _vp, _vp, ctypes.c_int64, _int, _vp, _vp]),
    "pdb200_elementwise_grad_workspace_bytes": (_sz, [_int, _c_i64p, _int, _c_i64p, _int, _int, _int, _int]),
    "pdb200_add_grad": (_int, [_c_i64p, _int, _c_i64p, _int, _vp, _int, _vp, _vp, _vp, _sz, _vp]),
    "pdb200_multiply_grad": (_int, [_vp, _c_i64p, _int, _vp, _c_i64p, _int, _vp, _int, _vp, _vp, _vp, _sz, _vp]),
    "pdb200_sum_grad": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_mean_grad": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_max_grad": (_int, [_vp, _c_i64p, _int, _vp, _vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_min_grad": (_int, [_vp, _c_i64p, _int, _vp, _vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_prod_grad": (_int, [_vp, _c_i64p, _int, _vp, _vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_peer_window_bytes": (_sz, [_sz]),
    "pdb200_peer_window_create": (_int, [_sz, ctypes.POINTER(_vp), _vp]),
    "pdb200_peer_window_open": (_int, [_vp, ctypes.POINTER(_vp)]),
    "pdb200_peer_window_close": (_int, [_vp]),
    "pdb200_peer_window_destroy": (_int, [_vp]),
    "pdb200_peer_slot": (_vp, [_vp, _sz, ctypes.c_uint64]),
    "pdb200_peer_error": (_int, [_vp]),
    "pdb200_peer_all_reduce": (_int, [_int, ctypes.POINTER(_vp), _int, _int, _sz, ctypes.c_uint64, ctypes.c_int64, _int, _int, _vp,
                               _vp]),
    "pdb200_peer_all_reduce_emulated": (_int, [_int, ctypes.POINTER(_vp), _int, _sz, ctypes.c_uint64, ctypes.c_int64, _int, _int,
                                        ctypes.POINTER(_vp), _vp]),
    "pdb200_reduce_workspace_bytes": (_sz, [_int, _c_i64p, _int, _c_i64p, _int, _int, _int, _int]),
    "pdb200_sum_raw": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_mean_raw": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_max_raw": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_min_raw": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_prod": (_int, [_vp, _c_i64p, _int, _c_i64p, _int, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_sum_axis0_static": (_int, [_vp, ctypes.c_int64, ctypes.c_int64, _int, _int, _vp, _vp]),
    "pdb200_argmax_workspace_bytes": (_sz, [_c_i64p, _int, ctypes.c_int64, _int, _int]),
    "pdb200_argmax": (_int, [_vp, _c_i64p, _int, ctypes.c_int64, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_argmin": (_int, [_vp, _c_i64p, _int, ctypes.c_int64, _int, _int, _int, _vp, _vp, _sz, _vp]),
    "pdb200_softmax": (_int, [_vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_log_softmax": (_int, [_vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_fused_softmax_mask": (_int, [_vp, _c_i64p, _vp, _c_i64p, _int, _int, _vp, _vp]),
    "pdb200_fused_softmax_mask_upper_triangle": (_int, [_vp, _c_i64p, _int, _vp, _vp]),
    "pdb200_fused_bias_dropout_residual_layer_norm": (_int, [_vp, _vp, _c_i64p, _int, _vp, _vp, _vp, _int, ctypes.c_float, _int, _int,
                                                      ctypes.c_float, _int, _vp, _vp, _vp, _vp, _vp]),
    "pdb200_fused_bias_residual_layernorm": (_int, [_vp, _c_i64p, _int, _vp, _vp, _vp, _vp, ctypes.c_float, ctypes.c_float, _int, _int,
                                             _vp, _vp, _vp, _vp, _vp]),
    "pdb200_cross_entropy_workspace_bytes": (_sz, [_c_i64p, _int, _int, _int]),
    "pdb200_cross_entropy_with_softmax": (_int, [_vp, _c_i64p, _int, _vp, _int, _int, _int, _int, _int, _int, _vp, _vp, _vp, _sz,
                                                 _vp]),
    "pdb200_cross_entropy_with_softmax_grad": (_int, [_vp, _int, _vp, _vp, _c_i64p, _int, _int, _int, _int, _int, _int, _vp, _vp]),
    "pdb200_rms_norm": (_int, [_vp, _c_i64p, _int, _vp, _vp, _vp, _vp, ctypes.c_float, _int, _int, _vp, _vp, _vp, _vp]),
    "pdb200_softmax_grad": (_int, [_vp, _vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_log_softmax_grad": (_int, [_vp, _vp, _c_i64p, _int, _int, _int, _vp, _vp]),
    "pdb200_layer_norm_grad_workspace_bytes": (_sz, [_c_i64p, _int, _int, _int]),
    "pdb200_layer_norm_grad": (_int, [_vp, _c_i64p, _int, _vp, _int, _vp, _vp, _vp, ctypes.c_float, _int, _int, _vp, _vp,
                                       _vp, _vp, _sz, _vp]),
    "pdb200_layer_norm": (_int, [_vp, _c_i64p, _int, _vp, _vp, _int, ctypes.c_float, _int, _int, _vp, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Load libpdb200.so once; raise if it is not built (run `make` or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: the CUDA extension is not built. Run `make` at the repo root "
                "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no fallback path.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is not exported
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def dims_array(dims):
    n = len(dims)
    arr = (ctypes.c_int64 * max(n, 1))(*[int(d) for d in dims])
    return arr, n


class PDB200Error(RuntimeError):
    pass


def check(rc, lib=None):
    """Translate a C-ABI status into the exception Paddle's pybind would raise
    (InvalidArgument -> ValueError, everything else -> RuntimeError)."""
    if rc == OK:
        return
    lib = lib or load()
    msg = lib.pdb200_last_error().decode("utf-8", "replace")
    if rc == INVALID_ARGUMENT:
        raise ValueError("(InvalidArgument) " + msg)
    if rc == UNSUPPORTED:
        raise NotImplementedError("(Unimplemented) " + msg)
    raise PDB200Error(f"(External) pdb200 status {rc}: {msg}")
