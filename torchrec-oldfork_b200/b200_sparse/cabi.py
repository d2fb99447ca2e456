"""ctypes binding of include/b200_sparse.h (libb200sparse.so).

The library is the product; there is no CPU fallback.  Importing this module fails loudly
when the shared object is missing or does not export a declared symbol.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libb200sparse.so")

# dtype / mode codes of b200_sparse.h
I32, I64, F32, F16, BF16, U8 = range(6)
POOL_SUM, POOL_MEAN = 0, 1
OPT_SGD, OPT_ROWWISE_ADAGRAD = 0, 1
BOUNDS_FATAL, BOUNDS_WARNING, BOUNDS_IGNORE = 0, 1, 2
CUMSUM_COMPLETE, CUMSUM_INCLUSIVE, CUMSUM_EXCLUSIVE = 0, 1, 2

_p, _i32, _i64, _f32, _sz, _int = C.c_void_p, C.c_int32, C.c_int64, C.c_float, C.c_size_t, C.c_int

# name -> (restype, argtypes); mirrors the header one to one
SIGNATURES = {
    "b200_error_string": (C.c_char_p, [_int]),
    "b200_abi_version": (_int, []),
    "b200_scan_workspace_bytes": (_sz, []),
    "b200_cumsum": (_int, [_p, _p, _i64, _int, _int, _p, _p]),
    "b200_permute_lengths": (_int, [_p, _p, _int, _i32, _i32, _i32, _p, _p, _p, _p, _p]),
    "b200_permute_data": (_int, [_p, _i32, _i32, _p, _p, _p, _p, _int, _p, _p, _int, _i64, _p]),
    "b200_permute_2d_fused": (_int, [_p, _p, _int, _i32, _i32, _i32, _p, _p, _p, _int, _p, _p, _int, _i64, _p, _p]),
    "b200_expand_into_jagged_permute": (_int, [_p, _p, _p, _int, _i64, _p, _i64, _p]),
    "b200_offsets_range": (_int, [_p, _int, _i64, _p, _i64, _p]),
    "b200_block_bucketize": (_int, [_p, _int, _p, _int, _p, _i32, _i32, _i32, _i64, _p, _p, _p, _p,
                                    _p, _p, _p, _p, _p, _p]),
    "b200_bounds_check": (_int, [_p, _p, _p, _int, _i32, _i32, _i64, _int, _p, _p, _p]),
    "b200_tbe_forward": (_int, [_p, _int, _p, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _int, _p,
                                _int, _p, _int, _p]),
    "b200_tbe_forward_peer": (_int, [_p, _int, _p, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _int, _p,
                                     _int, _p, _i32, _i32, _i64, _p, _int, _i32, _p, _i64, _p, _p]),
    "b200_tbe_backward_peer": (_int, [_p, _i32, _i32, _i64, _p, _f32, _p, _p, _p, _i64, _int, _p, _int, _p, _p,
                                      _p, _i32, _i32, _i32, _i32, _p, _int, _p, _int, _int, _p, _p, _f32, _f32,
                                      _p, _i32, C.c_uint64, C.c_uint64, _p]),
    "b200_peer_stage_reduce": (_int, [_p, _i32, _i32, _i32, _p, _i64, _p, _i32, _p]),
    "b200_peer_barrier": (_int, [_p, _i32, _i32, _p, _p]),
    "b200_peer_stage_reduce_flagged": (_int, [_p, _p, _i32, _i32, _i32, _i32, _p, _p, _i64, _p, _p]),
    "b200_tbe_sort_workspace_bytes": (_sz, [_i64, _int, _i32]),
    "b200_tbe_sort_indices": (_int, [_p, _p, _int, _p, _i32, _i32, _i64, _int, _int, _int, _p, _p, _p, _p,
                                     _sz, _p]),
    "b200_tbe_backward_scratch_bytes": (_sz, [_i64, _i32]),
    "b200_tbe_backward": (_int, [_p, _i64, _p, _p, _p, _i64, _int, _p, _int, _p, _p, _p, _i32, _i32,
                                 _i32, _i32, _p, _int, _p, _int, _int, _p, _p, _f32, _f32, _p, _p]),
    "b200_tbe_backward_static_max_scratch_bytes": (_sz, []),
    "b200_tbe_backward_static": (_int, [_p, _i64, _p, _p, _p, _i64, _int, _p, _int, _p, _p, _p, _i32, _i32,
                                        _i32, _i32, _p, _int, _p, _int, _int, _p, _p, _f32, _f32, _p]),
    "b200_tbe_backward_sr": (_int, [_p, _i64, _p, _p, _p, _i64, _int, _p, _int, _p, _p, _p, _i32, _i32,
                                    _i32, _i32, _p, _int, _p, _int, _int, _p, _p, _f32, _f32, _p,
                                    C.c_uint64, C.c_uint64, _p]),
    "b200_tbe_grad_indice_weights": (_int, [_p, _i64, _p, _int, _p, _p, _i32, _i32, _p, _p, _int,
                                            _p, _p]),
    "b200_quantize_rowwise_8bit_sbhalf": (_int, [_p, _i64, _i32, _p, _p]),
    "b200_tbe_int8_forward": (_int, [_p, _p, _p, _i32, _i32, _i32, _i32, _i32, _p, _p, _int, _p,
                                     _int, _p, _int, _p]),
    "b200_interaction_forward": (_int, [_p, _p, _i32, _i32, _i32, _p, _i64, _p]),
    "b200_interaction_forward_tma": (_int, [_p, _p, _i32, _i32, _i32, _p, _i64, _p]),
    "b200_interaction_backward": (_int, [_p, _p, _p, _i64, _i32, _i32, _i32, _p, _p, _p]),
    "b200_interaction_backward_scatter": (_int, [_p, _p, _p, _i64, _i32, _i32, _i32, _p, _p, _p, _p]),
    "b200_linearize_seq_indices": (_int, [_p, _p, _int, _p, _i32, _i32, _i64, _p, _p]),
    "b200_relu_bias_grad_scratch_bytes": (_sz, [_i32]),
    "b200_relu_bias_grad": (_int, [_p, _i64, _p, _i64, _i64, _i32, _p, _p, _p, _p]),
    "b200_pooled_a2a_recat": (_int, [_p, _p, _p, _i32, _i32, _int, _f32, _p]),
}


class B200Error(RuntimeError):
    pass


def load(path: str = LIB_PATH) -> C.CDLL:
    if not os.path.exists(path):
        raise ImportError(
            f"{path} not found: the CUDA extension is the only implementation of this package. "
            "Build it with `python torchrec-oldfork_b200/build.py` (needs nvcc, sm_100a)."
        )
    lib = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name, None)
        if fn is None:
            raise ImportError(f"{path} does not export {name}; rebuild the extension")
        fn.restype = res
        fn.argtypes = args
    return lib


lib = load()
ABI_VERSION = lib.b200_abi_version()


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib.b200_error_string(code)
        raise B200Error(f"{what} failed ({code}): {msg.decode() if msg else '?'}")
