"""TEST INFRASTRUCTURE ONLY: ctypes access to oracle/_build/liboracle.so (tbe_ref.c) and, when it
was built in the authoring container, oracle/_ref/libfbgemm_ref.so (the reference's unmodified
RefImplementations.cc behind ref_shim.cc)."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "liboracle.so")
REF_LIB = os.path.join(HERE, "_ref", "libfbgemm_ref.so")


def build() -> None:
    """Compile the C restatement (and oracle/_ref when /root/reference is present)."""
    subprocess.run(["make", "-C", HERE], check=True, capture_output=True)


def _ptr(a, ty):
    return None if a is None else a.ctypes.data_as(C.POINTER(ty))


_lib = None
_ref = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        _lib = C.CDLL(LIB)
    return _lib


def ref_lib():
    """The reference's own compiled code, or None when it was not built (no /root/reference)."""
    global _ref
    if _ref is None and os.path.exists(REF_LIB):
        _ref = C.CDLL(REF_LIB)
    return _ref


def _c(a, dt):
    return None if a is None else np.ascontiguousarray(a, dtype=dt)


def tbe_forward(weights, weights_offsets, D_offsets, indices, offsets, mean=False, psw=None):
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    w, wo, Do = _c(weights, np.float32), _c(weights_offsets, np.int64), _c(D_offsets, np.int32)
    idx, off, pw = _c(indices, np.int64), _c(offsets, np.int64), _c(psw, np.float32)
    out = np.empty((B, int(Do[-1])), dtype=np.float32)
    lib().oracle_tbe_forward_f32(_ptr(w, C.c_float), _ptr(wo, C.c_int64), _ptr(Do, C.c_int32), C.c_int32(T),
                                 C.c_int32(B), _ptr(idx, C.c_int64), _ptr(off, C.c_int64), _ptr(pw, C.c_float),
                                 C.c_int(int(mean)), _ptr(out, C.c_float))
    return out


def tbe_backward(grad, weights, weights_offsets, D_offsets, hash_size_cumsum, indices, offsets, optimizer, lr,
                 eps=0.0, momentum1=None, momentum1_offsets=None, mean=False, psw=None):
    """In place on `weights` (float32 C-contiguous) and `momentum1`."""
    assert weights.dtype == np.float32 and weights.flags.c_contiguous
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    g = _c(grad, np.float32)
    wo, Do, hsc = _c(weights_offsets, np.int64), _c(D_offsets, np.int32), _c(hash_size_cumsum, np.int64)
    idx, off, pw = _c(indices, np.int64), _c(offsets, np.int64), _c(psw, np.float32)
    mo = _c(momentum1_offsets, np.int64)
    if momentum1 is not None:
        assert momentum1.dtype == np.float32 and momentum1.flags.c_contiguous
    lib().oracle_tbe_backward_f32(
        _ptr(g, C.c_float), C.c_int64(g.shape[1]), _ptr(weights, C.c_float), _ptr(wo, C.c_int64),
        _ptr(Do, C.c_int32), _ptr(hsc, C.c_int64), C.c_int32(T), C.c_int32(B), _ptr(idx, C.c_int64),
        _ptr(off, C.c_int64), _ptr(pw, C.c_float), C.c_int(int(mean)),
        C.c_int(1 if optimizer == "rowwise_adagrad" else 0), _ptr(momentum1, C.c_float), _ptr(mo, C.c_int64),
        C.c_float(lr), C.c_float(eps))


def interaction_forward(dense, sparse):
    d, s = _c(dense, np.float32), _c(sparse, np.float32)
    B, D = d.shape
    F = s.shape[1]
    out = np.empty((B, D + (F + 1) * F // 2), dtype=np.float32)
    lib().oracle_interaction_forward(_ptr(d, C.c_float), _ptr(s, C.c_float), C.c_int32(B), C.c_int32(F),
                                     C.c_int32(D), _ptr(out, C.c_float))
    return out


# ---- the reference's own code (only where oracle/_ref exists)
def ref_spmdm_f32(table, indices, offsets, psw=None, mean=False):
    r = ref_lib()
    tab, idx, off, pw = _c(table, np.float32), _c(indices, np.int64), _c(offsets, np.int64), _c(psw, np.float32)
    rows, D = tab.shape
    n_out = len(off) - 1
    out = np.zeros((n_out, D), dtype=np.float32)
    ok = r.ref_spmdm_f32(C.c_int64(D), C.c_int64(n_out), C.c_int64(len(idx)), C.c_int64(rows),
                         _ptr(tab, C.c_float), _ptr(idx, C.c_int64), _ptr(off, C.c_int64), _ptr(pw, C.c_float),
                         C.c_int(int(mean)), _ptr(out, C.c_float), C.c_int64(D))
    assert ok == 1
    return out


def ref_spmdm_u8(packed_rows, D, indices, offsets, psw=None, mean=False):
    r = ref_lib()
    tab, idx, off, pw = _c(packed_rows, np.uint8), _c(indices, np.int64), _c(offsets, np.int64), _c(psw, np.float32)
    rows, stride = tab.shape
    n_out = len(off) - 1
    out = np.zeros((n_out, D), dtype=np.float32)
    ok = r.ref_spmdm_u8_sbfront(C.c_int64(D), C.c_int64(n_out), C.c_int64(len(idx)), C.c_int64(rows),
                                _ptr(tab, C.c_uint8), _ptr(idx, C.c_int64), _ptr(off, C.c_int64),
                                _ptr(pw, C.c_float), C.c_int(int(mean)), _ptr(out, C.c_float), C.c_int64(D),
                                C.c_int64(stride))
    assert ok == 1
    return out


def ref_rowwise_adagrad(w, g, h, indices, eps, lr):
    """w [rows, D] and h [rows] updated in place with gradient rows g [n, D] for `indices`."""
    r = ref_lib()
    idx = _c(indices, np.int64)
    gg = _c(g, np.float32)
    n = r.ref_rowwise_adagrad(C.c_int(len(idx)), C.c_int(w.shape[1]), C.c_uint64(w.size), _ptr(w, C.c_float),
                              _ptr(gg, C.c_float), _ptr(h, C.c_float), _ptr(idx, C.c_int64), C.c_float(eps),
                              C.c_float(lr))
    assert n == len(idx)


QUANT_REF_LIB = os.path.join(HERE, "_ref", "libfbgemm_quant_ref.so")
_quant = None


def quant_ref_lib():
    """The reference's own QuantUtils.cc (oracle/quant_shim.cc), or None when oracle/_ref was not built."""
    global _quant
    if _quant is None and os.path.exists(QUANT_REF_LIB):
        _quant = C.CDLL(QUANT_REF_LIB)
        for f in (_quant.ref_quantize_sbhalf_ref, _quant.ref_quantize_sbhalf):
            f.restype = None
            f.argtypes = [C.c_int, C.POINTER(C.c_float), C.c_size_t, C.c_int, C.POINTER(C.c_uint8)]
    return _quant


def ref_quantize_sbhalf(w, bit_rate=8, dispatch=False):
    """[rows, cols] fp32 -> [rows, cols + 4] uint8 ([q...][fp16 scale][fp16 bias]) by the reference's
    FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef (dispatch=False) or by the dispatcher its torch op calls
    (dispatch=True: the AVX2 kernel on an AVX2 host)."""
    q = quant_ref_lib()
    x = _c(w, np.float32)
    rows, cols = x.shape
    out = np.zeros((rows, cols * bit_rate // 8 + 4), dtype=np.uint8)
    (q.ref_quantize_sbhalf if dispatch else q.ref_quantize_sbhalf_ref)(
        bit_rate, _ptr(x, C.c_float), rows, cols, _ptr(out, C.c_uint8))
    return out


TBE_REF_LIB = os.path.join(HERE, "_ref", "libfbgemm_tbe_ref.so")
_tbe_ref = None


def tbe_ref_lib():
    """The reference's own JIT CPU embedding kernels behind oracle/ref_tbe_harness.cc, or None when not built."""
    global _tbe_ref
    if _tbe_ref is None and os.path.exists(TBE_REF_LIB):
        _tbe_ref = C.CDLL(TBE_REF_LIB)
    return _tbe_ref


def ref_tbe_forward(weights, weights_offsets, D_offsets, rows, indices, offsets, psw=None, mean=False):
    """embedding_forward_split_cpu.cpp:71-116 through FBGEMM's GenerateEmbeddingSpMDMWithStrides kernels."""
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    w, wo, Do = _c(weights, np.float32), _c(weights_offsets, np.int64), _c(D_offsets, np.int32)
    idx, off, pw, rr = _c(indices, np.int64), _c(offsets, np.int64), _c(psw, np.float32), _c(rows, np.int64)
    out = np.zeros((B, int(Do[-1])), dtype=np.float32)
    ok = tbe_ref_lib().ref_tbe_forward_f32(
        _ptr(w, C.c_float), _ptr(wo, C.c_int64), _ptr(Do, C.c_int32), _ptr(rr, C.c_int64), C.c_int32(T),
        C.c_int32(B), _ptr(idx, C.c_int64), _ptr(off, C.c_int64), _ptr(pw, C.c_float), C.c_int(int(mean)),
        _ptr(out, C.c_float), C.c_int64(out.shape[1]))
    assert ok == 1, "the reference kernel reported an out-of-range index"
    return out


def ref_tbe_backward_rowwise_adagrad(grad, weights, weights_offsets, D_offsets, rows, indices, offsets, lr, eps,
                                     momentum1, momentum1_offsets):
    """embedding_backward_split_cpu_template.cpp:125-182 (rowwise Adagrad) through FBGEMM's SpMDM +
    GenerateSparseAdaGrad kernels; in place on `weights` and `momentum1`."""
    assert weights.dtype == np.float32 and weights.flags.c_contiguous
    assert momentum1.dtype == np.float32 and momentum1.flags.c_contiguous
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    g = _c(grad, np.float32)
    wo, Do, rr = _c(weights_offsets, np.int64), _c(D_offsets, np.int32), _c(rows, np.int64)
    idx, off, mo = _c(indices, np.int64), _c(offsets, np.int64), _c(momentum1_offsets, np.int64)
    ok = tbe_ref_lib().ref_tbe_backward_rowwise_adagrad_f32(
        _ptr(g, C.c_float), C.c_int64(g.shape[1]), _ptr(weights, C.c_float), _ptr(wo, C.c_int64),
        _ptr(Do, C.c_int32), _ptr(rr, C.c_int64), C.c_int32(T), C.c_int32(B), _ptr(idx, C.c_int64),
        _ptr(off, C.c_int64), _ptr(momentum1, C.c_float), _ptr(mo, C.c_int64), C.c_float(lr), C.c_float(eps))
    assert ok == 1
