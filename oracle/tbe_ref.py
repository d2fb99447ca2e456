"""TEST INFRASTRUCTURE ONLY — CPU restatement (numpy) of the reference's pooled lookup,
its fused backward+optimizer, the int8 rowwise path and the DLRM dot interaction.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this.
Pinned against: the reference's own scalar C++ (third_party/fbgemm/src/RefImplementations.cc
compiled unmodified into oracle/_ref/libfbgemm_ref.so, see oracle/Makefile), torch.nn.EmbeddingBag
as the reference's tests do (fbgemm_gpu/test/split_table_batched_embeddings_test.py:322-415),
and the InteractionArch golden tensor of torchrec/models/tests/test_dlrm.py:247-327.
"""
from __future__ import annotations

import numpy as np


def tbe_forward(weights, weights_offsets, D_offsets, indices, offsets, pooling_mean=False,
                per_sample_weights=None, out_dtype=np.float32):
    """fbgemm_gpu/codegen/embedding_forward_split_cpu.cpp:117-156 (scalar branch), same
    arithmetic as EmbeddingSpMDM_ref (third_party/fbgemm/src/RefImplementations.cc:1186-1227):
    acc = fma(w, row, acc) in fp32, MEAN scales by 1/L when L > 0; bag (t,b) at offsets[t*B+b]."""
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    total_D = int(D_offsets[-1])
    out = np.zeros((B, total_D), dtype=np.float32)
    for t in range(T):
        d0, D = int(D_offsets[t]), int(D_offsets[t + 1] - D_offsets[t])
        base = int(weights_offsets[t])
        for b in range(B):
            s, e = int(offsets[t * B + b]), int(offsets[t * B + b + 1])
            acc = np.zeros(D, dtype=np.float32)
            for p in range(s, e):
                row = weights[base + int(indices[p]) * D: base + (int(indices[p]) + 1) * D].astype(np.float32)
                if per_sample_weights is not None:
                    # std::fma(w, x, acc): the float64 product of two float32 is exact
                    acc = (np.float64(np.float32(per_sample_weights[p])) * row.astype(np.float64)
                           + acc.astype(np.float64)).astype(np.float32)
                else:
                    acc = acc + row
            if pooling_mean and e > s:
                acc = acc * (np.float32(1.0) / np.float32(e - s))
            out[b, d0:d0 + D] = acc
    return out.astype(out_dtype)


def tbe_backward(grad_output, weights, weights_offsets, D_offsets, hash_size_cumsum, indices, offsets,
                 optimizer, lr, eps=0.0, momentum1=None, momentum1_offsets=None, pooling_mean=False,
                 per_sample_weights=None):
    """fbgemm_gpu/codegen/embedding_backward_split_cpu_template.cpp:193-224 (scalar branch) with
    split_weight_update_cpu of rowwise_adagrad (embedding_backward_code_generator.py:402-415)
    and sgd (:633-637); MEAN divides the bag gradient by L (grad_mean_kernel,
    embedding_backward_split_template.cu:41-77).  Duplicates of one (table,row) are summed
    first, then ONE update is applied.  Updates `weights` / `momentum1` in place (fp32 math,
    result cast back to the weight dtype with round-to-nearest)."""
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    groups = {}
    for t in range(T):
        for b in range(B):
            s, e = int(offsets[t * B + b]), int(offsets[t * B + b + 1])
            for p in range(s, e):
                key = int(hash_size_cumsum[t]) + int(indices[p])
                groups.setdefault(key, []).append((t, b, p, e - s))
    for key in sorted(groups):
        occ = groups[key]
        t0 = occ[0][0]
        D = int(D_offsets[t0 + 1] - D_offsets[t0])
        g = np.zeros(D, dtype=np.float32)
        for (t, b, p, L) in occ:  # stable order: t-major, then b, then in-bag position
            row = grad_output[b, int(D_offsets[t]): int(D_offsets[t]) + D].astype(np.float32)
            sc = np.float32(1.0)
            if pooling_mean:
                sc = np.float32(1.0 / L)
            if per_sample_weights is not None:
                sc = np.float32(sc * np.float32(per_sample_weights[p]))
            g = g + sc * row
        idx = key - int(hash_size_cumsum[t0])
        base = int(weights_offsets[t0]) + idx * D
        w = weights[base: base + D].astype(np.float32)
        if optimizer == "rowwise_adagrad":
            g2 = np.float32(np.sum(g.astype(np.float64) ** 2) / D)
            mi = int(momentum1_offsets[t0]) + idx
            m = np.float32(momentum1[mi] + g2)
            momentum1[mi] = m
            mult = np.float32(lr) / (np.sqrt(m, dtype=np.float32) + np.float32(eps))
            w = w - g * mult
        elif optimizer == "sgd":
            w = w - np.float32(lr) * g
        else:
            raise ValueError(optimizer)
        weights[base: base + D] = w.astype(weights.dtype)


def grad_indice_weights(grad_output, weights, weights_offsets, D_offsets, indices, offsets):
    """fbgemm_gpu/codegen/embedding_forward_split_cpu.cpp:229-300
    split_embedding_grad_indice_weights_cpu_kernel: g[p] = <grad_out[b, D_t], W_t[idx_p]>."""
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    out = np.zeros(len(indices), dtype=np.float32)
    for t in range(T):
        d0, D = int(D_offsets[t]), int(D_offsets[t + 1] - D_offsets[t])
        base = int(weights_offsets[t])
        for b in range(B):
            for p in range(int(offsets[t * B + b]), int(offsets[t * B + b + 1])):
                row = weights[base + int(indices[p]) * D: base + (int(indices[p]) + 1) * D].astype(np.float32)
                out[p] = np.dot(grad_output[b, d0:d0 + D].astype(np.float32), row)
    return out


def quantize_rowwise_8bit_sbhalf(w):
    """third_party/fbgemm/src/QuantUtils.cc:499-571 FloatOrHalfToFusedNBitRowwiseQuantizedSBHalfRef
    with bit_rate 8: row = [ncols x uint8][fp16 scale][fp16 bias(min)]."""
    w = np.asarray(w, dtype=np.float32)
    nrows, ncols = w.shape
    out = np.zeros((nrows, ncols + 4), dtype=np.uint8)
    for r in range(nrows):
        mn16 = np.float16(w[r].min())
        mn = np.float32(mn16)
        rng = np.float32(w[r].max()) - mn
        scale = np.float32(1.0) if rng == 0 else np.float32(rng / np.float32(255))
        scale = np.float32(np.float16(scale))
        if scale == 0:
            scale = np.float32(1.0)
        with np.errstate(divide="ignore", over="ignore"):
            inv = np.float32(1.0) / scale
        if np.isinf(inv):
            scale, inv = np.float32(1.0), np.float32(1.0)
        q = np.rint((w[r] - mn) * inv)  # lrintf: round half to even
        out[r, :ncols] = np.clip(q, 0, 255).astype(np.uint8)
        out[r, ncols:ncols + 2] = np.frombuffer(np.float16(scale).tobytes(), dtype=np.uint8)
        out[r, ncols + 2:] = np.frombuffer(mn16.tobytes(), dtype=np.uint8)
    return out


def int8_row_bytes(D, row_alignment=16):
    """split_table_batched_embeddings_ops.py:1469-1485 rounded_row_size_in_bytes(INT8)."""
    return (D + 4 + row_alignment - 1) // row_alignment * row_alignment


def tbe_int8_forward(weights_u8, weights_offsets, D_offsets, indices, offsets, row_alignment=16,
                     pooling_mean=False, per_sample_weights=None, out_dtype=np.float16):
    """EmbeddingSpMDM_ref 8-bit branch with scale_bias_last=false
    (third_party/fbgemm/src/RefImplementations.cc:1114-1179): row = [fp16 scale, fp16 bias, q...];
    buf = fma(scale*w, q, buf + bias*w); idx == -1 is a pruned row (zeros,
    embedding_forward_quantized_split_template.cu:276-278)."""
    T = len(D_offsets) - 1
    B = (len(offsets) - 1) // T
    out = np.zeros((B, int(D_offsets[-1])), dtype=np.float32)
    for t in range(T):
        d0, D = int(D_offsets[t]), int(D_offsets[t + 1] - D_offsets[t])
        rb = int8_row_bytes(D, row_alignment)
        base = int(weights_offsets[t])
        for b in range(B):
            s, e = int(offsets[t * B + b]), int(offsets[t * B + b + 1])
            acc = np.zeros(D, dtype=np.float32)
            for p in range(s, e):
                idx = int(indices[p])
                if idx == -1:
                    continue
                row = weights_u8[base + idx * rb: base + idx * rb + 4 + D]
                sb = np.frombuffer(row[:4].tobytes(), dtype=np.float16).astype(np.float32)
                pw = np.float32(1.0) if per_sample_weights is None else np.float32(per_sample_weights[p])
                sc, bi = np.float32(sb[0] * pw), np.float32(sb[1] * pw)
                acc = (np.float64(sc) * row[4:4 + D].astype(np.float64)
                       + (acc + bi).astype(np.float64)).astype(np.float32)  # fma(scale, q, buf + bias)
            if pooling_mean and e > s:
                acc = acc * (np.float32(1.0) / np.float32(e - s))
            out[b, d0:d0 + D] = acc
    return out.astype(out_dtype)


def interaction_forward(dense, sparse):
    """torchrec/models/dlrm.py:193-219 InteractionArch.forward: X = cat(dense[:,None], sparse);
    Z = bmm(X, X^T); out = cat(dense, Z[:, triu_indices(F+1, F+1, offset=1)]) (row-major)."""
    B, D = dense.shape
    X = np.concatenate([dense[:, None, :], sparse], axis=1).astype(np.float64)
    Z = np.einsum("bid,bjd->bij", X, X)
    iu = np.triu_indices(X.shape[1], k=1)
    return np.concatenate([dense.astype(np.float64), Z[:, iu[0], iu[1]]], axis=1).astype(np.float32)


def interaction_backward(dense, sparse, grad_out):
    """Autograd of the above: dX = (G + G^T) X with G the strict upper triangle; the dense row
    also receives grad_out[:, :D]."""
    B, D = dense.shape
    X = np.concatenate([dense[:, None, :], sparse], axis=1).astype(np.float64)
    n = X.shape[1]
    iu = np.triu_indices(n, k=1)
    G = np.zeros((B, n, n), dtype=np.float64)
    G[:, iu[0], iu[1]] = grad_out[:, D:]
    Gs = G + G.transpose(0, 2, 1)
    dX = np.einsum("bij,bjd->bid", Gs, X)
    gd = dX[:, 0, :] + grad_out[:, :D]
    return gd.astype(np.float32), dX[:, 1:, :].astype(np.float32)


# ---------------------------------------------------------------------------------------------
# Stochastic rounding of fp32 -> fp16 / bf16 weight rows (reference: stochastic_rounding_vector /
# stochastic_rounding_scalar, fbgemm_gpu/include/fbgemm_gpu/fbgemm_cuda_utils.cuh, driven by
# embedding_backward_split_template.cu:367-380 with a Philox4x32-10 stream).  The generator is the
# published Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11;
# Random123 / cuRAND), pinned in tests/test_oracle_golden.py by the Random123 known-answer vectors.
# ---------------------------------------------------------------------------------------------
_PHILOX_M0, _PHILOX_M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
_PHILOX_W0, _PHILOX_W1 = 0x9E3779B9, 0xBB67AE85


def philox4x32_10(ctr, key):
    """ctr: uint32 array [..., 4], key: two uint32 scalars (or arrays broadcastable to ctr[..., 0]).
    Returns uint32 [..., 4]."""
    c = [np.asarray(ctr)[..., i].astype(np.uint64) for i in range(4)]
    k0 = np.asarray(key[0]).astype(np.uint64)
    k1 = np.asarray(key[1]).astype(np.uint64)
    mask = np.uint64(0xFFFFFFFF)
    for r in range(10):
        p0 = _PHILOX_M0 * c[0]
        p1 = _PHILOX_M1 * c[2]
        hi0, lo0 = p0 >> np.uint64(32), p0 & mask
        hi1, lo1 = p1 >> np.uint64(32), p1 & mask
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0 = (k0 + np.uint64(_PHILOX_W0)) & mask
        k1 = (k1 + np.uint64(_PHILOX_W1)) & mask
    return np.stack([x.astype(np.uint32) for x in c], axis=-1)


def stochastic_round(values, dtype, seed, offset, first_element, epv):
    """What b200_tbe_backward_sr stores for the fp32 `values` (1-D, a whole number of EPV-element
    vectors starting at element `first_element` of the weight buffer): fp32 bit patterns that are
    exactly representable in `dtype` ("fp16": 13 low mantissa bits dropped, "bf16": 16).

    One Philox call per vector: counter (vector index lo, hi, offset lo, hi), key (seed lo, hi).
    EPV = 8: element 2i of the vector uses the low, 2i+1 the high 16 bits of word i; EPV = 4: element i
    uses the high 16 bits of word i.  Random bits (masked to the dropped width) are added to the fp32
    bit pattern and the dropped bits cleared; inf / nan pass through."""
    v = np.ascontiguousarray(values, dtype=np.float32)
    assert v.ndim == 1 and v.size % epv == 0 and first_element % epv == 0
    nvec = v.size // epv
    vec = np.arange(nvec, dtype=np.uint64) + np.uint64(first_element // epv)
    ctr = np.stack([(vec & np.uint64(0xFFFFFFFF)).astype(np.uint32), (vec >> np.uint64(32)).astype(np.uint32),
                    np.full(nvec, offset & 0xFFFFFFFF, dtype=np.uint32),
                    np.full(nvec, (offset >> 32) & 0xFFFFFFFF, dtype=np.uint32)], axis=-1)
    r = philox4x32_10(ctr, (np.uint32(seed & 0xFFFFFFFF), np.uint32((seed >> 32) & 0xFFFFFFFF)))  # [nvec, 4]
    if epv == 8:
        r16 = np.stack([r & np.uint32(0xFFFF), r >> np.uint32(16)], axis=-1).reshape(nvec, 8)
    else:
        assert epv == 4
        r16 = r >> np.uint32(16)
    mask = np.uint32(0xFFFF if dtype == "bf16" else 0x1FFF)
    u = v.view(np.uint32).reshape(nvec, epv)
    special = (u & np.uint32(0x7F800000)) == np.uint32(0x7F800000)
    out = (u + (r16.astype(np.uint32) & mask)) & ~mask
    return np.where(special, u, out).reshape(-1).view(np.float32)
