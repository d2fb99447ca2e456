"""TEST INFRASTRUCTURE ONLY — CPU restatement (numpy) of the reference's jagged integer ops.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this; the
product (torchrec-oldfork_b200/) never does.  Each function cites the reference code it
follows (paths relative to the reference checkout; fbgemm_gpu/ = third_party/fbgemm/fbgemm_gpu/).
Pinned against the reference's literal test vectors in tests/test_oracle_golden.py and against
the reference's own CPU ops compiled into oracle/_ref (tests/golden/*.npz).
"""
from __future__ import annotations


import numpy as np


def complete_cumsum(x: np.ndarray) -> np.ndarray:
    """fbgemm_gpu/src/sparse_ops_cpu.cpp:1075 asynchronous_complete_cumsum_cpu: out[0]=0, out[i+1]=sum(x[:i+1])."""
    out = np.zeros(x.size + 1, dtype=x.dtype)
    np.cumsum(x, out=out[1:], dtype=x.dtype)
    return out


def inclusive_cumsum(x: np.ndarray) -> np.ndarray:
    """sparse_ops_cpu.cpp:1059 asynchronous_inclusive_cumsum_cpu."""
    return np.cumsum(x, dtype=x.dtype)


def exclusive_cumsum(x: np.ndarray) -> np.ndarray:
    """sparse_ops_cpu.cpp:1028 asynchronous_exclusive_cumsum_cpu."""
    out = np.zeros_like(x)
    if x.size > 1:
        np.cumsum(x[:-1], out=out[1:], dtype=x.dtype)
    return out


def permute_2D_sparse_data(permute, lengths, values, weights=None):
    """sparse_ops_cpu.cpp:512 permute_2D_sparse_data_cpu with kernels :81-118 and :153-175.

    lengths [T, B]; out_lengths[t, b] = lengths[permute[t], b]; segment (t, b) of the output
    is the input segment (permute[t], b)."""
    T, B = lengths.shape
    Tp = permute.size
    out_lengths = lengths[permute.astype(np.int64), :].copy() if Tp else np.zeros((0, B), lengths.dtype)
    in_off = np.zeros(T * B + 1, dtype=np.int64)
    np.cumsum(lengths.reshape(-1), out=in_off[1:])
    pieces_v, pieces_w = [], []
    for t in range(Tp):
        for b in range(B):
            s = in_off[int(permute[t]) * B + b]
            n = int(out_lengths[t, b])
            pieces_v.append(values[s:s + n])
            if weights is not None:
                pieces_w.append(weights[s:s + n])
    ov = np.concatenate(pieces_v) if pieces_v else values[:0].copy()
    ow = None
    if weights is not None:
        ow = np.concatenate(pieces_w) if pieces_w else weights[:0].copy()
    return out_lengths, ov.astype(values.dtype), ow


def permute_1D_sparse_data(permute, lengths, values, weights=None):
    """sparse_ops_cpu.cpp:645 permute_1D_sparse_data_cpu (kernel :119-152): B == 1 case."""
    ol, ov, ow = permute_2D_sparse_data(permute, lengths.reshape(-1, 1), values, weights)
    return ol.reshape(-1), ov, ow


def expand_into_jagged_permute(permute, input_offsets, output_offsets, output_size):
    """sparse_ops_cpu.cpp:749-797 expand_into_jagged_permute_cpu:
    out[output_offsets[i] + k] = input_offsets[permute[i]] + k."""
    out = np.zeros(output_size, dtype=input_offsets.dtype)
    for i in range(permute.size):
        n = int(output_offsets[i + 1] - output_offsets[i])
        src = int(input_offsets[int(permute[i])])
        dst = int(output_offsets[i])
        out[dst:dst + n] = np.arange(src, src + n, dtype=input_offsets.dtype)
    return out


def offsets_range(offsets, range_size):
    """sparse_ops_cpu.cpp:1221 offsets_range_cpu: out[offsets[i] + k] = k within segment i."""
    out = np.zeros(range_size, dtype=offsets.dtype)
    n = offsets.size
    for i in range(n):
        s = int(offsets[i])
        e = int(offsets[i + 1]) if i + 1 < n else int(range_size)
        out[s:e] = np.arange(0, e - s, dtype=offsets.dtype)
    return out


def block_bucketize_sparse_features(lengths, indices, bucketize_pos, sequence, block_sizes, my_size,
                                    weights=None):
    """fbgemm_gpu/src/sparse_ops.cu:736-832 (_block_bucketize_sparse_features_cuda_kernel1/2).

    Follows the CUDA kernels' UNSIGNED arithmetic (:759-761, :806-809), which is what the
    reference's own known-answer test pins for negative / >2^46 ids
    (fbgemm_gpu/test/sparse_ops_test.py:446-511); for in-range ids it equals the CPU op
    (sparse_ops_cpu.cpp:240-343)."""
    W = int(my_size)
    T = block_sizes.size
    TB = lengths.size
    B = TB // T
    idt = indices.dtype
    udt = np.uint64 if idt == np.int64 else np.uint32
    bits = 64 if idt == np.int64 else 32
    mask = (1 << bits) - 1
    off = np.zeros(TB + 1, dtype=np.int64)
    np.cumsum(lengths, out=off[1:])
    new_lengths = np.zeros(TB * W, dtype=lengths.dtype)
    ps = np.zeros(indices.size, dtype=np.int64)
    nis = np.zeros(indices.size, dtype=udt)
    for bt in range(TB):
        blk = int(block_sizes[bt // B])
        ublk = blk & mask
        bound = (blk * W) & mask
        for i in range(off[bt], off[bt + 1]):
            idx = int(indices[i]) & mask
            if idx < bound:
                p, ni = idx // ublk, idx % ublk
            else:
                p, ni = idx % W, idx // W
            ps[i], nis[i] = p, ni
            new_lengths[p * TB + bt] += 1
    noff = np.zeros(TB * W + 1, dtype=np.int64)
    np.cumsum(new_lengths, out=noff[1:])
    cursor = noff[:-1].copy()
    new_indices = np.zeros_like(indices)
    new_weights = np.zeros_like(weights) if weights is not None else None
    new_pos = np.zeros_like(indices) if bucketize_pos else None
    unbucketize = np.zeros_like(indices) if sequence else None
    for bt in range(TB):
        for i in range(off[bt], off[bt + 1]):
            slot = ps[i] * TB + bt
            pos = cursor[slot]
            cursor[slot] += 1
            new_indices[pos] = np.array(nis[i], dtype=udt).astype(idt)
            if weights is not None:
                new_weights[pos] = weights[i]
            if bucketize_pos:
                new_pos[pos] = i - off[bt]
            if sequence:
                unbucketize[i] = pos
    return new_lengths, new_indices, new_weights, new_pos, unbucketize


def bounds_check_indices(rows_per_table, indices, offsets, mode):
    """fbgemm_gpu/codegen/embedding_bounds_check.cu:12-110 applied bag after bag, i.e. the
    sequential CPU order of embedding_bounds_check_host_cpu.cpp.  mode 1 = WARNING, 2 = IGNORE
    (FATAL asserts instead).  Returns (indices, offsets, n_warnings) — the inputs are copied."""
    indices = indices.copy()
    offsets = offsets.copy()
    T = rows_per_table.size
    B = (offsets.size - 1) // T
    N = indices.size
    warn = 0
    for t in range(T):
        for b in range(B):
            bt = t * B + b
            s, e = int(offsets[bt]), int(offsets[bt + 1])
            if s < 0 or s > e or e > N:
                warn += 1
                s = max(0, min(s, N))
                e = max(s, min(e, N))
                offsets[bt], offsets[bt + 1] = s, e
            for i in range(s, e):
                idx = int(indices[i])
                if idx == -1:
                    continue
                if idx < 0 or idx >= int(rows_per_table[t]):
                    warn += 1
                    indices[i] = 0
    return indices, offsets, warn


def get_recat(local_split: int, num_splits: int, stagger: int = 1):
    """torchrec/distributed/dist_data.py:40-81 _get_recat (docstring example
    _get_recat(2, 4, 1) == [0, 2, 4, 6, 1, 3, 5, 7])."""
    recat = []
    feature_order = [x + num_splits // stagger * y for x in range(num_splits // stagger) for y in range(stagger)]
    for i in range(local_split):
        for j in feature_order:
            recat.append(i + j * local_split)
    return recat
