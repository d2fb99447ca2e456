"""Generates tests/golden/jagged_ops.npz by running the REFERENCE's own CPU kernels
(fbgemm_gpu/src/sparse_ops_cpu.cpp and codegen/embedding_bounds_check_host_cpu.cpp, compiled
unmodified into oracle/_ref/libfbgemm_gpu_cpu_ref.so by oracle/Makefile) on seeded inputs.

Run here (the container with /root/reference):   python tests/golden/make_golden.py
It must run in its own process: the .so registers the same fbgemm::* schemas the product does.
The fixtures are small and committed; the tests never need /root/reference."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
SO = os.path.join(ROOT, "oracle", "_ref", "libfbgemm_gpu_cpu_ref.so")


def main() -> None:
    assert "fbgemm_gpu" not in sys.modules
    torch.ops.load_library(SO)
    F = torch.ops.fbgemm
    out = {}
    rng = np.random.default_rng(20211104)
    n_case = 0

    def put(prefix, **arrs):
        for k, v in arrs.items():
            if v is None:
                continue
            out[f"{prefix}/{k}"] = v.numpy() if isinstance(v, torch.Tensor) else np.asarray(v)

    # ---- cumsums (sparse_ops_cpu.cpp:1028-1090)
    for i, (n, dt) in enumerate([(0, np.int64), (1, np.int32), (257, np.int32), (5000, np.int64)]):
        x = torch.from_numpy(rng.integers(0, 1000, size=n).astype(dt))
        put(f"cumsum{i}", x=x, complete=F.asynchronous_complete_cumsum(x),
            inclusive=F.asynchronous_inclusive_cumsum(x), exclusive=F.asynchronous_exclusive_cumsum(x))
        n_case += 1

    # ---- permute_2D / permute_1D (sparse_ops_cpu.cpp:512-650), duplicates and dropped rows in permute
    cases = [(3, 4, 5, [2, 0, 1], False, np.int64), (5, 7, 3, [4, 4, 0, 2], True, np.int64),
             (26, 16, 2, list(rng.permutation(26)), False, np.int32), (4, 1, 0, [3, 2, 1, 0], False, np.int64),
             (6, 33, 9, [5, 0, 0, 3, 1, 2, 4, 4], True, np.int64)]
    for i, (T, B, maxl, perm, weighted, dt) in enumerate(cases):
        lengths = torch.from_numpy(rng.integers(0, maxl + 1, size=(T, B)).astype(np.int32))
        N = int(lengths.sum())
        values = torch.from_numpy(rng.integers(0, 1 << 40 if dt == np.int64 else 1 << 30, size=N).astype(dt))
        weights = torch.from_numpy(rng.random(N).astype(np.float32)) if weighted else None
        p = torch.tensor([int(x) for x in perm], dtype=torch.int32)
        l2, v2, w2 = F.permute_2D_sparse_data(p, lengths, values, weights, None)
        put(f"permute2d{i}", permute=p, lengths=lengths, values=values, weights=weights, out_lengths=l2,
            out_values=v2, out_weights=w2)
        l1 = lengths.reshape(-1)
        p1 = torch.from_numpy(rng.integers(0, l1.numel(), size=max(1, l1.numel() // 2)).astype(np.int32))
        o1, ov1, ow1 = F.permute_1D_sparse_data(p1, l1, values, weights, None)
        put(f"permute1d{i}", permute=p1, lengths=l1, values=values, weights=weights, out_lengths=o1,
            out_values=ov1, out_weights=ow1)
        n_case += 2

    # ---- block_bucketize_sparse_features (sparse_ops_cpu.cpp:240-343,806-880).  Ids stay inside
    # [0, blk*W): outside it the reference's CPU kernel (signed arithmetic, idx % blk) and its CUDA
    # kernel (unsigned, idx / W; sparse_ops.cu:759-760,806-809) disagree and the CPU one indexes out of
    # bounds for negative ids.  The out-of-range behaviour is pinned on the CUDA kernel by the literal
    # vectors of fbgemm_gpu/test/sparse_ops_test.py:602-750 (tests/test_oracle_golden.py).
    bcases = [(2, 4, 3, 2, False, False, False), (3, 5, 4, 4, True, True, True), (26, 8, 2, 8, True, False, False),
              (1, 1, 0, 2, True, False, False), (4, 16, 6, 3, True, True, False), (2, 3, 5, 2, False, True, True)]
    for i, (T, B, maxl, W, seq, pos, weighted) in enumerate(bcases):
        lengths = torch.from_numpy(rng.integers(0, maxl + 1, size=T * B).astype(np.int32))
        N = int(lengths.sum())
        block = rng.integers(1, 50, size=T).astype(np.int64)
        per_t = lengths.view(T, B).sum(1).numpy()
        ids = np.concatenate([rng.integers(0, int(block[t]) * W, size=int(per_t[t])) for t in range(T)]
                             ).astype(np.int64) if N else np.zeros(0, np.int64)
        weights = torch.from_numpy(rng.random(N).astype(np.float32)) if weighted else None
        r = F.block_bucketize_sparse_features(lengths, torch.from_numpy(ids), pos, seq, torch.from_numpy(block), W,
                                              weights)
        put(f"bucketize{i}", lengths=lengths, indices=ids, block_sizes=block, my_size=np.int64(W),
            sequence=np.int64(seq), bucketize_pos=np.int64(pos), weights=weights, new_lengths=r[0],
            new_indices=r[1], new_weights=r[2], new_pos=r[3], unbucketize_permute=r[4])
        n_case += 1

    # ---- expand_into_jagged_permute / offsets_range (sparse_ops_cpu.cpp:749-797, 81-152)
    for i, (n, dt) in enumerate([(1, np.int32), (5, np.int64), (40, np.int32)]):
        lens = rng.integers(0, 7, size=n).astype(np.int64)
        perm = rng.permutation(n).astype(dt)  # the reference requires permute and offsets of one dtype
        in_off = np.concatenate([[0], np.cumsum(lens)]).astype(dt)
        out_off = np.concatenate([[0], np.cumsum(lens[perm])]).astype(dt)
        e = F.expand_into_jagged_permute(torch.from_numpy(perm), torch.from_numpy(in_off), torch.from_numpy(out_off),
                                         int(out_off[-1]))
        orng = F.offsets_range(torch.from_numpy(in_off[:-1].copy()), int(in_off[-1]))
        put(f"expand{i}", permute=perm, input_offsets=in_off, output_offsets=out_off, out=e)
        put(f"offsets_range{i}", offsets=in_off[:-1], range_size=np.int64(in_off[-1]), out=orng)
        n_case += 2

    # ---- bounds_check_indices (embedding_bounds_check_host_cpu.cpp), modes WARNING(1) and IGNORE(2)
    for i, (T, B, dt) in enumerate([(3, 4, np.int64), (5, 9, np.int32), (2, 1, np.int64)]):
        rows = rng.integers(1, 30, size=T).astype(np.int64)
        lengths = rng.integers(0, 5, size=T * B)
        offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(dt)
        N = int(offsets[-1])
        ids = np.concatenate([rng.integers(0, rows[t], size=int(lengths[t * B:(t + 1) * B].sum())) for t in range(T)]
                             ).astype(dt) if N else np.zeros(0, dt)
        bad = rng.integers(0, max(N, 1), size=min(N, 3))
        for k, b in enumerate(bad):
            ids[b] = -1 - k if k % 2 else 1000 + k
        for corrupt in (False, True):
            off = offsets.copy()
            if corrupt and off.size > 3:
                off[2], off[1] = off[1], off[2] + 1   # start > end somewhere
                off[-1] = N + 3                      # last offset beyond numel
            for mode in (1, 2):
                ti, to = torch.from_numpy(ids.copy()), torch.from_numpy(off.copy())
                warn = torch.zeros(1, dtype=torch.int64)
                F.bounds_check_indices(torch.from_numpy(rows), ti, to, mode, warn)
                put(f"bounds{i}_{int(corrupt)}_{mode}", rows=rows, indices=ids, offsets=off, out_indices=ti,
                    out_offsets=to, warning=warn, B=np.int64(B))
                n_case += 1

    path = os.path.join(HERE, "jagged_ops.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {n_case} cases, {len(out)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
