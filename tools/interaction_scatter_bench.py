"""Times the dot-interaction backward at the Criteo shape (B 65536, F 26, D 128) with its two store
paths, both writing LOCAL memory: the dense [B, F, D] layout (8-byte stores from the accumulator
fragments) and the per-feature scatter (quarter samples staged in shared memory, 128-byte row
segments).  CUDA events, L2 flushed between iterations, median of 20."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))


def main() -> None:
    from b200_sparse import ops
    from b200_sparse.cabi import load
    lib = load()
    dev = torch.device("cuda", 0)
    B, F, D = 65536, 26, 128
    width = D + (F + 1) * F // 2
    pitch = (width + 3) // 4 * 4
    dense = torch.randn(B, D, device=dev)
    sparse = torch.randn(B, F, D, device=dev)
    grad = torch.randn(B, pitch, device=dev)
    gd = torch.empty_like(dense)
    gs = torch.empty_like(sparse)
    gs2 = torch.empty_like(sparse)
    fb = torch.tensor([gs2.data_ptr() + 4 * f * D for f in range(F)], dtype=torch.int64, device=dev)
    fp = torch.full((F,), F * D, dtype=torch.int64, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream

    def plain():
        assert lib.b200_interaction_backward(dense.data_ptr(), sparse.data_ptr(), grad.data_ptr(), pitch, B, F, D,
                                             gd.data_ptr(), gs.data_ptr(), st) == 0

    def scatter():
        assert lib.b200_interaction_backward_scatter(dense.data_ptr(), sparse.data_ptr(), grad.data_ptr(), pitch, B,
                                                     F, D, gd.data_ptr(), fb.data_ptr(), fp.data_ptr(), st) == 0

    out = {}
    for name, fn in (("dense_layout", plain), ("scatter_staged", scatter)):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(20):
            flush.fill_(1)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        out[name] = {"us": round(float(np.median(ts)) * 1e3, 1)}
    assert torch.equal(gs, gs2), "the two store paths must produce identical gradients"
    bytes_alg = 4 * (B * pitch + 2 * B * (F + 1) * D + B * D)  # grad_out + X read + grad X written (+ dense)
    for v in out.values():
        v["GBps"] = round(bytes_alg / v["us"] / 1e3, 1)
    print(json.dumps({"interaction_backward_store_paths": out, "algorithmic_bytes": bytes_alg}))


if __name__ == "__main__":
    main()
