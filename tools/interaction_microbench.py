"""Times the dot-interaction kernels at the Criteo shape (B=65536, F=26, D=128) with CUDA events."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))
from b200_sparse import ops  # noqa: E402


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    B, F, D = 65536, 26, 128
    n = F + 1
    dense = torch.randn(B, D, device="cuda")
    sparse = torch.randn(B, F, D, device="cuda")
    out = ops.interaction_forward(dense, sparse)
    go = torch.randn_like(out)
    fwd_bytes = 4 * B * (n * D + out.shape[1])
    bwd_bytes = 4 * B * (2 * n * D + out.shape[1])
    us = timeit(lambda: ops.interaction_forward(dense, sparse))
    print(f"interaction forward : {us:8.1f} us  {fwd_bytes / us / 1e3:7.1f} GB/s algorithmic")
    us = timeit(lambda: ops.interaction_backward(dense, sparse, go))
    print(f"interaction backward: {us:8.1f} us  {bwd_bytes / us / 1e3:7.1f} GB/s algorithmic")
    X = torch.cat([dense.unsqueeze(1), sparse], 1)
    us = timeit(lambda: torch.bmm(X, X.transpose(1, 2)))
    print(f"torch.bmm alone     : {us:8.1f} us")


if __name__ == "__main__":
    main()
