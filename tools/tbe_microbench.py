"""TBE micro-benchmark through the C ABI wrappers: forward, sort, fused backward at the
Criteo-1TB (MLPerf) shape.  CUDA-event timing on the current stream, median of `iters`."""
import argparse
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))
from b200_sparse import ops  # noqa: E402

CRITEO_ROWS = [40000000, 39060, 17295, 7424, 20265, 3, 7122, 1543, 63, 40000000, 3067956, 405282, 10, 2209,
               11938, 155, 4, 976, 14, 40000000, 40000000, 40000000, 590152, 12973, 108, 36]


def timeit(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    return float(np.median(ts)), float(np.min(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=65536)
    ap.add_argument("--D", type=int, default=128)
    ap.add_argument("--L", type=int, default=1)
    ap.add_argument("--cap", type=int, default=40000000)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--dtype", default="fp32")
    ap.add_argument("--opt", default="rowwise_adagrad")
    a = ap.parse_args()
    dev = torch.device("cuda:0")
    rows = [min(r, a.cap) for r in CRITEO_ROWS]
    T, B, D, L = len(rows), a.B, a.D, a.L
    wdt = {"fp32": torch.float32, "fp16": torch.float16}[a.dtype]
    total_rows = sum(rows)
    weights = torch.empty(total_rows * D, dtype=wdt, device=dev).uniform_(-0.01, 0.01)
    hsc = torch.tensor(np.concatenate([[0], np.cumsum(rows)]), dtype=torch.int64, device=dev)
    wo = (hsc[:-1] * D).contiguous()
    Do = torch.arange(0, (T + 1) * D, D, dtype=torch.int32, device=dev)
    mom = torch.zeros(total_rows, dtype=torch.float32, device=dev)
    g = torch.Generator(device=dev).manual_seed(1)
    idx = torch.cat([torch.randint(0, r, (B * L,), generator=g, device=dev, dtype=torch.int64) for r in rows])
    offsets = torch.arange(0, T * B * L + 1, L, dtype=torch.int64, device=dev)
    N = idx.numel()
    bits = int(np.log2(total_rows) + 1)
    grad = torch.randn(B, T * D, device=dev)
    wbytes = weights.element_size()
    U = int(torch.unique(torch.repeat_interleave(hsc[:-1], B * L) + idx).numel())

    fwd = lambda: ops.tbe_forward(weights, wo, Do, T * D, D, D % 8 == 0, idx, offsets, None, 0)
    srt_holder = {}

    def do_sort():
        srt_holder["s"] = ops.tbe_sort_indices(idx, offsets, hsc, bits, B, False)

    do_sort()
    optc = 1 if a.opt == "rowwise_adagrad" else 0
    bwd = lambda: ops.tbe_backward(grad, srt_holder["s"], weights, wo, Do, hsc, B, D, D % 8 == 0, offsets, None, 0,
                                   optc, mom, hsc[:-1].contiguous(), 0.01, 1e-8)
    t_f, t_fmin = timeit(fwd, a.iters)
    t_s, t_smin = timeit(do_sort, a.iters)
    t_b, t_bmin = timeit(bwd, a.iters)
    fwd_bytes = N * D * wbytes + B * T * D * 4 + N * 8 + (T * B + 1) * 8
    bwd_bytes = N * D * 4 + 2 * U * D * wbytes + (2 * U * 4 if optc else 0) + N * 8 + (T * B + 1) * 8
    print(f"T={T} B={B} D={D} L={L} N={N} U={U} dtype={a.dtype}")
    print(f"forward : {t_f*1e3:8.1f} us (min {t_fmin*1e3:.1f})  {fwd_bytes/t_f/1e6:8.1f} GB/s algorithmic")
    print(f"sort    : {t_s*1e3:8.1f} us (min {t_smin*1e3:.1f})")
    print(f"backward: {t_b*1e3:8.1f} us (min {t_bmin*1e3:.1f})  {bwd_bytes/t_b/1e6:8.1f} GB/s algorithmic")
    print(f"fwd+bwd : {(t_f+t_b)*1e3:8.1f} us  {(fwd_bytes+bwd_bytes)/(t_f+t_b)/1e6:8.1f} GB/s ; with sort {(t_f+t_b+t_s)*1e3:.1f} us")


if __name__ == "__main__":
    main()
