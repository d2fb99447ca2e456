"""Times the index transpose (b200_tbe_sort_indices) alone at the cfg2 shape; prints microseconds."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "torchrec-oldfork_b200")]
from b200_sparse import ops  # noqa: E402

MAXB = 0
ROWS = [40000000, 39060, 17295, 7424, 20265, 3, 7122, 1543, 63, 40000000, 3067956, 405282, 10, 2209,
        11938, 155, 4, 976, 14, 40000000, 40000000, 40000000, 590152, 12973, 108, 36]


def main():
    B = int(os.environ.get("B", 65536))
    zipf = os.environ.get("ZIPF") == "1"
    dev = torch.device("cuda")
    rng = np.random.default_rng(0)
    T = len(ROWS)
    hsc = np.concatenate([[0], np.cumsum(ROWS)]).astype(np.int64)
    if zipf:
        ids = [((rng.zipf(1.05, size=B) - 1) % r).astype(np.int64) for r in ROWS]
    else:
        ids = [rng.integers(0, r, size=B).astype(np.int64) for r in ROWS]
    idx = torch.from_numpy(np.concatenate(ids)).to(dev)
    off = torch.arange(T * B + 1, dtype=torch.int64, device=dev)
    d_hsc = torch.from_numpy(hsc).to(dev)
    bits = int(np.log2(hsc[-1]) + 1)
    global MAXB
    MAXB = max((r - 1).bit_length() for r in ROWS) if os.environ.get('HINT', '1') == '1' else 0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    for _ in range(5):
        ops.tbe_sort_indices(idx, off, d_hsc, bits, B, False, MAXB)
    ts = []
    for _ in range(30):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.tbe_sort_indices(idx, off, d_hsc, bits, B, False, MAXB)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    print(f"tbe_sort B={B} zipf={zipf} N={idx.numel()}: median {np.median(ts):.1f} us, min {min(ts):.1f} us")
    if os.environ.get("PROFILE") == "1":
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(5):
                flush.zero_()
                ops.tbe_sort_indices(idx, off, d_hsc, bits, B, False, MAXB)
            torch.cuda.synchronize()
        ev = sorted((e for e in prof.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
        last = [e for e in ev]
        # the last call: kernels after the last memset-sized fill
        t_prev = None
        rows = []
        for e in ev:
            rows.append((e.time_range.start, e.time_range.end, e.name[:60]))
        fills = [i for i, r in enumerate(rows) if "FillFunctor" in r[2] or "fill" in r[2].lower()]
        start = fills[-1] + 1 if fills else 0
        t0 = rows[start][0]
        for a, b, n in rows[start:]:
            print(f"{a - t0:9.1f} {b - a:8.1f}  {n}")


if __name__ == "__main__":
    main()
