"""Per-row micro-benchmarks of SURVEY §8's hot path through the C ABI wrappers (one B200).

Covers cfg2 (Criteo-1TB shape, fp32), a Zipf id variant, cfg4 (512 tables x 1M rows, fp16, pooling
sweep), cfg5 (int8 rowwise-quantized lookup, batch 4096) and the jagged integer ops at cfg2 sizes.
Each kernel is timed alone with CUDA events on the current stream after warm-up, with a 256 MB
write between iterations to flush the 126 MB L2; the median is reported next to the algorithmic
bytes of SURVEY §8(d) and the fraction of the measured copy bandwidth (MEASURED_PEAKS.json).

  python tools/microbench_all.py [--sections cfg2,zipf,cfg4,cfg5,jagged,interaction] [--out FILE.md]
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))
from b200_sparse import cabi, ops  # noqa: E402

CRITEO_ROWS = [40000000, 39060, 17295, 7424, 20265, 3, 7122, 1543, 63, 40000000, 3067956, 405282, 10, 2209,
               11938, 155, 4, 976, 14, 40000000, 40000000, 40000000, 590152, 12973, 108, 36]
DEV = torch.device("cuda:0")
ROWS_OUT = []
_FLUSH = None


def peak_gbs() -> float:
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6650.0


PEAK = peak_gbs()


def timeit(fn, iters=15, warmup=3, flush=True):
    global _FLUSH
    if _FLUSH is None:
        _FLUSH = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        if flush:
            _FLUSH.fill_(1)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        e.synchronize()
        ts.append(s.elapsed_time(e))
    return float(np.median(ts))


def report(section, name, shape, ms, nbytes):
    gbs = nbytes / ms / 1e6 if nbytes else None
    row = {"section": section, "kernel": name, "shape": shape, "us": round(ms * 1e3, 1),
           "algorithmic_MB": round(nbytes / 1e6, 2) if nbytes else None,
           "GBps": round(gbs, 1) if gbs else None, "frac_of_measured_peak": round(gbs / PEAK, 3) if gbs else None}
    ROWS_OUT.append(row)
    print(json.dumps(row), flush=True)


def tbe_setup(rows, D, wdt, B, L, zipf=False, seed=1):
    T = len(rows)
    total_rows = sum(rows)
    weights = torch.empty(total_rows * D, dtype=wdt, device=DEV).uniform_(-0.01, 0.01)
    hsc = torch.tensor(np.concatenate([[0], np.cumsum(rows)]), dtype=torch.int64, device=DEV)
    wo = (hsc[:-1] * D).contiguous()
    Do = torch.arange(0, (T + 1) * D, D, dtype=torch.int32, device=DEV)
    mom = torch.zeros(total_rows, dtype=torch.float32, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(seed)
    if zipf:
        # Zipf(1.05) over each table as fbgemm's bench does (generate_requests alpha): ranks drawn on the host
        rng = np.random.default_rng(seed)
        parts = []
        for r in rows:
            z = rng.zipf(1.05, size=B * L)
            parts.append(torch.from_numpy(((z - 1) % r).astype(np.int64)))
        idx = torch.cat(parts).to(DEV)
    else:
        idx = torch.cat([torch.randint(0, r, (B * L,), generator=g, device=DEV, dtype=torch.int64) for r in rows])
    offsets = torch.arange(0, T * B * L + 1, L, dtype=torch.int64, device=DEV)
    return dict(T=T, B=B, D=D, L=L, weights=weights, hsc=hsc, wo=wo, Do=Do, mom=mom, idx=idx, offsets=offsets,
                bits=int(np.log2(total_rows) + 1))


def tbe_bench(section, s, label, opt="rowwise_adagrad", iters=15):
    T, B, D, L = s["T"], s["B"], s["D"], s["L"]
    w, idx, offsets = s["weights"], s["idx"], s["offsets"]
    N = idx.numel()
    wb = w.element_size()
    keys = torch.repeat_interleave(s["hsc"][:-1], B * L) + idx
    U = int(torch.unique(keys).numel())
    del keys
    grad = torch.randn(B, T * D, device=DEV)
    holder = {}

    def do_sort():
        holder["s"] = ops.tbe_sort_indices(idx, offsets, s["hsc"], s["bits"], B, False)

    do_sort()
    optc = 1 if opt == "rowwise_adagrad" else 0
    mo = s["hsc"][:-1].contiguous()
    fwd = lambda: ops.tbe_forward(w, s["wo"], s["Do"], T * D, D, D % 8 == 0, idx, offsets, None, 0)  # noqa: E731
    bwd = lambda: ops.tbe_backward(grad, holder["s"], w, s["wo"], s["Do"], s["hsc"], B, D, D % 8 == 0, offsets,  # noqa: E731
                                   None, 0, optc, s["mom"], mo, 0.01, 1e-8)
    fwd_bytes = N * D * wb + B * T * D * 4 + N * 8 + (T * B + 1) * 8
    bwd_bytes = N * D * 4 + 2 * U * D * wb + (2 * U * 4 if optc else 0) + N * 8 + (T * B + 1) * 8
    shape = f"{label} T={T} B={B} D={D} L={L} N={N} U={U} w={wb}B"
    t_f = timeit(fwd, iters)
    report(section, "tbe_forward", shape, t_f, fwd_bytes)
    t_s = timeit(do_sort, iters)
    report(section, "tbe_sort_indices (linearize + segmented LSD radix sort)", shape, t_s, 0)
    t_b = timeit(bwd, iters)
    report(section, "tbe_backward (+rowwise Adagrad)", shape, t_b, bwd_bytes)
    report(section, "tbe fwd+bwd", shape, t_f + t_b, fwd_bytes + bwd_bytes)


def sec_cfg2():
    s = tbe_setup(CRITEO_ROWS, 128, torch.float32, 65536, 1)
    tbe_bench("cfg2", s, "criteo-1TB uniform")


def sec_zipf():
    s = tbe_setup(CRITEO_ROWS, 128, torch.float32, 65536, 1, zipf=True)
    tbe_bench("cfg2-zipf", s, "criteo-1TB zipf(1.05)")


def sec_cfg4(quick=False, only=""):
    rows1 = 1_000_000
    if only:  # "T,D,B,L": one point of the sweep (for ncu captures)
        T, D, B, L = (int(x) for x in only.split(","))
        s = tbe_setup([rows1] * T, D, torch.float16, B, L)
        s["idx"] = s["idx"].view(T * B, L).sort(dim=1).values.reshape(-1).contiguous()
        tbe_bench("cfg4", s, "pooling sweep fp16", iters=3)
        return
    for (T, D) in ((512, 128), (512, 32), (256, 256), (512, 64)):
        for B in (512, 4096):
            for L in ((1, 10, 50) if B == 4096 else (1, 10, 50, 200)):
                if quick and (L not in (1, 50) or D not in (128, 32)):
                    continue
                s = tbe_setup([rows1] * T, D, torch.float16, B, L)
                # ids sorted within each bag, as fbgemm's bench (generate_requests) does
                s["idx"] = s["idx"].view(T * B, L).sort(dim=1).values.reshape(-1).contiguous()
                tbe_bench("cfg4", s, "pooling sweep fp16", iters=8)
                del s
                torch.cuda.empty_cache()


def sec_cfg5():
    B, D = 4096, 128
    rows = [min(r, 4_000_000) for r in CRITEO_ROWS]
    T = len(rows)
    g = torch.Generator(device=DEV).manual_seed(3)
    idxs = [torch.randint(0, r, (B,), generator=g, device=DEV, dtype=torch.int32) for r in rows]
    from fbgemm_gpu.split_table_batched_embeddings_ops import (EmbeddingLocation, IntNBitTableBatchedEmbeddingBagsCodegen,
                                                                PoolingMode)
    from fbgemm_gpu.split_embedding_configs import SparseType
    # (a) the reference's arrangement: one T=1 module per table (torchrec/quant/embedding_modules.py:190)
    mods = []
    for t, r in enumerate(rows):
        m = IntNBitTableBatchedEmbeddingBagsCodegen([(f"t{t}", r, D, SparseType.INT8, EmbeddingLocation.DEVICE)],
                                                    pooling_mode=PoolingMode.SUM, device=DEV)
        m.initialize_weights()
        mods.append(m)
    off1 = torch.arange(0, B + 1, dtype=torch.int32, device=DEV)

    def per_table():
        return [m(idxs[t], off1) for t, m in enumerate(mods)]

    nbytes = T * B * 144 + B * T * D * 2 + T * B * 4 + (T * B + 1) * 4
    shape = f"int8 T={T} B={B} D={D} L=1 row=144B"
    report("cfg5", "int8 lookup, 26 separate T=1 modules (reference arrangement)", shape, timeit(per_table, 20), nbytes)
    # (b) one table-batched launch over all 26 tables
    mb = IntNBitTableBatchedEmbeddingBagsCodegen([(f"t{t}", r, D, SparseType.INT8, EmbeddingLocation.DEVICE)
                                                  for t, r in enumerate(rows)], pooling_mode=PoolingMode.SUM, device=DEV)
    mb.initialize_weights()
    idx_all = torch.cat(idxs)
    off_all = torch.arange(0, T * B + 1, dtype=torch.int32, device=DEV)
    report("cfg5", "int8 lookup, one table-batched launch", shape, timeit(lambda: mb(idx_all, off_all), 20), nbytes)
    # (c) the same through a CUDA graph of the 26 launches
    sgr = torch.cuda.Stream()
    with torch.cuda.stream(sgr):
        per_table()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr, stream=sgr):
            outs = per_table()  # noqa: F841
    torch.cuda.synchronize()
    report("cfg5", "int8 lookup, 26 modules replayed from one CUDA graph", shape, timeit(gr.replay, 20), nbytes)


def sec_jagged():
    T, B, W = 26, 65536, 8
    N = T * B
    g = torch.Generator(device=DEV).manual_seed(5)
    lengths = torch.ones(N, dtype=torch.int32, device=DEV)
    values = torch.cat([torch.randint(0, r, (B,), generator=g, device=DEV, dtype=torch.int64) for r in CRITEO_ROWS])
    shape = f"T={T} B={B} N={N}"
    report("jagged", "asynchronous_complete_cumsum (int32)", shape,
           timeit(lambda: ops.cumsum(lengths, cabi.CUMSUM_COMPLETE)), N * 4 + (N + 1) * 4)
    perm = torch.randperm(T, generator=torch.Generator().manual_seed(0)).to(torch.int32).to(DEV)
    l2d = lengths.view(T, B)
    report("jagged", "permute_2D_sparse_data (lengths + values)", shape,
           timeit(lambda: ops.permute_sparse_data(perm, l2d, values, None, N, True)), 2 * N * 8 + 2 * N * 4)
    wts = torch.rand(N, device=DEV)
    report("jagged", "permute_2D_sparse_data (+ weights)", shape,
           timeit(lambda: ops.permute_sparse_data(perm, l2d, values, wts, N, True)), 2 * N * 12 + 2 * N * 4)
    blk = torch.tensor([(r + W - 1) // W for r in CRITEO_ROWS], dtype=torch.int64, device=DEV)
    report("jagged", f"block_bucketize_sparse_features (W={W})", shape,
           timeit(lambda: ops.block_bucketize_sparse_features(lengths, values, False, False, blk, W)),
           2 * N * 8 + N * 4 + W * N * 4)
    report("jagged", f"block_bucketize_sparse_features (W={W}, pos + unbucketize)", shape,
           timeit(lambda: ops.block_bucketize_sparse_features(lengths, values, True, True, blk, W)),
           4 * N * 8 + N * 4 + W * N * 4)
    rows_t = torch.tensor(CRITEO_ROWS, dtype=torch.int64, device=DEV)
    offs = torch.arange(0, N + 1, dtype=torch.int64, device=DEV)
    warn = torch.zeros(1, dtype=torch.int64, device=DEV)
    v2 = values.clone()
    report("jagged", "bounds_check_indices (WARNING mode)", shape,
           timeit(lambda: ops.bounds_check_indices(rows_t, v2, offs, 1, warn)), N * 8 + (N + 1) * 8)
    report("jagged", "offsets_range", shape, timeit(lambda: ops.offsets_range(offs[:-1], N)), 2 * N * 8)


def sec_interaction():
    B, F, D = 65536, 26, 128
    dense = torch.randn(B, D, device=DEV)
    sparse = torch.randn(B, F, D, device=DEV)
    width = D + F * (F + 1) // 2
    go = torch.randn(B, width, device=DEV)
    fb = B * (F + 1) * D * 4 + B * width * 4
    shape = f"B={B} F={F} D={D}"
    report("interaction", "interaction_forward", shape, timeit(lambda: ops.interaction_forward(dense, sparse, True)), fb)
    report("interaction", "interaction_backward", shape,
           timeit(lambda: ops.interaction_backward(dense, sparse, go)), 2 * B * (F + 1) * D * 4 + B * width * 4)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sections", default="cfg2,zipf,jagged,interaction,cfg5,cfg4")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--out", default="")
    ap.add_argument("--cfg4-only", default="", help="T,D,B,L: run one point of the cfg4 sweep")
    a = ap.parse_args()
    secs = {"cfg2": sec_cfg2, "zipf": sec_zipf, "cfg4": lambda: sec_cfg4(a.quick, a.cfg4_only), "cfg5": sec_cfg5,
            "jagged": sec_jagged, "interaction": sec_interaction}
    for name in a.sections.split(","):
        try:
            secs[name]()
        except Exception as e:  # one section failing must not lose the others
            print(json.dumps({"section": name, "error": repr(e)}), flush=True)
        torch.cuda.empty_cache()
    if a.out:
        with open(a.out, "w") as f:
            f.write(f"peak = {PEAK} GB/s (MEASURED_PEAKS.json hbm_gbs); median of CUDA-event timings, L2 flushed "
                    "between iterations\n\n| section | kernel | shape | us | algorithmic MB | GB/s | of measured peak |\n"
                    "|---|---|---|---:|---:|---:|---:|\n")
            for r in ROWS_OUT:
                f.write(f"| {r['section']} | {r['kernel']} | {r['shape']} | {r['us']} | {r['algorithmic_MB'] or ''} | "
                        f"{r['GBps'] or ''} | {r['frac_of_measured_peak'] or ''} |\n")


if __name__ == "__main__":
    main()
