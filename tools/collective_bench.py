"""Bus bandwidth of the sharded data movement at the cfg3 shapes (run under torchrun, W = 2/4/8):

  torchrun --nnodes=1 --nproc-per-node W --master-addr 127.0.0.1 tools/collective_bench.py

NCCL legs (what the reference runs, comm_ops.py): pooled all_to_all_single, reduce_scatter_tensor,
all_gather_into_tensor and the two-phase KJT all-to-all (lengths + values) at local batch 65536.
Peer legs (this repo): b200_tbe_forward_peer storing pooled rows straight into the peers' buffers
and b200_tbe_backward_peer gathering gradient rows from them, against the same kernels with all
destinations local.  busbw = bytes_sent_per_rank * (W-1)/W / t  (nccl-tests convention for
all-to-all / reduce-scatter / all-gather).  CUDA events, max over ranks, median of `iters`.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))


def timed(fn, iters=20, warmup=3):
    for _ in range(warmup):
        fn()
    ts = []
    for _ in range(iters):
        dist.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        t = torch.tensor([s.elapsed_time(e)], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t.item()))
    return float(np.median(ts))


def main():
    rank, W = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from b200_sparse import ops
    B_local, D = 65536, 128
    B = B_local * W
    out = []

    def rep(name, nbytes_sent, ms, note=""):
        if rank == 0:
            row = {"op": name, "world": W, "bytes_per_rank_MB": round(nbytes_sent / 1e6, 1), "us": round(ms * 1e3, 1),
                   "busbw_GBps": round(nbytes_sent * (W - 1) / W / ms / 1e6, 1), "note": note}
            out.append(row)
            print(json.dumps(row), flush=True)

    # ---- NCCL: pooled all-to-all of a table-wise group with F_local features per rank
    for F_local in (max(1, 20 // W), 26 // W + 1):
        src = torch.randn(B, F_local * D, device=dev)
        dst = torch.empty(B_local, W * F_local * D, device=dev)
        ms = timed(lambda: dist.all_to_all_single(dst.view(-1), src.view(-1)))
        rep(f"nccl all_to_all_single pooled [B_global={B}, {F_local}x{D}]", src.numel() * 4, ms)
    # ---- NCCL: reduce-scatter / all-gather of the row-wise group (6 features at Criteo)
    F_rw = 6
    src = torch.randn(B, F_rw * D, device=dev)
    dst = torch.empty(B_local, F_rw * D, device=dev)
    ms = timed(lambda: dist.reduce_scatter_tensor(dst, src))
    rep(f"nccl reduce_scatter_tensor [B_global={B}, {F_rw}x{D}]", src.numel() * 4, ms)
    ms = timed(lambda: dist.all_gather_into_tensor(src, dst))
    rep(f"nccl all_gather_into_tensor (backward of the above)", src.numel() * 4, ms)
    # ---- NCCL: KJT all-to-all payloads (lengths int32 + values int64, 26 features, 1 id per bag)
    lens = torch.ones(26 * B_local, dtype=torch.int32, device=dev)
    vals = torch.randint(0, 1 << 20, (26 * B_local,), device=dev)
    lo, vo = torch.empty_like(lens), torch.empty_like(vals)

    def kjt_a2a():
        dist.all_to_all_single(lo, lens)
        dist.all_to_all_single(vo, vals)

    ms = timed(kjt_a2a)
    rep("nccl KJT all-to-all (lengths + values, 26 x 65536 ids)", lens.numel() * 4 + vals.numel() * 8, ms)

    # ---- peer-memory legs: forward scatter / backward gather over NVLink
    import torch.distributed._symmetric_memory as symm_mem
    gname = dist.group.WORLD.group_name
    try:
        symm_mem.enable_symm_mem_for_group(gname)
    except Exception:
        pass
    T = max(1, 20 // W)
    rows = [1_000_000] * T
    hsc = torch.tensor(np.concatenate([[0], np.cumsum(rows)]), dtype=torch.int64, device=dev)
    wo = (hsc[:-1] * D).contiguous()
    Do = torch.arange(0, (T + 1) * D, D, dtype=torch.int32, device=dev)
    weights = torch.empty(sum(rows) * D, device=dev).uniform_(-0.01, 0.01)
    mom = torch.zeros(sum(rows), device=dev)
    idx = torch.cat([torch.randint(0, r, (B,), device=dev) for r in rows])
    offs = torch.arange(0, T * B + 1, dtype=torch.int64, device=dev)
    total_D = W * T * D
    obuf = symm_mem.empty((B_local, total_D), dtype=torch.float32, device=dev)
    h = symm_mem.rendezvous(obuf, gname)
    gbuf = symm_mem.empty((B_local, total_D), dtype=torch.float32, device=dev)
    hg = symm_mem.rendezvous(gbuf, gname)
    gbuf.normal_()
    peer_bases = torch.tensor([int(p) for p in h.buffer_ptrs], dtype=torch.int64, device=dev)
    self_bases = torch.tensor([int(h.buffer_ptrs[rank])] * W, dtype=torch.int64, device=dev)
    gpeer = torch.tensor([int(p) for p in hg.buffer_ptrs], dtype=torch.int64, device=dev)
    gself = torch.tensor([int(hg.buffer_ptrs[rank])] * W, dtype=torch.int64, device=dev)
    col = (torch.arange(T, dtype=torch.int32, device=dev) * D + rank * T * D).contiguous()
    sent = B * T * D * 4

    def fwd(bases):
        ops.tbe_forward_peer(weights, wo, Do, T * D, D, True, idx, offs, None, 0, bases, B_local, total_D, col)

    h.barrier()
    ms_p = timed(lambda: fwd(peer_bases))
    h.barrier()
    ms_l = timed(lambda: fwd(self_bases))
    rep(f"peer forward: lookup + scatter to {W} ranks ({T} tables, B_global={B})", sent, ms_p,
        f"same kernel with all-local destinations: {ms_l * 1e3:.1f} us; the NCCL path adds an all_to_all_single "
        "and a re-layout copy after the lookup")
    srt = ops.tbe_sort_indices(idx, offs, hsc, int(np.log2(sum(rows)) + 1), B, False)
    mo = hsc[:-1].contiguous()

    def bwd(bases):
        ops.tbe_backward_peer(bases, B_local, total_D, col, 1.0 / W, srt, weights, wo, Do, hsc, B, D, True, offs,
                              None, 0, 1, mom, mo, 0.01, 1e-8)

    h.barrier()
    ms_p = timed(lambda: bwd(gpeer))
    h.barrier()
    ms_l = timed(lambda: bwd(gself))
    rep(f"peer backward: gradient gather from {W} ranks + fused rowwise Adagrad", sent, ms_p,
        f"same kernel with all-local sources: {ms_l * 1e3:.1f} us")
    if rank == 0:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", f"collectives_w{W}.json"), "w") as f:
            json.dump(out, f, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
