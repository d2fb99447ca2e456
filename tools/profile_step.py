"""torch.profiler view of a few DLRM training steps under torchrun (diagnostic; not a bench).
usage: torchrun --nproc-per-node N tools/profile_step.py"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "torchrec-oldfork_b200")]
import bench  # noqa: E402


def main():
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    lr = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29577")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    sys.argv = [sys.argv[0]]
    args = bench.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = True
    dmp, opt, rows = bench.build_gpu_model(args, rank, world, dev, dist.group.WORLD)
    host = bench.make_batches(rows, args.batch, rank, 2)
    for hb in host:
        hb.sparse_features.length_per_key()
    batches = [b.to(dev) for b in host]

    import itertools
    from torchrec.distributed.train_pipeline import TrainPipelineSparseDist
    use_pipe = os.environ.get("PIPE", "1") == "1"
    pipe = TrainPipelineSparseDist(dmp, opt, dev)
    it = itertools.cycle(batches)

    def step(b):
        if use_pipe:
            pipe.progress(it)
            return
        opt.zero_grad(set_to_none=True)
        loss, _ = dmp(b)
        loss.backward()
        opt.step()

    for i in range(5):
        step(batches[i % 2])
    torch.cuda.synchronize()
    if os.environ.get("CPROFILE") == "1":
        import cProfile
        import pstats
        import time
        t0 = time.perf_counter()
        for i in range(20):
            step(batches[i % 2])
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        pr = cProfile.Profile()
        pr.enable()
        for i in range(20):
            step(batches[i % 2])
        pr.disable()
        torch.cuda.synchronize()
        if rank == 0:
            print(f"host enqueue/step {1e3 * (t1 - t0) / 20:.2f} ms, wall/step {1e3 * (t2 - t0) / 20:.2f} ms")
            st = pstats.Stats(pr)
            st.sort_stats("tottime").print_stats(28)
        dist.destroy_process_group()
        return
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for i in range(4):
            step(batches[i % 2])
        torch.cuda.synchronize()
    if os.environ.get("TRACE") == "1":
        prof.export_chrome_trace(os.path.join(ROOT, "gpurun_out", f"trace_n{world}_r{rank}.json"))
    if os.environ.get("TIMELINE", "1") == "1" and rank in (0, world - 1):
        # compact per-rank device timeline: every kernel / memcpy of the 4 profiled steps with stream, start, duration
        import csv
        import gzip
        import json
        import tempfile
        with tempfile.TemporaryDirectory() as td:
            tp = os.path.join(td, "t.json")
            prof.export_chrome_trace(tp)
            ev = json.load(open(tp))["traceEvents"]
        launch = {e["args"]["correlation"]: e["ts"] for e in ev
                  if e.get("cat") == "cuda_runtime" and "correlation" in e.get("args", {})}
        rows = [(e["ts"], e.get("dur", 0), e.get("args", {}).get("stream", -1), e.get("cat"), e["name"][:120],
                 e["ts"] - launch.get(e.get("args", {}).get("correlation"), e["ts"]))
                for e in ev if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
        rows.sort()
        t0 = rows[0][0] if rows else 0
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with gzip.open(os.path.join(ROOT, "gpurun_out", f"timeline_n{world}_r{rank}.csv.gz"), "wt", newline="") as f:
            w = csv.writer(f)
            w.writerow(["start_us", "dur_us", "stream", "cat", "name", "queued_us"])
            for ts, dur, st, cat, name, q in rows:
                w.writerow([round(ts - t0, 1), round(dur, 1), st, cat, name, round(q, 1)])
    if rank == 0:
        print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=32, max_name_column_width=60))
        print(prof.key_averages().table(sort_by="self_cpu_time_total", row_limit=18, max_name_column_width=60))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
