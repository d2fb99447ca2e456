"""2-GPU probe: NCCL all_reduce of the dense-gradient bucket (9.6 MB fp32) vs symmetric-memory all-reduce."""
import os
import sys

import torch
import torch.distributed as dist


def t(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def main():
    rank = int(os.environ["RANK"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    for numel in (2_400_000, 262_144, 16_777_216):
        x = torch.randn(numel, device=dev)
        us = t(lambda: dist.all_reduce(x))
        if rank == 0:
            print(f"nccl all_reduce {numel * 4 / 1e6:8.1f} MB: {us:8.1f} us  ({numel * 4 / us / 1e3:7.1f} GB/s algbw)  "
                  f"NCCL_PROTO={os.environ.get('NCCL_PROTO')} NCCL_ALGO={os.environ.get('NCCL_ALGO')}")
    try:
        import torch.distributed._symmetric_memory as symm_mem
        g = dist.group.WORLD.group_name
        for numel in (2_400_000, 262_144):
            y = symm_mem.empty(numel, dtype=torch.float32, device=dev)
            symm_mem.rendezvous(y, g)
            y.normal_()
            for op in ("one_shot_all_reduce", "two_shot_all_reduce_"):
                f = getattr(torch.ops.symm_mem, op)
                us = t(lambda: f(y, "sum", g))
                if rank == 0:
                    print(f"symm_mem {op} {numel * 4 / 1e6:8.1f} MB: {us:8.1f} us")
    except Exception as e:  # diagnostic only
        if rank == 0:
            print("symm_mem probe failed:", repr(e)[:300])
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
