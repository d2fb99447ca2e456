"""cProfile of the host side of a DLRM training step (N=1): where the Python time between launches goes."""
import cProfile
import os
import pstats
import sys
import time

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "torchrec-oldfork_b200")]
import bench  # noqa: E402


def main():
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29578")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    sys.argv = [sys.argv[0]]
    args = bench.parse_args()
    torch.backends.cuda.matmul.allow_tf32 = True
    dmp, opt, rows = bench.build_gpu_model(args, 0, 1, dev, dist.group.WORLD)
    batches = [b.to(dev) for b in bench.make_batches(rows, args.batch, 0, 2)]

    def step(b):
        opt.zero_grad(set_to_none=True)
        loss, _ = dmp(b)
        loss.backward()
        opt.step()

    for i in range(5):
        step(batches[i % 2])
    torch.cuda.synchronize()
    # host time per step with the GPU queue never drained
    t0 = time.perf_counter()
    for i in range(20):
        step(batches[i % 2])
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    print(f"host enqueue time/step {1e3 * (t1 - t0) / 20:.2f} ms; wall/step incl. drain {1e3 * (t2 - t0) / 20:.2f} ms")
    pr = cProfile.Profile()
    pr.enable()
    for i in range(20):
        step(batches[i % 2])
    pr.disable()
    torch.cuda.synchronize()
    st = pstats.Stats(pr)
    st.sort_stats("tottime").print_stats(28)
    st.sort_stats("cumulative").print_stats(45)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
