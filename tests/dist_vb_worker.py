"""Multi-rank worker (torchrun, also runs at world 1): VARIABLE BATCH SIZE PER RANK on the pooled
table-wise path - EmbeddingBagCollection sharded with EmbeddingBagCollectionSharder(variable_batch_size=True)
(KJTAllToAll with per-rank batch sizes -> permute_1D recat -> table-batched lookup at stride sum_r B_r ->
pooled all-to-all with ragged send splits).  Two steps with different batch sizes on every rank;
after each, outputs equal the unsharded model on my samples and every local table equals the table
of nn.EmbeddingBag + SGD trained on the union of the ranks' batches (gradients averaged over ranks,
comm_ops.GRADIENT_DIVISION).  Pattern: torchrec/distributed/tests/test_dist_data.py with
variable_batch_size=True, lifted to the sharded module."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))


class Sparse(torch.nn.Module):
    def __init__(self, ebc):
        super().__init__()
        self.ebc = ebc

    def forward(self, kjt):
        return self.ebc(kjt)


def main() -> None:
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29651")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig, KeyedJaggedTensor
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder, ShardedEmbeddingBagCollection
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology

    rows = [50, 7, 1001, 5, 300]
    dims = [16, 8, 16, 32, 8]
    keys = [f"f{i}" for i in range(len(rows))]
    L, lr = 4, 0.1

    def tables():
        return [EmbeddingBagConfig(name=f"t{i}", embedding_dim=d, num_embeddings=r, feature_names=[keys[i]])
                for i, (r, d) in enumerate(zip(rows, dims))]

    torch.manual_seed(13)
    glob = Sparse(EmbeddingBagCollection(tables(), device=dev))
    local = Sparse(EmbeddingBagCollection(tables(), device=torch.device("meta")))
    sharders = [EmbeddingBagCollectionSharder(
        fused_params={"learning_rate": lr, "optimizer": EmbOptimType.EXACT_SGD}, variable_batch_size=True)]
    plan = EmbeddingShardingPlanner(Topology(world), {f"t{i}": ParameterConstraints(sharding_types=["table_wise"])
                                                      for i in range(len(rows))}).plan(local, sharders)
    dmp = DistributedModelParallel(local, env=ShardingEnv.from_process_group(dist.group.WORLD), device=dev,
                                   plan=plan, sharders=sharders)
    assert isinstance(dmp.module.ebc, ShardedEmbeddingBagCollection)
    ebc_plan = plan.plan["ebc"]
    from torchrec.distributed.embedding_kernel import local_tensor
    gsd = glob.state_dict()
    for k, v in dmp.state_dict().items():
        local_tensor(v).copy_(gsd[k])
    opt_g = torch.optim.SGD(glob.parameters(), lr=lr)

    def batch(step: int, r: int):
        B = [[5, 9, 3, 6], [7, 2, 8, 1]][step][r % 4] + (r // 4)
        g = torch.Generator().manual_seed(1000 * step + r)
        lengths = torch.randint(0, L + 1, (len(keys), B), generator=g)
        lengths[:, 0] = 0
        values = torch.cat([torch.randint(0, rows[i], (int(lengths[i].sum()),), generator=g)
                            for i in range(len(keys))])
        return KeyedJaggedTensor(keys=keys, values=values.to(dev), lengths=lengths.view(-1).int().to(dev), stride=B)

    for step in range(2):
        kjt = batch(step, rank)
        torch.manual_seed(50 + step)
        coef = {k: torch.randn(d, device=dev) for k, d in zip(keys, dims)}
        out_l = dmp(kjt)
        out_l = out_l.wait() if hasattr(out_l, "wait") else out_l
        out_g = glob(kjt)
        # the sharded module lists its keys in sharding order (owner rank by owner rank): compare per key
        dl, dg = out_l.to_dict(), out_g.to_dict()
        assert sorted(dl) == sorted(dg) == sorted(keys), (out_l.keys(), out_g.keys())
        for k, d in zip(keys, dims):
            assert dl[k].shape == dg[k].shape == (kjt.stride(), d), (k, dl[k].shape, dg[k].shape)
            torch.testing.assert_close(dl[k], dg[k], rtol=1e-5, atol=1e-6)
        # one step; the sharded backward divides by W, so the global model averages over ranks
        opt_g.zero_grad()
        sum((dg[k] * coef[k]).sum() for k in keys).backward()
        for p in glob.parameters():
            if p.grad is None:
                p.grad = torch.zeros_like(p)
            dist.all_reduce(p.grad)
            p.grad.div_(world)
        opt_g.step()
        sum((dl[k] * coef[k]).sum() for k in keys).backward()
        gsd = glob.state_dict()
        bad = []
        for k, v in dmp.state_dict().items():
            assert ebc_plan[k.split(".")[-2]].ranks[0] == rank
            err = float((local_tensor(v) - gsd[k]).abs().max())
            if err > 1e-5:
                bad.append(f"{k}: max abs diff {err:.3e}")
        assert not bad, f"rank {rank} step {step}: tables differ: " + "; ".join(bad)
    dist.barrier()
    if rank == 0:
        print(f"VB PARITY OK world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
