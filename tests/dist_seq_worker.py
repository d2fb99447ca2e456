"""Multi-rank worker (torchrun, also runs at world 1): sharded-vs-unsharded one-step parity of the
SEQUENCE (non-pooled) embedding path - EmbeddingCollection under DistributedModelParallel with
table-wise and row-wise tables, jagged bags (empty ones included), a feature shared by two tables,
fused EXACT_SGD - against nn.Embedding + torch.optim.SGD on the concatenated global batch
(pattern of torchrec/distributed/tests/test_sequence_model_parallel.py)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))


class SeqModel(torch.nn.Module):
    def __init__(self, ec):
        super().__init__()
        self.ec = ec

    def forward(self, kjt):
        return self.ec(kjt)


def main() -> None:
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29641")
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import KeyedJaggedTensor
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embedding import EmbeddingCollectionSharder, ShardedEmbeddingCollection
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    from torchrec.distributed.sharding.rw_sharding import row_wise_shard_sizes
    from torchrec.modules.embedding_configs import EmbeddingConfig
    from torchrec.modules.embedding_modules import EmbeddingCollection

    rows = [50, 7, 1001, 5, 300]
    sharding = ["table_wise", "row_wise", "row_wise", "table_wise", "table_wise"]
    feats = [["f0"], ["f1"], ["f2", "f0"], ["f3"], ["f4"]]  # f0 is served by t0 and t2
    D, B, L, lr = 16, 24, 4, 0.1
    keys = ["f3", "f0", "f1", "f4", "f2"]  # input order differs from the table order
    rows_of = {"f0": 50, "f1": 7, "f2": 1001, "f3": 5, "f4": 300}

    def tables():
        return [EmbeddingConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r, feature_names=list(feats[i]))
                for i, r in enumerate(rows)]

    torch.manual_seed(11)
    glob = SeqModel(EmbeddingCollection(tables(), device=dev))
    local = SeqModel(EmbeddingCollection(tables(), device=torch.device("meta")))
    sharders = [EmbeddingCollectionSharder(fused_params={"learning_rate": lr, "optimizer": EmbOptimType.EXACT_SGD})]
    plan = EmbeddingShardingPlanner(Topology(world), {f"t{i}": ParameterConstraints(sharding_types=[s])
                                                      for i, s in enumerate(sharding)}).plan(local, sharders)
    dmp = DistributedModelParallel(local, env=ShardingEnv.from_process_group(dist.group.WORLD), device=dev,
                                   plan=plan, sharders=sharders)
    assert isinstance(dmp.module.ec, ShardedEmbeddingCollection)
    ec_plan = plan.plan["ec"]

    def shard_of(name: str, full: torch.Tensor):
        ps = ec_plan[name]
        if ps.sharding_type == "table_wise":
            return full if ps.ranks[0] == rank else None
        sizes = row_wise_shard_sizes(full.shape[0], world)
        o = sum(sizes[:rank])
        return full[o:o + sizes[rank]]

    from torchrec.distributed.embedding_kernel import local_tensor
    gsd = glob.state_dict()
    for k, v in dmp.state_dict().items():
        local_tensor(v).copy_(shard_of(k.split(".")[-2], gsd[k]))

    # per-rank jagged batch; min(rows) = 5 keeps f0's ids valid for both of its tables
    g = torch.Generator().manual_seed(100 + rank)
    lengths = torch.randint(0, L + 1, (len(keys), B), generator=g)
    lengths[:, 3] = 0
    values = torch.cat([torch.randint(0, min(rows_of[k], 50) if k == "f0" else rows_of[k],
                                      (int(lengths[i].sum()),), generator=g) for i, k in enumerate(keys)])
    kjt = KeyedJaggedTensor(keys=keys, values=values.to(dev), lengths=lengths.view(-1).int().to(dev), stride=B)

    out_l = dmp(kjt)
    out_l = out_l.wait() if hasattr(out_l, "wait") else out_l
    out_g = glob(kjt)  # the unsharded model on MY samples: same rows, before any update
    names = sorted(out_g.keys())
    assert sorted(out_l.keys()) == names, (sorted(out_l.keys()), names)
    for n in names:
        torch.testing.assert_close(out_l[n].values(), out_g[n].values(), rtol=1e-5, atol=1e-6)
        assert torch.equal(out_l[n].lengths().long(), out_g[n].lengths().long())

    # one training step: loss = sum_f <values_f, c_f>; the global model sees every rank's batch
    torch.manual_seed(5)
    coef = {n: torch.randn(D, device=dev) for n in names}

    def loss_of(out):
        return sum((out[n].values() * coef[n]).sum() for n in names)

    opt_g = torch.optim.SGD(glob.parameters(), lr=lr)
    opt_g.zero_grad()
    loss_of(glob(kjt)).backward()
    for p in glob.parameters():  # sum of the per-rank gradients = gradient of the global batch
        if p.grad is None:
            p.grad = torch.zeros_like(p)
        dist.all_reduce(p.grad)
    opt_g.step()
    res = dmp(kjt)
    res = res.wait() if hasattr(res, "wait") else res
    loss_of(res).backward()
    gsd = glob.state_dict()
    bad = []
    for k, v in dmp.state_dict().items():
        exp = shard_of(k.split(".")[-2], gsd[k])
        v = local_tensor(v)
        err = float((v - exp).abs().max()) if v.numel() else 0.0
        if err > 1e-5:
            bad.append(f"{k}: max abs diff {err:.3e}")
    assert not bad, f"rank {rank}: sequence shards differ after one step: " + "; ".join(bad)
    dist.barrier()
    if rank == 0:
        print(f"SEQ PARITY OK world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
