"""Drop-in check (torchrun, PYTHONPATH = baseline/_ref : torchrec-oldfork_b200/dropin): the REFERENCE's own
torchrec - its DistributedModelParallel, planner, ShardedEmbeddingBagCollection, KJTAllToAll /
PooledEmbeddingsAllToAll / ReduceScatter and BatchedFusedEmbeddingBag
(torchrec/distributed/batched_embedding_kernel.py:604-640, embedding_lookup.py:287) - runs unchanged over THIS
repo's `fbgemm_gpu` package and `torch.ops.fbgemm.*` kernels, and one fused-SGD step matches the reference's
unsharded EmbeddingBagCollection (nn.EmbeddingBag) on the global batch.  Nothing of this repo's torchrec mirror is
imported: the first assertion checks where `torchrec` came from."""
import os

import torch
import torch.distributed as dist


def main() -> None:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    import fbgemm_gpu
    import torchrec
    assert os.sep + os.path.join("baseline", "_ref", "torchrec") in torchrec.__file__, torchrec.__file__
    assert "torchrec-oldfork_b200" in os.path.realpath(fbgemm_gpu.__file__), fbgemm_gpu.__file__
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec.distributed.batched_embedding_kernel import BatchedFusedEmbeddingBag
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder, ShardedEmbeddingBagCollection
    from torchrec.distributed.model_parallel import DistributedModelParallel
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    from torchrec.distributed.types import ShardingEnv
    from torchrec.modules.embedding_configs import EmbeddingBagConfig
    from torchrec.modules.embedding_modules import EmbeddingBagCollection
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

    rows = [50, 1001, 7, 300]
    sharding = ["table_wise", "row_wise", "row_wise", "table_wise"]
    D, B, L, lr = 32, 16, 2, 0.1
    keys = [f"f{i}" for i in range(len(rows))]

    def tables():
        return [EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r, feature_names=[keys[i]])
                for i, r in enumerate(rows)]

    class Model(torch.nn.Module):  # sparse arch + a dense head, so that DDP has something to wrap
        def __init__(self, ebc):
            super().__init__()
            self.ebc = ebc
            self.head = torch.nn.Linear(D * len(rows), 1, device=dev)

        def forward(self, kjt):
            kt = self.ebc(kjt)
            d = kt.to_dict()
            return self.head(torch.cat([d[k] for k in keys], dim=1))

    torch.manual_seed(3)
    glob = Model(EmbeddingBagCollection(tables=tables(), device=dev))
    torch.manual_seed(3)
    local = Model(EmbeddingBagCollection(tables=tables(), device=torch.device("meta")))
    local.head.load_state_dict(glob.head.state_dict())
    sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": lr, "optimizer": EmbOptimType.EXACT_SGD})]
    constraints = {f"t{i}": ParameterConstraints(sharding_types=[s], compute_kernels=["batched_fused"])
                   for i, s in enumerate(sharding)}
    pg = dist.group.WORLD
    plan = EmbeddingShardingPlanner(topology=Topology(world_size=world, compute_device="cuda"),
                                    constraints=constraints).collective_plan(local, sharders, pg)
    dmp = DistributedModelParallel(local, env=ShardingEnv.from_process_group(pg), device=dev, plan=plan,
                                   sharders=sharders)
    sebc = [m for m in dmp.modules() if isinstance(m, ShardedEmbeddingBagCollection)]
    fused = [m for lookup in sebc[0]._lookups for m in lookup.modules() if isinstance(m, BatchedFusedEmbeddingBag)]
    assert len(sebc) == 1 and len(fused) >= 1, (len(sebc), len(fused))
    from fbgemm_gpu.split_table_batched_embeddings_ops import SplitTableBatchedEmbeddingBagsCodegen
    assert all(isinstance(m.emb_module, SplitTableBatchedEmbeddingBagsCodegen) for m in fused)

    # same weights on both sides: copy the unsharded tables into the live shards of the fused modules (views
    # returned by split_embedding_weights(); each table's shard offsets come from the reference's own config)
    with torch.no_grad():
        for m in fused:
            for cfg, w in zip(m.config.embedding_tables, m.emb_module.split_embedding_weights()):
                full = glob.ebc.embedding_bags[cfg.name].weight
                o = cfg.local_metadata.shard_offsets if cfg.local_metadata is not None else [0, 0]
                w.copy_(full[o[0]:o[0] + w.shape[0], o[1]:o[1] + w.shape[1]])

    g = torch.Generator().manual_seed(100 + rank)
    lengths = torch.full((len(rows) * B,), L, dtype=torch.int32)
    values = torch.cat([torch.randint(0, r, (B * L,), generator=g) for r in rows])
    kjt = KeyedJaggedTensor(keys=keys, values=values.to(dev), lengths=lengths.to(dev))
    vals_all = [torch.empty_like(kjt.values()) for _ in range(world)]
    dist.all_gather(vals_all, kjt.values())
    gvalues = torch.cat([v.view(len(rows), B * L) for v in vals_all], dim=1).reshape(-1)
    gkjt = KeyedJaggedTensor(keys=keys, values=gvalues,
                             lengths=torch.full((len(rows) * B * world,), L, dtype=torch.int32, device=dev))

    opt_g = torch.optim.SGD(glob.parameters(), lr=lr)
    opt_l = torch.optim.SGD([p for p in dmp.parameters() if p.requires_grad], lr=lr)
    for _ in range(2):
        opt_g.zero_grad()
        (glob(gkjt).sum() / world).backward()
        opt_g.step()
        opt_l.zero_grad()
        dmp(kjt).sum().backward()
        opt_l.step()
    with torch.no_grad():
        pred_g = glob(gkjt)[rank * B:(rank + 1) * B]
        pred_l = dmp(kjt)
    torch.testing.assert_close(pred_l, pred_g, rtol=1e-4, atol=1e-4)
    dist.barrier()
    if rank == 0:
        print(f"DROPIN OK world={world}: reference torchrec {torchrec.__file__} over {fbgemm_gpu.__file__}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
