"""Multi-rank worker (torchrun): sharded-vs-unsharded one-step parity on real GPUs, the pattern of
torchrec/distributed/test_utils/test_model_parallel_base.py:144-290.  Every rank builds the same
unsharded DLRM (nn.EmbeddingBag tables, plain SGD) and a DistributedModelParallel DLRM with a mix of
table-wise and row-wise tables and the fused EXACT_SGD; after SIX steps on fresh per-rank batches (B200_PARITY_STEPS) the
sharded predictions of every step must equal the unsharded model's predictions on the same samples, and every
local weight shard must equal the corresponding slice of the unsharded weights."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))


def main() -> None:
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig, KeyedJaggedTensor
    from torchrec.datasets.random import RandomRecDataset
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    from torchrec.distributed.sharding.rw_sharding import row_wise_shard_sizes
    from torchrec.models.dlrm import DLRM
    torch.backends.cuda.matmul.allow_tf32 = False
    rows = [50, 7, 1001, 3, 5, 300]          # 5 rows over W ranks exercises short / empty row shards
    sharding = ["table_wise", "row_wise", "row_wise", "table_wise", "row_wise", "table_wise"]
    D, B, lr = 32, 32, 0.1  # D % 32 == 0: the tensor-core interaction kernels (and their gradient scatter) run
    keys = [f"f{i}" for i in range(len(rows))]

    def tables():
        return [EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r, feature_names=[keys[i]])
                for i, r in enumerate(rows)]

    # B200_PARITY_WEIGHTED=1: a weighted collection (per-id weights travel with the KJT through the input dist and
    # scale the pooled rows / their gradients), which the peer exchange serves as well
    weighted = os.environ.get("B200_PARITY_WEIGHTED", "0") == "1"
    torch.manual_seed(7)  # identical on every rank
    glob = DLRM(EmbeddingBagCollection(tables(), is_weighted=weighted, device=dev), 13, [32, D], [32, 16, 1],
                dense_device=dev)
    torch.manual_seed(7)
    local = DLRM(EmbeddingBagCollection(tables(), is_weighted=weighted, device=torch.device("meta")), 13, [32, D],
                 [32, 16, 1], dense_device=dev)
    sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": lr, "optimizer": EmbOptimType.EXACT_SGD})]
    plan = EmbeddingShardingPlanner(Topology(world), {f"t{i}": ParameterConstraints(sharding_types=[s])
                                                      for i, s in enumerate(sharding)}).plan(local, sharders)
    dmp = DistributedModelParallel(local, env=ShardingEnv.from_process_group(dist.group.WORLD), device=dev,
                                   plan=plan, sharders=sharders)
    ebc_plan = plan.plan["sparse_arch.embedding_bag_collection"]

    def shard_of(name: str, full: torch.Tensor):
        ps = ebc_plan[name]
        if ps.sharding_type == "table_wise":
            return full if ps.ranks[0] == rank else None
        sizes = row_wise_shard_sizes(full.shape[0], world)
        o = sum(sizes[:rank])
        return full[o:o + sizes[rank]]

    from torch.distributed._shard.sharded_tensor import ShardedTensor
    from torchrec.distributed.embedding_kernel import local_tensor
    gsd = glob.state_dict()
    lsd = dmp.state_dict()
    # checkpoint format: every sharded table is a ShardedTensor of the GLOBAL shape whose
    # metadata lists one shard per owning rank and whose local shard views the live table
    for i, (r, s) in enumerate(zip(rows, sharding)):
        key = f"sparse_arch.embedding_bag_collection.embedding_bags.t{i}.weight"
        if s == "table_wise" and ebc_plan[f"t{i}"].ranks[0] != rank:
            assert key not in lsd  # a table-wise table is listed by its owner only
            continue
        st = lsd[key]
        assert isinstance(st, ShardedTensor) and tuple(st.size()) == (r, D), (type(st), st.size())
        metas = st.metadata().shards_metadata
        assert len(metas) == (1 if s == "table_wise" else world), (s, len(metas))
        assert sum(m.shard_sizes[0] for m in metas) == r
        own = [m for m in metas if m.placement.rank() == rank]
        assert len(st.local_shards()) == len(own)
    # load: the unsharded weights cut to my shards, through DistributedModelParallel.load_state_dict
    to_load = {}
    for k, v in lsd.items():
        if "embedding_bags" in k:
            sh = shard_of(k.split(".")[-2], gsd[k])
            if sh is not None:
                to_load[k] = sh.clone()
        else:
            to_load[k] = gsd[k].clone()
    res = dmp.load_state_dict(to_load)
    assert not res.missing_keys and not res.unexpected_keys, res
    steps = int(os.environ.get("B200_PARITY_STEPS", "6"))
    F = len(keys)
    opt_g = torch.optim.SGD(glob.parameters(), lr=lr)
    opt_l = torch.optim.SGD(list(dmp.parameters()), lr=lr)

    def global_view(batch):
        """The global batch = concatenation of the per-rank batches along the batch dim."""
        dense_all = [torch.empty_like(batch.dense_features) for _ in range(world)]
        dist.all_gather(dense_all, batch.dense_features)
        vals_all = [torch.empty_like(batch.sparse_features.values()) for _ in range(world)]
        dist.all_gather(vals_all, batch.sparse_features.values())
        gvalues = torch.cat([v.view(F, B * 2) for v in vals_all], dim=1).reshape(-1)
        gweights = None
        if weighted:
            w_all = [torch.empty_like(batch.sparse_features.weights()) for _ in range(world)]
            dist.all_gather(w_all, batch.sparse_features.weights())
            gweights = torch.cat([v.view(F, B * 2) for v in w_all], dim=1).reshape(-1)
        gkjt = KeyedJaggedTensor(keys=keys, values=gvalues, weights=gweights,
                                 lengths=torch.full((F * B * world,), 2, dtype=torch.int32, device=dev),
                                 stride=B * world)
        return torch.cat(dense_all, dim=0), gkjt

    # SEVERAL steps on fresh batches: the peer exchange skips its first barrier from the second step on
    # ("quiesced"), persistent symmetric buffers are reused, and errors would compound in the weights
    for step in range(steps):
        batch = next(iter(RandomRecDataset(keys=keys, batch_size=B, hash_sizes=rows, ids_per_feature=2,
                                           num_dense=13, manual_seed=100 + rank + 1000 * step))).to(dev)
        if weighted:
            kj = batch.sparse_features
            gen = torch.Generator(device=dev).manual_seed(5000 + rank + 1000 * step)
            batch.sparse_features = KeyedJaggedTensor(
                keys=kj.keys(), values=kj.values(), lengths=kj.lengths(), stride=kj.stride(),
                weights=torch.rand(kj.values().numel(), device=dev, generator=gen) + 0.5)
        gdense, gkjt = global_view(batch)
        opt_g.zero_grad()
        # DDP averages dense grads over ranks and the pooled collectives divide by W (GRADIENT_DIVISION):
        # the equivalent unsharded objective is the mean over ranks of the per-rank sums.
        out_g = glob(gdense, gkjt)
        (torch.tanh(out_g).sum() / world).backward()  # a bounded objective: six steps at lr 0.1 must not diverge
        opt_g.step()
        opt_l.zero_grad()
        out_l = dmp(batch.dense_features, batch.sparse_features)
        torch.tanh(out_l).sum().backward()
        opt_l.step()
        # the training-time predictions of every step (the loss trajectory) agree as well
        assert bool(torch.isfinite(out_g).all()), f"step {step}: the unsharded model diverged"
        torch.testing.assert_close(out_l.detach(), out_g.detach()[rank * B:(rank + 1) * B], rtol=2e-4, atol=2e-4,
                                   msg=lambda m: f"rank {rank} step {step}: {m}")
    with torch.no_grad():
        pred_g = glob(gdense, gkjt)[rank * B:(rank + 1) * B]
        pred_l = dmp(batch.dense_features, batch.sparse_features)
    gsd = glob.state_dict()
    bad = []
    for k, v in dmp.state_dict().items():
        exp = shard_of(k.split(".")[-2], gsd[k]) if "embedding_bags" in k else gsd[k]
        v = local_tensor(v)
        err = float((v - exp).abs().max()) if v.numel() else 0.0
        if err > 3e-4:  # six steps of fp32 reassociation between nn.EmbeddingBag + SGD and the fused kernels
            bad.append(f"{k}: max abs diff {err:.3e}")
    assert not bad, f"rank {rank}: weights differ from the unsharded model after one step: " + "; ".join(bad)
    torch.testing.assert_close(pred_l, pred_g, rtol=3e-4, atol=3e-4)
    from torchrec.distributed.embeddingbag import ShardedEmbeddingBagCollection
    sebc = next(m for m in dmp.module.modules() if isinstance(m, ShardedEmbeddingBagCollection))
    mode = "peer" if sebc._peer is not None else "nccl"
    want = "nccl" if os.environ.get("B200_PEER_DIST", "1") == "0" or world == 1 else "peer"
    assert mode == want, f"exchange ran through {mode}, expected {want}"
    if mode == "peer":  # table-wise gradients: pushed by the interaction backward, or pulled by the owners
        pushed = sum(px.pushed_steps for px in sebc._peer._exchanges.values())
        want_push = os.environ.get("B200_PEER_PUSH", "1") != "0"
        assert (pushed > 0) == want_push, f"gradient push ran {pushed} times, expected push={want_push}"
        mode = "peer" if want_push else "peer_pull"
    dist.barrier()
    if rank == 0:
        print(f"DIST PARITY OK world={world} exchange={mode} steps={steps} weighted={int(weighted)}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
