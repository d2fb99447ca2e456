"""World-size-2 gloo worker (CPU): the N>1 host logic of the sharded data movement —
KJTAllToAll, bucketize + row-wise redistribution, pooled all-to-all and reduce-scatter with their
backward collectives — checked against what a single global view predicts.  Patterned on
torchrec/distributed/tests/test_dist_data.py (KJTAllToAllTest / PooledEmbeddingsAllToAllTest).
The jagged ops run on the test-only CPU kernels of tests/cpu_ops.py."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path[:0] = [os.path.join(ROOT, "torchrec-oldfork_b200"), ROOT, HERE]


def make_kjt(rank: int, keys, B: int, hash_size: int, weighted: bool = False):
    """Deterministic per-rank input every rank can regenerate."""
    from torchrec import KeyedJaggedTensor
    rng = np.random.default_rng(1000 + rank)
    lengths = rng.integers(0, 4, size=len(keys) * B).astype(np.int32)
    values = rng.integers(0, hash_size, size=int(lengths.sum())).astype(np.int64)
    w = torch.from_numpy(rng.random(int(lengths.sum())).astype(np.float32)) if weighted else None
    return KeyedJaggedTensor(keys=list(keys), values=torch.from_numpy(values), weights=w,
                             lengths=torch.from_numpy(lengths), stride=B)


def check_kjt_all2all(rank: int, world: int) -> None:
    from torchrec.distributed.dist_data import KJTAllToAll
    keys, splits, B = ["f0", "f1", "f2"], [2, 1], 3
    for weighted in (False, True):
        mine = make_kjt(rank, keys, B, 100, weighted)
        got = KJTAllToAll(dist.group.WORLD, splits)(mine).wait().wait()
        lo = sum(splits[:rank])
        my_keys = keys[lo:lo + splits[rank]]
        assert got.keys() == my_keys and got.stride() == B * world
        inputs = [make_kjt(r, keys, B, 100, weighted) for r in range(world)]
        for k in my_keys:
            exp_l = torch.cat([i[k].lengths() for i in inputs])
            exp_v = torch.cat([i[k].values() for i in inputs])
            assert torch.equal(got[k].lengths(), exp_l), (k, got[k].lengths(), exp_l)
            assert torch.equal(got[k].values(), exp_v)
            if weighted:
                assert torch.equal(got[k].weights(), torch.cat([i[k].weights() for i in inputs]))


def check_kjt_all2all_variable_batch(rank: int, world: int) -> None:
    """Each rank brings its own batch size (test_dist_data.py KJTAllToAllTest, variable_batch_size=True):
    key k of the output holds rank 0's samples, then rank 1's, ..."""
    from torchrec.distributed.dist_data import KJTAllToAll
    keys, splits = ["f0", "f1", "f2"], [2, 1]
    for batch_sizes in ([3, 5], [4, 1], [2, 2], [0, 3]):  # a rank may arrive with an empty batch
        for weighted in (False, True):
            mine = make_kjt(rank, keys, batch_sizes[rank], 100, weighted)
            got = KJTAllToAll(dist.group.WORLD, splits, variable_batch_size=True)(mine).wait().wait()
            lo = sum(splits[:rank])
            my_keys = keys[lo:lo + splits[rank]]
            assert got.keys() == my_keys and got.stride() == sum(batch_sizes), (got.stride(), batch_sizes)
            inputs = [make_kjt(r, keys, batch_sizes[r], 100, weighted) for r in range(world)]
            for k in my_keys:
                exp_l = torch.cat([i[k].lengths() for i in inputs])
                assert torch.equal(got[k].lengths(), exp_l), (k, got[k].lengths(), exp_l)
                assert torch.equal(got[k].values(), torch.cat([i[k].values() for i in inputs]))
                if weighted:
                    assert torch.equal(got[k].weights(), torch.cat([i[k].weights() for i in inputs]))


def check_pooled_all2all_variable_batch(rank: int, world: int) -> None:
    """PooledEmbeddingsAllToAllTest with batch_size_per_rank: rank r receives its own B_r samples."""
    from torchrec.distributed import comm_ops
    from torchrec.distributed.dist_data import PooledEmbeddingsAllToAll
    dims, batch_sizes = [8, 4], [3, 5]
    B_global, offs = sum(batch_sizes), [0, 3, 8]

    def inp(r):
        g = torch.Generator().manual_seed(150 + r)
        return torch.randn(B_global, dims[r], generator=g)

    x = inp(rank).requires_grad_(True)
    out = PooledEmbeddingsAllToAll(dist.group.WORLD, dims)(x, batch_size_per_rank=batch_sizes).wait()
    exp = torch.cat([inp(r)[offs[rank]:offs[rank + 1]] for r in range(world)], dim=1)
    torch.testing.assert_close(out, exp, rtol=0, atol=0)

    def wgt(r):
        g = torch.Generator().manual_seed(170 + r)
        return torch.randn(batch_sizes[r], sum(dims), generator=g)

    (out * wgt(rank)).sum().backward()
    lo = sum(dims[:rank])
    exp_g = torch.cat([wgt(r)[:, lo:lo + dims[rank]] for r in range(world)], dim=0)
    if comm_ops.GRADIENT_DIVISION:
        exp_g = exp_g / world
    torch.testing.assert_close(x.grad, exp_g, rtol=1e-6, atol=1e-7)


def check_vb_tw_dists(rank: int, world: int) -> None:
    """The variable-batch table-wise dist modules end to end (vb_tw_sharding.py), with a stand-in lookup:
    the "pooled embedding" of a bag is [sum of its ids, its length], so the output every rank gets back for
    its own samples can be computed from its local input alone."""
    from torchrec.distributed.embedding_types import SparseFeatures
    from torchrec.distributed.sharding.vb_sharding import VariableBatchShardingContext
    from torchrec.distributed.sharding.vb_tw_sharding import (
        VariableBatchTwPooledEmbeddingDist, VariableBatchTwSparseFeaturesDist)
    keys, splits, D = ["f0", "f1", "f2"], [2, 1], 2
    for batch_sizes in ([3, 5], [6, 0]):
        mine = make_kjt(rank, keys, batch_sizes[rank], 100)
        in_dist = VariableBatchTwSparseFeaturesDist(dist.group.WORLD, splits, [0, 0])
        out_dist = VariableBatchTwPooledEmbeddingDist(dist.group.WORLD, [s * D for s in splits])
        lengths_aw = in_dist(SparseFeatures(id_list_features=mine, id_score_list_features=None))
        sctx = VariableBatchShardingContext.from_lengths_awaitable(lengths_aw._a)
        assert sctx.batch_size_per_rank == batch_sizes and sctx.global_batch_size == sum(batch_sizes)
        assert sctx.rows_of(rank) == (sum(batch_sizes[:rank]), sum(batch_sizes[:rank + 1]))
        got = lengths_aw.wait().wait().id_list_features
        Bg = sum(batch_sizes)
        assert got.stride() == Bg

        def pooled(kjt, B):  # [B, n_keys * D]
            cols = []
            for k in kjt.keys():
                jt = kjt[k]
                off = jt.offsets().tolist()
                v = jt.values().double()
                cols.append(torch.stack([torch.stack([v[off[b]:off[b + 1]].sum(), torch.tensor(float(off[b + 1] - off[b]),
                                                                                               dtype=torch.float64)])
                                         for b in range(B)]) if B else torch.zeros(0, D, dtype=torch.float64))
            return torch.cat(cols, dim=1).float() if cols else torch.zeros(B, 0)

        out = out_dist(pooled(got, Bg), sctx).wait()
        exp = pooled(mine, batch_sizes[rank])  # keys in destination-rank order == model order here
        assert out.shape == exp.shape == (batch_sizes[rank], len(keys) * D)
        torch.testing.assert_close(out, exp, rtol=0, atol=0)


def check_vb_tw_sharding_factories(rank: int, world: int) -> None:
    """VariableBatchTwPooledEmbeddingSharding builds the variable-batch dists from a table-wise plan
    (the lookup it shares with TwPooledEmbeddingSharding needs the CUDA kernels and is covered on GPUs)."""
    from torchrec.distributed.embeddingbag import create_embedding_bag_sharding
    from torchrec.distributed.sharding.vb_tw_sharding import (
        VariableBatchTwPooledEmbeddingDist, VariableBatchTwPooledEmbeddingSharding, VariableBatchTwSparseFeaturesDist)
    from torchrec.distributed.types import ParameterSharding, ShardingEnv
    from torchrec.modules.embedding_configs import EmbeddingTableConfig, PoolingType
    infos = []
    for i, (rows, dim, owner) in enumerate([(10, 8, 0), (20, 4, 1), (30, 8, 0)]):
        cfg = EmbeddingTableConfig(num_embeddings=rows, embedding_dim=dim, name=f"t{i}", feature_names=[f"f{i}"],
                                   pooling=PoolingType.SUM, is_weighted=False, has_feature_processor=False,
                                   embedding_names=[f"f{i}"])
        infos.append((cfg, ParameterSharding(sharding_type="table_wise", compute_kernel="batched_fused",
                                             ranks=[owner]), None))
    env = ShardingEnv.from_process_group(dist.group.WORLD)
    sh = create_embedding_bag_sharding("table_wise", infos, env, torch.device("cpu"), variable_batch_size=True)
    assert isinstance(sh, VariableBatchTwPooledEmbeddingSharding)
    assert isinstance(sh.create_input_dist(), VariableBatchTwSparseFeaturesDist)
    out = sh.create_output_dist()
    assert isinstance(out, VariableBatchTwPooledEmbeddingDist) and out._dist._dim_sum_per_rank == [16, 4]
    assert sh.id_list_feature_names() == ["f0", "f2", "f1"] and sh.embedding_dims() == [8, 8, 4]
    try:
        create_embedding_bag_sharding("row_wise", infos, env, torch.device("cpu"), variable_batch_size=True)
    except ValueError as e:
        assert "table_wise" in str(e)
    else:
        raise AssertionError("variable batch size with row-wise sharding must be refused")


def check_row_wise_input_dist(rank: int, world: int) -> None:
    from torchrec.distributed.dist_data import KJTAllToAll
    from torchrec.distributed.embedding_sharding import bucketize_kjt_before_all2all
    keys, B, rows = ["a", "b"], 4, [10, 7]
    block = torch.tensor([(r + world - 1) // world for r in rows])
    mine = make_kjt(rank, keys, B, 7)
    bucketized, _ = bucketize_kjt_before_all2all(mine, world, block)
    got = KJTAllToAll(dist.group.WORLD, [len(keys)] * world)(bucketized).wait().wait()
    assert got.keys() == keys and got.stride() == B * world
    inputs = [make_kjt(r, keys, B, 7) for r in range(world)]
    for f, k in enumerate(keys):
        bs = int(block[f])
        exp_v, exp_l = [], []
        for src in inputs:
            jt = src[k]
            off = jt.offsets().tolist()
            vals = jt.values().tolist()
            for b in range(B):
                sel = [v - rank * bs for v in vals[off[b]:off[b + 1]] if v // bs == rank]
                exp_v += sel
                exp_l.append(len(sel))
        assert got[k].lengths().tolist() == exp_l, (k, got[k].lengths().tolist(), exp_l)
        assert got[k].values().tolist() == exp_v


def check_pooled_all2all(rank: int, world: int) -> None:
    from torchrec.distributed import comm_ops
    from torchrec.distributed.dist_data import PooledEmbeddingsAllToAll
    dims, B_local = [8, 4], 3
    B_global = B_local * world

    def inp(r):
        g = torch.Generator().manual_seed(50 + r)
        return torch.randn(B_global, dims[r], generator=g)

    x = inp(rank).requires_grad_(True)
    out = PooledEmbeddingsAllToAll(dist.group.WORLD, dims)(x).wait()
    exp = torch.cat([inp(r)[rank * B_local:(rank + 1) * B_local] for r in range(world)], dim=1)
    torch.testing.assert_close(out, exp, rtol=0, atol=0)

    def wgt(r):  # the upstream gradient rank r applies to its output
        g = torch.Generator().manual_seed(70 + r)
        return torch.randn(B_local, sum(dims), generator=g)

    (out * wgt(rank)).sum().backward()
    lo = sum(dims[:rank])
    exp_g = torch.cat([wgt(r)[:, lo:lo + dims[rank]] for r in range(world)], dim=0)
    if comm_ops.GRADIENT_DIVISION:
        exp_g = exp_g / world
    torch.testing.assert_close(x.grad, exp_g, rtol=1e-6, atol=1e-7)


def check_reduce_scatter(rank: int, world: int) -> None:
    from torchrec.distributed import comm_ops
    from torchrec.distributed.dist_data import PooledEmbeddingsReduceScatter
    B_local, D = 2, 8

    def inp(r):
        g = torch.Generator().manual_seed(90 + r)
        return torch.randn(B_local * world, D, generator=g)

    x = inp(rank).requires_grad_(True)
    out = PooledEmbeddingsReduceScatter(dist.group.WORLD)(x).wait()
    exp = sum(inp(r)[rank * B_local:(rank + 1) * B_local] for r in range(world))
    torch.testing.assert_close(out, exp, rtol=1e-6, atol=1e-6)

    def wgt(r):
        g = torch.Generator().manual_seed(110 + r)
        return torch.randn(B_local, D, generator=g)

    (out * wgt(rank)).sum().backward()
    exp_g = torch.cat([wgt(r) for r in range(world)], dim=0)
    if comm_ops.GRADIENT_DIVISION:
        exp_g = exp_g / world
    torch.testing.assert_close(x.grad, exp_g, rtol=1e-6, atol=1e-7)


def check_sequence_all2all(rank: int, world: int) -> None:
    """Non-pooled rows go back to the rank that owns the sample, in its local (key, sample, id) order, and
    gradients make the reverse trip (torchrec/distributed/tests/test_dist_data.py SeqEmbeddingsAllToAllTest)."""
    from torchrec.distributed.dist_data import KJTAllToAll, SequenceEmbeddingAllToAll
    keys, splits, B, D = ["f0", "f1", "f2"], [2, 1], 3, 4
    scale = torch.arange(1, D + 1, dtype=torch.float32)
    mine = make_kjt(rank, keys, B, 100)
    indices_aw = KJTAllToAll(dist.group.WORLD, splits)(mine).wait()
    got = indices_aw.wait()
    # "lookup": row of id v = v * [1..D]; rows are in the received (key, source rank, sample) order
    local_embs = (got.values().float().unsqueeze(1) * scale).requires_grad_(True)
    m = SequenceEmbeddingAllToAll(dist.group.WORLD, splits)
    out = m(local_embs, got.lengths().view(-1, got.stride()), indices_aw._in_lengths_per_worker,
            indices_aw._out_lengths_per_worker).wait()
    exp = mine.values().float().unsqueeze(1) * scale  # my own ids, key-major: exactly the local KJT order
    torch.testing.assert_close(out, exp, rtol=0, atol=0)
    # upstream gradient of the row of id v on rank r: v * [1..D] + 1000 r
    (out * (exp + 1000.0 * rank)).sum().backward()
    lo = sum(splits[:rank])
    inputs = [make_kjt(r, keys, B, 100) for r in range(world)]
    rows = []
    for k in keys[lo:lo + splits[rank]]:
        for r in range(world):
            rows.append(inputs[r][k].values().float().unsqueeze(1) * scale + 1000.0 * r)
    torch.testing.assert_close(local_embs.grad, torch.cat(rows), rtol=0, atol=0)


def check_sharded_tensor_state(rank: int, world: int) -> None:
    """Checkpoint format (reference batched_embedding_kernel.py:160-246, embedding_kernel.get_state_dict):
    a row-wise table becomes a ShardedTensor of the global shape whose local shard VIEWS the live rows;
    momentum1 is sharded along the same rows; an empty shard (fewer rows than ranks) is legal; loading
    copies shard to shard."""
    from torch.distributed._shard.sharded_tensor import ShardedTensor
    from torchrec.distributed.embedding_kernel import copy_state, local_tensor, sharded_tensor_from_local
    from torchrec.distributed.sharding.rw_sharding import row_wise_shard_sizes
    from torchrec.distributed.types import ShardMetadata
    from torchrec.optim.keyed import KeyedOptimizer
    D = 4
    for rows in (5, 1):
        sizes = row_wise_shard_sizes(rows, world)
        offs = [sum(sizes[:r]) for r in range(world)]
        metas = [ShardMetadata([offs[r], 0], [sizes[r], D], f"rank:{r}/cuda:{r}") for r in range(world)]
        live = torch.full((sizes[rank], D), float(rank + 1))
        mom = torch.arange(sizes[rank], dtype=torch.float32) + 10 * rank
        w = sharded_tensor_from_local(live, metas[rank], metas, [rows, D], dist.group.WORLD)
        m = sharded_tensor_from_local(mom, metas[rank], metas, [rows, D], dist.group.WORLD, rowwise=True)
        assert isinstance(w, ShardedTensor) and tuple(w.size()) == (rows, D) and tuple(m.size()) == (rows,)
        assert [tuple(x.shard_sizes) for x in w.metadata().shards_metadata] == [(n, D) for n in sizes]
        assert [tuple(x.shard_offsets) for x in m.metadata().shards_metadata] == [(o,) for o in offs]
        assert local_tensor(w).data_ptr() == live.data_ptr() or live.numel() == 0  # a view, not a copy
        local_tensor(w).mul_(2.0)
        assert torch.equal(live, torch.full((sizes[rank], D), 2.0 * (rank + 1)))
        # load from a tensor of the local shape and from another ShardedTensor
        copy_state(w, torch.zeros(sizes[rank], D))
        assert float(live.abs().sum()) == 0.0
        src = sharded_tensor_from_local(torch.ones(sizes[rank], D), metas[rank], metas, [rows, D], dist.group.WORLD)
        copy_state(w, src)
        assert torch.equal(live, torch.ones(sizes[rank], D))
        # optimizer checkpoint: state keyed by the parameter's name, momentum loaded shard to shard
        opt = KeyedOptimizer({"t.weight": w}, {w: {"t.momentum1": m}}, [{"params": [w], "lr": 0.1}])
        sd = opt.state_dict()
        assert list(sd["state"]) == ["t.weight"] and isinstance(sd["state"]["t.weight"]["t.momentum1"], ShardedTensor)
        new_m = sharded_tensor_from_local(torch.full((sizes[rank],), 7.0), metas[rank], metas, [rows, D],
                                          dist.group.WORLD, rowwise=True)
        opt.save_param_groups(True)  # reference default: param_groups are neither saved nor loaded
        opt.load_state_dict({"state": {"t.weight": {"t.momentum1": new_m}},
                             "param_groups": [{"params": ["t.weight"], "lr": 0.5}]})
        assert torch.equal(mom, torch.full((sizes[rank],), 7.0)) and opt.param_groups[0]["lr"] == 0.5


def main() -> None:
    import cpu_ops
    cpu_ops.register()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("gloo", rank=rank, world_size=world)
    check_kjt_all2all(rank, world)
    check_kjt_all2all_variable_batch(rank, world)
    check_pooled_all2all_variable_batch(rank, world)
    check_vb_tw_dists(rank, world)
    check_vb_tw_sharding_factories(rank, world)
    check_row_wise_input_dist(rank, world)
    check_pooled_all2all(rank, world)
    check_reduce_scatter(rank, world)
    check_sequence_all2all(rank, world)
    check_sharded_tensor_state(rank, world)
    dist.barrier()
    if rank == 0:
        print(f"GLOO DIST OK world={world}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
