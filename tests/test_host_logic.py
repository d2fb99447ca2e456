"""CPU tests of the host layer that mirrors the reference API: KeyedJaggedTensor bookkeeping
(torchrec/sparse/tests/test_jagged_tensor.py vectors), KJTAllToAll index helpers
(torchrec/distributed/tests/test_dist_data.py), row-wise shard geometry, the planner, the TBE module's
bookkeeping, and the world-size-2 gloo run of the sharded data movement.
Jagged ops run on tests/cpu_ops.py (oracle-backed, test-only); the product registers CUDA only."""
import os
import subprocess
import sys

import pytest
import torch

import cpu_ops

cpu_ops.register()

HERE = os.path.dirname(os.path.abspath(__file__))


def test_get_recat_reference_vectors():
    from torchrec.distributed.dist_data import _get_recat
    assert _get_recat(2, 4, 1).tolist() == [0, 2, 4, 6, 1, 3, 5, 7]
    assert _get_recat(2, 4, 2).tolist() == [0, 4, 2, 6, 1, 5, 3, 7]
    assert _get_recat(0, 4, 1).numel() == 0
    from oracle.jagged_ref import get_recat
    for ls, ns, st in [(1, 2, 1), (3, 8, 4), (5, 4, 2)]:
        assert _get_recat(ls, ns, st).tolist() == list(get_recat(ls, ns, st))


def test_split_lengths():
    from torchrec.distributed.dist_data import _split_lengths
    assert _split_lengths([2, 1, 0], ["a", "b", "c"], [0, 3, 5, 9]) == [5, 4, 0]


def test_row_wise_shard_sizes():
    from torchrec.distributed.sharding.rw_sharding import row_wise_shard_sizes
    assert row_wise_shard_sizes(5, 4) == [2, 2, 1, 0]
    assert row_wise_shard_sizes(8, 4) == [2, 2, 2, 2]
    assert row_wise_shard_sizes(1, 2) == [1, 0]
    for rows in (1, 7, 100, 40_000_000):
        for w in (1, 2, 8):
            s = row_wise_shard_sizes(rows, w)
            assert sum(s) == rows and len(s) == w and all(a >= b for a, b in zip(s, s[1:]))


def _kjt():
    from torchrec import KeyedJaggedTensor
    # the tensor of torchrec/sparse/tests/test_jagged_tensor.py:394-415
    return KeyedJaggedTensor(keys=["index_0", "index_1"],
                             values=torch.tensor([1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0, 8.0]),
                             weights=torch.tensor([1.0, 0.5, 1.5, 1.0, 0.5, 1.0, 1.0, 1.5]),
                             lengths=torch.tensor([2, 0, 1, 1, 1, 3], dtype=torch.int32))


def test_kjt_metadata_and_views():
    k = _kjt()
    assert k.stride() == 3 and k.keys() == ["index_0", "index_1"]
    assert k.offsets().tolist() == [0, 2, 2, 3, 4, 5, 8]
    assert k.length_per_key() == [3, 5] and k.offset_per_key() == [0, 3, 8]
    j0, j1 = k["index_0"], k["index_1"]
    assert j0.values().tolist() == [1.0, 2.0, 3.0] and j0.lengths().tolist() == [2, 0, 1]
    assert j1.values().tolist() == [4.0, 5.0, 6.0, 7.0, 8.0] and j1.offsets().tolist() == [0, 1, 2, 5]
    assert j1.weights().tolist() == [1.0, 0.5, 1.0, 1.0, 1.5]
    d = k.to_dict()
    assert list(d) == ["index_0", "index_1"] and d["index_1"].values().tolist() == j1.values().tolist()


def test_kjt_split_and_permute():
    k = _kjt()
    a, b = k.split([1, 1])
    assert a.keys() == ["index_0"] and b.keys() == ["index_1"]
    assert a.values().tolist() == [1.0, 2.0, 3.0] and b.lengths().tolist() == [1, 1, 3]
    p = k.permute([1, 0])
    assert p.keys() == ["index_1", "index_0"]
    assert p.values().tolist() == [4.0, 5.0, 6.0, 7.0, 8.0, 1.0, 2.0, 3.0]
    assert p.lengths().tolist() == [1, 1, 3, 2, 0, 1]
    assert p.weights().tolist() == [1.0, 0.5, 1.0, 1.0, 1.5, 1.0, 0.5, 1.5]
    # duplicates are allowed in permute (test_jagged_tensor.py:560-600)
    q = k.permute([1, 0, 1])
    assert q.keys() == ["index_1", "index_0", "index_1"] and q.values().numel() == 13


def test_kjt_empty_and_from_offsets():
    from torchrec import KeyedJaggedTensor
    e = KeyedJaggedTensor(keys=[], values=torch.tensor([], dtype=torch.int64),
                          lengths=torch.tensor([], dtype=torch.int32), stride=0)
    assert e.keys() == [] and e.values().numel() == 0
    k = KeyedJaggedTensor.from_offsets_sync(keys=["a", "b"], values=torch.arange(5),
                                            offsets=torch.tensor([0, 2, 2, 3, 5]))
    assert k.lengths().tolist() == [2, 0, 1, 2] and k.stride() == 2
    z = KeyedJaggedTensor.from_lengths_sync(keys=["a"], values=torch.tensor([], dtype=torch.int64),
                                            lengths=torch.zeros(4, dtype=torch.int32))
    assert z["a"].values().numel() == 0 and z.offsets().tolist() == [0, 0, 0, 0, 0]


def test_keyed_tensor_regroup():
    from torchrec import KeyedTensor
    kt = KeyedTensor(keys=["a", "b", "c"], length_per_key=[2, 3, 1], values=torch.arange(12.0).view(2, 6))
    d = kt.to_dict()
    assert d["b"].tolist() == [[2.0, 3.0, 4.0], [8.0, 9.0, 10.0]] and kt.offset_per_key() == [0, 2, 5, 6]
    assert kt["c"].tolist() == [[5.0], [11.0]]


def test_unsharded_ebc_matches_embedding_bag():
    """EmbeddingBagCollection on CPU tensors (torchrec/modules/tests/test_embedding_modules.py:20-80)."""
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig, KeyedJaggedTensor
    torch.manual_seed(0)
    ebc = EmbeddingBagCollection([
        EmbeddingBagConfig(name="t1", embedding_dim=4, num_embeddings=10, feature_names=["f1"]),
        EmbeddingBagConfig(name="t2", embedding_dim=8, num_embeddings=10, feature_names=["f2", "f3"])])
    kjt = KeyedJaggedTensor.from_offsets_sync(keys=["f1", "f2", "f3"], values=torch.tensor([1, 2, 4, 5, 4, 3, 2, 9, 0]),
                                              offsets=torch.tensor([0, 2, 4, 6, 8, 9, 9]))
    out = ebc(kjt)
    assert out.keys() == ["f1", "f2", "f3"] and out.values().shape == (2, 20)
    w1, w2 = ebc.embedding_bags["t1"].weight, ebc.embedding_bags["t2"].weight
    torch.testing.assert_close(out["f1"][0], w1[1] + w1[2])
    torch.testing.assert_close(out["f2"][1], w2[2] + w2[9])
    torch.testing.assert_close(out["f3"][1], torch.zeros(8))


def test_planner_is_deterministic_and_honours_constraints():
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    rows = [1000, 10, 500, 20_000, 7]
    ebc = EmbeddingBagCollection([EmbeddingBagConfig(name=f"t{i}", embedding_dim=16, num_embeddings=r,
                                                     feature_names=[f"f{i}"]) for i, r in enumerate(rows)],
                                 device=torch.device("meta"))
    cons = {"t3": ParameterConstraints(sharding_types=["row_wise"])}
    plans = [EmbeddingShardingPlanner(Topology(4), cons).plan(ebc, [EmbeddingBagCollectionSharder()])
             for _ in range(2)]
    p0, p1 = (next(iter(p.plan.values())) for p in plans)
    assert set(p0) == {f"t{i}" for i in range(5)}
    for name in p0:
        assert p0[name].sharding_type == p1[name].sharding_type and p0[name].ranks == p1[name].ranks
    assert p0["t3"].sharding_type == "row_wise" and sorted(p0["t3"].ranks) == [0, 1, 2, 3]
    tw = [n for n in p0 if p0[n].sharding_type == "table_wise"]
    assert all(len(p0[n].ranks) == 1 and 0 <= p0[n].ranks[0] < 4 for n in tw)
    # greedy balance: the two biggest table-wise tables land on different ranks
    assert p0["t0"].ranks != p0["t2"].ranks


def test_tbe_module_requires_cuda():
    """No CPU path: the fused module refuses a CPU device instead of computing on the host."""
    from fbgemm_gpu.split_table_batched_embeddings_ops import (ComputeDevice, EmbeddingLocation,
                                                               SplitTableBatchedEmbeddingBagsCodegen)
    with pytest.raises((NotImplementedError, RuntimeError, AssertionError, ValueError)):
        SplitTableBatchedEmbeddingBagsCodegen([(10, 4, EmbeddingLocation.HOST, ComputeDevice.CPU)])


def test_row_size_helpers():
    from fbgemm_gpu.split_embedding_configs import SparseType
    from fbgemm_gpu.split_table_batched_embeddings_ops import rounded_row_size_in_bytes, unpadded_row_size_in_bytes
    # split_table_batched_embeddings_ops.py:1469-1485: int8 rows carry a 4-byte fp16 scale/bias header
    assert unpadded_row_size_in_bytes(128, SparseType.INT8) == 132
    assert rounded_row_size_in_bytes(128, SparseType.INT8, 16) == 144
    assert rounded_row_size_in_bytes(12, SparseType.INT8, 16) == 16


def test_world2_gloo_sharded_data_movement():
    """KJTAllToAll, bucketize + row-wise input dist, pooled all-to-all and reduce-scatter (fwd + bwd)
    at world_size 2 over gloo."""
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29741",
                        os.path.join(HERE, "dist_gloo_worker.py")], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0 and "GLOO DIST OK world=2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


# ---------------------------------------------------------------- binary Criteo datapipe
def test_binary_criteo_utils_known_answers(tmp_path):
    """get_file_idx_to_row_range / load_npy_range against the reference's own expectations
    (torchrec/datasets/tests/test_criteo.py:164-193) and the fixture the reference class produced."""
    import numpy as np
    from golden_cases import CRITEO_PATH
    from torchrec.datasets.criteo import BinaryCriteoUtils
    expected = [{0: (0, 13), 1: (0, 2)}, {1: (3, 16), 2: (0, 2)}, {2: (3, 19)}]
    z = np.load(CRITEO_PATH)
    for r in range(3):
        got = BinaryCriteoUtils.get_file_idx_to_row_range(lengths=[14, 17, 20], rank=r, world_size=3)
        assert got == expected[r]
        assert [[k, a, b] for k, (a, b) in sorted(got.items())] == z[f"row_range/rank{r}"].tolist()
    full = np.random.default_rng(0).random((10, 13), dtype=np.float32)
    path = str(tmp_path / "dense.npy")
    np.save(path, full)
    assert BinaryCriteoUtils.get_shape_from_npy(path) == (10, 13)
    for mmap in (False, True):
        np.testing.assert_array_equal(BinaryCriteoUtils.load_npy_range(path, 2, 4, mmap_mode=mmap), full[2:6])
    with pytest.raises(ValueError, match="out of bounds"):
        BinaryCriteoUtils.load_npy_range(path, 10, 1)
    with pytest.raises(ValueError, match="exceeds number of available rows"):
        BinaryCriteoUtils.load_npy_range(path, 8, 3)


@pytest.mark.parametrize("mmap_mode", [False, True])
def test_binary_criteo_datapipe_matches_reference_fixture(tmp_path, mmap_mode):
    """Every batch of every rank equals, bit for bit, what the reference's InMemoryBinaryCriteoIterDataPipe
    produced from the same seeded day files (fixture: tests/golden/make_golden_criteo.py)."""
    import numpy as np
    from golden_cases import CRITEO_CASES, CRITEO_PATH, write_criteo_days
    from torchrec.datasets.criteo import CAT_FEATURE_COUNT, INT_FEATURE_COUNT, InMemoryBinaryCriteoIterDataPipe
    z = np.load(CRITEO_PATH)
    for ci, (rows_per_file, batch_size, world, use_hashes, seed) in enumerate(CRITEO_CASES):
        folder = tmp_path / f"case{ci}"
        folder.mkdir()
        dense, sparse, labels = write_criteo_days(str(folder), rows_per_file, seed)
        hashes = [3 * i + 2 for i in range(CAT_FEATURE_COUNT)] if use_hashes else None
        lens = set()
        for rank in range(world):
            pipe = InMemoryBinaryCriteoIterDataPipe(dense, sparse, labels, batch_size=batch_size, rank=rank,
                                                    world_size=world, hashes=hashes, mmap_mode=mmap_mode)
            batches = list(pipe)
            pre = f"case{ci}/rank{rank}"
            assert len(batches) == len(pipe) == int(z[f"{pre}/num_batches"])
            lens.add(len(batches))
            for b, batch in enumerate(batches):
                kjt = batch.sparse_features
                assert tuple(batch.dense_features.shape) == (batch_size, INT_FEATURE_COUNT)
                assert tuple(batch.labels.shape) == (batch_size,) and kjt.stride() == batch_size
                for got, name in ((batch.dense_features, "dense"), (kjt.values(), "values"),
                                  (batch.labels, "labels")):
                    exp = z[f"{pre}/{name}"][b]
                    assert got.numpy().dtype == exp.dtype, (name, got.dtype, exp.dtype)
                    np.testing.assert_array_equal(got.numpy(), exp, err_msg=f"{pre} batch {b} {name}")
                np.testing.assert_array_equal(kjt.lengths().numpy(), z[f"{pre}/lengths"])
                np.testing.assert_array_equal(kjt.offsets().numpy(), z[f"{pre}/offsets"])
                assert kjt.keys() == z[f"{pre}/keys"].tolist()
                if hashes is not None:
                    v = kjt.values().view(CAT_FEATURE_COUNT, batch_size)
                    assert bool((v < torch.tensor(hashes).view(-1, 1)).all())
        assert len(lens) == 1  # every rank serves the same number of batches


def test_binary_criteo_datapipe_shuffle_batches(tmp_path):
    """shuffle_batches permutes the rows INSIDE each batch, dense / ids / labels in unison."""
    import numpy as np
    from golden_cases import write_criteo_days
    from torchrec.datasets.criteo import CAT_FEATURE_COUNT, InMemoryBinaryCriteoIterDataPipe
    dense, sparse, labels = write_criteo_days(str(tmp_path), [11, 30, 7], 9)
    plain = list(InMemoryBinaryCriteoIterDataPipe(dense, sparse, labels, batch_size=12, rank=0, world_size=1))
    np.random.seed(0)
    shuf = list(InMemoryBinaryCriteoIterDataPipe(dense, sparse, labels, batch_size=12, rank=0, world_size=1,
                                                 shuffle_batches=True))
    assert len(plain) == len(shuf) == 4
    moved = False
    for p, s in zip(plain, shuf):
        rows_p = np.concatenate([p.dense_features.numpy(), p.sparse_features.values().view(CAT_FEATURE_COUNT, -1).T.numpy(),
                                 p.labels.numpy()[:, None]], axis=1)
        rows_s = np.concatenate([s.dense_features.numpy(), s.sparse_features.values().view(CAT_FEATURE_COUNT, -1).T.numpy(),
                                 s.labels.numpy()[:, None]], axis=1)
        order = np.lexsort(rows_p.T[::-1]), np.lexsort(rows_s.T[::-1])
        np.testing.assert_array_equal(rows_p[order[0]], rows_s[order[1]])
        moved |= not np.array_equal(rows_p, rows_s)
    assert moved


# ---------------------------------------------------------------- KJT / KeyedTensor vs the reference's own classes
def test_kjt_ops_match_reference_fixture():
    """offsets / per-key totals / permute (duplicates, drops) / split / to_dict / KeyedTensor.regroup equal what
    the reference's own KeyedJaggedTensor and KeyedTensor produced on the same seeded inputs
    (tests/golden/make_golden_kjt.py ran them over the reference's CPU kernels)."""
    import numpy as np
    from golden_cases import KJT_CASES, kjt_case_inputs
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor, KeyedTensor
    z = np.load(os.path.join(HERE, "golden", "kjt_ops.npz"))

    def eq(got, key):
        exp = z[key]
        got = got.numpy() if isinstance(got, torch.Tensor) else np.asarray(got)
        assert got.dtype == exp.dtype or exp.dtype.kind in "US", (key, got.dtype, exp.dtype)
        np.testing.assert_array_equal(got, exp, err_msg=key)

    for ci, case in enumerate(KJT_CASES):
        keys, lengths, values, weights, stride = kjt_case_inputs(case)
        k = KeyedJaggedTensor(keys=keys, values=torch.from_numpy(values), lengths=torch.from_numpy(lengths),
                              weights=None if weights is None else torch.from_numpy(weights), stride=stride)
        pre = f"case{ci}"
        eq(k.offsets(), f"{pre}/offsets")
        eq(np.array(k.length_per_key(), dtype=np.int64), f"{pre}/length_per_key")
        eq(np.array(k.offset_per_key(), dtype=np.int64), f"{pre}/offset_per_key")
        for pi, perm in enumerate(case["permutes"]):
            p = k.permute(perm, torch.tensor(perm, dtype=torch.int32))
            assert p.keys() == z[f"{pre}/permute{pi}/keys"].tolist()
            eq(p.values(), f"{pre}/permute{pi}/values")
            eq(p.lengths(), f"{pre}/permute{pi}/lengths")
            eq(np.array(p.length_per_key(), dtype=np.int64), f"{pre}/permute{pi}/length_per_key")
            if weights is not None:
                eq(p.weights(), f"{pre}/permute{pi}/weights")
        for si, seg in enumerate(case["splits"]):
            parts = k.split(seg)
            assert len(parts) == len(seg)
            for part_i, part in enumerate(parts):
                q = f"{pre}/split{si}/part{part_i}"
                assert part.keys() == z[f"{q}/keys"].tolist() and part.stride() == int(z[f"{q}/stride"])
                eq(part.values(), f"{q}/values")
                eq(part.lengths(), f"{q}/lengths")
                if weights is not None:
                    eq(part.weights(), f"{q}/weights")
        d = k.to_dict()
        assert list(d) == keys
        for key, j in d.items():
            eq(j.values(), f"{pre}/dict/{key}/values")
            eq(j.lengths(), f"{pre}/dict/{key}/lengths")
            eq(j.offsets(), f"{pre}/dict/{key}/offsets")
    kts = [KeyedTensor(keys=names, length_per_key=dims, values=torch.from_numpy(z[f"regroup/in_{names[0]}"]))
           for names, dims in ((["a", "b", "c"], [3, 1, 4]), (["d", "e"], [2, 5]))]
    outs = KeyedTensor.regroup(kts, [["b", "e"], ["a", "d", "c"], ["e", "a"]])
    for gi, t in enumerate(outs):
        eq(t, f"regroup/out{gi}")


def test_recat_and_split_lengths_match_reference_fixture():
    """_get_recat (fixed and variable batch sizes, staggered) and _split_lengths equal what the reference's
    own functions returned (tests/golden/make_golden_recat.py executed them from /root/reference)."""
    import numpy as np
    from golden_cases import RECAT_CASES, SPLIT_LENGTHS_CASES
    from torchrec.distributed.dist_data import _get_recat, _split_lengths
    z = np.load(os.path.join(HERE, "golden", "recat.npz"))
    for i, (local_split, num_splits, stagger, batch_sizes) in enumerate(RECAT_CASES):
        got = _get_recat(local_split, num_splits, stagger, torch.device("cpu"), batch_sizes)
        exp = z[f"recat{i}"]
        assert got.numpy().dtype == exp.dtype, (i, got.dtype, exp.dtype)
        np.testing.assert_array_equal(got.numpy(), exp, err_msg=f"recat case {i}")
    for i, (splits, keys, offset_per_key) in enumerate(SPLIT_LENGTHS_CASES):
        assert _split_lengths(splits, keys, offset_per_key) == z[f"split_lengths{i}"].tolist()


def test_bucketize_kjt_matches_reference_fixture():
    """Row-wise input dist, step 1: bucketize_kjt_before_all2all (plain / weighted / unbucketize permutation /
    positions) equals the reference's own function on its own KJT class and CPU kernel
    (tests/golden/make_golden_bucketize_kjt.py)."""
    import numpy as np
    from golden_cases import BUCKETIZE_KJT_CASES, bucketize_kjt_inputs
    from torchrec.distributed.embedding_sharding import bucketize_kjt_before_all2all
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor
    z = np.load(os.path.join(HERE, "golden", "bucketize_kjt.npz"))
    for ci, case in enumerate(BUCKETIZE_KJT_CASES):
        keys, lengths, values, weights, B, block_sizes = bucketize_kjt_inputs(case)
        k = KeyedJaggedTensor(keys=keys, values=torch.from_numpy(values), lengths=torch.from_numpy(lengths),
                              weights=None if weights is None else torch.from_numpy(weights), stride=B)
        got, unbucketize = bucketize_kjt_before_all2all(k, case["W"], torch.from_numpy(block_sizes),
                                                        output_permute=case["permute"], bucketize_pos=case["pos"])
        pre = f"case{ci}"
        assert got.keys() == z[f"{pre}/keys"].tolist() and got.stride() == int(z[f"{pre}/stride"])
        for name, t in (("values", got.values()), ("lengths", got.lengths()), ("weights", got.weights_or_none()),
                        ("unbucketize", unbucketize)):
            key = f"{pre}/{name}"
            assert (t is not None) == (key in z.files), key
            if t is not None:
                assert t.numpy().dtype == z[key].dtype, (key, t.dtype, z[key].dtype)
                np.testing.assert_array_equal(t.numpy(), z[key], err_msg=key)


def test_combined_optimizer_checkpoint_is_flat_and_strict():
    """Two fused optimizers combined under empty keys (one per kernel group / sharding type, as
    ShardedEmbeddingBagCollection builds them) must BOTH survive state_dict()/load_state_dict(); layout and
    strictness follow the reference's KeyedOptimizer (torchrec/optim/keyed.py:66-196): flat
    {"state": {<param key>: {<state key>: tensor}}}, no param_groups by default, errors on a missing
    parameter / state key / count mismatch."""
    from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer

    def fused(name, rows):
        w, m = torch.zeros(rows, 4), torch.arange(rows, dtype=torch.float32)
        return KeyedOptimizer({f"embedding_bags.{name}.weight": w}, {w: {f"{name}.momentum1": m}},
                              [{"params": [w], "lr": 0.1}]), m

    o0, m0 = fused("t0", 3)
    o1, m1 = fused("t1", 5)
    comb = CombinedOptimizer([("", o0), ("", o1)])
    sd = comb.state_dict()
    assert set(sd) == {"state"}
    assert set(sd["state"]) == {"embedding_bags.t0.weight", "embedding_bags.t1.weight"}
    assert sd["state"]["embedding_bags.t0.weight"]["t0.momentum1"] is m0
    new = {"state": {"embedding_bags.t0.weight": {"t0.momentum1": torch.full((3,), 7.0)},
                     "embedding_bags.t1.weight": {"t1.momentum1": torch.full((5,), 9.0)}}}
    comb.load_state_dict(new)
    assert torch.equal(m0, torch.full((3,), 7.0)) and torch.equal(m1, torch.full((5,), 9.0))
    with pytest.raises(ValueError, match="Different parameter count"):
        comb.load_state_dict({"state": {"embedding_bags.t0.weight": new["state"]["embedding_bags.t0.weight"]}})
    with pytest.raises(ValueError, match="not found"):
        comb.load_state_dict({"state": {"embedding_bags.t0.weight": new["state"]["embedding_bags.t0.weight"],
                                        "embedding_bags.tX.weight": {"t1.momentum1": torch.zeros(5)}}})
    with pytest.raises(ValueError, match="State key"):
        comb.load_state_dict({"state": {"embedding_bags.t0.weight": {"wrong": torch.zeros(3)},
                                        "embedding_bags.t1.weight": new["state"]["embedding_bags.t1.weight"]}})
    with pytest.raises(ValueError, match="Duplicate param key"):
        CombinedOptimizer([("", o0), ("", fused("t0", 3)[0])])
    # param_groups travel only when asked for
    comb.save_param_groups(True)
    sd = comb.state_dict()
    assert [g["params"] for g in sd["param_groups"]] == [["embedding_bags.t0.weight"], ["embedding_bags.t1.weight"]]
    sd["param_groups"][1]["lr"] = 0.5
    comb.load_state_dict(sd)
    assert o1.param_groups[0]["lr"] == 0.5 and o0.param_groups[0]["lr"] == 0.1


def test_cuda_graph_step_signature_and_static_buffers():
    """Host logic of TrainPipelineSparseDist's CUDA-graph mode (runs on CPU tensors): the step signature covers
    shapes, dtypes, keys and the per-key id counts (they become pointer offsets inside a captured graph), the
    static copy has the same structure with its own storage, and the flattened tensor lists line up one to one."""
    from torchrec.datasets.utils import Batch
    from torchrec.distributed.embedding_types import SparseFeatures, SparseFeaturesList
    from torchrec.distributed.train_pipeline import _sig, _static_copy, _tensors
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

    def kjt(lengths):
        lengths = torch.tensor(lengths, dtype=torch.int32)
        return KeyedJaggedTensor(keys=["a", "b"], values=torch.arange(int(lengths.sum())), lengths=lengths)

    def step(lengths):
        k = kjt(lengths)
        batch = Batch(dense_features=torch.zeros(2, 3), sparse_features=k, labels=torch.zeros(2))
        return (batch, [SparseFeaturesList([SparseFeatures(id_list_features=kjt(lengths))])])

    a, b, c = step([1, 2, 0, 1]), step([1, 2, 0, 1]), step([1, 1, 1, 1])  # c: same total, other per-key counts
    assert _sig(a) == _sig(b)
    assert _sig(a) != _sig(c)
    assert _sig(a) != _sig(step([1, 2, 0, 2]))
    s = _static_copy(a)
    ta, ts = _tensors(a, []), _tensors(s, [])
    assert len(ta) == len(ts) > 0
    for x, y in zip(ta, ts):
        assert x.shape == y.shape and x.dtype == y.dtype and torch.equal(x, y) and x.data_ptr() != y.data_ptr()
    assert s[0].sparse_features.length_per_key() == a[0].sparse_features.length_per_key()
    with pytest.raises(TypeError):
        _sig(object())


def test_interaction_supported_shapes():
    """InteractionArch routes shapes the kernels refuse (D % 4 != 0, or a 4-sample tile beyond 200 KB of shared
    memory) to the plain torch formulation instead of failing at run time."""
    import importlib
    ops = importlib.import_module("b200_sparse.ops")
    assert ops.interaction_supported(26, 128) and ops.interaction_supported(1, 4)
    assert not ops.interaction_supported(26, 130)
    assert not ops.interaction_supported(200, 512)


def test_bench_reference_arm_line_and_isolation():
    """`bench.py --impl reference` (the CPU arm the driver times beside ours): prints ONE JSON line with the
    contract's keys, runs the reference's own CPU kernels when oracle/_ref is built (else the C port, and says so),
    and loads neither the product package nor libb200sparse.so."""
    import json
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    probe = ("import runpy, sys; sys.argv = ['bench.py', '--impl', 'reference', '--steps', '1', '--warmup', '1', "
             "'--cpu-batch', '512', '--cpu-row-cap', '20000']; "
             "runpy.run_path('bench.py', run_name='__main__'); "
             "bad = [m for m in sys.modules if m.split('.')[0] in ('b200_sparse', 'fbgemm_gpu', 'torchrec')]; "
             "maps = open('/proc/self/maps').read(); "
             "print('LOADED', bad, 'libb200sparse' in maps)")
    out = subprocess.run([sys.executable, "-c", probe], cwd=root, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "dlrm_train_samples_per_sec" and d["unit"] == "samples/s"
    assert d["higher_is_better"] is True and d["steps"] == 1 and d["value"] > 0
    assert d["e2e"] == {"value": d["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["value"] == d["value"] and cb["cores"] >= 1 and cb["kind"] in ("reference", "port")
    if os.path.exists(os.path.join(root, "oracle", "_ref", "libfbgemm_tbe_ref.so")):
        assert cb["kind"] == "reference"
    assert d["gpu_launches"] == 0
    loaded = [ln for ln in out.stdout.splitlines() if ln.startswith("LOADED")][0]
    assert loaded == "LOADED [] False", loaded
