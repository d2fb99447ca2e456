"""GPU parity at the module boundary: TBE module vs nn.EmbeddingBag + closed-form optimizer (the
reference's own oracle pattern, fbgemm_gpu/test/split_table_batched_embeddings_test.py:322-415,
1756-1785), KJT ops, sharded-vs-unsharded one-step parity
(torchrec/distributed/test_utils/test_model_parallel_base.py:144-290), quantized EBC, DLRM."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def pg():
    import torch.distributed as dist
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29533")
        dist.init_process_group("nccl", rank=0, world_size=1)
    yield dist.group.WORLD


def _tbe(rows, dims, opt, **kw):
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from fbgemm_gpu.split_table_batched_embeddings_ops import (
        ComputeDevice, EmbeddingLocation, SplitTableBatchedEmbeddingBagsCodegen)
    return SplitTableBatchedEmbeddingBagsCodegen(
        [(r, d, EmbeddingLocation.DEVICE, ComputeDevice.CUDA) for r, d in zip(rows, dims)],
        optimizer=getattr(EmbOptimType, opt), device=torch.device("cuda"), **kw)


@pytest.mark.parametrize("opt", ["EXACT_SGD", "EXACT_ROWWISE_ADAGRAD"])
@pytest.mark.parametrize("pooling", ["SUM", "MEAN"])
@pytest.mark.parametrize("weighted", [False, True])
def test_tbe_module_vs_embedding_bag(opt, pooling, weighted):
    from fbgemm_gpu.split_table_batched_embeddings_ops import PoolingMode
    if weighted and pooling == "MEAN":
        pytest.skip("MEAN pooling ignores per-sample weights in nn.EmbeddingBag")
    torch.manual_seed(1)
    rows, dims, B, L = [40, 9, 300], [16, 32, 8], 50, 4
    lr, eps = 0.1, 1e-4
    m = _tbe(rows, dims, opt, learning_rate=lr, eps=eps, pooling_mode=getattr(PoolingMode, pooling))
    m.init_embedding_weights_uniform(-1.0, 1.0)
    bags = [torch.nn.EmbeddingBag(r, d, mode=pooling.lower(), sparse=False, include_last_offset=True).cuda()
            for r, d in zip(rows, dims)]
    for b, w in zip(bags, m.split_embedding_weights()):
        b.weight.data.copy_(w)
    lengths = torch.randint(0, L + 1, (len(rows), B))
    idx = [torch.randint(0, r, (int(lengths[t].sum()),)) for t, r in enumerate(rows)]
    offs = [torch.cat([torch.zeros(1, dtype=torch.long), lengths[t].cumsum(0)]) for t in range(len(rows))]
    psw = [torch.rand(i.numel()) for i in idx] if weighted else None
    indices = torch.cat(idx).cuda()
    offsets = torch.cat([torch.zeros(1, dtype=torch.long), lengths.view(-1).cumsum(0)]).cuda()
    psw_all = torch.cat(psw).cuda().requires_grad_(True) if weighted else None
    out = m(indices, offsets, psw_all)
    ref_psw = [p.cuda().requires_grad_(True) for p in psw] if weighted else [None] * len(rows)
    ref = torch.cat([b(i.cuda(), o.cuda(), per_sample_weights=p) for b, i, o, p in zip(bags, idx, offs, ref_psw)], 1)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)
    go = torch.randn_like(out)
    out.backward(go)
    ref.backward(go)
    for t, (b, w) in enumerate(zip(bags, m.split_embedding_weights())):
        g = b.weight.grad
        if opt == "EXACT_SGD":
            exp = b.weight.data - lr * g
        else:
            state = g.pow(2).mean(dim=1)
            exp = torch.addcdiv(b.weight.data, g, (state.sqrt() + eps).unsqueeze(1), value=-lr)
            touched = torch.unique(idx[t]).cuda()
            torch.testing.assert_close(m.split_optimizer_states()[t][0][touched], state[touched],
                                       rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(w, exp, rtol=1e-4, atol=1e-4)
    if weighted:
        torch.testing.assert_close(psw_all.grad, torch.cat([p.grad for p in ref_psw]), rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("opt", ["EXACT_SGD", "EXACT_ROWWISE_ADAGRAD"])
@pytest.mark.parametrize("wp", ["FP32", "FP16"])
@pytest.mark.parametrize("shared", [False, True])
def test_tbe_sequence_mode_vs_embedding(opt, wp, shared):
    """PoolingMode.NONE (sequence embeddings): [N, D] rows and the fused update against nn.Embedding +
    autograd + the closed-form optimizer (fbgemm nobag kernels; torchrec EmbeddingCollection's lookup)."""
    from fbgemm_gpu.split_embedding_configs import SparseType
    from fbgemm_gpu.split_table_batched_embeddings_ops import PoolingMode
    torch.manual_seed(5)
    rows, D, B, L = [40, 9, 300], 32, 37, 5
    ftm = [0, 1, 1, 2] if shared else [0, 1, 2]  # feature -> table (a shared table serves two features)
    T = len(ftm)
    lr, eps = 0.1, 1e-4
    m = _tbe(rows, [D] * len(rows), opt, learning_rate=lr, eps=eps, pooling_mode=PoolingMode.NONE,
             feature_table_map=ftm, weights_precision=getattr(SparseType, wp), stochastic_rounding=False)
    m.init_embedding_weights_uniform(-1.0, 1.0)
    embs = [torch.nn.Embedding(r, D).cuda() for r in rows]
    for e, w in zip(embs, m.split_embedding_weights()):
        e.weight.data.copy_(w.float())
    lengths = torch.randint(0, L + 1, (T, B))
    lengths[1] = 0  # an empty feature in the middle
    idx = [torch.randint(0, rows[ftm[f]], (int(lengths[f].sum()),)) for f in range(T)]
    indices = torch.cat(idx).cuda()
    offsets = torch.cat([torch.zeros(1, dtype=torch.long), lengths.view(-1).cumsum(0)]).cuda()
    out = m(indices, offsets)
    assert out.shape == (indices.numel(), D)
    ref = torch.cat([embs[ftm[f]](idx[f].cuda()) for f in range(T)], 0)
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)
    go = torch.randn_like(out)
    out.backward(go)
    ref.backward(go)
    tol = 1e-4 if wp == "FP32" else 1e-2
    for t, (e, w) in enumerate(zip(embs, m.split_embedding_weights())):
        g = e.weight.grad if e.weight.grad is not None else torch.zeros_like(e.weight)
        if opt == "EXACT_SGD":
            exp = e.weight.data - lr * g
        else:
            state = g.pow(2).mean(dim=1)
            exp = torch.addcdiv(e.weight.data, g, (state.sqrt() + eps).unsqueeze(1), value=-lr)
            touched = torch.unique(torch.cat([idx[f] for f in range(T) if ftm[f] == t])).cuda()
            torch.testing.assert_close(m.split_optimizer_states()[t][0][touched], state[touched], rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(w.float(), exp, rtol=tol, atol=tol)


@pytest.mark.parametrize("opt", ["EXACT_SGD", "EXACT_ROWWISE_ADAGRAD"])
@pytest.mark.parametrize("weighted", [False, True])
def test_tbe_pooled_shared_tables_and_feature_requires_grad(opt, weighted):
    """f2: POOLED lookups where several features share a table (feature_table_map with T > T_,
    split_table_batched_embeddings_ops.py:580-640): the forward pools each feature from its table, and the fused
    backward must coalesce the gradients of ALL features of a table into one update per row (one index-transpose
    segment per table).  With per-sample weights, feature_requires_grad masks the weight gradient of the features
    that do not need it (embedding_backward_split_indice_weights_template.cu)."""
    from fbgemm_gpu.split_table_batched_embeddings_ops import PoolingMode
    torch.manual_seed(11)
    rows, dims, B, L = [40, 9, 300], [16, 32, 8], 50, 4
    ftm = [0, 0, 1, 2, 2, 2]  # features 0-1 share table 0, features 3-5 share table 2
    T = len(ftm)
    lr, eps = 0.1, 1e-4
    m = _tbe(rows, dims, opt, learning_rate=lr, eps=eps, pooling_mode=PoolingMode.SUM, feature_table_map=ftm)
    m.init_embedding_weights_uniform(-1.0, 1.0)
    bags = [torch.nn.EmbeddingBag(r, d, mode="sum", sparse=False, include_last_offset=True).cuda()
            for r, d in zip(rows, dims)]
    for b, w in zip(bags, m.split_embedding_weights()):
        b.weight.data.copy_(w)
    lengths = torch.randint(0, L + 1, (T, B))
    idx = [torch.randint(0, rows[ftm[f]], (int(lengths[f].sum()),)) for f in range(T)]
    offs = [torch.cat([torch.zeros(1, dtype=torch.long), lengths[f].cumsum(0)]) for f in range(T)]
    indices = torch.cat(idx).cuda()
    offsets = torch.cat([torch.zeros(1, dtype=torch.long), lengths.view(-1).cumsum(0)]).cuda()
    psw = [torch.rand(i.numel()) for i in idx] if weighted else None
    psw_all = torch.cat(psw).cuda().requires_grad_(True) if weighted else None
    frg = torch.tensor([1, 0, 1, 0, 0, 1], dtype=torch.int32, device="cuda") if weighted else None
    out = m(indices, offsets, psw_all, frg)
    ref_psw = [p.cuda().requires_grad_(True) for p in psw] if weighted else [None] * T
    ref = torch.cat([bags[ftm[f]](idx[f].cuda(), offs[f].cuda(), per_sample_weights=ref_psw[f]) for f in range(T)], 1)
    assert out.shape == (B, sum(dims[t] for t in ftm))
    torch.testing.assert_close(out, ref, rtol=1e-5, atol=1e-5)
    go = torch.randn_like(out)
    out.backward(go)
    ref.backward(go)
    for t, (b, w) in enumerate(zip(bags, m.split_embedding_weights())):
        g = b.weight.grad  # autograd already summed the contributions of every feature of the table
        if opt == "EXACT_SGD":
            exp = b.weight.data - lr * g
        else:
            state = g.pow(2).mean(dim=1)
            exp = torch.addcdiv(b.weight.data, g, (state.sqrt() + eps).unsqueeze(1), value=-lr)
            touched = torch.unique(torch.cat([idx[f] for f in range(T) if ftm[f] == t])).cuda()
            torch.testing.assert_close(m.split_optimizer_states()[t][0][touched], state[touched],
                                       rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(w, exp, rtol=1e-4, atol=1e-4)
    if weighted:
        exp_g = torch.cat([p.grad if int(frg[f]) else torch.zeros_like(p) for f, p in enumerate(ref_psw)])
        torch.testing.assert_close(psw_all.grad, exp_g, rtol=1e-4, atol=1e-4)
        assert float(psw_all.grad[: idx[0].numel()].abs().sum()) > 0  # feature 0 does get its gradient


@pytest.mark.parametrize("wp", ["FP16", "BF16"])
def test_tbe_module_stochastic_rounding_default(wp):
    """16-bit tables round stochastically by default, like the reference's module (:200): many SGD steps of
    1/8 ulp are lost entirely under round-to-nearest and accumulate to the fp32 result (in the mean) here;
    torch.cuda.manual_seed makes the run reproducible."""
    from fbgemm_gpu.split_embedding_configs import SparseType
    rows, D, steps = 512, 64, 64
    ulp = 2.0 ** -10 if wp == "FP16" else 2.0 ** -7
    idx = torch.arange(rows, device="cuda")
    offsets = torch.arange(rows + 1, device="cuda")

    def run(sr, seed):
        torch.cuda.manual_seed(seed)
        m = _tbe([rows], [D], "EXACT_SGD", learning_rate=1.0, weights_precision=getattr(SparseType, wp),
                 stochastic_rounding=sr)
        m.split_embedding_weights()[0].fill_(1.5)
        for _ in range(steps):
            out = m(idx, offsets)
            out.backward(torch.full_like(out, ulp / 8))
        return m.split_embedding_weights()[0].float().clone()

    assert bool((run(False, 1) == 1.5).all())            # round-to-nearest: every update vanishes
    a = run(True, 1)
    exp = 1.5 - steps * ulp / 8                           # the fp32 trajectory: 8 ulps down
    assert abs(float(a.mean()) - exp) < 0.05 * ulp, (float(a.mean()), exp)
    assert float((a - exp).abs().max()) < 20 * ulp           # Binomial(64, 1/8): sigma = 2.6 ulps
    assert torch.equal(run(True, 1), a) and not torch.equal(run(True, 2), a)


def test_kjt_reference_vectors():
    # torchrec/sparse/tests/test_jagged_tensor.py: test_permute :671, test_split :577, test_to_dict :549
    from torchrec.sparse.jagged_tensor import KeyedJaggedTensor
    values = torch.tensor([1.0, 2.0, 3.0, 4.0, 5.0, 6.0, 7.0, 8.0]).cuda()
    lengths = torch.tensor([0, 2, 0, 1, 1, 1, 0, 3, 0], dtype=torch.int32).cuda()
    kjt = KeyedJaggedTensor.from_lengths_sync(keys=["index_0", "index_1", "index_2"], values=values, lengths=lengths)
    p = kjt.permute([1, 0, 2])
    assert p.keys() == ["index_1", "index_0", "index_2"]
    assert p.values().tolist() == [3, 4, 5, 1, 2, 6, 7, 8]
    assert p.lengths().tolist() == [1, 1, 1, 0, 2, 0, 0, 3, 0]
    assert kjt.offsets().tolist() == [0, 0, 2, 2, 3, 4, 5, 5, 8, 8]
    d = kjt.to_dict()
    assert d["index_1"].values().tolist() == [3, 4, 5] and d["index_1"].offsets().tolist() == [0, 1, 2, 3]
    a, b = kjt.split([1, 2])
    assert a.keys() == ["index_0"] and a.values().tolist() == [1, 2] and b.lengths().tolist() == [1, 1, 1, 0, 3, 0]
    dup = kjt.permute([2, 0, 0])  # test_permute_duplicates :704
    assert dup.values().tolist() == [6, 7, 8, 1, 2, 1, 2]


def _make_dlrm(tables, dense_in, dev, meta=True):
    from torchrec import EmbeddingBagCollection
    from torchrec.models.dlrm import DLRM
    ebc = EmbeddingBagCollection(tables, device=torch.device("meta") if meta else dev)
    return DLRM(ebc, dense_in, [32, tables[0].embedding_dim], [32, 16, 1], dense_device=dev)


@pytest.mark.parametrize("sharding", ["table_wise", "row_wise"])
def test_sharded_vs_unsharded_one_step(pg, sharding):
    """global (unsharded nn.EmbeddingBag, SGD) vs local (DMP + fused EXACT_SGD) after ONE step."""
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagConfig
    from torchrec.datasets.random import RandomRecDataset
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    torch.manual_seed(0)
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda")
    rows = [50, 7, 1000, 3]
    D, B, lr = 16, 64, 0.1
    keys = [f"f{i}" for i in range(len(rows))]
    mk = lambda: [EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r, feature_names=[keys[i]])  # noqa
                  for i, r in enumerate(rows)]
    glob = _make_dlrm(mk(), 13, dev, meta=False)
    local = _make_dlrm(mk(), 13, dev, meta=True)
    sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": lr, "optimizer": EmbOptimType.EXACT_SGD})]
    plan = EmbeddingShardingPlanner(Topology(1), {f"t{i}": ParameterConstraints(sharding_types=[sharding])
                                                  for i in range(len(rows))}).plan(local, sharders)
    dmp = DistributedModelParallel(local, env=ShardingEnv.from_process_group(pg), device=dev, plan=plan,
                                   sharders=sharders)
    # copy weights global -> local
    gsd = glob.state_dict()
    lsd = dmp.state_dict()
    from torchrec.distributed.embedding_kernel import local_tensor
    for k, v in lsd.items():
        local_tensor(v).data.copy_(gsd[k])
    batch = next(iter(RandomRecDataset(keys=keys, batch_size=B, hash_sizes=rows, ids_per_feature=3, num_dense=13,
                                       manual_seed=3))).to(dev)
    dense_opt_g = torch.optim.SGD(glob.parameters(), lr=lr)
    dense_opt_l = torch.optim.SGD([p for p in dmp.parameters()], lr=lr)
    for model, opt in ((glob, dense_opt_g), (dmp, dense_opt_l)):
        opt.zero_grad()
        out = model(batch.dense_features, batch.sparse_features)
        out.sum().backward()
        opt.step()
    with torch.no_grad():
        pg_, pl_ = glob(batch.dense_features, batch.sparse_features), dmp(batch.dense_features, batch.sparse_features)
    torch.testing.assert_close(pl_, pg_, rtol=1e-4, atol=1e-4)
    for k, v in dmp.state_dict().items():
        torch.testing.assert_close(local_tensor(v), glob.state_dict()[k], rtol=1e-4, atol=1e-5, msg=k)


def test_quantized_ebc():
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig, KeyedJaggedTensor
    from torchrec.quant.embedding_modules import EmbeddingBagCollection as QuantEBC
    torch.manual_seed(2)
    dev = torch.device("cuda")
    rows, D, B = [100, 30, 500], 128, 33
    keys = ["a", "b", "c"]
    ebc = EmbeddingBagCollection([EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r,
                                                     feature_names=[keys[i]]) for i, r in enumerate(rows)], device=dev)
    ebc.qconfig = torch.quantization.QConfig(activation=torch.quantization.PlaceholderObserver,
                                             weight=torch.quantization.PlaceholderObserver.with_args(dtype=torch.qint8))
    q = QuantEBC.from_float(ebc)
    lengths = torch.randint(0, 4, (len(rows) * B,), dtype=torch.int32)
    vals = torch.cat([torch.randint(0, r, (int(lengths[i * B:(i + 1) * B].sum()),)) for i, r in enumerate(rows)])
    kjt = KeyedJaggedTensor.from_lengths_sync(keys=keys, values=vals.cuda(), lengths=lengths.cuda())
    ref = ebc(kjt).values()
    got = q(kjt)
    assert got.keys() == keys
    # int8 rowwise error: |x - deq(x)| <= scale/2 = range/510 per element, times L <= 3 ids
    torch.testing.assert_close(got.values(), ref, rtol=0.0, atol=3 * 2.0 / 510 * 4 + 2e-2)
    # state_dict layout: [rows, 16-byte-aligned 4 + D] uint8 per table
    sd = q.state_dict()
    assert sd["embedding_bags.t0.weight"].shape == (100, 144)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("exchange", ["peer", "peer_pull", "nccl"])
def test_multi_gpu_sharded_vs_unsharded(world, exchange):
    """TW + RW over `world` GPUs (torchrun): the fused peer-memory exchange and the NCCL collectives both
    reproduce the unsharded model; skipped on a single-GPU box."""
    import subprocess
    import sys
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, B200_PEER_DIST="0" if exchange == "nccl" else "1",
               B200_PEER_PUSH="0" if exchange == "peer_pull" else "1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port",
                        {"peer": "29611", "peer_pull": "29616", "nccl": "29621"}[exchange],
                        os.path.join(here, "dist_parity_worker.py")], capture_output=True, text=True, timeout=300,
                       env=env)
    assert r.returncode == 0 and f"exchange={exchange}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("exchange", ["peer", "nccl"])
def test_multi_gpu_weighted_collection_vs_unsharded(exchange):
    """A WEIGHTED EmbeddingBagCollection (per-id weights) sharded table-wise + row-wise over 2 GPUs: the peer
    exchange carries the weights into the fused lookup and the fused backward (weighted index transpose), six steps
    against the unsharded model; the NCCL path likewise."""
    import subprocess
    import sys
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    env = dict(os.environ, B200_PEER_DIST="0" if exchange == "nccl" else "1", B200_PARITY_WEIGHTED="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29671",
                        os.path.join(here, "dist_parity_worker.py")], capture_output=True, text=True, timeout=300,
                       env=env)
    assert r.returncode == 0 and f"exchange={exchange}" in r.stdout and "weighted=1" in r.stdout, \
        r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world", [1, 2])
def test_reference_torchrec_runs_over_this_fbgemm_gpu(world):
    """The drop-in claim shown, not asserted: the REFERENCE's torchrec (copied unmodified into the git-ignored
    baseline/_ref by tools/install_reference_torchrec.py) builds its own DistributedModelParallel +
    ShardedEmbeddingBagCollection + BatchedFusedEmbeddingBag over this repo's `fbgemm_gpu` (table-wise + row-wise,
    NCCL collectives at world 2) and two fused-SGD steps match its unsharded EmbeddingBagCollection."""
    import subprocess
    import sys
    here = os.path.dirname(os.path.abspath(__file__))
    root = os.path.dirname(here)
    ref = os.path.join(root, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref, "torchrec")):
        pytest.skip("baseline/_ref/torchrec not installed (python tools/install_reference_torchrec.py)")
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    pkg = os.path.join(root, "torchrec-oldfork_b200")
    dropin = os.path.join(pkg, "dropin")
    # reference torchrec first; this repo contributes fbgemm_gpu + b200_sparse only
    path = [ref, dropin if os.path.isdir(os.path.join(dropin, "fbgemm_gpu")) else pkg]
    env = dict(os.environ, PYTHONPATH=os.pathsep.join(path))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29661",
                        os.path.join(here, "dropin_worker.py")], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0 and f"DROPIN OK world={world}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world", [1, 2])
def test_sharded_sequence_embeddings_vs_unsharded(world):
    """EmbeddingCollection under DistributedModelParallel (table-wise + row-wise sequence sharding,
    SequenceEmbeddingAllToAll, fused SGD) equals nn.Embedding + SGD on the global batch."""
    import subprocess
    import sys
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29631",
                        os.path.join(here, "dist_seq_worker.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and f"SEQ PARITY OK world={world}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


@pytest.mark.parametrize("world", [1, 2])
def test_sharded_variable_batch_size_vs_unsharded(world):
    """Table-wise sharded EmbeddingBagCollection with a different batch size on every rank and step
    (KJTAllToAll variable_batch_size, permute_1D recat, ragged pooled all-to-all) equals the unsharded model."""
    import subprocess
    import sys
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    here = os.path.dirname(os.path.abspath(__file__))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", "29651",
                        os.path.join(here, "dist_vb_worker.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and f"VB PARITY OK world={world}" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]


def test_interaction_row_pitch_and_pitched_linear():
    """InteractionArch output sits at a 16-byte row pitch with zero padding, and RowPitchedLinear consumes
    it with the same results/gradients as nn.Linear on a dense copy (fan-in 479 -> 480-wide GEMM)."""
    from b200_sparse import ops
    from torchrec.models.dlrm import InteractionArch
    from torchrec.modules.mlp import RowPitchedLinear
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(3)
    B, F, D, N = 70, 26, 128, 96
    dense = torch.randn(B, D, device="cuda", requires_grad=True)
    sparse = torch.randn(B, F, D, device="cuda", requires_grad=True)
    out = InteractionArch(F)(dense, sparse)
    width = D + (F + 1) * F // 2
    assert out.shape == (B, width) and out.stride() == (480, 1)
    full = out.detach().as_strided((B, 480), (480, 1), out.storage_offset())
    assert torch.count_nonzero(full[:, width:]) == 0
    lin = RowPitchedLinear(width, N, device=torch.device("cuda"))
    assert lin.weight.shape == (N, width) and lin.weight.stride() == (480, 1)
    assert lin.state_dict()["weight"].shape == (N, width)
    ref = torch.nn.Linear(width, N, device="cuda")
    with torch.no_grad():
        ref.weight.copy_(lin.weight)
        ref.bias.copy_(lin.bias)
    y = lin(out)
    gy = torch.randn_like(y)
    y.backward(gy)
    d2 = dense.detach().clone().requires_grad_(True)
    s2 = sparse.detach().clone().requires_grad_(True)
    out2 = ops.interaction_forward(d2.detach(), s2.detach())  # dense layout
    assert out2.is_contiguous()
    torch.testing.assert_close(out.detach(), out2, rtol=0, atol=0)
    x2 = out2.clone().requires_grad_(True)
    y2 = ref(x2)
    y2.backward(gy)
    torch.testing.assert_close(y, y2, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(lin.weight.grad, ref.weight.grad, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(lin.bias.grad, ref.bias.grad, rtol=1e-5, atol=1e-4)
    gd, gs = ops.interaction_backward(d2.detach(), s2.detach(), x2.grad)
    torch.testing.assert_close(dense.grad, gd, rtol=1e-5, atol=1e-4)
    torch.testing.assert_close(sparse.grad, gs, rtol=1e-5, atol=1e-4)
    # an unmarked / dense input takes the plain path
    torch.testing.assert_close(lin(out2), ref(out2), rtol=1e-5, atol=1e-4)
    # an SGD step keeps the pad column at zero and the parameter a view of the pitched buffer
    torch.optim.SGD(lin.parameters(), lr=0.1).step()
    wf = lin.weight.detach().as_strided((N, 480), (480, 1), lin.weight.storage_offset())
    assert torch.count_nonzero(wf[:, width:]) == 0 and lin.weight.stride() == (480, 1)


@pytest.mark.parametrize("B,N", [(1, 4), (70, 96), (4097, 128), (3000, 1024), (513, 512), (100, 20)])
def test_relu_bias_grad_kernel(B, N):
    """One-pass ReLU backward + bias gradient against threshold_backward + sum(0)."""
    from b200_sparse import ops
    torch.manual_seed(B + N)
    y = torch.relu(torch.randn(B, N, device="cuda"))
    gy = torch.randn(B, N, device="cuda")
    g, gb = ops.relu_bias_grad(gy, y)
    ref = gy * (y > 0)
    assert torch.equal(g, ref)  # a select, no arithmetic: bit-exact
    torch.testing.assert_close(gb, ref.sum(0), rtol=1e-5, atol=1e-4 * max(1.0, B ** 0.5))
    g2, gb2 = ops.relu_bias_grad(gy, y)
    assert torch.equal(gb, gb2)  # deterministic
    # strided (row-pitched) inputs are consumed in place
    pad = torch.randn(B, N + 4, device="cuda")
    g3, gb3 = ops.relu_bias_grad(pad[:, :N], y)
    assert torch.equal(g3, pad[:, :N] * (y > 0))


def test_fused_perceptron_matches_linear_relu():
    """MLP with the fused bias+ReLU epilogue / one-pass backward against nn.Linear + relu + autograd
    (reference torchrec/modules/mlp.py semantics), including the row-pitched 479-wide first layer."""
    from torchrec.modules.mlp import MLP, mark_zero_padded
    torch.backends.cuda.matmul.allow_tf32 = False
    torch.manual_seed(11)
    B, K, sizes = 257, 479, [64, 128, 32]
    mlp = MLP(K, sizes, device=torch.device("cuda"))
    refs = []
    k = K
    for i, n in enumerate(sizes):
        lin = torch.nn.Linear(k, n, device="cuda")
        with torch.no_grad():
            lin.weight.copy_(mlp._mlp[i]._linear.weight)
            lin.bias.copy_(mlp._mlp[i]._linear.bias)
        refs.append(lin)
        k = n
    buf = torch.zeros(B, 480, device="cuda")
    buf[:, :K].normal_()
    for pitched in (False, True):
        x = buf[:, :K] if pitched else buf[:, :K].contiguous()
        x = x.detach().requires_grad_(True)
        xin = mark_zero_padded(x) if pitched else x
        y = mlp(xin)
        xr = buf[:, :K].contiguous().requires_grad_(True)
        h = xr
        for lin in refs:
            h = torch.relu(lin(h))
        gy = torch.randn_like(y)
        for m in list(mlp.parameters()) + [p for lin in refs for p in lin.parameters()]:
            m.grad = None
        y.backward(gy)
        h.backward(gy)
        torch.testing.assert_close(y, h, rtol=1e-5, atol=1e-5)
        torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-5)
        for i, lin in enumerate(refs):
            torch.testing.assert_close(mlp._mlp[i]._linear.weight.grad, lin.weight.grad, rtol=1e-4, atol=1e-4)
            torch.testing.assert_close(mlp._mlp[i]._linear.bias.grad, lin.bias.grad, rtol=1e-4, atol=1e-4)
    torch.backends.cuda.matmul.allow_tf32 = True


def test_train_pipeline_sparse_dist_matches_sequential(pg):
    """TrainPipelineSparseDist (H2D | input dist | step, three batches in flight) produces the same losses
    and weights as stepping through the same pinned host batches one by one."""
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig
    from torchrec.datasets.random import RandomRecDataset
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
    from torchrec.distributed.train_pipeline import TrainPipelineSparseDist
    from torchrec.models.dlrm import DLRM, DLRMTrain
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda")
    rows, D, B = [50, 7, 1000, 3], 16, 64
    keys = [f"f{i}" for i in range(len(rows))]

    def build():
        torch.manual_seed(11)
        ebc = EmbeddingBagCollection([EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r,
                                                         feature_names=[keys[i]]) for i, r in enumerate(rows)],
                                     device=torch.device("meta"))
        model = DLRMTrain(ebc, 13, [32, D], [32, 1], dense_device=dev)
        sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": 0.05,
                                                                "optimizer": EmbOptimType.EXACT_ROWWISE_ADAGRAD})]
        dmp = DistributedModelParallel(model, env=ShardingEnv.from_process_group(pg), device=dev, sharders=sharders)
        return dmp, torch.optim.SGD(list(dmp.parameters()), lr=0.05)

    a, opt_a = build()
    b, opt_b = build()
    from torchrec.distributed.embedding_kernel import local_tensor
    res = b.load_state_dict(a.state_dict())  # ShardedTensor -> ShardedTensor for the tables
    assert not res.missing_keys and not res.unexpected_keys, res
    ds = iter(RandomRecDataset(keys=keys, batch_size=B, hash_sizes=rows, ids_per_feature=2, num_dense=13,
                               manual_seed=5))
    host = [next(ds).pin_memory() for _ in range(5)]
    seq_losses = []
    for hb in host:
        opt_a.zero_grad()
        loss, _ = a(hb.to(dev))
        loss.backward()
        opt_a.step()
        seq_losses.append(float(loss))
    pipe = TrainPipelineSparseDist(b, opt_b, dev)
    it = iter(host)
    pipe_losses = []
    while True:
        try:
            out = pipe.progress(it)
        except StopIteration:
            break
        pipe_losses.append(float(out[0]))
    assert len(pipe_losses) == len(host)
    np.testing.assert_allclose(pipe_losses, seq_losses, rtol=1e-5, atol=1e-6)
    for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        torch.testing.assert_close(local_tensor(vb), local_tensor(va), rtol=1e-5, atol=1e-6, msg=k)


def test_train_pipeline_cuda_graph_replay_matches_eager(pg):
    """TrainPipelineSparseDist(cuda_graph=True): after the eager warm-up the step (forward, backward with the fused
    table update) is captured once and replayed with the inputs copied into the graph's buffers; losses and final
    weights equal the eager pipeline's on the same batches, and the replays really happened."""
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig
    from torchrec.datasets.random import RandomRecDataset
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embedding_kernel import local_tensor
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
    from torchrec.distributed.train_pipeline import TrainPipelineSparseDist
    from torchrec.models.dlrm import DLRMTrain
    torch.backends.cuda.matmul.allow_tf32 = False
    dev = torch.device("cuda")
    rows, D, B = [50, 7, 1000, 3], 32, 64
    keys = [f"f{i}" for i in range(len(rows))]

    def build():
        torch.manual_seed(13)
        ebc = EmbeddingBagCollection([EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r,
                                                         feature_names=[keys[i]]) for i, r in enumerate(rows)],
                                     device=torch.device("meta"))
        model = DLRMTrain(ebc, 13, [32, D], [32, 1], dense_device=dev)
        sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": 0.05,
                                                                "optimizer": EmbOptimType.EXACT_ROWWISE_ADAGRAD})]
        dmp = DistributedModelParallel(model, env=ShardingEnv.from_process_group(pg), device=dev, sharders=sharders)
        return dmp, torch.optim.SGD(list(dmp.parameters()), lr=0.05)

    a, opt_a = build()
    b, opt_b = build()
    res = b.load_state_dict(a.state_dict())
    assert not res.missing_keys and not res.unexpected_keys, res
    ds = iter(RandomRecDataset(keys=keys, batch_size=B, hash_sizes=rows, ids_per_feature=2, num_dense=13,
                               manual_seed=9))
    host = [next(ds).pin_memory() for _ in range(12)]  # fixed pooling factor: every step has the same signature

    def run(model, opt, graph):
        pipe = TrainPipelineSparseDist(model, opt, dev, cuda_graph=graph, graph_warmup=3)
        it, losses = iter(host), []
        while True:
            try:
                out = pipe.progress(it)
            except StopIteration:
                break
            losses.append(float(out[0]))
        return losses, pipe

    eager, _ = run(a, opt_a, False)
    graphed, pipe = run(b, opt_b, True)
    assert len(eager) == len(graphed) == len(host)
    assert pipe.graph_replays >= 6, pipe.graph_replays
    np.testing.assert_allclose(graphed, eager, rtol=1e-5, atol=1e-6)
    for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        torch.testing.assert_close(local_tensor(vb), local_tensor(va), rtol=1e-5, atol=1e-6, msg=k)
