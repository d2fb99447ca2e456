"""2-GPU diagnostic: pooled outputs and one-step updates of the peer-memory exchange vs the NCCL path."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "torchrec-oldfork_b200"))
from torchrec.distributed.embedding_kernel import local_tensor  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    from fbgemm_gpu.split_embedding_configs import EmbOptimType
    from torchrec import EmbeddingBagCollection, EmbeddingBagConfig
    from torchrec.datasets.random import RandomRecDataset
    from torchrec.distributed import DistributedModelParallel, ShardingEnv
    from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder, ShardedEmbeddingBagCollection
    from torchrec.distributed.planner import EmbeddingShardingPlanner, ParameterConstraints, Topology
    rows = [50, 7, 1001, 3, 5, 300]
    sharding = ["table_wise", "row_wise", "row_wise", "table_wise", "row_wise", "table_wise"]
    D, B = 16, 32
    keys = [f"f{i}" for i in range(len(rows))]

    def build(peer):
        os.environ["B200_PEER_DIST"] = "1" if peer else "0"
        torch.manual_seed(7)
        ebc = EmbeddingBagCollection([EmbeddingBagConfig(name=f"t{i}", embedding_dim=D, num_embeddings=r,
                                                         feature_names=[keys[i]]) for i, r in enumerate(rows)],
                                     device=torch.device("meta"))
        sharders = [EmbeddingBagCollectionSharder(fused_params={"learning_rate": 0.1,
                                                                "optimizer": EmbOptimType.EXACT_SGD})]
        plan = EmbeddingShardingPlanner(Topology(world), {f"t{i}": ParameterConstraints(sharding_types=[s])
                                                          for i, s in enumerate(sharding)}).plan(ebc, sharders)
        m = DistributedModelParallel(ebc, env=ShardingEnv.from_process_group(dist.group.WORLD), device=dev,
                                     plan=plan, sharders=sharders)
        return m

    a, b = build(True), build(False)
    with torch.no_grad():
        for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
            local_tensor(vb).copy_(local_tensor(va))
    batch = next(iter(RandomRecDataset(keys=keys, batch_size=B, hash_sizes=rows, ids_per_feature=2, num_dense=13,
                                       manual_seed=100 + rank))).to(dev)
    os.environ["B200_PEER_DIST"] = "1"   # read at the first forward
    oa = a(batch.sparse_features)
    os.environ["B200_PEER_DIST"] = "0"
    ob = b(batch.sparse_features)
    oa = oa.wait() if hasattr(oa, "wait") else oa
    ob = ob.wait() if hasattr(ob, "wait") else ob
    sa = next(m for m in a.module.modules() if isinstance(m, ShardedEmbeddingBagCollection))
    print(rank, "peer used:", sa._peer is not None, "keys", oa.keys() == ob.keys(), oa.keys())
    for k in oa.keys():
        da = oa[k]
        db = ob[k]
        print(rank, k, "max diff", float((da - db).detach().abs().max()), "peer abs max", float(da.detach().abs().max()))
    g = torch.randn_like(ob.values())
    oa.values().backward(g)
    ob.values().backward(g)
    for (k, va), (_, vb) in zip(a.state_dict().items(), b.state_dict().items()):
        print(rank, k, "weight diff after step", float((va - vb).abs().max()))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
