"""Row-wise sharding (reference torchrec/distributed/sharding/rw_sharding.py:
BaseRwEmbeddingSharding :45, RwSparseFeaturesDist :193, RwPooledEmbeddingDist :314,
RwPooledEmbeddingSharding :344): every table is cut in W contiguous row blocks of
ceil(rows / W); ids are bucketized by id // block (block_bucketize_sparse_features), exchanged by
KJT all-to-all with splits [F] * W, looked up against the local block, and the partial pooled rows
are summed across ranks by reduce-scatter along the batch."""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from torchrec.distributed.dist_data import PooledEmbeddingsReduceScatter
from torchrec.distributed.embedding_lookup import GroupedPooledEmbeddingsLookup
from torchrec.distributed.embedding_sharding import (
    BaseEmbeddingDist, BaseEmbeddingLookup, BaseSparseFeaturesDist, EmbeddingSharding,
    SparseFeaturesAllToAll, bucketize_kjt_before_all2all, group_tables)
from torchrec.distributed.embedding_types import (
    EmbeddingComputeKernel, ShardedEmbeddingTable, SparseFeatures)
from torchrec.distributed.types import Awaitable, ParameterSharding, ShardMetadata, ShardingEnv
from torchrec.modules.embedding_configs import EmbeddingTableConfig

EmbeddingInfo = Tuple[EmbeddingTableConfig, ParameterSharding, torch.Tensor]


def row_wise_shard_sizes(rows: int, world_size: int) -> List[int]:
    """ceil blocks with a short last block and possibly zero-row trailing ranks
    (planner/enumerators.py:277-311; 5 rows over 4 ranks -> 2, 2, 1, 0)."""
    block = (rows + world_size - 1) // world_size
    return [max(0, min(block, rows - r * block)) for r in range(world_size)]


class BaseRwEmbeddingSharding(EmbeddingSharding[SparseFeatures, torch.Tensor]):
    def __init__(self, embedding_configs: List[EmbeddingInfo], env: ShardingEnv,
                 device: Optional[torch.device] = None) -> None:
        self._env, self._device = env, device
        self._pg: Optional[dist.ProcessGroup] = env.process_group
        self._world_size, self._rank = env.world_size, env.rank
        per_rank = self._shard(embedding_configs)
        self._grouped_embedding_configs_per_rank, self._score_grouped_embedding_configs_per_rank = \
            group_tables(per_rank)
        self._grouped_embedding_configs = self._grouped_embedding_configs_per_rank[self._rank]
        self._score_grouped_embedding_configs = self._score_grouped_embedding_configs_per_rank[self._rank]

    def _shard(self, embedding_configs: List[EmbeddingInfo]) -> List[List[ShardedEmbeddingTable]]:
        W = self._world_size
        tables_per_rank: List[List[ShardedEmbeddingTable]] = [[] for _ in range(W)]
        for cfg, ps, _param in embedding_configs:
            if ps.sharding_spec is not None:
                sizes = [s.shard_sizes[0] for s in ps.sharding_spec.shards]
            else:
                sizes = row_wise_shard_sizes(cfg.num_embeddings, W)
            assert len(sizes) == W and sum(sizes) == cfg.num_embeddings
            offs = [sum(sizes[:r]) for r in range(W)]
            metas = [ShardMetadata([offs[r], 0], [sizes[r], cfg.embedding_dim], f"rank:{r}/cuda:{r}")
                     for r in range(W)]
            for r in range(W):
                tables_per_rank[r].append(ShardedEmbeddingTable(
                    num_embeddings=cfg.num_embeddings, embedding_dim=cfg.embedding_dim, name=cfg.name,
                    embedding_names=cfg.embedding_names, data_type=cfg.data_type, feature_names=cfg.feature_names,
                    pooling=cfg.pooling, is_weighted=cfg.is_weighted,
                    has_feature_processor=cfg.has_feature_processor, local_rows=sizes[r],
                    local_cols=cfg.embedding_dim, compute_kernel=EmbeddingComputeKernel(ps.compute_kernel),
                    local_metadata=metas[r], global_metadata=metas, weight_init_max=cfg.weight_init_max,
                    weight_init_min=cfg.weight_init_min))
        return tables_per_rank

    def embedding_dims(self) -> List[int]:
        return [d for c in self._grouped_embedding_configs + self._score_grouped_embedding_configs
                for d in c.embedding_dims()]

    def embedding_names(self) -> List[str]:
        return [n for c in self._grouped_embedding_configs + self._score_grouped_embedding_configs
                for n in c.embedding_names()]

    def id_list_feature_names(self) -> List[str]:
        return [f for c in self._grouped_embedding_configs for f in c.feature_names()]

    def id_score_list_feature_names(self) -> List[str]:
        return [f for c in self._score_grouped_embedding_configs for f in c.feature_names()]

    def _get_id_list_features_num(self) -> int:
        return sum(c.num_features() for c in self._grouped_embedding_configs)

    def _get_id_score_list_features_num(self) -> int:
        return sum(c.num_features() for c in self._score_grouped_embedding_configs)

    def _get_id_list_features_hash_sizes(self) -> List[int]:
        return [h for c in self._grouped_embedding_configs for h in c.feature_hash_sizes()]

    def _get_id_score_list_features_hash_sizes(self) -> List[int]:
        return [h for c in self._score_grouped_embedding_configs for h in c.feature_hash_sizes()]


class RwSparseFeaturesDist(BaseSparseFeaturesDist[SparseFeatures]):
    def __init__(self, pg: dist.ProcessGroup, num_id_list_features: int, num_id_score_list_features: int,
                 id_list_feature_hash_sizes: List[int], id_score_list_feature_hash_sizes: List[int],
                 device: Optional[torch.device] = None, is_sequence: bool = False,
                 has_feature_processor: bool = False) -> None:
        super().__init__()
        W = dist.get_world_size(pg)
        self._world_size = W
        self._n_list, self._n_score = num_id_list_features, num_id_score_list_features
        blocks = lambda hs: torch.tensor([(h + W - 1) // W for h in hs], device=device, dtype=torch.int32)  # noqa: E731
        self.register_buffer("_id_list_feature_block_sizes_tensor", blocks(id_list_feature_hash_sizes))
        self.register_buffer("_id_score_list_feature_block_sizes_tensor", blocks(id_score_list_feature_hash_sizes))
        self._dist = SparseFeaturesAllToAll(pg, W * [num_id_list_features], W * [num_id_score_list_features], device)
        self._is_sequence, self._has_feature_processor = is_sequence, has_feature_processor
        self.unbucketize_permute_tensor: Optional[torch.Tensor] = None

    def forward(self, sparse_features: SparseFeatures) -> Awaitable[Awaitable[SparseFeatures]]:
        id_list = id_score = None
        if self._n_list > 0:
            assert sparse_features.id_list_features is not None
            id_list, self.unbucketize_permute_tensor = bucketize_kjt_before_all2all(
                sparse_features.id_list_features, self._world_size, self._id_list_feature_block_sizes_tensor,
                output_permute=self._is_sequence, bucketize_pos=self._has_feature_processor)
        if self._n_score > 0:
            assert sparse_features.id_score_list_features is not None
            id_score, _ = bucketize_kjt_before_all2all(
                sparse_features.id_score_list_features, self._world_size,
                self._id_score_list_feature_block_sizes_tensor)
        return self._dist(SparseFeatures(id_list_features=id_list, id_score_list_features=id_score))


class RwPooledEmbeddingDist(BaseEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup) -> None:
        super().__init__()
        self._dist = PooledEmbeddingsReduceScatter(pg)

    def forward(self, local_embs: torch.Tensor) -> Awaitable[torch.Tensor]:
        return self._dist(local_embs)


class RwPooledEmbeddingSharding(BaseRwEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[SparseFeatures]:
        return RwSparseFeaturesDist(self._pg, self._get_id_list_features_num(),
                                    self._get_id_score_list_features_num(),
                                    self._get_id_list_features_hash_sizes(),
                                    self._get_id_score_list_features_hash_sizes(), device or self._device)

    def create_lookup(self, device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None,
                      feature_processor=None) -> BaseEmbeddingLookup:
        return GroupedPooledEmbeddingsLookup(self._grouped_embedding_configs, self._score_grouped_embedding_configs,
                                             device=device or self._device, pg=self._pg,
                                             fused_params=fused_params, feature_processor=feature_processor)

    def create_output_dist(self, device: Optional[torch.device] = None) -> BaseEmbeddingDist[torch.Tensor]:
        return RwPooledEmbeddingDist(self._pg)
