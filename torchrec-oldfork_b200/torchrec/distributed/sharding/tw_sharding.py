"""Table-wise sharding (reference torchrec/distributed/sharding/tw_sharding.py:
BaseTwEmbeddingSharding :53, TwSparseFeaturesDist :231, TwPooledEmbeddingDist :278,
TwPooledEmbeddingSharding :313): a table lives whole on ranks[0]; ids travel by KJT all-to-all,
pooled rows come back by PooledEmbeddingsAllToAll."""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from torchrec.distributed.dist_data import PooledEmbeddingsAllToAll
from torchrec.distributed.embedding_lookup import GroupedPooledEmbeddingsLookup
from torchrec.distributed.embedding_sharding import (
    BaseEmbeddingDist, BaseEmbeddingLookup, BaseSparseFeaturesDist, EmbeddingSharding,
    SparseFeaturesAllToAll, group_tables)
from torchrec.distributed.embedding_types import (
    EmbeddingComputeKernel, GroupedEmbeddingConfig, ShardedEmbeddingTable, SparseFeatures)
from torchrec.distributed.types import Awaitable, ParameterSharding, ShardMetadata, ShardingEnv
from torchrec.modules.embedding_configs import EmbeddingTableConfig

EmbeddingInfo = Tuple[EmbeddingTableConfig, ParameterSharding, torch.Tensor]


class BaseTwEmbeddingSharding(EmbeddingSharding[SparseFeatures, torch.Tensor]):
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
        tables_per_rank: List[List[ShardedEmbeddingTable]] = [[] for _ in range(self._world_size)]
        for cfg, ps, _param in embedding_configs:
            assert ps.ranks, f"table-wise sharding of {cfg.name} needs a rank"
            owner = ps.ranks[0]
            meta = ShardMetadata([0, 0], [cfg.num_embeddings, cfg.embedding_dim], f"rank:{owner}/cuda:{owner}")
            tables_per_rank[owner].append(ShardedEmbeddingTable(
                num_embeddings=cfg.num_embeddings, embedding_dim=cfg.embedding_dim, name=cfg.name,
                embedding_names=cfg.embedding_names, data_type=cfg.data_type, feature_names=cfg.feature_names,
                pooling=cfg.pooling, is_weighted=cfg.is_weighted, has_feature_processor=cfg.has_feature_processor,
                local_rows=cfg.num_embeddings, local_cols=cfg.embedding_dim,
                compute_kernel=EmbeddingComputeKernel(ps.compute_kernel), local_metadata=meta,
                global_metadata=[meta], weight_init_max=cfg.weight_init_max, weight_init_min=cfg.weight_init_min))
        return tables_per_rank

    def _dim_sum_per_rank(self) -> List[int]:
        return [sum(c.dim_sum() for c in plain) + sum(c.dim_sum() for c in scored)
                for plain, scored in zip(self._grouped_embedding_configs_per_rank,
                                         self._score_grouped_embedding_configs_per_rank)]

    def embedding_dims(self) -> List[int]:
        return [d for plain, scored in zip(self._grouped_embedding_configs_per_rank,
                                           self._score_grouped_embedding_configs_per_rank)
                for c in plain + scored for d in c.embedding_dims()]

    def embedding_names(self) -> List[str]:
        return [n for plain, scored in zip(self._grouped_embedding_configs_per_rank,
                                           self._score_grouped_embedding_configs_per_rank)
                for c in plain + scored for n in c.embedding_names()]

    def id_list_feature_names(self) -> List[str]:
        return [f for plain in self._grouped_embedding_configs_per_rank for c in plain for f in c.feature_names()]

    def id_score_list_feature_names(self) -> List[str]:
        return [f for scored in self._score_grouped_embedding_configs_per_rank for c in scored
                for f in c.feature_names()]

    def _id_list_features_per_rank(self) -> List[int]:
        return [sum(c.num_features() for c in plain) for plain in self._grouped_embedding_configs_per_rank]

    def _id_score_list_features_per_rank(self) -> List[int]:
        return [sum(c.num_features() for c in scored) for scored in self._score_grouped_embedding_configs_per_rank]


class TwSparseFeaturesDist(BaseSparseFeaturesDist[SparseFeatures]):
    def __init__(self, pg: dist.ProcessGroup, id_list_features_per_rank: List[int],
                 id_score_list_features_per_rank: List[int], device: Optional[torch.device] = None,
                 variable_batch_size: bool = False) -> None:
        super().__init__()
        self._dist = SparseFeaturesAllToAll(pg, id_list_features_per_rank, id_score_list_features_per_rank, device,
                                            variable_batch_size=variable_batch_size)

    def forward(self, sparse_features: SparseFeatures) -> Awaitable[Awaitable[SparseFeatures]]:
        return self._dist(sparse_features)


class TwPooledEmbeddingDist(BaseEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup, dim_sum_per_rank: List[int],
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._dist = PooledEmbeddingsAllToAll(pg, dim_sum_per_rank, device)

    def forward(self, local_embs: torch.Tensor) -> Awaitable[torch.Tensor]:
        return self._dist(local_embs)


class TwPooledEmbeddingSharding(BaseTwEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[SparseFeatures]:
        return TwSparseFeaturesDist(self._pg, self._id_list_features_per_rank(),
                                    self._id_score_list_features_per_rank(), device or self._device)

    def create_lookup(self, device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None,
                      feature_processor=None) -> BaseEmbeddingLookup:
        return GroupedPooledEmbeddingsLookup(self._grouped_embedding_configs, self._score_grouped_embedding_configs,
                                             device=device or self._device, pg=self._pg,
                                             fused_params=fused_params, feature_processor=feature_processor)

    def create_output_dist(self, device: Optional[torch.device] = None) -> BaseEmbeddingDist[torch.Tensor]:
        return TwPooledEmbeddingDist(self._pg, self._dim_sum_per_rank(), device or self._device)
