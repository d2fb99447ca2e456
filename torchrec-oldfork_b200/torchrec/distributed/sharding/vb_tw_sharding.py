"""Table-wise sharding with a different batch size on every rank (reference
torchrec/distributed/sharding/vb_tw_sharding.py: VariableBatchTwSparseFeaturesDist :33,
VariableBatchTwPooledEmbeddingDist :83, VariableBatchTwPooledEmbeddingSharding :103).

Input dist: KJTAllToAll(variable_batch_size=True) - batch sizes exchanged first, recat expanded by
expand_into_jagged_permute, values reordered by permute_1D_sparse_data.  The lookup sees a KJT of
stride sum_r B_r.  Output dist: pooled all-to-all whose send splits follow batch_size_per_rank."""
from __future__ import annotations

from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist

from torchrec.distributed.dist_data import PooledEmbeddingsAllToAll
from torchrec.distributed.embedding_lookup import GroupedPooledEmbeddingsLookup
from torchrec.distributed.embedding_sharding import (
    BaseEmbeddingLookup, BaseSparseFeaturesDist, SparseFeaturesAllToAll)
from torchrec.distributed.embedding_types import SparseFeatures
from torchrec.distributed.sharding.tw_sharding import BaseTwEmbeddingSharding
from torchrec.distributed.sharding.vb_sharding import BaseVariableBatchEmbeddingDist, VariableBatchShardingContext
from torchrec.distributed.types import Awaitable


class VariableBatchTwSparseFeaturesDist(BaseSparseFeaturesDist[SparseFeatures]):
    def __init__(self, pg: dist.ProcessGroup, id_list_features_per_rank: List[int],
                 id_score_list_features_per_rank: List[int], device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._dist = SparseFeaturesAllToAll(pg, id_list_features_per_rank, id_score_list_features_per_rank, device,
                                            variable_batch_size=True)

    def forward(self, sparse_features: SparseFeatures) -> Awaitable[Awaitable[SparseFeatures]]:
        return self._dist(sparse_features)


class VariableBatchTwPooledEmbeddingDist(BaseVariableBatchEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup, dim_sum_per_rank: List[int],
                 device: Optional[torch.device] = None, callbacks=None) -> None:
        super().__init__()
        self._dist = PooledEmbeddingsAllToAll(pg, dim_sum_per_rank, device, callbacks)

    def forward(self, local_embs: torch.Tensor, sharding_ctx: VariableBatchShardingContext
                ) -> Awaitable[torch.Tensor]:
        return self._dist(local_embs, batch_size_per_rank=sharding_ctx.batch_size_per_rank)


class VariableBatchTwPooledEmbeddingSharding(BaseTwEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[SparseFeatures]:
        return VariableBatchTwSparseFeaturesDist(self._pg, self._id_list_features_per_rank(),
                                                 self._id_score_list_features_per_rank(), device or self._device)

    def create_lookup(self, device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None,
                      feature_processor=None) -> BaseEmbeddingLookup:
        return GroupedPooledEmbeddingsLookup(self._grouped_embedding_configs, self._score_grouped_embedding_configs,
                                             device=device or self._device, pg=self._pg,
                                             fused_params=fused_params, feature_processor=feature_processor)

    def create_output_dist(self, device: Optional[torch.device] = None
                           ) -> BaseVariableBatchEmbeddingDist[torch.Tensor]:
        return VariableBatchTwPooledEmbeddingDist(self._pg, self._dim_sum_per_rank(), device or self._device)
