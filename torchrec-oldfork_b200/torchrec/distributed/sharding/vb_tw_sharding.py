"""Table-wise sharding when every rank may bring a different batch size each step (reference
torchrec/distributed/sharding/vb_tw_sharding.py: VariableBatchTwSparseFeaturesDist :33,
VariableBatchTwPooledEmbeddingDist :83, VariableBatchTwPooledEmbeddingSharding :103).

Same tables, lookup and kernels as TwPooledEmbeddingSharding; only the two exchanges change.
in : KJTAllToAll(variable_batch_size=True) - batch sizes first, recat expanded to one slot per
      (feature, sample) by expand_into_jagged_permute, values reordered by permute_1D_sparse_data;
      the lookup then sees a KJT of stride sum_r B_r.
out: alltoall_pooled with send splits D_local * B_r (rank r gets back exactly its B_r samples)."""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist

from torchrec.distributed.dist_data import PooledEmbeddingsAllToAll
from torchrec.distributed.sharding.tw_sharding import TwPooledEmbeddingSharding, TwSparseFeaturesDist
from torchrec.distributed.sharding.vb_sharding import BaseVariableBatchEmbeddingDist, VariableBatchShardingContext
from torchrec.distributed.types import Awaitable


class VariableBatchTwSparseFeaturesDist(TwSparseFeaturesDist):
    def __init__(self, pg: dist.ProcessGroup, id_list_features_per_rank: List[int],
                 id_score_list_features_per_rank: List[int], device: Optional[torch.device] = None) -> None:
        super().__init__(pg, id_list_features_per_rank, id_score_list_features_per_rank, device,
                         variable_batch_size=True)


class VariableBatchTwPooledEmbeddingDist(BaseVariableBatchEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup, dim_sum_per_rank: List[int],
                 device: Optional[torch.device] = None, callbacks=None) -> None:
        super().__init__()
        self._dist = PooledEmbeddingsAllToAll(pg, dim_sum_per_rank, device, callbacks)

    def forward(self, local_embs: torch.Tensor, sharding_ctx: VariableBatchShardingContext
                ) -> Awaitable[torch.Tensor]:
        assert local_embs.shape[0] == sharding_ctx.global_batch_size, (
            f"{local_embs.shape[0]} pooled rows for batch sizes {sharding_ctx.batch_size_per_rank}")
        return self._dist(local_embs, batch_size_per_rank=sharding_ctx.batch_size_per_rank)


class VariableBatchTwPooledEmbeddingSharding(TwPooledEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> VariableBatchTwSparseFeaturesDist:
        return VariableBatchTwSparseFeaturesDist(self._pg, self._id_list_features_per_rank(),
                                                 self._id_score_list_features_per_rank(), device or self._device)

    def create_output_dist(self, device: Optional[torch.device] = None) -> VariableBatchTwPooledEmbeddingDist:
        return VariableBatchTwPooledEmbeddingDist(self._pg, self._dim_sum_per_rank(), device or self._device)
