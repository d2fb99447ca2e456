"""Sequence (non-pooled) embedding sharding (reference torchrec/distributed/sharding/
sequence_sharding.py: SequenceShardingContext :19, BaseSequenceEmbeddingDist :45;
tw_sequence_sharding.py: TwSequenceEmbeddingDist :74, TwSequenceEmbeddingSharding :118;
rw_sequence_sharding.py: RwSequenceEmbeddingDist :33, RwSequenceEmbeddingSharding :77).

Ids travel exactly as for pooled lookups (KJT all-to-all; row-wise adds block_bucketize with the
unbucketize permutation kept); every looked-up row then goes back to the rank that owns its sample
through SequenceEmbeddingAllToAll, and row-wise results are put back into the original id order."""
from __future__ import annotations

import abc
from dataclasses import dataclass, field
from typing import Any, Dict, Generic, List, Optional, TypeVar

import torch
import torch.distributed as dist
from torch import nn

from torchrec.distributed.dist_data import SequenceEmbeddingAllToAll
from torchrec.distributed.embedding_lookup import GroupedEmbeddingsLookup
from torchrec.distributed.embedding_sharding import BaseEmbeddingLookup, BaseSparseFeaturesDist
from torchrec.distributed.embedding_types import SparseFeatures
from torchrec.distributed.sharding.rw_sharding import BaseRwEmbeddingSharding, RwSparseFeaturesDist
from torchrec.distributed.sharding.tw_sharding import BaseTwEmbeddingSharding, TwSparseFeaturesDist
from torchrec.distributed.types import Awaitable
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

T = TypeVar("T")


@dataclass
class SequenceShardingContext:
    """What the output dist needs from the input dist of the same batch."""
    features_before_input_dist: Optional[KeyedJaggedTensor] = None
    input_splits: List[int] = field(default_factory=list)    # ids sent to each rank
    output_splits: List[int] = field(default_factory=list)   # ids received from each rank
    unbucketize_permute_tensor: Optional[torch.Tensor] = None
    lengths_after_input_dist: Optional[torch.Tensor] = None  # [F_local, W * B_local]

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        if self.features_before_input_dist is not None:
            self.features_before_input_dist.record_stream(stream)
        for t in (self.unbucketize_permute_tensor, self.lengths_after_input_dist):
            if t is not None:
                t.record_stream(stream)


class BaseSequenceEmbeddingDist(abc.ABC, nn.Module, Generic[T]):
    @abc.abstractmethod
    def forward(self, local_embs: T, sharding_ctx: SequenceShardingContext) -> Awaitable[torch.Tensor]:
        ...


class TwSequenceEmbeddingDist(BaseSequenceEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup, features_per_rank: List[int],
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._dist = SequenceEmbeddingAllToAll(pg, features_per_rank, device)

    def forward(self, local_embs: torch.Tensor, sharding_ctx: SequenceShardingContext) -> Awaitable[torch.Tensor]:
        return self._dist(local_embs, lengths=sharding_ctx.lengths_after_input_dist,
                          input_splits=sharding_ctx.input_splits, output_splits=sharding_ctx.output_splits,
                          unbucketize_permute_tensor=None)


class TwSequenceEmbeddingSharding(BaseTwEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[SparseFeatures]:
        return TwSparseFeaturesDist(self._pg, self._id_list_features_per_rank(),
                                    self._id_score_list_features_per_rank(), device or self._device)

    def create_lookup(self, device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None,
                      feature_processor=None) -> BaseEmbeddingLookup:
        assert feature_processor is None
        return GroupedEmbeddingsLookup(self._grouped_embedding_configs, device=device or self._device, pg=self._pg,
                                       fused_params=fused_params)

    def create_output_dist(self, device: Optional[torch.device] = None) -> BaseSequenceEmbeddingDist[torch.Tensor]:
        return TwSequenceEmbeddingDist(self._pg, self._id_list_features_per_rank(), device or self._device)


class RwSequenceEmbeddingDist(BaseSequenceEmbeddingDist[torch.Tensor]):
    def __init__(self, pg: dist.ProcessGroup, num_features: int, device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._dist = SequenceEmbeddingAllToAll(pg, [num_features] * dist.get_world_size(pg), device)

    def forward(self, local_embs: torch.Tensor, sharding_ctx: SequenceShardingContext) -> Awaitable[torch.Tensor]:
        return self._dist(local_embs, lengths=sharding_ctx.lengths_after_input_dist,
                          input_splits=sharding_ctx.input_splits, output_splits=sharding_ctx.output_splits,
                          unbucketize_permute_tensor=sharding_ctx.unbucketize_permute_tensor)


class RwSequenceEmbeddingSharding(BaseRwEmbeddingSharding):
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[SparseFeatures]:
        return RwSparseFeaturesDist(self._pg, self._get_id_list_features_num(), 0,
                                    self._get_id_list_features_hash_sizes(), [], device or self._device,
                                    is_sequence=True)

    def create_lookup(self, device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None,
                      feature_processor=None) -> BaseEmbeddingLookup:
        assert feature_processor is None
        return GroupedEmbeddingsLookup(self._grouped_embedding_configs, device=device or self._device, pg=self._pg,
                                       fused_params=fused_params)

    def create_output_dist(self, device: Optional[torch.device] = None) -> BaseSequenceEmbeddingDist[torch.Tensor]:
        return RwSequenceEmbeddingDist(self._pg, self._get_id_list_features_num(), device or self._device)
