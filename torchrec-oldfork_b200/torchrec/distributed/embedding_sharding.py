"""Sparse-feature distribution helpers (reference torchrec/distributed/embedding_sharding.py:
bucketize_kjt_before_all2all :121, SparseFeaturesAllToAll :187, group_tables :393, ABCs :581-654)."""
from __future__ import annotations

import abc
from typing import Dict, Generic, List, Optional, Tuple, TypeVar

import torch
import torch.distributed as dist
from torch import nn

from torchrec.distributed.dist_data import KJTAllToAll
from torchrec.distributed.embedding_types import (
    GroupedEmbeddingConfig, ShardedEmbeddingTable, SparseFeatures, SparseFeaturesList)
from torchrec.distributed.types import Awaitable, NoWait
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

F = TypeVar("F")
T = TypeVar("T")


def bucketize_kjt_before_all2all(kjt: KeyedJaggedTensor, num_buckets: int, block_sizes: torch.Tensor,
                                 output_permute: bool = False, bucketize_pos: bool = False
                                 ) -> Tuple[KeyedJaggedTensor, Optional[torch.Tensor]]:
    """Row-wise sharding: route every id to bucket id // block_size and rebase it to id % block_size.
    The result has keys * num_buckets in bucket-major order, ready for KJTAllToAll(splits=[F] * W)."""
    n_feat = len(kjt.keys())
    assert block_sizes.numel() == n_feat, \
        f"Expecting block sizes for {n_feat} features, but {block_sizes.numel()} received."
    new_lengths, new_ids, new_w, pos, unbucketize = torch.ops.fbgemm.block_bucketize_sparse_features(
        kjt.lengths().view(-1), kjt.values(), bucketize_pos, output_permute,
        block_sizes.to(kjt.values().dtype), num_buckets, kjt.weights_or_none())
    out = KeyedJaggedTensor(keys=kjt.keys() * num_buckets, values=new_ids,
                            weights=pos if bucketize_pos else new_w, lengths=new_lengths.view(-1),
                            stride=kjt.stride())
    return out, unbucketize


class _SparseFeaturesIndicesAwaitable(Awaitable[SparseFeatures]):
    def __init__(self, id_list, id_score_list) -> None:
        super().__init__()
        self._a, self._b = id_list, id_score_list

    def _wait_impl(self) -> SparseFeatures:
        return SparseFeatures(id_list_features=self._a.wait() if self._a is not None else None,
                              id_score_list_features=self._b.wait() if self._b is not None else None)


class _SparseFeaturesLengthsAwaitable(Awaitable[Awaitable[SparseFeatures]]):
    def __init__(self, id_list, id_score_list) -> None:
        super().__init__()
        self._a, self._b = id_list, id_score_list

    def _wait_impl(self) -> Awaitable[SparseFeatures]:
        return _SparseFeaturesIndicesAwaitable(self._a.wait() if self._a is not None else None,
                                               self._b.wait() if self._b is not None else None)


class SparseFeaturesAllToAll(nn.Module):
    """KJTAllToAll for the unweighted (id_list) and weighted (id_score_list) halves of SparseFeatures."""

    def __init__(self, pg: dist.ProcessGroup, id_list_features_per_rank: List[int],
                 id_score_list_features_per_rank: List[int], device: Optional[torch.device] = None,
                 stagger: int = 1, variable_batch_size: bool = False) -> None:
        super().__init__()
        self._id_list = KJTAllToAll(pg, id_list_features_per_rank, device, stagger, variable_batch_size)
        self._id_score_list = KJTAllToAll(pg, id_score_list_features_per_rank, device, stagger,
                                          variable_batch_size)

    def forward(self, sparse_features: SparseFeatures) -> Awaitable[Awaitable[SparseFeatures]]:
        a = sparse_features.id_list_features
        b = sparse_features.id_score_list_features
        return _SparseFeaturesLengthsAwaitable(self._id_list(a) if a is not None else None,
                                               self._id_score_list(b) if b is not None else None)


def group_tables(tables_per_rank: List[List[ShardedEmbeddingTable]]
                 ) -> Tuple[List[List[GroupedEmbeddingConfig]], List[List[GroupedEmbeddingConfig]]]:
    """Per rank: group tables that can share one table-batched kernel, i.e. equal
    (data_type, pooling, is_weighted, has_feature_processor, compute_kernel).  Unweighted groups
    come first, score (weighted) groups second, as in the reference (:412-476)."""
    grouped: List[List[GroupedEmbeddingConfig]] = []
    score_grouped: List[List[GroupedEmbeddingConfig]] = []
    for tables in tables_per_rank:
        buckets: Dict[tuple, List[ShardedEmbeddingTable]] = {}
        for t in tables:
            key = (t.is_weighted, t.data_type.value, t.pooling.value, t.has_feature_processor,
                   t.compute_kernel.value)
            buckets.setdefault(key, []).append(t)
        plain, scored = [], []
        for key in sorted(buckets):
            ts = buckets[key]
            cfg = GroupedEmbeddingConfig(data_type=ts[0].data_type, pooling=ts[0].pooling,
                                         is_weighted=ts[0].is_weighted,
                                         has_feature_processor=ts[0].has_feature_processor,
                                         compute_kernel=ts[0].compute_kernel, embedding_tables=ts)
            (scored if ts[0].is_weighted else plain).append(cfg)
        grouped.append(plain)
        score_grouped.append(scored)
    return grouped, score_grouped


class BaseSparseFeaturesDist(abc.ABC, nn.Module, Generic[F]):
    @abc.abstractmethod
    def forward(self, sparse_features: SparseFeatures) -> Awaitable[Awaitable[F]]:
        ...


class BaseEmbeddingDist(abc.ABC, nn.Module, Generic[T]):
    @abc.abstractmethod
    def forward(self, local_embs: T) -> Awaitable[torch.Tensor]:
        ...


class BaseEmbeddingLookup(abc.ABC, nn.Module, Generic[F, T]):
    @abc.abstractmethod
    def forward(self, sparse_features: F) -> T:
        ...


class EmbeddingSharding(abc.ABC, Generic[F, T]):
    """input dist / lookup / output dist factories of one sharding type."""

    @abc.abstractmethod
    def create_input_dist(self, device: Optional[torch.device] = None) -> BaseSparseFeaturesDist[F]:
        ...

    @abc.abstractmethod
    def create_output_dist(self, device: Optional[torch.device] = None) -> BaseEmbeddingDist[T]:
        ...

    @abc.abstractmethod
    def create_lookup(self, device: Optional[torch.device] = None, fused_params=None,
                      feature_processor=None) -> BaseEmbeddingLookup[F, T]:
        ...

    @abc.abstractmethod
    def embedding_dims(self) -> List[int]:
        ...

    @abc.abstractmethod
    def embedding_names(self) -> List[str]:
        ...

    @abc.abstractmethod
    def id_list_feature_names(self) -> List[str]:
        ...

    @abc.abstractmethod
    def id_score_list_feature_names(self) -> List[str]:
        ...
