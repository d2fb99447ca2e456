"""Embedding-sharding data types (reference torchrec/distributed/embedding_types.py:
EmbeddingComputeKernel :45, ShardedEmbeddingTable, GroupedEmbeddingConfig :160, SparseFeatures)."""
from __future__ import annotations

from dataclasses import dataclass
from enum import Enum, unique
from typing import List, Optional

import torch

from torchrec.modules.embedding_configs import DataType, EmbeddingTableConfig, PoolingType
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor
from torchrec.distributed.types import ShardMetadata


@unique
class EmbeddingComputeKernel(Enum):
    DENSE = "dense"
    SPARSE = "sparse"
    BATCHED_DENSE = "batched_dense"
    BATCHED_FUSED = "batched_fused"
    BATCHED_FUSED_UVM = "batched_fused_uvm"
    BATCHED_FUSED_UVM_CACHING = "batched_fused_uvm_caching"
    BATCHED_QUANT = "batched_quant"


@dataclass
class SparseFeatures:
    id_list_features: Optional[KeyedJaggedTensor] = None
    id_score_list_features: Optional[KeyedJaggedTensor] = None

    def record_stream(self, stream) -> None:
        for f in (self.id_list_features, self.id_score_list_features):
            if f is not None:
                f.record_stream(stream)


@dataclass
class SparseFeaturesList:
    features: List[SparseFeatures]

    def __len__(self) -> int:
        return len(self.features)

    def __getitem__(self, i: int) -> SparseFeatures:
        return self.features[i]

    def __iter__(self):
        return iter(self.features)

    def record_stream(self, stream) -> None:
        for f in self.features:
            f.record_stream(stream)


@dataclass
class ShardedEmbeddingTable(EmbeddingTableConfig):
    """One table (or the local shard of one) as placed on this rank."""
    compute_kernel: EmbeddingComputeKernel = EmbeddingComputeKernel.BATCHED_FUSED
    local_rows: int = 0
    local_cols: int = 0
    local_metadata: Optional[ShardMetadata] = None
    global_metadata: Optional[List[ShardMetadata]] = None
    weight_init_max: Optional[float] = None
    weight_init_min: Optional[float] = None


@dataclass
class GroupedEmbeddingConfig:
    """Tables served by ONE table-batched kernel instance."""
    data_type: DataType
    pooling: PoolingType
    is_weighted: bool
    has_feature_processor: bool
    compute_kernel: EmbeddingComputeKernel
    embedding_tables: List[ShardedEmbeddingTable]

    def feature_hash_sizes(self) -> List[int]:
        return [t.num_embeddings for t in self.embedding_tables for _ in t.feature_names]

    def num_features(self) -> int:
        return sum(t.num_features() for t in self.embedding_tables)

    def dim_sum(self) -> int:
        return sum(t.num_features() * t.local_cols for t in self.embedding_tables)

    def feature_names(self) -> List[str]:
        return [f for t in self.embedding_tables for f in t.feature_names]

    def embedding_dims(self) -> List[int]:
        return [t.local_cols for t in self.embedding_tables for _ in t.feature_names]

    def embedding_names(self) -> List[str]:
        return [n for t in self.embedding_tables for n in t.embedding_names]
