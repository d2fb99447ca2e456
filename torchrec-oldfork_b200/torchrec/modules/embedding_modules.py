"""Unsharded EmbeddingBagCollection (reference torchrec/modules/embedding_modules.py:67-231).

This is the semantic reference of the pooled output layout — one nn.EmbeddingBag per table,
features in table order, [B, sum_f D_f] — and the module DistributedModelParallel replaces by
ShardedEmbeddingBagCollection, which is where the sm_100a kernels run.
"""
import abc
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from torchrec.modules.embedding_configs import (
    DataType, EmbeddingBagConfig, EmbeddingConfig, PoolingType, pooling_type_to_str)
from torchrec.sparse.jagged_tensor import JaggedTensor, KeyedJaggedTensor, KeyedTensor


def ebc_get_embedding_names(tables: List[EmbeddingBagConfig]) -> List[str]:
    """feature name, suffixed by @table when a feature is served by several tables (:47-64)."""
    counts: Dict[str, int] = {}
    for t in tables:
        for f in t.feature_names:
            counts[f] = counts.get(f, 0) + 1
    return [f + "@" + t.name if counts[f] > 1 else f for t in tables for f in t.feature_names]


class EmbeddingBagCollectionInterface(abc.ABC, nn.Module):
    @abc.abstractmethod
    def forward(self, features: KeyedJaggedTensor) -> KeyedTensor:
        ...

    @property
    @abc.abstractmethod
    def embedding_bag_configs(self) -> List[EmbeddingBagConfig]:
        ...

    @property
    @abc.abstractmethod
    def is_weighted(self) -> bool:
        ...


class EmbeddingBagCollection(EmbeddingBagCollectionInterface):
    def __init__(self, tables: List[EmbeddingBagConfig], is_weighted: bool = False,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._is_weighted = is_weighted
        self.embedding_bags: nn.ModuleDict = nn.ModuleDict()
        self._embedding_bag_configs = tables
        self._lengths_per_embedding: List[int] = []
        for cfg in tables:
            if cfg.name in self.embedding_bags:
                raise ValueError(f"Duplicate table name {cfg.name}")
            self.embedding_bags[cfg.name] = nn.EmbeddingBag(
                num_embeddings=cfg.num_embeddings, embedding_dim=cfg.embedding_dim,
                mode=pooling_type_to_str(cfg.pooling), device=device, include_last_offset=True)
            if not cfg.feature_names:
                cfg.feature_names = [cfg.name]
            self._lengths_per_embedding.extend(len(cfg.feature_names) * [cfg.embedding_dim])
        self._embedding_names: List[str] = ebc_get_embedding_names(tables)

    def forward(self, features: KeyedJaggedTensor) -> KeyedTensor:
        per_feature = features.to_dict()
        pooled: List[torch.Tensor] = []
        for cfg in self._embedding_bag_configs:
            bag = self.embedding_bags[cfg.name]
            for name in cfg.feature_names:
                jt = per_feature[name]
                pooled.append(bag(input=jt.values(), offsets=jt.offsets(),
                                  per_sample_weights=jt.weights() if self._is_weighted else None).float())
        return KeyedTensor(keys=self._embedding_names, values=torch.cat(pooled, dim=1),
                           length_per_key=self._lengths_per_embedding)

    @property
    def embedding_bag_configs(self) -> List[EmbeddingBagConfig]:
        return self._embedding_bag_configs

    @property
    def is_weighted(self) -> bool:
        return self._is_weighted


class EmbeddingCollection(nn.Module):
    """Collection of NON-pooled (sequence) embeddings (reference :204-348): KJT [F x B x L] in,
    Dict[feature, JaggedTensor] out, each JaggedTensor holding one [D] row per id.  One nn.Embedding per
    table; all tables share one embedding dim.  This is the semantic reference that
    DistributedModelParallel replaces by ShardedEmbeddingCollection (sm_100a sequence lookup)."""

    def __init__(self, tables: List[EmbeddingConfig], device: Optional[torch.device] = None) -> None:
        super().__init__()
        self.embeddings: nn.ModuleDict = nn.ModuleDict()
        self.embedding_configs = tables
        self.embedding_dim: int = -1
        uses: Dict[str, int] = {}
        for cfg in tables:
            if cfg.name in self.embeddings:
                raise ValueError(f"Duplicate table name {cfg.name}")
            if self.embedding_dim < 0:
                self.embedding_dim = cfg.embedding_dim
            if self.embedding_dim != cfg.embedding_dim:
                raise ValueError("All tables in a EmbeddingCollection are required to have same embedding dimension.")
            dtype = torch.float32 if cfg.data_type == DataType.FP32 else torch.float16
            self.embeddings[cfg.name] = nn.Embedding(num_embeddings=cfg.num_embeddings,
                                                     embedding_dim=cfg.embedding_dim, device=device, dtype=dtype)
            if not cfg.feature_names:
                cfg.feature_names = [cfg.name]
            for f in cfg.feature_names:
                uses[f] = uses.get(f, 0) + 1
        self.embedding_names_by_table: List[List[str]] = [
            [f + "@" + cfg.name if uses[f] > 1 else f for f in cfg.feature_names] for cfg in tables]

    def forward(self, features: KeyedJaggedTensor) -> Dict[str, JaggedTensor]:
        out: Dict[str, JaggedTensor] = {}
        for cfg, names, emb in zip(self.embedding_configs, self.embedding_names_by_table, self.embeddings.values()):
            for feature, name in zip(cfg.feature_names, names):
                f = features[feature]
                out[name] = JaggedTensor(values=emb(f.values()), lengths=f.lengths())
        return out
