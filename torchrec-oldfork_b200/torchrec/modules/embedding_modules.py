"""Unsharded EmbeddingBagCollection (reference torchrec/modules/embedding_modules.py:67-231).

This is the semantic reference of the pooled output layout — one nn.EmbeddingBag per table,
features in table order, [B, sum_f D_f] — and the module DistributedModelParallel replaces by
ShardedEmbeddingBagCollection, which is where the sm_100a kernels run.
"""
import abc
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from torchrec.modules.embedding_configs import EmbeddingBagConfig, PoolingType, pooling_type_to_str
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor, KeyedTensor


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
