"""Grouped pooled lookups (reference torchrec/distributed/embedding_lookup.py:
GroupedPooledEmbeddingsLookup :209, forward :287-337): the local KJT is split per kernel group,
each group runs ONE table-batched lookup, outputs are concatenated along the embedding dim."""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn

from torchrec.distributed.batched_embedding_kernel import BatchedFusedEmbedding, BatchedFusedEmbeddingBag
from torchrec.distributed.embedding_sharding import BaseEmbeddingLookup
from torchrec.distributed.embedding_types import (
    EmbeddingComputeKernel, GroupedEmbeddingConfig, SparseFeatures)
from torchrec.optim.fused import FusedOptimizerModule
from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer


def prefixed_fused_optimizers(lookups, prefix: str) -> List[Tuple[str, KeyedOptimizer]]:
    """The fused optimizers of every table-batched kernel under `lookups`, their parameter keys re-rooted at
    the unsharded module's attribute (`embedding_bags.` / `embedding_modules.`) so optimizer checkpoints carry
    the same keys as the reference's (embeddingbag.py:272-283, embedding.py:242-253)."""
    optims: List[Tuple[str, KeyedOptimizer]] = []
    for lookup in lookups:
        for m in lookup.modules():
            if isinstance(m, (BatchedFusedEmbeddingBag, BatchedFusedEmbedding)):
                opt = m.fused_optimizer
                if not getattr(opt, "_b200_prefixed", False):
                    opt.params = {prefix + k: w for k, w in opt.params.items()}
                    opt._b200_prefixed = True
                optims.append(("", opt))
    return optims


class GroupedPooledEmbeddingsLookup(BaseEmbeddingLookup[SparseFeatures, torch.Tensor], FusedOptimizerModule):
    def __init__(self, grouped_configs: List[GroupedEmbeddingConfig],
                 grouped_score_configs: List[GroupedEmbeddingConfig],
                 device: Optional[torch.device] = None, pg: Optional[dist.ProcessGroup] = None,
                 fused_params: Optional[Dict[str, Any]] = None, feature_processor=None) -> None:
        super().__init__()
        if feature_processor is not None:
            raise NotImplementedError("feature processors are out of scope")

        def make(cfg: GroupedEmbeddingConfig) -> nn.Module:
            if cfg.compute_kernel != EmbeddingComputeKernel.BATCHED_FUSED:
                raise NotImplementedError(
                    f"compute kernel {cfg.compute_kernel} is not built; only BATCHED_FUSED (SURVEY.md §8 a11)")
            return BatchedFusedEmbeddingBag(cfg, pg=pg, device=device, fused_params=fused_params)

        self._emb_modules = nn.ModuleList([make(c) for c in grouped_configs])
        self._score_emb_modules = nn.ModuleList([make(c) for c in grouped_score_configs])
        self._id_list_feature_splits = [c.num_features() for c in grouped_configs]
        self._id_score_list_feature_splits = [c.num_features() for c in grouped_score_configs]
        self.grouped_configs = grouped_configs
        self.grouped_score_configs = grouped_score_configs
        self._device = device
        self._dim_sum = sum(c.dim_sum() for c in grouped_configs) + sum(c.dim_sum() for c in grouped_score_configs)

    def forward(self, sparse_features: SparseFeatures) -> torch.Tensor:
        outs: List[torch.Tensor] = []
        if self._emb_modules:
            assert sparse_features.id_list_features is not None
            parts = sparse_features.id_list_features.split(self._id_list_feature_splits)
            outs.extend(m(f) for m, f in zip(self._emb_modules, parts))
        if self._score_emb_modules:
            assert sparse_features.id_score_list_features is not None
            parts = sparse_features.id_score_list_features.split(self._id_score_list_feature_splits)
            outs.extend(m(f) for m, f in zip(self._score_emb_modules, parts))
        if len(outs) == 0:
            kjt = sparse_features.id_list_features or sparse_features.id_score_list_features
            B = kjt.stride() if kjt is not None else 0
            # a rank without tables of this sharding type still takes part in the output collective and
            # in its backward: the empty block must be part of the autograd graph
            return torch.empty((B, 0), device=self._device, requires_grad=torch.is_grad_enabled())
        return outs[0] if len(outs) == 1 else torch.cat(outs, dim=1)

    def _all_modules(self) -> List[nn.Module]:
        return list(self._emb_modules) + list(self._score_emb_modules)

    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        if destination is None:
            destination = {}
        for m in self._all_modules():
            m.state_dict(destination, prefix, keep_vars)
        return destination

    def load_state_dict(self, state_dict, strict: bool = True):
        """Copies into the live tables; entries may be ShardedTensors or tensors of the local shard shape
        (reference _load_state_dict :50-74).  Returns (missing_keys, unexpected_keys)."""
        from torchrec.distributed.embedding_kernel import load_state_dict_into
        return load_state_dict_into(self._all_modules(), state_dict)

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        yield from ()

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        for m in self._all_modules():
            yield from m.named_buffers(prefix, recurse)

    @property
    def fused_optimizer(self) -> KeyedOptimizer:
        return CombinedOptimizer([m.fused_optimizer for m in self._all_modules()
                                  if isinstance(m, FusedOptimizerModule)])


class GroupedEmbeddingsLookup(BaseEmbeddingLookup[SparseFeatures, torch.Tensor], FusedOptimizerModule):
    """Sequence (non-pooled) lookups (reference GroupedEmbeddingsLookup :59-207): one table-batched
    PoolingMode.NONE module per kernel group; rows are concatenated in feature order, [sum lengths, D]."""

    def __init__(self, grouped_configs: List[GroupedEmbeddingConfig], device: Optional[torch.device] = None,
                 pg: Optional[dist.ProcessGroup] = None, fused_params: Optional[Dict[str, Any]] = None) -> None:
        super().__init__()
        for cfg in grouped_configs:
            if cfg.compute_kernel != EmbeddingComputeKernel.BATCHED_FUSED:
                raise NotImplementedError(f"compute kernel {cfg.compute_kernel} is not built; only BATCHED_FUSED")
        self._emb_modules = nn.ModuleList(
            [BatchedFusedEmbedding(c, pg=pg, device=device, fused_params=fused_params) for c in grouped_configs])
        self._id_list_feature_splits = [c.num_features() for c in grouped_configs]
        self.grouped_configs = grouped_configs
        self._device = device
        self._dim = grouped_configs[0].embedding_tables[0].local_cols if grouped_configs else 0

    def forward(self, sparse_features: SparseFeatures) -> torch.Tensor:
        kjt = sparse_features.id_list_features
        if not self._emb_modules:
            # a rank without tables still joins the output all-to-all and its backward
            return torch.empty((0, self._dim), device=self._device, requires_grad=torch.is_grad_enabled())
        assert kjt is not None
        parts = kjt.split(self._id_list_feature_splits) if len(self._emb_modules) > 1 else [kjt]
        outs = [m(f) for m, f in zip(self._emb_modules, parts)]
        return outs[0] if len(outs) == 1 else torch.cat(outs, dim=0)

    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        if destination is None:
            destination = {}
        for m in self._emb_modules:
            m.state_dict(destination, prefix, keep_vars)
        return destination

    def load_state_dict(self, state_dict, strict: bool = True):
        from torchrec.distributed.embedding_kernel import load_state_dict_into
        return load_state_dict_into(list(self._emb_modules), state_dict)

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        yield from ()

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        for m in self._emb_modules:
            yield from m.named_buffers(prefix, recurse)

    @property
    def fused_optimizer(self) -> KeyedOptimizer:
        return CombinedOptimizer([m.fused_optimizer for m in self._emb_modules])
