"""ShardedEmbeddingCollection: sharded NON-pooled (sequence) embeddings (reference
torchrec/distributed/embedding.py: create_embedding_sharding :52, _construct_jagged_tensors :140,
EmbeddingCollectionAwaitable :170, ShardedEmbeddingCollection :196, EmbeddingCollectionSharder :497).

input_dist (permute keys by sharding type -> KJT all-to-all, row-wise adds block_bucketize with the
unbucketize permutation) -> compute (table-batched PoolingMode.NONE lookup on the sm_100a kernels,
one [D] row per id) -> output_dist (SequenceEmbeddingAllToAll sends every row back to the rank that
owns its sample) -> Dict[feature, JaggedTensor] with the lengths of the ORIGINAL local features."""
from __future__ import annotations

import copy
from collections import OrderedDict
from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Tuple, Type

import torch
from torch import nn

from torchrec.distributed.embedding_sharding import EmbeddingSharding
from torchrec.distributed.embedding_types import EmbeddingComputeKernel, SparseFeatures, SparseFeaturesList
from torchrec.distributed.sharding.rw_sharding import RwSparseFeaturesDist
from torchrec.distributed.sharding.sequence_sharding import (
    RwSequenceEmbeddingSharding, SequenceShardingContext, TwSequenceEmbeddingSharding)
from torchrec.distributed.types import (
    Awaitable, LazyAwaitable, ModuleSharder, ParameterSharding, ShardedModule, ShardedModuleContext, ShardingEnv,
    ShardingType)
from torchrec.modules.embedding_configs import EmbeddingTableConfig, PoolingType
from torchrec.modules.embedding_modules import EmbeddingCollection
from torchrec.distributed.embedding_lookup import prefixed_fused_optimizers
from torchrec.optim.fused import FusedOptimizerModule
from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer
from torchrec.sparse.jagged_tensor import JaggedTensor, KeyedJaggedTensor

EmbeddingInfo = Tuple[EmbeddingTableConfig, ParameterSharding, torch.Tensor]


def create_embedding_sharding(sharding_type: str, embedding_configs: List[EmbeddingInfo], env: ShardingEnv,
                              device: Optional[torch.device] = None) -> EmbeddingSharding:
    if sharding_type == ShardingType.TABLE_WISE.value:
        return TwSequenceEmbeddingSharding(embedding_configs, env, device)
    if sharding_type == ShardingType.ROW_WISE.value:
        return RwSequenceEmbeddingSharding(embedding_configs, env, device)
    raise ValueError(f"Sharding not supported {sharding_type}: table_wise and row_wise are built")


def _create_embedding_configs_by_sharding(module: EmbeddingCollection,
                                          table_name_to_parameter_sharding: Dict[str, ParameterSharding]
                                          ) -> Dict[str, List[EmbeddingInfo]]:
    by_type: Dict[str, List[EmbeddingInfo]] = OrderedDict()
    state = module.state_dict()
    valid = {k.value for k in EmbeddingComputeKernel}
    for cfg, names in zip(module.embedding_configs, module.embedding_names_by_table):
        assert cfg.name in table_name_to_parameter_sharding, f"no sharding plan for table {cfg.name}"
        ps = table_name_to_parameter_sharding[cfg.name]
        if ps.compute_kernel not in valid:
            raise ValueError(f"Compute kernel not supported {ps.compute_kernel}")
        by_type.setdefault(ps.sharding_type, []).append((
            EmbeddingTableConfig(num_embeddings=cfg.num_embeddings, embedding_dim=cfg.embedding_dim, name=cfg.name,
                                 data_type=cfg.data_type, feature_names=copy.deepcopy(cfg.feature_names),
                                 pooling=PoolingType.NONE, is_weighted=False, has_feature_processor=False,
                                 embedding_names=list(names), weight_init_max=cfg.weight_init_max,
                                 weight_init_min=cfg.weight_init_min),
            ps, state["embeddings." + cfg.name + ".weight"]))
    return by_type


def _construct_jagged_tensors(embeddings: torch.Tensor, features: KeyedJaggedTensor,
                              embedding_names: List[str]) -> Dict[str, JaggedTensor]:
    """Rows arrive in the key order of `features` (the local KJT before the input dist)."""
    ret: Dict[str, JaggedTensor] = {}
    opk = features.offset_per_key()
    lengths = features.lengths().view(-1, features.stride())
    for i, name in enumerate(embedding_names):
        ret[name] = JaggedTensor(values=embeddings[opk[i]:opk[i + 1], :], lengths=lengths[i])
    return ret


@dataclass
class EmbeddingCollectionContext(ShardedModuleContext):
    sharding_contexts: List[SequenceShardingContext] = field(default_factory=list)

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        for c in self.sharding_contexts:
            c.record_stream(stream)


class _SeqFeaturesAwaitable(Awaitable[SparseFeaturesList]):
    def __init__(self, awaitables: List[Awaitable[SparseFeatures]]) -> None:
        super().__init__()
        self._awaitables = awaitables

    def _wait_impl(self) -> SparseFeaturesList:
        return SparseFeaturesList([a.wait() for a in self._awaitables])


class EmbeddingCollectionAwaitable(LazyAwaitable[Dict[str, JaggedTensor]]):
    def __init__(self, awaitables: List[Awaitable[torch.Tensor]], features: List[KeyedJaggedTensor],
                 embedding_names: List[List[str]]) -> None:
        super().__init__()
        self._awaitables, self._features, self._embedding_names = awaitables, features, embedding_names

    def _wait_impl(self) -> Dict[str, JaggedTensor]:
        out: Dict[str, JaggedTensor] = {}
        for w, f, names in zip(self._awaitables, self._features, self._embedding_names):
            out.update(_construct_jagged_tensors(w.wait(), f, names))
        return out


class ShardedEmbeddingCollection(ShardedModule[SparseFeaturesList, List[torch.Tensor], Dict[str, JaggedTensor]],
                                 FusedOptimizerModule):
    def __init__(self, module: EmbeddingCollection, table_name_to_parameter_sharding: Dict[str, ParameterSharding],
                 env: ShardingEnv, fused_params: Optional[Dict[str, Any]] = None,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        by_type = _create_embedding_configs_by_sharding(module, table_name_to_parameter_sharding)
        self._sharding_type_to_sharding: Dict[str, EmbeddingSharding] = OrderedDict(
            (st, create_embedding_sharding(st, cfgs, env, device)) for st, cfgs in by_type.items())
        self._device = device
        self._embedding_dim: int = module.embedding_dim
        self._embedding_configs = module.embedding_configs
        self._input_dists = nn.ModuleList()
        self._lookups = nn.ModuleList(
            [s.create_lookup(fused_params=fused_params) for s in self._sharding_type_to_sharding.values()])
        self._output_dists = nn.ModuleList(
            [s.create_output_dist(device=device) for s in self._sharding_type_to_sharding.values()])
        self._embedding_names_per_sharding: List[List[str]] = [
            s.embedding_names() for s in self._sharding_type_to_sharding.values()]
        self._feature_splits: List[int] = []
        self._features_order: List[int] = []
        self._has_uninitialized_input_dist = True
        self._optim = CombinedOptimizer(prefixed_fused_optimizers(self._lookups, "embedding_modules."))

    def create_context(self) -> EmbeddingCollectionContext:
        return EmbeddingCollectionContext()

    def _create_input_dist(self, input_feature_names: List[str]) -> None:
        wanted: List[str] = []
        for sharding in self._sharding_type_to_sharding.values():
            self._input_dists.append(sharding.create_input_dist())
            names = sharding.id_list_feature_names()
            wanted.extend(names)
            self._feature_splits.append(len(names))
        order = [input_feature_names.index(f) for f in wanted]
        self._features_order = [] if order == list(range(len(input_feature_names))) else order
        self.register_buffer("_features_order_tensor",
                             torch.tensor(self._features_order, device=self._device, dtype=torch.int32),
                             persistent=False)

    def input_dist(self, ctx: EmbeddingCollectionContext, features: KeyedJaggedTensor
                   ) -> Awaitable[SparseFeaturesList]:
        if self._has_uninitialized_input_dist:
            self._create_input_dist(list(features.keys()))
            self._has_uninitialized_input_dist = False
        with torch.no_grad():
            if self._features_order:
                features = features.permute(self._features_order, self._features_order_tensor)
            awaitables = []
            for dist_module, f in zip(self._input_dists, features.split(self._feature_splits)):
                indices_aw = dist_module(SparseFeatures(id_list_features=f, id_score_list_features=None)).wait()
                kjt_aw = getattr(indices_aw, "_a", None)  # KJTAllToAllIndicesAwaitable of the id_list half
                ctx.sharding_contexts.append(SequenceShardingContext(
                    features_before_input_dist=f,
                    input_splits=list(getattr(kjt_aw, "_in_lengths_per_worker", [])),
                    output_splits=list(getattr(kjt_aw, "_out_lengths_per_worker", [])),
                    unbucketize_permute_tensor=(dist_module.unbucketize_permute_tensor
                                                if isinstance(dist_module, RwSparseFeaturesDist) else None)))
                awaitables.append(indices_aw)
            return _SeqFeaturesAwaitable(awaitables)

    def compute(self, ctx: EmbeddingCollectionContext, dist_input: SparseFeaturesList) -> List[torch.Tensor]:
        out: List[torch.Tensor] = []
        for lookup, sf, sctx in zip(self._lookups, dist_input, ctx.sharding_contexts):
            kjt = sf.id_list_features
            sctx.lengths_after_input_dist = kjt.lengths().view(-1, kjt.stride())
            out.append(lookup(sf).view(-1, self._embedding_dim))
        return out

    def output_dist(self, ctx: EmbeddingCollectionContext, output: List[torch.Tensor]
                    ) -> LazyAwaitable[Dict[str, JaggedTensor]]:
        return EmbeddingCollectionAwaitable(
            [d(e, sctx) for d, e, sctx in zip(self._output_dists, output, ctx.sharding_contexts)],
            [sctx.features_before_input_dist for sctx in ctx.sharding_contexts],
            self._embedding_names_per_sharding)

    # ---- state / optimizer
    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        if destination is None:
            destination = OrderedDict()
        for lookup in self._lookups:
            lookup.state_dict(destination, prefix + "embeddings.", keep_vars)
        return destination

    def load_state_dict(self, state_dict, strict: bool = True):
        """Loads `embeddings.<table>.weight` entries (ShardedTensor or local-shard tensors) into the live
        tables (reference load_state_dict: per-lookup copy of the matching keys)."""
        from torch.nn.modules.module import _IncompatibleKeys
        from torchrec.distributed.embedding_kernel import filter_state_dict
        sub = filter_state_dict(state_dict, "embeddings")
        missing, seen = [], set()
        for lookup in self._lookups:
            m, u = lookup.load_state_dict(sub, strict)
            missing.extend("embeddings." + k for k in m)
            seen.update(k for k in sub if k not in u)
        unexpected = ["embeddings." + k for k in sub if k not in seen]
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing keys {missing}, unexpected keys {unexpected}")
        return _IncompatibleKeys(missing_keys=missing, unexpected_keys=unexpected)

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        for lookup in self._lookups:
            yield from lookup.named_parameters(prefix + ("." if prefix else "") + "embeddings", recurse)

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        for lookup in self._lookups:
            yield from lookup.named_buffers(prefix + ("." if prefix else "") + "embeddings", recurse)

    @property
    def fused_optimizer(self) -> KeyedOptimizer:
        return self._optim


class EmbeddingCollectionSharder(ModuleSharder[EmbeddingCollection]):
    def __init__(self, fused_params: Optional[Dict[str, Any]] = None) -> None:
        self._fused_params = fused_params

    def shard(self, module: EmbeddingCollection, params: Dict[str, ParameterSharding], env: ShardingEnv,
              device: Optional[torch.device] = None) -> ShardedEmbeddingCollection:
        return ShardedEmbeddingCollection(module, params, env, self._fused_params, device)

    def shardable_parameters(self, module: EmbeddingCollection) -> Dict[str, nn.Parameter]:
        return {name.split(".")[0]: p for name, p in module.embeddings.named_parameters()}

    @property
    def module_type(self) -> Type[EmbeddingCollection]:
        return EmbeddingCollection

    def sharding_types(self, compute_device_type: str) -> List[str]:
        return [ShardingType.TABLE_WISE.value, ShardingType.ROW_WISE.value]

    def compute_kernels(self, sharding_type: str, compute_device_type: str) -> List[str]:
        return [EmbeddingComputeKernel.BATCHED_FUSED.value]

    @property
    def fused_params(self) -> Optional[Dict[str, Any]]:
        return self._fused_params
