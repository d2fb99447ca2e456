"""ShardedEmbeddingBagCollection (reference torchrec/distributed/embeddingbag.py:
create_embedding_bag_sharding :94, EmbeddingBagCollectionAwaitable :200,
ShardedEmbeddingBagCollection :226, EmbeddingBagCollectionSharder :489).

input_dist (permute keys by sharding type -> split -> KJT all-to-all / bucketize)
-> compute (table-batched sm_100a lookup per sharding type)
-> output_dist (pooled all-to-all / reduce-scatter) -> KeyedTensor.
Output keys are concatenated per sharding type in plan order, as in the reference
(_create_output_dist :324-328); callers reorder by name."""
from __future__ import annotations

import copy
from collections import OrderedDict
from typing import Any, Dict, List, Optional, Tuple, Type

import torch
from torch import nn

from torchrec.distributed.embedding_sharding import EmbeddingSharding
from torchrec.distributed.embedding_types import (
    EmbeddingComputeKernel, SparseFeatures, SparseFeaturesList)
from torchrec.distributed.sharding.rw_sharding import RwPooledEmbeddingSharding
from torchrec.distributed.sharding.tw_sharding import TwPooledEmbeddingSharding
from torchrec.distributed.sharding.vb_sharding import VariableBatchShardingContext
from torchrec.distributed.sharding.vb_tw_sharding import VariableBatchTwPooledEmbeddingSharding
from torchrec.distributed.types import (
    Awaitable, LazyAwaitable, ModuleSharder, NoWait, ParameterSharding, ShardedModule, ShardedModuleContext,
    ShardingEnv, ShardingType)
from torchrec.modules.embedding_configs import EmbeddingTableConfig
from torchrec.modules.embedding_modules import EmbeddingBagCollection, EmbeddingBagCollectionInterface
from torchrec.distributed.embedding_lookup import prefixed_fused_optimizers
from torchrec.optim.fused import FusedOptimizerModule
from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor, KeyedTensor

EmbeddingInfo = Tuple[EmbeddingTableConfig, ParameterSharding, torch.Tensor]


def create_embedding_bag_sharding(sharding_type: str, embedding_configs: List[EmbeddingInfo], env: ShardingEnv,
                                  device: Optional[torch.device] = None,
                                  variable_batch_size: bool = False) -> EmbeddingSharding:
    if variable_batch_size:
        if sharding_type == ShardingType.TABLE_WISE.value:
            return VariableBatchTwPooledEmbeddingSharding(embedding_configs, env, device)
        raise ValueError(f"variable batch size per rank is built for table_wise sharding, not {sharding_type} "
                         "(the reference has vb_tw_sharding / vb_cw_sharding only)")
    if sharding_type == ShardingType.TABLE_WISE.value:
        return TwPooledEmbeddingSharding(embedding_configs, env, device)
    if sharding_type == ShardingType.ROW_WISE.value:
        return RwPooledEmbeddingSharding(embedding_configs, env, device)
    raise ValueError(
        f"Sharding type not supported {sharding_type}: this package builds table_wise and row_wise "
        "(SURVEY.md §8; data_parallel / column_wise / table_row_wise are out of scope)")


def create_embedding_configs_by_sharding(module: EmbeddingBagCollectionInterface,
                                         table_name_to_parameter_sharding: Dict[str, ParameterSharding],
                                         prefix: str) -> Dict[str, List[EmbeddingInfo]]:
    uses: Dict[str, int] = {}
    for cfg in module.embedding_bag_configs:
        if not cfg.feature_names:
            cfg.feature_names = [cfg.name]
        for f in cfg.feature_names:
            uses[f] = uses.get(f, 0) + 1
    state = module.state_dict()
    by_type: Dict[str, List[EmbeddingInfo]] = OrderedDict()
    valid_kernels = {k.value for k in EmbeddingComputeKernel}
    for cfg in module.embedding_bag_configs:
        assert cfg.name in table_name_to_parameter_sharding, f"no sharding plan for table {cfg.name}"
        ps = table_name_to_parameter_sharding[cfg.name]
        if ps.compute_kernel not in valid_kernels:
            raise ValueError(f"Compute kernel not supported {ps.compute_kernel}")
        names = [f + "@" + cfg.name if uses[f] > 1 else f for f in cfg.feature_names]
        param = state[prefix + cfg.name + ".weight"]
        by_type.setdefault(ps.sharding_type, []).append((
            EmbeddingTableConfig(num_embeddings=cfg.num_embeddings, embedding_dim=cfg.embedding_dim, name=cfg.name,
                                 data_type=cfg.data_type, feature_names=copy.deepcopy(cfg.feature_names),
                                 pooling=cfg.pooling, is_weighted=module.is_weighted, has_feature_processor=False,
                                 embedding_names=names, weight_init_max=cfg.weight_init_max,
                                 weight_init_min=cfg.weight_init_min),
            ps, param))
    return by_type


class SparseFeaturesListAwaitable(Awaitable[SparseFeaturesList]):
    """Holds the phase-1 (lengths) awaitables of every sharding: building it only ENQUEUES the lengths
    all-to-alls; wait() reads the receive sizes (the one host sync per sharding), starts every values
    all-to-all, then waits for them.  A pipeline can therefore enqueue phase 1 early and take the host
    sync when the exchange has long finished."""

    def __init__(self, awaitables: List[Awaitable[Awaitable[SparseFeatures]]]) -> None:
        super().__init__()
        self.awaitables = awaitables

    def _wait_impl(self) -> SparseFeaturesList:
        phase2 = [a.wait() for a in self.awaitables]
        return SparseFeaturesList([a.wait() for a in phase2])


class EmbeddingBagCollectionAwaitable(LazyAwaitable[KeyedTensor]):
    def __init__(self, awaitables: List[Awaitable[torch.Tensor]], embedding_dims: List[int],
                 embedding_names: List[str]) -> None:
        super().__init__()
        self._awaitables, self._embedding_dims, self._embedding_names = awaitables, embedding_dims, embedding_names

    def _wait_impl(self) -> KeyedTensor:
        parts = [a.wait() for a in self._awaitables]
        values = parts[0] if len(parts) == 1 else torch.cat(parts, dim=1)
        return KeyedTensor(keys=self._embedding_names, length_per_key=self._embedding_dims, values=values,
                           key_dim=1)


class _JoinStream(LazyAwaitable[KeyedTensor]):
    """Finishes `inner` on the stream it was issued on and makes the consumer's stream wait for it."""

    def __init__(self, inner: Awaitable[KeyedTensor], side: "torch.cuda.Stream") -> None:
        super().__init__()
        self._inner, self._side = inner, side

    def _wait_impl(self) -> KeyedTensor:
        with torch.cuda.stream(self._side):
            kt = self._inner.wait()
        cur = torch.cuda.current_stream(self._side.device)
        cur.wait_stream(self._side)
        kt.values().record_stream(cur)
        return kt


class ShardedEmbeddingBagCollection(ShardedModule[SparseFeaturesList, List[torch.Tensor], KeyedTensor],
                                    FusedOptimizerModule):
    def __init__(self, module: EmbeddingBagCollectionInterface,
                 table_name_to_parameter_sharding: Dict[str, ParameterSharding], env: ShardingEnv,
                 fused_params: Optional[Dict[str, Any]] = None, device: Optional[torch.device] = None,
                 variable_batch_size: bool = False) -> None:
        super().__init__()
        by_type = create_embedding_configs_by_sharding(module, table_name_to_parameter_sharding, "embedding_bags.")
        # variable_batch_size: every rank may feed a different batch size each step (vb_tw_sharding.py)
        self._variable_batch_size = variable_batch_size
        self._sharding_type_to_sharding: Dict[str, EmbeddingSharding] = OrderedDict(
            (st, create_embedding_bag_sharding(st, cfgs, env, device, variable_batch_size))
            for st, cfgs in by_type.items())
        self._is_weighted: bool = module.is_weighted
        self._device = device
        self._embedding_bag_configs = module.embedding_bag_configs
        self._input_dists = nn.ModuleList()
        self._output_dists = nn.ModuleList()
        self._lookups = nn.ModuleList(
            [s.create_lookup(fused_params=fused_params) for s in self._sharding_type_to_sharding.values()])
        self._embedding_names: List[str] = []
        self._embedding_dims: List[int] = []
        self._feature_splits: List[int] = []
        self._features_order: List[int] = []
        self._has_uninitialized_input_dist = True
        self._has_uninitialized_output_dist = True
        self._has_features_permute = True
        self._env = env
        self._peer = None          # fused lookup + exchange over NVLink peer memory (peer_dist.py)
        self._lookup_stream_obj = None
        self._lookup_stream_checked = False
        self._peer_checked = False
        self._optim = CombinedOptimizer(prefixed_fused_optimizers(self._lookups, "embedding_bags."))

    # ---- lazily built distribution modules
    def _feature_names_of(self, sharding: EmbeddingSharding) -> List[str]:
        return sharding.id_score_list_feature_names() if self._is_weighted else sharding.id_list_feature_names()

    def _create_input_dist(self, input_feature_names: List[str]) -> None:
        wanted: List[str] = []
        for sharding in self._sharding_type_to_sharding.values():
            self._input_dists.append(sharding.create_input_dist())
            names = self._feature_names_of(sharding)
            wanted.extend(names)
            self._feature_splits.append(len(names))
        if wanted == list(input_feature_names):
            self._has_features_permute = False
        else:
            self._features_order = [input_feature_names.index(f) for f in wanted]
            self.register_buffer("_features_order_tensor",
                                 torch.tensor(self._features_order, device=self._device, dtype=torch.int32),
                                 persistent=False)

    def _create_output_dist(self) -> None:
        for sharding in self._sharding_type_to_sharding.values():
            self._output_dists.append(sharding.create_output_dist(device=self._device))
            self._embedding_names.extend(sharding.embedding_names())
            self._embedding_dims.extend(sharding.embedding_dims())
        self._has_uninitialized_output_dist = False

    # ---- ShardedModule protocol
    def input_dist(self, ctx: ShardedModuleContext, features: KeyedJaggedTensor) -> Awaitable[SparseFeaturesList]:
        if self._has_uninitialized_input_dist:
            self._create_input_dist(features.keys())
            self._has_uninitialized_input_dist = False
        ctx.local_batch = features.stride()
        ctx.vb_contexts = []
        with torch.no_grad():
            if self._has_features_permute:
                features = features.permute(self._features_order, self._features_order_tensor)
            shards = features.split(self._feature_splits)
            awaitables = []
            for dist_module, f in zip(self._input_dists, shards):
                sf = SparseFeatures(id_list_features=None if self._is_weighted else f,
                                    id_score_list_features=f if self._is_weighted else None)
                aw = dist_module(sf)
                if self._variable_batch_size:
                    # the batch sizes were exchanged while the lengths all-to-all was being set up
                    la = aw._b if self._is_weighted else aw._a
                    ctx.vb_contexts.append(VariableBatchShardingContext.from_lengths_awaitable(la))
                awaitables.append(aw)
            return SparseFeaturesListAwaitable(awaitables)

    def compute(self, ctx: ShardedModuleContext, dist_input: SparseFeaturesList) -> List[torch.Tensor]:
        return [lookup(f) for lookup, f in zip(self._lookups, dist_input)]

    def output_dist(self, ctx: ShardedModuleContext, output: List[torch.Tensor]) -> LazyAwaitable[KeyedTensor]:
        if self._has_uninitialized_output_dist:
            self._create_output_dist()
        if self._variable_batch_size:
            return EmbeddingBagCollectionAwaitable(
                [d(e, c) for d, e, c in zip(self._output_dists, output, ctx.vb_contexts)],
                self._embedding_dims, self._embedding_names)
        return EmbeddingBagCollectionAwaitable([d(e) for d, e in zip(self._output_dists, output)],
                                               self._embedding_dims, self._embedding_names)

    def _peer_dist(self):
        if not self._peer_checked:
            from torchrec.distributed import peer_dist
            self._peer_checked = True
            if not self._variable_batch_size:
                self._peer = peer_dist.try_build(self, self._env.process_group, self._device)
        return self._peer

    def _lookup_stream(self) -> Optional["torch.cuda.Stream"]:
        """The stream the lookups and their output distribution run on.  The pooled lookup depends only on the
        distributed ids and on the previous step's table update, not on the dense arch, so it gets a stream of
        its own: forward overlaps the dense-arch forward, and - autograd runs a node's backward on the stream
        of its forward - the fused optimizer backward (HBM / NVLink bound) overlaps the dense-arch backward
        GEMMs."""
        if self._lookup_stream_checked:
            return self._lookup_stream_obj
        self._lookup_stream_checked = True
        import os
        # default: on when the module is sharded over several ranks (that is where the lookup carries the NVLink
        # exchange: measured 5.41 vs 6.46 ms/step at N = 2); on one GPU the lookup is short and the overlap only
        # makes it compete with the GEMMs.  B200_LOOKUP_STREAM=1 / 0 forces it.
        want = os.environ.get("B200_LOOKUP_STREAM", "1" if self._env.world_size > 1 else "0")
        if self._device is not None and self._device.type == "cuda" and want != "0":
            self._lookup_stream_obj = torch.cuda.Stream(device=self._device)
        return self._lookup_stream_obj

    def _compute_and_output_dist(self, ctx: ShardedModuleContext, input: SparseFeaturesList
                                 ) -> LazyAwaitable[KeyedTensor]:
        if self._variable_batch_size:
            return self.output_dist(ctx, self.compute(ctx, input))
        peer = self._peer_dist()
        if peer is not None:
            values = peer(input, ctx.local_batch)
            return NoWait(KeyedTensor(keys=peer.embedding_names, length_per_key=peer.embedding_dims,
                                      values=values, key_dim=1))
        return EmbeddingBagCollectionAwaitable(
            [d(lookup(f)) for lookup, d, f in zip(self._lookups, self._output_dists, input)],
            self._embedding_dims, self._embedding_names)

    def compute_and_output_dist(self, ctx: ShardedModuleContext, input: SparseFeaturesList
                                ) -> LazyAwaitable[KeyedTensor]:
        if self._has_uninitialized_output_dist:
            self._create_output_dist()
        side = self._lookup_stream()
        if side is None:
            return self._compute_and_output_dist(ctx, input)
        main = torch.cuda.current_stream(self._device)
        ready = getattr(ctx, "inputs_ready_event", None)
        if ready is not None:
            side.wait_event(ready)  # pipelined: the ids were distributed on another stream a step ago
        else:
            side.wait_stream(main)
        with torch.cuda.stream(side):
            inner = self._compute_and_output_dist(ctx, input)
        return _JoinStream(inner, side)

    # ---- state / optimizer
    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        if destination is None:
            destination = OrderedDict()
        for lookup in self._lookups:
            lookup.state_dict(destination, prefix + "embedding_bags.", keep_vars)
        return destination

    def load_state_dict(self, state_dict, strict: bool = True):
        """Loads `embedding_bags.<table>.weight` entries (ShardedTensor or local-shard tensors) into the live
        tables (reference load_state_dict: per-lookup copy of the matching keys)."""
        from torch.nn.modules.module import _IncompatibleKeys
        from torchrec.distributed.embedding_kernel import filter_state_dict
        sub = filter_state_dict(state_dict, "embedding_bags")
        missing, seen = [], set()
        for lookup in self._lookups:
            m, u = lookup.load_state_dict(sub, strict)
            missing.extend("embedding_bags." + k for k in m)
            seen.update(k for k in sub if k not in u)
        unexpected = ["embedding_bags." + k for k in sub if k not in seen]
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing keys {missing}, unexpected keys {unexpected}")
        return _IncompatibleKeys(missing_keys=missing, unexpected_keys=unexpected)

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        for lookup in self._lookups:
            yield from lookup.named_parameters(prefix + ("." if prefix else "") + "embedding_bags", recurse)

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        for lookup in self._lookups:
            yield from lookup.named_buffers(prefix + ("." if prefix else "") + "embedding_bags", recurse)

    def sparse_grad_parameter_names(self, destination: Optional[List[str]] = None, prefix: str = "") -> List[str]:
        return [] if destination is None else destination

    @property
    def fused_optimizer(self) -> KeyedOptimizer:
        return self._optim

    @property
    def embedding_bag_configs(self):
        return self._embedding_bag_configs

    @property
    def is_weighted(self) -> bool:
        return self._is_weighted


class EmbeddingBagCollectionSharder(ModuleSharder[EmbeddingBagCollection]):
    def __init__(self, fused_params: Optional[Dict[str, Any]] = None, variable_batch_size: bool = False) -> None:
        self._fused_params = fused_params
        self._variable_batch_size = variable_batch_size

    def shard(self, module: EmbeddingBagCollection, params: Dict[str, ParameterSharding], env: ShardingEnv,
              device: Optional[torch.device] = None) -> ShardedEmbeddingBagCollection:
        return ShardedEmbeddingBagCollection(module, params, env, self._fused_params, device,
                                             variable_batch_size=self._variable_batch_size)

    def shardable_parameters(self, module: EmbeddingBagCollection) -> Dict[str, nn.Parameter]:
        return {name.split(".")[0]: p for name, p in module.embedding_bags.named_parameters()}

    @property
    def module_type(self) -> Type[EmbeddingBagCollection]:
        return EmbeddingBagCollection

    def sharding_types(self, compute_device_type: str) -> List[str]:
        if self._variable_batch_size:
            return [ShardingType.TABLE_WISE.value]
        return [ShardingType.TABLE_WISE.value, ShardingType.ROW_WISE.value]

    def compute_kernels(self, sharding_type: str, compute_device_type: str) -> List[str]:
        return [EmbeddingComputeKernel.BATCHED_FUSED.value]

    @property
    def fused_params(self) -> Optional[Dict[str, Any]]:
        return self._fused_params
