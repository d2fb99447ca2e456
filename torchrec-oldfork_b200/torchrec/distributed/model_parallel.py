"""DistributedModelParallel (reference torchrec/distributed/model_parallel.py:
DefaultDataParallelWrapper :65, DistributedModelParallel :127, _shard_modules_impl :310):
replaces every module that has a sharding-plan entry by its sharded counterpart, materialises the
remaining (dense) parameters, wraps them in DistributedDataParallel and collects the fused
optimizers of the sharded modules."""
from __future__ import annotations

from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn
from torch.nn.parallel import DistributedDataParallel

from torchrec.distributed.embeddingbag import EmbeddingBagCollectionSharder
from torchrec.distributed.planner import EmbeddingShardingPlanner, Topology
from torchrec.distributed.types import ModuleSharder, ShardedModule, ShardingEnv, ShardingPlan
from torchrec.optim.fused import FusedOptimizerModule
from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer


def get_default_sharders() -> List[ModuleSharder]:
    return [EmbeddingBagCollectionSharder()]


class DataParallelWrapper:
    def wrap(self, dmp: "DistributedModelParallel", env: ShardingEnv, device: torch.device) -> None:
        raise NotImplementedError


class DefaultDataParallelWrapper(DataParallelWrapper):
    def __init__(self, bucket_cap_mb: int = 25) -> None:
        self._bucket_cap_mb = bucket_cap_mb

    def wrap(self, dmp: "DistributedModelParallel", env: ShardingEnv, device: torch.device) -> None:
        if isinstance(dmp._dmp_wrapped_module, DistributedDataParallel):
            return
        pg = env.process_group
        if pg is None or env.world_size <= 1:
            return
        # Everything under a ShardedModule is model-parallel: the table shards are updated by the fused
        # backward, and the number of lookup modules (hence of their autograd placeholders) differs from
        # rank to rank.  DDP must see the same dense parameter list everywhere (reference :96-104).
        ignore: List[str] = []
        for mod_name, mod in dmp._dmp_wrapped_module.named_modules():
            if isinstance(mod, ShardedModule):
                for sub_name, sub in nn.Module.named_modules(mod):
                    full = ".".join(x for x in (mod_name, sub_name) if x)
                    for n, _ in list(sub._parameters.items()) + list(sub._buffers.items()):
                        ignore.append(f"{full}.{n}" if full else n)
        ignored = set(ignore)
        if not any(p.requires_grad for n, p in dmp._dmp_wrapped_module.named_parameters() if n not in ignored):
            return  # sparse-only model (e.g. a bare EmbeddingCollection): nothing is data-parallel
        DistributedDataParallel._set_params_and_buffers_to_ignore_for_model(dmp._dmp_wrapped_module, ignore)
        def build() -> DistributedDataParallel:
            return DistributedDataParallel(
                module=dmp._dmp_wrapped_module.to(device),
                device_ids=None if device.type == "cpu" else [device], process_group=pg,
                gradient_as_bucket_view=True, broadcast_buffers=False, bucket_cap_mb=self._bucket_cap_mb)

        if device.type == "cuda":
            # DDP keeps every parameter's AccumulateGrad node alive, and a node runs on the stream that was
            # current when it was created.  Built on the default (legacy) stream those nodes could never take part
            # in a CUDA-graph capture of the step (TrainPipelineSparseDist(cuda_graph=True)), so DDP is built on a
            # side stream, as PyTorch's CUDA-graph notes prescribe.
            cur = torch.cuda.current_stream(device)
            side = torch.cuda.Stream(device=device)
            side.wait_stream(cur)
            with torch.cuda.stream(side):
                dmp._dmp_wrapped_module = build()
            cur.wait_stream(side)
            dmp._ddp_stream = side
            try:  # the stream mismatch between those nodes and their producers is intended
                torch.autograd.graph.set_warn_on_accumulate_grad_stream_mismatch(False)
            except AttributeError:
                pass
        else:
            dmp._dmp_wrapped_module = build()


class DistributedModelParallel(nn.Module, FusedOptimizerModule):
    def __init__(self, module: nn.Module, env: Optional[ShardingEnv] = None,
                 device: Optional[torch.device] = None, plan: Optional[ShardingPlan] = None,
                 sharders: Optional[List[ModuleSharder]] = None, init_data_parallel: bool = True,
                 init_parameters: bool = True, data_parallel_wrapper: Optional[DataParallelWrapper] = None) -> None:
        super().__init__()
        torch._C._log_api_usage_once(f"torchrec.distributed.{self.__class__.__name__}")
        self.init_parameters = init_parameters
        self._ddp_wrapped = False
        if env is None:
            pg = dist.distributed_c10d._get_default_group() if dist.is_initialized() else None
            env = ShardingEnv.from_process_group(pg) if pg is not None else ShardingEnv.from_local(1, 0)
        self._env = env
        self.device = device if device is not None else torch.device("cpu")
        sharders = sharders if sharders is not None else get_default_sharders()
        self._sharder_map: Dict[type, ModuleSharder] = {s.module_type: s for s in sharders}
        if data_parallel_wrapper is None:
            data_parallel_wrapper = DefaultDataParallelWrapper()
        self._data_parallel_wrapper = data_parallel_wrapper
        if plan is None:
            plan = EmbeddingShardingPlanner(Topology(env.world_size, self.device.type)).collective_plan(
                module, sharders, env.process_group)
        self._plan = plan
        self._dmp_wrapped_module: nn.Module = self._shard_modules(module)
        if init_parameters:
            self._init_parameters(self._dmp_wrapped_module)
        if init_data_parallel:
            self.init_data_parallel()
        fused = [(name, m.fused_optimizer) for name, m in self._named_sharded_modules(self.module)
                 if isinstance(m, FusedOptimizerModule)]
        self._optim = CombinedOptimizer(fused)

    @property
    def module(self) -> nn.Module:
        m = self._dmp_wrapped_module
        return m.module if isinstance(m, DistributedDataParallel) else m

    @module.setter
    def module(self, value: nn.Module) -> None:
        if isinstance(self._dmp_wrapped_module, DistributedDataParallel):
            raise RuntimeError("module can't be set after calling init_data_parallel(...)")
        self._dmp_wrapped_module = value

    def forward(self, *args, **kwargs) -> Any:
        return self._dmp_wrapped_module(*args, **kwargs)

    def init_data_parallel(self) -> None:
        if not self._ddp_wrapped:
            self._data_parallel_wrapper.wrap(self, self._env, self.device)
            self._ddp_wrapped = True

    def copy(self, device: torch.device) -> "DistributedModelParallel":
        raise NotImplementedError("copy() (inference multi-device) is out of scope")

    # ---- sharding
    def _shard_modules(self, module: nn.Module, path: str = "") -> nn.Module:
        if isinstance(module, ShardedModule):
            return module
        entry = self._plan.get_plan_for_module(path)
        if entry is not None:
            sharder = self._sharder_map.get(type(module))
            if sharder is None:
                raise RuntimeError(f"no sharder registered for {type(module).__name__} at '{path}'")
            return sharder.shard(module, entry, self._env, self.device)
        for name, child in list(module.named_children()):
            new = self._shard_modules(child, f"{path}.{name}" if path else name)
            if new is not child:
                setattr(module, name, new)
        return module

    def _init_parameters(self, module: nn.Module) -> None:
        @torch.no_grad()
        def init(m: nn.Module) -> None:
            if isinstance(m, ShardedModule):
                return
            meta = any(t.device.type == "meta" for t in list(m.parameters(recurse=False)) + list(m.buffers(recurse=False)))
            if meta:
                m.to_empty(device=self.device, recurse=False)
                if hasattr(m, "reset_parameters"):
                    m.reset_parameters()
            else:
                for key, p in list(m._parameters.items()):
                    if p is not None and p.device != self.device:
                        m._parameters[key] = nn.Parameter(p.to(self.device), requires_grad=p.requires_grad)
                for key, b in list(m._buffers.items()):
                    if b is not None and b.device != self.device:
                        m._buffers[key] = b.to(self.device)
        for m in module.modules():
            init(m)

    @staticmethod
    def _named_sharded_modules(module: nn.Module, prefix: str = "") -> Iterator[Tuple[str, ShardedModule]]:
        if isinstance(module, ShardedModule):
            yield prefix, module
            return
        for name, child in module.named_children():
            yield from DistributedModelParallel._named_sharded_modules(child, f"{prefix}.{name}" if prefix else name)

    # ---- delegation
    def sparse_grad_parameter_names(self, destination: Optional[List[str]] = None, prefix: str = "") -> List[str]:
        return [] if destination is None else destination

    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        return self.module.state_dict(destination=destination, prefix=prefix, keep_vars=keep_vars) \
            if destination is not None else self.module.state_dict(prefix=prefix, keep_vars=keep_vars)

    def load_state_dict(self, state_dict, prefix: str = "", strict: bool = True):
        """Sharded modules load their tables from ShardedTensors / local-shard tensors, everything else
        goes through nn.Module._load_from_state_dict (reference :421-455)."""
        from torch.nn.modules.module import _IncompatibleKeys
        missing, unexpected = self._load_state_dict(self.module, state_dict, prefix, strict)
        return _IncompatibleKeys(missing_keys=missing, unexpected_keys=unexpected)

    def _load_state_dict(self, module: nn.Module, state_dict, prefix: str, strict: bool):
        from torchrec.distributed.embedding_kernel import filter_state_dict
        if isinstance(module, ShardedModule):
            r = module.load_state_dict(state_dict, strict=strict)
            return list(r.missing_keys), list(r.unexpected_keys)
        missing, unexpected, errors = [], [], []
        module._load_from_state_dict(state_dict, prefix, {}, strict, missing, unexpected, errors)
        if errors:
            raise RuntimeError("load_state_dict: " + "; ".join(errors))
        for name, child in module.named_children():
            m, u = self._load_state_dict(child, filter_state_dict(state_dict, prefix + name), "", strict)
            missing.extend(m)
            unexpected.extend(u)
        return missing, unexpected

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        yield from self.module.named_parameters(prefix, recurse)

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        yield from self.module.named_buffers(prefix, recurse)

    @property
    def fused_optimizer(self) -> KeyedOptimizer:
        return self._optim

    @property
    def plan(self) -> ShardingPlan:
        return self._plan
