"""Keyed optimizers (reference torchrec/optim/keyed.py: KeyedOptimizer :30, CombinedOptimizer :236,
KeyedOptimizerWrapper :328): torch.optim.Optimizer subclasses addressed by parameter name so
fused (in-backward) and dense optimizers can be combined and checkpointed together."""
from __future__ import annotations

from typing import Any, Callable, Collection, Dict, List, Mapping, Optional, Tuple, Union

import torch
from torch import optim

OptimizerFactory = Callable[[List[torch.Tensor]], optim.Optimizer]


def _copy_state(dst: torch.Tensor, src: torch.Tensor) -> None:
    """In-place load of one state tensor; ShardedTensors are copied shard by shard (reference :188-205:
    same number of local shards, same metadata)."""
    from torch.distributed._shard.sharded_tensor import ShardedTensor
    if isinstance(dst, ShardedTensor) or isinstance(src, ShardedTensor):
        d = dst.local_shards() if isinstance(dst, ShardedTensor) else None
        s = src.local_shards() if isinstance(src, ShardedTensor) else None
        if d is not None and s is not None:
            assert len(d) == len(s), f"{len(d)} local shards to load from {len(s)}"
            for ds, ss in zip(d, s):
                assert ds.metadata.shard_offsets == ss.metadata.shard_offsets and \
                    ds.metadata.shard_sizes == ss.metadata.shard_sizes, (ds.metadata, ss.metadata)
                ds.tensor.detach().copy_(ss.tensor)
        elif d is not None:
            assert len(d) == 1
            d[0].tensor.detach().copy_(src)
        else:
            assert len(s) == 1
            dst.detach().copy_(s[0].tensor)
    else:
        dst.detach().copy_(src)


class KeyedOptimizer(optim.Optimizer):
    def __init__(self, params: Mapping[str, torch.Tensor], state: Mapping[Any, Any],
                 param_groups: Collection[Mapping[str, Any]]) -> None:
        torch._C._log_api_usage_once(f"torchrec.optim.{self.__class__.__name__}")
        self._optimizer_step_pre_hooks = {}
        self._optimizer_step_post_hooks = {}
        self.state = state
        self.param_groups = param_groups
        self.params = params
        self.defaults: Dict[str, Any] = {}
        flat = {id(p) for g in param_groups for p in g["params"]}
        assert flat == {id(p) for p in params.values()}, "param_groups and params must hold the same tensors"

    def state_dict(self) -> Dict[str, Any]:
        name_of = {id(p): n for n, p in self.params.items()}
        groups = []
        for g in self.param_groups:
            d = {k: v for k, v in g.items() if k != "params"}
            d["params"] = sorted(name_of[id(p)] for p in g["params"])
            groups.append(d)
        state = {name_of[id(p)]: s for p, s in self.state.items() if id(p) in name_of}
        return {"state": state, "param_groups": groups}

    def load_state_dict(self, state_dict: Mapping[str, Any]) -> None:
        new_state = state_dict["state"]
        for name, p in self.params.items():
            if name not in new_state or p not in self.state:
                continue
            for k, v in new_state[name].items():
                cur = self.state[p].get(k)
                if isinstance(cur, torch.Tensor) and isinstance(v, torch.Tensor):
                    _copy_state(cur, v)
                else:
                    self.state[p][k] = v
        for g, ng in zip(self.param_groups, state_dict["param_groups"]):
            for k, v in ng.items():
                if k != "params":
                    g[k] = v

    def add_param_group(self, param_group: Any) -> None:
        raise NotImplementedError()

    def init_state(self, sparse_grad_parameter_names: Optional[Collection[str]] = None) -> None:
        for name, p in self.params.items():
            if p.requires_grad:
                p.grad = torch.zeros_like(p)
        self.step(closure=None)


class CombinedOptimizer(KeyedOptimizer):
    """Runs several KeyedOptimizers as one; parameter names get the optimizer's key as prefix."""

    def __init__(self, optims: List[Union[KeyedOptimizer, Tuple[str, KeyedOptimizer]]]) -> None:
        self._optims: List[Tuple[str, KeyedOptimizer]] = []
        for o in optims:
            self._optims.append(o if isinstance(o, tuple) else ("", o))
        self._optimizer_step_pre_hooks = {}
        self._optimizer_step_post_hooks = {}
        self.defaults = {}
        seen = set()
        for key, o in self._optims:
            for n in o.params:
                full = self.prepend_opt_key(n, key)
                if full in seen:
                    raise ValueError(f"Duplicate param key {full}")
                seen.add(full)

    @staticmethod
    def prepend_opt_key(name: str, opt_key: str) -> str:
        return f"{opt_key}.{name}" if opt_key else name

    @property
    def optimizers(self) -> List[Tuple[str, KeyedOptimizer]]:
        return self._optims

    @property
    def param_groups(self):
        return [g for _, o in self._optims for g in o.param_groups]

    @property
    def params(self) -> Mapping[str, torch.Tensor]:
        return {self.prepend_opt_key(n, k): p for k, o in self._optims for n, p in o.params.items()}

    @property
    def state(self):
        out = {}
        for _, o in self._optims:
            out.update(o.state)
        return out

    def zero_grad(self, set_to_none: bool = False) -> None:
        for _, o in self._optims:
            o.zero_grad(set_to_none=set_to_none)

    def step(self, closure: Any = None) -> None:
        for _, o in self._optims:
            o.step(closure=closure)

    def state_dict(self) -> Dict[str, Any]:
        return {"optimizers": {k: o.state_dict() for k, o in self._optims}}

    def load_state_dict(self, state_dict: Mapping[str, Any]) -> None:
        for k, o in self._optims:
            o.load_state_dict(state_dict["optimizers"][k])


class KeyedOptimizerWrapper(KeyedOptimizer):
    """Wraps a plain torch optimizer built from a name->parameter mapping."""

    def __init__(self, params: Mapping[str, torch.Tensor], optim_factory: OptimizerFactory) -> None:
        self._optimizer: optim.Optimizer = optim_factory(list(params.values()))
        super().__init__(params, self._optimizer.state, self._optimizer.param_groups)

    def zero_grad(self, set_to_none: bool = False) -> None:
        self._optimizer.zero_grad(set_to_none=set_to_none)

    def step(self, closure: Any = None) -> None:
        self._optimizer.step(closure=closure)
