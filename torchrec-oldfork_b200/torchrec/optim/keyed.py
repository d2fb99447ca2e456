"""Keyed optimizers (reference torchrec/optim/keyed.py: KeyedOptimizer :30, CombinedOptimizer :236,
KeyedOptimizerWrapper :328): torch.optim.Optimizer subclasses addressed by parameter name so
fused (in-backward) and dense optimizers can be combined and checkpointed together."""
from __future__ import annotations

import copy
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
        self.defaults: Dict[str, Any] = {"_save_param_groups": False}
        ids = {id(p) for p in params.values()}
        stray = [k for k in state if id(k) not in ids]
        if stray:
            raise ValueError(f"All state keys must be params. The following keys are not: {stray}.")

    def state_dict(self) -> Dict[str, Any]:
        """Flat `{"state": {<param key>: {...}}}` keyed by parameter name; `param_groups` only after
        save_param_groups(True) (reference :66-98)."""
        name_of = {id(p): n for n, p in self.params.items()}
        ret: Dict[str, Any] = {"state": {name_of[id(p)]: s for p, s in self.state.items()}}
        if self.defaults.get("_save_param_groups", False):
            groups = []
            for g in self.param_groups:
                d = {"params": sorted(name_of[id(p)] for p in g["params"])}
                d.update({k: copy.deepcopy(v) for k, v in g.items() if k != "params"})
                groups.append(d)
            ret["param_groups"] = groups
        return ret

    def post_load_state_dict(self) -> None:
        pass

    def load_state_dict(self, state_dict: Mapping[str, Any]) -> None:
        """Strict in-place load (reference :103-196): the same parameters, the same state keys per
        parameter; tensors are copied into the live state (ShardedTensors shard by shard)."""
        new_state = state_dict["state"]
        state, params = self.state, self.params
        if len(state) != len(new_state):
            raise ValueError(f"Different parameter count: {len(state)} vs {len(new_state)}")
        for name, p in params.items():
            if p not in state:
                continue
            if name not in new_state:
                raise ValueError(f"Parameter {name} not found")
            if len(state[p]) != len(new_state[name]):
                raise ValueError(f"Different state size: {len(state[p])} vs {len(new_state[name])}")
            for k, cur in state[p].items():
                if k not in new_state[name]:
                    raise ValueError(f"State key {k} not found for param {name}")
                v = new_state[name][k]
                if isinstance(cur, torch.Tensor):
                    assert isinstance(v, torch.Tensor), f"{name}/{k}: tensor state loaded from {type(v)}"
                    _copy_state(cur, v)
                else:
                    state[p][k] = copy.deepcopy(v)
        if self.defaults.get("_save_param_groups", False):
            new_groups = state_dict["param_groups"]
            if len(self.param_groups) != len(new_groups):
                raise ValueError(f"Different param_groups count: {len(self.param_groups)} vs {len(new_groups)}")
            name_of = {id(p): n for n, p in params.items()}
            mine = {"/".join(sorted(name_of[id(p)] for p in g["params"])): g for g in self.param_groups}
            theirs = {"/".join(sorted(g["params"])): g for g in new_groups}
            for key, g in mine.items():
                if key not in theirs:
                    raise ValueError(f"Group {key} not found")
                ng = theirs[key]
                if len(g) != len(ng):
                    raise ValueError(f"Different param_group size: {len(g)} vs {len(ng)}")
                for k in g:
                    if k not in ng:
                        raise ValueError(f"Group key {k} not found for group {key}")
                    if k != "params":
                        g[k] = copy.deepcopy(ng[k])
        self.post_load_state_dict()

    def save_param_groups(self, save: bool) -> None:
        self.defaults["_save_param_groups"] = save

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
        self.defaults = {"_save_param_groups": self._optims[0][1].defaults.get("_save_param_groups", False)
                         if self._optims else False}
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

    def post_load_state_dict(self) -> None:
        for _, o in self._optims:
            o.post_load_state_dict()

    def save_param_groups(self, save: bool) -> None:
        self.defaults["_save_param_groups"] = save
        for _, o in self._optims:
            o.save_param_groups(save)


class KeyedOptimizerWrapper(KeyedOptimizer):
    """Wraps a plain torch optimizer built from a name->parameter mapping."""

    def __init__(self, params: Mapping[str, torch.Tensor], optim_factory: OptimizerFactory) -> None:
        self._optimizer: optim.Optimizer = optim_factory(list(params.values()))
        super().__init__(params, self._optimizer.state, self._optimizer.param_groups)

    def zero_grad(self, set_to_none: bool = False) -> None:
        self._optimizer.zero_grad(set_to_none=set_to_none)

    def step(self, closure: Any = None) -> None:
        self._optimizer.step(closure=closure)
