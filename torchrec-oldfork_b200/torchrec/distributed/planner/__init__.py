"""Minimal sharding planner (the reference's cost-model planner, torchrec/distributed/planner/, is
out of scope — SURVEY.md §2a T14; its A100-era constants would mis-plan B200 anyway).

Plans are explicit: `ParameterConstraints(sharding_types=[...], compute_kernels=[...])` per table
decide the sharding type; table-wise tables are placed greedily on the rank with the least LOOKUP
traffic (bytes gathered per sample = embedding_dim x element size: a 3-row table costs a step as much
as a 40M-row one), ties broken by stored bytes, subject to the HBM capacity; the result is
deterministic, so every rank computes the same plan without a broadcast."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

import torch.distributed as dist
from torch import nn

from torchrec.distributed.embedding_types import EmbeddingComputeKernel
from torchrec.distributed.types import ModuleSharder, ParameterSharding, ShardingPlan, ShardingType

HBM_CAP: int = 180 * 1024 ** 3  # B200


@dataclass
class ParameterConstraints:
    sharding_types: Optional[List[str]] = None
    compute_kernels: Optional[List[str]] = None
    min_partition: Optional[int] = None
    caching_ratio: Optional[float] = None


@dataclass
class Topology:
    world_size: int
    compute_device: str = "cuda"
    hbm_cap: int = HBM_CAP
    local_world_size: Optional[int] = None


class EmbeddingShardingPlanner:
    def __init__(self, topology: Topology, constraints: Optional[Dict[str, ParameterConstraints]] = None,
                 **_ignored) -> None:
        self._topology = topology
        self._constraints = constraints or {}

    def collective_plan(self, module: nn.Module, sharders: List[ModuleSharder],
                        pg: Optional[dist.ProcessGroup] = None) -> ShardingPlan:
        return self.plan(module, sharders)

    def plan(self, module: nn.Module, sharders: List[ModuleSharder]) -> ShardingPlan:
        W = self._topology.world_size
        by_type = {s.module_type: s for s in sharders}
        load = [0] * W      # stored bytes per rank
        traffic = [0] * W   # bytes gathered per sample per rank
        plan: Dict[str, Dict[str, ParameterSharding]] = {}
        for path, child in module.named_modules():
            sharder = by_type.get(type(child))
            if sharder is None:
                continue
            entry: Dict[str, ParameterSharding] = {}
            params = sharder.shardable_parameters(child)
            # place big tables first so the greedy balance is tight
            for name in sorted(params, key=lambda n: -params[n].numel()):
                p = params[name]
                c = self._constraints.get(name, ParameterConstraints())
                allowed = c.sharding_types or sharder.sharding_types(self._topology.compute_device)
                kernels = c.compute_kernels or sharder.compute_kernels(allowed[0], self._topology.compute_device)
                nbytes = p.numel() * p.element_size()
                st = allowed[0]
                if c.sharding_types is None and nbytes > self._topology.hbm_cap // 2 and \
                        ShardingType.ROW_WISE.value in allowed and W > 1:
                    st = ShardingType.ROW_WISE.value
                row_bytes = (p.shape[1] if p.dim() == 2 else 1) * p.element_size()
                if st == ShardingType.TABLE_WISE.value:
                    fits = [i for i in range(W) if load[i] + nbytes <= self._topology.hbm_cap] or list(range(W))
                    r = min(fits, key=lambda i: (traffic[i], load[i], i))
                    load[r] += nbytes
                    traffic[r] += row_bytes
                    entry[name] = ParameterSharding(st, kernels[0], ranks=[r])
                elif st == ShardingType.ROW_WISE.value:
                    for i in range(W):
                        load[i] += nbytes // W
                        traffic[i] += row_bytes  # every rank looks every sample up in its row block
                    entry[name] = ParameterSharding(st, kernels[0], ranks=list(range(W)))
                else:
                    raise ValueError(f"sharding type {st} is not supported by this package")
            plan[path] = entry
        return ShardingPlan(plan)
