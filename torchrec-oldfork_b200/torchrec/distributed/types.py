"""Core sharding abstractions (reference torchrec/distributed/types.py: ShardingType :64,
Awaitable :108, LazyAwaitable :149, ShardingEnv :328, ShardedModule :366, ParameterSharding,
ShardingPlan, ModuleSharder)."""
from __future__ import annotations

import abc
from dataclasses import dataclass
from enum import Enum, unique
from typing import Dict, Generic, List, Optional, Type, TypeVar

import torch
import torch.distributed as dist
from torch import nn

W = TypeVar("W")
M = TypeVar("M", bound=nn.Module)
CompIn = TypeVar("CompIn")
DistOut = TypeVar("DistOut")
Out = TypeVar("Out")


@unique
class ShardingType(Enum):
    DATA_PARALLEL = "data_parallel"
    TABLE_WISE = "table_wise"
    COLUMN_WISE = "column_wise"
    ROW_WISE = "row_wise"
    TABLE_ROW_WISE = "table_row_wise"
    TABLE_COLUMN_WISE = "table_column_wise"


class ParameterStorage(Enum):
    HBM = "hbm"
    DDR = "ddr"


@unique
class ComputeKernel(Enum):
    DEFAULT = "default"


class Awaitable(abc.ABC, Generic[W]):
    """Result of an asynchronous step; wait() blocks the calling stream, not the host."""

    def __init__(self) -> None:
        self._callbacks: List = []

    @abc.abstractmethod
    def _wait_impl(self) -> W:
        ...

    def wait(self) -> W:
        ret = self._wait_impl()
        for cb in self._callbacks:
            ret = cb(ret)
        return ret

    @property
    def callbacks(self) -> List:
        return self._callbacks


class NoWait(Awaitable[W]):
    def __init__(self, obj: W) -> None:
        super().__init__()
        self._obj = obj

    def _wait_impl(self) -> W:
        return self._obj


class LazyAwaitable(Awaitable[W]):
    """Awaitable whose result is computed on first wait() and cached (reference :149-325 also
    forwards tensor operators; callers here wait explicitly)."""

    def __init__(self) -> None:
        super().__init__()
        self._result: Optional[W] = None
        self._done = False

    def wait(self) -> W:
        if not self._done:
            self._result = super().wait()
            self._done = True
        return self._result  # type: ignore[return-value]

    def __getattr__(self, name: str):
        # transparently wait when the result is used like the awaited object (reference :192-325)
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.wait(), name)


@dataclass
class ShardMetadata:
    shard_offsets: List[int]
    shard_sizes: List[int]
    placement: Optional[str] = None  # "rank:R/cuda:R"


@dataclass
class EnumerableShardingSpec:
    shards: List[ShardMetadata]


@dataclass
class ParameterSharding:
    """How one embedding table is sharded (reference types.py ParameterSharding)."""
    sharding_type: str
    compute_kernel: str
    ranks: Optional[List[int]] = None
    sharding_spec: Optional[EnumerableShardingSpec] = None


@dataclass
class ShardingPlan:
    plan: Dict[str, Dict[str, ParameterSharding]]

    def get_plan_for_module(self, module_path: str) -> Optional[Dict[str, ParameterSharding]]:
        return self.plan.get(module_path, None)

    def __str__(self) -> str:
        return str(self.plan)


class ShardingEnv:
    def __init__(self, world_size: int, rank: int, pg: Optional[dist.ProcessGroup] = None) -> None:
        self.world_size, self.rank, self.process_group = world_size, rank, pg

    @classmethod
    def from_process_group(cls, pg: dist.ProcessGroup) -> "ShardingEnv":
        return cls(dist.get_world_size(pg), dist.get_rank(pg), pg)

    @classmethod
    def from_local(cls, world_size: int, rank: int) -> "ShardingEnv":
        return cls(world_size, rank, None)


class ShardedModuleContext:
    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        pass


class EmptyContext(ShardedModuleContext):
    pass


class ShardedModule(abc.ABC, nn.Module, Generic[CompIn, DistOut, Out]):
    """input_dist -> compute -> output_dist (reference types.py:366-440)."""

    def create_context(self) -> ShardedModuleContext:
        return EmptyContext()

    @abc.abstractmethod
    def input_dist(self, ctx, *input, **kwargs) -> Awaitable[CompIn]:
        ...

    @abc.abstractmethod
    def compute(self, ctx, dist_input: CompIn) -> DistOut:
        ...

    @abc.abstractmethod
    def output_dist(self, ctx, output: DistOut) -> LazyAwaitable[Out]:
        ...

    def compute_and_output_dist(self, ctx, input: CompIn) -> LazyAwaitable[Out]:
        return self.output_dist(ctx, self.compute(ctx, input))

    def forward(self, *input, **kwargs) -> LazyAwaitable[Out]:
        pipelined = getattr(self, "_pipelined_inputs", None)
        if pipelined:
            # TrainPipelineSparseDist ran input_dist for this batch on its data-dist stream already
            ctx, dist_input = pipelined.popleft()
            return self.compute_and_output_dist(ctx, dist_input)
        ctx = self.create_context()
        dist_input = self.input_dist(ctx, *input, **kwargs).wait()
        return self.compute_and_output_dist(ctx, dist_input)

    def sparse_grad_parameter_names(self, destination: Optional[List[str]] = None, prefix: str = "") -> List[str]:
        return [] if destination is None else destination


class ModuleSharder(abc.ABC, Generic[M]):
    @abc.abstractmethod
    def shard(self, module: M, params: Dict[str, ParameterSharding], env: ShardingEnv,
              device: Optional[torch.device] = None) -> ShardedModule:
        ...

    @property
    @abc.abstractmethod
    def module_type(self) -> Type[M]:
        ...

    def shardable_parameters(self, module: M) -> Dict[str, nn.Parameter]:
        return dict(module.named_parameters())

    def sharding_types(self, compute_device_type: str) -> List[str]:
        return [ShardingType.DATA_PARALLEL.value]

    def compute_kernels(self, sharding_type: str, compute_device_type: str) -> List[str]:
        return [ComputeKernel.DEFAULT.value]
