"""Variable batch size per rank (reference torchrec/distributed/sharding/vb_sharding.py:
VariableBatchShardingContext :19, BaseVariableBatchEmbeddingDist :39): the pooled output dist has
to know how many samples every rank brought, which only the input dist learns."""
from __future__ import annotations

import abc
from dataclasses import dataclass, field
from typing import Generic, List, Optional, TypeVar

import torch

from torchrec.distributed.embedding_sharding import BaseEmbeddingDist
from torchrec.distributed.types import Awaitable

T = TypeVar("T")


@dataclass
class VariableBatchShardingContext:
    """batch_size_per_rank[r]: samples rank r contributed to this step (host list, and the int32
    device tensor the exchange produced)."""
    batch_size_per_rank: List[int] = field(default_factory=list)
    batch_size_per_rank_tensor: Optional[torch.Tensor] = None

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        if self.batch_size_per_rank_tensor is not None and self.batch_size_per_rank_tensor.is_cuda:
            self.batch_size_per_rank_tensor.record_stream(stream)


class BaseVariableBatchEmbeddingDist(BaseEmbeddingDist[T], Generic[T]):
    """Model-parallel -> data-parallel output dist that takes the step's batch sizes."""

    @abc.abstractmethod
    def forward(self, local_embs: T, sharding_ctx: VariableBatchShardingContext) -> Awaitable[torch.Tensor]:
        ...
