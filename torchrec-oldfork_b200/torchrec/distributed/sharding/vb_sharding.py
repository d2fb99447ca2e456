"""Per-step batch sizes of a variable-batch exchange (reference
torchrec/distributed/sharding/vb_sharding.py: VariableBatchShardingContext :19,
BaseVariableBatchEmbeddingDist :39).  Only the input dist learns how many samples every rank brought;
the pooled output dist needs the same numbers to size its ragged all-to-all."""
from __future__ import annotations

from typing import Generic, List, Optional, Sequence, Tuple, TypeVar

import torch

from torchrec.distributed.embedding_sharding import BaseEmbeddingDist
from torchrec.distributed.types import Awaitable

T = TypeVar("T")


class VariableBatchShardingContext:
    """batch_size_per_rank[r] = samples rank r contributed this step; batch_size_per_rank_tensor is the
    int32 device tensor the batch-size exchange produced (kept alive across streams by record_stream)."""

    __slots__ = ("batch_size_per_rank", "batch_size_per_rank_tensor")

    def __init__(self, batch_size_per_rank: Optional[Sequence[int]] = None,
                 batch_size_per_rank_tensor: Optional[torch.Tensor] = None) -> None:
        self.batch_size_per_rank: List[int] = list(batch_size_per_rank or [])
        self.batch_size_per_rank_tensor = batch_size_per_rank_tensor

    @classmethod
    def from_lengths_awaitable(cls, lengths_awaitable) -> "VariableBatchShardingContext":
        """From a KJTAllToAllLengthsAwaitable: the sizes were exchanged when it was constructed."""
        return cls(lengths_awaitable._batch_size_per_rank, lengths_awaitable._batch_size_per_rank_tensor)

    @property
    def global_batch_size(self) -> int:
        return sum(self.batch_size_per_rank)

    def rows_of(self, rank: int) -> Tuple[int, int]:
        """[start, stop) of rank's samples inside a [global batch, ...] tensor."""
        start = sum(self.batch_size_per_rank[:rank])
        return start, start + self.batch_size_per_rank[rank]

    def record_stream(self, stream: "torch.cuda.streams.Stream") -> None:
        t = self.batch_size_per_rank_tensor
        if t is not None and t.is_cuda:
            t.record_stream(stream)


class BaseVariableBatchEmbeddingDist(BaseEmbeddingDist[T], Generic[T]):
    """Output dist whose forward also takes the step's VariableBatchShardingContext."""

    def forward(self, local_embs: T, sharding_ctx: VariableBatchShardingContext) -> Awaitable[torch.Tensor]:
        raise NotImplementedError
