"""Sharded data movement modules (reference torchrec/distributed/dist_data.py: _get_recat :40,
_split_lengths :121, KJTAllToAllIndicesAwaitable :137, KJTAllToAllLengthsAwaitable :276,
KJTAllToAll :414, PooledEmbeddingsAllToAll :602, PooledEmbeddingsReduceScatter :745).

Same two-phase protocol (lengths, then values/weights, NCCL all_to_all_single over NVLink) and
the same rank-major -> feature-major recat through permute_2D_sparse_data.  One difference in
host traffic: the single D2H read after the lengths exchange fetches the [W, F_local] table of
per-(source rank, key) totals, which yields the receive split sizes AND the output KJT's
length_per_key, so the second sync of the reference (from_lengths_sync, :266) disappears.
"""
from __future__ import annotations

import os

from typing import List, Optional

import torch
import torch.distributed as dist
from torch import nn

from torchrec.distributed.comm_ops import alltoall_pooled, alltoall_sequence, reduce_scatter_pooled
from torchrec.distributed.types import Awaitable, NoWait
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor


def _get_recat(local_split: int, num_splits: int, stagger: int = 1,
               device: Optional[torch.device] = None,
               batch_size_per_rank: Optional[List[int]] = None) -> torch.Tensor:
    """Permutation taking (source rank, local feature) order to (local feature, source rank):
    _get_recat(2, 4, 1) == [0, 2, 4, 6, 1, 3, 5, 7]; _get_recat(2, 4, 2) == [0, 4, 2, 6, 1, 5, 3, 7].

    With `batch_size_per_rank` (variable batch size, reference :84-116) the segment permutation is
    expanded to one entry per (feature, sample) length slot - segment (r, f) holds
    batch_size_per_rank[r] slots - so permute_1D_sparse_data can apply it."""
    if local_split == 0:
        return torch.tensor([], device=device, dtype=torch.int32)
    per = num_splits // stagger
    rank_order = [x + per * y for x in range(per) for y in range(stagger)]
    recat = [f + r * local_split for f in range(local_split) for r in rank_order]
    recat_t = torch.tensor(recat, device=device, dtype=torch.int32)
    if batch_size_per_rank is None:
        return recat_t
    seg = [b for b in batch_size_per_rank for _ in range(local_split)]
    in_off, out_off = [0], [0]
    for b in seg:
        in_off.append(in_off[-1] + b)
    for r in recat:
        out_off.append(out_off[-1] + seg[r])
    return torch.ops.fbgemm.expand_into_jagged_permute(
        recat_t, torch.tensor(in_off, device=device, dtype=torch.int32),
        torch.tensor(out_off, device=device, dtype=torch.int32), out_off[-1])


def _split_lengths(splits: List[int], keys: List[str], offset_per_key: List[int]) -> List[int]:
    """Number of ids destined to each rank: keys are ordered by destination, `splits` keys per rank."""
    out, k = [], 0
    for s in splits:
        out.append(offset_per_key[k + s] - offset_per_key[k])
        k += s
    return out


class KJTAllToAllIndicesAwaitable(Awaitable[KeyedJaggedTensor]):
    """Phase 2: values (+weights) exchange, then recat to feature-major."""

    def __init__(self, pg, input: KeyedJaggedTensor, lengths: torch.Tensor, splits: List[int], keys: List[str],
                 recat: torch.Tensor, in_lengths_per_worker: List[int], out_lengths_per_worker: List[int],
                 batch_size_per_rank: List[int], length_per_key: Optional[List[int]] = None,
                 variable_batch_size: bool = False) -> None:
        super().__init__()
        self._variable_batch_size = variable_batch_size
        self._pg, self._input, self._lengths, self._splits = pg, input, lengths, splits
        self._keys, self._recat = keys, recat
        self._workers = dist.get_world_size(pg)
        self._batch_size_per_rank = batch_size_per_rank
        self._length_per_key = length_per_key
        # ids sent to / received from each rank: the sequence output dist sends one row back per id
        self._in_lengths_per_worker = list(in_lengths_per_worker)
        self._out_lengths_per_worker = list(out_lengths_per_worker)
        self._works: List = []
        self._values: Optional[torch.Tensor] = None
        self._weights: Optional[torch.Tensor] = None
        if self._workers == 1:
            return
        total = sum(out_lengths_per_worker)

        def exchange(t: torch.Tensor) -> torch.Tensor:
            src = t.contiguous().view(-1)
            dst = src.new_empty(total)
            self._works.append(dist.all_to_all_single(dst, src, output_split_sizes=out_lengths_per_worker,
                                                      input_split_sizes=in_lengths_per_worker, group=pg,
                                                      async_op=True))
            return dst

        self._values = exchange(input.values())
        if input.weights_or_none() is not None:
            self._weights = exchange(input.weights())

    def _wait_impl(self) -> KeyedJaggedTensor:
        if self._workers == 1:
            return self._input.sync()
        for w in self._works:
            w.wait()
        lengths, values, weights = self._lengths, self._values, self._weights
        if self._recat.numel() > 0:
            if self._variable_batch_size:  # per-slot permutation (reference :248-255)
                lengths, values, weights = torch.ops.fbgemm.permute_1D_sparse_data(
                    self._recat, lengths.view(-1), values, weights, values.numel())
            else:
                rows = self._workers * self._splits[dist.get_rank(self._pg)]
                lengths, values, weights = torch.ops.fbgemm.permute_2D_sparse_data(
                    self._recat, lengths.view(rows, -1), values, weights, values.numel())
                lengths = lengths.view(-1)
        return KeyedJaggedTensor(keys=self._keys, values=values, weights=weights, lengths=lengths,
                                 stride=sum(self._batch_size_per_rank), length_per_key=self._length_per_key)


class KJTAllToAllLengthsAwaitable(Awaitable[KJTAllToAllIndicesAwaitable]):
    """Phase 1: lengths exchange; wait() reads the receive sizes and starts phase 2.

    variable_batch_size (reference :321-345): the ranks first exchange their batch sizes (one int
    each, read on the host - it sizes every later buffer), the recat permutation is rebuilt per
    call with one slot per (feature, sample), and the lengths exchange has ragged receive splits."""

    def __init__(self, pg, input: KeyedJaggedTensor, splits: List[int], keys: List[str], recat: torch.Tensor,
                 stagger: int = 1, variable_batch_size: bool = False) -> None:
        super().__init__()
        self._pg, self._input, self._splits, self._keys, self._recat = pg, input, splits, keys, recat
        self._workers = dist.get_world_size(pg)
        self._local_split = splits[dist.get_rank(pg)]
        self._variable_batch_size = variable_batch_size
        B = input.stride()
        self._batch_size_per_rank = [B] * self._workers
        self._batch_size_per_rank_tensor: Optional[torch.Tensor] = None
        self._lengths = input.lengths()
        self._in_lengths_per_worker: List[int] = []
        self._work = None
        if self._workers == 1:
            return
        src = input.lengths().contiguous().view(-1)
        if variable_batch_size:
            mine = torch.full((self._workers,), B, device=src.device, dtype=torch.int32)
            theirs = torch.empty_like(mine)
            dist.all_to_all_single(theirs, mine, output_split_sizes=[1] * self._workers,
                                   input_split_sizes=[1] * self._workers, group=pg)
            self._batch_size_per_rank_tensor = theirs
            self._batch_size_per_rank = theirs.cpu().tolist()
            self._recat = _get_recat(self._local_split, len(splits), stagger, src.device,
                                     self._batch_size_per_rank)
        self._output_split_sizes = [self._local_split * b for b in self._batch_size_per_rank]
        self._lengths = src.new_empty(sum(self._output_split_sizes))
        # the lengths exchange has fixed split sizes; the (data-dependent) send sizes of the values
        # exchange are read in wait(), so issuing phase 1 never blocks the host
        self._work = dist.all_to_all_single(
            self._lengths, src, output_split_sizes=self._output_split_sizes,
            input_split_sizes=[s * B for s in splits], group=pg, async_op=True)
        # Host reads, PREFETCHED: the two small tables wait() needs on the host (per-key totals of what I send,
        # [W, F_local] totals of what I receive) are reduced and copied to pinned memory right behind the
        # exchange, on the issuing stream.  wait() then only synchronises an event that has usually long fired,
        # instead of queueing reduction kernels + a blocking copy behind whatever the GPU is busy with by then
        # (measured at N = 8: 1.15 ms of host stall per step in the blocking reads).
        # OFF by default (B200_KJT_PREFETCH=1): allocating the pinned buffers can synchronise the device, and a
        # host blocked on a device that is spinning in a cross-rank barrier deadlocked the 4-rank bench; it needs
        # preallocated pinned rings before it can be the default.
        self._send_lpk_host = self._recv_table_host = self._prefetch_event = None
        if src.is_cuda and not variable_batch_size and os.environ.get("B200_KJT_PREFETCH", "0") == "1":
            if input._length_per_key is None and len(input.keys()) > 0:
                lpk = input.lengths().view(len(input.keys()), -1).sum(dim=1)
                self._send_lpk_host = torch.empty(lpk.shape, dtype=lpk.dtype, pin_memory=True)
                self._send_lpk_host.copy_(lpk, non_blocking=True)
            if self._local_split > 0:
                self._work.wait()  # orders the current STREAM behind the exchange; the host does not block
                table = self._lengths.view(self._workers, self._local_split, -1).sum(dim=2)
                self._recv_table_host = torch.empty(table.shape, dtype=table.dtype, pin_memory=True)
                self._recv_table_host.copy_(table, non_blocking=True)
            self._prefetch_event = torch.cuda.Event()
            self._prefetch_event.record()

    def _wait_impl(self) -> KJTAllToAllIndicesAwaitable:
        out_lengths_per_worker: List[int] = []
        length_per_key: Optional[List[int]] = None
        if self._workers > 1:
            if self._prefetch_event is not None:
                self._prefetch_event.synchronize()
                if self._send_lpk_host is not None and self._input._length_per_key is None:
                    self._input._length_per_key = self._send_lpk_host.tolist()
            self._in_lengths_per_worker = _split_lengths(self._splits, self._input.keys(),
                                                         self._input.offset_per_key())
            self._work.wait()
            if self._local_split > 0:
                # one D2H read: [W, F_local] totals -> receive splits and per-key totals
                if self._recv_table_host is not None:
                    table = self._recv_table_host
                elif self._variable_batch_size:
                    table = torch.stack([
                        (c.view(self._local_split, -1).sum(dim=1) if c.numel()
                         else c.new_zeros(self._local_split))
                        for c in self._lengths.split(self._output_split_sizes)]).cpu()
                else:
                    table = self._lengths.view(self._workers, self._local_split, -1).sum(dim=2).cpu()
                out_lengths_per_worker = table.sum(dim=1).tolist()
                length_per_key = table.sum(dim=0).tolist()
            else:
                out_lengths_per_worker = [0] * self._workers
                length_per_key = []
        return KJTAllToAllIndicesAwaitable(
            self._pg, self._input, self._lengths, self._splits, self._keys, self._recat,
            self._in_lengths_per_worker, out_lengths_per_worker, self._batch_size_per_rank, length_per_key,
            self._variable_batch_size and self._workers > 1)


class KJTAllToAll(nn.Module):
    """Redistributes a KJT whose keys are ordered by destination rank (`splits[r]` keys to rank r).
    variable_batch_size: every rank may bring a different batch size; the output KJT has
    stride sum_r B_r with rank r's samples at columns [sum_{q<r} B_q, +B_r) of every key."""

    def __init__(self, pg: dist.ProcessGroup, splits: List[int], device: Optional[torch.device] = None,
                 stagger: int = 1, variable_batch_size: bool = False) -> None:
        super().__init__()
        assert len(splits) == dist.get_world_size(pg)
        self._pg, self._splits = pg, splits
        self._stagger, self._variable_batch_size = stagger, variable_batch_size
        self._no_dist = all(s == 0 for s in splits)
        self._splits_cumsum = [0]
        for s in splits:
            self._splits_cumsum.append(self._splits_cumsum[-1] + s)
        self.register_buffer("_recat", _get_recat(splits[dist.get_rank(pg)], len(splits), stagger, device),
                             persistent=False)

    def forward(self, input: KeyedJaggedTensor) -> Awaitable[KJTAllToAllIndicesAwaitable]:
        rank = dist.get_rank(self._pg)
        keys = input.keys()[self._splits_cumsum[rank]:self._splits_cumsum[rank + 1]]
        return KJTAllToAllLengthsAwaitable(self._pg, input, self._splits, keys, self._recat, self._stagger,
                                           self._variable_batch_size)


class PooledEmbeddingsAwaitable(Awaitable[torch.Tensor]):
    def __init__(self, tensor_awaitable: Awaitable[torch.Tensor]) -> None:
        super().__init__()
        self._tensor_awaitable = tensor_awaitable

    def _wait_impl(self) -> torch.Tensor:
        return self._tensor_awaitable.wait()


class PooledEmbeddingsAllToAll(nn.Module):
    """[B_global, D_local_sum] -> [B_local, sum_r dim_sum_per_rank[r]] (table-wise output dist)."""

    def __init__(self, pg: dist.ProcessGroup, dim_sum_per_rank: List[int],
                 device: Optional[torch.device] = None, callbacks=None) -> None:
        super().__init__()
        self._pg, self._dim_sum_per_rank = pg, dim_sum_per_rank
        self._callbacks = callbacks or []

    def forward(self, local_embs: torch.Tensor, batch_size_per_rank: Optional[List[int]] = None
                ) -> PooledEmbeddingsAwaitable:
        if batch_size_per_rank is None:
            W = dist.get_world_size(self._pg)
            batch_size_per_rank = [local_embs.shape[0] // W] * W
        aw = PooledEmbeddingsAwaitable(
            alltoall_pooled(local_embs, batch_size_per_rank, self._dim_sum_per_rank, group=self._pg))
        aw.callbacks.extend(self._callbacks)
        return aw

    @property
    def callbacks(self):
        return self._callbacks


class PooledEmbeddingsReduceScatter(nn.Module):
    """Sum over ranks of [B_global, D], scattered along the batch (row-wise output dist)."""

    def __init__(self, pg: dist.ProcessGroup) -> None:
        super().__init__()
        self._pg = pg

    def forward(self, local_embs: torch.Tensor) -> PooledEmbeddingsAwaitable:
        return PooledEmbeddingsAwaitable(reduce_scatter_pooled(local_embs, self._pg))


class SequenceEmbeddingsAwaitable(Awaitable[torch.Tensor]):
    def __init__(self, tensor_awaitable: Awaitable[torch.Tensor], unbucketize_permute_tensor: Optional[torch.Tensor],
                 embedding_dim: int) -> None:
        super().__init__()
        self._tensor_awaitable = tensor_awaitable
        self._unbucketize_permute_tensor = unbucketize_permute_tensor
        self._embedding_dim = embedding_dim

    def _wait_impl(self) -> torch.Tensor:
        ret = self._tensor_awaitable.wait()
        if self._unbucketize_permute_tensor is not None:  # row-wise: back to the pre-bucketize id order
            ret = torch.index_select(ret.view(-1, self._embedding_dim), 0, self._unbucketize_permute_tensor)
        return ret


class SequenceEmbeddingAllToAll(nn.Module):
    """Sends every looked-up row back to the rank that owns its sample (reference :841-932)."""

    def __init__(self, pg: dist.ProcessGroup, features_per_rank: List[int],
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._pg = pg
        W, mine = dist.get_world_size(pg), features_per_rank[dist.get_rank(pg)]
        self.register_buffer("_forward_recat_tensor",
                             torch.tensor([j + i * W for j in range(W) for i in range(mine)], device=device,
                                          dtype=torch.int32), persistent=False)
        self.register_buffer("_backward_recat_tensor",
                             torch.tensor([i + j * mine for i in range(mine) for j in range(W)], device=device,
                                          dtype=torch.int32), persistent=False)

    def forward(self, local_embs: torch.Tensor, lengths: torch.Tensor, input_splits: List[int],
                output_splits: List[int], unbucketize_permute_tensor: Optional[torch.Tensor] = None
                ) -> SequenceEmbeddingsAwaitable:
        aw = alltoall_sequence(local_embs, self._forward_recat_tensor, self._backward_recat_tensor, lengths,
                               input_splits, output_splits, self._pg)
        return SequenceEmbeddingsAwaitable(aw, unbucketize_permute_tensor, local_embs.shape[1])
