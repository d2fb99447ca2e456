"""Pooled-embedding collectives as autograd ops (reference torchrec/distributed/comm_ops.py:
alltoall_pooled :203, All2All_Pooled_Req :462 / _Wait :533, reduce_scatter_pooled :382,
ReduceScatter_Req :848 / _Wait :890, GRADIENT_DIVISION :35).

Each collective is a Req/Wait pair so forward issues the NCCL op asynchronously and backward
issues the transposed collective asynchronously.  The layout changes around the all-to-all —
the split+view+cat of :555-560 and _recat_pooled_embedding_grad_out :418-428 with the 1/W
scaling of :527-528 — run as one sm_100a kernel (b200_pooled_a2a_recat) on CUDA tensors.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.distributed as dist
from torch.autograd import Function

from torchrec.distributed.types import Awaitable, NoWait

GRADIENT_DIVISION: bool = True


def set_gradient_division(val: bool) -> None:
    global GRADIENT_DIVISION
    GRADIENT_DIVISION = val


def _recat(src: torch.Tensor, dims: List[int], B_local: int, forward: bool, scale: float = 1.0) -> torch.Tensor:
    """forward: W contiguous blocks [B_local, dims[r]] -> [B_local, sum dims]; backward: inverse (x scale)."""
    if src.is_cuda and src.dtype == torch.float32 and all(d % 4 == 0 for d in dims):
        from b200_sparse import ops as _b200
        cum = [0]
        for d in dims:
            cum.append(cum[-1] + d)
        cum_t = torch.tensor(cum, dtype=torch.int32, device=src.device)
        return _b200.pooled_a2a_recat(src.reshape(-1), cum_t, B_local, forward, scale)
    if forward:
        parts = src.reshape(-1).split([B_local * d for d in dims])
        return torch.cat([p.view(B_local, d) for p, d in zip(parts, dims)], dim=1).reshape(-1)
    parts = src.view(B_local, -1).split(dims, dim=1)
    out = torch.cat([p.contiguous().view(-1) for p in parts])
    return out * scale if scale != 1.0 else out


class _Request:
    """State shared by a Req/Wait pair."""

    def __init__(self, pg: dist.ProcessGroup) -> None:
        self.pg = pg
        self.work = None
        self.tensor: Optional[torch.Tensor] = None
        self.info = None
        self.dummy: Optional[torch.Tensor] = None


class _AwaitTensor(Awaitable[torch.Tensor]):
    def __init__(self, wait_fn, req: _Request) -> None:
        super().__init__()
        self._fn, self._req = wait_fn, req

    def _wait_impl(self) -> torch.Tensor:
        return self._fn.apply(self._req.pg, self._req, self._req.dummy)


# ------------------------------------------------------------------ all-to-all (table-wise)
class _A2AInfo:
    def __init__(self, dim_sum_per_rank: List[int], batch_size_per_rank: List[int], B_local: int,
                 D_local_sum: int) -> None:
        self.dims, self.batch_sizes = dim_sum_per_rank, batch_size_per_rank
        self.B_global, self.B_local, self.D_local_sum = sum(batch_size_per_rank), B_local, D_local_sum


class All2All_Pooled_Req(Function):
    @staticmethod
    def forward(ctx, pg, req: _Request, info: _A2AInfo, embs: torch.Tensor) -> torch.Tensor:
        send = embs.contiguous().view(-1)
        in_splits = [info.D_local_sum * b for b in info.batch_sizes]  # rank r gets its B_r samples
        out_splits = [d * info.B_local for d in info.dims]
        recv = send.new_empty(sum(out_splits))
        req.work = dist.all_to_all_single(recv, send, output_split_sizes=out_splits,
                                          input_split_sizes=in_splits, group=pg, async_op=True)
        req.tensor, req.info = recv, info
        ctx.req, ctx.pg = req, pg
        return torch.zeros(1, device=embs.device)

    @staticmethod
    def backward(ctx, _unused):
        req = ctx.req
        req.work.wait()
        info: _A2AInfo = req.info
        grad = req.tensor.view(info.B_global, info.D_local_sum)
        req.tensor = None
        if GRADIENT_DIVISION:
            grad.div_(dist.get_world_size(ctx.pg))
        return None, None, None, grad


class All2All_Pooled_Wait(Function):
    @staticmethod
    def forward(ctx, pg, req: _Request, dummy: torch.Tensor) -> torch.Tensor:
        req.work.wait()
        info: _A2AInfo = req.info
        ctx.req, ctx.pg = req, pg
        out = _recat(req.tensor, info.dims, info.B_local, True).view(info.B_local, sum(info.dims))
        req.tensor = None
        return out

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        req, pg = ctx.req, ctx.pg
        info: _A2AInfo = req.info
        send = _recat(grad_out.contiguous(), info.dims, info.B_local, False)
        in_splits = [d * info.B_local for d in info.dims]
        out_splits = [info.D_local_sum * b for b in info.batch_sizes]
        recv = send.new_empty(sum(out_splits))
        req.work = dist.all_to_all_single(recv, send, output_split_sizes=out_splits,
                                          input_split_sizes=in_splits, group=pg, async_op=True)
        req.tensor = recv
        return None, None, req.dummy.new_zeros(1)


def alltoall_pooled(a2a_pooled_embs_tensor: torch.Tensor, batch_size_per_rank: List[int],
                    dim_sum_per_rank: List[int], dim_sum_per_rank_tensor: Optional[torch.Tensor] = None,
                    cumsum_dim_sum_per_rank_tensor: Optional[torch.Tensor] = None,
                    group: Optional[dist.ProcessGroup] = None) -> Awaitable[torch.Tensor]:
    """[B_global, D_local_sum] on every rank -> [B_local, sum_r D_r] (reference :203-257, same
    argument order).  batch_size_per_rank[r] samples go to rank r (B_global = their sum; the sizes
    may differ between ranks).  The two *_tensor arguments feed the reference's fast recat kernel;
    b200_pooled_a2a_recat builds its own column table, they are accepted and unused."""
    group = group if group is not None else dist.distributed_c10d._get_default_group()
    W = dist.get_world_size(group)
    if W <= 1:
        return NoWait(a2a_pooled_embs_tensor)
    B_global, D_local_sum = a2a_pooled_embs_tensor.shape
    assert len(batch_size_per_rank) == W and B_global == sum(batch_size_per_rank), (
        f"batch_size_per_rank {batch_size_per_rank} does not add up to {B_global} rows")
    req = _Request(group)
    # the dummy output links Req and Wait in the autograd graph
    req.dummy = All2All_Pooled_Req.apply(
        group, req, _A2AInfo(dim_sum_per_rank, list(batch_size_per_rank),
                             batch_size_per_rank[dist.get_rank(group)], D_local_sum), a2a_pooled_embs_tensor)
    return _AwaitTensor(All2All_Pooled_Wait, req)


# ------------------------------------------------------------------ reduce-scatter (row-wise)
class ReduceScatter_Req(Function):
    @staticmethod
    def forward(ctx, pg, req: _Request, embs: torch.Tensor) -> torch.Tensor:
        W = dist.get_world_size(pg)
        B_global = embs.shape[0]
        B_local = B_global // W
        send = embs.contiguous()
        out = send.new_empty((B_local,) + tuple(send.shape[1:]))
        if dist.get_backend(pg) == "gloo":
            # gloo has no reduce_scatter: same result through all_reduce + slice (host-logic tests)
            tmp = send.clone()
            req.work = dist.all_reduce(tmp, group=pg, async_op=True)
            req.info = (tmp, dist.get_rank(pg), B_local)
        else:
            req.work = dist.reduce_scatter_tensor(out, send, group=pg, async_op=True)
            req.info = None
        req.tensor = out
        ctx.req, ctx.pg = req, pg
        return torch.zeros(1, device=embs.device)

    @staticmethod
    def backward(ctx, _unused):
        req = ctx.req
        req.work.wait()
        grad = req.tensor
        req.tensor = None
        if GRADIENT_DIVISION:
            grad.div_(dist.get_world_size(ctx.pg))
        return None, None, grad


class ReduceScatter_Wait(Function):
    @staticmethod
    def forward(ctx, pg, req: _Request, dummy: torch.Tensor) -> torch.Tensor:
        req.work.wait()
        ctx.req, ctx.pg = req, pg
        out = req.tensor
        if req.info is not None:
            tmp, rank, B_local = req.info
            out.copy_(tmp[rank * B_local:(rank + 1) * B_local])
            req.info = None
        req.tensor = None
        return out

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        req, pg = ctx.req, ctx.pg
        W = dist.get_world_size(pg)
        g = grad_out.contiguous()
        full = g.new_empty((g.shape[0] * W,) + tuple(g.shape[1:]))
        req.work = dist.all_gather_into_tensor(full, g, group=pg, async_op=True)
        req.tensor = full
        return None, None, req.dummy.new_zeros(1)


def reduce_scatter_pooled(inputs, group: Optional[dist.ProcessGroup] = None) -> Awaitable[torch.Tensor]:
    """Sum over ranks of [W * B_local, D] (or a list of W [B_local, D] chunks), scattered along batch."""
    group = group if group is not None else dist.distributed_c10d._get_default_group()
    embs = torch.cat(inputs, dim=0) if isinstance(inputs, (list, tuple)) else inputs
    if dist.get_world_size(group) <= 1:
        return NoWait(embs)
    req = _Request(group)
    req.dummy = ReduceScatter_Req.apply(group, req, embs)
    return _AwaitTensor(ReduceScatter_Wait, req)


# ------------------------------------------------------------------ sequence all-to-all (non-pooled)
class All2All_Seq(Function):
    """Sequence-embedding all-to-all (reference alltoall_sequence :259, All2All_Seq_Req :608 / _Wait :701).

    Input rows [sum lengths, D] are feature-major ((local feature, source rank, sample) bags, the order the
    KJT all-to-all delivered the ids in); they are regrouped rank-major with ONE permute_2D_sparse_data
    launch over float payloads (lengths x D), sent back to the ranks that own the samples, and arrive as
    (owner rank, feature, sample) = global feature order.  Backward is the transposed exchange followed by
    the inverse regrouping.  No 1/W scaling: a sequence row has exactly one consumer."""

    @staticmethod
    def forward(ctx, pg, embs: torch.Tensor, forward_recat: torch.Tensor, backward_recat: torch.Tensor,
                lengths_after: torch.Tensor, input_splits: List[int], output_splits: List[int]) -> torch.Tensor:
        W = dist.get_world_size(pg)
        D = embs.shape[1]
        local_T = lengths_after.shape[0]
        send = embs.contiguous().view(-1)
        permuted_lengths = None
        if local_T > 0 and forward_recat.numel() > 0:
            lens_d = (lengths_after * D).view(forward_recat.numel(), -1)
            permuted_lengths, send, _ = torch.ops.fbgemm.permute_2D_sparse_data(
                forward_recat, lens_d, send, None, send.numel())
        # rows this rank RECEIVED ids for go back where they came from: splits are swapped
        in_splits = [n * D for n in output_splits]
        out_splits = [n * D for n in input_splits]
        recv = send.new_empty(sum(out_splits))
        dist.all_to_all_single(recv, send, output_split_sizes=out_splits, input_split_sizes=in_splits, group=pg)
        ctx.pg, ctx.D, ctx.permuted_lengths, ctx.backward_recat = pg, D, permuted_lengths, backward_recat
        ctx.in_splits, ctx.out_splits = in_splits, out_splits
        return recv.view(-1, D)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        send = grad_out.contiguous().view(-1)
        recv = send.new_empty(sum(ctx.in_splits))
        dist.all_to_all_single(recv, send, output_split_sizes=ctx.in_splits, input_split_sizes=ctx.out_splits,
                               group=ctx.pg)
        if ctx.permuted_lengths is not None:
            _, recv, _ = torch.ops.fbgemm.permute_2D_sparse_data(
                ctx.backward_recat, ctx.permuted_lengths, recv, None, recv.numel())
        return None, recv.view(-1, ctx.D), None, None, None, None, None


def alltoall_sequence(a2a_sequence_embs_tensor: torch.Tensor, forward_recat_tensor: torch.Tensor,
                      backward_recat_tensor: torch.Tensor, lengths_after_sparse_data_all2all: torch.Tensor,
                      input_splits: List[int], output_splits: List[int],
                      group: Optional[dist.ProcessGroup] = None) -> Awaitable[torch.Tensor]:
    """[T_local * B_global * L, D] rows on every rank -> the rows of MY samples, global feature order."""
    group = group if group is not None else dist.distributed_c10d._get_default_group()
    if dist.get_world_size(group) <= 1:
        return NoWait(a2a_sequence_embs_tensor)
    return NoWait(All2All_Seq.apply(group, a2a_sequence_embs_tensor, forward_recat_tensor, backward_recat_tensor,
                                    lengths_after_sparse_data_all2all, input_splits, output_splits))
