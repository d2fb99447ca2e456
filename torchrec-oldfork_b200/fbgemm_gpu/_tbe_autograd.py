"""Autograd glue of the fused pooled lookup: the role of SplitLookupFunction_<opt>_Op
(third_party/fbgemm/fbgemm_gpu/codegen/embedding_backward_split_host_template.cpp:137-423).

forward  -> fbgemm::b200_tbe_forward          (pooled gather, tbe_forward.cu)
            fbgemm::b200_tbe_sort_indices     (index transpose, issued on a side stream so it
                                               overlaps the gather; it only depends on the ids)
backward -> fbgemm::b200_tbe_backward         (segment-reduce + optimizer in place, tbe_backward.cu)
            fbgemm::b200_tbe_grad_indice_weights (only when per-sample weights need a gradient)

The four ops are dispatcher ops with CUDA kernels only, so the module works under
torch.ops.fbgemm like the reference's generated invokers.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
from torch import Tensor

from b200_sparse import cabi, ops
from . import _torch_ops  # noqa: F401  (registers the public fbgemm ops)
from .split_embedding_configs import EmbOptimType, SparseType

_lib = torch.library.Library("fbgemm", "FRAGMENT")

_lib.define(
    "b200_tbe_forward(Tensor dev_weights, Tensor weights_offsets, Tensor D_offsets, int total_D, int max_D, "
    "bool all_D_mult8, Tensor indices, Tensor offsets, int pooling_mode, Tensor? indice_weights, "
    "int output_dtype) -> Tensor")
_lib.define(
    "b200_tbe_sort_indices(Tensor indices, Tensor offsets, Tensor hash_size_cumsum, int total_hash_size_bits, "
    "int B, bool weighted, int max_table_bits=0) -> (Tensor, Tensor, Tensor)")
_lib.define(
    "b200_tbe_backward(Tensor grad_output, Tensor sorted_keys, Tensor sorted_payload, Tensor bag_of_pos, "
    "int total_hash_size_bits, Tensor(a!) dev_weights, Tensor weights_offsets, Tensor D_offsets, "
    "Tensor hash_size_cumsum, int B, int max_D, bool all_D_mult8, Tensor offsets, Tensor? indice_weights, "
    "int pooling_mode, int optimizer, Tensor(b!) momentum1, Tensor momentum1_offsets, float learning_rate, "
    "float eps, bool stochastic_rounding=False, int sr_seed=0, int sr_offset=0) -> ()")
_lib.define(
    "b200_tbe_grad_indice_weights(Tensor grad_output, Tensor dev_weights, Tensor weights_offsets, "
    "Tensor D_offsets, Tensor indices, Tensor offsets) -> Tensor")

_lib.define("b200_linearize_seq_indices(Tensor indices, Tensor offsets, Tensor hash_size_cumsum, int B) -> Tensor")

_OUT_DT = {SparseType.FP32.as_int(): torch.float32, SparseType.FP16.as_int(): torch.float16}


def _fwd_cuda(dev_weights, weights_offsets, D_offsets, total_D, max_D, all_D_mult8, indices, offsets,
              pooling_mode, indice_weights, output_dtype):
    return ops.tbe_forward(dev_weights, weights_offsets, D_offsets, total_D, max_D, all_D_mult8, indices,
                           offsets, indice_weights, pooling_mode, _OUT_DT[output_dtype])


def _sort_cuda(indices, offsets, hash_size_cumsum, total_hash_size_bits, B, weighted, max_table_bits=0):
    s = ops.tbe_sort_indices(indices, offsets, hash_size_cumsum, total_hash_size_bits, B, weighted,
                             max_table_bits)
    bag = s.bag_of_pos if s.bag_of_pos is not None else s.payload.new_empty(0)
    return s.keys, s.payload, bag


def _bwd_cuda(grad_output, sorted_keys, sorted_payload, bag_of_pos, total_hash_size_bits, dev_weights,
              weights_offsets, D_offsets, hash_size_cumsum, B, max_D, all_D_mult8, offsets, indice_weights,
              pooling_mode, optimizer, momentum1, momentum1_offsets, learning_rate, eps,
              stochastic_rounding=False, sr_seed=0, sr_offset=0):
    srt = ops.SortedIndices(sorted_keys, sorted_payload, bag_of_pos if bag_of_pos.numel() else None,
                            sorted_keys.numel(), max(1, total_hash_size_bits), None)
    ops.tbe_backward(grad_output, srt, dev_weights, weights_offsets, D_offsets, hash_size_cumsum, B, max_D,
                     all_D_mult8, offsets, indice_weights, pooling_mode, optimizer,
                     momentum1 if momentum1.numel() else None, momentum1_offsets, learning_rate, eps,
                     stochastic_rounding, sr_seed, sr_offset)


def _giw_cuda(grad_output, dev_weights, weights_offsets, D_offsets, indices, offsets):
    return ops.tbe_grad_indice_weights(grad_output, dev_weights, weights_offsets, D_offsets, indices, offsets)


_lib.impl("b200_linearize_seq_indices", lambda i, o, h, B: ops.linearize_seq_indices(i, o, h, B), "CUDA")
_lib.impl("b200_tbe_forward", _fwd_cuda, "CUDA")
_lib.impl("b200_tbe_sort_indices", _sort_cuda, "CUDA")
_lib.impl("b200_tbe_backward", _bwd_cuda, "CUDA")
_lib.impl("b200_tbe_grad_indice_weights", _giw_cuda, "CUDA")

_OPT_CODE = {EmbOptimType.EXACT_SGD: cabi.OPT_SGD, EmbOptimType.EXACT_ROWWISE_ADAGRAD: cabi.OPT_ROWWISE_ADAGRAD}
_SIDE_STREAMS: Dict[int, "torch.cuda.Stream"] = {}


def _side_stream(dev: torch.device) -> "torch.cuda.Stream":
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    s = _SIDE_STREAMS.get(idx)
    if s is None:
        s = torch.cuda.Stream(device=dev, priority=-1)  # small kernels: let them slip in between big ones
        _SIDE_STREAMS[idx] = s
    return s


class _Lookup(torch.autograd.Function):
    @staticmethod
    def forward(ctx, placeholder, module, indices, offsets, indice_weights, feature_requires_grad):
        m = module
        T = m.D_offsets.numel() - 1
        B = (offsets.numel() - 1) // T
        need_grad = ctx.needs_input_grad[0]  # placeholder requires grad and grad mode was on at apply()
        weighted = indice_weights is not None
        sort_out = None
        ev = None
        if need_grad and indices.numel() > 0:
            if indices.is_cuda:
                main = torch.cuda.current_stream(indices.device)
                side = _side_stream(indices.device)
                side.wait_stream(main)
                with torch.cuda.stream(side):
                    sort_out = torch.ops.fbgemm.b200_tbe_sort_indices(
                        indices, offsets, m.hash_size_cumsum, m.total_hash_size_bits, B, weighted,
                        getattr(m, "max_table_bits", 0))
                    ev = torch.cuda.Event()
                    ev.record(side)
                for t in (indices, offsets):
                    t.record_stream(side)
                for t in sort_out:
                    t.record_stream(main)
            else:
                sort_out = torch.ops.fbgemm.b200_tbe_sort_indices(
                    indices, offsets, m.hash_size_cumsum, m.total_hash_size_bits, B, weighted,
                        getattr(m, "max_table_bits", 0))
        out = torch.ops.fbgemm.b200_tbe_forward(
            m.weights_dev, m.weights_offsets, m.D_offsets, m.total_D, m.max_D, m._all_D_mult8, indices, offsets,
            int(m.pooling_mode), indice_weights, m.output_dtype)
        ctx.module = m
        ctx.B = B
        ctx.sort_out = sort_out
        ctx.sort_event = ev
        ctx.weights_need_grad = weighted and indice_weights.requires_grad
        ctx.save_for_backward(indices, offsets, indice_weights if weighted else None, feature_requires_grad)
        return out

    @staticmethod
    def backward(ctx, grad_output):
        m = ctx.module
        indices, offsets, indice_weights, feature_requires_grad = ctx.saved_tensors
        grad_iw = None
        if ctx.sort_out is not None:
            a = m.optimizer_args
            if a.gradient_clipping:
                grad_output = grad_output.clamp(-a.max_gradient, a.max_gradient)
            if ctx.weights_need_grad:
                # uses the weights BEFORE the update, like the reference (host_template.cpp:300-330)
                grad_iw = torch.ops.fbgemm.b200_tbe_grad_indice_weights(
                    grad_output, m.weights_dev, m.weights_offsets, m.D_offsets, indices, offsets)
                if feature_requires_grad is not None:
                    T = m.D_offsets.numel() - 1
                    lengths = offsets[1:] - offsets[:-1]
                    mask = torch.repeat_interleave(
                        feature_requires_grad.to(grad_iw.dtype).repeat_interleave(ctx.B)[: T * ctx.B], lengths)
                    grad_iw = grad_iw * mask
            if ctx.sort_event is not None:
                torch.cuda.current_stream(grad_output.device).wait_event(ctx.sort_event)
            keys, payload, bag = ctx.sort_out
            torch.ops.fbgemm.b200_tbe_backward(
                grad_output, keys, payload, bag, m.total_hash_size_bits, m.weights_dev, m.weights_offsets,
                m.D_offsets, m.hash_size_cumsum, ctx.B, m.max_D, m._all_D_mult8, offsets, indice_weights,
                int(m.pooling_mode), _OPT_CODE[m.optimizer], m.momentum1_dev, m.momentum1_offsets,
                a.learning_rate, a.eps, *m._next_sr_state())
            ctx.sort_out = None
        return None, None, None, None, grad_iw, None


def lookup(module, indices: Tensor, offsets: Tensor, per_sample_weights: Optional[Tensor],
           feature_requires_grad: Optional[Tensor]) -> Tensor:
    return _Lookup.apply(module.placeholder_autograd_tensor, module, indices, offsets, per_sample_weights,
                         feature_requires_grad)


_lib.define(
    "b200_tbe_int8_forward(Tensor dev_weights, Tensor weights_offsets, Tensor D_offsets, int total_D, int max_D, "
    "int row_alignment, Tensor indices, Tensor offsets, Tensor? indice_weights, int pooling_mode, "
    "int output_dtype) -> Tensor")


def _int8_fwd_cuda(dev_weights, weights_offsets, D_offsets, total_D, max_D, row_alignment, indices, offsets,
                   indice_weights, pooling_mode, output_dtype):
    return ops.tbe_int8_forward(dev_weights, weights_offsets, D_offsets, total_D, max_D, row_alignment, indices,
                                offsets, indice_weights, pooling_mode, _OUT_DT[output_dtype])


_lib.impl("b200_tbe_int8_forward", _int8_fwd_cuda, "CUDA")
