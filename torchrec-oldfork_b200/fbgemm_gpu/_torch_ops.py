"""Registers the torch.ops.fbgemm.* operators torchrec calls, CUDA-only, over the C ABI.

Schemas are the reference's (third_party/fbgemm/fbgemm_gpu/src/sparse_ops_cpu.cpp:1967-2007,
quantize_ops_cpu.cpp, codegen/embedding_bounds_check_host.cpp).  No CPU kernels are
registered: a CPU tensor reaches the dispatcher and raises NotImplementedError — there is
no CPU fallback in the product (the CPU oracle lives in tests/ only).
"""

import torch

from b200_sparse import cabi, ops

_lib = torch.library.Library("fbgemm", "FRAGMENT")
_DEFINED = set()


def _define(schema: str) -> str:
    name = schema.split("(")[0]
    if name not in _DEFINED:
        _lib.define(schema)
        _DEFINED.add(name)
    return name


def _cuda(schema: str):
    def deco(fn):
        _lib.impl(_define(schema), fn, "CUDA")
        return fn
    return deco


@_cuda("asynchronous_complete_cumsum(Tensor t_in) -> Tensor")
def _complete_cumsum(t_in):
    return ops.cumsum(t_in, cabi.CUMSUM_COMPLETE)


@_cuda("asynchronous_inclusive_cumsum(Tensor t_in) -> Tensor")
def _inclusive_cumsum(t_in):
    return ops.cumsum(t_in, cabi.CUMSUM_INCLUSIVE)


@_cuda("asynchronous_exclusive_cumsum(Tensor t_in) -> Tensor")
def _exclusive_cumsum(t_in):
    return ops.cumsum(t_in, cabi.CUMSUM_EXCLUSIVE)


_PERM = "(Tensor permute, Tensor lengths, Tensor values, Tensor? weights=None, int? permuted_lengths_sum=None) -> (Tensor, Tensor, Tensor?)"


@_cuda("permute_2D_sparse_data" + _PERM)
def _permute_2d(permute, lengths, values, weights=None, permuted_lengths_sum=None):
    return ops.permute_sparse_data(permute, lengths, values, weights, permuted_lengths_sum, True)


@_cuda("permute_sparse_data" + _PERM)
def _permute_legacy(permute, lengths, values, weights=None, permuted_lengths_sum=None):
    return ops.permute_sparse_data(permute, lengths, values, weights, permuted_lengths_sum, True)


@_cuda("permute_1D_sparse_data" + _PERM)
def _permute_1d(permute, lengths, values, weights=None, permuted_lengths_sum=None):
    return ops.permute_sparse_data(permute, lengths, values, weights, permuted_lengths_sum, False)


@_cuda("expand_into_jagged_permute(Tensor permute, Tensor input_offset, Tensor output_offset, int output_size) -> Tensor")
def _expand(permute, input_offset, output_offset, output_size):
    return ops.expand_into_jagged_permute(permute, input_offset, output_offset, output_size)


@_cuda("block_bucketize_sparse_features(Tensor lengths, Tensor indices, bool bucketize_pos, bool sequence, Tensor block_sizes, int my_size, Tensor? weights=None) -> (Tensor, Tensor, Tensor?, Tensor?, Tensor?)")
def _bucketize(lengths, indices, bucketize_pos, sequence, block_sizes, my_size, weights=None):
    return ops.block_bucketize_sparse_features(lengths, indices, bucketize_pos, sequence, block_sizes,
                                               my_size, weights)


@_cuda("offsets_range(Tensor offsets, int range_size) -> Tensor")
def _offsets_range(offsets, range_size):
    return ops.offsets_range(offsets, range_size)


@_cuda("bounds_check_indices(Tensor rows_per_table, Tensor(a!) indices, Tensor(b!) offsets, int bounds_check_mode, Tensor(c!) warning) -> ()")
def _bounds_check(rows_per_table, indices, offsets, bounds_check_mode, warning):
    ops.bounds_check_indices(rows_per_table, indices, offsets, bounds_check_mode, warning)


@_cuda("FloatToFusedNBitRowwiseQuantizedSBHalf(Tensor input, int bit_rate) -> Tensor")
def _quantize_nbit(input, bit_rate):
    if bit_rate != 8:
        raise NotImplementedError("only bit_rate 8 (int8 rowwise) is built; int4/int2 are out of scope")
    return ops.quantize_rowwise_8bit_sbhalf(input)
