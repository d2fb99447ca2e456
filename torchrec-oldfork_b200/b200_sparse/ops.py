"""Tensor-level wrappers over the C ABI: argument checks, allocation, stream lookup.

Every function requires CUDA tensors and launches on torch's *current* stream (the reference
ops use at::cuda::getCurrentCUDAStream(), embedding_forward_split_template.cu:400), never
the legacy default stream, and never synchronises the device.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from . import cabi
from .cabi import check, lib

_DT = {torch.int32: cabi.I32, torch.int64: cabi.I64, torch.float32: cabi.F32,
       torch.float16: cabi.F16, torch.bfloat16: cabi.BF16, torch.uint8: cabi.U8}

_WS: Dict[Tuple[int, int], torch.Tensor] = {}


class KernelProfile:
    """Launch accounting and optional CUDA-event timing of the hot kernels (used by bench.py:
    events are recorded on the launching stream around the C-ABI call, so the elapsed time is
    the kernel's own duration)."""
    launches = 0
    enabled = False
    events: Dict[str, List[Tuple["torch.cuda.Event", "torch.cuda.Event"]]] = {}
    last_sort_args: Optional[tuple] = None  # arguments of the latest index transpose (bench.py re-times it alone)

    @classmethod
    def reset(cls, enabled: bool = False) -> None:
        cls.launches, cls.enabled, cls.events = 0, enabled, {}

    @classmethod
    def summary_ms(cls) -> Dict[str, Tuple[int, float]]:
        """name -> (count, mean milliseconds); call after a device synchronize."""
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v) / max(1, len(v))) for k, v in cls.events.items()}


class _timed:
    def __init__(self, name: str, dev: torch.device, launches: int) -> None:
        self.name, self.dev, self.n = name, dev, launches

    def __enter__(self):
        KernelProfile.launches += self.n
        if KernelProfile.enabled:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record(torch.cuda.current_stream(self.dev))
        return self

    def __exit__(self, *exc):
        if KernelProfile.enabled:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record(torch.cuda.current_stream(self.dev))
            KernelProfile.events.setdefault(self.name, []).append((self.e0, e1))
        return False


def _need_cuda(*ts: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in ts:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError(
                "b200_sparse ops have a CUDA (sm_100a) implementation only; got a tensor on "
                f"{t.device}. There is no CPU fallback by design."
            )
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise RuntimeError(f"tensors on different devices: {dev} vs {t.device}")
    assert dev is not None
    return dev


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_get_dev = getattr(torch._C, "_cuda_getDevice", None)
_set_dev = getattr(torch._C, "_cuda_setDevice", None)


class _on:
    """`with _on(dev)`: makes `dev` the current CUDA device for the C-ABI call; when it already is (the usual
    case: one process per GPU) this is two attribute reads instead of torch.cuda.device's full guard - the
    training step makes ~30 of these calls and is bound by host time at 8 ranks."""
    __slots__ = ("idx", "prev")

    def __init__(self, dev: torch.device) -> None:
        self.idx = dev.index
        self.prev = -1

    def __enter__(self):
        if self.idx is not None and _get_dev is not None:
            cur = _get_dev()
            if cur != self.idx:
                _set_dev(self.idx)
                self.prev = cur
        return self

    def __exit__(self, *exc):
        if self.prev >= 0:
            _set_dev(self.prev)
        return False


def _stream(dev: torch.device) -> int:
    """cudaStream_t of torch's current stream on `dev` (every launch goes there)."""
    if _raw_stream is not None:
        return _raw_stream(dev.index if dev.index is not None else torch.cuda.current_device())
    return torch.cuda.current_stream(dev).cuda_stream


def _ws(dev: torch.device) -> torch.Tensor:
    """Scan scratch, one per (device, stream) so ops on side streams never share it."""
    key = (dev.index if dev.index is not None else torch.cuda.current_device(), _stream(dev))
    w = _WS.get(key)
    if w is None:
        w = torch.zeros(lib.b200_scan_workspace_bytes() // 8, dtype=torch.int64, device=dev)
        _WS[key] = w
    return w


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _idx_dt(t: torch.Tensor, what: str) -> int:
    if t.dtype not in (torch.int32, torch.int64):
        raise TypeError(f"{what} must be int32 or int64, got {t.dtype}")
    return _DT[t.dtype]


# ------------------------------------------------------------------------ jagged ops
def cumsum(t_in: torch.Tensor, mode: int) -> torch.Tensor:
    dev = _need_cuda(t_in)
    if t_in.dim() != 1:
        # reference: TORCH_CHECK(t_in.dim() == 1) for complete; others flatten
        if mode == cabi.CUMSUM_COMPLETE:
            raise RuntimeError("asynchronous_complete_cumsum expects a 1-D tensor")
    x = t_in.contiguous()
    n = x.numel()
    if n >= 2 ** 31:
        raise RuntimeError("cumsum: numel must be < 2^31")
    dt = _idx_dt(x, "cumsum input")
    shape = (n + 1,) if mode == cabi.CUMSUM_COMPLETE else x.shape
    out = torch.empty(shape, dtype=x.dtype, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_cumsum(_p(x), _p(out), n, dt, mode, _p(_ws(dev)), _stream(dev)), "b200_cumsum")
    return out


def permute_sparse_data(
    permute: torch.Tensor,
    lengths: torch.Tensor,
    values: torch.Tensor,
    weights: Optional[torch.Tensor],
    permuted_lengths_sum: Optional[int],
    two_d: bool,
) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
    dev = _need_cuda(permute, lengths, values, weights)
    if permute.dtype != torch.int32:
        raise TypeError("permute must be int32")
    permute = permute.contiguous()
    lengths_c = lengths.contiguous()
    values_c = values.contiguous()
    weights_c = weights.contiguous() if weights is not None else None
    if two_d:
        if lengths_c.dim() != 2:
            raise RuntimeError("permute_2D_sparse_data: lengths must be [T, B]")
        T, B = lengths_c.shape
    else:
        T, B = lengths_c.numel(), 1
    Tp = permute.numel()
    ldt = _idx_dt(lengths_c, "lengths")
    out_shape = (Tp, B) if two_d else (Tp,)
    out_lengths = torch.empty(out_shape, dtype=lengths_c.dtype, device=dev)
    if Tp == 0 or B == 0:
        return out_lengths, values_c.new_empty(0), (None if weights_c is None else weights_c.new_empty(0))
    if two_d and permuted_lengths_sum is not None and T <= 1024 and Tp <= 1024:
        # the hot callers (KJT.permute, the input-dist recat) know the total: one fused launch
        total = int(permuted_lengths_sum)
        out_values = torch.empty(total, dtype=values_c.dtype, device=dev)
        out_weights = torch.empty(total, dtype=weights_c.dtype, device=dev) if weights_c is not None else None
        with _on(dev):
            KernelProfile.launches += 1
            check(lib.b200_permute_2d_fused(_p(permute), _p(lengths_c), ldt, T, Tp, B, _p(out_lengths),
                                            _p(values_c), _p(out_values), values_c.element_size(),
                                            _p(weights_c), _p(out_weights),
                                            weights_c.element_size() if weights_c is not None else 4,
                                            total, _p(_ws(dev)), _stream(dev)), "b200_permute_2d_fused")
        return out_lengths, out_values, out_weights
    in_off = torch.empty(T * B + 1, dtype=torch.int64, device=dev)
    out_off = torch.empty(Tp * B + 1, dtype=torch.int64, device=dev)
    with _on(dev):
        st = _stream(dev)
        KernelProfile.launches += 1
        check(lib.b200_permute_lengths(_p(permute), _p(lengths_c), ldt, T, Tp, B, _p(out_lengths),
                                       _p(in_off), _p(out_off), _p(_ws(dev)), st),
              "b200_permute_lengths")
        if permuted_lengths_sum is None:
            # same host sync as the reference when the caller does not pass the total
            # (sparse_ops.cu:412)
            permuted_lengths_sum = int(out_off[-1].item())
        total = int(permuted_lengths_sum)
        out_values = torch.empty(total, dtype=values_c.dtype, device=dev)
        out_weights = None
        if weights_c is not None:
            out_weights = torch.empty(total, dtype=weights_c.dtype, device=dev)
        KernelProfile.launches += 1
        check(lib.b200_permute_data(_p(permute), Tp, B, _p(in_off), _p(out_off), _p(values_c),
                                    _p(out_values), values_c.element_size(), _p(weights_c),
                                    _p(out_weights),
                                    weights_c.element_size() if weights_c is not None else 4,
                                    total, st),
              "b200_permute_data")
    return out_lengths, out_values, out_weights


def expand_into_jagged_permute(permute: torch.Tensor, input_offsets: torch.Tensor,
                               output_offsets: torch.Tensor, output_size: int) -> torch.Tensor:
    dev = _need_cuda(permute, input_offsets, output_offsets)
    if permute.dtype != torch.int32:
        permute = permute.to(torch.int32)
    if input_offsets.dtype != output_offsets.dtype:
        raise TypeError("input_offsets and output_offsets must share a dtype")
    n = permute.numel()
    if input_offsets.numel() != n + 1 or output_offsets.numel() != n + 1:
        raise RuntimeError("offsets must have permute.numel() + 1 entries")
    dt = _idx_dt(input_offsets, "offsets")
    out = torch.empty(int(output_size), dtype=input_offsets.dtype, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_expand_into_jagged_permute(_p(permute.contiguous()), _p(input_offsets.contiguous()),
                                                  _p(output_offsets.contiguous()), dt, n, _p(out),
                                                  int(output_size), _stream(dev)),
              "b200_expand_into_jagged_permute")
    return out


def offsets_range(offsets: torch.Tensor, range_size: int) -> torch.Tensor:
    dev = _need_cuda(offsets)
    dt = _idx_dt(offsets, "offsets")
    out = torch.empty(int(range_size), dtype=offsets.dtype, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_offsets_range(_p(offsets.contiguous()), dt, offsets.numel(), _p(out),
                                     int(range_size), _stream(dev)), "b200_offsets_range")
    return out


def block_bucketize_sparse_features(
    lengths: torch.Tensor, indices: torch.Tensor, bucketize_pos: bool, sequence: bool,
    block_sizes: torch.Tensor, my_size: int, weights: Optional[torch.Tensor] = None,
) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor], Optional[torch.Tensor], Optional[torch.Tensor]]:
    dev = _need_cuda(lengths, indices, block_sizes, weights)
    lengths = lengths.contiguous()
    indices = indices.contiguous()
    block_sizes = block_sizes.contiguous()
    if block_sizes.dtype != indices.dtype:
        raise TypeError("block_sizes must have the dtype of indices (sparse_ops.cu:855)")
    if weights is not None and weights.dtype != torch.float32:
        raise TypeError("bucketize weights must be float32")
    odt, idt = _idx_dt(lengths, "lengths"), _idx_dt(indices, "indices")
    T = block_sizes.numel()
    TB = lengths.numel()
    if T == 0 or TB % T != 0:
        raise RuntimeError("lengths.numel() must be a multiple of block_sizes.numel()")
    B = TB // T
    N = indices.numel()
    W = int(my_size)
    new_lengths = torch.empty(TB * W, dtype=lengths.dtype, device=dev)
    new_indices = torch.empty_like(indices)
    new_weights = torch.empty_like(weights) if weights is not None else None
    new_pos = torch.empty_like(indices) if bucketize_pos else None
    unbucketize = torch.empty_like(indices) if sequence else None
    if TB == 0:
        return new_lengths, new_indices, new_weights, new_pos, unbucketize
    off = torch.empty(TB + 1, dtype=torch.int64, device=dev)
    noff = torch.empty(TB * W + 1, dtype=torch.int64, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_block_bucketize(_p(lengths), odt, _p(indices), idt, _p(block_sizes), T, B, W, N,
                                       _p(weights.contiguous() if weights is not None else None),
                                       _p(new_lengths), _p(new_indices), _p(new_weights), _p(new_pos),
                                       _p(unbucketize), _p(off), _p(noff), _p(_ws(dev)), _stream(dev)),
              "b200_block_bucketize")
    return new_lengths, new_indices, new_weights, new_pos, unbucketize


def bounds_check_indices(rows_per_table: torch.Tensor, indices: torch.Tensor, offsets: torch.Tensor,
                         mode: int, warning: torch.Tensor) -> None:
    dev = _need_cuda(rows_per_table, indices, offsets, warning)
    if indices.dtype != offsets.dtype:
        raise TypeError("indices and offsets must share a dtype")
    if not (indices.is_contiguous() and offsets.is_contiguous()):
        raise RuntimeError("bounds_check_indices works in place and needs contiguous tensors")
    dt = _idx_dt(indices, "indices")
    T = rows_per_table.numel()
    if T == 0:
        return
    B = (offsets.numel() - 1) // T
    with _on(dev):
        KernelProfile.launches += 2
        check(lib.b200_bounds_check(_p(rows_per_table), _p(indices), _p(offsets), dt, T, B,
                                    indices.numel(), int(mode), _p(warning), _p(_ws(dev)), _stream(dev)),
              "b200_bounds_check")


# ------------------------------------------------------------------------ pooled lookup
def tbe_forward(weights: torch.Tensor, weights_offsets: torch.Tensor, D_offsets: torch.Tensor,
                total_D: int, max_D: int, all_D_mult8: bool, indices: torch.Tensor,
                offsets: torch.Tensor, per_sample_weights: Optional[torch.Tensor], pooling_mode: int,
                output_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    dev = _need_cuda(weights, weights_offsets, D_offsets, indices, offsets, per_sample_weights)
    if indices.dtype != offsets.dtype:
        raise TypeError("indices and offsets must share a dtype")
    idt = _idx_dt(indices, "indices")
    T = D_offsets.numel() - 1
    if T <= 0:
        raise RuntimeError("TBE forward needs at least one table")
    if (offsets.numel() - 1) % T != 0:
        raise RuntimeError("offsets must have T*B+1 entries")
    B = (offsets.numel() - 1) // T
    if per_sample_weights is not None:
        if per_sample_weights.dtype != torch.float32:
            per_sample_weights = per_sample_weights.float()
        per_sample_weights = per_sample_weights.contiguous()
    out = torch.empty((B, total_D), dtype=output_dtype, device=dev)
    with _on(dev), _timed("tbe_forward", dev, 1):
        check(lib.b200_tbe_forward(_p(weights), _DT[weights.dtype], _p(weights_offsets), _p(D_offsets),
                                   T, B, total_D, max_D, int(all_D_mult8), _p(indices.contiguous()),
                                   _p(offsets.contiguous()), idt, _p(per_sample_weights),
                                   int(pooling_mode), _p(out), _DT[output_dtype], _stream(dev)),
              "b200_tbe_forward")
    return out


class SortedIndices:
    """Result of the index transpose: sorted (table,row) keys with their bag payload."""

    __slots__ = ("keys", "payload", "bag_of_pos", "N", "key_bits", "ws")

    def __init__(self, keys, payload, bag_of_pos, N, key_bits, ws):
        self.keys, self.payload, self.bag_of_pos = keys, payload, bag_of_pos
        self.N, self.key_bits, self.ws = N, key_bits, ws


def tbe_sort_indices(indices: torch.Tensor, offsets: torch.Tensor, hash_size_cumsum: torch.Tensor,
                     total_hash_size_bits: int, B: int, weighted: bool, max_table_bits: int = 0) -> SortedIndices:
    """max_table_bits: ceil(log2(rows)) of the largest table (0 = unknown); only for features grouped by table."""
    dev = _need_cuda(indices, offsets, hash_size_cumsum)
    idt = _idx_dt(indices, "indices")
    T = hash_size_cumsum.numel() - 1
    N = indices.numel()
    key_bits = max(1, int(total_hash_size_bits))
    kdt = torch.int64 if key_bits > 32 else torch.int32
    keys = torch.empty(N, dtype=kdt, device=dev)
    payload = torch.empty(N, dtype=torch.int32, device=dev)
    bag_of_pos = torch.empty(N, dtype=torch.int32, device=dev) if weighted else None
    ws_bytes = lib.b200_tbe_sort_workspace_bytes(N, key_bits, T)
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
    host_bits = max_table_bits if 0 < max_table_bits < key_bits else key_bits
    n_pass = (host_bits + 8) // 9
    if KernelProfile.enabled:
        KernelProfile.last_sort_args = (indices, offsets, hash_size_cumsum, total_hash_size_bits, B, weighted,
                                        max_table_bits)
    with _on(dev), _timed("tbe_sort", dev, 2 + 3 * n_pass):  # plan, linearize, 3 kernels per pass
        check(lib.b200_tbe_sort_indices(_p(indices), _p(offsets), idt, _p(hash_size_cumsum), T, B, N,
                                        key_bits, int(max_table_bits), int(weighted), _p(keys), _p(payload), _p(bag_of_pos),
                                        _p(ws), ws_bytes, _stream(dev)), "b200_tbe_sort_indices")
    return SortedIndices(keys, payload, bag_of_pos, N, key_bits, ws)


def tbe_backward(grad_output: torch.Tensor, srt: SortedIndices, weights: torch.Tensor,
                 weights_offsets: torch.Tensor, D_offsets: torch.Tensor, hash_size_cumsum: torch.Tensor,
                 B: int, max_D: int, all_D_mult8: bool, offsets: torch.Tensor,
                 per_sample_weights: Optional[torch.Tensor], pooling_mode: int, optimizer: int,
                 momentum1: Optional[torch.Tensor], momentum1_offsets: Optional[torch.Tensor],
                 learning_rate: float, eps: float, stochastic_rounding: bool = False, sr_seed: int = 0,
                 sr_offset: int = 0) -> None:
    """stochastic_rounding: fp16 / bf16 rows are rounded stochastically with the Philox stream
    (sr_seed, sr_offset); the caller moves sr_offset every call.  fp32 tables ignore it."""
    dev = _need_cuda(grad_output, weights, weights_offsets, D_offsets, hash_size_cumsum, offsets)
    if grad_output.dtype != torch.float32:
        grad_output = grad_output.float()
    # the reference forces a contiguous, 16-byte aligned gradient
    # (embedding_backward_split_host_template.cpp:274-278)
    if not grad_output.is_contiguous() or grad_output.data_ptr() % 16 != 0 or grad_output.stride(0) % 4 != 0:
        grad_output = grad_output.contiguous().clone(memory_format=torch.contiguous_format)
    T = D_offsets.numel() - 1
    scratch_bytes = lib.b200_tbe_backward_scratch_bytes(srt.N, max_D)
    scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device=dev)
    if per_sample_weights is not None:
        per_sample_weights = per_sample_weights.float().contiguous()
    if stochastic_rounding and weights.dtype != torch.float32:
        with _on(dev), _timed("tbe_backward", dev, 2):
            check(lib.b200_tbe_backward_sr(_p(grad_output), grad_output.stride(0), _p(srt.keys), _p(srt.payload),
                                           _p(srt.bag_of_pos), srt.N, srt.key_bits, _p(weights),
                                           _DT[weights.dtype], _p(weights_offsets), _p(D_offsets),
                                           _p(hash_size_cumsum), T, B, max_D, int(all_D_mult8), _p(offsets),
                                           _idx_dt(offsets, "offsets"), _p(per_sample_weights),
                                           int(pooling_mode), int(optimizer), _p(momentum1),
                                           _p(momentum1_offsets), float(learning_rate), float(eps),
                                           _p(scratch), int(sr_seed) & (2 ** 64 - 1),
                                           int(sr_offset) & (2 ** 64 - 1), _stream(dev)), "b200_tbe_backward_sr")
        return
    with _on(dev), _timed("tbe_backward", dev, 2):  # tile kernel + cross-CTA combine
        check(lib.b200_tbe_backward(_p(grad_output), grad_output.stride(0), _p(srt.keys), _p(srt.payload),
                                    _p(srt.bag_of_pos), srt.N, srt.key_bits, _p(weights),
                                    _DT[weights.dtype], _p(weights_offsets), _p(D_offsets),
                                    _p(hash_size_cumsum), T, B, max_D, int(all_D_mult8), _p(offsets),
                                    _idx_dt(offsets, "offsets"), _p(per_sample_weights),
                                    int(pooling_mode), int(optimizer), _p(momentum1),
                                    _p(momentum1_offsets), float(learning_rate), float(eps),
                                    _p(scratch), _stream(dev)), "b200_tbe_backward")


def tbe_grad_indice_weights(grad_output: torch.Tensor, weights: torch.Tensor,
                            weights_offsets: torch.Tensor, D_offsets: torch.Tensor,
                            indices: torch.Tensor, offsets: torch.Tensor) -> torch.Tensor:
    dev = _need_cuda(grad_output, weights, indices, offsets)
    grad_output = grad_output.float().contiguous()
    T = D_offsets.numel() - 1
    B = (offsets.numel() - 1) // T
    out = torch.empty(indices.numel(), dtype=torch.float32, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_tbe_grad_indice_weights(_p(grad_output), grad_output.stride(0), _p(weights),
                                               _DT[weights.dtype], _p(weights_offsets), _p(D_offsets),
                                               T, B, _p(indices), _p(offsets),
                                               _idx_dt(indices, "indices"), _p(out), _stream(dev)),
              "b200_tbe_grad_indice_weights")
    return out


# ------------------------------------------------------------------------ int8 inference
def quantize_rowwise_8bit_sbhalf(w: torch.Tensor) -> torch.Tensor:
    dev = _need_cuda(w)
    if w.dim() != 2:
        raise RuntimeError("expected a 2-D weight")
    w = w.float().contiguous()
    out = torch.empty((w.shape[0], w.shape[1] + 4), dtype=torch.uint8, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_quantize_rowwise_8bit_sbhalf(_p(w), w.shape[0], w.shape[1], _p(out), _stream(dev)),
              "b200_quantize_rowwise_8bit_sbhalf")
    return out


def tbe_int8_forward(weights: torch.Tensor, weights_offsets: torch.Tensor, D_offsets: torch.Tensor,
                     total_D: int, max_D: int, row_alignment: int, indices: torch.Tensor,
                     offsets: torch.Tensor, per_sample_weights: Optional[torch.Tensor],
                     pooling_mode: int, output_dtype: torch.dtype) -> torch.Tensor:
    dev = _need_cuda(weights, weights_offsets, D_offsets, indices, offsets, per_sample_weights)
    if indices.dtype != offsets.dtype:
        raise TypeError("indices and offsets must share a dtype")
    T = D_offsets.numel() - 1
    B = (offsets.numel() - 1) // T
    out = torch.empty((B, total_D), dtype=output_dtype, device=dev)
    if per_sample_weights is not None:
        per_sample_weights = per_sample_weights.float().contiguous()
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_tbe_int8_forward(_p(weights), _p(weights_offsets), _p(D_offsets), T, B, total_D,
                                        max_D, int(row_alignment), _p(indices.contiguous()),
                                        _p(offsets.contiguous()), _idx_dt(indices, "indices"),
                                        _p(per_sample_weights), int(pooling_mode), _p(out),
                                        _DT[output_dtype], _stream(dev)), "b200_tbe_int8_forward")
    return out


# ------------------------------------------------------------------------ fused lookup + exchange
def tbe_forward_peer(weights: torch.Tensor, weights_offsets: torch.Tensor, D_offsets: torch.Tensor,
                     total_D: int, max_D: int, all_D_mult8: bool, indices: torch.Tensor, offsets: torch.Tensor,
                     per_sample_weights: Optional[torch.Tensor], pooling_mode: int, dst_bases: torch.Tensor,
                     B_dst: int, dst_pitch: int, dst_col: torch.Tensor, skip_empty: bool = False,
                     flag_bases: Optional[torch.Tensor] = None, flag_pitch: int = 0,
                     flag_col: Optional[torch.Tensor] = None) -> None:
    """Pooled lookup whose rows go to dst_bases[b // B_dst] (int64 device array of fp32 buffer
    addresses, usually NVLink peer mappings) at row b % B_dst, column dst_col[t].  skip_empty: empty
    bags store nothing.  flag_bases (int64 addresses of uint8 buffers) / flag_pitch / flag_col (int32 [T]):
    non-empty bags also set a presence byte for peer_stage_reduce_flagged."""
    dev = _need_cuda(weights, weights_offsets, D_offsets, indices, offsets, per_sample_weights, dst_bases, dst_col)
    if indices.dtype != offsets.dtype:
        raise TypeError("indices and offsets must share a dtype")
    if dst_bases.dtype != torch.int64 or dst_col.dtype != torch.int32:
        raise TypeError("dst_bases must be int64 addresses and dst_col int32 columns")
    idt = _idx_dt(indices, "indices")
    T = D_offsets.numel() - 1
    B = (offsets.numel() - 1) // T
    if per_sample_weights is not None:
        per_sample_weights = per_sample_weights.float().contiguous()
    with _on(dev), _timed("tbe_forward", dev, 1):
        check(lib.b200_tbe_forward_peer(_p(weights), _DT[weights.dtype], _p(weights_offsets), _p(D_offsets), T, B,
                                        total_D, max_D, int(all_D_mult8), _p(indices.contiguous()),
                                        _p(offsets.contiguous()), idt, _p(per_sample_weights), int(pooling_mode),
                                        _p(dst_bases), dst_bases.numel(), int(B_dst), int(dst_pitch), _p(dst_col),
                                        _DT[torch.float32], int(bool(skip_empty)), _p(flag_bases),
                                        int(flag_pitch), _p(flag_col), _stream(dev)),
              "b200_tbe_forward_peer")


def tbe_backward_peer(grad_bases: torch.Tensor, B_src: int, grad_stride: int, src_col: torch.Tensor,
                      grad_scale: float, srt: SortedIndices, weights: torch.Tensor, weights_offsets: torch.Tensor,
                      D_offsets: torch.Tensor, hash_size_cumsum: torch.Tensor, B: int, max_D: int,
                      all_D_mult8: bool, offsets: torch.Tensor, per_sample_weights: Optional[torch.Tensor],
                      pooling_mode: int, optimizer: int, momentum1: Optional[torch.Tensor],
                      momentum1_offsets: Optional[torch.Tensor], learning_rate: float, eps: float,
                      stochastic_rounding: bool = False, sr_seed: int = 0, sr_offset: int = 0) -> None:
    """Fused backward + optimizer reading the gradient row of sample b from grad_bases[b // B_src].
    per_sample_weights needs a weighted sort (payload = position, srt.bag_of_pos); stochastic_rounding as in
    tbe_backward."""
    dev = _need_cuda(grad_bases, src_col, weights, weights_offsets, D_offsets, hash_size_cumsum, offsets)
    if grad_bases.dtype != torch.int64 or src_col.dtype != torch.int32:
        raise TypeError("grad_bases must be int64 addresses and src_col int32 columns")
    T = D_offsets.numel() - 1
    scratch_bytes = lib.b200_tbe_backward_scratch_bytes(srt.N, max_D)
    scratch = torch.empty(scratch_bytes, dtype=torch.uint8, device=dev)
    if per_sample_weights is not None:
        per_sample_weights = per_sample_weights.float().contiguous()
    with _on(dev), _timed("tbe_backward", dev, 2):
        check(lib.b200_tbe_backward_peer(_p(grad_bases), grad_bases.numel(), int(B_src), int(grad_stride),
                                         _p(src_col), float(grad_scale), _p(srt.keys), _p(srt.payload),
                                         _p(srt.bag_of_pos), srt.N, srt.key_bits, _p(weights), _DT[weights.dtype],
                                         _p(weights_offsets), _p(D_offsets), _p(hash_size_cumsum), T, B, max_D,
                                         int(all_D_mult8), _p(offsets), _idx_dt(offsets, "offsets"),
                                         _p(per_sample_weights), int(pooling_mode), int(optimizer), _p(momentum1),
                                         _p(momentum1_offsets), float(learning_rate), float(eps), _p(scratch),
                                         int(bool(stochastic_rounding)), int(sr_seed) & (2 ** 64 - 1),
                                         int(sr_offset) & (2 ** 64 - 1), _stream(dev)), "b200_tbe_backward_peer")


def peer_stage_reduce(stage: torch.Tensor, W: int, B_local: int, Dw: int, out: torch.Tensor,
                      out_col4: torch.Tensor, clear_stage: bool = False) -> None:
    """Sum over the W rank slots of stage ([W, B_local, Dw] fp32); float4 column c of the stage goes to
    output column out_col4[c] (int32 device tensor [Dw // 4]).  clear_stage: zero what was read."""
    dev = _need_cuda(stage, out, out_col4)
    if out_col4.dtype != torch.int32 or out_col4.numel() != Dw // 4:
        raise TypeError("out_col4 must be int32 with Dw // 4 entries")
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_peer_stage_reduce(_p(stage), int(W), int(B_local), int(Dw), _p(out), out.stride(0),
                                         _p(out_col4), int(bool(clear_stage)), _stream(dev)),
              "b200_peer_stage_reduce")


def peer_barrier(pad_bases: torch.Tensor, rank: int, W: int, epoch_ctr: torch.Tensor) -> None:
    """Cross-rank barrier on the current stream: pad_bases (int64 [W]) are the ranks' uint32[W] signal pads,
    epoch_ctr a device-local int32/uint32 [1] counter of this barrier instance."""
    dev = _need_cuda(pad_bases, epoch_ctr)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_peer_barrier(_p(pad_bases), int(rank), int(W), _p(epoch_ctr), _stream(dev)),
              "b200_peer_barrier")


def peer_stage_reduce_flagged(stage: torch.Tensor, flags: torch.Tensor, W: int, B_local: int, Dw: int,
                              feat_col4: torch.Tensor, out: torch.Tensor, out_col4: torch.Tensor) -> None:
    """Sum of the FLAGGED rank slots of stage ([W, B_local, Dw] fp32); flags is uint8 [B_local, F, W] and is
    cleared where consumed; feat_col4 (int32 [F + 1]) gives every feature's staged float4 columns."""
    dev = _need_cuda(stage, flags, feat_col4, out, out_col4)
    F = feat_col4.numel() - 1
    if flags.dtype != torch.uint8 or flags.numel() != B_local * F * W:
        raise TypeError("flags must be uint8 [B_local, F, W]")
    if out_col4.dtype != torch.int32 or out_col4.numel() != Dw // 4 or feat_col4.dtype != torch.int32:
        raise TypeError("out_col4 / feat_col4 must be int32 with Dw // 4 / F + 1 entries")
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_peer_stage_reduce_flagged(_p(stage), _p(flags), int(W), int(B_local), int(Dw), int(F),
                                                 _p(feat_col4), _p(out), out.stride(0), _p(out_col4),
                                                 _stream(dev)), "b200_peer_stage_reduce_flagged")


# ------------------------------------------------------------------------ dot interaction
def interaction_forward(dense: torch.Tensor, sparse: torch.Tensor, pad_rows: bool = False) -> torch.Tensor:
    """[B, D + (F+1)F/2].  pad_rows=True returns a view of a buffer whose rows sit at a 16-byte
    pitch with zeros in the pad columns (see b200_sparse.h); the view's stride(0) is the pitch."""
    dev = _need_cuda(dense, sparse)
    B, D = dense.shape
    F = sparse.shape[1]
    dense = dense.float().contiguous()
    sparse = sparse.float().contiguous()
    width = D + (F + 1) * F // 2
    pitch = (width + 3) // 4 * 4 if pad_rows else width
    buf = torch.empty((B, pitch), dtype=torch.float32, device=dev)
    with _on(dev), _timed("interaction_forward", dev, 1):
        check(lib.b200_interaction_forward(_p(dense), _p(sparse), B, F, D, _p(buf), pitch, _stream(dev)),
              "b200_interaction_forward")
    return buf[:, :width] if pitch != width else buf


# Gradient sinks: a consumer that needs grad_sparse in a buffer of its own (the peer-memory exchange
# keeps the pooled-embedding gradient in symmetric memory so that other ranks can read it) registers
# that buffer here; interaction_backward then writes grad_sparse straight into it instead of a fresh
# tensor, which saves the consumer a [B, F*D] copy per step.
# Keyed by (device, data_ptr of the consumer's forward output): only the backward of THAT tensor writes
# into the sink; any other [B, F*D] input - a second InteractionArch, an unsharded parity model, a leaf in a
# unit test - gets a fresh gradient tensor.
_GRAD_SINKS: Dict[Tuple[int, int], "GradSink"] = {}


def interaction_supported(F: int, D: int) -> bool:
    """True when b200_interaction_forward/backward cover the shape (csrc/interaction.cu: D % 4 == 0 and
    the general kernel's 4-sample shared-memory tile fits); callers route other shapes to plain torch."""
    n = F + 1
    return D > 0 and D % 4 == 0 and 16 * ((n * (D + 4) + n * (n + 1) + 3) & ~3) <= 200 * 1024


class GradSink:
    """buf: the consumer's [B, F*D] gradient buffer.  feat_base / feat_pitch (int64 device tensors [F],
    optional): per-feature destinations for b200_interaction_backward_scatter - the consumer may point
    a feature at another GPU's buffer.  `scattered` tells the consumer which layout the last backward
    used (True: per-feature destinations; False: everything in buf)."""

    def __init__(self, buf: torch.Tensor, feat_base: Optional[torch.Tensor] = None,
                 feat_pitch: Optional[torch.Tensor] = None, src: Optional[torch.Tensor] = None) -> None:
        self.buf, self.feat_base, self.feat_pitch = buf, feat_base, feat_pitch
        # scatter only the gradient of THIS tensor (the consumer's forward output, features in the
        # order of feat_base): any other [B, F*D] input of the same size keeps the dense layout
        self.src_ptr = src.data_ptr() if src is not None else None
        self.numel = src.numel() if src is not None else buf.numel()
        self.scattered = False
        self.key: Optional[Tuple[int, int]] = None


def register_grad_sink(buf: torch.Tensor, feat_base: Optional[torch.Tensor] = None,
                       feat_pitch: Optional[torch.Tensor] = None, src: Optional[torch.Tensor] = None) -> GradSink:
    """src: the tensor whose gradient the sink receives (the consumer's persistent forward output)."""
    if src is None:
        raise ValueError("register_grad_sink needs the forward output whose gradient goes to the sink")
    sink = GradSink(buf, feat_base, feat_pitch, src)
    sink.key = (buf.device.index, src.data_ptr())
    _GRAD_SINKS[sink.key] = sink
    return sink


def unregister_grad_sink(sink: GradSink) -> None:
    if sink.key is not None and _GRAD_SINKS.get(sink.key) is sink:
        del _GRAD_SINKS[sink.key]


def interaction_backward(dense: torch.Tensor, sparse: torch.Tensor,
                         grad_out: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    dev = _need_cuda(dense, sparse, grad_out)
    B, D = dense.shape
    F = sparse.shape[1]
    grad_out = grad_out.float()
    if grad_out.dim() != 2 or grad_out.stride(1) != 1 or grad_out.stride(0) < grad_out.shape[1]:
        grad_out = grad_out.contiguous()  # a row-pitched view is consumed in place
    gd = torch.empty_like(dense)
    sink = _GRAD_SINKS.get((dev.index, sparse.data_ptr()))
    if sink is not None and (sink.buf.dtype != sparse.dtype or sink.numel != sparse.numel()):
        sink = None
    gs = sink.buf.view(sparse.shape) if sink is not None else torch.empty_like(sparse)
    with _on(dev), _timed("interaction_backward", dev, 1):
        if sink is not None:
            sink.scattered = False
            if sink.feat_base is not None and sink.feat_base.numel() == F and F + 1 <= 32 and D % 32 == 0:
                rc = lib.b200_interaction_backward_scatter(
                    _p(dense), _p(sparse), _p(grad_out), grad_out.stride(0), B, F, D, _p(gd),
                    _p(sink.feat_base), _p(sink.feat_pitch), _stream(dev))
                if rc == 0:
                    sink.scattered = True
                    return gd, gs
                if rc != -2:  # B200_EUNSUPPORTED: the dense layout below; anything else is an error
                    check(rc, "b200_interaction_backward_scatter")
        check(lib.b200_interaction_backward(_p(dense), _p(sparse), _p(grad_out), grad_out.stride(0), B, F, D,
                                            _p(gd), _p(gs), _stream(dev)), "b200_interaction_backward")
    return gd, gs


# ------------------------------------------------------------------------ sequence embeddings
def linearize_seq_indices(indices: torch.Tensor, offsets: torch.Tensor, hash_size_cumsum: torch.Tensor,
                          B: int) -> torch.Tensor:
    """int64 [N]: indices[p] + hash_size_cumsum[feature of position p] (table-major ids)."""
    dev = _need_cuda(indices, offsets, hash_size_cumsum)
    if indices.dtype != offsets.dtype:
        raise TypeError("indices and offsets must share a dtype")
    T = hash_size_cumsum.numel() - 1
    N = indices.numel()
    out = torch.empty(N, dtype=torch.int64, device=dev)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_linearize_seq_indices(_p(indices.contiguous()), _p(offsets.contiguous()),
                                             _idx_dt(indices, "indices"), _p(hash_size_cumsum), T, int(B), N,
                                             _p(out), _stream(dev)), "b200_linearize_seq_indices")
    return out


# ------------------------------------------------------------------------ dense-layer glue
def relu_bias_grad(grad_out: torch.Tensor, out: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(grad_out * (out > 0), its column sums) in one pass - backward of ReLU + bias of a Linear+ReLU layer."""
    dev = _need_cuda(grad_out, out)
    if grad_out.dtype != torch.float32 or out.dtype != torch.float32 or grad_out.dim() != 2:
        raise TypeError("relu_bias_grad expects 2-D float32 tensors")
    if grad_out.stride(1) != 1 or grad_out.stride(0) % 4 or grad_out.data_ptr() % 16:
        grad_out = grad_out.contiguous()
    if out.stride(1) != 1 or out.stride(0) % 4 or out.data_ptr() % 16:
        out = out.contiguous()
    B, N = grad_out.shape
    g = torch.empty((B, N), dtype=torch.float32, device=dev)
    gb = torch.empty(N, dtype=torch.float32, device=dev)
    scratch = torch.empty(lib.b200_relu_bias_grad_scratch_bytes(N) // 4, dtype=torch.float32, device=dev)
    with _on(dev):
        KernelProfile.launches += 2
        check(lib.b200_relu_bias_grad(_p(grad_out), grad_out.stride(0), _p(out), out.stride(0), B, N, _p(g), _p(gb),
                                      _p(scratch), _stream(dev)), "b200_relu_bias_grad")
    return g, gb


# ------------------------------------------------------------------------ pooled a2a layout
def pooled_a2a_recat(src: torch.Tensor, dim_cumsum: torch.Tensor, B_local: int, forward: bool,
                     scale: float = 1.0) -> torch.Tensor:
    dev = _need_cuda(src, dim_cumsum)
    W = dim_cumsum.numel() - 1
    src = src.contiguous()
    dst = torch.empty_like(src)
    with _on(dev):
        KernelProfile.launches += 1
        check(lib.b200_pooled_a2a_recat(_p(src), _p(dst), _p(dim_cumsum), W, int(B_local), int(forward),
                                        float(scale), _stream(dev)), "b200_pooled_a2a_recat")
    return dst
