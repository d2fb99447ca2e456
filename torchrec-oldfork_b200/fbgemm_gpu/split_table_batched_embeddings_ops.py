"""fbgemm_gpu.split_table_batched_embeddings_ops — the module surface torchrec constructs
(torchrec/distributed/batched_embedding_kernel.py:18-25,604-647; quant/embedding_modules.py:14-17),
re-implemented over the B200 C ABI.

Mirrors the names, constructor arguments and methods of
third_party/fbgemm/fbgemm_gpu/fbgemm_gpu/split_table_batched_embeddings_ops.py
(SplitTableBatchedEmbeddingBagsCodegen :167, IntNBitTableBatchedEmbeddingBagsCodegen :1533,
rounded_row_size_in_bytes :1469) for device-resident tables; UVM / caching / host placements and
the other ten optimizers are out of scope and raise NotImplementedError.
"""
from __future__ import annotations

import enum
from dataclasses import dataclass
from itertools import accumulate
from math import log2
from typing import List, Optional, Tuple, Union

import torch
from torch import Tensor, nn

from . import _tbe_autograd
from .split_embedding_configs import EmbOptimType as OptimType, SparseType


class EmbeddingLocation(enum.IntEnum):
    DEVICE = 0
    MANAGED = 1
    MANAGED_CACHING = 2
    HOST = 3


class ComputeDevice(enum.IntEnum):
    CPU = 0
    CUDA = 1


class CacheAlgorithm(enum.Enum):
    LRU = 0
    LFU = 1


class PoolingMode(enum.IntEnum):
    SUM = 0
    MEAN = 1
    NONE = 2


class BoundsCheckMode(enum.IntEnum):
    FATAL = 0
    WARNING = 1
    IGNORE = 2
    NONE = 3


_TORCH_DTYPE = {SparseType.FP32: torch.float32, SparseType.FP16: torch.float16,
                SparseType.BF16: torch.bfloat16, SparseType.INT8: torch.uint8}


def round_up(a: int, b: int) -> int:
    return int((a + b - 1) // b) * b


def unpadded_row_size_in_bytes(dim: int, weight_ty: SparseType) -> int:
    return {"fp32": dim * 4, "fp16": dim * 2, "int8": dim + 4, "int4": dim // 2 + 4,
            "int2": dim // 4 + 4}[weight_ty.value]


def rounded_row_size_in_bytes(dim: int, weight_ty: SparseType, row_alignment: int) -> int:
    return round_up(unpadded_row_size_in_bytes(dim, weight_ty), row_alignment)


@dataclass
class _OptimizerArgs:
    """Subset of invokers.lookup_args.OptimizerArgs (codegen/lookup_args.py) this build uses."""
    stochastic_rounding: bool
    gradient_clipping: bool
    max_gradient: float
    learning_rate: float
    eps: float


class SplitTableBatchedEmbeddingBagsCodegen(nn.Module):
    """Table-batched nn.EmbeddingBag(sparse=True) with the optimizer fused into backward.

    forward(indices, offsets, per_sample_weights=None) -> [B, sum_t D_t]; offsets are
    table-major (bag (t, b) at offsets[t*B + b]).  backward applies EXACT_SGD or
    EXACT_ROWWISE_ADAGRAD in place and returns no weight gradient.
    """

    def __init__(
        self,
        embedding_specs: List[Tuple[int, int, EmbeddingLocation, ComputeDevice]],
        feature_table_map: Optional[List[int]] = None,
        cache_algorithm: CacheAlgorithm = CacheAlgorithm.LRU,
        cache_load_factor: float = 0.2,
        cache_sets: int = 0,
        cache_reserved_memory: float = 0.0,
        cache_precision: SparseType = SparseType.FP32,
        weights_precision: SparseType = SparseType.FP32,
        output_dtype: SparseType = SparseType.FP32,
        enforce_hbm: bool = False,
        optimizer: OptimType = OptimType.EXACT_SGD,
        record_cache_metrics=None,
        stochastic_rounding: bool = True,
        gradient_clipping: bool = False,
        max_gradient: float = 1.0,
        learning_rate: float = 0.01,
        eps: float = 1.0e-8,
        momentum: float = 0.9,
        weight_decay: float = 0.0,
        eta: float = 0.001,
        beta1: float = 0.9,
        beta2: float = 0.999,
        pooling_mode: PoolingMode = PoolingMode.SUM,
        device: Optional[torch.device] = None,
        bounds_check_mode: BoundsCheckMode = BoundsCheckMode.WARNING,
    ) -> None:
        super().__init__()
        self.pooling_mode = pooling_mode
        self.bounds_check_mode_int: int = bounds_check_mode.value
        self.weights_precision = weights_precision
        self.output_dtype: int = output_dtype.as_int()
        self.embedding_specs = embedding_specs
        rows, dims, locations, compute_devices = zip(*embedding_specs)
        T_ = len(embedding_specs)
        assert T_ > 0
        if any(loc != EmbeddingLocation.DEVICE for loc in locations) or any(
            cd != ComputeDevice.CUDA for cd in compute_devices
        ):
            raise NotImplementedError(
                "only EmbeddingLocation.DEVICE + ComputeDevice.CUDA tables are built (HBM-resident, "
                "sm_100a); HOST/MANAGED/MANAGED_CACHING placements and the CPU kernels are out of scope"
            )
        if pooling_mode == PoolingMode.NONE and len(set(dims)) != 1:
            raise ValueError("PoolingMode.NONE (sequence embeddings) needs one embedding dim for all tables")
        if optimizer not in (OptimType.EXACT_SGD, OptimType.EXACT_ROWWISE_ADAGRAD):
            raise NotImplementedError(f"optimizer {optimizer} is not built; use EXACT_SGD or EXACT_ROWWISE_ADAGRAD")
        if weights_precision not in (SparseType.FP32, SparseType.FP16, SparseType.BF16):
            raise NotImplementedError(f"weights_precision {weights_precision} is not built for training")
        if output_dtype not in (SparseType.FP32, SparseType.FP16):
            raise NotImplementedError("output_dtype must be FP32 or FP16")
        for d in dims:
            assert d % 4 == 0, f"embedding dim {d} must be a multiple of 4"
        if device is not None:
            self.current_device = torch.device(device)
        else:
            self.current_device = torch.device("cuda", torch.cuda.current_device())
        dev = self.current_device
        # fp16 / bf16 rows are written back with stochastic rounding when asked (the reference's default,
        # :200); the Philox state of a step is drawn from the device's CUDA generator, as the reference does
        self._sr_on = bool(stochastic_rounding) and weights_precision != SparseType.FP32
        self.optimizer = optimizer
        self.optimizer_args = _OptimizerArgs(stochastic_rounding, gradient_clipping, max_gradient,
                                             learning_rate, eps)
        self.placeholder_autograd_tensor = nn.Parameter(torch.zeros(0, device=dev, dtype=torch.float))

        self.feature_table_map: List[int] = (
            list(feature_table_map) if feature_table_map is not None else list(range(T_))
        )
        T = len(self.feature_table_map)
        assert T_ <= T
        assert set(self.feature_table_map) == set(range(T_)), "Each table must have at least one feature!"
        D_offsets = [0] + list(accumulate(dims[t] for t in self.feature_table_map))
        self.total_D: int = D_offsets[-1]
        self.max_D: int = max(dims)
        self._all_D_mult8 = all(d % 8 == 0 for d in dims)
        table_hash_cumsum = [0] + list(accumulate(rows))
        self.total_hash_size_bits: int = (
            0 if table_hash_cumsum[-1] == 0 else int(log2(float(table_hash_cumsum[-1])) + 1)
        )
        # key width of the largest table: a hint for the index transpose (how many radix passes to launch); only
        # valid when features are grouped by table, i.e. the per-feature hash bases never go back
        grouped = self.feature_table_map == sorted(self.feature_table_map)
        self.max_table_bits: int = max((max(r, 1) - 1).bit_length() for r in rows) if grouped and rows else 0
        table_weight_offsets = [0] + list(accumulate(r * d for r, d in zip(rows, dims)))
        i64 = dict(device=dev, dtype=torch.int64)
        self.register_buffer("D_offsets", torch.tensor(D_offsets, device=dev, dtype=torch.int32))
        self.register_buffer(
            "hash_size_cumsum",
            torch.tensor([table_hash_cumsum[t] for t in self.feature_table_map] + [table_hash_cumsum[-1]], **i64))
        self.register_buffer("rows_per_table", torch.tensor([rows[t] for t in self.feature_table_map], **i64))
        self.register_buffer("bounds_check_warning", torch.tensor([0], **i64))
        self.register_buffer("weights_offsets",
                             torch.tensor([table_weight_offsets[t] for t in self.feature_table_map], **i64))
        self.register_buffer("weights_placements",
                             torch.zeros(T, device=dev, dtype=torch.int32))  # all DEVICE
        self._table_weight_offsets = table_weight_offsets
        self._table_row_offsets = table_hash_cumsum
        wdt = _TORCH_DTYPE[weights_precision]
        self.register_buffer("weights_dev", torch.zeros(table_weight_offsets[-1], device=dev, dtype=wdt))
        # empty placeholders kept for attribute compatibility with the reference module
        self.register_buffer("weights_host", torch.zeros(0, device=dev, dtype=wdt))
        self.register_buffer("weights_uvm", torch.zeros(0, device=dev, dtype=wdt))
        if optimizer == OptimType.EXACT_ROWWISE_ADAGRAD:
            self.register_buffer("momentum1_dev", torch.zeros(table_hash_cumsum[-1], device=dev, dtype=torch.float32))
            self.register_buffer("momentum1_offsets",
                                 torch.tensor([table_hash_cumsum[t] for t in self.feature_table_map], **i64))
        else:
            self.register_buffer("momentum1_dev", torch.zeros(0, device=dev, dtype=torch.float32))
            self.register_buffer("momentum1_offsets", torch.zeros(T, **i64))
        self.step = 0
        self._seq_view: Optional["_SequenceView"] = None
        if pooling_mode == PoolingMode.NONE:
            self.total_D = dims[0]  # output is [N, D]
            self.register_buffer("seq_D_offsets", torch.tensor([0, dims[0]], device=dev, dtype=torch.int32))
            self.register_buffer("seq_hash_size_cumsum", torch.tensor([0, table_hash_cumsum[-1]], **i64))
            self.register_buffer("seq_zero_offset", torch.zeros(1, **i64))
            self._seq_view = _SequenceView(self)

    # ------------------------------------------------------------------ forward
    def forward(self, indices: Tensor, offsets: Tensor, per_sample_weights: Optional[Tensor] = None,
                feature_requires_grad: Optional[Tensor] = None) -> Tensor:
        indices, offsets = self.prepare_indices(indices, offsets)
        if self._seq_view is not None:
            # sequence embeddings: [N, D], one row per id (reference `nobag` kernels).  The ids are
            # rebased into ONE virtual table (all tables share D and are contiguous) and the pooled
            # kernels run with T = 1, B = N, L = 1: same gather, same sorted fused backward.
            if per_sample_weights is not None:
                raise ValueError("PoolingMode.NONE takes no per_sample_weights")
            T = self.D_offsets.numel() - 1
            B = (offsets.numel() - 1) // T
            lin = torch.ops.fbgemm.b200_linearize_seq_indices(indices, offsets, self.hash_size_cumsum, B)
            seq_offsets = torch.arange(lin.numel() + 1, device=lin.device, dtype=torch.int64)
            return _tbe_autograd.lookup(self._seq_view, lin, seq_offsets, None, None)
        return _tbe_autograd.lookup(self, indices, offsets, per_sample_weights, feature_requires_grad)

    def prepare_indices(self, indices: Tensor, offsets: Tensor) -> Tuple[Tensor, Tensor]:
        """Dtype normalisation + bounds check (reference :585-595) shared by forward() and the fused
        lookup+exchange path of torchrec.distributed.peer_dist."""
        if indices.dtype != offsets.dtype or indices.dtype not in (torch.int32, torch.int64):
            indices, offsets = indices.long(), offsets.long()
        if self.bounds_check_mode_int != BoundsCheckMode.NONE.value:
            torch.ops.fbgemm.bounds_check_indices(self.rows_per_table, indices, offsets,
                                                  self.bounds_check_mode_int, self.bounds_check_warning)
        self.step += 1
        return indices, offsets

    def prefetch(self, indices: Tensor, offsets: Tensor) -> None:
        """No UVM cache in this build: nothing to prefetch (reference :736-743 early-returns too)."""
        return None

    # ------------------------------------------------------------------ state access
    def split_embedding_weights(self) -> List[Tensor]:
        """Views [rows_t, D_t] into the live buffer, one per TABLE (reference :887-917)."""
        out = []
        for t, (r, d, _, _) in enumerate(self.embedding_specs):
            o = self._table_weight_offsets[t]
            out.append(self.weights_dev.detach()[o:o + r * d].view(r, d))
        return out

    def split_optimizer_states(self) -> List[Tuple[Tensor]]:
        """Per table tuple of optimizer-state views (reference :948-1024): (momentum1[rows],) for rowwise
        Adagrad; for SGD, which has no state, the reference's `list(zip(*[]))` - an EMPTY list (its torchrec
        wrapper walks it with zip_longest, batched_embedding_kernel.py:136-148)."""
        if self.optimizer != OptimType.EXACT_ROWWISE_ADAGRAD:
            return []
        out = []
        for t, (r, _, _, _) in enumerate(self.embedding_specs):
            o = self._table_row_offsets[t]
            out.append((self.momentum1_dev.detach()[o:o + r],))
        return out

    def get_optimizer_state(self):
        if self.optimizer == OptimType.EXACT_ROWWISE_ADAGRAD:
            return [{"sum": s[0]} for s in self.split_optimizer_states()]
        return [{} for _ in self.embedding_specs]

    def init_embedding_weights_uniform(self, min_val: float, max_val: float) -> None:
        for w in self.split_embedding_weights():
            w.uniform_(min_val, max_val)

    def _next_sr_state(self):
        """(stochastic_rounding, seed, offset) of this backward: the CUDA generator's Philox seed and a
        fresh offset (the generator is advanced like at::PhiloxCudaState does, so runs are reproducible
        under torch.cuda.manual_seed and independent between modules and steps)."""
        if not self._sr_on:
            return False, 0, 0
        dev = self.current_device
        gen = torch.cuda.default_generators[dev.index if dev.index is not None else torch.cuda.current_device()]
        offset = gen.get_offset()
        gen.set_offset(offset + 4)
        return True, gen.initial_seed() & (2 ** 63 - 1), offset // 4 + 1

    def set_learning_rate(self, lr: float) -> None:
        self.optimizer_args.learning_rate = float(lr)

    def flush(self) -> None:
        """No software cache to flush (reference :1050)."""
        return None

    def reset_cache_states(self) -> None:
        return None


class _SequenceView:
    """The attributes _tbe_autograd.lookup reads, describing a PoolingMode.NONE module as ONE table of
    sum(rows) rows pooled with one id per bag.  Buffers are read through the owner at call time so that
    .to() / load_state_dict on the owner stay visible; optimizer_args is the owner's object."""

    def __init__(self, owner: "SplitTableBatchedEmbeddingBagsCodegen") -> None:
        self._o = owner

    pooling_mode = PoolingMode.SUM

    def __getattr__(self, name: str):
        o = self.__dict__["_o"]
        if name in ("weights_offsets", "momentum1_offsets"):
            return o.seq_zero_offset
        if name == "D_offsets":
            return o.seq_D_offsets
        if name == "hash_size_cumsum":
            return o.seq_hash_size_cumsum
        if name == "max_table_bits":
            return 0  # one "table" spanning every row: the key width is total_hash_size_bits
        return getattr(o, name)


class DenseTableBatchedEmbeddingBagsCodegen(nn.Module):
    """Imported by torchrec (batched_embedding_kernel.py:21); the BATCHED_DENSE kernel is outside
    the accelerated path and is not built."""

    def __init__(self, *args, **kwargs) -> None:
        super().__init__()
        raise NotImplementedError("DenseTableBatchedEmbeddingBagsCodegen is out of scope; use the fused kernel")


class IntNBitTableBatchedEmbeddingBagsCodegen(nn.Module):
    """Inference lookup over rowwise-quantized tables (reference :1533).  INT8 rows only:
    [fp16 scale, fp16 bias, D x uint8] padded to `row_alignment` bytes; fp16 or fp32 output.
    Unlike the reference's per-precision kernel zoo, every table of the module is served by one
    launch (tbe_int8.cu)."""

    def __init__(
        self,
        embedding_specs: List[Tuple[str, int, int, SparseType, EmbeddingLocation]],
        feature_table_map: Optional[List[int]] = None,
        index_remapping: Optional[List[Tensor]] = None,
        pooling_mode: PoolingMode = PoolingMode.SUM,
        device: Optional[Union[str, int, torch.device]] = None,
        bounds_check_mode: BoundsCheckMode = BoundsCheckMode.WARNING,
        weight_lists: Optional[List[Tuple[Tensor, Tensor]]] = None,
        load_factor: float = 0.5,
        use_array_for_index_remapping: bool = True,
        output_dtype: SparseType = SparseType.FP16,
        cache_algorithm: CacheAlgorithm = CacheAlgorithm.LRU,
        cache_load_factor: float = 0.2,
        cache_sets: int = 0,
        cache_reserved_memory: float = 0.0,
        enforce_hbm: bool = False,
        record_cache_metrics=None,
        row_alignment: Optional[int] = None,
    ) -> None:
        super().__init__()
        if device is None:
            self.current_device = torch.device("cuda", torch.cuda.current_device())
        else:
            self.current_device = torch.device(device)
        if index_remapping is not None:
            raise NotImplementedError("pruned-row index remapping is out of scope")
        if pooling_mode == PoolingMode.NONE:
            raise NotImplementedError("PoolingMode.NONE is out of scope")
        if output_dtype not in (SparseType.FP16, SparseType.FP32):
            raise NotImplementedError("output_dtype must be FP16 or FP32")
        self.pooling_mode = pooling_mode
        self.bounds_check_mode_int: int = bounds_check_mode.value
        self.embedding_specs = embedding_specs
        self.output_dtype: int = output_dtype.as_int()
        self.row_alignment: int = 16 if row_alignment is None else int(row_alignment)
        names, rows, dims, tys, locs = zip(*embedding_specs)
        for ty, loc in zip(tys, locs):
            if ty != SparseType.INT8:
                raise NotImplementedError(f"{ty} rows are not built; only SparseType.INT8 (SURVEY.md §8 a16)")
            if loc != EmbeddingLocation.DEVICE:
                raise NotImplementedError("only EmbeddingLocation.DEVICE tables are built")
        T_ = len(embedding_specs)
        self.feature_table_map: List[int] = (
            list(feature_table_map) if feature_table_map is not None else list(range(T_))
        )
        T = len(self.feature_table_map)
        dev = self.current_device
        D_offsets = [0] + list(accumulate(dims[t] for t in self.feature_table_map))
        self.total_D: int = D_offsets[-1]
        self.max_D: int = max(dims)
        row_bytes = [rounded_row_size_in_bytes(d, ty, self.row_alignment) for d, ty in zip(dims, tys)]
        # tables start on 128-byte boundaries like the reference (align_to_cacheline :1488)
        offs, cur = [], 0
        for r, rb in zip(rows, row_bytes):
            offs.append(cur)
            cur = round_up(cur + r * rb, 128)
        self._table_byte_offsets = offs
        self._row_bytes = row_bytes
        self.register_buffer("D_offsets", torch.tensor(D_offsets, device=dev, dtype=torch.int32))
        self.register_buffer("weights_offsets",
                             torch.tensor([offs[t] for t in self.feature_table_map], device=dev, dtype=torch.int64))
        self.register_buffer("rows_per_table",
                             torch.tensor([rows[t] for t in self.feature_table_map], device=dev, dtype=torch.int64))
        self.register_buffer("bounds_check_warning", torch.tensor([0], device=dev, dtype=torch.int64))
        self.register_buffer("weights_dev", torch.zeros(cur, device=dev, dtype=torch.uint8))
        self.weight_initialized = True
        if weight_lists is not None:
            self.assign_embedding_weights(weight_lists)

    def initialize_weights(self) -> None:
        self.weight_initialized = True

    def split_embedding_weights(self, split_scale_shifts: bool = True) -> List[Tuple[Tensor, Optional[Tensor]]]:
        out: List[Tuple[Tensor, Optional[Tensor]]] = []
        for t, (_, r, d, ty, _) in enumerate(self.embedding_specs):
            o, rb = self._table_byte_offsets[t], self._row_bytes[t]
            ws = self.weights_dev.detach()[o:o + r * rb].view(r, rb)
            if split_scale_shifts:
                ws = ws[:, : unpadded_row_size_in_bytes(d, ty)]
                out.append((ws[:, 4:], ws[:, :4]))
            else:
                out.append((ws, None))
        return out

    def assign_embedding_weights(self, q_weight_list: List[Tuple[Tensor, Optional[Tensor]]]) -> None:
        dst = self.split_embedding_weights()
        assert len(q_weight_list) == len(dst)
        for (dw, dss), (sw, sss) in zip(dst, q_weight_list):
            dw.copy_(sw.to(dw.device).view(torch.uint8) if sw.dtype != torch.uint8 else sw.to(dw.device))
            if sss is not None:
                dss.copy_(sss.to(dss.device))

    def fill_random_weights(self) -> None:
        for w, _ in self.split_embedding_weights():
            w.copy_(torch.randint(0, 255, size=w.shape, dtype=torch.uint8, device=self.current_device))

    def forward(self, indices: Tensor, offsets: Tensor, per_sample_weights: Optional[Tensor] = None) -> Tensor:
        if indices.dtype != offsets.dtype or indices.dtype not in (torch.int32, torch.int64):
            indices, offsets = indices.int(), offsets.int()
        if self.bounds_check_mode_int != BoundsCheckMode.NONE.value:
            torch.ops.fbgemm.bounds_check_indices(self.rows_per_table, indices, offsets,
                                                  self.bounds_check_mode_int, self.bounds_check_warning)
        return torch.ops.fbgemm.b200_tbe_int8_forward(
            self.weights_dev, self.weights_offsets, self.D_offsets, self.total_D, self.max_D,
            self.row_alignment, indices, offsets, per_sample_weights, int(self.pooling_mode), self.output_dtype)
