"""Enums of fbgemm_gpu.split_embedding_configs as imported by torchrec
(torchrec/modules/embedding_configs.py:14, distributed/batched_embedding_kernel.py:17).
Value strings and integer codes follow
third_party/fbgemm/fbgemm_gpu/fbgemm_gpu/split_embedding_configs.py:14-112."""
import enum
from typing import Dict


@enum.unique
class EmbOptimType(enum.Enum):
    SGD = "sgd"
    EXACT_SGD = "exact_sgd"
    LAMB = "lamb"
    ADAM = "adam"
    EXACT_ADAGRAD = "exact_adagrad"
    EXACT_ROWWISE_ADAGRAD = "exact_row_wise_adagrad"
    LARS_SGD = "lars_sgd"
    PARTIAL_ROWWISE_ADAM = "partial_row_wise_adam"
    PARTIAL_ROWWISE_LAMB = "partial_row_wise_lamb"
    ROWWISE_ADAGRAD = "row_wise_adagrad"
    MADGRAD = "madgrad"
    EXACT_ROWWISE_WEIGHTED_ADAGRAD = "exact_row_wise_weighted_adagrad"

    def __str__(self) -> str:
        return self.value


_ORDER = ["fp32", "fp16", "int8", "int4", "int2", "bf16"]
_BITS = {"fp32": 32, "fp16": 16, "int8": 8, "int4": 4, "int2": 2, "bf16": 16}
_ALIGN = {"fp32": 1, "fp16": 2, "int8": 4, "int4": 8, "int2": 16, "bf16": 2}


@enum.unique
class SparseType(enum.Enum):
    FP32 = "fp32"
    FP16 = "fp16"
    INT8 = "int8"
    INT4 = "int4"
    INT2 = "int2"
    BF16 = "bf16"

    def __str__(self) -> str:
        return self.value

    @staticmethod
    def from_int(ty: int) -> "SparseType":
        if not 0 <= ty < len(_ORDER):
            raise ValueError(f"Unsupported sparse type: {ty}")
        return SparseType(_ORDER[ty])

    def as_int(self) -> int:
        return _ORDER.index(self.value)

    def bit_rate(self) -> int:
        return _BITS[self.value]

    def align_size(self) -> int:
        return _ALIGN[self.value]

    def is_float(self) -> bool:
        return self.value in ("fp32", "fp16", "bf16")


ELEMENT_SIZE: Dict[SparseType, int] = {
    SparseType.FP32: 4,
    SparseType.FP16: 2,
    SparseType.INT8: 1,
    SparseType.BF16: 2,
}
