"""Embedding table configuration types (reference torchrec/modules/embedding_configs.py:19-145)."""
from dataclasses import dataclass, field
from enum import Enum, unique
from math import sqrt
from typing import Dict, List, Optional

import torch
from fbgemm_gpu.split_embedding_configs import SparseType
from fbgemm_gpu.split_table_batched_embeddings_ops import PoolingMode


@unique
class PoolingType(Enum):
    SUM = "SUM"
    MEAN = "MEAN"
    NONE = "NONE"


@unique
class DataType(Enum):
    FP32 = "FP32"
    FP16 = "FP16"
    INT8 = "INT8"
    INT4 = "INT4"
    INT2 = "INT2"


DATA_TYPE_NUM_BITS: Dict[DataType, int] = {
    DataType.FP32: 32, DataType.FP16: 16, DataType.INT8: 8, DataType.INT4: 4, DataType.INT2: 2}

_POOL = {PoolingType.SUM: (PoolingMode.SUM, "sum"), PoolingType.MEAN: (PoolingMode.MEAN, "mean")}


def dtype_to_data_type(dtype: torch.dtype) -> DataType:
    table = {torch.quint8: DataType.INT8, torch.qint8: DataType.INT8, torch.quint4x2: DataType.INT4,
             torch.quint2x4: DataType.INT2}
    if dtype not in table:
        raise Exception(f"Invalid data type {dtype}")
    return table[dtype]


def pooling_type_to_pooling_mode(pooling_type: PoolingType) -> PoolingMode:
    if pooling_type == PoolingType.NONE:
        return PoolingMode.NONE
    if pooling_type not in _POOL:
        raise Exception(f"Invalid pooling type {pooling_type}")
    return _POOL[pooling_type][0]


def pooling_type_to_str(pooling_type: PoolingType) -> str:
    if pooling_type not in _POOL:
        raise ValueError(f"Unsupported pooling type {pooling_type}")
    return _POOL[pooling_type][1]


def data_type_to_sparse_type(data_type: DataType) -> SparseType:
    try:
        return SparseType(data_type.value.lower())
    except (ValueError, AttributeError):
        raise ValueError(f"Invalid DataType {data_type}")


@dataclass
class BaseEmbeddingConfig:
    num_embeddings: int
    embedding_dim: int
    name: str = ""
    data_type: DataType = DataType.FP32
    feature_names: List[str] = field(default_factory=list)
    weight_init_max: Optional[float] = None
    weight_init_min: Optional[float] = None

    def get_weight_init_max(self) -> float:
        return sqrt(1 / self.num_embeddings) if self.weight_init_max is None else self.weight_init_max

    def get_weight_init_min(self) -> float:
        return -sqrt(1 / self.num_embeddings) if self.weight_init_min is None else self.weight_init_min

    def num_features(self) -> int:
        return len(self.feature_names)


@dataclass
class EmbeddingTableConfig(BaseEmbeddingConfig):
    pooling: PoolingType = PoolingType.SUM
    is_weighted: bool = False
    has_feature_processor: bool = False
    embedding_names: List[str] = field(default_factory=list)


@dataclass
class EmbeddingBagConfig(BaseEmbeddingConfig):
    pooling: PoolingType = PoolingType.SUM


@dataclass
class EmbeddingConfig(BaseEmbeddingConfig):
    pass
