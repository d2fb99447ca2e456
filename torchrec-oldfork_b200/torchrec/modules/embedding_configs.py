"""Embedding table configuration types (reference torchrec/modules/embedding_configs.py:19-145)."""
from dataclasses import dataclass, field
from enum import Enum, unique
from math import sqrt
from typing import Dict, List, Optional

import torch
from fbgemm_gpu.split_embedding_configs import SparseType
from fbgemm_gpu.split_table_batched_embeddings_ops import PoolingMode


# name -> (bits per element, fbgemm SparseType value); INT4 / INT2 exist for configuration parity only
_DATA_TYPES = {"FP32": (32, "fp32"), "FP16": (16, "fp16"), "INT8": (8, "int8"), "INT4": (4, "int4"), "INT2": (2, "int2")}
# name -> (table-batched kernel mode, nn.EmbeddingBag mode string)
_POOLINGS = {"SUM": (PoolingMode.SUM, "sum"), "MEAN": (PoolingMode.MEAN, "mean"), "NONE": (PoolingMode.NONE, None)}

PoolingType = unique(Enum("PoolingType", {n: n for n in _POOLINGS}, module=__name__))
DataType = unique(Enum("DataType", {n: n for n in _DATA_TYPES}, module=__name__))
DATA_TYPE_NUM_BITS: Dict[DataType, int] = {DataType[n]: bits for n, (bits, _) in _DATA_TYPES.items()}
_QUANTIZED_DTYPES = {torch.quint8: DataType.INT8, torch.qint8: DataType.INT8, torch.quint4x2: DataType.INT4,
                     torch.quint2x4: DataType.INT2}


def dtype_to_data_type(dtype: torch.dtype) -> DataType:
    """Quantized torch dtype of a state_dict entry -> DataType (quant/embedding_modules.py)."""
    found = _QUANTIZED_DTYPES.get(dtype)
    if found is None:
        raise Exception(f"Invalid data type {dtype}")
    return found


def pooling_type_to_pooling_mode(pooling_type: PoolingType) -> PoolingMode:
    entry = _POOLINGS.get(getattr(pooling_type, "name", None))
    if entry is None:
        raise Exception(f"Invalid pooling type {pooling_type}")
    return entry[0]


def pooling_type_to_str(pooling_type: PoolingType) -> str:
    entry = _POOLINGS.get(getattr(pooling_type, "name", None))
    if entry is None or entry[1] is None:
        raise ValueError(f"Unsupported pooling type {pooling_type}")
    return entry[1]


def data_type_to_sparse_type(data_type: DataType) -> SparseType:
    entry = _DATA_TYPES.get(getattr(data_type, "name", None))
    if entry is None:
        raise ValueError(f"Invalid DataType {data_type}")
    return SparseType(entry[1])


def _default_init_bound(num_embeddings: int) -> float:
    return sqrt(1.0 / num_embeddings)


@dataclass
class BaseEmbeddingConfig:
    num_embeddings: int
    embedding_dim: int
    name: str = ""
    data_type: DataType = DataType.FP32
    feature_names: List[str] = field(default_factory=list)
    weight_init_max: Optional[float] = None
    weight_init_min: Optional[float] = None

    # uniform(-1/sqrt(rows), 1/sqrt(rows)) unless the config says otherwise (reference :95-106)
    def get_weight_init_max(self) -> float:
        return self.weight_init_max if self.weight_init_max is not None else _default_init_bound(self.num_embeddings)

    def get_weight_init_min(self) -> float:
        return self.weight_init_min if self.weight_init_min is not None else -_default_init_bound(self.num_embeddings)

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
