"""Quantized (int8 rowwise) EmbeddingBagCollection for inference
(reference torchrec/quant/embedding_modules.py: quantize_state_dict :49, EmbeddingBagCollection :84,
forward :190, from_float :245).

Same API and per-table state_dict layout as the reference; the difference is the launch shape:
the reference builds one IntNBit module PER TABLE (26 T=1 launches per forward at Criteo shape,
launch-latency bound); here all tables live in ONE table-batched module and a forward is one
bounds-check + one lookup launch over the whole KJT."""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
import torch.nn as nn
from fbgemm_gpu.split_table_batched_embeddings_ops import (
    EmbeddingLocation, IntNBitTableBatchedEmbeddingBagsCodegen, PoolingMode)
from torch import Tensor

from torchrec.modules.embedding_configs import (
    DATA_TYPE_NUM_BITS, DataType, EmbeddingBagConfig, data_type_to_sparse_type, dtype_to_data_type,
    pooling_type_to_pooling_mode)
from torchrec.modules.embedding_modules import (
    EmbeddingBagCollection as OriginalEmbeddingBagCollection, EmbeddingBagCollectionInterface,
    ebc_get_embedding_names)
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor, KeyedTensor


def quantize_state_dict(module: nn.Module, table_name_to_quantized_weights: Dict[str, Tuple[Tensor, Tensor]],
                        data_type: DataType) -> torch.device:
    device = torch.device("cpu")
    for key, tensor in module.state_dict().items():
        parts = key.split(".")
        assert parts[-1] == "weight"
        table = parts[-2]
        device = tensor.device
        bits = DATA_TYPE_NUM_BITS[data_type]
        if tensor.is_meta:
            q = torch.empty((tensor.shape[0], tensor.shape[1] * bits // 8), device="meta", dtype=torch.uint8)
            ss = torch.empty((tensor.shape[0], 4), device="meta", dtype=torch.uint8)
        else:
            res = torch.ops.fbgemm.FloatToFusedNBitRowwiseQuantizedSBHalf(tensor, bits)
            q, ss = res[:, :-4], res[:, -4:]
        table_name_to_quantized_weights[table] = (q, ss)
    return device


class EmbeddingBagCollection(EmbeddingBagCollectionInterface):
    def __init__(self, table_name_to_quantized_weights: Dict[str, Tuple[Tensor, Tensor]],
                 embedding_configs: List[EmbeddingBagConfig], is_weighted: bool, device: torch.device) -> None:
        super().__init__()
        self._is_weighted = is_weighted
        self._embedding_bag_configs = embedding_configs
        names = set()
        for cfg in embedding_configs:
            if cfg.name in names:
                raise ValueError(f"Duplicate table name {cfg.name}")
            names.add(cfg.name)
            if not cfg.feature_names:
                cfg.feature_names = [cfg.name]
        poolings = {cfg.pooling for cfg in embedding_configs}
        if len(poolings) != 1:
            raise NotImplementedError("all tables of a quantized collection must share one pooling type")
        feature_table_map = [t for t, cfg in enumerate(embedding_configs) for _ in cfg.feature_names]
        self._feature_names: List[str] = [f for cfg in embedding_configs for f in cfg.feature_names]
        self._lengths_per_embedding: List[int] = [cfg.embedding_dim for cfg in embedding_configs
                                                  for _ in cfg.feature_names]
        self._emb_module = IntNBitTableBatchedEmbeddingBagsCodegen(
            embedding_specs=[("", cfg.num_embeddings, cfg.embedding_dim, data_type_to_sparse_type(cfg.data_type),
                              EmbeddingLocation.DEVICE) for cfg in embedding_configs],
            feature_table_map=feature_table_map,
            pooling_mode=pooling_type_to_pooling_mode(next(iter(poolings))),
            weight_lists=[table_name_to_quantized_weights[cfg.name] for cfg in embedding_configs],
            device=device, row_alignment=16)
        self._embedding_names: List[str] = ebc_get_embedding_names(embedding_configs)
        self._order_cache: Dict[Tuple[str, ...], Optional[Tensor]] = {}

    def forward(self, features: KeyedJaggedTensor) -> KeyedTensor:
        keys = tuple(features.keys())
        if keys != tuple(self._feature_names):
            if keys not in self._order_cache:
                order = [features.keys().index(f) for f in self._feature_names]
                self._order_cache[keys] = torch.tensor(order, dtype=torch.int32, device=features.device())
            order_t = self._order_cache[keys]
            features = features.permute(order_t.tolist(), order_t)
        out = self._emb_module(indices=features.values(), offsets=features.offsets(),
                               per_sample_weights=features.weights() if self._is_weighted else None)
        return KeyedTensor(keys=self._embedding_names, values=out.float(),
                           length_per_key=self._lengths_per_embedding)

    def state_dict(self, destination: Optional[Dict[str, Any]] = None, prefix: str = "",
                   keep_vars: bool = False) -> Dict[str, Any]:
        if destination is None:
            destination = OrderedDict()
        for cfg, (w, _) in zip(self._embedding_bag_configs,
                               self._emb_module.split_embedding_weights(split_scale_shifts=False)):
            destination[prefix + f"embedding_bags.{cfg.name}.weight"] = w
        return destination

    def named_buffers(self, prefix: str = "", recurse: bool = True) -> Iterator[Tuple[str, Tensor]]:
        yield from self.state_dict(prefix=prefix, keep_vars=True).items()

    def _get_name(self) -> str:
        return "QuantizedEmbeddingBagCollection"

    @classmethod
    def from_float(cls, module: OriginalEmbeddingBagCollection) -> "EmbeddingBagCollection":
        assert hasattr(module, "qconfig"), "EmbeddingBagCollection input float module must have qconfig defined"
        data_type = dtype_to_data_type(module.qconfig.weight().dtype)
        cfgs = [EmbeddingBagConfig(num_embeddings=c.num_embeddings, embedding_dim=c.embedding_dim, name=c.name,
                                   data_type=data_type, feature_names=list(c.feature_names), pooling=c.pooling,
                                   weight_init_max=c.weight_init_max, weight_init_min=c.weight_init_min)
                for c in module.embedding_bag_configs]
        table_q: Dict[str, Tuple[Tensor, Tensor]] = {}
        device = quantize_state_dict(module, table_q, data_type)
        return cls(table_q, cfgs, module.is_weighted, device=device)

    @property
    def embedding_bag_configs(self) -> List[EmbeddingBagConfig]:
        return self._embedding_bag_configs

    @property
    def is_weighted(self) -> bool:
        return self._is_weighted
