"""Table-batched fused lookup wrapper (reference torchrec/distributed/batched_embedding_kernel.py:
EmbeddingFusedOptimizer :53, BaseBatchedEmbeddingBag :498, BatchedFusedEmbeddingBag :604).

One GroupedEmbeddingConfig -> one SplitTableBatchedEmbeddingBagsCodegen (sm_100a kernels).  The
module owns the flat weight / optimizer-state buffers; state_dict and the fused optimizer expose
views of them."""
from __future__ import annotations

import copy
import itertools
from typing import Any, Dict, Iterator, List, Optional, Tuple

import torch
import torch.distributed as dist
from fbgemm_gpu.split_embedding_configs import EmbOptimType, SparseType
from fbgemm_gpu.split_table_batched_embeddings_ops import (
    ComputeDevice, EmbeddingLocation, PoolingMode, SplitTableBatchedEmbeddingBagsCodegen)
from torch import nn

from torchrec.distributed.embedding_kernel import get_state_dict, load_state_dict_into, table_sharded_tensor
from torchrec.distributed.embedding_types import GroupedEmbeddingConfig, ShardedEmbeddingTable
from torchrec.modules.embedding_configs import data_type_to_sparse_type, pooling_type_to_pooling_mode
from torchrec.optim.fused import FusedOptimizer, FusedOptimizerModule
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor


class EmbeddingFusedOptimizer(FusedOptimizer):
    """Exposes the TBE module's in-place optimizer: params are the per-table weight views, state is
    `<table>.momentum1` for rowwise Adagrad; step()/zero_grad() only push the learning rate
    (reference :250-257).  With a process group both are ShardedTensors over the table's shards
    (reference :160-246: weights with the table's metadata, momentum1 sharded along the rows) whose
    local shard is the live view; without one (unsharded use) they are the views themselves."""

    def __init__(self, config: GroupedEmbeddingConfig, emb_module: SplitTableBatchedEmbeddingBagsCodegen,
                 pg: Optional[dist.ProcessGroup] = None) -> None:
        self._emb_module = emb_module
        self._pg = pg
        params: Dict[str, torch.Tensor] = {}
        state: Dict[Any, Any] = {}
        weights = emb_module.split_embedding_weights()
        opt_states = emb_module.split_optimizer_states()
        for table, w, st in itertools.zip_longest(config.embedding_tables, weights, opt_states):  # SGD: no states
            name = f"{table.name}.weight"
            sharded = pg is not None and table.global_metadata is not None and table.local_metadata is not None
            if sharded:
                w = table_sharded_tensor(table, w, pg)
            params[name] = w
            state[w] = {}
            if emb_module.optimizer == EmbOptimType.EXACT_ROWWISE_ADAGRAD:
                m1 = st[0]
                state[w][f"{table.name}.momentum1"] = (table_sharded_tensor(table, m1, pg, rowwise=True)
                                                       if sharded else m1)
        param_group = {"params": list(params.values()), "lr": emb_module.optimizer_args.learning_rate}
        super().__init__(params, state, [param_group])

    def zero_grad(self, set_to_none: bool = False) -> None:
        self._emb_module.set_learning_rate(self.param_groups[0]["lr"])

    def step(self, closure: Any = None) -> None:
        self._emb_module.set_learning_rate(self.param_groups[0]["lr"])


class BaseBatchedEmbeddingBag(nn.Module):
    def __init__(self, config: GroupedEmbeddingConfig, pg: Optional[dist.ProcessGroup] = None,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        torch._C._log_api_usage_once(f"torchrec.distributed.{self.__class__.__name__}")
        self._config = config
        self._pg = pg
        self._pooling: PoolingMode = pooling_type_to_pooling_mode(config.pooling)
        self._local_rows = [t.local_rows for t in config.embedding_tables]
        self._local_cols = [t.local_cols for t in config.embedding_tables]
        self._weight_init_mins = [t.get_weight_init_min() for t in config.embedding_tables]
        self._weight_init_maxs = [t.get_weight_init_max() for t in config.embedding_tables]
        self._feature_table_map: List[int] = []
        self._emb_names: List[str] = []
        self._lengths_per_emb: List[int] = []
        for idx, t in enumerate(config.embedding_tables):
            self._feature_table_map.extend([idx] * t.num_features())
            self._emb_names.extend(t.embedding_names)
            self._lengths_per_emb.extend([t.local_cols] * t.num_features())

    def init_parameters(self) -> None:
        """uniform(weight_init_min, weight_init_max) per table (reference :530-544)."""
        for rows, cols, lo, hi, w in zip(self._local_rows, self._local_cols, self._weight_init_mins,
                                         self._weight_init_maxs, self.emb_module.split_embedding_weights()):
            assert w.shape == (rows, cols)
            w.uniform_(lo, hi)

    def forward(self, features: KeyedJaggedTensor) -> torch.Tensor:
        # the kernels read int32 or int64 ids directly: no .long() copies (reference :551-552)
        return self.emb_module(indices=features.values(), offsets=features.offsets(),
                               per_sample_weights=features.weights_or_none())

    def state_dict(self, destination=None, prefix: str = "", keep_vars: bool = False) -> Dict[str, Any]:
        """`<table>.weight` -> ShardedTensor whose local shard views the live table (reference :557-570 ->
        embedding_kernel.get_state_dict); plain views without a process group."""
        self.flush()
        return get_state_dict(self._config.embedding_tables, self.emb_module.split_embedding_weights(), self._pg,
                              destination, prefix)

    def load_state_dict(self, state_dict, strict: bool = True):
        missing, unexpected = load_state_dict_into([self], state_dict)
        if strict and (missing or unexpected):
            raise RuntimeError(f"load_state_dict: missing keys {missing}, unexpected keys {unexpected}")
        return missing, unexpected

    def split_embedding_weights(self) -> List[torch.Tensor]:
        return self.emb_module.split_embedding_weights()

    def named_split_embedding_weights(self, prefix: str = "") -> Iterator[Tuple[str, torch.Tensor]]:
        for t, w in zip(self._config.embedding_tables, self.emb_module.split_embedding_weights()):
            yield f"{prefix}{t.name}.weight" if not prefix or prefix.endswith(".") else f"{prefix}.{t.name}.weight", w

    def flush(self) -> None:
        self.emb_module.flush()

    @property
    def config(self) -> GroupedEmbeddingConfig:
        return self._config


class BatchedFusedEmbeddingBag(BaseBatchedEmbeddingBag, FusedOptimizerModule):
    _sequence = False  # BatchedFusedEmbedding: PoolingMode.NONE, output [sum lengths, D]

    def __init__(self, config: GroupedEmbeddingConfig, pg: Optional[dist.ProcessGroup] = None,
                 device: Optional[torch.device] = None, fused_params: Optional[Dict[str, Any]] = None) -> None:
        super().__init__(config, pg, device)
        if self._sequence:
            self._pooling = PoolingMode.NONE
        if device is None or device.type != "cuda":
            raise RuntimeError(
                "BatchedFusedEmbeddingBag runs on the sm_100a kernels only; got device "
                f"{device}. There is no CPU embedding kernel in this package.")
        fused_params = copy.copy(fused_params) if fused_params else {}
        self._emb_module = SplitTableBatchedEmbeddingBagsCodegen(
            embedding_specs=[(r, c, EmbeddingLocation.DEVICE, ComputeDevice.CUDA)
                             for r, c in zip(self._local_rows, self._local_cols)],
            feature_table_map=self._feature_table_map, pooling_mode=self._pooling,
            weights_precision=data_type_to_sparse_type(config.data_type), device=device, **fused_params)
        self._optim = EmbeddingFusedOptimizer(config, self._emb_module, pg)
        self.init_parameters()

    @property
    def emb_module(self) -> SplitTableBatchedEmbeddingBagsCodegen:
        return self._emb_module

    @property
    def fused_optimizer(self) -> FusedOptimizer:
        return self._optim

    def named_parameters(self, prefix: str = "", recurse: bool = True):
        # fused tables are not nn.Parameters of the dense optimizer (reference :655-664 yields nothing)
        yield from ()

    def named_buffers(self, prefix: str = "", recurse: bool = True):
        for t, w in zip(self._config.embedding_tables, self.emb_module.split_embedding_weights()):
            yield (f"{prefix}.{t.name}.weight" if prefix else f"{t.name}.weight"), w


class BatchedFusedEmbedding(BatchedFusedEmbeddingBag):
    """Non-pooled (sequence) twin (reference BatchedFusedEmbedding :459-496): same buffers, fused optimizer
    and state_dict; the lookup returns one [D] row per id."""
    _sequence = True

    def forward(self, features: KeyedJaggedTensor) -> torch.Tensor:
        return self.emb_module(indices=features.values(), offsets=features.offsets())
