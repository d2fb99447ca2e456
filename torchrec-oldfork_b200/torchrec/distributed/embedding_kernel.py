"""Checkpoint view of sharded tables (reference torchrec/distributed/embedding_kernel.py
get_state_dict :63-125; batched_embedding_kernel.py to_rowwise_sharded_metadata :71-125 and the
ShardedTensor construction :160-246; embedding_lookup.py _load_state_dict :50-74).

A sharded module's state_dict holds one `torch.distributed._shard.sharded_tensor.ShardedTensor` per
table: the local shard is a VIEW of the table-batched weight buffer (writing into it updates the
live table), the metadata lists every rank's (offset, size, placement), so the reference's
checkpoint tooling sees the same objects.  Optimizer state (`<table>.momentum1`, one float per
row) is sharded along the rows the same way.  Without a process group (unsharded / inference)
plain tensors are returned."""
from __future__ import annotations

from collections import OrderedDict
from typing import Any, Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
from torch.distributed._shard.metadata import ShardMetadata as TorchShardMetadata
from torch.distributed._shard.sharded_tensor import Shard, ShardedTensor, ShardedTensorMetadata, TensorProperties

from torchrec.distributed.embedding_types import ShardedEmbeddingTable
from torchrec.distributed.types import ShardMetadata


def _placement(meta: ShardMetadata, device: torch.device) -> str:
    """`rank:R/cuda:R` as planned, or `rank:R/cpu` when the tables live on the host (gloo tests)."""
    assert meta.placement is not None, "a sharded table needs a placement per shard"
    return meta.placement if device.type == "cuda" else meta.placement.split("/")[0] + "/cpu"


def _torch_meta(meta: ShardMetadata, device: torch.device, rowwise: bool) -> TorchShardMetadata:
    offs, sizes = list(meta.shard_offsets), list(meta.shard_sizes)
    if rowwise:
        offs, sizes = offs[:1], sizes[:1]
    return TorchShardMetadata(shard_offsets=offs, shard_sizes=sizes, placement=_placement(meta, device))


def sharded_tensor_from_local(local: torch.Tensor, local_metadata: ShardMetadata,
                              global_metadata: Sequence[ShardMetadata], size: Sequence[int],
                              pg: Optional[dist.ProcessGroup], rowwise: bool = False) -> ShardedTensor:
    """ShardedTensor of global shape `size` whose one local shard is `local` (no copy, no collective:
    every rank builds the same metadata from the sharding plan).  rowwise: 1-D per-row state
    (momentum1) of a table sharded along rows - only dim 0 of the table's metadata applies."""
    dev = local.device
    shards = [_torch_meta(m, dev, rowwise) for m in global_metadata]
    mine = _torch_meta(local_metadata, dev, rowwise)
    assert list(local.shape) == mine.shard_sizes, (tuple(local.shape), mine.shard_sizes)
    meta = ShardedTensorMetadata(
        shards_metadata=shards, size=torch.Size(size[:1] if rowwise else size),
        tensor_properties=TensorProperties(dtype=local.dtype, layout=local.layout,
                                           requires_grad=local.requires_grad,
                                           memory_format=torch.contiguous_format, pin_memory=False))
    return ShardedTensor._init_from_local_shards_and_global_metadata(
        local_shards=[Shard(local, mine)], sharded_tensor_metadata=meta, process_group=pg)


def table_sharded_tensor(table: ShardedEmbeddingTable, local: torch.Tensor,
                         pg: Optional[dist.ProcessGroup], rowwise: bool = False) -> ShardedTensor:
    assert table.local_metadata is not None and table.global_metadata is not None
    return sharded_tensor_from_local(local, table.local_metadata, table.global_metadata,
                                     [table.num_embeddings, table.embedding_dim], pg, rowwise)


def get_state_dict(embedding_tables: List[ShardedEmbeddingTable], params: Sequence[torch.Tensor],
                   pg: Optional[dist.ProcessGroup] = None, destination: Optional[Dict[str, Any]] = None,
                   prefix: str = "") -> Dict[str, Any]:
    """`<prefix><table>.weight` -> ShardedTensor (pg given and the table carries shard metadata) or
    the local tensor."""
    if destination is None:
        destination = OrderedDict()
    for table, param in zip(embedding_tables, params):
        assert table.local_rows == param.size(0), (table.name, table.local_rows, tuple(param.shape))
        key = f"{prefix}{table.name}.weight"
        if pg is not None and table.global_metadata is not None and table.local_metadata is not None:
            destination[key] = table_sharded_tensor(table, param, pg)
        else:
            destination[key] = param
    return destination


def local_tensor(t: Any) -> torch.Tensor:
    """The rank's own data of a state_dict entry: the single local shard of a ShardedTensor."""
    if isinstance(t, ShardedTensor):
        shards = t.local_shards()
        assert len(shards) == 1, f"expected one local shard, found {len(shards)}"
        return shards[0].tensor
    return t


@torch.no_grad()
def copy_state(dst: Any, src: Any) -> None:
    """dst <- src on the local shards (either side may be a ShardedTensor or a tensor of the local shape)."""
    d, s = local_tensor(dst), local_tensor(src)
    assert d.shape == s.shape, f"local shard shape {tuple(d.shape)} != source {tuple(s.shape)}"
    d.detach().copy_(s)


def load_state_dict_into(modules: Sequence[torch.nn.Module], state_dict: Dict[str, Any]
                         ) -> Tuple[List[str], List[str]]:
    """Copy matching entries into the modules' live tables; returns (missing, unexpected) keys."""
    missing: List[str] = []
    unexpected = list(state_dict.keys())
    for m in modules:
        for key, param in m.state_dict().items():
            if key in state_dict:
                copy_state(param, state_dict[key])
                unexpected.remove(key)
            else:
                missing.append(key)
    return missing, unexpected


def filter_state_dict(state_dict: Dict[str, Any], name: str) -> Dict[str, Any]:
    """Entries under `name.` with that prefix removed (reference torchrec/distributed/utils.py:45)."""
    out: Dict[str, Any] = OrderedDict()
    pre = name + "."
    for k, v in state_dict.items():
        if k.startswith(pre):
            out[k[len(pre):]] = v
    return out
