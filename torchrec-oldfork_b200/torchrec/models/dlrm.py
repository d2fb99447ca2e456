"""DLRM (reference torchrec/models/dlrm.py: SparseArch :36, DenseArch :116, InteractionArch :155,
OverArch :222, DLRM :270) plus the DLRMTrain wrapper of examples/dlrm/modules/dlrm_train.py:58-83.

InteractionArch runs the fused sm_100a dot-interaction kernel on CUDA inputs
(b200_interaction_forward/backward) instead of cat + bmm + triu gather + cat; results follow
the reference layout [dense | triu(X X^T, offset=1) row-major].
"""
from typing import Dict, List, Optional, Tuple

import torch
from torch import nn

from b200_sparse import ops as _b200
from torchrec.datasets.utils import Batch
from torchrec.modules.embedding_modules import EmbeddingBagCollection
from torchrec.modules.mlp import MLP, mark_zero_padded
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor, KeyedTensor


def choose(n: int, k: int) -> int:
    if not 0 <= k <= n:
        return 0
    out = 1
    for i in range(1, min(k, n - k) + 1):
        out = out * (n - i + 1) // i
    return out


class SparseArch(nn.Module):
    """Pooled lookups for every sparse feature -> [B, F, D]."""

    def __init__(self, embedding_bag_collection: EmbeddingBagCollection) -> None:
        super().__init__()
        self.embedding_bag_collection = embedding_bag_collection
        cfgs = embedding_bag_collection.embedding_bag_configs
        assert cfgs, "Embedding bag collection cannot be empty!"
        self.D: int = cfgs[0].embedding_dim
        self._sparse_feature_names: List[str] = [n for c in cfgs for n in c.feature_names]
        self.F: int = len(self._sparse_feature_names)

    def forward(self, features: KeyedJaggedTensor) -> torch.Tensor:
        pooled = self.embedding_bag_collection(features)
        if not isinstance(pooled, KeyedTensor):  # sharded collection returns an awaitable
            pooled = pooled.wait()
        B = features.stride()
        if pooled.keys() == self._sparse_feature_names:
            # already in model order: the [B, F*D] buffer IS the [B, F, D] tensor
            return pooled.values().reshape(B, self.F, self.D)
        by_name: Dict[str, torch.Tensor] = pooled.to_dict()
        return torch.cat([by_name[n] for n in self._sparse_feature_names], dim=1).reshape(B, self.F, self.D)

    @property
    def sparse_feature_names(self) -> List[str]:
        return self._sparse_feature_names


class DenseArch(nn.Module):
    def __init__(self, in_features: int, layer_sizes: List[int], device: Optional[torch.device] = None) -> None:
        super().__init__()
        self.model: nn.Module = MLP(in_features, layer_sizes, bias=True, activation="relu", device=device)

    def forward(self, features: torch.Tensor) -> torch.Tensor:
        return self.model(features)


class _DotInteraction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dense: torch.Tensor, sparse: torch.Tensor) -> torch.Tensor:
        dense_c, sparse_c = dense.float().contiguous(), sparse.float().contiguous()
        ctx.save_for_backward(dense_c, sparse_c)
        # rows at a 16-byte pitch, zeros in the pad: the next Linear can run an aligned GEMM on it
        return _b200.interaction_forward(dense_c, sparse_c, pad_rows=True)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        dense_c, sparse_c = ctx.saved_tensors
        return _b200.interaction_backward(dense_c, sparse_c, grad_out)


class InteractionArch(nn.Module):
    """Pairwise dot products of the dense and sparse embeddings.

    dense [B, D], sparse [B, F, D] -> [B, D + F + choose(F, 2)]."""

    def __init__(self, num_sparse_features: int) -> None:
        super().__init__()
        self.F: int = num_sparse_features
        self.register_buffer("triu_indices", torch.triu_indices(self.F + 1, self.F + 1, offset=1), persistent=False)

    def forward(self, dense_features: torch.Tensor, sparse_features: torch.Tensor) -> torch.Tensor:
        if self.F <= 0:
            return dense_features
        if dense_features.is_cuda and _b200.interaction_supported(self.F, dense_features.shape[1]):
            return mark_zero_padded(_DotInteraction.apply(dense_features, sparse_features))
        # shapes the kernels do not cover (D % 4 != 0, or F * D beyond the shared-memory tile): the reference
        # formulation in plain torch
        combined = torch.cat((dense_features.unsqueeze(1), sparse_features), dim=1)
        inter = torch.bmm(combined, combined.transpose(1, 2))
        flat = inter[:, self.triu_indices[0], self.triu_indices[1]]
        return torch.cat((dense_features, flat), dim=1)


class OverArch(nn.Module):
    def __init__(self, in_features: int, layer_sizes: List[int], device: Optional[torch.device] = None) -> None:
        super().__init__()
        if len(layer_sizes) <= 1:
            raise ValueError("OverArch must have multiple layers.")
        self.model: nn.Module = nn.Sequential(
            MLP(in_features, layer_sizes[:-1], bias=True, activation="relu", device=device),
            nn.Linear(layer_sizes[-2], layer_sizes[-1], bias=True, device=device))

    def forward(self, features: torch.Tensor) -> torch.Tensor:
        return self.model(features)


class DLRM(nn.Module):
    def __init__(self, embedding_bag_collection: EmbeddingBagCollection, dense_in_features: int,
                 dense_arch_layer_sizes: List[int], over_arch_layer_sizes: List[int],
                 dense_device: Optional[torch.device] = None) -> None:
        super().__init__()
        cfgs = embedding_bag_collection.embedding_bag_configs
        assert len(cfgs) > 0, "At least one embedding bag is required"
        for c in cfgs[1:]:
            assert c.embedding_dim == cfgs[0].embedding_dim, "All EmbeddingBagConfigs must have the same dimension"
        D = cfgs[0].embedding_dim
        if dense_arch_layer_sizes[-1] != D:
            raise ValueError(
                f"embedding_bag_collection dimension ({D}) and final dense arch layer size "
                f"({dense_arch_layer_sizes[-1]}) must match.")
        self.sparse_arch = SparseArch(embedding_bag_collection)
        F = len(self.sparse_arch.sparse_feature_names)
        self.dense_arch = DenseArch(dense_in_features, dense_arch_layer_sizes, device=dense_device)
        self.inter_arch = InteractionArch(num_sparse_features=F)
        self.over_arch = OverArch(D + choose(F, 2) + F, over_arch_layer_sizes, device=dense_device)

    def forward(self, dense_features: torch.Tensor, sparse_features: KeyedJaggedTensor) -> torch.Tensor:
        embedded_dense = self.dense_arch(dense_features)
        embedded_sparse = self.sparse_arch(sparse_features)
        return self.over_arch(self.inter_arch(dense_features=embedded_dense, sparse_features=embedded_sparse))


class DLRMTrain(nn.Module):
    """DLRM + BCEWithLogits loss, returning (loss, (loss, logits, labels)) like the example trainer."""

    def __init__(self, embedding_bag_collection: EmbeddingBagCollection, dense_in_features: int,
                 dense_arch_layer_sizes: List[int], over_arch_layer_sizes: List[int],
                 dense_device: Optional[torch.device] = None) -> None:
        super().__init__()
        self.model = DLRM(embedding_bag_collection, dense_in_features, dense_arch_layer_sizes,
                          over_arch_layer_sizes, dense_device)
        self.loss_fn: nn.Module = nn.BCEWithLogitsLoss()

    def forward(self, batch: Batch):
        logits = self.model(batch.dense_features, batch.sparse_features).squeeze()
        loss = self.loss_fn(logits, batch.labels.float())
        return loss, (loss.detach(), logits.detach(), batch.labels.detach())
