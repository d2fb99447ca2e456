"""Perceptron / MLP (reference torchrec/modules/mlp.py).  Dense layers stay on cuBLAS through
PyTorch (BASELINE.json north_star); nothing custom here."""
from typing import Callable, List, Optional, Union

import torch
from torch import nn


class Perceptron(nn.Module):
    def __init__(self, in_size: int, out_size: int, bias: bool = True,
                 activation: Union[nn.Module, Callable[[torch.Tensor], torch.Tensor]] = torch.relu,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        self._out_size, self._in_size = out_size, in_size
        self._linear = nn.Linear(in_size, out_size, bias=bias, device=device)
        self._activation_fn = activation

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return self._activation_fn(self._linear(input))


class MLP(nn.Module):
    def __init__(self, in_size: int, layer_sizes: List[int], bias: bool = True,
                 activation: Union[str, Callable[[], nn.Module], nn.Module,
                                   Callable[[torch.Tensor], torch.Tensor]] = torch.relu,
                 device: Optional[torch.device] = None) -> None:
        super().__init__()
        if activation == "relu":
            activation = torch.relu
        elif activation == "sigmoid":
            activation = torch.sigmoid
        if not callable(activation) and not isinstance(activation, nn.Module):
            raise ValueError("activation must be 'relu', 'sigmoid', a callable or an nn.Module")
        sizes = [in_size] + list(layer_sizes)
        self._mlp = nn.Sequential(*[
            Perceptron(sizes[i], sizes[i + 1], bias=bias, activation=activation, device=device)
            for i in range(len(layer_sizes))])

    def forward(self, input: torch.Tensor) -> torch.Tensor:
        return self._mlp(input)
