"""Contracts of optimizers that run INSIDE the embedding backward (reference torchrec/optim/fused.py:15-45).

b200_tbe_backward applies SGD / rowwise Adagrad to a row the moment its gradient has been reduced, so by
the time autograd returns there is nothing left for an optimizer object to do: a FusedOptimizer exists to
(1) carry the learning rate into the kernel's arguments before the next backward (its step / zero_grad),
(2) expose the tables and optimizer state by name for checkpoints (the KeyedOptimizer part), and
(3) let DistributedModelParallel collect every sharded module's optimizer through FusedOptimizerModule."""
import abc
from typing import Any

import torch

from torchrec.optim.keyed import KeyedOptimizer


class FusedOptimizer(KeyedOptimizer, abc.ABC):
    """step() and zero_grad() must not touch parameters; implementations push hyper-parameters only."""

    @abc.abstractmethod
    def step(self, closure: Any = None) -> None:
        raise NotImplementedError

    @abc.abstractmethod
    def zero_grad(self, set_to_none: bool = False) -> None:
        raise NotImplementedError

    def __repr__(self) -> str:  # KeyedOptimizer skips Optimizer.__init__, the stock repr still works
        return torch.optim.Optimizer.__repr__(self)


class FusedOptimizerModule(abc.ABC):
    """A module whose parameters are updated by its own backward."""

    @property
    @abc.abstractmethod
    def fused_optimizer(self) -> KeyedOptimizer:
        raise NotImplementedError
