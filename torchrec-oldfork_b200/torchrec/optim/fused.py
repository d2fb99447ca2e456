"""Fused-optimizer contracts (reference torchrec/optim/fused.py:15-45): the update happens inside
the lookup's backward, so step()/zero_grad() of a FusedOptimizer do no tensor work."""
import abc
from typing import Any

from torch import optim

from torchrec.optim.keyed import KeyedOptimizer


class FusedOptimizer(KeyedOptimizer, abc.ABC):
    @abc.abstractmethod
    def step(self, closure: Any = None) -> None:
        ...

    @abc.abstractmethod
    def zero_grad(self, set_to_none: bool = False) -> None:
        ...

    def __repr__(self) -> str:
        return optim.Optimizer.__repr__(self)


class FusedOptimizerModule(abc.ABC):
    @property
    @abc.abstractmethod
    def fused_optimizer(self) -> KeyedOptimizer:
        ...
