from torchrec.optim.fused import FusedOptimizer, FusedOptimizerModule  # noqa: F401
from torchrec.optim.keyed import CombinedOptimizer, KeyedOptimizer, KeyedOptimizerWrapper  # noqa: F401
