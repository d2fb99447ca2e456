from torchrec.distributed.model_parallel import DistributedModelParallel  # noqa: F401
from torchrec.distributed.types import (  # noqa: F401
    Awaitable, LazyAwaitable, ParameterSharding, ShardingEnv, ShardingPlan, ShardingType)
