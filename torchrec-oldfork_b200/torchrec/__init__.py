"""torchrec API mirror for the B200 sparse hot path.

Same module paths, class names and call signatures as samiwilf/torchrec-oldfork for the path
SURVEY.md §8 scopes (KeyedJaggedTensor in, KeyedTensor out, EmbeddingBagConfig,
EmbeddingBagCollection, DLRM, DistributedModelParallel with table-wise / row-wise sharding and
the fused rowwise-Adagrad / SGD optimizer hooks); everything below it runs on the sm_100a
kernels of libb200sparse.so through the fbgemm_gpu boundary.
"""
import fbgemm_gpu  # noqa: F401  -- loads the CUDA extension and registers torch.ops.fbgemm.*

from torchrec.sparse.jagged_tensor import JaggedTensor, KeyedJaggedTensor, KeyedTensor  # noqa: F401
from torchrec.modules.embedding_configs import (  # noqa: F401
    DataType, EmbeddingBagConfig, EmbeddingConfig, PoolingType)
from torchrec.modules.embedding_modules import EmbeddingBagCollection  # noqa: F401
