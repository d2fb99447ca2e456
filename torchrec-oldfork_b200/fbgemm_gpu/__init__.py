"""fbgemm_gpu — drop-in module surface for torchrec, backed by libb200sparse.so (sm_100a).

Importing this package loads the CUDA extension (and fails loudly if it is missing, like the
reference's `torch.ops.load_library(fbgemm_gpu_py.so)`, fbgemm_gpu/__init__.py:11) and registers
the torch.ops.fbgemm.* operators.
"""
import b200_sparse  # noqa: F401  -- raises ImportError when the extension is not built

from . import _torch_ops, _tbe_autograd  # noqa: F401,E402
from . import split_embedding_configs, split_table_batched_embeddings_ops, permute_pooled_embedding_modules  # noqa: F401,E402
