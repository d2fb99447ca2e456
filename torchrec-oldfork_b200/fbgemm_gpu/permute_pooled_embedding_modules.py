"""fbgemm_gpu.permute_pooled_embedding_modules — imported by torchrec's column-wise sharding
(torchrec/distributed/sharding/cw_sharding.py:12).  Column-wise sharding is outside the hot
path this package accelerates; the class exists so imports resolve and fails
loudly when constructed."""
from typing import List, Optional

import torch


class PermutePooledEmbeddings:
    def __init__(self, embs_dims: List[int], permute: List[int], device: Optional[torch.device] = None) -> None:
        raise NotImplementedError(
            "PermutePooledEmbeddings (column-wise sharding) is out of scope of the B200 sparse path; "
            "use table_wise or row_wise sharding"
        )
