"""RandomRecDataset (reference torchrec/datasets/random.py:20-208): Criteo-shaped synthetic batches.

Per batch, in this order on one torch.Generator (so a seed yields the reference's data): for every key
`randint(high=hash_size, size=(ids_per_feature * B,))` int64 ids; then `randn(B, num_dense)`; then
`randint(0, 2, (B,))` labels.  Lengths are the constant ids_per_feature (int32), so the KJT's per-key
totals are known on the host and no device sync is ever needed to read them.

Like the reference (:53-55) the dataset generates ONE batch up front and serves it
`num_generated_batches` times per cycle; a negative num_generated_batches draws a fresh batch each step.
"""
from itertools import count
from typing import Iterator, List, Optional

import torch
from torch.utils.data.dataset import IterableDataset

from torchrec.datasets.utils import Batch
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor


class RandomRecDataset(IterableDataset):
    def __init__(self, keys: List[str], batch_size: int, hash_size: Optional[int] = 100,
                 hash_sizes: Optional[List[int]] = None, ids_per_feature: Optional[int] = 2,
                 ids_per_features: Optional[List[int]] = None, num_dense: int = 50,
                 manual_seed: Optional[int] = None, num_batches: Optional[int] = None,
                 num_generated_batches: int = 10) -> None:
        super().__init__()
        n_keys = len(keys)
        self.keys, self.batch_size, self.num_dense = list(keys), batch_size, num_dense
        self.hash_sizes = list(hash_sizes) if hash_sizes is not None else [hash_size] * n_keys
        self.ids_per_features = (list(ids_per_features) if ids_per_features is not None
                                 else [ids_per_feature] * n_keys)
        assert len(self.hash_sizes) == n_keys, "length of hash_sizes must be equal to the number of keys"
        assert len(self.ids_per_features) == n_keys, "length of ids_per_features must be equal to the number of keys"
        self.num_batches, self.num_generated_batches = num_batches, num_generated_batches
        self.generator: Optional[torch.Generator] = None
        if manual_seed is not None:
            self.generator = torch.Generator().manual_seed(manual_seed)
        B = batch_size
        self._lengths = torch.tensor(self.ids_per_features, dtype=torch.int32).repeat_interleave(B)
        self._length_per_key = [n * B for n in self.ids_per_features]
        first = self.make_batch()  # drawn even when it is not served (the reference does the same, :53-55)
        self._cached: Optional[Batch] = first if num_generated_batches >= 0 else None

    def make_batch(self) -> Batch:
        """One batch from the generator's current state (call order documented above)."""
        B, g = self.batch_size, self.generator
        ids = torch.cat([torch.randint(high=h, size=(n * B,), generator=g)
                         for h, n in zip(self.hash_sizes, self.ids_per_features)])
        dense = torch.randn(B, self.num_dense, generator=g)
        labels = torch.randint(low=0, high=2, size=(B,), generator=g)
        sparse = KeyedJaggedTensor(keys=self.keys, values=ids, lengths=self._lengths, stride=B,
                                   length_per_key=self._length_per_key)
        return Batch(dense_features=dense, sparse_features=sparse, labels=labels)

    def __iter__(self) -> Iterator[Batch]:
        steps = count() if self.num_batches is None else range(self.num_batches)
        for _ in steps:
            yield self._cached if self._cached is not None else self.make_batch()
