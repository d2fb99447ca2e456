"""RandomRecDataset (reference torchrec/datasets/random.py:20-208): Criteo-shaped synthetic
batches — per key `randint(high=hash_size, size=(ids_per_feature * B,))` int64 ids, constant
int32 lengths, randn dense features, randint(0, 2) labels, same generator call order."""
from typing import Iterator, List, Optional

import torch
from torch.utils.data.dataset import IterableDataset

from torchrec.datasets.utils import Batch
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor


class _RandomRecBatch:
    def __init__(self, keys: List[str], batch_size: int, hash_sizes: List[int], ids_per_features: List[int],
                 num_dense: int, manual_seed: Optional[int] = None, num_generated_batches: int = 10,
                 num_batches: Optional[int] = None) -> None:
        self.keys, self.batch_size = keys, batch_size
        self.hash_sizes, self.ids_per_features = hash_sizes, ids_per_features
        self.num_dense, self.num_batches = num_dense, num_batches
        self.num_generated_batches = num_generated_batches
        self.generator = None
        if manual_seed is not None:
            self.generator = torch.Generator()
            self.generator.manual_seed(manual_seed)
        # the reference caches ONE generated batch num_generated_batches times (random.py:53-55)
        self._generated_batches: List[Batch] = [self._generate_batch()] * num_generated_batches
        self.batch_index = 0

    def __iter__(self) -> "_RandomRecBatch":
        self.batch_index = 0
        return self

    def __next__(self) -> Batch:
        if self.batch_index == self.num_batches:
            raise StopIteration
        if self.num_generated_batches >= 0:
            batch = self._generated_batches[self.batch_index % len(self._generated_batches)]
        else:
            batch = self._generate_batch()
        self.batch_index += 1
        return batch

    def _generate_batch(self) -> Batch:
        B = self.batch_size
        values = [torch.randint(high=h, size=(n * B,), generator=self.generator)
                  for h, n in zip(self.hash_sizes, self.ids_per_features)]
        lengths = torch.tensor([n for n in self.ids_per_features for _ in range(B)], dtype=torch.int32)
        sparse = KeyedJaggedTensor(
            keys=self.keys, values=torch.cat(values), lengths=lengths, stride=B,
            length_per_key=[n * B for n in self.ids_per_features])  # known on the host: no sync needed
        dense = torch.randn(B, self.num_dense, generator=self.generator)
        labels = torch.randint(low=0, high=2, size=(B,), generator=self.generator)
        return Batch(dense_features=dense, sparse_features=sparse, labels=labels)


class RandomRecDataset(IterableDataset):
    def __init__(self, keys: List[str], batch_size: int, hash_size: Optional[int] = 100,
                 hash_sizes: Optional[List[int]] = None, ids_per_feature: Optional[int] = 2,
                 ids_per_features: Optional[List[int]] = None, num_dense: int = 50,
                 manual_seed: Optional[int] = None, num_batches: Optional[int] = None,
                 num_generated_batches: int = 10) -> None:
        super().__init__()
        if hash_sizes is None:
            hash_sizes = [hash_size] * len(keys)
        if ids_per_features is None:
            ids_per_features = [ids_per_feature] * len(keys)
        assert len(hash_sizes) == len(keys), "length of hash_sizes must be equal to the number of keys"
        assert len(ids_per_features) == len(keys), "length of ids_per_features must be equal to the number of keys"
        self.batch_generator = _RandomRecBatch(keys=keys, batch_size=batch_size, hash_sizes=hash_sizes,
                                               ids_per_features=ids_per_features, num_dense=num_dense,
                                               manual_seed=manual_seed, num_batches=num_batches,
                                               num_generated_batches=num_generated_batches)

    def __iter__(self) -> Iterator[Batch]:
        return iter(self.batch_generator)
