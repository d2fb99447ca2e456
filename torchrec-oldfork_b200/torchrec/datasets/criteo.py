"""Criteo column names (reference torchrec/datasets/criteo.py:29-43) and the binary (npy) datapipe
the DLRM example trains from (:256-400, :638-842)."""
import os
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np
import torch
from torch.utils.data import IterableDataset

from torchrec.datasets.utils import Batch
from torchrec.sparse.jagged_tensor import KeyedJaggedTensor

INT_FEATURE_COUNT = 13
CAT_FEATURE_COUNT = 26
DEFAULT_LABEL_NAME = "label"
DEFAULT_INT_NAMES = [f"int_{i}" for i in range(INT_FEATURE_COUNT)]
DEFAULT_CAT_NAMES = [f"cat_{i}" for i in range(CAT_FEATURE_COUNT)]


# ---------------------------------------------------------------------------------------------
# Binary (npy) Criteo: the on-disk format the MLPerf DLRM run trains from (reference
# torchrec/datasets/criteo.py: BinaryCriteoUtils.get_shape_from_npy :256, get_file_idx_to_row_range
# :277, load_npy_range :338; InMemoryBinaryCriteoIterDataPipe :638-842).  Per day three files:
# `day_D_dense.npy` float32 [rows, 13] (already log(x+3)), `day_D_sparse.npy` int32 [rows, 26],
# `day_D_labels.npy` int32 [rows, 1].  Every rank owns one contiguous run of total_rows // world
# rows of the day files laid end to end and serves floor(own_rows / batch_size) batches.
#
# Difference in mechanism, not in results: the reference grows a list of slices with
# np.concatenate and then transposes; here each batch is assembled by writing the slices straight
# into one staging block (pinned when CUDA is present, so Batch.to(device, non_blocking=True)
# is a real asynchronous copy) with the ids stored key-major in the same pass.
# ---------------------------------------------------------------------------------------------

class BinaryCriteoUtils:
    @staticmethod
    def _read_header(fin) -> Tuple[Tuple[int, ...], np.dtype]:
        version = np.lib.format.read_magic(fin)
        if version == (1, 0):
            shape, _order, dtype = np.lib.format.read_array_header_1_0(fin)
        else:
            shape, _order, dtype = np.lib.format.read_array_header_2_0(fin)
        return shape, dtype

    @staticmethod
    def get_shape_from_npy(path: str, path_manager_key: str = "torchrec") -> Tuple[int, ...]:
        """Shape of an npy file from its header only."""
        with open(path, "rb") as fin:
            return BinaryCriteoUtils._read_header(fin)[0]

    @staticmethod
    def get_file_idx_to_row_range(lengths: List[int], rank: int, world_size: int) -> Dict[int, Tuple[int, int]]:
        """{file index: (first row, last row)} (both inclusive) of the rows `rank` owns when the files
        are laid end to end and every rank takes sum(lengths) // world_size consecutive rows."""
        per_rank = sum(lengths) // world_size
        lo, hi = rank * per_rank, (rank + 1) * per_rank - 1
        out: Dict[int, Tuple[int, int]] = {}
        start = 0
        for idx, n in enumerate(lengths):
            a, b = max(lo, start), min(hi, start + n - 1)
            if a <= b:
                out[idx] = (a - start, b - start)
            start += n
        return out

    @staticmethod
    def load_npy_range(fname: str, start_row: int, num_rows: int, path_manager_key: str = "torchrec",
                       mmap_mode: bool = False) -> np.ndarray:
        """Rows [start_row, start_row + num_rows) of a 2-D npy file without reading the rest."""
        with open(fname, "rb") as fin:
            shape, dtype = BinaryCriteoUtils._read_header(fin)
            if len(shape) != 2:
                raise ValueError("Cannot load range for npy with ndim != 2.")
            total_rows, row_size = shape
            if not 0 <= start_row < total_rows:
                raise ValueError(f"start_row ({start_row}) is out of bounds. It must be between 0 "
                                 f"and {total_rows - 1}, inclusive.")
            if start_row + num_rows > total_rows:
                raise ValueError(f"num_rows ({num_rows}) exceeds number of available rows "
                                 f"({total_rows}) for the given start_row ({start_row}).")
            if mmap_mode:
                return np.load(fname, mmap_mode="r")[start_row:start_row + num_rows]
            fin.seek(start_row * row_size * dtype.itemsize, os.SEEK_CUR)
            data = np.fromfile(fin, dtype=dtype, count=num_rows * row_size)
        return data.reshape((num_rows, row_size))


class InMemoryBinaryCriteoIterDataPipe(IterableDataset):
    """Batches of the binary Criteo files, this rank's rows held in memory (or memory-mapped).

    Same arguments and batches as the reference class: batch b holds rows [b*batch_size, (b+1)*batch_size)
    of the rank's run; dense [B, 13] as stored, sparse ids key-major in a KeyedJaggedTensor with one id per
    (feature, sample) (`% hashes[f]` when hashes are given), labels [B].  shuffle_batches permutes the rows
    inside each batch."""

    def __init__(self, dense_paths: List[str], sparse_paths: List[str], labels_paths: List[str], batch_size: int,
                 rank: int, world_size: int, shuffle_batches: bool = False, mmap_mode: bool = False,
                 hashes: Optional[List[int]] = None, path_manager_key: str = "torchrec") -> None:
        assert len(dense_paths) == len(sparse_paths) == len(labels_paths)
        self.dense_paths, self.sparse_paths, self.labels_paths = dense_paths, sparse_paths, labels_paths
        self.batch_size, self.rank, self.world_size = batch_size, rank, world_size
        self.shuffle_batches, self.mmap_mode, self.hashes = shuffle_batches, mmap_mode, hashes
        self.path_manager_key = path_manager_key
        self._hashes_np = None if hashes is None else np.asarray(hashes).reshape((1, CAT_FEATURE_COUNT))
        self._load_data_for_rank()
        self.num_rows_per_file: List[int] = [a.shape[0] for a in self.dense_arrs]
        self.num_batches: int = sum(self.num_rows_per_file) // batch_size
        # identical for every batch: built once, shared by all KeyedJaggedTensors
        n_ids = CAT_FEATURE_COUNT * batch_size
        self.keys: List[str] = DEFAULT_CAT_NAMES
        self.lengths = torch.ones((n_ids,), dtype=torch.int32)
        self.offsets = torch.arange(0, n_ids + 1, dtype=torch.int32)
        self.stride = batch_size
        self.length_per_key: List[int] = CAT_FEATURE_COUNT * [batch_size]
        self.offset_per_key: List[int] = [batch_size * i for i in range(CAT_FEATURE_COUNT + 1)]
        self.index_per_key: Dict[str, int] = {k: i for i, k in enumerate(self.keys)}
        self._pin = torch.cuda.is_available()

    def _load_data_for_rank(self) -> None:
        ranges = BinaryCriteoUtils.get_file_idx_to_row_range(
            [BinaryCriteoUtils.get_shape_from_npy(p)[0] for p in self.dense_paths], self.rank, self.world_size)
        self.dense_arrs, self.sparse_arrs, self.labels_arrs = [], [], []
        for arrs, paths in ((self.dense_arrs, self.dense_paths), (self.sparse_arrs, self.sparse_paths),
                            (self.labels_arrs, self.labels_paths)):
            for idx, (first, last) in ranges.items():
                arrs.append(BinaryCriteoUtils.load_npy_range(paths[idx], first, last - first + 1,
                                                             mmap_mode=self.mmap_mode))
        if not self.mmap_mode and self._hashes_np is not None:  # memory-mapped files are hashed per batch
            for a in self.sparse_arrs:
                a %= self._hashes_np.astype(a.dtype)

    def _empty(self, shape, dtype: np.dtype) -> torch.Tensor:
        t = torch.empty(shape, dtype=torch.from_numpy(np.empty(0, dtype=dtype)).dtype)
        return t.pin_memory() if self._pin else t

    def __iter__(self) -> Iterator[Batch]:
        B = self.batch_size
        if self.num_batches == 0:
            return
        d_dt, s_dt, l_dt = self.dense_arrs[0].dtype, self.sparse_arrs[0].dtype, self.labels_arrs[0].dtype
        file_idx, row_idx = 0, 0
        for _ in range(self.num_batches):
            dense = self._empty((B, INT_FEATURE_COUNT), d_dt)
            ids = self._empty((CAT_FEATURE_COUNT, B), s_dt)  # key-major: values = ids.view(-1)
            labels = self._empty((B,), l_dt)
            dn, sn, ln = dense.numpy(), ids.numpy(), labels.numpy()
            perm = np.random.permutation(B) if self.shuffle_batches else None
            filled = 0
            while filled < B:
                take = min(B - filled, self.num_rows_per_file[file_idx] - row_idx)
                rows = slice(row_idx, row_idx + take)
                sp = self.sparse_arrs[file_idx][rows, :]
                if self.mmap_mode and self._hashes_np is not None:
                    sp = sp % self._hashes_np.astype(sp.dtype)
                if perm is None:
                    dn[filled:filled + take] = self.dense_arrs[file_idx][rows, :]
                    sn[:, filled:filled + take] = sp.T
                    ln[filled:filled + take] = self.labels_arrs[file_idx][rows, 0]
                else:  # output row j holds gathered row perm[j]: gathered row i lands at inv[i]
                    if filled == 0:
                        inv = np.empty(B, dtype=np.int64)
                        inv[perm] = np.arange(B)
                    dst = inv[filled:filled + take]
                    dn[dst] = self.dense_arrs[file_idx][rows, :]
                    sn[:, dst] = sp.T
                    ln[dst] = self.labels_arrs[file_idx][rows, 0]
                filled += take
                row_idx += take
                if row_idx >= self.num_rows_per_file[file_idx]:
                    file_idx, row_idx = file_idx + 1, 0
            yield Batch(
                dense_features=dense,
                sparse_features=KeyedJaggedTensor(
                    keys=self.keys, values=ids.view(-1), lengths=self.lengths, offsets=self.offsets,
                    stride=self.stride, length_per_key=self.length_per_key, offset_per_key=self.offset_per_key,
                    index_per_key=self.index_per_key),
                labels=labels)

    def __len__(self) -> int:
        return self.num_batches
