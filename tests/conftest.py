"""pytest configuration: import paths, the `gpu` marker, shared generators."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "torchrec-oldfork_b200")
for p in (PKG, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    import torch

    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container; GPU tests run under gpurun")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


def make_jagged(rng, T, B, max_L, rows, idx_dtype=np.int64, empty_frac=0.2):
    """Random table-major jagged batch: lengths [T*B], offsets [T*B+1], indices [N]."""
    lengths = rng.integers(0, max_L + 1, size=T * B)
    lengths[rng.random(T * B) < empty_frac] = 0
    offsets = np.zeros(T * B + 1, dtype=idx_dtype)
    np.cumsum(lengths, out=offsets[1:])
    rows = np.broadcast_to(np.asarray(rows), (T,))
    idx = np.concatenate([rng.integers(0, rows[t], size=int(lengths[t * B:(t + 1) * B].sum())) for t in range(T)]
                         ) if T else np.zeros(0)
    return lengths.astype(idx_dtype), offsets, idx.astype(idx_dtype)
