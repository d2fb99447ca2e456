"""Generates tests/golden/criteo_binary.npz by running the REFERENCE's own
InMemoryBinaryCriteoIterDataPipe (torchrec/datasets/criteo.py:638, loaded from /root/reference by
file path; `iopath` / `pyre_extensions` are replaced by the few-line stand-ins below) on seeded npy
day files, for every rank of several (rows per file, batch size, world size) cases - the two of
torchrec/datasets/tests/test_criteo.py:399-404 among them.

Run here (the container with /root/reference):   python tests/golden/make_golden_criteo.py
The input day files are regenerated from the seed by tests/test_host_logic.py; the fixture holds only
what the reference produced, so the tests never need /root/reference."""
import importlib.util
import os
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from golden_cases import CRITEO_CASES, write_criteo_days  # noqa: E402

REF = "/root/reference"


def _stub(name: str, **attrs) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    m.__path__ = []  # a package, so sub-imports resolve through sys.modules
    sys.modules[name] = m
    return m


def _load(name: str, rel: str) -> types.ModuleType:
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    m = importlib.util.module_from_spec(spec)
    sys.modules[name] = m
    spec.loader.exec_module(m)
    return m


def load_reference_criteo() -> types.ModuleType:
    import typing

    class PathManager:
        def open(self, path, mode="r", **kw):
            return open(path, mode)

    class PathManagerFactory:
        def get(self, key=None):
            return PathManager()

    def none_throws(x, msg=""):
        assert x is not None, msg
        return x

    _stub("iopath")
    _stub("iopath.common")
    _stub("iopath.common.file_io", PathManager=PathManager, PathManagerFactory=PathManagerFactory)
    _stub("pyre_extensions", none_throws=none_throws, ParameterSpecification=typing.ParamSpec)
    for pkg in ("torchrec", "torchrec.datasets", "torchrec.sparse"):
        _stub(pkg)
    _load("torchrec.streamable", "torchrec/streamable.py")
    _load("torchrec.sparse.jagged_tensor", "torchrec/sparse/jagged_tensor.py")
    _load("torchrec.datasets.utils", "torchrec/datasets/utils.py")
    return _load("torchrec.datasets.criteo", "torchrec/datasets/criteo.py")


def main() -> None:
    ref = load_reference_criteo()
    out = {}
    for ci, (rows_per_file, batch_size, world, use_hashes, seed) in enumerate(CRITEO_CASES):
        with tempfile.TemporaryDirectory() as tmp:
            dense, sparse, labels = write_criteo_days(tmp, rows_per_file, seed)
            hashes = [3 * i + 2 for i in range(ref.CAT_FEATURE_COUNT)] if use_hashes else None
            for rank in range(world):
                pipe = ref.InMemoryBinaryCriteoIterDataPipe(
                    dense_paths=dense, sparse_paths=sparse, labels_paths=labels, batch_size=batch_size,
                    rank=rank, world_size=world, hashes=hashes)
                batches = list(pipe)
                assert len(batches) == len(pipe)
                pre = f"case{ci}/rank{rank}"
                out[f"{pre}/num_batches"] = np.int64(len(batches))
                if batches:
                    out[f"{pre}/dense"] = np.stack([b.dense_features.numpy() for b in batches])
                    out[f"{pre}/values"] = np.stack([b.sparse_features.values().numpy() for b in batches])
                    out[f"{pre}/lengths"] = batches[0].sparse_features.lengths().numpy()
                    out[f"{pre}/offsets"] = batches[0].sparse_features.offsets().numpy()
                    out[f"{pre}/labels"] = np.stack([b.labels.numpy() for b in batches])
                    out[f"{pre}/keys"] = np.array(batches[0].sparse_features.keys())
    # BinaryCriteoUtils known answers of test_criteo.py:164-177
    for r in range(3):
        got = ref.BinaryCriteoUtils.get_file_idx_to_row_range(lengths=[14, 17, 20], rank=r, world_size=3)
        out[f"row_range/rank{r}"] = np.array([[k, a, b] for k, (a, b) in sorted(got.items())], dtype=np.int64)
    path = os.path.join(HERE, "criteo_binary.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
