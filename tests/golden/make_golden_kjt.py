"""Generates tests/golden/kjt_ops.npz by running the REFERENCE's own KeyedJaggedTensor / KeyedTensor
(torchrec/sparse/jagged_tensor.py, loaded from /root/reference by file path) over the reference's own CPU
kernels (oracle/_ref/libfbgemm_gpu_cpu_ref.so: sparse_ops_cpu.cpp compiled unmodified) on seeded inputs:
offsets / length_per_key / offset_per_key, permute (with duplicates and drops), split, to_dict and
KeyedTensor.regroup.

Run here (the container with /root/reference):   python tests/golden/make_golden_kjt.py
Own process: the .so registers the same fbgemm::* schemas the product does."""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
from golden_cases import KJT_CASES, kjt_case_inputs  # noqa: E402
from make_golden_criteo import _load, _stub  # noqa: E402


def load_reference_kjt():
    import typing
    assert "fbgemm_gpu" not in sys.modules
    torch.ops.load_library(os.path.join(ROOT, "oracle", "_ref", "libfbgemm_gpu_cpu_ref.so"))
    _stub("pyre_extensions", none_throws=lambda x, m="": x, ParameterSpecification=typing.ParamSpec)
    for pkg in ("torchrec", "torchrec.sparse"):
        _stub(pkg)
    _load("torchrec.streamable", "torchrec/streamable.py")
    return _load("torchrec.sparse.jagged_tensor", "torchrec/sparse/jagged_tensor.py")


def main() -> None:
    jt = load_reference_kjt()
    out = {}
    for ci, case in enumerate(KJT_CASES):
        keys, lengths, values, weights, stride = kjt_case_inputs(case)
        k = jt.KeyedJaggedTensor(keys=keys, values=torch.from_numpy(values), lengths=torch.from_numpy(lengths),
                                 weights=None if weights is None else torch.from_numpy(weights), stride=stride)
        pre = f"case{ci}"
        out[f"{pre}/offsets"] = k.offsets().numpy()
        out[f"{pre}/length_per_key"] = np.array(k.length_per_key(), dtype=np.int64)
        out[f"{pre}/offset_per_key"] = np.array(k.offset_per_key(), dtype=np.int64)
        for pi, perm in enumerate(case["permutes"]):
            p = k.permute(perm, torch.tensor(perm, dtype=torch.int32))
            out[f"{pre}/permute{pi}/keys"] = np.array(p.keys())
            out[f"{pre}/permute{pi}/values"] = p.values().numpy()
            out[f"{pre}/permute{pi}/lengths"] = p.lengths().numpy()
            out[f"{pre}/permute{pi}/length_per_key"] = np.array(p.length_per_key(), dtype=np.int64)
            if weights is not None:
                out[f"{pre}/permute{pi}/weights"] = p.weights().numpy()
        for si, seg in enumerate(case["splits"]):
            for part_i, part in enumerate(k.split(seg)):
                q = f"{pre}/split{si}/part{part_i}"
                out[f"{q}/keys"] = np.array(part.keys())
                out[f"{q}/values"] = part.values().numpy()
                out[f"{q}/lengths"] = part.lengths().numpy()
                out[f"{q}/stride"] = np.int64(part.stride())
                if weights is not None:
                    out[f"{q}/weights"] = part.weights().numpy()
        for key, j in k.to_dict().items():
            out[f"{pre}/dict/{key}/values"] = j.values().numpy()
            out[f"{pre}/dict/{key}/lengths"] = j.lengths().numpy()
            out[f"{pre}/dict/{key}/offsets"] = j.offsets().numpy()
    # KeyedTensor.regroup (jagged_tensor.py:1172-1209)
    rng = np.random.default_rng(77)
    kts, groups = [], [["b", "e"], ["a", "d", "c"], ["e", "a"]]
    for names, dims in ((["a", "b", "c"], [3, 1, 4]), (["d", "e"], [2, 5])):
        vals = rng.standard_normal((6, sum(dims))).astype(np.float32)
        out[f"regroup/in_{names[0]}"] = vals
        kts.append(jt.KeyedTensor(keys=names, length_per_key=dims, values=torch.from_numpy(vals)))
    for gi, t in enumerate(jt.KeyedTensor.regroup(kts, groups)):
        out[f"regroup/out{gi}"] = t.numpy()
    path = os.path.join(HERE, "kjt_ops.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
