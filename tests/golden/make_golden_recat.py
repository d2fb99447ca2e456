"""Generates tests/golden/recat.npz by executing the REFERENCE's own `_get_recat` and `_split_lengths`
(torchrec/distributed/dist_data.py:40-134; the two function definitions are compiled straight from the file
under /root/reference with `ast`, nothing else of that module is imported) with the reference's own CPU
`expand_into_jagged_permute` (oracle/_ref/libfbgemm_gpu_cpu_ref.so) - fixed and variable batch sizes.

Run here (the container with /root/reference):   python tests/golden/make_golden_recat.py"""
import ast
import itertools
import os
import sys
from typing import List, Optional

import numpy as np
import torch
from torch.autograd.profiler import record_function

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(HERE))
from golden_cases import RECAT_CASES, SPLIT_LENGTHS_CASES  # noqa: E402


def reference_functions(names):
    path = "/root/reference/torchrec/distributed/dist_data.py"
    tree = ast.parse(open(path).read(), filename=path)
    picked = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    assert {n.name for n in picked} == set(names)
    ns = {"torch": torch, "itertools": itertools, "List": List, "Optional": Optional,
          "record_function": record_function}
    exec(compile(ast.Module(body=picked, type_ignores=[]), path, "exec"), ns)
    return [ns[n] for n in names]


def main() -> None:
    assert "fbgemm_gpu" not in sys.modules
    torch.ops.load_library(os.path.join(ROOT, "oracle", "_ref", "libfbgemm_gpu_cpu_ref.so"))
    get_recat, split_lengths = reference_functions(["_get_recat", "_split_lengths"])
    out = {}
    for i, (local_split, num_splits, stagger, batch_sizes) in enumerate(RECAT_CASES):
        r = get_recat(local_split, num_splits, stagger, torch.device("cpu"), batch_sizes)
        out[f"recat{i}"] = r.numpy()
    for i, (splits, keys, offset_per_key) in enumerate(SPLIT_LENGTHS_CASES):
        out[f"split_lengths{i}"] = np.array(split_lengths(splits, keys, offset_per_key), dtype=np.int64)
    path = os.path.join(HERE, "recat.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
