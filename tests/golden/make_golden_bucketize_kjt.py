"""Generates tests/golden/bucketize_kjt.npz by executing the REFERENCE's own
`bucketize_kjt_before_all2all` (torchrec/distributed/embedding_sharding.py:121-186; the function definition
is compiled straight from that file with `ast`) on the reference's own KeyedJaggedTensor class and its CPU
`block_bucketize_sparse_features` kernel (oracle/_ref/libfbgemm_gpu_cpu_ref.so), for the row-wise input dist:
plain, with the unbucketize permutation (sequence path), with positions, with weights.

Run here (the container with /root/reference):   python tests/golden/make_golden_bucketize_kjt.py"""
import ast
import os
import sys
from typing import Optional, Tuple

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
from golden_cases import BUCKETIZE_KJT_CASES, bucketize_kjt_inputs  # noqa: E402
from make_golden_kjt import load_reference_kjt  # noqa: E402


def main() -> None:
    jt = load_reference_kjt()
    path = "/root/reference/torchrec/distributed/embedding_sharding.py"
    tree = ast.parse(open(path).read(), filename=path)
    fn = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name == "bucketize_kjt_before_all2all"]
    ns = {"torch": torch, "KeyedJaggedTensor": jt.KeyedJaggedTensor, "Tuple": Tuple, "Optional": Optional}
    exec(compile(ast.Module(body=fn, type_ignores=[]), path, "exec"), ns)
    bucketize = ns["bucketize_kjt_before_all2all"]
    out = {}
    for ci, case in enumerate(BUCKETIZE_KJT_CASES):
        keys, lengths, values, weights, B, block_sizes = bucketize_kjt_inputs(case)
        k = jt.KeyedJaggedTensor(keys=keys, values=torch.from_numpy(values), lengths=torch.from_numpy(lengths),
                                 weights=None if weights is None else torch.from_numpy(weights), stride=B)
        got, unbucketize = bucketize(k, case["W"], torch.from_numpy(block_sizes), output_permute=case["permute"],
                                     bucketize_pos=case["pos"])
        pre = f"case{ci}"
        out[f"{pre}/keys"] = np.array(got.keys())
        out[f"{pre}/values"] = got.values().numpy()
        out[f"{pre}/lengths"] = got.lengths().numpy()
        out[f"{pre}/stride"] = np.int64(got.stride())
        if got.weights_or_none() is not None:
            out[f"{pre}/weights"] = got.weights().numpy()
        if unbucketize is not None:
            out[f"{pre}/unbucketize"] = unbucketize.numpy()
    dst = os.path.join(HERE, "bucketize_kjt.npz")
    np.savez_compressed(dst, **out)
    print(f"wrote {dst}: {len(out)} arrays, {os.path.getsize(dst)} bytes")


if __name__ == "__main__":
    main()
