"""TEST INFRASTRUCTURE ONLY.  Registers CPU kernels for the torch.ops.fbgemm.* jagged ops from the
numpy oracle so that the HOST logic above them (KeyedJaggedTensor bookkeeping, KJTAllToAll,
bucketize + row-wise redistribution, planner) can run in this GPU-less container and under gloo.
The product registers CUDA kernels only (fbgemm_gpu/_torch_ops.py) and never imports this file."""
import numpy as np
import torch

import fbgemm_gpu  # noqa: F401  (defines the schemas)
from oracle import jagged_ref as J

_lib = torch.library.Library("fbgemm", "IMPL", "CPU")
_done = False


def _t(a, like=None):
    t = torch.from_numpy(np.ascontiguousarray(a))
    return t.to(like.dtype) if like is not None else t


def register() -> None:
    global _done
    if _done:
        return
    _done = True
    _lib.impl("asynchronous_complete_cumsum", lambda x: _t(J.complete_cumsum(x.numpy()), x))
    _lib.impl("asynchronous_inclusive_cumsum", lambda x: _t(J.inclusive_cumsum(x.numpy()), x))
    _lib.impl("asynchronous_exclusive_cumsum", lambda x: _t(J.exclusive_cumsum(x.numpy()), x))

    def perm(fn):
        def impl(permute, lengths, values, weights=None, permuted_lengths_sum=None):
            l, v, w = fn(permute.numpy(), lengths.numpy(), values.numpy(),
                         None if weights is None else weights.numpy())
            return _t(l, lengths), _t(v, values), (None if w is None else _t(w, weights))
        return impl

    _lib.impl("permute_2D_sparse_data", perm(J.permute_2D_sparse_data))
    _lib.impl("permute_sparse_data", perm(J.permute_2D_sparse_data))
    _lib.impl("permute_1D_sparse_data", perm(J.permute_1D_sparse_data))
    _lib.impl("expand_into_jagged_permute",
              lambda p, i, o, n: _t(J.expand_into_jagged_permute(p.numpy(), i.numpy(), o.numpy(), n), p))
    _lib.impl("offsets_range", lambda o, n: _t(J.offsets_range(o.numpy(), n), o))

    def bucketize(lengths, indices, bucketize_pos, sequence, block_sizes, my_size, weights=None):
        r = J.block_bucketize_sparse_features(lengths.numpy(), indices.numpy(), bucketize_pos, sequence,
                                              block_sizes.numpy(), my_size,
                                              None if weights is None else weights.numpy())
        nl, ni, nw, pos, unb = r
        return (_t(nl, lengths), _t(ni, indices), None if nw is None else _t(nw, weights),
                None if pos is None else _t(pos, indices), None if unb is None else _t(unb, indices))

    _lib.impl("block_bucketize_sparse_features", bucketize)
