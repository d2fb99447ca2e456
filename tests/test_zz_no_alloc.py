"""b200_tbe_backward_static: the fused backward with its scratch in a static device array (no allocation).

No counterpart in the reference.  The file is named so that
pytest collects it last.  The new entry point is called through the C ABI (ctypes) and compared with the
existing b200_tbe_backward on the same inputs, twice in a row (the second call finds the first call's
contents in the shared scratch), bit for bit: both entry points run the same deterministic kernels."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _case(seed: int, rows, dims, B: int, L: int):
    rng = np.random.default_rng(seed)
    T = len(rows)
    Do = np.concatenate([[0], np.cumsum(dims)]).astype(np.int32)
    hsc = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    wo = np.concatenate([[0], np.cumsum([r * d for r, d in zip(rows, dims)])[:-1]]).astype(np.int64)
    n_w = int(sum(r * d for r, d in zip(rows, dims)))
    w = rng.standard_normal(n_w).astype(np.float32)
    lengths = rng.integers(0, L + 1, size=T * B)
    offsets = np.concatenate([[0], np.cumsum(lengths)]).astype(np.int64)
    idx = np.concatenate([rng.integers(0, rows[t], size=int(lengths[t * B:(t + 1) * B].sum()))
                          for t in range(T)]).astype(np.int64)
    grad = rng.standard_normal((B, int(Do[-1]))).astype(np.float32)
    return T, Do, hsc, wo, w, offsets, idx, grad


@pytest.mark.parametrize("rows,dims,B,L", [([50, 3, 400], [128, 128, 128], 96, 3),
                                           ([1000, 7, 20000, 64], [64, 32, 128, 16], 700, 5)])
def test_tbe_backward_static_matches_tbe_backward(rows, dims, B, L):
    from b200_sparse import ops
    from b200_sparse.cabi import load
    lib = load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    assert lib.b200_tbe_backward_static_max_scratch_bytes() == 16 << 20
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    stream = torch.cuda.current_stream(dev).cuda_stream
    for call in range(2):  # two calls in a row share the static scratch
        T, Do, hsc, wo, w, offsets, idx, grad = _case(10 * call + len(rows), rows, dims, B, L)
        max_D, mult8 = int(max(dims)), int(all(d % 8 == 0 for d in dims))
        d_idx, d_off, d_hsc, d_wo, d_Do, d_grad = t(idx), t(offsets), t(hsc), t(wo), t(Do), t(grad)
        d_mo = t(hsc[:-1].copy())
        srt = ops.tbe_sort_indices(d_idx, d_off, d_hsc, int(np.log2(hsc[-1]) + 1), B, False)
        N = int(idx.size)
        for opt in (1, 0):  # rowwise Adagrad, SGD
            w_ref, m_ref = t(w), torch.zeros(int(hsc[-1]), device=dev)
            w_new, m_new = t(w), torch.zeros(int(hsc[-1]), device=dev)
            scratch = torch.empty(lib.b200_tbe_backward_scratch_bytes(N, max_D), dtype=torch.uint8, device=dev)

            def common(wt, mt):
                return (d_grad.data_ptr(), d_grad.stride(0), srt.keys.data_ptr(), srt.payload.data_ptr(), None, N,
                        srt.key_bits, wt.data_ptr(), 2, d_wo.data_ptr(), d_Do.data_ptr(), d_hsc.data_ptr(), T, B,
                        max_D, mult8, d_off.data_ptr(), 1, None, 0, opt, mt.data_ptr(), d_mo.data_ptr(), 0.05, 1e-8)
            assert lib.b200_tbe_backward(*common(w_ref, m_ref), scratch.data_ptr(), stream) == 0
            mem_before = torch.cuda.memory_allocated(dev)
            assert lib.b200_tbe_backward_static(*common(w_new, m_new), stream) == 0
            assert torch.cuda.memory_allocated(dev) == mem_before
            torch.cuda.synchronize(dev)
            assert not torch.equal(w_new, t(w)), "the update must have touched the tables"
            assert torch.equal(w_new, w_ref), f"call {call} optimizer {opt}: weights differ"
            assert torch.equal(m_new, m_ref), f"call {call} optimizer {opt}: momentum differs"


def test_tbe_backward_static_rejects_oversized_problems():
    from b200_sparse.cabi import load
    lib = load()
    torch.cuda.set_device(0)
    limit = lib.b200_tbe_backward_static_max_scratch_bytes()
    N = 1
    while lib.b200_tbe_backward_scratch_bytes(N, 128) <= limit:
        N *= 2
    one = torch.zeros(4, device="cuda:0")
    rc = lib.b200_tbe_backward_static(one.data_ptr(), 4, None, None, None, N, 20, None, 2, None, None, None, 1, 1,
                                      128, 1, None, 1, None, 0, 0, None, None, 0.1, 1e-8, None)
    assert rc == -3, rc  # B200_ETOOBIG, before any pointer is touched
    assert b"scratch" in lib.b200_error_string(rc)
    # the BASELINE cfg2 shape (1,703,936 ids, dim 128) must fit
    assert lib.b200_tbe_backward_scratch_bytes(1703936, 128) <= limit
