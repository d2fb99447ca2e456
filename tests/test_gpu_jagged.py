"""GPU parity of the bit-exact jagged integer ops against the oracle and the reference's
literal test vectors (through the C ABI via b200_sparse.ops)."""
import numpy as np
import pytest
import torch

from conftest import make_jagged
from oracle import jagged_ref as R

pytestmark = pytest.mark.gpu


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


@pytest.fixture(scope="module")
def ops():
    from b200_sparse import ops as o
    return o


@pytest.mark.parametrize("n", [0, 1, 31, 100, 2048, 2049, 70001, 1703936])
@pytest.mark.parametrize("dt", [np.int32, np.int64])
def test_cumsum(ops, n, dt):
    rng = np.random.default_rng(n)
    x = rng.integers(0, 100, size=n).astype(dt)
    for mode, ref in ((0, R.complete_cumsum), (1, R.inclusive_cumsum), (2, R.exclusive_cumsum)):
        got = ops.cumsum(dev(x), mode).cpu().numpy()
        np.testing.assert_array_equal(got, ref(x))


def test_permute_2d_reference_vector(ops):
    # torchrec/sparse/tests/test_jagged_tensor.py:671-702 (test_permute)
    values = np.array([1, 2, 3, 4, 5, 6, 7, 8], dtype=np.float32)
    lengths = np.array([0, 2, 0, 1, 1, 1, 0, 3, 0], dtype=np.int32).reshape(3, 3)
    perm = np.array([1, 0, 2], dtype=np.int32)
    ol, ov, _ = ops.permute_sparse_data(dev(perm), dev(lengths), dev(values), None, None, True)
    np.testing.assert_array_equal(ov.cpu().numpy(), [3, 4, 5, 1, 2, 6, 7, 8])
    np.testing.assert_array_equal(ol.cpu().numpy().reshape(-1), [1, 1, 1, 0, 2, 0, 0, 3, 0])


@pytest.mark.parametrize("T,B,L", [(3, 5, 4), (26, 512, 1), (7, 33, 40), (1, 1, 0), (4, 1000, 3), (26, 5000, 2),
                                   (700, 7, 3), (1100, 3, 2)])
@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("vdt", [np.int64, np.int32])
def test_permute_2d_random(ops, T, B, L, weighted, vdt):
    rng = np.random.default_rng(T * 1000 + B)
    lengths, _, values = make_jagged(rng, T, B, L, 1000, idx_dtype=vdt, empty_frac=0.3 if L else 1.0)
    lengths = lengths.astype(np.int32).reshape(T, B)
    weights = rng.random(values.size).astype(np.float32) if weighted else None
    perm = rng.integers(0, T, size=T + 2).astype(np.int32)  # with duplicates
    el, ev, ew = R.permute_2D_sparse_data(perm, lengths, values, weights)
    for total in (None, int(ev.size)):
        ol, ov, ow = ops.permute_sparse_data(dev(perm), dev(lengths), dev(values),
                                             dev(weights) if weighted else None, total, True)
        np.testing.assert_array_equal(ol.cpu().numpy(), el)
        np.testing.assert_array_equal(ov.cpu().numpy(), ev)
        if weighted:
            np.testing.assert_array_equal(ow.cpu().numpy(), ew)


def test_permute_1d(ops):
    rng = np.random.default_rng(5)
    lengths = rng.integers(0, 6, size=50).astype(np.int64)
    values = rng.integers(0, 1 << 40, size=int(lengths.sum())).astype(np.int64)
    perm = rng.permutation(50).astype(np.int32)[:37]
    el, ev, _ = R.permute_1D_sparse_data(perm, lengths, values)
    ol, ov, _ = ops.permute_sparse_data(dev(perm), dev(lengths), dev(values), None, None, False)
    np.testing.assert_array_equal(ol.cpu().numpy(), el)
    np.testing.assert_array_equal(ov.cpu().numpy(), ev)


@pytest.mark.parametrize("dt", [np.int32, np.int64])
def test_expand_into_jagged_permute(ops, dt):
    rng = np.random.default_rng(11)
    n = 40
    lens = rng.integers(0, 9, size=n)
    perm = rng.permutation(n).astype(np.int32)
    in_off = np.zeros(n + 1, dtype=dt)
    np.cumsum(lens, out=in_off[1:])
    out_off = np.zeros(n + 1, dtype=dt)
    np.cumsum(lens[perm], out=out_off[1:])
    exp = R.expand_into_jagged_permute(perm, in_off, out_off, int(out_off[-1]))
    got = ops.expand_into_jagged_permute(dev(perm), dev(in_off), dev(out_off), int(out_off[-1]))
    np.testing.assert_array_equal(got.cpu().numpy(), exp)


@pytest.mark.parametrize("dt", [np.int32, np.int64])
def test_offsets_range(ops, dt):
    rng = np.random.default_rng(3)
    lens = rng.integers(0, 50, size=300)
    off = np.zeros(300, dtype=dt)
    np.cumsum(lens[:-1], out=off[1:])
    total = int(lens.sum())
    np.testing.assert_array_equal(ops.offsets_range(dev(off), total).cpu().numpy(), R.offsets_range(off, total))


def test_bucketize_reference_long_indices(ops):
    # fbgemm_gpu/test/sparse_ops_test.py:446-511
    bs = np.array([3, 4, 5], dtype=np.int64)
    lengths = np.array([0, 3, 2, 0, 1, 4], dtype=np.int64)
    for indices, nl, ni in (
        ([1, 2, 3, 4, 5, 6, 7, 8, 9, 10],
         [0, 2, 0, 0, 0, 0, 0, 1, 2, 0, 1, 3, 0, 0, 0, 0, 0, 1], [1, 2, 0, 0, 1, 1, 2, 3, 4, 0]),
        ([1, 2, 3, 100061827127359, 5, 6, 7, -8, 100058153792324, 10],
         [0, 2, 0, 0, 0, 0, 0, 1, 2, 0, 1, 1, 0, 0, 0, 0, 0, 3],
         [1, 2, 0, 33353942375786, 1, 1, 2, 6148914691236517202, 33352717930774, 0]),
    ):
        out = ops.block_bucketize_sparse_features(dev(lengths), dev(np.array(indices, dtype=np.int64)), False,
                                                  False, dev(bs), 3, None)
        np.testing.assert_array_equal(out[0].cpu().numpy(), nl)
        np.testing.assert_array_equal(out[1].cpu().numpy(), ni)


@pytest.mark.parametrize("T,B,L,W", [(3, 7, 5, 2), (26, 256, 1, 8), (2, 9, 64, 4), (5, 100, 3, 3), (26, 4096, 1, 8),
                                     (6, 3000, 9, 5)])
@pytest.mark.parametrize("idt", [np.int32, np.int64])
@pytest.mark.parametrize("extras", [False, True])
def test_bucketize_random(ops, T, B, L, W, idt, extras):
    rng = np.random.default_rng(T * 100 + W)
    rows = rng.integers(1, 5000, size=T)
    lengths, _, indices = make_jagged(rng, T, B, L, rows, idx_dtype=idt, empty_frac=0.1)
    odt = np.int32
    lengths = lengths.astype(odt)
    bs = ((rows + W - 1) // W).astype(idt)
    weights = rng.random(indices.size).astype(np.float32) if extras else None
    exp = R.block_bucketize_sparse_features(lengths, indices, extras, extras, bs, W, weights)
    got = ops.block_bucketize_sparse_features(dev(lengths), dev(indices), extras, extras, dev(bs), W,
                                              dev(weights) if extras else None)
    for g, e in zip(got, exp):
        if e is None:
            assert g is None
        else:
            np.testing.assert_array_equal(g.cpu().numpy(), e)


@pytest.mark.parametrize("dt", [np.int32, np.int64])
@pytest.mark.parametrize("mode", [1, 2])
@pytest.mark.parametrize("L", [1, 20])
def test_bounds_check(ops, dt, mode, L):
    rng = np.random.default_rng(L)
    T, B = 4, 37
    rows = np.array([10, 200, 3, 50], dtype=np.int64)
    lengths, offsets, indices = make_jagged(rng, T, B, L, rows, idx_dtype=dt, empty_frac=0.1)
    bad = rng.random(indices.size) < 0.2
    indices[bad] = rng.integers(-3, 400, size=int(bad.sum()))
    ei, eo, ew = R.bounds_check_indices(rows, indices, offsets, mode)
    gi, go = dev(indices), dev(offsets)
    warn = torch.zeros(1, dtype=torch.int64, device="cuda")
    ops.bounds_check_indices(dev(rows), gi, go, mode, warn)
    np.testing.assert_array_equal(gi.cpu().numpy(), ei)
    np.testing.assert_array_equal(go.cpu().numpy(), eo)
    if mode == 1:
        assert int(warn.item()) == ew
    # malformed offsets: repaired with the reference's sequential rule
    offsets2 = offsets.copy()
    offsets2[5] = offsets2[-1] + 7
    offsets2[9] = -2
    ei, eo, _ = R.bounds_check_indices(rows, indices, offsets2, 2)
    gi, go = dev(indices), dev(offsets2)
    ops.bounds_check_indices(dev(rows), gi, go, 2, warn)
    np.testing.assert_array_equal(go.cpu().numpy(), eo)
    np.testing.assert_array_equal(gi.cpu().numpy(), ei)


# ---------------------------------------------------------------------------------------------
# CUDA path vs fixtures produced by the reference's own CPU kernels (tests/golden/make_golden.py)
# ---------------------------------------------------------------------------------------------
class _CudaBackend:
    @staticmethod
    def _d(x):
        return None if x is None else torch.from_numpy(np.ascontiguousarray(x)).cuda()

    @staticmethod
    def _h(x):
        return None if x is None else x.cpu().numpy()

    @classmethod
    def cumsum(cls, x, mode):
        from b200_sparse import cabi, ops
        m = {"complete": cabi.CUMSUM_COMPLETE, "inclusive": cabi.CUMSUM_INCLUSIVE, "exclusive": cabi.CUMSUM_EXCLUSIVE}
        return cls._h(ops.cumsum(cls._d(x), m[mode]))

    @classmethod
    def permute(cls, perm, lengths, values, weights, two_d):
        from b200_sparse import ops
        r = ops.permute_sparse_data(cls._d(perm), cls._d(lengths), cls._d(values), cls._d(weights), None, two_d)
        return tuple(cls._h(t) for t in r)

    @classmethod
    def expand(cls, perm, in_off, out_off, n):
        from b200_sparse import ops
        return cls._h(ops.expand_into_jagged_permute(cls._d(perm), cls._d(in_off), cls._d(out_off), n))

    @classmethod
    def offsets_range(cls, off, n):
        from b200_sparse import ops
        return cls._h(ops.offsets_range(cls._d(off), n))

    @classmethod
    def bucketize(cls, lengths, indices, pos, seq, block, W, weights):
        from b200_sparse import ops
        r = ops.block_bucketize_sparse_features(cls._d(lengths), cls._d(indices), pos, seq, cls._d(block), W,
                                                cls._d(weights))
        return tuple(cls._h(t) for t in r)

    @classmethod
    def bounds_check(cls, rows, indices, offsets, mode):
        from b200_sparse import ops
        i, o = cls._d(indices.copy()), cls._d(offsets.copy())
        warn = torch.zeros(1, dtype=torch.int64, device="cuda")
        ops.bounds_check_indices(cls._d(rows), i, o, mode, warn)
        return cls._h(i), cls._h(o), int(warn.item())


def test_cuda_matches_reference_cpu_kernel_fixtures():
    import golden_cases
    assert golden_cases.run_all(_CudaBackend) == 38
