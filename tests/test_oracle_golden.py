"""CPU: pins the oracle (a) to the reference's literal test vectors, (b) to the reference's own
compiled scalar code in oracle/_ref (when built), (c) to torch.nn.EmbeddingBag + the closed-form
optimizer the reference's tests use, and (d) checks the C restatement against the numpy one."""
import numpy as np
import pytest
import torch

from conftest import make_jagged
from oracle import c_oracle, jagged_ref as J, tbe_ref as R

have_ref = pytest.mark.skipif(c_oracle.ref_lib() is None, reason="oracle/_ref not built (no reference checkout)")


def test_recat_docstring_vectors():
    # torchrec/distributed/dist_data.py:60-66
    assert J.get_recat(2, 4, 1) == [0, 2, 4, 6, 1, 3, 5, 7]
    assert J.get_recat(2, 4, 2) == [0, 4, 2, 6, 1, 5, 3, 7]


def test_permute_reference_vectors():
    # torchrec/sparse/tests/test_jagged_tensor.py:671-702 and :704 (duplicates)
    values = np.array([1, 2, 3, 4, 5, 6, 7, 8], dtype=np.float32)
    lengths = np.array([0, 2, 0, 1, 1, 1, 0, 3, 0], dtype=np.int32).reshape(3, 3)
    ol, ov, _ = J.permute_2D_sparse_data(np.array([1, 0, 2], dtype=np.int32), lengths, values)
    assert ov.tolist() == [3, 4, 5, 1, 2, 6, 7, 8] and ol.reshape(-1).tolist() == [1, 1, 1, 0, 2, 0, 0, 3, 0]
    _, ov, _ = J.permute_2D_sparse_data(np.array([2, 0, 0], dtype=np.int32), lengths, values)
    assert ov.tolist() == [6, 7, 8, 1, 2, 1, 2]


def test_cumsum_vectors():
    x = np.array([3, 0, 2, 5], dtype=np.int32)
    assert J.complete_cumsum(x).tolist() == [0, 3, 3, 5, 10]
    assert J.inclusive_cumsum(x).tolist() == [3, 3, 5, 10]
    assert J.exclusive_cumsum(x).tolist() == [0, 3, 3, 5]


def test_bucketize_reference_vectors():
    # fbgemm_gpu/test/sparse_ops_test.py:446-511 (incl. negative and > 2^46 ids)
    bs = np.array([3, 4, 5], dtype=np.int64)
    lengths = np.array([0, 3, 2, 0, 1, 4], dtype=np.int64)
    out = J.block_bucketize_sparse_features(lengths, np.arange(1, 11, dtype=np.int64), False, False, bs, 3)
    assert out[0].tolist() == [0, 2, 0, 0, 0, 0, 0, 1, 2, 0, 1, 3, 0, 0, 0, 0, 0, 1]
    assert out[1].tolist() == [1, 2, 0, 0, 1, 1, 2, 3, 4, 0]
    idx = np.array([1, 2, 3, 100061827127359, 5, 6, 7, -8, 100058153792324, 10], dtype=np.int64)
    out = J.block_bucketize_sparse_features(lengths, idx, False, False, bs, 3)
    assert out[0].tolist() == [0, 2, 0, 0, 0, 0, 0, 1, 2, 0, 1, 1, 0, 0, 0, 0, 0, 3]
    assert out[1].tolist() == [1, 2, 0, 33353942375786, 1, 1, 2, 6148914691236517202, 33352717930774, 0]


def test_bucketize_block_vectors():
    # fbgemm_gpu/test/sparse_ops_test.py:602-680 (test_block_bucketize_sparse_features, 2 buckets)
    lengths = np.array([0, 2, 1, 3, 2, 3, 3, 1], dtype=np.int32)
    indices = np.array([3, 4, 15, 11, 28, 29, 1, 10, 11, 12, 13, 11, 22, 20, 20], dtype=np.int32)
    bs = np.array([5, 15, 10, 20], dtype=np.int32)
    nl, ni, _, npos, _ = J.block_bucketize_sparse_features(lengths, indices, True, False, bs, 2)
    assert nl.tolist() == [0, 2, 0, 1, 1, 0, 1, 0, 0, 0, 1, 2, 1, 3, 2, 1]
    assert ni.tolist() == [3, 4, 11, 1, 11, 0, 13, 14, 0, 1, 2, 3, 2, 0, 0]
    assert npos.tolist() == [0, 1, 0, 0, 0, 0, 1, 2, 1, 0, 1, 2, 1, 2, 0]


def test_interaction_reference_golden():
    # torchrec/models/tests/test_dlrm.py:247-327: seed 0, B=3, D=4, F=4, the reference's whole expected tensor
    # at the reference's own tolerance (rtol = atol = 1e-4, :312-319)
    from golden_cases import INTERACTION_GOLDEN
    torch.manual_seed(0)
    dense, sparse = torch.rand((3, 4)), torch.rand((3, 4, 4))
    exp = np.array(INTERACTION_GOLDEN, dtype=np.float32)
    out = R.interaction_forward(dense.numpy(), sparse.numpy())
    assert out.shape == (3, 14)
    np.testing.assert_allclose(out, exp, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(c_oracle.interaction_forward(dense.numpy(), sparse.numpy()), exp, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(out, c_oracle.interaction_forward(dense.numpy(), sparse.numpy()), rtol=1e-5)


def _tables(rng, rows, dims):
    Do = np.concatenate([[0], np.cumsum(dims)]).astype(np.int32)
    sizes = [r * d for r, d in zip(rows, dims)]
    wo = np.concatenate([[0], np.cumsum(sizes)[:-1]]).astype(np.int64)
    hsc = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    return rng.standard_normal(sum(sizes)).astype(np.float32), wo, Do, hsc


@pytest.mark.parametrize("mode", ["sum", "mean", "weighted"])
def test_forward_vs_embedding_bag_and_c(mode):
    rng = np.random.default_rng(0)
    rows, dims, B = [30, 7, 200], [16, 8, 32], 25
    w, wo, Do, _ = _tables(rng, rows, dims)
    _, offsets, indices = make_jagged(rng, 3, B, 5, rows)
    psw = rng.random(indices.size).astype(np.float32) if mode == "weighted" else None
    got = R.tbe_forward(w, wo, Do, indices, offsets, mode == "mean", psw)
    np.testing.assert_allclose(c_oracle.tbe_forward(w, wo, Do, indices, offsets, mode == "mean", psw), got,
                               rtol=1e-6, atol=1e-6)
    for t in range(3):
        tab = torch.from_numpy(w[wo[t]: wo[t] + rows[t] * dims[t]].reshape(rows[t], dims[t]))
        off_t = offsets[t * B:(t + 1) * B + 1]
        idx_t = torch.from_numpy(indices[off_t[0]:off_t[-1]])
        ref = torch.nn.functional.embedding_bag(
            idx_t, tab, torch.from_numpy(off_t - off_t[0]), mode="mean" if mode == "mean" else "sum",
            per_sample_weights=torch.from_numpy(psw[off_t[0]:off_t[-1]]) if psw is not None else None,
            include_last_offset=True)
        np.testing.assert_allclose(got[:, Do[t]:Do[t + 1]], ref.numpy(), rtol=1e-5, atol=1e-5)


@have_ref
def test_forward_vs_reference_code():
    rng = np.random.default_rng(1)
    rows, D, B = 50, 24, 40
    tab = rng.standard_normal((rows, D)).astype(np.float32)
    _, offsets, indices = make_jagged(rng, 1, B, 6, [rows])
    psw = rng.random(indices.size).astype(np.float32)
    Do, wo = np.array([0, D], np.int32), np.array([0], np.int64)
    for mean, pw in ((False, None), (True, None), (False, psw)):
        ref = c_oracle.ref_spmdm_f32(tab, indices, offsets, pw, mean)
        np.testing.assert_array_equal(R.tbe_forward(tab.reshape(-1), wo, Do, indices, offsets, mean, pw), ref)


@have_ref
def test_int8_forward_and_quantize_vs_reference_code():
    rng = np.random.default_rng(2)
    rows, D, B = 60, 20, 30
    wf = rng.standard_normal((rows, D)).astype(np.float32)
    q = R.quantize_rowwise_8bit_sbhalf(wf)
    # torch's own FBGEMM-backed op is the same reference routine (FloatToFusedNBitRowwiseQuantizedSBHalf)
    rb = R.int8_row_bytes(D)
    packed = np.zeros((rows, rb), np.uint8)
    packed[:, :4], packed[:, 4:4 + D] = q[:, D:], q[:, :D]
    _, offsets, indices = make_jagged(rng, 1, B, 4, [rows])
    ref = c_oracle.ref_spmdm_u8(packed, D, indices, offsets)
    got = R.tbe_int8_forward(packed.reshape(-1), np.array([0], np.int64), np.array([0, D], np.int32), indices,
                             offsets, 16, False, None, np.float32)
    np.testing.assert_array_equal(got, ref)
    deq = q[:, :D].astype(np.float32) * np.frombuffer(q[:, D:D + 2].tobytes(), np.float16).astype(np.float32)[:, None] \
        + np.frombuffer(q[:, D + 2:].tobytes(), np.float16).astype(np.float32)[:, None]
    assert np.abs(deq - wf).max() <= (wf.max(1) - wf.min(1)).max() / 255 * 0.51 + 2e-3


def test_quantizer_oracle_matches_reference_fixture():
    """a17 pinned: the numpy restatement reproduces, bit for bit, what the reference's own compiled
    QuantUtils.cc produced (tests/golden/quantize_sbhalf.npz, tests/golden/make_golden_quant.py) - constant
    rows, all-zero rows, fp16-overflowing rows and exact ties included."""
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "quantize_sbhalf.npz"))
    names = sorted({k.split(".")[0] for k in g.files})
    assert len(names) == 5
    for n in names:
        np.testing.assert_array_equal(R.quantize_rowwise_8bit_sbhalf(g[n + ".w"]), g[n + ".q"], err_msg=n)


@pytest.mark.skipif(c_oracle.quant_ref_lib() is None, reason="oracle/_ref not built (no reference checkout)")
def test_quantizer_oracle_vs_reference_code():
    """Random shapes against the reference's scalar definition AND the dispatcher its torch op calls."""
    rng = np.random.default_rng(5)
    for rows, cols in ((200, 128), (31, 64), (7, 20), (64, 256), (3, 1)):
        w = (rng.standard_normal((rows, cols)) * rng.uniform(1e-3, 1e3)).astype(np.float32)
        mine = R.quantize_rowwise_8bit_sbhalf(w)
        np.testing.assert_array_equal(mine, c_oracle.ref_quantize_sbhalf(w, dispatch=False))
        np.testing.assert_array_equal(mine, c_oracle.ref_quantize_sbhalf(w, dispatch=True))


@pytest.mark.parametrize("opt", ["sgd", "rowwise_adagrad"])
@pytest.mark.parametrize("mode", ["sum", "mean", "weighted"])
def test_backward_closed_form_and_c(opt, mode):
    """weights == addcdiv(w, -lr, grad, sqrt(state)+eps), state == mean(grad^2)
    (fbgemm_gpu/test/split_table_batched_embeddings_test.py:1756-1785)."""
    rng = np.random.default_rng(3)
    rows, dims, B = [30, 7, 200], [16, 8, 32], 25
    lr, eps = 0.1, 1e-4
    w, wo, Do, hsc = _tables(rng, rows, dims)
    _, offsets, indices = make_jagged(rng, 3, B, 5, rows)
    psw = rng.random(indices.size).astype(np.float32) if mode == "weighted" else None
    grad = rng.standard_normal((B, int(Do[-1]))).astype(np.float32)
    w1, m1 = w.copy(), np.zeros(int(hsc[-1]), np.float32)
    R.tbe_backward(grad, w1, wo, Do, hsc, indices, offsets, opt, lr, eps, m1, hsc[:-1], mode == "mean", psw)
    w2, m2 = w.copy(), np.zeros(int(hsc[-1]), np.float32)
    c_oracle.tbe_backward(grad, w2, wo, Do, hsc, indices, offsets, opt, lr, eps, m2, hsc[:-1], mode == "mean", psw)
    np.testing.assert_allclose(w2, w1, rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(m2, m1, rtol=1e-5, atol=1e-6)
    for t in range(3):
        tab = torch.from_numpy(w[wo[t]: wo[t] + rows[t] * dims[t]].reshape(rows[t], dims[t]).copy()).requires_grad_()
        off_t = offsets[t * B:(t + 1) * B + 1]
        out = torch.nn.functional.embedding_bag(
            torch.from_numpy(indices[off_t[0]:off_t[-1]]), tab, torch.from_numpy(off_t - off_t[0]),
            mode="mean" if mode == "mean" else "sum", include_last_offset=True,
            per_sample_weights=torch.from_numpy(psw[off_t[0]:off_t[-1]]) if psw is not None else None)
        out.backward(torch.from_numpy(grad[:, Do[t]:Do[t + 1]]))
        g = tab.grad
        if opt == "sgd":
            exp = tab.detach() - lr * g
        else:
            state = g.pow(2).mean(1)
            exp = torch.addcdiv(tab.detach(), g, (state.sqrt() + eps).unsqueeze(1), value=-lr)
            np.testing.assert_allclose(m1[hsc[t]:hsc[t + 1]], state.numpy(), rtol=1e-4, atol=1e-6)
        np.testing.assert_allclose(w1[wo[t]: wo[t] + rows[t] * dims[t]].reshape(rows[t], dims[t]), exp.numpy(),
                                   rtol=1e-4, atol=1e-5)


@have_ref
def test_rowwise_adagrad_vs_reference_code():
    """rowwise_sparse_adagrad_ref on coalesced rows == the oracle's update (lr passed negated,
    embedding_backward_split_cpu_template.cpp:171)."""
    rng = np.random.default_rng(4)
    rows, D, B = 40, 16, 64
    lr, eps = 0.05, 1e-6
    tab = rng.standard_normal((rows, D)).astype(np.float32)
    idx = rng.permutation(rows)[:B % rows or rows].astype(np.int64)  # unique ids, one per bag
    B = idx.size
    offsets = np.arange(B + 1, dtype=np.int64)
    grad = rng.standard_normal((B, D)).astype(np.float32)
    w_ref, h_ref = tab.copy(), np.zeros(rows, np.float32)
    c_oracle.ref_rowwise_adagrad(w_ref, grad, h_ref, idx, eps, -lr)
    w, h = tab.copy().reshape(-1), np.zeros(rows, np.float32)
    R.tbe_backward(grad, w, np.array([0], np.int64), np.array([0, D], np.int32), np.array([0, rows], np.int64), idx,
                   offsets, "rowwise_adagrad", lr, eps, h, np.array([0], np.int64))
    np.testing.assert_allclose(w.reshape(rows, D), w_ref, rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(h, h_ref, rtol=1e-6, atol=1e-7)


# ---------------------------------------------------------------------------------------------
# Fixtures produced by the reference's own CPU kernels (tests/golden/make_golden.py)
# ---------------------------------------------------------------------------------------------
class _OracleBackend:
    @staticmethod
    def cumsum(x, mode):
        from oracle import jagged_ref as J
        return getattr(J, mode + "_cumsum")(x)

    @staticmethod
    def permute(perm, lengths, values, weights, two_d):
        from oracle import jagged_ref as J
        return (J.permute_2D_sparse_data if two_d else J.permute_1D_sparse_data)(perm, lengths, values, weights)

    @staticmethod
    def expand(perm, in_off, out_off, n):
        from oracle import jagged_ref as J
        return J.expand_into_jagged_permute(perm, in_off, out_off, n)

    @staticmethod
    def offsets_range(off, n):
        from oracle import jagged_ref as J
        return J.offsets_range(off, n)

    @staticmethod
    def bucketize(lengths, indices, pos, seq, block, W, weights):
        from oracle import jagged_ref as J
        return J.block_bucketize_sparse_features(lengths, indices, pos, seq, block, W, weights)

    @staticmethod
    def bounds_check(rows, indices, offsets, mode):
        from oracle import jagged_ref as J
        return J.bounds_check_indices(rows, indices, offsets, mode)


def test_oracle_matches_reference_cpu_kernel_fixtures():
    import golden_cases
    assert golden_cases.run_all(_OracleBackend) == 38


def test_philox4x32_10_known_answers():
    """Random123's known-answer vectors for philox4x32-10 (examples/kat_vectors: zero, all-ones and
    digits-of-pi inputs) pin the generator the stochastic-rounding oracle - and through it the CUDA
    kernel (tests/test_gpu_tbe.py::test_backward_stochastic_rounding_bits) - is built on."""
    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
        ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
        ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
         (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1)),
    ]
    for ctr, key, exp in kat:
        got = R.philox4x32_10(np.array(ctr, dtype=np.uint32), (np.uint32(key[0]), np.uint32(key[1])))
        assert [int(x) for x in got] == list(exp), ([hex(int(x)) for x in got], [hex(x) for x in exp])
    # vectorised call == one call per counter
    ctrs = np.array([k[0] for k in kat], dtype=np.uint32)
    many = R.philox4x32_10(ctrs, (np.uint32(7), np.uint32(9)))
    for i in range(3):
        np.testing.assert_array_equal(many[i], R.philox4x32_10(ctrs[i], (np.uint32(7), np.uint32(9))))


def test_stochastic_round_properties():
    """Results are representable in the 16-bit type, bracket the input by one ulp, are unbiased, and
    reduce to the identity for values that are already representable."""
    import torch
    rng = np.random.default_rng(0)
    x = rng.standard_normal(1 << 16).astype(np.float32)
    x = (np.sign(x) * np.maximum(np.abs(x), 2.0 ** -10)).astype(np.float32)  # fp16 normal range (see oracle)
    for dtype, tdt, keep in (("fp16", torch.float16, 13), ("bf16", torch.bfloat16, 16)):
        for epv in (8, 4):
            y = R.stochastic_round(x, dtype, seed=123, offset=5, first_element=64, epv=epv)
            t = torch.from_numpy(y.copy())
            assert torch.equal(t.to(tdt).float(), t)  # exactly representable: the later RN pack is exact
            lo = (x.view(np.uint32) & ~np.uint32((1 << keep) - 1)).view(np.float32)   # truncation
            ulp = np.abs((lo.view(np.uint32) + np.uint32(1 << keep)).view(np.float32) - lo)
            assert np.all((y == lo) | (np.abs(y - lo) == ulp))
            # unbiased: E[y] = x, error of the mean ~ ulp / sqrt(n)
            assert abs(float(np.mean((y - x) / ulp))) < 0.01
            again = R.stochastic_round(y, dtype, seed=9, offset=1, first_element=0, epv=epv)
            np.testing.assert_array_equal(again, y)
    z = np.array([np.inf, -np.inf, np.nan, 0.0, -0.0, 1.0, 65504.0, 1e-30], dtype=np.float32)
    out = R.stochastic_round(z, "fp16", 1, 1, 0, 8)
    assert np.isinf(out[0]) and np.isinf(out[1]) and np.isnan(out[2]) and out[3] == 0 and out[5] == 1.0
