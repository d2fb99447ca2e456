"""Parity at the sizes BASELINE.json names.

cfg2: the 26 Criteo-1TB row counts (five 40 M-row tables, 97.4 GiB of fp32 rows at dim 128), B = 65536.
cfg4: 512 tables x 1 M rows, dim 128, fp16, pooling factor 50, B = 512 (131 GB).
cfg5: 26 int8 rowwise-quantized tables, B = 4096, int32 ids, fp16 output.

The oracle cannot hold 100 GB tables, so the check runs on the TOUCHED rows only: they are gathered off the
GPU (before the update), ids are remapped to their rank among the touched ids of their table, and
oracle/tbe_ref.c runs the same batch on that compact problem.  Big-table ids are drawn from the LAST rows of
each table so that every row offset is past 2^31 elements (64-bit addressing is what these sizes add).
Untouched neighbour rows are checked to be bit-identical after the update."""
import numpy as np
import pytest
import torch

from oracle import c_oracle, tbe_ref as R

pytestmark = pytest.mark.gpu

CRITEO_ROWS = [40000000, 39060, 17295, 7424, 20265, 3, 7122, 1543, 63, 40000000, 3067956, 405282, 10, 2209,
               11938, 155, 4, 976, 14, 40000000, 40000000, 40000000, 590152, 12973, 108, 36]
EPS32 = 2.0 ** -24


def _need_free(gib: float) -> None:
    torch.cuda.empty_cache()
    free, _total = torch.cuda.mem_get_info()
    if free < gib * 2 ** 30:
        pytest.skip(f"needs {gib} GiB of free HBM, {free / 2 ** 30:.0f} GiB available")


def _draw_ids(rng, rows, B, L, tail=200_000):
    """[T][B*L] ids: small tables uniform, big tables from their last `tail` rows (duplicates included)."""
    out = []
    for r in rows:
        if r > 4 * tail:
            out.append(r - 1 - rng.integers(0, tail, size=B * L))
        else:
            out.append(rng.integers(0, r, size=B * L))
    return [a.astype(np.int64) for a in out]


def _compact(ids_per_table):
    """Touched ids per table (sorted unique) and the ids remapped to their rank."""
    uniq, remapped = [], []
    for a in ids_per_table:
        u, inv = np.unique(a, return_inverse=True)
        uniq.append(u)
        remapped.append(inv.astype(np.int64))
    return uniq, remapped


def _gather_rows(weights, wo, uniq, D):
    """Host fp32 copy of the touched rows, table after table (the compact problem's weight buffer)."""
    parts = []
    for t, u in enumerate(uniq):
        base = int(wo[t])
        idx = torch.from_numpy(u).cuda()
        # 64-bit row gather from a [rows_t, D] view of the flat buffer
        view = weights[base: base + (int(u[-1]) + 1) * D].view(-1, D)
        parts.append(view.index_select(0, idx).float().cpu().numpy())
    return np.concatenate(parts).reshape(-1)


def test_cfg2_criteo_1tb_shape_forward_and_fused_backward():
    from b200_sparse import ops
    _need_free(112)
    dev = torch.device("cuda")
    rows, D, B, L = CRITEO_ROWS, 128, 65536, 1
    T = len(rows)
    hsc = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    wo = (hsc[:-1] * D).astype(np.int64)
    assert int(wo[-1]) > 2 ** 31 and int(hsc[-1]) * D * 4 > 96 * 2 ** 30
    Do = np.arange(0, (T + 1) * D, D, dtype=np.int32)
    weights = torch.empty(int(hsc[-1]) * D, dtype=torch.float32, device=dev)
    for t in range(T):  # table by table: bounded temporaries
        weights[int(wo[t]): int(wo[t]) + rows[t] * D].normal_(0.0, 0.05)
    mom = torch.zeros(int(hsc[-1]), dtype=torch.float32, device=dev)
    rng = np.random.default_rng(2024)
    ids = _draw_ids(rng, rows, B, L)
    uniq, remapped = _compact(ids)
    U = sum(len(u) for u in uniq)
    assert min(int(ids[0].min()), int(ids[9].min())) * D > 2 ** 31  # every big-table row offset needs 64 bits
    idx_full = np.concatenate(ids)
    offsets = np.arange(T * B * L + 1, dtype=np.int64)[:: L].copy()
    assert offsets.size == T * B + 1
    c_rows = [len(u) for u in uniq]
    c_hsc = np.concatenate([[0], np.cumsum(c_rows)]).astype(np.int64)
    c_wo = (c_hsc[:-1] * D).astype(np.int64)
    c_w = _gather_rows(weights, wo, uniq, D)
    c_m = np.zeros(U, dtype=np.float32)
    c_idx = np.concatenate(remapped)

    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    d_idx, d_off, d_wo, d_Do, d_hsc = t(idx_full), t(offsets), t(wo), t(Do), t(hsc)
    # ---- forward: L = 1 is a pure gather, bit-exact
    out = ops.tbe_forward(weights, d_wo, d_Do, T * D, D, True, d_idx, d_off, None, 0)
    exp = c_oracle.tbe_forward(c_w, c_wo, Do, c_idx, offsets)
    got = out.cpu().numpy()
    assert got.shape == (B, T * D)
    np.testing.assert_array_equal(got, exp)
    del out, got

    # ---- one fused rowwise-Adagrad step
    grad = (rng.standard_normal((B, T * D)) * 0.1).astype(np.float32)
    # untouched neighbours of touched rows in a 40 M-row table (last table block past 2^31 elements)
    tb = 21
    cand = np.setdiff1d(np.unique(uniq[tb][:2000] - 1), uniq[tb])
    cand = cand[cand >= 0][:500]
    nb_idx = t(cand)
    nb_view = weights[int(wo[tb]): int(wo[tb]) + rows[tb] * D].view(-1, D)
    nb_before = nb_view.index_select(0, nb_idx).clone()
    total_bits = int(np.log2(hsc[-1]) + 1)
    srt = ops.tbe_sort_indices(d_idx, d_off, d_hsc, total_bits, B, False)
    # the transpose itself: keys ascending, payload a stable argsort (bit-exact integer work)
    keys = srt.keys.cpu().numpy().astype(np.int64) & 0xFFFFFFFF if srt.key_bits <= 32 else srt.keys.cpu().numpy()
    lin = np.concatenate([ids[k] + hsc[k] for k in range(T)])
    order = np.argsort(lin, kind="stable")
    np.testing.assert_array_equal(keys, lin[order])
    np.testing.assert_array_equal(srt.payload.cpu().numpy().astype(np.int64), order)  # L = 1: bag == position
    lr, eps = 0.01, 1e-8
    ops.tbe_backward(t(grad), srt, weights, d_wo, d_Do, d_hsc, B, D, True, d_off, None, 0, 1, mom, t(hsc[:-1].copy()),
                     lr, eps)
    c_oracle.tbe_backward(grad, c_w, c_wo, Do, c_hsc, c_idx, offsets, "rowwise_adagrad", lr, eps, c_m, c_hsc[:-1])
    got_w = _gather_rows(weights, wo, uniq, D)
    got_m = np.concatenate([mom[int(hsc[k]): int(hsc[k + 1])].index_select(0, t(uniq[k])).cpu().numpy()
                            for k in range(T)])
    # Tolerance per table, derived: a run of n gradient rows summed in a different order differs by at most
    # n * eps32 * sum|g| ~ relative n * eps32 of the sum; momentum is quadratic in it (x2), and the weight step
    # lr * G / (sqrt(m) + eps) has magnitude <= lr * sqrt(D), so its absolute error is lr * sqrt(D) * that.
    for k in range(T):
        n_run = int(np.bincount(remapped[k]).max())
        rel = max(1e-5, 4.0 * n_run * EPS32)
        sl_w = slice(int(c_wo[k]), int(c_wo[k]) + c_rows[k] * D)
        sl_m = slice(int(c_hsc[k]), int(c_hsc[k + 1]))
        np.testing.assert_allclose(got_m[sl_m], c_m[sl_m], rtol=2 * rel, atol=1e-12, err_msg=f"momentum1, table {k}")
        np.testing.assert_allclose(got_w[sl_w], c_w[sl_w], rtol=1e-5, atol=lr * np.sqrt(D) * rel,
                                   err_msg=f"weights, table {k} (longest run {n_run})")
    assert torch.equal(nb_view.index_select(0, nb_idx), nb_before), "rows next to updated rows must not change"
    # momentum of untouched rows stays zero (checked on the whole 0.8 GB state)
    touched = torch.zeros_like(mom, dtype=torch.bool)
    for k in range(T):
        touched[int(hsc[k]) + t(uniq[k])] = True
    assert float(mom[~touched].abs().max()) == 0.0
    del weights, mom, nb_view, touched
    torch.cuda.empty_cache()


def test_cfg4_512_tables_fp16_pooling_50():
    from b200_sparse import ops
    _need_free(140)
    dev = torch.device("cuda")
    T, rows_t, D, B, L = 512, 1_000_000, 128, 512, 50
    rows = [rows_t] * T
    hsc = np.concatenate([[0], np.cumsum(rows)]).astype(np.int64)
    wo = (hsc[:-1] * D).astype(np.int64)
    Do = np.arange(0, (T + 1) * D, D, dtype=np.int32)
    weights = torch.empty(int(hsc[-1]) * D, dtype=torch.float16, device=dev)
    chunk = 64 * rows_t * D
    for s in range(0, weights.numel(), chunk):
        weights[s: s + chunk].normal_(0.0, 0.05)
    mom = torch.zeros(int(hsc[-1]), dtype=torch.float32, device=dev)
    rng = np.random.default_rng(4)
    # ids sorted within a bag as fbgemm's benchmark does (bench/split_table_batched_embeddings_benchmark.py:180-189)
    ids = [np.sort(rng.integers(0, rows_t, size=(B, L)), axis=1).reshape(-1).astype(np.int64) for _ in range(T)]
    uniq, remapped = _compact(ids)
    c_rows = [len(u) for u in uniq]
    c_hsc = np.concatenate([[0], np.cumsum(c_rows)]).astype(np.int64)
    c_wo = (c_hsc[:-1] * D).astype(np.int64)
    c_w = _gather_rows(weights, wo, uniq, D)  # fp16 rows widened to fp32: exact
    c_m = np.zeros(int(c_hsc[-1]), dtype=np.float32)
    c_idx = np.concatenate(remapped)
    idx_full = np.concatenate(ids)
    offsets = (np.arange(T * B + 1, dtype=np.int64) * L)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    d_idx, d_off, d_wo, d_Do, d_hsc = t(idx_full), t(offsets), t(wo), t(Do), t(hsc)
    out = ops.tbe_forward(weights, d_wo, d_Do, T * D, D, True, d_idx, d_off, None, 0)
    exp = c_oracle.tbe_forward(c_w, c_wo, Do, c_idx, offsets)
    # fp32 accumulation of 50 exact fp16 values in a different order: rtol 1e-5 with the n*eps*sum|x| floor
    np.testing.assert_allclose(out.cpu().numpy(), exp, rtol=1e-5, atol=L * EPS32 * 0.05 * 8)
    del out
    grad = (rng.standard_normal((B, T * D)) * 0.1).astype(np.float32)
    srt = ops.tbe_sort_indices(d_idx, d_off, d_hsc, int(np.log2(hsc[-1]) + 1), B, False)
    lr, eps = 0.01, 1e-8
    ops.tbe_backward(t(grad), srt, weights, d_wo, d_Do, d_hsc, B, D, True, d_off, None, 0, 1, mom, t(hsc[:-1].copy()),
                     lr, eps, stochastic_rounding=False)
    c_oracle.tbe_backward(grad, c_w, c_wo, Do, c_hsc, c_idx, offsets, "rowwise_adagrad", lr, eps, c_m, c_hsc[:-1])
    got_w = _gather_rows(weights, wo, uniq, D)
    got_m = np.concatenate([mom[int(hsc[k]): int(hsc[k + 1])].index_select(0, t(uniq[k])).cpu().numpy()
                            for k in range(T)])
    np.testing.assert_allclose(got_m, c_m, rtol=1e-5, atol=1e-12)
    # updated rows are rounded to fp16 (round to nearest): half an fp16 ulp of the value + the fp32 slack
    exp_w16 = c_w.astype(np.float16).astype(np.float32)
    np.testing.assert_allclose(got_w, exp_w16, rtol=2.0 ** -10, atol=1e-7)  # north_star: rtol 1e-2 for fp16 tables
    assert float(np.mean(got_w == exp_w16)) > 0.999  # and all but rounding-boundary cases are bit-identical
    del weights, mom
    torch.cuda.empty_cache()


def test_cfg5_int8_26_tables_batch_4096():
    from b200_sparse import ops
    _need_free(20)
    dev = torch.device("cuda")
    D, B = 128, 4096
    rows = [min(r, 2_000_000) for r in CRITEO_ROWS]  # quantised from fp32 on the host: cap the 40 M tables
    T = len(rows)
    rb = R.int8_row_bytes(D)
    assert rb == 144
    rng = np.random.default_rng(5)
    ids = [rng.integers(0, r, size=B).astype(np.int32) for r in rows]
    uniq, remapped = _compact([a.astype(np.int64) for a in ids])
    # the device tables: random bytes everywhere, real quantised rows (the reference's layout: fp16 scale, fp16
    # bias, D x uint8, pad to 16) at the touched ids
    wo = np.concatenate([[0], np.cumsum([(r * rb + 127) // 128 * 128 for r in rows])]).astype(np.int64)
    dev_w = torch.randint(0, 255, (int(wo[-1]),), dtype=torch.uint8, device=dev)
    c_parts, c_wo, pos = [], [], 0
    for k in range(T):
        wf = rng.standard_normal((len(uniq[k]), D)).astype(np.float32)
        q = R.quantize_rowwise_8bit_sbhalf(wf)
        packed = np.zeros((len(uniq[k]), rb), dtype=np.uint8)
        packed[:, :4], packed[:, 4:4 + D] = q[:, D:], q[:, :D]
        view = dev_w[int(wo[k]): int(wo[k]) + rows[k] * rb].view(-1, rb)
        view[torch.from_numpy(uniq[k]).to(dev)] = torch.from_numpy(packed).to(dev)
        c_parts.append(packed.reshape(-1))
        c_wo.append(pos)
        pos += packed.size
    Do = np.arange(0, (T + 1) * D, D, dtype=np.int32)
    offsets = np.arange(T * B + 1, dtype=np.int32)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    exp = R.tbe_int8_forward(np.concatenate(c_parts), np.array(c_wo, np.int64), Do,
                             np.concatenate(remapped).astype(np.int32), offsets, 16, False, None, np.float16)
    got = ops.tbe_int8_forward(dev_w, t(wo[:-1].copy()), t(Do), T * D, D, 16, t(np.concatenate(ids)), t(offsets),
                               None, 0, torch.float16)
    assert got.shape == (B, T * D) and got.dtype == torch.float16
    # L = 1: scale * q + bias in fp32, rounded once to fp16 (one fp16 ulp covers fma-vs-mul+add)
    np.testing.assert_allclose(got.float().cpu().numpy(), exp.astype(np.float32), rtol=2.0 ** -10, atol=1e-4)
