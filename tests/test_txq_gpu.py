"""GPU parity: fused transform + quantise (+ block error) through the C ABI vs the
oracle, bit-exact.  Inputs replay the reference tests' distributions
(test/av1_fwd_txfm2d_test.cc:243-297,518-576; test/quantize_func_test.cc:209-259)."""
import numpy as np
import pytest

import harness as H
import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def _make_batch(rng, ts, n_per_type, bd, mode, qindexes):
    import b200av1
    W, Hh = H.TX_W[ts], H.TX_H[ts]
    n = H.max_eob(ts)
    types = [t for t in range(16) if H.tx_valid(ts, t)]
    nb = n_per_type * len(types)
    stride = W * 3 + 5  # deliberately odd pitch: blocks side by side with gaps
    rows_per_block = Hh + 1
    hi = (1 << bd) - 1
    src = np.zeros(nb // 3 * rows_per_block * stride + rows_per_block * stride * 2 + 64, np.int16)
    blocks = np.zeros(nb, b200av1.TXBLOCK_DTYPE)
    for i in range(nb):
        brow, bcol = divmod(i, 3)
        off = brow * rows_per_block * stride + bcol * W + (i % 2)
        blk = np.lib.stride_tricks.as_strided(src[off:], (Hh, W), (stride * 2, 2))
        if mode == "max":
            blk[:] = hi if i % 2 == 0 else -hi
        elif mode == "small":
            blk[:] = rng.integers(-40, 41, (Hh, W))
        elif mode == "wild":  # non-conformant: any int16 (exercises the exact-path guard)
            blk[:] = rng.integers(-32768, 32768, (Hh, W))
        elif mode == "edge":  # just inside / just outside the guard
            blk[:] = rng.choice(np.array([-4095, -4094, 4094, 4095, 0, 1]), (Hh, W))
        else:
            blk[:] = rng.integers(-hi, hi + 1, (Hh, W))
        blocks[i] = (off, i * n, stride, i % len(qindexes), ts, types[i % len(types)], 0)
    rng.shuffle(blocks)  # tx types and offsets in arbitrary order
    return src, blocks, nb * n


CASES = [(8, "rand", 1), (8, "max", 1), (8, "small", 1), (10, "rand", 2), (12, "rand", 2),
         (12, "max", 2), (8, "rand", 3), (10, "small", 4), (12, "rand", 4)]


@pytest.mark.parametrize("ts", range(19), ids=H.TX_NAMES)
def test_txq_all_types(ctx, ts):
    import b200av1
    rng = np.random.default_rng(1000 + ts)
    qidx = [0, 1, 20, 60, 128, 200, 255]
    for bd, mode, q_kind in CASES:
        rows = [H.oracle_qrow(bd, q) for q in qidx]
        qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
        src, blocks, out_elems = _make_batch(rng, ts, 3 if H.TX_W[ts] * H.TX_H[ts] >= 1024 else 6,
                                             bd, mode, qidx)
        for fast in (0, b200av1.TXQ_FORCE_EXACT):
            coeff = np.full(out_elems, 123, np.int32)
            qc = np.full(out_elems, 123, np.int32)
            dq = np.full(out_elems, 123, np.int32)
            eob = np.full(len(blocks), 9999, np.uint16)
            dist = np.full((len(blocks), 2), -1, np.int64)
            flags = b200av1.TXQ_STORE_COEFF | b200av1.TXQ_BLOCK_ERROR | fast
            ctx.txq_batch_host(b200av1.IN_RESIDUAL, q_kind, bd, flags, blocks, None, src, None,
                               qrows, coeff, qc, dq, eob, dist)
            r = H.oracle_txq_batch(src, blocks, rows, q_kind, bd, out_elems)
            tag = (H.TX_NAMES[ts], bd, mode, q_kind, fast)
            assert np.array_equal(coeff, r[0]), tag
            assert np.array_equal(qc, r[1]), tag
            assert np.array_equal(dq, r[2]), tag
            assert np.array_equal(eob, r[3]), tag
            assert np.array_equal(dist, r[4]), tag


@pytest.mark.parametrize("ts", range(19), ids=H.TX_NAMES)
def test_txfm_nonconformant_input_takes_exact_path(ctx, ts):
    """any int16 residual (far outside what a bd-bit encoder produces) must still match the
    reference's wrapping 32-bit products / 64-bit sums: the in-kernel guard routes such tiles
    to the literal half_btf form.  Also +-4094/4095 around the guard threshold."""
    import b200av1
    rng = np.random.default_rng(4000 + ts)
    for mode in ("wild", "edge"):
        src, blocks, out_elems = _make_batch(rng, ts, 4, 12, mode, [0])
        coeff = np.zeros(out_elems, np.int32)
        ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_NONE, 12, b200av1.TXQ_STORE_COEFF,
                           blocks, None, src, None, None, coeff, None, None, None, None)
        r = H.oracle_txq_batch(src, blocks, [], 0, 12, out_elems)
        assert np.array_equal(coeff, r[0]), (H.TX_NAMES[ts], mode)


def test_txq_unaligned_sources(ctx):
    """source offsets / strides that defeat the 16-byte cp.async path (element-wise copy)"""
    import b200av1
    rng = np.random.default_rng(77)
    for ts in (0, 1, 2, 3, 6, 8, 12):
        W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
        stride = W + 3
        src = rng.integers(-255, 256, (Hh + 2) * stride * 6 + 16).astype(np.int16)
        blocks = np.zeros(6, b200av1.TXBLOCK_DTYPE)
        for i in range(6):
            blocks[i] = (i * (Hh + 2) * stride + 1 + i, i * n, stride, 0, ts, 0, 0)
        rows = [H.oracle_qrow(8, 77)]
        qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
        coeff, qc, dq = (np.zeros(6 * n, np.int32) for _ in range(3))
        eob = np.zeros(6, np.uint16)
        ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_FP, 8, b200av1.TXQ_STORE_COEFF, blocks,
                           None, src, None, qrows, coeff, qc, dq, eob, None)
        r = H.oracle_txq_batch(src, blocks, rows, 1, 8, 6 * n, with_err=False)
        for a, b in zip((coeff, qc, dq, eob), r):
            assert np.array_equal(a, b), H.TX_NAMES[ts]


def test_txq_transform_only_and_empty(ctx):
    import b200av1
    rng = np.random.default_rng(5)
    src, blocks, out_elems = _make_batch(rng, 2, 2, 8, "rand", [0])
    coeff = np.zeros(out_elems, np.int32)
    ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_NONE, 8, b200av1.TXQ_STORE_COEFF, blocks,
                       None, src, None, None, coeff, None, None, None, None)
    r = H.oracle_txq_batch(src, blocks, [], 0, 8, out_elems)
    assert np.array_equal(coeff, r[0])
    # empty batch is a no-op
    ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_NONE, 8, b200av1.TXQ_STORE_COEFF,
                       blocks[:0], None, src, None, None, coeff, None, None, None, None)


def test_txq_mixed_sizes_one_call(ctx):
    """b200_txq_batch_host groups a ragged batch of mixed tx sizes into per-size launches"""
    import b200av1
    rng = np.random.default_rng(6)
    parts, srcs, off_src, off_out = [], [], 0, 0
    for ts in (0, 5, 1, 13, 2, 9, 3, 17, 4):
        src, blocks, out_elems = _make_batch(rng, ts, 1, 8, "rand", [0, 1, 2])
        blocks["src_offset"] += off_src
        blocks["out_offset"] += off_out
        off_src += len(src)
        off_out += out_elems
        parts.append(blocks)
        srcs.append(src)
    blocks = np.concatenate(parts)
    rng.shuffle(blocks)
    src = np.concatenate(srcs)
    rows = [H.oracle_qrow(8, q) for q in (10, 90, 220)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    coeff, qc, dq = (np.zeros(off_out, np.int32) for _ in range(3))
    eob = np.zeros(len(blocks), np.uint16)
    dist = np.zeros((len(blocks), 2), np.int64)
    ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_FP, 8,
                       b200av1.TXQ_STORE_COEFF | b200av1.TXQ_BLOCK_ERROR, blocks, None, src, None,
                       qrows, coeff, qc, dq, eob, dist)
    r = H.oracle_txq_batch(src, blocks, rows, 1, 8, off_out)
    for a, b in zip((coeff, qc, dq, eob, dist), r):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("bd", [8, 10])
def test_txq_fused_subtract_from_planes(ctx, bd):
    """pixels in (src, pred planes) -> residual formed in-kernel (aom_subtract_block)"""
    import b200av1
    rng = np.random.default_rng(7 + bd)
    w, h, stride = 192, 128, 224
    srcp = np.zeros((h, stride), np.uint8 if bd == 8 else np.uint16)
    srcp[:, :w] = synth.smooth_frame(rng, w, h, bd)
    predp = np.zeros((h, stride + 32), srcp.dtype)
    predp[:, :w] = synth.pred_from(rng, srcp[:, :w], bd)
    src, pred = srcp.reshape(-1), predp.reshape(-1)
    rows = [H.oracle_qrow(bd, q) for q in (40, 120)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    for ts in (0, 1, 2, 3, 4, 7, 10, 12):
        W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
        offs = synth.tile_blocks(w, h, stride, W, Hh)
        poffs = synth.tile_blocks(w, h, stride + 32, W, Hh)
        types = [t for t in range(16) if H.tx_valid(ts, t)]
        blocks = np.zeros(len(offs), b200av1.TXBLOCK_DTYPE)
        preds = np.zeros(len(offs), b200av1.TXPRED_DTYPE)
        for i in range(len(offs)):
            blocks[i] = (offs[i], i * n, stride, i % 2, ts, types[i % len(types)], 0)
            preds[i] = (poffs[i], stride + 32, 0)
        out_elems = len(offs) * n
        coeff, qc, dq = (np.zeros(out_elems, np.int32) for _ in range(3))
        eob = np.zeros(len(blocks), np.uint16)
        dist = np.zeros((len(blocks), 2), np.int64)
        q_kind = b200av1.Q_FP if bd == 8 else b200av1.Q_FP_HBD
        in_kind = b200av1.IN_PIXELS8 if bd == 8 else b200av1.IN_PIXELS16
        ctx.txq_batch_host(in_kind, q_kind, bd, b200av1.TXQ_STORE_COEFF | b200av1.TXQ_BLOCK_ERROR,
                           blocks, preds, src, pred, qrows, coeff, qc, dq, eob, dist)
        r = H.oracle_txq_batch(src, blocks, rows, q_kind, bd, out_elems, in_kind, pred, preds)
        for a, b in zip((coeff, qc, dq, eob, dist), r):
            assert np.array_equal(a, b), (H.TX_NAMES[ts], bd)


@pytest.mark.parametrize("q_kind", [1, 2, 3, 4], ids=["fp", "fp_hbd", "b", "b_hbd"])
def test_txq_compact_expands_to_dense_oracle(ctx, q_kind):
    """b200_txq_batch_host_compact: eob + packed levels in scan order (int16, int32 for the
    high-bit-depth quantisers) + offsets; b200_txq_expand_block(compact) must equal the oracle's
    dense qcoeff / dqcoeff bit for bit (dqcoeff = (|q| * dequant[rc != 0]) >> log_scale,
    av1_quantize.c:61-63), shuffled mixed-size queue, with and without the dense coeff."""
    import b200av1
    bd = 8 if q_kind in (1, 3) else 12
    rng = np.random.default_rng(6100 + q_kind)
    qidx = [0, 40, 128, 255]
    parts, srcs, off_src, off_out = [], [], 0, 0
    for ts in range(19):
        mode = ("rand", "small", "max")[ts % 3]
        src, blocks, out_elems = _make_batch(rng, ts, 2, bd, mode, qidx)
        blocks["src_offset"] += off_src
        blocks["out_offset"] += off_out
        off_src += len(src)
        off_out += out_elems
        parts.append(blocks)
        srcs.append(src)
    blocks = np.concatenate(parts)
    rng.shuffle(blocks)
    src = np.concatenate(srcs)
    rows = [H.oracle_qrow(bd, q) for q in qidx]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    r = H.oracle_txq_batch(src, blocks, rows, q_kind, bd, off_out)
    wide = q_kind in (2, 4)
    for with_coeff in (0, 1):
        levels = np.full(off_out, 0x5a5a, np.int32 if wide else np.int16)
        loff = np.zeros(len(blocks) + 1, np.uint32)
        eob = np.zeros(len(blocks), np.uint16)
        dist = np.zeros((len(blocks), 2), np.int64)
        coeff = np.full(off_out, 7, np.int32) if with_coeff else None
        flags = b200av1.TXQ_BLOCK_ERROR | (b200av1.TXQ_STORE_COEFF if with_coeff else 0)
        d2h = ctx.txq_batch_host_compact(b200av1.IN_RESIDUAL, q_kind, bd, flags, blocks, None, src,
                                         None, qrows, off_out, levels, loff, eob, dist, coeff)
        assert np.array_equal(eob, r[3])
        assert np.array_equal(dist, r[4])
        assert int(loff[-1]) == int(eob.astype(np.int64).sum())
        qc, dq = b200av1.expand_compact(blocks, qrows, levels, loff, eob, off_out, wide)
        assert np.array_equal(qc, r[1])
        assert np.array_equal(dq, r[2])
        if with_coeff:
            assert np.array_equal(coeff, r[0])
        dense_bytes = off_out * 12 + len(blocks) * 2
        assert d2h < dense_bytes
    # a levels buffer that is too small is refused, not overrun
    small = np.zeros(max(1, int(loff[-1]) // 2), levels.dtype)
    with pytest.raises(b200av1.B200Error):
        ctx.txq_batch_host_compact(b200av1.IN_RESIDUAL, q_kind, bd, 0, blocks, None, src, None, qrows,
                                   off_out, small, loff, eob)


def test_txq_batch_host_rejects_out_of_bounds_descriptors(ctx):
    """descriptors that read outside src / qrows are refused on the host (never launched)"""
    import b200av1
    rng = np.random.default_rng(9)
    src, blocks, out_elems = _make_batch(rng, 1, 1, 8, "rand", [0])
    rows = [H.oracle_qrow(8, 50)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    coeff, qc, dq = (np.zeros(out_elems, np.int32) for _ in range(3))
    eob = np.zeros(len(blocks), np.uint16)
    for field, val in (("src_offset", len(src)), ("qrow", 3)):
        bad = blocks.copy()
        bad[field][0] = val
        with pytest.raises(b200av1.B200Error):
            ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_FP, 8, b200av1.TXQ_STORE_COEFF, bad, None,
                               src, None, qrows, coeff, qc, dq, eob, None)


def test_txq_batch_host_overlapping_blocks_take_the_safe_path(ctx):
    """two blocks sharing an out_offset plus one gap have the 'dense' range length; the slot
    bitmap must detect it so the gap reads back as zero, not as uninitialised scratch"""
    import b200av1
    rng = np.random.default_rng(10)
    n = H.max_eob(1)
    src = rng.integers(-255, 256, 8 * 64).astype(np.int16)
    blocks = np.zeros(3, b200av1.TXBLOCK_DTYPE)
    for i, oo in enumerate((0, 0, 2 * n)):   # slot 1 is never written
        blocks[i] = (i * 8, oo, 64, 0, 1, 0, 0)
    rows = [H.oracle_qrow(8, 50)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    for _ in range(2):  # second call: the device scratch holds the first call's data
        coeff, qc, dq = (np.full(3 * n, 99, np.int32) for _ in range(3))
        eob = np.zeros(3, np.uint16)
        ctx.txq_batch_host(b200av1.IN_RESIDUAL, b200av1.Q_FP, 8, b200av1.TXQ_STORE_COEFF, blocks, None,
                           src, None, qrows, coeff, qc, dq, eob, None)
        assert not coeff[n:2 * n].any() and not qc[n:2 * n].any() and not dq[n:2 * n].any()
