"""GPU parity of b200_itx_launch (inverse transform + reconstruction + pixel-domain distortion,
SURVEY.md section 8f rank 3) against the oracle restatement of av1_inverse_transform_block
(av1/common/idct.c:303) and dist_block_px_domain (av1/encoder/tx_search.c:969), bit-exact, for all
19 sizes x valid types at 8 / 10 / 12 bit, uint8 and uint16 pixels; and chained behind the fused
forward launch on the same descriptors (txq -> itx on the device)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
import synth
from harness import i32p, ptr

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def _oracle_itx(blocks, preds, dq, eob, src, pred, highbd, bd, want_dist=True):
    O = H.oracle()
    O.av1o_dist_block_px.restype = C.c_int64
    recon = pred.copy()
    dist = np.zeros(len(blocks), np.int64)
    for i, (b, p) in enumerate(zip(blocks, preds)):
        ts, tt = int(b["tx_size"]), int(b["tx_type"])
        w, h = H.TX_W[ts], H.TX_H[ts]
        c = np.ascontiguousarray(dq[int(b["out_offset"]):int(b["out_offset"]) + H.max_eob(ts)])
        po, ps = int(p["pred_offset"]), int(p["pred_stride"])
        isz = pred.itemsize
        O.av1o_recon_block(ptr(c, i32p), C.c_void_p(pred.ctypes.data + po * isz), ps,
                           C.c_void_p(recon.ctypes.data + po * isz), ps, ts, tt,
                           int(eob[i]) if eob is not None else 1, 0, highbd, bd)
        vis = int(b["reserved"])
        vw, vh = (vis & 255, vis >> 8) if vis else (w, h)
        if want_dist:
            dist[i] = O.av1o_dist_block_px(C.c_void_p(src.ctypes.data + int(b["src_offset"]) * isz),
                                           int(b["src_stride"]), C.c_void_p(recon.ctypes.data + po * isz), ps,
                                           vw, vh, highbd, bd)
    return recon, dist


def _run_itx(ctx, ts, highbd, bd, blocks, preds, dq, eob, src, pred, in_place=False, want_recon=True):
    import b200av1
    d_blocks, d_preds = b200av1.to_device(ctx, blocks), b200av1.to_device(ctx, preds)
    d_dq = b200av1.to_device(ctx, dq)
    d_eob = b200av1.to_device(ctx, eob) if eob is not None else None
    d_src, d_pred = b200av1.to_device(ctx, src), b200av1.to_device(ctx, pred)
    d_recon = d_pred if in_place else b200av1.to_device(ctx, np.zeros_like(pred))
    d_dist = b200av1.to_device(ctx, np.full(len(blocks), -1, np.int64))
    l = b200av1.ItxLaunch(ts, highbd, bd, len(blocks), d_blocks.ptr, d_preds.ptr, d_dq.ptr,
                          d_eob.ptr if d_eob else None, d_src.ptr, d_pred.ptr,
                          d_recon.ptr if want_recon else None, d_dist.ptr)
    b200av1.check(b200av1.lib().b200_itx_launch(ctx.h, C.byref(l), None))
    ctx.sync()
    recon = b200av1.to_host(ctx, d_recon, pred.dtype, pred.size).reshape(pred.shape)
    dist = b200av1.to_host(ctx, d_dist, np.int64, len(blocks))
    return recon, dist


def _coeff_block(rng, ts, tt, bd, mode):
    n = H.max_eob(ts)
    w, h = H.TX_W[ts], H.TX_H[ts]
    if mode == 0:
        resid = rng.integers(-(1 << bd) + 1, 1 << bd, (h, w)).astype(np.int16)
        c = H.oracle_fwd_txfm2d(resid, w, ts, tt, bd)
        q = int(rng.integers(1, 200))
        return ((c // q) * q).astype(np.int32)
    if mode == 1:
        lim = 1 << (bd + 7)
        return rng.integers(-lim, lim, n).astype(np.int32)
    if mode == 2:
        c = np.zeros(n, np.int32)
        k = int(rng.integers(1, 4))
        c[rng.integers(0, n, k)] = rng.integers(-(1 << (bd + 6)), 1 << (bd + 6), k)
        return c
    return rng.integers(-(1 << 31), (1 << 31) - 1, n).astype(np.int32)  # any int32: exercises the clamps


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
@pytest.mark.parametrize("ts", range(H.N_TX_SIZES))
def test_itx_launch_all_types_vs_oracle(ctx, ts, highbd, bd):
    import b200av1
    rng = np.random.default_rng(8100 + ts * 4 + bd + highbd)
    w, h, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
    types = [t for t in range(16) if H.tx_valid(ts, t)]
    nb = 3 * len(types) + 5  # ragged against the warp tile
    dt = np.uint16 if highbd else np.uint8
    stride_s, stride_p = w * 3 + 16, w * 3 + 48
    rows = (nb + 2) // 3
    src = rng.integers(0, 1 << bd, (rows * h, stride_s)).astype(dt)
    pred = rng.integers(0, 1 << bd, (rows * h, stride_p)).astype(dt)
    pred[: h] = (1 << bd) - 1  # saturating rows
    pred[h: 2 * h] = 0
    blocks = np.zeros(nb, b200av1.TXBLOCK_DTYPE)
    preds = np.zeros(nb, b200av1.TXPRED_DTYPE)
    dq = np.zeros(nb * n + 7, np.int32)
    eob = np.zeros(nb, np.uint16)
    for i in range(nb):
        tt = types[i % len(types)]
        r, c = i // 3, i % 3
        vis = 0
        if i % 5 == 4:
            vis = max(w - 4 * (i % 3), 4) | (max(h - 4 * ((i // 3) % 3), 4) << 8)
        blocks[i] = (r * h * stride_s + c * w, i * n + 3, stride_s, 0, ts, tt, vis)
        preds[i] = (r * h * stride_p + c * w, stride_p, 0)
        cb = _coeff_block(rng, ts, tt, bd, i % 4)
        eob[i] = 0 if i % 7 == 6 else max(1, int(np.count_nonzero(cb)))
        dq[i * n + 3: i * n + 3 + n] = cb
    want_recon, want_dist = _oracle_itx(blocks, preds, dq, eob, src.reshape(-1), pred.reshape(-1), highbd, bd)
    got_recon, got_dist = _run_itx(ctx, ts, highbd, bd, blocks, preds, dq, eob, src, pred)
    mask = np.zeros(pred.shape, bool)
    for i in range(nb):
        r, c = i // 3, i % 3
        mask[r * h:(r + 1) * h, c * w:(c + 1) * w] = True
    assert np.array_equal(got_recon[mask], want_recon.reshape(pred.shape)[mask]), H.TX_NAMES[ts]
    assert not got_recon[~mask].any(), "wrote outside the blocks"
    assert np.array_equal(got_dist, want_dist), H.TX_NAMES[ts]
    # in place over the prediction plane, no eob array, distortion only
    want2, dist2 = _oracle_itx(blocks, preds, dq, None, src.reshape(-1), pred.reshape(-1), highbd, bd)
    got2, gdist2 = _run_itx(ctx, ts, highbd, bd, blocks, preds, dq, None, src, pred, in_place=True)
    assert np.array_equal(got2[mask], want2.reshape(pred.shape)[mask])
    assert np.array_equal(gdist2, dist2)


@pytest.mark.parametrize("bd", [8, 10])
def test_txq_then_itx_chained_on_device(ctx, bd):
    """forward + quantise launch, then the inverse launch on the SAME descriptor arrays and the dqcoeff /
    eob the first launch left on the device: dist = dist_block_px_domain of every block."""
    import b200av1
    rng = np.random.default_rng(8400 + bd)
    w, h, stride = 128, 64, 160
    dt = np.uint8 if bd == 8 else np.uint16
    srcp = np.zeros((h, stride), dt)
    srcp[:, :w] = synth.smooth_frame(rng, w, h, bd)
    predp = np.zeros((h, stride), dt)
    predp[:, :w] = synth.pred_from(rng, srcp[:, :w], bd)
    rows = [H.oracle_qrow(bd, q) for q in (60, 200)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    q_kind = b200av1.Q_FP if bd == 8 else b200av1.Q_FP_HBD
    in_kind = b200av1.IN_PIXELS8 if bd == 8 else b200av1.IN_PIXELS16
    for ts in (0, 2, 3, 4, 6, 9, 17):
        W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
        offs = synth.tile_blocks(w, h, stride, W, Hh)
        types = [t for t in range(16) if H.tx_valid(ts, t)]
        blocks = np.zeros(len(offs), b200av1.TXBLOCK_DTYPE)
        preds = np.zeros(len(offs), b200av1.TXPRED_DTYPE)
        for i in range(len(offs)):
            blocks[i] = (offs[i], i * n, stride, i % 2, ts, types[i % len(types)], 0)
            preds[i] = (offs[i], stride, 0)
        out_elems = len(offs) * n
        d_blocks, d_preds = b200av1.to_device(ctx, blocks), b200av1.to_device(ctx, preds)
        d_src, d_pred = b200av1.to_device(ctx, srcp), b200av1.to_device(ctx, predp)
        d_q = b200av1.to_device(ctx, qrows)
        d_qc, d_dq = b200av1.DevBuf(ctx, out_elems * 4), b200av1.DevBuf(ctx, out_elems * 4)
        d_eob = b200av1.DevBuf(ctx, len(blocks) * 2)
        d_dist = b200av1.DevBuf(ctx, len(blocks) * 8)
        d_recon = b200av1.to_device(ctx, np.zeros_like(predp))
        fl = b200av1.TxqLaunch(ts, in_kind, q_kind, bd, 0, len(blocks), d_blocks.ptr, d_preds.ptr,
                               d_src.ptr, d_pred.ptr, d_q.ptr, None, d_qc.ptr, d_dq.ptr, d_eob.ptr, None)
        ctx.txq_launch(fl)
        il = b200av1.ItxLaunch(ts, int(bd > 8), bd, len(blocks), d_blocks.ptr, d_preds.ptr, d_dq.ptr,
                               d_eob.ptr, d_src.ptr, d_pred.ptr, d_recon.ptr, d_dist.ptr)
        b200av1.check(b200av1.lib().b200_itx_launch(ctx.h, C.byref(il), None))
        ctx.sync()
        got_dist = b200av1.to_host(ctx, d_dist, np.int64, len(blocks))
        got_recon = b200av1.to_host(ctx, d_recon, dt, predp.size).reshape(predp.shape)
        r = H.oracle_txq_batch(srcp.reshape(-1), blocks, rows, q_kind, bd, out_elems, in_kind,
                               predp.reshape(-1), preds)
        want_recon, want_dist = _oracle_itx(blocks, preds, r[2], r[3], srcp.reshape(-1), predp.reshape(-1),
                                            int(bd > 8), bd)
        assert np.array_equal(got_dist, want_dist), H.TX_NAMES[ts]
        assert np.array_equal(got_recon[:, :w], want_recon.reshape(predp.shape)[:, :w]), H.TX_NAMES[ts]
        # sanity: reconstruction error is far below the prediction error at these quantisers
        assert got_dist.sum() < 16 * ((srcp.astype(np.int64) - predp) ** 2).sum() or bd > 8


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 10), (1, 12)])
def test_itx_lossless_wht_vs_oracle(ctx, highbd, bd):
    """lossless blocks: av1_highbd_iwht4x4_16_add / _1_add by eob (av1_inv_txfm2d.c:20-109)"""
    import b200av1
    O = H.oracle()
    O.av1o_dist_block_px.restype = C.c_int64
    rng = np.random.default_rng(8600 + bd)
    nb, stride = 301, 4 * 301 + 8
    dt = np.uint16 if highbd else np.uint8
    src = rng.integers(0, 1 << bd, (4, stride)).astype(dt)
    pred = rng.integers(0, 1 << bd, (4, stride)).astype(dt)
    blocks = np.zeros(nb, b200av1.TXBLOCK_DTYPE)
    preds = np.zeros(nb, b200av1.TXPRED_DTYPE)
    dq = rng.integers(-(1 << (bd + 3)), 1 << (bd + 3), nb * 16).astype(np.int32)
    eob = rng.integers(0, 17, nb).astype(np.uint16)
    eob[:3] = (0, 1, 2)
    want = pred.copy()
    want_dist = np.zeros(nb, np.int64)
    for i in range(nb):
        blocks[i] = (4 * i, 16 * i, stride, 0, 0, 0, 0)
        preds[i] = (4 * i, stride, 0)
        c = np.ascontiguousarray(dq[16 * i:16 * i + 16])
        isz = pred.itemsize
        O.av1o_recon_block(ptr(c, i32p), C.c_void_p(pred.ctypes.data + 4 * i * isz), stride,
                           C.c_void_p(want.ctypes.data + 4 * i * isz), stride, 0, 0, int(eob[i]), 1, highbd, bd)
        want_dist[i] = O.av1o_dist_block_px(C.c_void_p(src.ctypes.data + 4 * i * isz), stride,
                                            C.c_void_p(want.ctypes.data + 4 * i * isz), stride, 4, 4, highbd, bd)
    d = [b200av1.to_device(ctx, x) for x in (blocks, preds, dq, eob, src, pred)]
    d_recon = b200av1.to_device(ctx, np.zeros_like(pred))
    d_dist = b200av1.DevBuf(ctx, nb * 8)
    l = b200av1.ItxLaunch(0, highbd, bd, nb, d[0].ptr, d[1].ptr, d[2].ptr, d[3].ptr, d[4].ptr, d[5].ptr,
                          d_recon.ptr, d_dist.ptr, 1)
    b200av1.check(b200av1.lib().b200_itx_launch(ctx.h, C.byref(l), None))
    ctx.sync()
    got = b200av1.to_host(ctx, d_recon, dt, pred.size).reshape(pred.shape)
    assert np.array_equal(got[:, :4 * nb], want[:, :4 * nb])
    assert np.array_equal(b200av1.to_host(ctx, d_dist, np.int64, nb), want_dist)


class TxfmParam(C.Structure):  # aom_dsp/txfm_common.h:89-101 == B200TxfmParam
    _fields_ = [("tx_type", C.c_uint8), ("tx_size", C.c_uint8), ("lossless", C.c_int), ("bd", C.c_int),
                ("is_hbd", C.c_int), ("tx_set_type", C.c_uint8), ("eob", C.c_int)]


def test_inverse_dropins_vs_oracle(ctx):
    """av1_inv_txfm2d_add_WxH_b200, av1_inv_txfm_add_b200, av1_highbd_inv_txfm_add_b200 (one block per
    call, host buffers) and the lossless route of av1_lowbd_fwd_txfm_b200"""
    import b200av1
    from harness import i16p, u8p, u16p
    L, O = b200av1.lib(), H.oracle()
    rng = np.random.default_rng(8700)
    for ts in range(H.N_TX_SIZES):
        w, h = H.TX_W[ts], H.TX_H[ts]
        stride = w + 8
        types = [t for t in range(16) if H.tx_valid(ts, t)]
        for tt in (types[0], types[-1], types[len(types) // 2]):
            for bd in (8, 10, 12):
                c = _coeff_block(rng, ts, tt, bd, int(rng.integers(0, 3)))
                pred = rng.integers(0, 1 << bd, (h, stride)).astype(np.uint16)
                a, b = pred.copy(), pred.copy()
                getattr(L, "av1_inv_txfm2d_add_%dx%d_b200" % (w, h))(ptr(c, i32p), ptr(a, u16p), stride,
                                                                     C.c_uint8(tt), bd)
                O.av1o_inv_txfm2d_add(ptr(c, i32p), ptr(b, u16p), stride, ts, tt, bd)
                assert np.array_equal(a, b), (ts, tt, bd)
            c = _coeff_block(rng, ts, tt, 8, 0)
            pred8 = rng.integers(0, 256, (h, stride)).astype(np.uint8)
            a8, b8 = pred8.copy(), pred8.copy()
            p = TxfmParam(tt, ts, 0, 8, 0, 0, int(np.count_nonzero(c)) or 1)
            L.av1_inv_txfm_add_b200(ptr(c, i32p), ptr(a8, u8p), stride, C.byref(p))
            O.av1o_recon_block(ptr(c, i32p), pred8.ctypes.data_as(C.c_void_p), stride,
                               b8.ctypes.data_as(C.c_void_p), stride, ts, tt, 1, 0, 0, 8)
            assert np.array_equal(a8[:, :w], b8[:, :w]) and np.array_equal(a8[:, w:], pred8[:, w:])
            pred16 = rng.integers(0, 1024, (h, stride)).astype(np.uint16)
            a16, b16 = pred16.copy(), pred16.copy()
            c = _coeff_block(rng, ts, tt, 10, 0)
            p = TxfmParam(tt, ts, 0, 10, 1, 0, 1)
            L.av1_highbd_inv_txfm_add_b200(ptr(c, i32p), C.cast(C.c_void_p(a16.ctypes.data >> 1), u8p), stride,
                                           C.byref(p))
            O.av1o_inv_txfm2d_add(ptr(c, i32p), ptr(b16, u16p), stride, ts, tt, 10)
            assert np.array_equal(a16, b16)
    # lossless: forward WHT through av1_lowbd_fwd_txfm_b200, inverse through av1_inv_txfm_add_b200
    resid = rng.integers(-255, 256, (4, 12)).astype(np.int16)
    co, cw = np.zeros(16, np.int32), np.zeros(16, np.int32)
    p = TxfmParam(0, 0, 1, 8, 0, 0, 16)
    L.av1_lowbd_fwd_txfm_b200(ptr(resid, i16p), ptr(co, i32p), 12, C.byref(p))
    O.av1o_fwht4x4(ptr(resid, i16p), ptr(cw, i32p), 12)
    assert np.array_equal(co, cw)
    pred8 = rng.integers(0, 256, (4, 12)).astype(np.uint8)
    a8, b8 = pred8.copy(), pred8.copy()
    L.av1_inv_txfm_add_b200(ptr(co, i32p), ptr(a8, u8p), 12, C.byref(p))
    O.av1o_recon_block(ptr(co, i32p), pred8.ctypes.data_as(C.c_void_p), 12, b8.ctypes.data_as(C.c_void_p), 12,
                       0, 0, 16, 1, 0, 8)
    assert np.array_equal(a8[:, :4], b8[:, :4])


@pytest.mark.parametrize("bd", [8, 10])
def test_txsearch_batch_host_vs_oracle(ctx, bd):
    """b200_txsearch_batch_host: every valid tx type of every block as one descriptor; eob, tx-domain
    error, Laplacian rate estimate and pixel-domain distortion against the oracle, caller's order kept."""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(8800 + bd)
    w, h, stride = 128, 64, 144
    dt = np.uint8 if bd == 8 else np.uint16
    srcp = np.zeros((h, stride), dt)
    srcp[:, :w] = synth.smooth_frame(rng, w, h, bd)
    predp = np.zeros((h, stride), dt)
    predp[:, :w] = synth.pred_from(rng, srcp[:, :w], bd)
    rows = [H.oracle_qrow(bd, q) for q in (30, 110, 220)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    q_kind = b200av1.Q_B if bd == 8 else b200av1.Q_B_HBD  # prune_txk_type_separ quantises with _B
    in_kind = b200av1.IN_PIXELS8 if bd == 8 else b200av1.IN_PIXELS16
    recs, precs = [], []
    for ts in (0, 5, 6, 1, 7, 8, 2, 13, 14, 3, 4):
        W, Hh = H.TX_W[ts], H.TX_H[ts]
        for k, off in enumerate(synth.tile_blocks(w, h, stride, W, Hh)[:6]):
            for tt in range(16):
                if H.tx_valid(ts, tt):
                    recs.append((off, 0, stride, (k + tt) % 3, ts, tt, 0))
                    precs.append((off, stride, 0))
    blocks = np.array(recs, b200av1.TXBLOCK_DTYPE)
    preds = np.array(precs, b200av1.TXPRED_DTYPE)
    perm = rng.permutation(len(blocks))
    blocks, preds = blocks[perm], preds[perm]
    got, d2h = b200av1.txsearch_batch_host(ctx, in_kind, q_kind, bd, 7, blocks, preds, srcp, predp, qrows)
    assert d2h == len(blocks) * 30
    # oracle: dense outputs at offsets of our own choosing
    ob = blocks.copy()
    o = 0
    for i in range(len(ob)):
        ob[i]["out_offset"] = o
        o += H.max_eob(int(ob[i]["tx_size"]))
    r = H.oracle_txq_batch(srcp.reshape(-1), ob, rows, q_kind, bd, o, in_kind, predp.reshape(-1), preds)
    coeff, qc, dq, eob, dist = r
    assert np.array_equal(got["eob"], eob)
    assert np.array_equal(got["dist_tx"], dist[:, 0]) and np.array_equal(got["ssz"], dist[:, 1])
    _, want_px = _oracle_itx(ob, preds, dq, eob, srcp.reshape(-1), predp.reshape(-1), int(bd > 8), bd)
    assert np.array_equal(got["dist_px"], want_px)
    for i in range(len(ob)):
        ts, tt = int(ob[i]["tx_size"]), int(ob[i]["tx_type"])
        want = 0
        if eob[i]:
            q = np.ascontiguousarray(qc[int(ob[i]["out_offset"]):int(ob[i]["out_offset"]) + H.max_eob(ts)])
            want = O.av1o_cost_coeffs_txb_estimate(ptr(q, i32p), int(eob[i]), ts, tt)
        assert got["rate_est"][i] == want, (i, ts, tt)
    assert (eob > 0).any() and (eob == 0).any() or bd > 8
