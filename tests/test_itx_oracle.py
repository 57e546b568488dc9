"""Pin the inverse-transform / reconstruction / pixel-domain-distortion oracle (oracle/av1o_itxfm.c)
against the reference's own C functions (oracle/_ref): av1_idct*/av1_iadst*/av1_iidentity*
(av1/common/av1_inv_txfm1d.c), av1_inv_txfm2d_add_WxH_c (av1_inv_txfm2d.c), av1_inv_txfm_add_c /
av1_highbd_inv_txfm_add_c incl. the lossless WHT (idct.c), and the SSE behind dist_block_px_domain
(tx_search.c:969).  Input distributions follow test/av1_inv_txfm2d_test.cc / av1_inv_txfm1d_test.cc
(forward-transformed random residuals) plus full-range coefficients for the clamp rules."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, ptr, u8p, u16p

pytestmark = [pytest.mark.ref,
              pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]


class TxfmParam(C.Structure):  # aom_dsp/txfm_common.h:89-101
    _fields_ = [("tx_type", C.c_uint8), ("tx_size", C.c_uint8), ("lossless", C.c_int), ("bd", C.c_int),
                ("is_hbd", C.c_int), ("tx_set_type", C.c_uint8), ("eob", C.c_int)]


def byteptr(a16):
    return C.cast(C.c_void_p(a16.ctypes.data >> 1), u8p)


@pytest.mark.parametrize("name,kind,n", [("av1_idct4", 0, 4), ("av1_idct8", 0, 8), ("av1_idct16", 0, 16),
                                         ("av1_idct32", 0, 32), ("av1_idct64", 0, 64),
                                         ("av1_iadst4", 1, 4), ("av1_iadst8", 1, 8), ("av1_iadst16", 1, 16),
                                         ("av1_iidentity4_c", 2, 4), ("av1_iidentity8_c", 2, 8),
                                         ("av1_iidentity16_c", 2, 16), ("av1_iidentity32_c", 2, 32)])
def test_inv_txfm1d(name, kind, n):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(7000 + n + kind)
    for it in range(400):
        rb = [16, 18, 20][it % 3]
        # inputs of a pass are clamped to the stage range before the kernel runs (20 bits = bd 12
        # rows: the 64-bit sum of half_btf is needed there, single products still fit int32)
        lim = 1 << (rb - 1)
        x = rng.integers(-lim, lim, n).astype(np.int32)
        if it < 3:
            x[:] = lim - 1
        elif it < 6:
            x[:] = -lim
        elif it < 9:
            x[:] = 0
        sr = (C.c_int8 * 12)(*([rb] * 12))
        a = np.zeros(n, np.int32)
        b = np.zeros(n, np.int32)
        getattr(R, name)(ptr(x, i32p), ptr(a, i32p), C.c_int8(12), sr)
        O.av1o_inv_txfm1d(kind, n, ptr(x, i32p), ptr(b, i32p), rb)
        assert np.array_equal(a, b), (name, rb, it)


def _coeffs(rng, tx_size, tx_type, bd, mode):
    """dqcoeff-like input: mode 0 = forward transform of a random residual, coarsely quantised;
    mode 1 = full range of the row-pass clamp (bd + 8 bits); mode 2 = sparse."""
    n = H.max_eob(tx_size)
    w, h = H.TX_W[tx_size], H.TX_H[tx_size]
    if mode == 0:
        resid = rng.integers(-(1 << bd) + 1, 1 << bd, (h, w)).astype(np.int16)
        c = H.oracle_fwd_txfm2d(resid, w, tx_size, tx_type, bd)
        q = int(rng.integers(1, 200))
        return ((c // q) * q).astype(np.int32)
    if mode == 1:
        lim = 1 << (bd + 7)
        return rng.integers(-lim, lim, n).astype(np.int32)
    c = np.zeros(n, np.int32)
    k = int(rng.integers(1, 4))
    c[rng.integers(0, n, k)] = rng.integers(-(1 << (bd + 6)), 1 << (bd + 6), k)
    return c


@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("tx_size", range(H.N_TX_SIZES))
def test_inv_txfm2d_add_all_types(tx_size, bd):
    R, O = H.ref(), H.oracle()
    w, h = H.TX_W[tx_size], H.TX_H[tx_size]
    fn = getattr(R, "av1_inv_txfm2d_add_%dx%d_c" % (w, h))
    rng = np.random.default_rng(7100 + tx_size * 3 + bd)
    stride = w + 8
    for tx_type in range(16):
        if not H.tx_valid(tx_size, tx_type):
            continue
        for it in range(6 if w * h <= 256 else 3):
            c = _coeffs(rng, tx_size, tx_type, bd, it % 3)
            pred = rng.integers(0, 1 << bd, (h, stride)).astype(np.uint16)
            if it == 1:
                pred[:] = (1 << bd) - 1
            a, b = pred.copy(), pred.copy()
            fn(ptr(c, i32p), ptr(a, u16p), stride, C.c_uint8(tx_type), bd)
            O.av1o_inv_txfm2d_add(ptr(c, i32p), ptr(b, u16p), stride, tx_size, tx_type, bd)
            assert np.array_equal(a, b), (tx_size, tx_type, bd, it)


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
def test_inverse_transform_block_and_px_dist(highbd, bd):
    """av1_inv_txfm_add_c / av1_highbd_inv_txfm_add_c on pixel buffers (lossless 4x4 included) and the
    distortion of dist_block_px_domain: aom_sse / aom_highbd_sse rounded as the variance functions do."""
    R, O = H.ref(), H.oracle()
    R.aom_sse_c.restype = C.c_int64
    R.aom_highbd_sse_c.restype = C.c_int64
    O.av1o_dist_block_px.restype = C.c_int64
    rng = np.random.default_rng(7300 + bd + highbd)
    dt = np.uint16 if highbd else np.uint8
    for tx_size in range(H.N_TX_SIZES):
        w, h = H.TX_W[tx_size], H.TX_H[tx_size]
        stride = w + 16
        for tx_type in (0, 1, 5, 9, 10, 15):
            for lossless in ((0, 1) if tx_size == 0 and tx_type == 0 else (0,)):
                if not H.tx_valid(tx_size, tx_type):
                    continue
                c = _coeffs(rng, tx_size, tx_type, bd, int(rng.integers(0, 3)))
                eob = int(rng.integers(0, 3)) if lossless else int(np.count_nonzero(c))
                if lossless:
                    c = rng.integers(-2000, 2000, 16).astype(np.int32)
                    if eob <= 1:
                        c[1:] = 0
                pred = rng.integers(0, 1 << bd, (h, stride)).astype(dt)
                src = rng.integers(0, 1 << bd, (h, stride)).astype(dt)
                a = pred.copy()
                p = TxfmParam(tx_type, tx_size, lossless, bd, highbd, 0, eob)
                if eob:  # av1_inverse_transform_block returns early on eob == 0
                    if highbd:
                        R.av1_highbd_inv_txfm_add_c(ptr(c, i32p), byteptr(a), stride, C.byref(p))
                    else:
                        R.av1_inv_txfm_add_c(ptr(c, i32p), ptr(a, u8p), stride, C.byref(p))
                b = np.zeros_like(pred)
                O.av1o_recon_block(ptr(c, i32p), pred.ctypes.data_as(C.c_void_p), stride,
                                   b.ctypes.data_as(C.c_void_p), stride, tx_size, tx_type, eob, lossless,
                                   highbd, bd)
                assert np.array_equal(a[:, :w], b[:, :w]), (tx_size, tx_type, lossless, eob)
                for vw, vh in ((w, h), (max(w - 3, 1), max(h - 2, 1))):
                    if highbd:
                        s = R.aom_highbd_sse_c(byteptr(src), stride, byteptr(a), stride, vw, vh)
                        s = (s + ((1 << (2 * (bd - 8))) >> 1)) >> (2 * (bd - 8))
                    else:
                        s = R.aom_sse_c(ptr(src, u8p), stride, ptr(a, u8p), stride, vw, vh)
                    d = O.av1o_dist_block_px(src.ctypes.data_as(C.c_void_p), stride,
                                             b.ctypes.data_as(C.c_void_p), stride, vw, vh, highbd, bd)
                    assert d == 16 * (s & 0xffffffff), (tx_size, tx_type, vw, vh)


def test_cost_coeffs_txb_estimate():
    """av1_cost_coeffs_txb_estimate (txb_rdopt.c:547) through a bare MACROBLOCK (oracle/ref_shim.c)"""
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(7500)
    for ts in range(H.N_TX_SIZES):
        n = H.max_eob(ts)
        for tt in range(16):
            if not H.tx_valid(ts, tt):
                continue
            scan, _ = H.oracle_scan(ts, tt)
            for it in range(4):
                q = np.zeros(n, np.int32)
                eob = int(rng.integers(1, n + 1)) if it else 1
                lv = rng.integers(-3, 4, eob) if it < 3 else rng.integers(-40000, 40000, eob)
                q[scan[:eob]] = lv
                if q[scan[eob - 1]] == 0:
                    q[scan[eob - 1]] = 1 + it
                a = R.refshim_cost_coeffs_txb_estimate(ptr(q, i32p), eob, ts, tt)
                b = O.av1o_cost_coeffs_txb_estimate(ptr(q, i32p), eob, ts, tt)
                assert a == b, (ts, tt, it)
