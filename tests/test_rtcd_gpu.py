"""GPU parity of the drop-in tier (include/b200av1_rtcd.h): every `_b200` flavour is
called with the reference's own argument list and compared with the oracle, replaying
the reference unit tests' inputs (test/quantize_func_test.cc, error_block_test.cc,
sad_test.cc, variance_test.cc, subtract_test.cc)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, ptr, u8p, u16p, u32p

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    import b200av1
    lib = b200av1.lib()
    for n in ("av1_block_error_b200", "av1_block_error_lp_b200", "av1_highbd_block_error_b200",
              "aom_sse_b200", "aom_highbd_sse_b200"):
        getattr(lib, n).restype = C.c_int64
    lib.b200_default_ctx()
    return lib


def byteptr(a16, byte_off=0):
    return C.cast(C.c_void_p((a16.ctypes.data + byte_off) >> 1), u8p)


@pytest.mark.parametrize("ts", range(19), ids=H.TX_NAMES)
def test_fwd_txfm2d_dropin(L, ts):
    W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
    fn = getattr(L, "av1_fwd_txfm2d_%dx%d_b200" % (W, Hh))
    rng = np.random.default_rng(ts)
    stride = 80
    for tt in range(16):
        if not H.tx_valid(ts, tt):
            continue
        for bd in (8, 12):
            hi = (1 << bd) - 1
            inp = np.zeros((Hh, stride), np.int16)
            inp[:, :W] = rng.integers(-hi, hi + 1, (Hh, W))
            out = np.zeros(W * Hh, np.int32)
            fn(ptr(inp, i16p), ptr(out, i32p), stride, C.c_uint8(tt), bd)
            assert np.array_equal(out[:n], H.oracle_fwd_txfm2d(inp, stride, ts, tt, bd)), (tt, bd)


def test_lowbd_fwd_txfm_dropin(L):
    class TxfmParam(C.Structure):
        _fields_ = [("tx_type", C.c_uint8), ("tx_size", C.c_uint8), ("lossless", C.c_int),
                    ("bd", C.c_int), ("is_hbd", C.c_int), ("tx_set_type", C.c_uint8), ("eob", C.c_int)]
    rng = np.random.default_rng(2)
    for ts in (0, 2, 7, 12):
        W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
        inp = rng.integers(-255, 256, (Hh, W)).astype(np.int16)
        out = np.zeros(W * Hh, np.int32)
        p = TxfmParam(0, ts, 0, 8, 0, 0, 0)
        L.av1_lowbd_fwd_txfm_b200(ptr(inp, i16p), ptr(out, i32p), W, C.byref(p))
        assert np.array_equal(out[:n], H.oracle_fwd_txfm2d(inp, W, ts, 0, 8))


def _fill(rng, n, bd, mode, highbd):
    c = np.zeros(n, np.int32)
    lim = (1 << (7 + bd)) if highbd else 32767
    if mode == "neg":
        c[:] = -8191
    elif mode == "dc":
        c[0] = rng.integers(-lim, lim)
    elif mode == "half":
        c[:] = 16
    elif mode == "sparse":
        k = rng.integers(1, n + 1)
        c[:k] = rng.integers(-lim, lim, k)
    elif mode == "rand":
        c[:] = rng.integers(-lim, lim, n)
    elif mode == "near":  # around the dead zone: exercises the adaptive trailing cut / lone +-1
        c[:] = rng.integers(-3, 4, n)
        c[rng.integers(0, n)] = rng.integers(-200, 200)
    return c


LS = {k: [t for t in range(19) if H.log_scale(t) == k] for k in (0, 1, 2)}
QCASES = [("av1_quantize_fp_b200", "av1o_quantize_fp", "fp", 0, 0),
          ("av1_quantize_fp_32x32_b200", "av1o_quantize_fp", "fp", 1, 0),
          ("av1_quantize_fp_64x64_b200", "av1o_quantize_fp", "fp", 2, 0),
          ("av1_highbd_quantize_fp_b200", "av1o_highbd_quantize_fp", "hfp", None, 1),
          ("aom_quantize_b_b200", "av1o_quantize_b", "b", 0, 0),
          ("aom_quantize_b_32x32_b200", "av1o_quantize_b", "b", 1, 0),
          ("aom_quantize_b_64x64_b200", "av1o_quantize_b", "b", 2, 0),
          ("aom_highbd_quantize_b_b200", "av1o_highbd_quantize_b", "b", 0, 1),
          ("aom_highbd_quantize_b_32x32_b200", "av1o_highbd_quantize_b", "b", 1, 1),
          ("aom_highbd_quantize_b_64x64_b200", "av1o_highbd_quantize_b", "b", 2, 1),
          ("aom_quantize_b_adaptive_b200", "av1o_quantize_b_adaptive", "ba", 0, 0),
          ("aom_quantize_b_32x32_adaptive_b200", "av1o_quantize_b_adaptive", "ba", 1, 0),
          ("aom_quantize_b_64x64_adaptive_b200", "av1o_quantize_b_adaptive", "ba", 2, 0),
          ("aom_highbd_quantize_b_adaptive_b200", "av1o_highbd_quantize_b_adaptive", "ba", 0, 1),
          ("aom_highbd_quantize_b_32x32_adaptive_b200", "av1o_highbd_quantize_b_adaptive", "ba", 1, 1),
          ("aom_highbd_quantize_b_64x64_adaptive_b200", "av1o_highbd_quantize_b_adaptive", "ba", 2, 1)]


@pytest.mark.parametrize("case", QCASES, ids=[c[0] for c in QCASES])
def test_quantizer_dropins(L, case):
    pname, oname, kind, ls_fixed, highbd = case
    O = H.oracle()
    pfn, ofn = getattr(L, pname), getattr(O, oname)
    rng = np.random.default_rng(abs(hash(pname)) % 997)
    sizes = list(range(0, 19, 2)) if ls_fixed is None else LS[ls_fixed]
    for ts in sizes[:6]:
        n = H.max_eob(ts)
        ls = H.log_scale(ts)
        for bd in ((10, 12) if highbd else (8,)):
            for qi, mode in [(0, "rand"), (255, "rand"), (100, "neg"), (60, "dc"), (25, "half"),
                             (140, "sparse"), (200, "near"), (33, "near"), (90, "zero")]:
                row = H.oracle_qrow(bd, qi)
                scan, iscan = H.oracle_scan(ts, int(rng.integers(0, 16)))
                coeff = _fill(rng, n, bd, mode, highbd)
                res = []
                for which in (0, 1):
                    qc = np.full(n, 77, np.int32)
                    dq = np.full(n, 77, np.int32)
                    eob = C.c_uint16(999)
                    fp = kind in ("fp", "hfp")
                    args = [ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round_fp if fp else row.round,
                            row.quant_fp if fp else row.quant, row.quant_shift, ptr(qc, i32p),
                            ptr(dq, i32p), row.dequant, C.byref(eob), ptr(scan, i16p), ptr(iscan, i16p)]
                    if which == 0:
                        if kind == "hfp":
                            args.append(ls)
                        pfn(*args)
                    else:
                        if kind in ("fp", "hfp", "b"):
                            args += [None, None]
                        args.append(ls)
                        ofn(*args)
                    res.append((qc, dq, eob.value))
                tag = (pname, H.TX_NAMES[ts], bd, qi, mode)
                assert res[0][2] == res[1][2], tag
                assert np.array_equal(res[0][0], res[1][0]), tag
                assert np.array_equal(res[0][1], res[1][1]), tag


def test_quantize_b_helper_qmatrix_and_lp(L):
    O = H.oracle()
    rng = np.random.default_rng(5)
    for highbd, pname, oname in ((0, "aom_quantize_b_helper_b200", "av1o_quantize_b"),
                                 (1, "aom_highbd_quantize_b_helper_b200", "av1o_highbd_quantize_b")):
        for ts in (0, 2, 3, 4):
            n, ls = H.max_eob(ts), H.log_scale(ts)
            for it in range(4):
                bd = 10 if highbd else 8
                row = H.oracle_qrow(bd, int(rng.integers(0, 256)))
                scan, iscan = H.oracle_scan(ts, 0)
                coeff = _fill(rng, n, bd, "rand" if it % 2 else "sparse", highbd)
                qm = rng.integers(20, 120, n).astype(np.uint8)
                iqm = rng.integers(20, 120, n).astype(np.uint8)
                res = []
                for fn in (getattr(L, pname), getattr(O, oname)):
                    qc, dq, eob = np.zeros(n, np.int32), np.zeros(n, np.int32), C.c_uint16(0)
                    fn(ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round, row.quant,
                       row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob),
                       ptr(scan, i16p), ptr(iscan, i16p), ptr(qm, u8p), ptr(iqm, u8p), ls)
                    res.append((qc, dq, eob.value))
                assert res[0][2] == res[1][2]
                assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
    for ts in (0, 1, 2, 7):
        n = H.max_eob(ts)
        row = H.oracle_qrow(8, int(rng.integers(0, 256)))
        scan, iscan = H.oracle_scan(ts, 0)
        coeff = rng.integers(-4000, 4000, n).astype(np.int16)
        res = []
        for fn in (L.av1_quantize_lp_b200, O.av1o_quantize_lp):
            qc, dq, eob = np.zeros(n, np.int16), np.zeros(n, np.int16), C.c_uint16(0)
            fn(ptr(coeff, i16p), C.c_ssize_t(n), row.round_fp, row.quant_fp, ptr(qc, i16p),
               ptr(dq, i16p), row.dequant, C.byref(eob), ptr(scan, i16p), ptr(iscan, i16p))
            res.append((qc, dq, eob.value))
        assert res[0][2] == res[1][2]
        assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


def test_block_error_dropins(L):
    O = H.oracle()
    rng = np.random.default_rng(13)
    for it in range(27):
        n = 16 << (it % 9) if it % 9 < 7 else 1024
        for bd in (8, 10, 12):
            msb = bd + 8 - 1
            coeff = rng.integers(0, 1 << msb, n).astype(np.int32)
            dq = rng.integers(0, 1 << msb, n).astype(np.int32)
            if it % 3 == 1:
                coeff, dq = -coeff, -dq
            s0, s1 = C.c_int64(0), C.c_int64(0)
            if bd == 8:
                a = L.av1_block_error_b200(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s0))
                b = O.av1o_block_error(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s1))
                assert (a, s0.value) == (b, s1.value)
                c16, d16 = (coeff >> 2).astype(np.int16), (dq >> 2).astype(np.int16)
                assert L.av1_block_error_lp_b200(ptr(c16, i16p), ptr(d16, i16p), C.c_ssize_t(n)) == \
                    O.av1o_block_error_lp(ptr(c16, i16p), ptr(d16, i16p), C.c_ssize_t(n))
            a = L.av1_highbd_block_error_b200(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s0), bd)
            b = O.av1o_highbd_block_error(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s1), bd)
            assert (a, s0.value) == (b, s1.value)


def test_subtract_and_sse_dropins(L):
    O = H.oracle()
    rng = np.random.default_rng(14)
    for (w, h) in H.BLOCK_SIZES[::3]:
        src = rng.integers(0, 256, (h, w + 16)).astype(np.uint8)
        pred = rng.integers(0, 256, (h, w + 8)).astype(np.uint8)
        a, b = np.zeros((h, w + 4), np.int16), np.zeros((h, w + 4), np.int16)
        L.aom_subtract_block_b200(h, w, ptr(a, i16p), C.c_ssize_t(w + 4), ptr(src, u8p),
                                  C.c_ssize_t(w + 16), ptr(pred, u8p), C.c_ssize_t(w + 8))
        O.av1o_subtract_block(h, w, ptr(b, i16p), C.c_ssize_t(w + 4), ptr(src, u8p),
                              C.c_ssize_t(w + 16), ptr(pred, u8p), C.c_ssize_t(w + 8))
        assert np.array_equal(a, b)
        assert L.aom_sse_b200(ptr(src, u8p), w + 16, ptr(pred, u8p), w + 8, w, h) == \
            O.av1o_sse(ptr(src, u8p), w + 16, ptr(pred, u8p), w + 8, w, h)
        s16 = rng.integers(0, 4096, (h, w + 16)).astype(np.uint16)
        p16 = rng.integers(0, 4096, (h, w + 8)).astype(np.uint16)
        L.aom_highbd_subtract_block_b200(h, w, ptr(a, i16p), C.c_ssize_t(w + 4), byteptr(s16),
                                         C.c_ssize_t(w + 16), byteptr(p16), C.c_ssize_t(w + 8))
        O.av1o_highbd_subtract_block(h, w, ptr(b, i16p), C.c_ssize_t(w + 4), ptr(s16, u16p),
                                     C.c_ssize_t(w + 16), ptr(p16, u16p), C.c_ssize_t(w + 8))
        assert np.array_equal(a, b)
        assert L.aom_highbd_sse_b200(byteptr(s16), w + 16, byteptr(p16), w + 8, w, h) == \
            O.av1o_highbd_sse(ptr(s16, u16p), w + 16, ptr(p16, u16p), w + 8, w, h)


@pytest.mark.parametrize("wh", H.BLOCK_SIZES, ids=["%dx%d" % s for s in H.BLOCK_SIZES])
def test_sad_and_variance_dropins(L, wh):
    w, h = wh
    O = H.oracle()
    rng = np.random.default_rng(w * 131 + h)
    nm = "%dx%d" % (w, h)
    sstride, rstride = (w + 31) & ~31, 2 * w + 3
    for it in range(3):
        src = rng.integers(0, 256, (h, sstride)).astype(np.uint8)
        refs = rng.integers(0, 256, (4, h + 2, rstride)).astype(np.uint8)
        if it == 0:
            src[:] = 0
            refs[:] = 255
        rp = [C.cast(C.c_void_p(refs[i].ctypes.data + 1 + it), u8p) for i in range(4)]
        s = ptr(src, u8p)
        assert getattr(L, "aom_sad%s_b200" % nm)(s, sstride, rp[0], rstride) == \
            O.av1o_sad(s, sstride, rp[0], rstride, w, h)
        assert getattr(L, "aom_sad_skip_%s_b200" % nm)(s, sstride, rp[1], rstride) == \
            O.av1o_sad_skip(s, sstride, rp[1], rstride, w, h)
        second = rng.integers(0, 256, w * h).astype(np.uint8)
        assert getattr(L, "aom_sad%s_avg_b200" % nm)(s, sstride, rp[2], rstride, ptr(second, u8p)) == \
            O.av1o_sad_avg(s, sstride, rp[2], rstride, ptr(second, u8p), w, h)
        arr = (u8p * 4)(*rp)
        for sym, skip in (("aom_sad%sx4d_b200" % nm, 0), ("aom_sad_skip_%sx4d_b200" % nm, 1),
                          ("aom_sad%sx3d_b200" % nm, 0)):
            a, b = np.zeros(4, np.uint32), np.zeros(4, np.uint32)
            getattr(L, sym)(s, sstride, arr, rstride, ptr(a, u32p))
            O.av1o_sad_x4d(s, sstride, arr, rstride, ptr(b, u32p), w, h, 4, skip)
            assert np.array_equal(a, b), sym
        # variance family: a = first pointer (interpolated operand for sub-pel)
        a8 = rng.integers(0, 256, (h + 1, w + 9)).astype(np.uint8)
        b8 = rng.integers(0, 256, (h, w + 3)).astype(np.uint8)
        if it == 0:  # OneQuarter KAT (test/variance_test.cc:821): var = N*255^2/4
            a8[:] = 0
            b8[:] = 0
            b8[: h // 2, :] = 255
        s0, s1 = C.c_uint(0), C.c_uint(0)
        v0 = getattr(L, "aom_variance%s_b200" % nm)(ptr(a8, u8p), w + 9, ptr(b8, u8p), w + 3, C.byref(s0))
        v1 = O.av1o_variance(ptr(a8, u8p), w + 9, ptr(b8, u8p), w + 3, w, h, C.byref(s1))
        assert (C.c_uint(v0).value, s0.value) == (v1, s1.value)
        if it == 0:
            assert v1 == w * h * 255 * 255 // 4
        for xo, yo in ((0, 0), (4, 0), (0, 3), (7, 7), (int(rng.integers(0, 8)), int(rng.integers(0, 8)))):
            v0 = getattr(L, "aom_sub_pixel_variance%s_b200" % nm)(
                ptr(a8, u8p), w + 9, xo, yo, ptr(b8, u8p), w + 3, C.byref(s0))
            v1 = O.av1o_sub_pixel_variance(ptr(a8, u8p), w + 9, xo, yo, ptr(b8, u8p), w + 3, w, h, C.byref(s1))
            assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), (xo, yo)
        for bd in (8, 10, 12):
            a16 = rng.integers(0, 1 << bd, (h + 1, w + 9)).astype(np.uint16)
            b16 = rng.integers(0, 1 << bd, (h, w + 3)).astype(np.uint16)
            if it == 0:
                a16[:] = (1 << bd) - 1
                b16[:] = 0
            v0 = getattr(L, "aom_highbd_%d_variance%s_b200" % (bd, nm))(
                byteptr(a16), w + 9, byteptr(b16), w + 3, C.byref(s0))
            v1 = O.av1o_highbd_variance(ptr(a16, u16p), w + 9, ptr(b16, u16p), w + 3, w, h, bd, C.byref(s1))
            assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), bd
            xo, yo = int(rng.integers(0, 8)), int(rng.integers(0, 8))
            v0 = getattr(L, "aom_highbd_%d_sub_pixel_variance%s_b200" % (bd, nm))(
                byteptr(a16), w + 9, xo, yo, byteptr(b16), w + 3, C.byref(s0))
            v1 = O.av1o_highbd_sub_pixel_variance(ptr(a16, u16p), w + 9, xo, yo, ptr(b16, u16p), w + 3,
                                                  w, h, bd, C.byref(s1))
            assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), (bd, xo, yo)
            if bd > 8:
                assert getattr(L, "aom_highbd_sad%s_b200" % nm)(byteptr(a16), w + 9, byteptr(b16), w + 3) == \
                    O.av1o_highbd_sad(ptr(a16, u16p), w + 9, ptr(b16, u16p), w + 3, w, h)
                assert getattr(L, "aom_highbd_sad_skip_%s_b200" % nm)(byteptr(a16), w + 9, byteptr(b16), w + 3) == \
                    O.av1o_highbd_sad_skip(ptr(a16, u16p), w + 9, ptr(b16, u16p), w + 3, w, h)


def test_mse_dropins(L):
    O = H.oracle()
    rng = np.random.default_rng(3)
    for (w, h) in ((16, 16), (16, 8), (8, 16), (8, 8)):
        a = rng.integers(0, 256, (h, w + 5)).astype(np.uint8)
        b = rng.integers(0, 256, (h, w + 1)).astype(np.uint8)
        s0, s1 = C.c_uint(0), C.c_uint(0)
        v0 = getattr(L, "aom_mse%dx%d_b200" % (w, h))(ptr(a, u8p), w + 5, ptr(b, u8p), w + 1, C.byref(s0))
        v1 = O.av1o_mse(ptr(a, u8p), w + 5, ptr(b, u8p), w + 1, w, h, C.byref(s1))
        assert (C.c_uint(v0).value, s0.value) == (v1, s1.value)


def test_upsampled_pred_dropins(L):
    O = H.oracle()
    rng = np.random.default_rng(15)
    for (w, h) in [(4, 4), (8, 8), (16, 16), (64, 64), (16, 8), (128, 128)]:
        stride = w + 24
        refimg = rng.integers(0, 256, (h + 16, stride)).astype(np.uint8)
        base = refimg.ctypes.data + 8 * stride + 8
        for sx, sy in ((0, 0), (3, 0), (0, 5), (7, 7), (4, 2), (1, 6)):
            for search, taps in ((1, 2), (2, 4), (3, 8)):
                a, b = np.zeros(w * h, np.uint8), np.zeros(w * h, np.uint8)
                L.aom_upsampled_pred_b200(None, None, 0, 0, None, ptr(a, u8p), w, h, sx, sy,
                                          C.cast(C.c_void_p(base), u8p), stride, search)
                O.av1o_upsampled_pred(ptr(b, u8p), w, h, sx, sy, C.cast(C.c_void_p(base), u8p), stride, taps)
                assert np.array_equal(a, b), (w, h, sx, sy, taps)
        for bd in (10, 12):
            ref16 = rng.integers(0, 1 << bd, (h + 16, stride)).astype(np.uint16)
            base16 = ref16.ctypes.data + 2 * (8 * stride + 8)
            for sx, sy in ((0, 0), (3, 0), (0, 5), (7, 7)):
                for search, taps in ((1, 2), (3, 8)):
                    a, b = np.zeros(w * h, np.uint16), np.zeros(w * h, np.uint16)
                    L.aom_highbd_upsampled_pred_b200(None, None, 0, 0, None, byteptr(a), w, h, sx, sy,
                                                     C.cast(C.c_void_p(base16 >> 1), u8p), stride, bd, search)
                    O.av1o_highbd_upsampled_pred(ptr(b, u16p), w, h, sx, sy,
                                                 C.cast(C.c_void_p(base16), u16p), stride, bd, taps)
                    assert np.array_equal(a, b), (w, h, sx, sy, taps, bd)


def test_pix_batch_many_blocks(L):
    """batched tier: thousands of SAD / variance blocks of mixed sizes in ONE launch"""
    import b200av1
    O = H.oracle()
    ctx = b200av1.Context(0)
    rng = np.random.default_rng(21)
    W, Hh = 640, 360
    cur = rng.integers(0, 256, (Hh, W)).astype(np.uint8)
    ref = rng.integers(0, 256, (Hh + 8, W + 8)).astype(np.uint8)
    n = 3000
    blocks = np.zeros(n, b200av1.PIXBLOCK_DTYPE)   # a = cur, b = ref
    blocks_sp = np.zeros(n, b200av1.PIXBLOCK_DTYPE)  # sub-pel: a = ref (interpolated), b = cur
    for i in range(n):
        w, h = H.BLOCK_SIZES[int(rng.integers(0, len(H.BLOCK_SIZES)))]
        x, y = int(rng.integers(0, W - w)), int(rng.integers(0, Hh - h))
        rx, ry = int(rng.integers(0, W + 8 - w - 1)), int(rng.integers(0, Hh + 8 - h - 1))
        blocks[i] = (y * W + x, ry * (W + 8) + rx, 0, W, W + 8, w, h, 0, 0)
        blocks_sp[i] = (ry * (W + 8) + rx, y * W + x, 0, W + 8, W, w, h, rng.integers(0, 8),
                        rng.integers(0, 8))
    d_cur, d_ref = b200av1.to_device(ctx, cur), b200av1.to_device(ctx, ref)
    d_blk, d_blk_sp = b200av1.to_device(ctx, blocks), b200av1.to_device(ctx, blocks_sp)
    d_out = b200av1.DevBuf(ctx, n * 8)
    for op in (b200av1.PIX_SAD, b200av1.PIX_SAD_SKIP, b200av1.PIX_VAR, b200av1.PIX_SUBPEL_VAR):
        sp = op == b200av1.PIX_SUBPEL_VAR
        l = b200av1.PixLaunch(op, 0, 8, n, (d_blk_sp if sp else d_blk).ptr,
                              (d_ref if sp else d_cur).ptr, (d_cur if sp else d_ref).ptr, None,
                              d_out.ptr)
        b200av1.check(L.b200_pix_launch(ctx.h, C.byref(l), None))
        out = b200av1.to_host(ctx, d_out, np.uint32, 2 * n).reshape(n, 2)
        for i in range(0, n, 7):
            b = blocks_sp[i] if sp else blocks[i]
            w, h = int(b["w"]), int(b["h"])
            pa = C.cast(C.c_void_p((ref if sp else cur).ctypes.data + int(b["a_offset"])), u8p)
            pb = C.cast(C.c_void_p((cur if sp else ref).ctypes.data + int(b["b_offset"])), u8p)
            sa, sb = int(b["a_stride"]), int(b["b_stride"])
            sse = C.c_uint(0)
            if op == b200av1.PIX_SAD:
                exp = (O.av1o_sad(pa, sa, pb, sb, w, h), 0)
            elif op == b200av1.PIX_SAD_SKIP:
                exp = (O.av1o_sad_skip(pa, sa, pb, sb, w, h), 0)
            elif op == b200av1.PIX_VAR:
                exp = (O.av1o_variance(pa, sa, pb, sb, w, h, C.byref(sse)), sse.value)
            else:
                exp = (O.av1o_sub_pixel_variance(pa, sa, int(b["xoff"]), int(b["yoff"]), pb, sb, w, h,
                                                 C.byref(sse)), sse.value)
            assert tuple(int(v) for v in out[i]) == tuple(int(v) for v in exp), (op, i)
    ctx.close()
