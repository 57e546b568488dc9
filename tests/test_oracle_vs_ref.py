"""Pin the oracle (oracle/*.c) against the reference's own C functions compiled from
/root/reference (oracle/_ref/libaomref.so), replaying the input distributions of the
reference's unit tests.  Skipped where the reference build is
absent; tests/test_oracle_golden.py covers those machines with committed vectors."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, i64p, ptr, u8p, u16p, u32p

pytestmark = [pytest.mark.ref,
              pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]


def byteptr(a16):
    """CONVERT_TO_BYTEPTR (aom_ports/mem.h:80): uint16 address >> 1, typed uint8*."""
    return C.cast(C.c_void_p(a16.ctypes.data >> 1), u8p)


def test_cospi_sinpi_tables():
    R, O = H.ref(), H.oracle()
    cos = (C.c_int32 * 64 * 7).in_dll(R, "av1_cospi_arr_data")
    sin = (C.c_int32 * 5 * 7).in_dll(R, "av1_sinpi_arr_data")
    for b in range(7):
        oc = np.ctypeslib.as_array(O.av1o_cospi(10 + b), shape=(64,))
        osn = np.ctypeslib.as_array(O.av1o_sinpi(10 + b), shape=(5,))
        assert list(cos[b]) == list(oc)
        assert list(sin[b]) == list(osn)


@pytest.mark.parametrize("name,kind,n", [("av1_fdct4", 0, 4), ("av1_fdct8", 0, 8),
                                         ("av1_fdct16", 0, 16), ("av1_fdct32", 0, 32),
                                         ("av1_fdct64", 0, 64), ("av1_fadst4", 1, 4),
                                         ("av1_fadst8", 1, 8), ("av1_fadst16", 1, 16),
                                         ("av1_fidentity4_c", 2, 4), ("av1_fidentity8_c", 2, 8),
                                         ("av1_fidentity16_c", 2, 16),
                                         ("av1_fidentity32_c", 2, 32)])
def test_txfm1d(name, kind, n):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(100 + n + kind)
    stage_range = (C.c_int8 * 12)(*([32] * 12))
    for it in range(300):
        bits = [8, 12, 16, 20][it % 4]
        x = rng.integers(-(1 << bits), 1 << bits, n).astype(np.int32)
        if it == 0:
            x[:] = 0
        if it == 1:
            x[:] = (1 << 18) - 1
        for cos_bit in (10, 11, 12, 13):
            a = np.zeros(n, np.int32)
            b = np.zeros(n, np.int32)
            getattr(R, name)(ptr(x, i32p), ptr(a, i32p), C.c_int8(cos_bit), stage_range)
            O.av1o_fwd_txfm1d(kind, n, ptr(x, i32p), ptr(b, i32p), cos_bit)
            assert np.array_equal(a, b), (name, cos_bit, it)


@pytest.mark.parametrize("tx_size", range(H.N_TX_SIZES))
def test_fwd_txfm2d_all_types(tx_size):
    """test/av1_fwd_txfm2d_test.cc:243-297,518-576: all-max block, then random."""
    R, O = H.ref(), H.oracle()
    W, Hh = H.TX_W[tx_size], H.TX_H[tx_size]
    fn = getattr(R, "av1_fwd_txfm2d_%dx%d_c" % (W, Hh))
    rng = np.random.default_rng(7 + tx_size)
    stride = 64 if max(W, Hh) <= 64 else 128
    n = H.max_eob(tx_size)
    iters = 40 if W * Hh <= 256 else 12
    for tx_type in range(16):
        if not H.tx_valid(tx_size, tx_type):
            continue
        for bd in (8, 10, 12):
            for it in range(iters):
                inp = np.zeros((Hh, stride), np.int16)
                hi = (1 << bd) - 1
                if it == 0:
                    inp[:, :W] = hi
                elif it == 1:
                    inp[:, :W] = -hi
                else:
                    inp[:, :W] = rng.integers(-hi, hi + 1, (Hh, W))
                a = np.zeros(W * Hh + 64, np.int32)
                fn(ptr(inp, i16p), ptr(a, i32p), stride, C.c_uint8(tx_type), bd)
                b = H.oracle_fwd_txfm2d(inp, stride, tx_size, tx_type, bd)
                assert np.array_equal(a[:n], b), (H.TX_NAMES[tx_size], tx_type, bd, it)


def test_fwht4x4():
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(3)
    for it in range(200):
        inp = rng.integers(-255, 256, (4, 8)).astype(np.int16)
        a = np.zeros(16, np.int32)
        b = np.zeros(16, np.int32)
        R.av1_fwht4x4_c(ptr(inp, i16p), ptr(a, i32p), 8)
        O.av1o_fwht4x4(ptr(inp, i16p), ptr(b, i32p), 8)
        assert np.array_equal(a, b)


def test_scan_tables():
    """all 19 x 16 entries of av1_scan_orders (scan.c:1666)"""
    R = H.ref()

    class ScanOrder(C.Structure):
        _fields_ = [("scan", i16p), ("iscan", i16p)]

    tab = (ScanOrder * 16 * 19).in_dll(R, "av1_scan_orders")
    for ts in range(19):
        n = H.max_eob(ts)
        for tt in range(16):
            s, i = H.oracle_scan(ts, tt)
            rs = np.ctypeslib.as_array(tab[ts][tt].scan, shape=(n,))
            ri = np.ctypeslib.as_array(tab[ts][tt].iscan, shape=(n,))
            assert np.array_equal(rs, s), (ts, tt)
            assert np.array_equal(ri, i), (ts, tt)


def test_qlookup_and_build_quantizer():
    R, O = H.ref(), H.oracle()
    R.av1_dc_quant_QTX.restype = C.c_int16
    R.av1_ac_quant_QTX.restype = C.c_int16

    class Quants(C.Structure):
        _fields_ = [(n, C.c_int16 * 8 * 256) for n in
                    ("y_quant", "y_quant_shift", "y_zbin", "y_round", "y_quant_fp",
                     "u_quant_fp", "v_quant_fp", "y_round_fp", "u_round_fp", "v_round_fp",
                     "u_quant", "v_quant", "u_quant_shift", "v_quant_shift", "u_zbin",
                     "v_zbin", "u_round", "v_round")]

    class Dequants(C.Structure):
        _fields_ = [(n, C.c_int16 * 8 * 256) for n in ("y", "u", "v")]

    for bd in (8, 10, 12):
        for q in range(256):
            assert R.av1_dc_quant_QTX(q, 0, bd) == O.av1o_dc_q(q, 0, bd)
            assert R.av1_ac_quant_QTX(q, 0, bd) == O.av1o_ac_q(q, 0, bd)
        for sharp in (0, 3, -2):
            qs, dq = Quants(), Dequants()
            R.av1_build_quantizer(bd, 0, 0, 0, 0, 0, C.byref(qs), C.byref(dq), sharp)
            for q in range(256):
                row = H.oracle_qrow(bd, q, 0, sharp)
                for mine, theirs in ((row.quant, qs.y_quant), (row.quant_shift, qs.y_quant_shift),
                                     (row.zbin, qs.y_zbin), (row.round, qs.y_round),
                                     (row.quant_fp, qs.y_quant_fp), (row.round_fp, qs.y_round_fp),
                                     (row.dequant, dq.y), (row.quant, qs.v_quant),
                                     (row.round_fp, qs.u_round_fp)):
                    assert list(mine) == list(theirs[q]), (bd, q, sharp)


def _fill_coeff(rng, n, bd, mode, highbd):
    """test/quantize_func_test.cc:209-259 input generators"""
    c = np.zeros(n, np.int32)
    lim = (1 << (7 + bd)) if highbd else 32767
    if mode == "zero":
        pass
    elif mode == "neg":
        c[:] = -8191
    elif mode == "dc":
        c[0] = rng.integers(-lim, lim)
    elif mode == "half":
        c[:] = 16
    elif mode == "sparse":
        k = rng.integers(1, n + 1)
        c[:k] = rng.integers(-lim, lim, k)
    else:
        c[:] = rng.integers(-lim, lim, n)
    return c


LS0 = [t for t in range(19) if H.log_scale(t) == 0]
LS1 = [t for t in range(19) if H.log_scale(t) == 1]
LS2 = [t for t in range(19) if H.log_scale(t) == 2]
QUANT_CASES = [
    # (ref symbol, oracle symbol, kind, tx sizes with the matching log_scale, highbd)
    ("av1_quantize_fp_c", "av1o_quantize_fp", "fp", LS0, 0),
    ("av1_quantize_fp_32x32_c", "av1o_quantize_fp", "fp", LS1, 0),
    ("av1_quantize_fp_64x64_c", "av1o_quantize_fp", "fp", LS2, 0),
    ("av1_highbd_quantize_fp_c", "av1o_highbd_quantize_fp", "hfp", list(range(19)), 1),
    ("aom_quantize_b_c", "av1o_quantize_b", "b", LS0, 0),
    ("aom_quantize_b_32x32_c", "av1o_quantize_b", "b", LS1, 0),
    ("aom_quantize_b_64x64_c", "av1o_quantize_b", "b", LS2, 0),
    ("aom_highbd_quantize_b_c", "av1o_highbd_quantize_b", "b", LS0[:5], 1),
    ("aom_highbd_quantize_b_32x32_c", "av1o_highbd_quantize_b", "b", LS1[:3], 1),
    ("aom_highbd_quantize_b_64x64_c", "av1o_highbd_quantize_b", "b", LS2, 1),
    ("aom_quantize_b_adaptive_c", "av1o_quantize_b_adaptive", "ba", LS0, 0),
    ("aom_quantize_b_32x32_adaptive_c", "av1o_quantize_b_adaptive", "ba", LS1, 0),
    ("aom_quantize_b_64x64_adaptive_c", "av1o_quantize_b_adaptive", "ba", LS2, 0),
    ("aom_highbd_quantize_b_adaptive_c", "av1o_highbd_quantize_b_adaptive", "ba", LS0[:4], 1),
    ("aom_highbd_quantize_b_32x32_adaptive_c", "av1o_highbd_quantize_b_adaptive", "ba", LS1[:2], 1),
    ("aom_highbd_quantize_b_64x64_adaptive_c", "av1o_highbd_quantize_b_adaptive", "ba", LS2[:2], 1),
]


@pytest.mark.parametrize("case", QUANT_CASES, ids=[c[0] for c in QUANT_CASES])
def test_quantizers(case):
    rname, oname, kind, sizes, highbd = case
    R, O = H.ref(), H.oracle()
    rfn, ofn = getattr(R, rname), getattr(O, oname)
    rng = np.random.default_rng(abs(hash(rname)) % 1000)
    for ts in sizes:
        n = H.max_eob(ts)
        ls = H.log_scale(ts)
        for bd in ((8, 10, 12) if highbd else (8,)):
            for qi, mode in [(0, "rand"), (255, "rand"), (32, "zero"), (100, "neg"),
                             (60, "dc"), (25, "half"), (140, "sparse"), (200, "rand"),
                             (1, "sparse"), (80, "rand"), (180, "sparse"), (230, "dc")]:
                row = H.oracle_qrow(bd, qi)
                tt = int(rng.integers(0, 16))
                scan, iscan = H.oracle_scan(ts, tt)
                coeff = _fill_coeff(rng, n, bd, mode, highbd)
                outs = []
                for which in (0, 1):
                    qc = np.full(n, 77, np.int32)
                    dq = np.full(n, 77, np.int32)
                    eob = C.c_uint16(999)
                    fp = kind in ("fp", "hfp")
                    rnd = row.round_fp if fp else row.round
                    qnt = row.quant_fp if fp else row.quant
                    args = [ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, rnd, qnt,
                            row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant,
                            C.byref(eob), ptr(scan, i16p), ptr(iscan, i16p)]
                    if which == 0:
                        if kind == "hfp":
                            args.append(ls)
                        rfn(*args)
                    else:
                        if kind in ("fp", "hfp", "b"):
                            args += [None, None]
                        args.append(ls)
                        ofn(*args)
                    outs.append((qc, dq, eob.value))
                assert outs[0][2] == outs[1][2], (rname, ts, bd, qi, mode)
                assert np.array_equal(outs[0][0], outs[1][0]), (rname, ts, bd, qi, mode)
                assert np.array_equal(outs[0][1], outs[1][1]), (rname, ts, bd, qi, mode)


def test_quantize_b_helper_qmatrix():
    """aom_quantize_b_helper_c / highbd with synthetic qmatrices (quantize.c:108,261)"""
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(11)
    for highbd, rname, oname in ((0, "aom_quantize_b_helper_c", "av1o_quantize_b"),
                                 (1, "aom_highbd_quantize_b_helper_c", "av1o_highbd_quantize_b")):
        for ts in (0, 1, 2, 3, 4, 9):
            n, ls = H.max_eob(ts), H.log_scale(ts)
            for it in range(12):
                bd = 10 if highbd else 8
                row = H.oracle_qrow(bd, int(rng.integers(0, 256)))
                scan, iscan = H.oracle_scan(ts, 0)
                coeff = _fill_coeff(rng, n, bd, "rand" if it % 2 else "sparse", highbd)
                qm = rng.integers(20, 120, n).astype(np.uint8)
                iqm = rng.integers(20, 120, n).astype(np.uint8)
                res = []
                for fn in (getattr(R, rname), getattr(O, oname)):
                    qc = np.zeros(n, np.int32)
                    dq = np.zeros(n, np.int32)
                    eob = C.c_uint16(0)
                    fn(ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round, row.quant,
                       row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob),
                       ptr(scan, i16p), ptr(iscan, i16p), ptr(qm, u8p), ptr(iqm, u8p), ls)
                    res.append((qc, dq, eob.value))
                assert res[0][2] == res[1][2]
                assert np.array_equal(res[0][0], res[1][0])
                assert np.array_equal(res[0][1], res[1][1])


def test_quantize_fp_qmatrix_via_facade():
    """quantize_fp_helper_c / highbd_quantize_fp_helper_c with a qmatrix (av1_quantize.c:96-120,
    :150-169) are static: reached through av1_quantize_fp_facade / av1_highbd_quantize_fp_facade
    with a hand-filled MACROBLOCK_PLANE + QUANT_PARAM (oracle/ref_shim.c)."""
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(12)
    for highbd, oname in ((0, "av1o_quantize_fp"), (1, "av1o_highbd_quantize_fp")):
        for ts in (0, 1, 2, 3, 4, 9, 13, 18):
            n, ls = H.max_eob(ts), H.log_scale(ts)
            for it in range(12):
                bd = (10, 12)[it % 2] if highbd else 8
                row = H.oracle_qrow(bd, int(rng.integers(0, 256)))
                scan, iscan = H.oracle_scan(ts, 0)
                coeff = _fill_coeff(rng, n, bd, ("rand", "sparse", "neg")[it % 3], highbd)
                qm = rng.integers(16, 200, n).astype(np.uint8)
                iqm = rng.integers(16, 200, n).astype(np.uint8)
                res = []
                for which in (0, 1):
                    qc = np.zeros(n, np.int32)
                    dq = np.zeros(n, np.int32)
                    eob = C.c_uint16(0)
                    args = [ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round_fp, row.quant_fp,
                            row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob),
                            ptr(scan, i16p), ptr(iscan, i16p), ptr(qm, u8p), ptr(iqm, u8p), ls]
                    if which == 0:
                        R.refshim_quantize_fp_qm(*args, highbd)
                    else:
                        getattr(O, oname)(*args)
                    res.append((qc, dq, eob.value))
                assert res[0][2] == res[1][2], (oname, ts, bd, it)
                assert np.array_equal(res[0][0], res[1][0]), (oname, ts, bd, it)
                assert np.array_equal(res[0][1], res[1][1]), (oname, ts, bd, it)


def test_quantize_lp():
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(12)
    for ts in (0, 1, 2, 7):
        n = H.max_eob(ts)
        for it in range(20):
            row = H.oracle_qrow(8, int(rng.integers(0, 256)))
            scan, iscan = H.oracle_scan(ts, 0)
            coeff = rng.integers(-4000, 4000, n).astype(np.int16)
            res = []
            for fn in (R.av1_quantize_lp_c, O.av1o_quantize_lp):
                qc = np.zeros(n, np.int16)
                dq = np.zeros(n, np.int16)
                eob = C.c_uint16(0)
                fn(ptr(coeff, i16p), C.c_ssize_t(n), row.round_fp, row.quant_fp, ptr(qc, i16p),
                   ptr(dq, i16p), row.dequant, C.byref(eob), ptr(scan, i16p), ptr(iscan, i16p))
                res.append((qc, dq, eob.value))
            assert res[0][2] == res[1][2]
            assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])


def test_block_error():
    """test/error_block_test.cc:84-160"""
    R, O = H.ref(), H.oracle()
    for f in (R.av1_block_error_c, R.av1_highbd_block_error_c, R.av1_block_error_lp_c):
        f.restype = C.c_int64
    rng = np.random.default_rng(13)
    for it in range(120):
        n = 16 << (it % 9) if it % 9 < 7 else 1024
        for bd in (8, 10, 12):
            msb = bd + 8 - 1
            coeff = rng.integers(0, 1 << msb, n).astype(np.int32)
            dq = rng.integers(0, 1 << msb, n).astype(np.int32)
            if it % 3 == 1:
                coeff, dq = -coeff, -dq
            if it == 5:
                coeff[:] = (1 << msb) - 1
                dq[:] = 0
            s0, s1 = C.c_int64(0), C.c_int64(0)
            if bd == 8:
                a = R.av1_block_error_c(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s0))
                b = O.av1o_block_error(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n), C.byref(s1))
                assert (a, s0.value) == (b, s1.value)
                c16, d16 = (coeff >> 2).astype(np.int16), (dq >> 2).astype(np.int16)
                assert R.av1_block_error_lp_c(ptr(c16, i16p), ptr(d16, i16p), C.c_ssize_t(n)) == \
                    O.av1o_block_error_lp(ptr(c16, i16p), ptr(d16, i16p), C.c_ssize_t(n))
            a = R.av1_highbd_block_error_c(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n),
                                           C.byref(s0), bd)
            b = O.av1o_highbd_block_error(ptr(coeff, i32p), ptr(dq, i32p), C.c_ssize_t(n),
                                          C.byref(s1), bd)
            assert (a, s0.value) == (b, s1.value)


def test_subtract_and_sse():
    R, O = H.ref(), H.oracle()
    R.aom_sse_c.restype = C.c_int64
    R.aom_highbd_sse_c.restype = C.c_int64
    rng = np.random.default_rng(14)
    for (w, h) in H.BLOCK_SIZES:
        src = rng.integers(0, 256, (h, w + 16)).astype(np.uint8)
        pred = rng.integers(0, 256, (h, w + 8)).astype(np.uint8)
        a = np.zeros((h, w), np.int16)
        b = np.zeros((h, w), np.int16)
        R.aom_subtract_block_c(h, w, ptr(a, i16p), C.c_ssize_t(w), ptr(src, u8p),
                               C.c_ssize_t(w + 16), ptr(pred, u8p), C.c_ssize_t(w + 8))
        O.av1o_subtract_block(h, w, ptr(b, i16p), C.c_ssize_t(w), ptr(src, u8p),
                              C.c_ssize_t(w + 16), ptr(pred, u8p), C.c_ssize_t(w + 8))
        assert np.array_equal(a, b)
        assert R.aom_sse_c(ptr(src, u8p), w + 16, ptr(pred, u8p), w + 8, w, h) == \
            O.av1o_sse(ptr(src, u8p), w + 16, ptr(pred, u8p), w + 8, w, h)
        s16 = rng.integers(0, 4096, (h, w + 16)).astype(np.uint16)
        p16 = rng.integers(0, 4096, (h, w + 8)).astype(np.uint16)
        R.aom_highbd_subtract_block_c(h, w, ptr(a, i16p), C.c_ssize_t(w), byteptr(s16),
                                      C.c_ssize_t(w + 16), byteptr(p16), C.c_ssize_t(w + 8))
        O.av1o_highbd_subtract_block(h, w, ptr(b, i16p), C.c_ssize_t(w), ptr(s16, u16p),
                                     C.c_ssize_t(w + 16), ptr(p16, u16p), C.c_ssize_t(w + 8))
        assert np.array_equal(a, b)
        assert R.aom_highbd_sse_c(byteptr(s16), w + 16, byteptr(p16), w + 8, w, h) == \
            O.av1o_highbd_sse(ptr(s16, u16p), w + 16, ptr(p16, u16p), w + 8, w, h)


@pytest.mark.parametrize("wh", H.BLOCK_SIZES, ids=["%dx%d" % s for s in H.BLOCK_SIZES])
def test_sad_family(wh):
    """test/sad_test.cc: MaxRef / MaxSrc / ShortRef / UnalignedRef / random"""
    w, h = wh
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(w * 131 + h)
    sstride, rstride = (w + 31) & ~31, 2 * w
    for it in range(8):
        src = rng.integers(0, 256, (h, sstride)).astype(np.uint8)
        refs = rng.integers(0, 256, (4, h + 1, rstride + 8)).astype(np.uint8)
        if it == 0:
            src[:] = 0
            refs[:] = 255
        if it == 1:
            src[:] = 255
            refs[:] = 0
        off = it % 3  # unaligned ref start
        rs = rstride + 8 if it % 2 == 0 else rstride - 1 if w > 4 else rstride + 8
        rp = [C.cast(C.c_void_p(refs[i].ctypes.data + off), u8p) for i in range(4)]
        s = ptr(src, u8p)
        nm = "%dx%d" % (w, h)
        assert getattr(R, "aom_sad%s_c" % nm)(s, sstride, rp[0], rs) == \
            O.av1o_sad(s, sstride, rp[0], rs, w, h)
        assert getattr(R, "aom_sad_skip_%s_c" % nm)(s, sstride, rp[1], rs) == \
            O.av1o_sad_skip(s, sstride, rp[1], rs, w, h)
        second = rng.integers(0, 256, w * h).astype(np.uint8)
        assert getattr(R, "aom_sad%s_avg_c" % nm)(s, sstride, rp[2], rs, ptr(second, u8p)) == \
            O.av1o_sad_avg(s, sstride, rp[2], rs, ptr(second, u8p), w, h)
        arr = (u8p * 4)(*rp)
        for sym, skip in (("aom_sad%sx4d_c" % nm, 0), ("aom_sad_skip_%sx4d_c" % nm, 1),
                          ("aom_sad%sx3d_c" % nm, 0)):
            a = np.zeros(4, np.uint32)
            b = np.zeros(4, np.uint32)
            getattr(R, sym)(s, sstride, arr, rs, ptr(a, u32p))
            O.av1o_sad_x4d(s, sstride, arr, rs, ptr(b, u32p), w, h, 4, skip)
            assert np.array_equal(a, b), sym
        for bd in (10, 12):
            s16 = rng.integers(0, 1 << bd, (h, sstride)).astype(np.uint16)
            r16 = rng.integers(0, 1 << bd, (h, rstride)).astype(np.uint16)
            if it == 0:
                s16[:] = 0
                r16[:] = (1 << bd) - 1
            assert getattr(R, "aom_highbd_sad%s_c" % nm)(byteptr(s16), sstride, byteptr(r16), rstride) == \
                O.av1o_highbd_sad(ptr(s16, u16p), sstride, ptr(r16, u16p), rstride, w, h)
            assert getattr(R, "aom_highbd_sad_skip_%s_c" % nm)(byteptr(s16), sstride, byteptr(r16), rstride) == \
                O.av1o_highbd_sad_skip(ptr(s16, u16p), sstride, ptr(r16, u16p), rstride, w, h)


@pytest.mark.parametrize("wh", H.BLOCK_SIZES, ids=["%dx%d" % s for s in H.BLOCK_SIZES])
def test_variance_family(wh):
    """test/variance_test.cc: Zero / Ref / RefStride / OneQuarter / Extreme"""
    w, h = wh
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(w * 17 + h)
    nm = "%dx%d" % (w, h)
    for it in range(6):
        a = rng.integers(0, 256, (h + 1, w + 9)).astype(np.uint8)
        b = rng.integers(0, 256, (h, w + 3)).astype(np.uint8)
        if it == 0:
            a[:] = 255
            b[:] = 0
        if it == 1:  # OneQuarter: half of b is 255, a is 0 -> var = N*255^2/4
            a[:] = 0
            b[:] = 0
            b[: h // 2, :] = 255
        s0, s1 = C.c_uint(0), C.c_uint(0)
        v0 = getattr(R, "aom_variance%s_c" % nm)(ptr(a, u8p), w + 9, ptr(b, u8p), w + 3, C.byref(s0))
        v1 = O.av1o_variance(ptr(a, u8p), w + 9, ptr(b, u8p), w + 3, w, h, C.byref(s1))
        assert (C.c_uint(v0).value, s0.value) == (v1, s1.value)
        if it == 1:
            assert v1 == w * h * 255 * 255 // 4
        for xo in range(8):
            for yo in range(8):
                if (xo * 8 + yo + it) % 3:
                    continue
                v0 = getattr(R, "aom_sub_pixel_variance%s_c" % nm)(
                    ptr(a, u8p), w + 9, xo, yo, ptr(b, u8p), w + 3, C.byref(s0))
                v1 = O.av1o_sub_pixel_variance(ptr(a, u8p), w + 9, xo, yo, ptr(b, u8p), w + 3,
                                               w, h, C.byref(s1))
                assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), (xo, yo)
        for bd in (8, 10, 12):
            a16 = rng.integers(0, 1 << bd, (h + 1, w + 9)).astype(np.uint16)
            b16 = rng.integers(0, 1 << bd, (h, w + 3)).astype(np.uint16)
            if it == 0:
                a16[:] = (1 << bd) - 1
                b16[:] = 0
            v0 = getattr(R, "aom_highbd_%d_variance%s_c" % (bd, nm))(
                byteptr(a16), w + 9, byteptr(b16), w + 3, C.byref(s0))
            v1 = O.av1o_highbd_variance(ptr(a16, u16p), w + 9, ptr(b16, u16p), w + 3, w, h, bd,
                                        C.byref(s1))
            assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), bd
            xo, yo = int(rng.integers(0, 8)), int(rng.integers(0, 8))
            v0 = getattr(R, "aom_highbd_%d_sub_pixel_variance%s_c" % (bd, nm))(
                byteptr(a16), w + 9, xo, yo, byteptr(b16), w + 3, C.byref(s0))
            v1 = O.av1o_highbd_sub_pixel_variance(ptr(a16, u16p), w + 9, xo, yo, ptr(b16, u16p),
                                                  w + 3, w, h, bd, C.byref(s1))
            assert (C.c_uint(v0).value, s0.value) == (v1, s1.value), (bd, xo, yo)


def test_upsampled_pred():
    """aom_upsampled_pred_c with xd == NULL ("expect xd == NULL only in tests",
    reconinter_enc.c:430): USE_2_TAPS=1?, see mcomp.h enum -> pass subpel_search ids"""
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(15)
    # SUBPEL_SEARCH_TYPE: USE_2_TAPS_ORIG=0, USE_2_TAPS=1, USE_4_TAPS=2, USE_8_TAPS=3
    for (w, h) in [(4, 4), (8, 8), (16, 16), (64, 64), (16, 8), (32, 64), (128, 128)]:
        refimg = rng.integers(0, 256, (h + 16, w + 24)).astype(np.uint8)
        stride = w + 24
        base = refimg.ctypes.data + 8 * stride + 8
        for sx in range(8):
            for sy in range(8):
                if (sx + 3 * sy) % 4 and w > 16:
                    continue
                for search, taps in ((1, 2), (2, 4), (3, 8)):
                    a = np.zeros(w * h, np.uint8)
                    b = np.zeros(w * h, np.uint8)
                    R.aom_upsampled_pred_c(None, None, 0, 0, None, ptr(a, u8p), w, h, sx, sy,
                                           C.cast(C.c_void_p(base), u8p), stride, search)
                    O.av1o_upsampled_pred(ptr(b, u8p), w, h, sx, sy,
                                          C.cast(C.c_void_p(base), u8p), stride, taps)
                    assert np.array_equal(a, b), (w, h, sx, sy, taps)
        for bd in (10, 12):
            ref16 = rng.integers(0, 1 << bd, (h + 16, w + 24)).astype(np.uint16)
            base16 = ref16.ctypes.data + 2 * (8 * stride + 8)
            for sx, sy in ((0, 0), (3, 0), (0, 5), (7, 7), (4, 2)):
                for search, taps in ((1, 2), (2, 4), (3, 8)):
                    a = np.zeros(w * h, np.uint16)
                    b = np.zeros(w * h, np.uint16)
                    R.aom_highbd_upsampled_pred_c(None, None, 0, 0, None, byteptr(a), w, h, sx, sy,
                                                  C.cast(C.c_void_p(base16 >> 1), u8p), stride,
                                                  bd, search)
                    O.av1o_highbd_upsampled_pred(ptr(b, u16p), w, h, sx, sy,
                                                 C.cast(C.c_void_p(base16), u16p), stride, bd, taps)
                    assert np.array_equal(a, b), (w, h, sx, sy, taps, bd)
