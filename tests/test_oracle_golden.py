"""Oracle vs committed golden vectors (tests/golden/hotpath_golden.npz, produced from the
reference build by tests/golden/gen_golden.py).  Runs anywhere -- no /root/reference."""
import ctypes as C
import os

import numpy as np

import harness as H
from harness import i16p, i32p, ptr, u8p

G = np.load(os.path.join(H.ROOT, "tests", "golden", "hotpath_golden.npz"))


def test_txfm_golden():
    for ts in range(19):
        W = H.TX_W[ts]
        inp, exp = G["txfm_in_%d" % ts], G["txfm_out_%d" % ts]
        for tt in range(16):
            if not H.tx_valid(ts, tt):
                continue
            got = H.oracle_fwd_txfm2d(np.ascontiguousarray(inp[tt]), W, ts, tt, 10)
            assert np.array_equal(got, exp[tt]), (ts, tt)


def test_quantize_fp_golden():
    O = H.oracle()
    row = H.oracle_qrow(8, 60)
    for ts in range(19):
        n, ls = H.max_eob(ts), H.log_scale(ts)
        coeff = np.ascontiguousarray(G["q_coeff_%d" % ts])
        scan, iscan = H.oracle_scan(ts, 0)
        qc, dq, eob = np.zeros(n, np.int32), np.zeros(n, np.int32), C.c_uint16(0)
        O.av1o_quantize_fp(ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round_fp, row.quant_fp,
                           row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob),
                           ptr(scan, i16p), ptr(iscan, i16p), None, None, ls)
        assert np.array_equal(qc, G["q_qc_%d" % ts]) and np.array_equal(dq, G["q_dq_%d" % ts])
        assert eob.value == int(G["q_eob_%d" % ts][0])


def test_pixel_golden():
    O = H.oracle()
    a, b = np.ascontiguousarray(G["pix_a"]), np.ascontiguousarray(G["pix_b"])
    for (w, h), exp in zip(H.BLOCK_SIZES, G["pix_results"]):
        sse, sse2 = C.c_uint(0), C.c_uint(0)
        got = [O.av1o_sad(ptr(a, u8p), 144, ptr(b, u8p), 144, w, h),
               O.av1o_variance(ptr(a, u8p), 144, ptr(b, u8p), 144, w, h, C.byref(sse)), 0,
               O.av1o_sub_pixel_variance(ptr(a, u8p), 144, 3, 5, ptr(b, u8p), 144, w, h, C.byref(sse2)), 0]
        got[2], got[4] = sse.value, sse2.value
        assert got == [int(v) for v in exp], (w, h)
