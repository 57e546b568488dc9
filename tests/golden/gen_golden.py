#!/usr/bin/env python3
"""Generate tests/golden/hotpath_golden.npz from the REFERENCE build (oracle/_ref): seeded
inputs and the outputs of the reference's own C functions, for machines where
/root/reference (and possibly oracle/_ref) is absent.  Run here: python tests/golden/gen_golden.py"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402
from harness import i16p, i32p, ptr, u8p  # noqa: E402

R = H.ref()
assert R is not None, "oracle/_ref not built"
rng = np.random.default_rng(20260101)
out = {}
# transforms: one block per (tx_size, tx_type), 10-bit random residual
for ts in range(19):
    W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
    inp = rng.integers(-1023, 1024, (16, Hh, W)).astype(np.int16)
    res = np.zeros((16, n), np.int32)
    fn = getattr(R, "av1_fwd_txfm2d_%dx%d_c" % (W, Hh))
    for tt in range(16):
        if not H.tx_valid(ts, tt):
            continue
        tmp = np.zeros(W * Hh + 64, np.int32)
        fn(ptr(np.ascontiguousarray(inp[tt]), i16p), ptr(tmp, i32p), W, C.c_uint8(tt), 10)
        res[tt] = tmp[:n]
    out["txfm_in_%d" % ts] = inp
    out["txfm_out_%d" % ts] = res
# quantize_fp on the DCT_DCT coefficients of every size at qindex 60 (8-bit tables)
row = H.oracle_qrow(8, 60)
for ts in range(19):
    n, ls = H.max_eob(ts), H.log_scale(ts)
    coeff = np.clip(out["txfm_out_%d" % ts][0], -32767, 32767).astype(np.int32)
    scan, iscan = H.oracle_scan(ts, 0)
    qc, dq, eob = np.zeros(n, np.int32), np.zeros(n, np.int32), C.c_uint16(0)
    fn = [R.av1_quantize_fp_c, R.av1_quantize_fp_32x32_c, R.av1_quantize_fp_64x64_c][ls]
    fn(ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round_fp, row.quant_fp, row.quant_shift,
       ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob), ptr(scan, i16p), ptr(iscan, i16p))
    out["q_coeff_%d" % ts], out["q_qc_%d" % ts], out["q_dq_%d" % ts] = coeff, qc, dq
    out["q_eob_%d" % ts] = np.array([eob.value], np.int32)
# SAD / variance / sub-pel variance on one pair of 8-bit images
a = rng.integers(0, 256, (136, 144)).astype(np.uint8)
b = rng.integers(0, 256, (136, 144)).astype(np.uint8)
out["pix_a"], out["pix_b"] = a, b
pix = []
for (w, h) in H.BLOCK_SIZES:
    nm = "%dx%d" % (w, h)
    sse, sse2 = C.c_uint(0), C.c_uint(0)
    sad = getattr(R, "aom_sad%s_c" % nm)(ptr(a, u8p), 144, ptr(b, u8p), 144)
    var = getattr(R, "aom_variance%s_c" % nm)(ptr(a, u8p), 144, ptr(b, u8p), 144, C.byref(sse))
    spv = getattr(R, "aom_sub_pixel_variance%s_c" % nm)(ptr(a, u8p), 144, 3, 5, ptr(b, u8p), 144, C.byref(sse2))
    pix.append([sad, C.c_uint(var).value, sse.value, C.c_uint(spv).value, sse2.value])
out["pix_results"] = np.array(pix, np.int64)
np.savez_compressed(os.path.join(HERE, "hotpath_golden.npz"), **out)
print("wrote", os.path.join(HERE, "hotpath_golden.npz"))
