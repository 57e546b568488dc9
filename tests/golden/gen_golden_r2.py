#!/usr/bin/env python3
"""Generate tests/golden/r2_golden.npz from the REFERENCE build (oracle/_ref): inverse transforms
(av1_inv_txfm2d_add_WxH_c, av1_inv_txfm_add_c incl. the lossless WHT), the Laplacian rate estimate
(av1_cost_coeffs_txb_estimate) and the compound / masked / OBMC SAD and variance families, for machines
where /root/reference (and possibly oracle/_ref) is absent.  Run here: python tests/golden/gen_golden_r2.py"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402
from harness import i32p, ptr, u8p, u16p  # noqa: E402
from test_compound_oracle import DistWtd, bp  # noqa: E402

R = H.ref()
assert R is not None, "oracle/_ref not built"
rng = np.random.default_rng(20261105)
out = {}
# inverse transforms: one block per (tx_size, tx_type, bd): coefficients = quantised forward transform
for ts in range(19):
    w, h, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
    fn = getattr(R, "av1_inv_txfm2d_add_%dx%d_c" % (w, h))
    for bd in ((8, 10, 12) if n <= 256 else (10,)):
        co = np.zeros((16, n), np.int32)
        pr = rng.integers(0, 1 << bd, (16, h, w)).astype(np.uint16)
        rc = np.zeros((16, h, w), np.uint16)
        for tt in (0, 1, 5, 6, 9, 10, 13, 15):
            if not H.tx_valid(ts, tt):
                continue
            resid = rng.integers(-(1 << bd) + 1, 1 << bd, (h, w)).astype(np.int16)
            c = H.oracle_fwd_txfm2d(resid, w, ts, tt, bd)
            q = int(rng.integers(1, 64))
            co[tt] = (c // q) * q
            if tt in (9, 13):
                co[tt] = rng.integers(-(1 << (bd + 7)), 1 << (bd + 7), n)
            a = pr[tt].copy()
            fn(ptr(np.ascontiguousarray(co[tt]), i32p), ptr(a, u16p), w, C.c_uint8(tt), bd)
            rc[tt] = a
        out["itx_co_%d_%d" % (ts, bd)], out["itx_pred_%d_%d" % (ts, bd)], out["itx_recon_%d_%d" % (ts, bd)] = co, pr, rc
# lossless WHT through av1_highbd_iwht4x4_add (eob switch)
R.av1_highbd_iwht4x4_add.restype = None
for eob in (1, 2, 16):
    c = rng.integers(-2000, 2000, 16).astype(np.int32)
    if eob == 1:
        c[1:] = 0
    p = rng.integers(0, 1024, (4, 4)).astype(np.uint16)
    a = p.copy()
    R.av1_highbd_iwht4x4_add(ptr(c, i32p), C.cast(C.c_void_p(a.ctypes.data >> 1), u8p), 4, eob, 10)
    out["iwht_co_%d" % eob], out["iwht_pred_%d" % eob], out["iwht_recon_%d" % eob] = c, p, a
# rate estimate
est = []
for ts in range(19):
    n = H.max_eob(ts)
    for tt in (0, 9, 10, 11):
        if not H.tx_valid(ts, tt):
            continue
        scan, _ = H.oracle_scan(ts, tt)
        q = np.zeros(n, np.int32)
        eob = int(rng.integers(1, n + 1))
        q[scan[:eob]] = rng.integers(-20, 21, eob)
        if q[scan[eob - 1]] == 0:
            q[scan[eob - 1]] = 3
        est.append((ts, tt, eob, R.refshim_cost_coeffs_txb_estimate(ptr(q, i32p), eob, ts, tt)))
        out["est_q_%d_%d" % (ts, tt)] = q
out["est"] = np.array(est, np.int64)
# compound families on one set of planes, sizes 8x8 16x32 64x64 4x16, 8 bit and 10 bit
for hb, bd in ((0, 8), (1, 10)):
    dt = np.uint16 if hb else np.uint8
    hi = (1 << bd) - 1
    pre = rng.integers(0, hi + 1, (70, 80)).astype(dt)
    tgt = rng.integers(0, hi + 1, (66, 72)).astype(dt)
    mask = rng.integers(0, 65, (66, 70)).astype(np.uint8)
    second = rng.integers(0, hi + 1, 64 * 64).astype(dt)
    wsrc = (rng.integers(0, hi + 1, 64 * 64) * rng.integers(0, 4097, 64 * 64)).astype(np.int32)
    omask = rng.integers(0, 4097, 64 * 64).astype(np.int32)
    tag = "cmp%d_" % bd
    for k, v in dict(pre=pre, tgt=tgt, mask=mask, second=second, wsrc=wsrc, omask=omask).items():
        out[tag + k] = v
    jcp = DistWtd(1, 11, 5)
    rows = []
    pfx = ("aom_highbd_%d_" % bd) if hb else "aom_"
    sp = "aom_highbd_" if hb else "aom_"
    op = ("aom_highbd_%d_obmc_" % bd) if hb else "aom_obmc_"
    for (w, h) in ((8, 8), (16, 32), (64, 64), (4, 16)):
        sse = C.c_uint(0)
        i32 = C.POINTER(C.c_int32)
        r = [w, h]
        r.append(getattr(R, sp + "masked_sad%dx%d_c" % (w, h))(bp(tgt, hb), 72, bp(pre, hb), 80, bp(second, hb),
                                                              ptr(mask, u8p), 70, 1))
        r.append(getattr(R, pfx + "masked_sub_pixel_variance%dx%d_c" % (w, h))(
            bp(pre, hb), 80, 3, 5, bp(tgt, hb), 72, bp(second, hb), ptr(mask, u8p), 70, 0, C.byref(sse)))
        r.append(sse.value)
        r.append(getattr(R, pfx + "sub_pixel_avg_variance%dx%d_c" % (w, h))(
            bp(pre, hb), 80, 3, 5, bp(tgt, hb), 72, C.byref(sse), bp(second, hb)))
        r.append(sse.value)
        r.append(getattr(R, pfx + "dist_wtd_sub_pixel_avg_variance%dx%d_c" % (w, h))(
            bp(pre, hb), 80, 3, 5, bp(tgt, hb), 72, C.byref(sse), bp(second, hb), C.byref(jcp)))
        r.append(sse.value)
        r.append(getattr(R, sp + "dist_wtd_sad%dx%d_avg_c" % (w, h))(bp(tgt, hb), 72, bp(pre, hb), 80,
                                                                    bp(second, hb), C.byref(jcp)))
        r.append(getattr(R, sp + "obmc_sad%dx%d_c" % (w, h))(bp(pre, hb), 80, wsrc.ctypes.data_as(i32),
                                                            omask.ctypes.data_as(i32)))
        r.append(getattr(R, op + "variance%dx%d_c" % (w, h))(bp(pre, hb), 80, wsrc.ctypes.data_as(i32),
                                                            omask.ctypes.data_as(i32), C.byref(sse)))
        r.append(sse.value)
        r.append(getattr(R, op + "sub_pixel_variance%dx%d_c" % (w, h))(bp(pre, hb), 80, 3, 5, wsrc.ctypes.data_as(i32),
                                                                      omask.ctypes.data_as(i32), C.byref(sse)))
        r.append(sse.value)
        rows.append([int(v) & 0xffffffff for v in r])
    out[tag + "expect"] = np.array(rows, np.int64)
np.savez_compressed(os.path.join(HERE, "r2_golden.npz"), **out)
print("wrote r2_golden.npz:", len(out), "arrays,", os.path.getsize(os.path.join(HERE, "r2_golden.npz")), "bytes")
