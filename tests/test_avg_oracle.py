"""Pin the oracle's Hadamard / SATD / average family (oracle/av1o_avg.c) against the
reference's aom_dsp/avg.c `_c` functions (oracle/_ref), on the input ranges of the
reference's own tests (test/hadamard_test.cc: 9-bit residuals for the low-bit-depth
functions, 13-bit for highbd, all-max / all-min extremes; test/avg_test.cc)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, ptr, u8p

pytestmark = [pytest.mark.ref,
              pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]

HADAMARDS = [("aom_hadamard_4x4_c", 4, 0), ("aom_hadamard_8x8_c", 8, 0),
             ("aom_hadamard_16x16_c", 16, 0), ("aom_hadamard_32x32_c", 32, 0),
             ("aom_hadamard_lp_8x8_c", 8, 1), ("aom_hadamard_lp_16x16_c", 16, 1),
             ("aom_highbd_hadamard_8x8_c", 8, 2), ("aom_highbd_hadamard_16x16_c", 16, 2),
             ("aom_highbd_hadamard_32x32_c", 32, 2)]


def residual(rng, n, stride, kind, it):
    lim = 4095 if kind == 2 else 255
    if it == 0:
        return np.full((n, stride), lim, np.int16)
    if it == 1:
        return np.full((n, stride), -lim, np.int16)
    if it == 2:  # beyond the documented range: int16 wrap-around must match too
        return rng.integers(-32768, 32768, (n, stride)).astype(np.int16)
    return rng.integers(-lim, lim + 1, (n, stride)).astype(np.int16)


@pytest.mark.parametrize("name,n,kind", HADAMARDS)
def test_hadamard(name, n, kind):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(1100 + n + kind)
    for it in range(40):
        stride = n + 8 * (it % 3)
        src = residual(rng, n, stride, kind, it)
        b = np.zeros(n * n, np.int32)
        O.av1o_hadamard(ptr(src, i16p), C.c_ssize_t(stride), ptr(b, i32p), n, kind)
        if kind == 1:
            a = np.zeros(n * n, np.int16)
            getattr(R, name)(ptr(src, i16p), C.c_ssize_t(stride), ptr(a, i16p))
        else:
            a = np.zeros(n * n, np.int32)
            getattr(R, name)(ptr(src, i16p), C.c_ssize_t(stride), ptr(a, i32p))
        assert np.array_equal(a.astype(np.int32), b), (name, it)
        assert R.aom_satd_c(ptr(b, i32p), n * n) == O.av1o_satd(ptr(b, i32p), n * n)
        if kind == 1:
            assert R.aom_satd_lp_c(ptr(a, i16p), n * n) == O.av1o_satd(ptr(b, i32p), n * n)


def test_hadamard_lp_8x8_dual():
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(1200)
    src = rng.integers(-255, 256, (8, 24)).astype(np.int16)
    a = np.zeros(128, np.int16)
    R.aom_hadamard_lp_8x8_dual_c(ptr(src, i16p), C.c_ssize_t(24), ptr(a, i16p))
    for i in range(2):
        b = np.zeros(64, np.int32)
        O.av1o_hadamard(C.cast(src.ctypes.data + 16 * i, i16p), C.c_ssize_t(24), ptr(b, i32p), 8, 1)
        assert np.array_equal(a[64 * i: 64 * i + 64].astype(np.int32), b)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_avg_minmax(bd):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(1300 + bd)
    from test_oracle_vs_ref import byteptr
    for it in range(30):
        stride = 16 + 8 * (it % 2)
        hi = (1 << bd) - 1
        if bd == 8:
            s = rng.integers(0, 256, (8, stride)).astype(np.uint8)
            d = rng.integers(0, 256, (8, stride)).astype(np.uint8)
            if it == 0:
                s[:], d[:] = 255, 0
            sp, dp = ptr(s, u8p), ptr(d, u8p)
            f4, f8, fm = R.aom_avg_4x4_c, R.aom_avg_8x8_c, R.aom_minmax_8x8_c
        else:
            s = rng.integers(0, hi + 1, (8, stride)).astype(np.uint16)
            d = np.clip(s.astype(np.int32) + rng.integers(-200, 201, s.shape), 0, hi).astype(np.uint16)
            if it == 0:
                s[:] = hi
            sp, dp = byteptr(s), byteptr(d)
            f4, f8, fm = R.aom_highbd_avg_4x4_c, R.aom_highbd_avg_8x8_c, R.aom_highbd_minmax_8x8_c
        assert f4(sp, stride) == O.av1o_avg(C.c_void_p(s.ctypes.data), stride, 4, int(bd > 8))
        assert f8(sp, stride) == O.av1o_avg(C.c_void_p(s.ctypes.data), stride, 8, int(bd > 8))
        mn, mx, mn2, mx2 = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        fm(sp, stride, dp, stride, C.byref(mn), C.byref(mx))
        O.av1o_minmax_8x8(C.c_void_p(s.ctypes.data), stride, C.c_void_p(d.ctypes.data), stride,
                          int(bd > 8), C.byref(mn2), C.byref(mx2))
        assert (mn.value, mx.value) == (mn2.value, mx2.value)


def test_avg_8x8_quad():
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(1400)
    s = rng.integers(0, 256, (64, 64)).astype(np.uint8)
    for (x16, y16) in ((0, 0), (16, 32), (48, 48)):
        avg = (C.c_int * 4)()
        R.aom_avg_8x8_quad_c(ptr(s, u8p), 64, x16, y16, avg)
        for k in range(4):
            x8, y8 = x16 + ((k & 1) << 3), y16 + ((k >> 1) << 3)
            assert avg[k] == O.av1o_avg(C.c_void_p(s.ctypes.data + y8 * 64 + x8), 64, 8, 0)


def test_int_pro_and_vector_var():
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(1500)
    for it in range(20):
        w = int(rng.choice([16, 32, 64, 128]))
        h = int(rng.choice([16, 32, 64, 128]))
        stride = w + 16
        ref = rng.integers(0, 256, (h, stride)).astype(np.uint8)
        if it == 0:
            ref[:] = 255
        for norm in (3, 4, 5, 6):
            a, b = np.zeros(w, np.int16), np.zeros(w, np.int16)
            R.aom_int_pro_row_c(ptr(a, i16p), ptr(ref, u8p), stride, w, h, norm)
            O.av1o_int_pro_row(ptr(b, i16p), ptr(ref, u8p), stride, w, h, norm)
            assert np.array_equal(a, b)
            a, b = np.zeros(h, np.int16), np.zeros(h, np.int16)
            R.aom_int_pro_col_c(ptr(a, i16p), ptr(ref, u8p), stride, w, h, norm)
            O.av1o_int_pro_col(ptr(b, i16p), ptr(ref, u8p), stride, w, h, norm)
            assert np.array_equal(a, b)
        for bwl in (2, 3, 4, 5):
            n = 4 << bwl
            x = rng.integers(0, 511, n).astype(np.int16)
            y = rng.integers(0, 511, n).astype(np.int16)
            if it == 0:
                x[:], y[:] = 510, 0
            assert R.aom_vector_var_c(ptr(x, i16p), ptr(y, i16p), bwl) == \
                O.av1o_vector_var(ptr(x, i16p), ptr(y, i16p), bwl)
