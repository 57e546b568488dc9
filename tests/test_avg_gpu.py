"""GPU parity of the Hadamard / SATD / WHT / average family (b200_hadamard_launch,
b200_avg_launch and the aom_*_b200 drop-in entries) against the oracle, which
tests/test_avg_oracle.py pins to the reference's aom_dsp/avg.c.  Bit-exact."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, ptr, u8p

pytestmark = pytest.mark.gpu

CASES = [(4, 0), (8, 0), (16, 0), (32, 0), (8, 1), (16, 1), (8, 2), (16, 2), (32, 2), (4, 3)]


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def oracle_block(O, plane, off, stride, n, kind):
    out = np.zeros(n * n, np.int32)
    src = C.cast(plane.ctypes.data + 2 * off, i16p)
    if kind == 3:
        O.av1o_fwht4x4(src, ptr(out, i32p), stride)
    else:
        O.av1o_hadamard(src, C.c_ssize_t(stride), ptr(out, i32p), n, kind)
    return out


def run_hadamard(ctx, plane, blocks, n, kind, want_coeff=True, want_satd=True):
    import b200av1
    nb = len(blocks)
    d_src, d_blk = b200av1.to_device(ctx, plane), b200av1.to_device(ctx, blocks)
    esz = 2 if kind == b200av1.HAD_LP else 4
    d_coeff = b200av1.DevBuf(ctx, nb * n * n * esz)
    d_satd = b200av1.DevBuf(ctx, nb * 4)
    l = b200av1.HadamardLaunch()
    l.n, l.kind, l.n_blocks = n, kind, nb
    l.blocks, l.src = d_blk.ptr, d_src.ptr
    l.coeff = d_coeff.ptr if want_coeff else None
    l.satd = d_satd.ptr if want_satd else None
    b200av1.check(b200av1.lib().b200_hadamard_launch(ctx.h, C.byref(l), None))
    coeff = b200av1.to_host(ctx, d_coeff, np.int16 if esz == 2 else np.int32, nb * n * n)
    satd = b200av1.to_host(ctx, d_satd, np.int32, nb)
    return coeff.astype(np.int32), satd


@pytest.mark.parametrize("n,kind", CASES)
def test_hadamard_batch_vs_oracle(ctx, n, kind):
    """a ragged batch (block count not a multiple of the warp tile), unaligned offsets,
    full-range int16 residuals in part of the plane so the int16 wrap-around is exercised"""
    O = H.oracle()
    rng = np.random.default_rng(2100 + n * 4 + kind)
    w, h = 200, 136
    lim = 4095 if kind == 2 else 255
    plane = rng.integers(-lim, lim + 1, (h, w)).astype(np.int16)
    plane[:40] = rng.integers(-32768, 32768, (40, w)).astype(np.int16)
    plane[40:48] = lim
    plane[48:56] = -lim
    nb = 333
    blocks = np.zeros(nb, __import__("b200av1").HADBLOCK_DTYPE)
    for i in range(nb):
        x, y = int(rng.integers(0, w - n + 1)), int(rng.integers(0, h - n + 1))
        blocks[i] = (y * w + x, i * n * n, w)
    coeff, satd = run_hadamard(ctx, plane, blocks, n, kind)
    for i in range(nb):
        e = oracle_block(O, plane, int(blocks[i]["src_offset"]), w, n, kind)
        assert np.array_equal(coeff[i * n * n:(i + 1) * n * n], e), (n, kind, i)
        assert int(satd[i]) == O.av1o_satd(ptr(e, i32p), n * n), (n, kind, i)
    # SATD-only launch (no coefficient stores): same sums
    _, satd2 = run_hadamard(ctx, plane, blocks, n, kind, want_coeff=False)
    assert np.array_equal(satd, satd2)


def test_hadamard_full_frame_linearity(ctx):
    """size-independent property at 1080p: below the int16 wrap the 8x8 Hadamard is linear,
    H(a + b) = H(a) + H(b) on every block of the frame, and SATD(0) = 0."""
    import b200av1
    rng = np.random.default_rng(2200)
    w, h, n = 1920, 1080, 8
    a = rng.integers(-120, 121, (h, w)).astype(np.int16)
    b = rng.integers(-120, 121, (h, w)).astype(np.int16)
    bx, by = np.meshgrid(np.arange(0, w, n), np.arange(0, h, n))
    nb = bx.size
    blocks = np.zeros(nb, b200av1.HADBLOCK_DTYPE)
    blocks["src_offset"] = (by * w + bx).ravel()
    blocks["out_offset"] = np.arange(nb) * n * n
    blocks["stride"] = w
    ca, _ = run_hadamard(ctx, a, blocks, n, 0)
    cb, _ = run_hadamard(ctx, b, blocks, n, 0)
    cs, _ = run_hadamard(ctx, (a + b).astype(np.int16), blocks, n, 0)
    assert np.array_equal(ca + cb, cs)
    _, z = run_hadamard(ctx, np.zeros((h, w), np.int16), blocks, n, 0)
    assert not z.any()


def test_hadamard_rejects_missing_sizes(ctx):
    import b200av1
    l = b200av1.HadamardLaunch()
    dummy = b200av1.DevBuf(ctx, 64)
    l.n, l.kind, l.n_blocks = 32, b200av1.HAD_LP, 1
    l.blocks = l.src = l.coeff = dummy.ptr
    assert b200av1.lib().b200_hadamard_launch(ctx.h, C.byref(l), None) != 0


@pytest.mark.parametrize("bd", [8, 10])
def test_avg_batch_vs_oracle(ctx, bd):
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(2300 + bd)
    w, h = 160, 96
    dt = np.uint8 if bd == 8 else np.uint16
    a = rng.integers(0, 1 << bd, (h, w)).astype(dt)
    b = np.clip(a.astype(np.int32) + rng.integers(-200, 201, a.shape), 0, (1 << bd) - 1).astype(dt)
    a[:8] = (1 << bd) - 1
    nb = 257
    blocks = np.zeros(nb, b200av1.AVGBLOCK_DTYPE)
    for i in range(nb):
        x, y = int(rng.integers(0, w - 8 + 1)), int(rng.integers(0, h - 8 + 1))
        blocks[i] = (y * w + x, y * w + x, w, w)
    d_a, d_b, d_blk = (b200av1.to_device(ctx, v) for v in (a, b, blocks))
    d_out = b200av1.DevBuf(ctx, nb * 8)
    for op in (b200av1.AVG_4X4, b200av1.AVG_8X8, b200av1.MINMAX_8X8):
        l = b200av1.AvgLaunch()
        l.op, l.highbd, l.n_blocks = op, int(bd > 8), nb
        l.blocks, l.a, l.b, l.out = d_blk.ptr, d_a.ptr, d_b.ptr, d_out.ptr
        b200av1.check(b200av1.lib().b200_avg_launch(ctx.h, C.byref(l), None))
        out = b200av1.to_host(ctx, d_out, np.int32, 2 * nb).reshape(nb, 2)
        for i in range(nb):
            off = int(blocks[i]["a_offset"]) * a.itemsize
            pa, pb = C.c_void_p(a.ctypes.data + off), C.c_void_p(b.ctypes.data + off)
            if op == b200av1.MINMAX_8X8:
                mn, mx = C.c_int(), C.c_int()
                O.av1o_minmax_8x8(pa, w, pb, w, int(bd > 8), C.byref(mn), C.byref(mx))
                assert (int(out[i, 0]), int(out[i, 1])) == (mn.value, mx.value), (op, i)
            else:
                assert int(out[i, 0]) == O.av1o_avg(pa, w, 4 if op == b200av1.AVG_4X4 else 8,
                                                    int(bd > 8)), (op, i)


def test_dropin_entries_vs_oracle():
    """the RTCD-shaped `_b200` functions with host pointers (one block per call)"""
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    rng = np.random.default_rng(2400)
    stride = 48
    names = {(4, 0): "aom_hadamard_4x4_b200", (8, 0): "aom_hadamard_8x8_b200",
             (16, 0): "aom_hadamard_16x16_b200", (32, 0): "aom_hadamard_32x32_b200",
             (8, 1): "aom_hadamard_lp_8x8_b200", (16, 1): "aom_hadamard_lp_16x16_b200",
             (8, 2): "aom_highbd_hadamard_8x8_b200", (16, 2): "aom_highbd_hadamard_16x16_b200",
             (32, 2): "aom_highbd_hadamard_32x32_b200"}
    for (n, kind), name in names.items():
        src = rng.integers(-255, 256, (n, stride)).astype(np.int16)
        e = oracle_block(O, src, 0, stride, n, kind)
        if kind == 1:
            g = np.zeros(n * n, np.int16)
            getattr(L, name)(ptr(src, i16p), C.c_ssize_t(stride), ptr(g, i16p))
            assert L.aom_satd_lp_b200(ptr(g, i16p), n * n) == O.av1o_satd(ptr(e, i32p), n * n)
        else:
            g = np.zeros(n * n, np.int32)
            getattr(L, name)(ptr(src, i16p), C.c_ssize_t(stride), ptr(g, i32p))
            assert L.aom_satd_b200(ptr(g, i32p), n * n) == O.av1o_satd(ptr(e, i32p), n * n)
        assert np.array_equal(g.astype(np.int32), e), name
    # dual 8x8 LP, WHT
    src = rng.integers(-255, 256, (8, stride)).astype(np.int16)
    g = np.zeros(128, np.int16)
    L.aom_hadamard_lp_8x8_dual_b200(ptr(src, i16p), C.c_ssize_t(stride), ptr(g, i16p))
    for i in range(2):
        assert np.array_equal(g[64 * i:64 * i + 64].astype(np.int32), oracle_block(O, src, 8 * i, stride, 8, 1))
    src = rng.integers(-255, 256, (4, stride)).astype(np.int16)
    g = np.zeros(16, np.int32)
    L.av1_fwht4x4_b200(ptr(src, i16p), ptr(g, i32p), stride)
    assert np.array_equal(g, oracle_block(O, src, 0, stride, 4, 3))
    # averages, min/max, quad
    s = rng.integers(0, 256, (32, stride)).astype(np.uint8)
    d = rng.integers(0, 256, (32, stride)).astype(np.uint8)
    L.aom_avg_4x4_b200.restype = L.aom_avg_8x8_b200.restype = C.c_uint
    assert L.aom_avg_4x4_b200(ptr(s, u8p), stride) == O.av1o_avg(C.c_void_p(s.ctypes.data), stride, 4, 0)
    assert L.aom_avg_8x8_b200(ptr(s, u8p), stride) == O.av1o_avg(C.c_void_p(s.ctypes.data), stride, 8, 0)
    mn, mx, mn2, mx2 = C.c_int(), C.c_int(), C.c_int(), C.c_int()
    L.aom_minmax_8x8_b200(ptr(s, u8p), stride, ptr(d, u8p), stride, C.byref(mn), C.byref(mx))
    O.av1o_minmax_8x8(C.c_void_p(s.ctypes.data), stride, C.c_void_p(d.ctypes.data), stride, 0,
                      C.byref(mn2), C.byref(mx2))
    assert (mn.value, mx.value) == (mn2.value, mx2.value)
    avg = (C.c_int * 4)()
    L.aom_avg_8x8_quad_b200(ptr(s, u8p), stride, 16, 8, avg)
    for k in range(4):
        x8, y8 = 16 + ((k & 1) << 3), 8 + ((k >> 1) << 3)
        assert avg[k] == O.av1o_avg(C.c_void_p(s.ctypes.data + y8 * stride + x8), stride, 8, 0)
    s16 = rng.integers(0, 1024, (8, stride)).astype(np.uint16)
    L.aom_highbd_avg_8x8_b200.restype = C.c_uint
    assert L.aom_highbd_avg_8x8_b200(C.cast(C.c_void_p(s16.ctypes.data >> 1), u8p), stride) == \
        O.av1o_avg(C.c_void_p(s16.ctypes.data), stride, 8, 1)
    # integer projections, vector variance
    ref = rng.integers(0, 256, (64, 80)).astype(np.uint8)
    for (w, h, norm) in ((64, 64, 5), (16, 32, 4), (32, 16, 3)):
        a, b = np.zeros(w, np.int16), np.zeros(w, np.int16)
        L.aom_int_pro_row_b200(ptr(a, i16p), ptr(ref, u8p), 80, w, h, norm)
        O.av1o_int_pro_row(ptr(b, i16p), ptr(ref, u8p), 80, w, h, norm)
        assert np.array_equal(a, b)
        a, b = np.zeros(h, np.int16), np.zeros(h, np.int16)
        L.aom_int_pro_col_b200(ptr(a, i16p), ptr(ref, u8p), 80, w, h, norm)
        O.av1o_int_pro_col(ptr(b, i16p), ptr(ref, u8p), 80, w, h, norm)
        assert np.array_equal(a, b)
    for bwl in (2, 3, 4, 5):
        x = rng.integers(0, 511, 4 << bwl).astype(np.int16)
        y = rng.integers(0, 511, 4 << bwl).astype(np.int16)
        assert L.aom_vector_var_b200(ptr(x, i16p), ptr(y, i16p), bwl) == \
            O.av1o_vector_var(ptr(x, i16p), ptr(y, i16p), bwl)
