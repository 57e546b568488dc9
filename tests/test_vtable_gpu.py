"""b200_variance_fn_ptr: the aom_variance_fn_ptr_t of a block size with all 16 members bound to `_b200`
drop-ins.  Every member is called through the table, with the reference's
pointer conventions (CONVERT_TO_BYTEPTR for high bit depth), and compared with the oracle."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import ptr, u8p, i32p
from test_compound_oracle import Compound, DistWtd

pytestmark = pytest.mark.gpu

P8, U, I = C.POINTER(C.c_uint8), C.c_uint, C.c_int
PU, PI32 = C.POINTER(C.c_uint), C.POINTER(C.c_int32)
PP8 = C.POINTER(P8)
F = C.CFUNCTYPE


class VarianceFnPtr(C.Structure):
    _fields_ = [("sdf", F(U, P8, I, P8, I)), ("sdsf", F(U, P8, I, P8, I)), ("sdaf", F(U, P8, I, P8, I, P8)),
                ("vf", F(U, P8, I, P8, I, PU)), ("svf", F(U, P8, I, I, I, P8, I, PU)),
                ("svaf", F(U, P8, I, I, I, P8, I, PU, P8)),
                ("sdx4df", F(None, P8, I, PP8, I, PU)), ("sdx3df", F(None, P8, I, PP8, I, PU)),
                ("sdsx4df", F(None, P8, I, PP8, I, PU)),
                ("msdf", F(U, P8, I, P8, I, P8, P8, I, I)), ("msvf", F(U, P8, I, I, I, P8, I, P8, P8, I, I, PU)),
                ("osdf", F(U, P8, I, PI32, PI32)), ("ovf", F(U, P8, I, PI32, PI32, PU)),
                ("osvf", F(U, P8, I, I, I, PI32, PI32, PU)),
                ("jsdaf", F(U, P8, I, P8, I, P8, C.POINTER(DistWtd))),
                ("jsvaf", F(U, P8, I, I, I, P8, I, PU, P8, C.POINTER(DistWtd)))]


def bp(a, highbd, off=0):
    addr = a.ctypes.data + off * a.itemsize
    return C.cast(C.c_void_p(addr >> 1 if highbd else addr), P8)


def cs(w, h, hb, bd, pre, ps, tgt, ts, blend=0, second=None, mask=None, ms=0, inv=0, subpel=0, xo=0, yo=0,
       fwd=0, bck=0, wsrc=None, omask=None, pre_off=0):
    return Compound(w, h, hb, bd, pre.ctypes.data + pre_off * pre.itemsize, ps, subpel, xo, yo,
                    tgt.ctypes.data if tgt is not None else None, ts, blend,
                    second.ctypes.data if second is not None else None,
                    mask.ctypes.data if mask is not None else None, ms, inv, fwd, bck,
                    wsrc.ctypes.data if wsrc is not None else None,
                    omask.ctypes.data if omask is not None else None)


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
@pytest.mark.parametrize("w,h", [(8, 8), (16, 32), (64, 64), (4, 16), (128, 64)])
def test_variance_vtable_members(w, h, highbd, bd):
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    vt = VarianceFnPtr()
    assert L.b200_variance_fn_ptr(w, h, highbd, bd, C.byref(vt)) == 0
    rng = np.random.default_rng(9700 + w * 3 + h + bd + highbd)
    dt = np.uint16 if highbd else np.uint8
    hi = (1 << bd) - 1
    ss, rs = w + 8, w + 24
    src = rng.integers(0, hi + 1, (h, ss)).astype(dt)
    refs = rng.integers(0, hi + 1, (h + 3, rs + 8)).astype(dt)
    second = rng.integers(0, hi + 1, (h, w)).astype(dt)
    mask = rng.integers(0, 65, (h, w + 4)).astype(np.uint8)
    wsrc = (rng.integers(0, hi + 1, h * w) * rng.integers(0, 4097, h * w)).astype(np.int32)
    omask = rng.integers(0, 4097, h * w).astype(np.int32)
    rstride = refs.shape[1]
    xo, yo = 3, 6
    jcp = DistWtd(1, 11, 5)
    sse_g, sse_w = C.c_uint(0), C.c_uint(0)
    c0 = cs(w, h, highbd, bd, refs, rstride, src, ss)
    assert vt.sdf(bp(src, highbd), ss, bp(refs, highbd), rstride) == O.av1o_compound_sad(C.byref(c0)) & 0xffffffff
    # sdsf: every other row, doubled (sad.c:65)
    ch = cs(w, h // 2, highbd, bd, refs, 2 * rstride, src, 2 * ss)
    assert vt.sdsf(bp(src, highbd), ss, bp(refs, highbd), rstride) == 2 * O.av1o_compound_sad(C.byref(ch))
    ca = cs(w, h, highbd, bd, refs, rstride, src, ss, 1, second)
    assert vt.sdaf(bp(src, highbd), ss, bp(refs, highbd), rstride, bp(second, highbd)) == O.av1o_compound_sad(C.byref(ca))
    got = vt.vf(bp(refs, highbd), rstride, bp(src, highbd), ss, C.byref(sse_g))
    assert (got, sse_g.value) == (O.av1o_compound_variance(C.byref(c0), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    csub = cs(w, h, highbd, bd, refs, rstride, src, ss, subpel=1, xo=xo, yo=yo)
    got = vt.svf(bp(refs, highbd), rstride, xo, yo, bp(src, highbd), ss, C.byref(sse_g))
    assert (got, sse_g.value) == (O.av1o_compound_variance(C.byref(csub), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    cav = cs(w, h, highbd, bd, refs, rstride, src, ss, 1, second, subpel=1, xo=xo, yo=yo)
    got = vt.svaf(bp(refs, highbd), rstride, xo, yo, bp(src, highbd), ss, C.byref(sse_g), bp(second, highbd))
    assert (got, sse_g.value) == (O.av1o_compound_variance(C.byref(cav), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    # x4d / x3d / skip x4d: four candidate positions inside refs
    offs = [0, 1, rstride + 2, 2 * rstride + 5]
    arr = (P8 * 4)(*[bp(refs, highbd, o) for o in offs])
    for member, skip in ((vt.sdx4df, 0), (vt.sdx3df, 0), (vt.sdsx4df, 1)):
        out = (C.c_uint * 4)(7, 7, 7, 7)
        member(bp(src, highbd), ss, arr, rstride, out)
        n = 3 if member is vt.sdx3df else 4
        for k in range(n):
            if skip:
                ck = cs(w, h // 2, highbd, bd, refs, 2 * rstride, src, 2 * ss, pre_off=offs[k])
                want = 2 * O.av1o_compound_sad(C.byref(ck))
            else:
                ck = cs(w, h, highbd, bd, refs, rstride, src, ss, pre_off=offs[k])
                want = O.av1o_compound_sad(C.byref(ck))
            assert out[k] == want, (k, skip)
    for inv in (0, 1):
        cm = cs(w, h, highbd, bd, refs, rstride, src, ss, 3, second, mask, mask.shape[1], inv)
        got = vt.msdf(bp(src, highbd), ss, bp(refs, highbd), rstride, bp(second, highbd), ptr(mask, u8p),
                      mask.shape[1], inv)
        assert got == O.av1o_compound_sad(C.byref(cm)) & 0xffffffff
        cm = cs(w, h, highbd, bd, refs, rstride, src, ss, 3, second, mask, mask.shape[1], inv, 1, xo, yo)
        got = vt.msvf(bp(refs, highbd), rstride, xo, yo, bp(src, highbd), ss, bp(second, highbd), ptr(mask, u8p),
                      mask.shape[1], inv, C.byref(sse_g))
        assert (got, sse_g.value) == (O.av1o_compound_variance(C.byref(cm), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    co = cs(w, h, highbd, bd, refs, rstride, None, 0, wsrc=wsrc, omask=omask)
    assert vt.osdf(bp(refs, highbd), rstride, ptr(wsrc, i32p), ptr(omask, i32p)) == O.av1o_obmc_sad(C.byref(co)) & 0xffffffff
    got = vt.ovf(bp(refs, highbd), rstride, ptr(wsrc, i32p), ptr(omask, i32p), C.byref(sse_g))
    assert (got, sse_g.value) == (O.av1o_obmc_variance(C.byref(co), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    co = cs(w, h, highbd, bd, refs, rstride, None, 0, wsrc=wsrc, omask=omask, subpel=1, xo=xo, yo=yo)
    got = vt.osvf(bp(refs, highbd), rstride, xo, yo, ptr(wsrc, i32p), ptr(omask, i32p), C.byref(sse_g))
    assert (got, sse_g.value) == (O.av1o_obmc_variance(C.byref(co), C.byref(sse_w)) & 0xffffffff, sse_w.value)
    cj = cs(w, h, highbd, bd, refs, rstride, src, ss, 2, second, fwd=11, bck=5)
    got = vt.jsdaf(bp(src, highbd), ss, bp(refs, highbd), rstride, bp(second, highbd), C.byref(jcp))
    assert got == O.av1o_compound_sad(C.byref(cj)) & 0xffffffff
    cj = cs(w, h, highbd, bd, refs, rstride, src, ss, 2, second, subpel=1, xo=xo, yo=yo, fwd=11, bck=5)
    got = vt.jsvaf(bp(refs, highbd), rstride, xo, yo, bp(src, highbd), ss, C.byref(sse_g), bp(second, highbd),
                   C.byref(jcp))
    assert (got, sse_g.value) == (O.av1o_compound_variance(C.byref(cj), C.byref(sse_w)) & 0xffffffff, sse_w.value)


def test_misc_dropins(ctx=None):
    """aom_get8x8var / 16x16var, av1_round_shift_array, the comp-pred builders"""
    import b200av1
    L, O = b200av1.lib(), H.oracle()
    rng = np.random.default_rng(9800)
    for n in (8, 16):
        a = rng.integers(0, 256, (n, n + 5)).astype(np.uint8)
        b = rng.integers(0, 256, (n, n + 9)).astype(np.uint8)
        sse, s = C.c_uint(0), C.c_int(0)
        getattr(L, "aom_get%dx%dvar_b200" % (n, n))(ptr(a, u8p), n + 5, ptr(b, u8p), n + 9, C.byref(sse), C.byref(s))
        d = a[:, :n].astype(np.int64) - b[:, :n]
        assert (sse.value, s.value) == (int((d * d).sum()), int(d.sum()))
    for bit in (3, -2, 1):
        arr = rng.integers(-(1 << 30), 1 << 30, 777).astype(np.int32)
        want = arr.copy()
        O.av1o_round_shift_array(ptr(want, i32p), 777, bit)
        L.av1_round_shift_array_b200(ptr(arr, i32p), 777, bit)
        assert np.array_equal(arr, want)
    for hb in (0, 1):
        dt = np.uint16 if hb else np.uint8
        w, h = 16, 8
        pred = rng.integers(0, 1024 if hb else 256, (h, w)).astype(dt)
        ref = rng.integers(0, 1024 if hb else 256, (h, w + 6)).astype(dt)
        mask = rng.integers(0, 65, (h, w + 2)).astype(np.uint8)
        jcp = DistWtd(1, 9, 7)
        pfx = "aom_highbd_" if hb else "aom_"
        for name, blend, extra in (("comp_avg_pred_b200", 1, ()), ("dist_wtd_comp_avg_pred_b200", 2, (C.byref(jcp),)),
                                   ("comp_mask_pred_b200", 3, (ptr(mask, u8p), w + 2, 1))):
            got, want = np.zeros((h, w), dt), np.zeros((h, w), dt)
            getattr(L, pfx + name)(bp(got, hb), bp(pred, hb), w, h, bp(ref, hb), w + 6, *extra)
            c = cs(w, h, hb, 10 if hb else 8, ref, w + 6, None, 0, blend, pred, mask, w + 2, 1, fwd=9, bck=7)
            O.av1o_compound_pred(C.byref(c), want.ctypes.data_as(C.c_void_p))
            assert np.array_equal(got, want), (name, hb)
