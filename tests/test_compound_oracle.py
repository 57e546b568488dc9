"""Pin oracle/av1o_compound.c against the reference's own per-size C functions (oracle/_ref): masked SAD,
masked sub-pel variance, compound-average and distance-weighted sub-pel variance, distance-weighted SAD,
OBMC SAD / variance / sub-pel variance (8 bit and high bit depth 8/10/12) and the compound predictor
builders, all 22 block sizes.  Inputs follow test/masked_sad_test.cc, masked_variance_test.cc,
obmc_sad_test.cc, obmc_variance_test.cc, comp_avg_pred_test.h (random pixels / masks 0..64, the
extreme-value cases)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import ptr, u8p, u16p, u32p, i32p

pytestmark = [pytest.mark.ref,
              pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]


class Compound(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("highbd", C.c_int), ("bd", C.c_int),
                ("pre", C.c_void_p), ("pre_stride", C.c_int),
                ("subpel", C.c_int), ("xoff", C.c_int), ("yoff", C.c_int),
                ("tgt", C.c_void_p), ("tgt_stride", C.c_int), ("blend", C.c_int),
                ("second_pred", C.c_void_p), ("mask", C.c_void_p), ("mask_stride", C.c_int),
                ("invert_mask", C.c_int), ("fwd_offset", C.c_int), ("bck_offset", C.c_int),
                ("wsrc", C.c_void_p), ("obmc_mask", C.c_void_p)]


class DistWtd(C.Structure):  # DIST_WTD_COMP_PARAMS, aom_dsp/aom_dsp_common.h
    _fields_ = [("use_dist_wtd_comp_avg", C.c_int), ("fwd_offset", C.c_int), ("bck_offset", C.c_int)]


NONE, AVG, DIST, MASK = range(4)
# quant_dist_lookup_table (av1/common/reconinter.c): the (fwd, bck) pairs av1_dist_wtd_comp_weight_assign yields
DIST_PAIRS = [(9, 7), (11, 5), (12, 4), (13, 3), (7, 9), (5, 11), (4, 12), (3, 13), (8, 8)]


def bp(a, highbd):
    """pixel pointer as the reference takes it: CONVERT_TO_BYTEPTR for uint16 buffers"""
    return C.cast(C.c_void_p(a.ctypes.data >> 1), u8p) if highbd else ptr(a, u8p)


def make(rng, w, h, highbd, bd, extreme=0):
    dt = np.uint16 if highbd else np.uint8
    hi = (1 << bd) - 1
    stride = w + 9
    def plane(rows, cols):
        if extreme == 1:
            return np.full((rows, cols), hi, dt)
        if extreme == 2:
            return np.zeros((rows, cols), dt)
        return rng.integers(0, hi + 1, (rows, cols)).astype(dt)
    pre = plane(h + 2, stride + 1)
    tgt = plane(h, stride) if extreme != 1 else np.zeros((h, stride), dt)
    second = plane(h, w)
    mask = rng.integers(0, 65, (h, w + 5)).astype(np.uint8)
    if extreme:
        mask[:] = 64 if extreme == 1 else 0
    return pre, tgt, second, mask, stride


def cstruct(w, h, highbd, bd, pre, pre_stride, tgt, tgt_stride, blend=NONE, second=None, mask=None,
            mask_stride=0, invert=0, subpel=0, xoff=0, yoff=0, fwd=0, bck=0, wsrc=None, omask=None):
    return Compound(w, h, highbd, bd, pre.ctypes.data, pre_stride, subpel, xoff, yoff,
                    tgt.ctypes.data if tgt is not None else None, tgt_stride, blend,
                    second.ctypes.data if second is not None else None,
                    mask.ctypes.data if mask is not None else None, mask_stride, invert, fwd, bck,
                    wsrc.ctypes.data if wsrc is not None else None,
                    omask.ctypes.data if omask is not None else None)


def hb_name(base, bd, highbd):
    """aom_X -> aom_highbd_X (SAD family) or aom_highbd_{8,10,12}_X (variance family)"""
    return base


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
def test_masked_sad_and_variance(highbd, bd):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(9100 + bd + highbd)
    for (w, h) in H.BLOCK_SIZES:
        for it in range(4):
            pre, tgt, second, mask, stride = make(rng, w, h, highbd, bd, extreme=it if it < 3 else 0)
            ms = mask.shape[1]
            inv = it & 1
            sad_fn = getattr(R, ("aom_highbd_masked_sad%dx%d_c" if highbd else "aom_masked_sad%dx%d_c") % (w, h))
            want = sad_fn(bp(tgt, highbd), stride, bp(pre, highbd), pre.shape[1], bp(second, highbd),
                          ptr(mask, u8p), ms, inv)
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], tgt, stride, MASK, second, mask, ms, inv)
            assert O.av1o_compound_sad(C.byref(c)) == want & 0xffffffff, (w, h, it)
            xo, yo = int(rng.integers(0, 8)), int(rng.integers(0, 8))
            vname = ("aom_highbd_%d_masked_sub_pixel_variance%dx%d_c" % (bd, w, h)) if highbd else \
                "aom_masked_sub_pixel_variance%dx%d_c" % (w, h)
            sse_a, sse_b = C.c_uint(0), C.c_uint(0)
            want = getattr(R, vname)(bp(pre, highbd), pre.shape[1], xo, yo, bp(tgt, highbd), stride,
                                     bp(second, highbd), ptr(mask, u8p), ms, inv, C.byref(sse_a))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], tgt, stride, MASK, second, mask, ms, inv, 1, xo, yo)
            got = O.av1o_compound_variance(C.byref(c), C.byref(sse_b))
            assert (got & 0xffffffff, sse_b.value) == (want & 0xffffffff, sse_a.value), (w, h, it, xo, yo)


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
def test_avg_and_dist_wtd(highbd, bd):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(9200 + bd + highbd)
    for (w, h) in H.BLOCK_SIZES:
        for it in range(4):
            pre, tgt, second, mask, stride = make(rng, w, h, highbd, bd, extreme=it if it < 3 else 0)
            xo, yo = int(rng.integers(0, 8)), int(rng.integers(0, 8))
            fwd, bck = DIST_PAIRS[int(rng.integers(0, len(DIST_PAIRS)))]
            jcp = DistWtd(1, fwd, bck)
            pfx = ("aom_highbd_%d_" % bd) if highbd else "aom_"
            sse_a, sse_b = C.c_uint(0), C.c_uint(0)
            want = getattr(R, pfx + "sub_pixel_avg_variance%dx%d_c" % (w, h))(
                bp(pre, highbd), pre.shape[1], xo, yo, bp(tgt, highbd), stride, C.byref(sse_a), bp(second, highbd))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], tgt, stride, AVG, second, subpel=1, xoff=xo, yoff=yo)
            got = O.av1o_compound_variance(C.byref(c), C.byref(sse_b))
            assert (got & 0xffffffff, sse_b.value) == (want & 0xffffffff, sse_a.value), ("avg", w, h, it)
            want = getattr(R, pfx + "dist_wtd_sub_pixel_avg_variance%dx%d_c" % (w, h))(
                bp(pre, highbd), pre.shape[1], xo, yo, bp(tgt, highbd), stride, C.byref(sse_a), bp(second, highbd),
                C.byref(jcp))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], tgt, stride, DIST, second, subpel=1, xoff=xo, yoff=yo,
                        fwd=fwd, bck=bck)
            got = O.av1o_compound_variance(C.byref(c), C.byref(sse_b))
            assert (got & 0xffffffff, sse_b.value) == (want & 0xffffffff, sse_a.value), ("dist", w, h, it)
            sname = ("aom_highbd_dist_wtd_sad%dx%d_avg_c" if highbd else "aom_dist_wtd_sad%dx%d_avg_c") % (w, h)
            want = getattr(R, sname)(bp(tgt, highbd), stride, bp(pre, highbd), pre.shape[1], bp(second, highbd),
                                     C.byref(jcp))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], tgt, stride, DIST, second, fwd=fwd, bck=bck)
            assert O.av1o_compound_sad(C.byref(c)) == want & 0xffffffff, ("dsad", w, h, it)
            # predictor builders
            dt = np.uint16 if highbd else np.uint8
            a, b = np.zeros((h, w), dt), np.zeros((h, w), dt)
            hp = "aom_highbd_" if highbd else "aom_"
            getattr(R, hp + "comp_avg_pred_c")(bp(a, highbd), bp(second, highbd), w, h, bp(pre, highbd), pre.shape[1])
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], None, 0, AVG, second)
            O.av1o_compound_pred(C.byref(c), b.ctypes.data_as(C.c_void_p))
            assert np.array_equal(a, b)
            getattr(R, hp + "dist_wtd_comp_avg_pred_c")(bp(a, highbd), bp(second, highbd), w, h, bp(pre, highbd),
                                                        pre.shape[1], C.byref(jcp))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], None, 0, DIST, second, fwd=fwd, bck=bck)
            O.av1o_compound_pred(C.byref(c), b.ctypes.data_as(C.c_void_p))
            assert np.array_equal(a, b)
            getattr(R, hp + "comp_mask_pred_c")(bp(a, highbd), bp(second, highbd), w, h, bp(pre, highbd),
                                                pre.shape[1], ptr(mask, u8p), mask.shape[1], it & 1)
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], None, 0, MASK, second, mask, mask.shape[1], it & 1)
            O.av1o_compound_pred(C.byref(c), b.ctypes.data_as(C.c_void_p))
            assert np.array_equal(a, b)


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
def test_obmc(highbd, bd):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(9300 + bd + highbd)
    for (w, h) in H.BLOCK_SIZES:
        for it in range(4):
            pre, _, _, _, _ = make(rng, w, h, highbd, bd, extreme=it if it < 3 else 0)
            # test/obmc_sad_test.cc: wsrc = pixel * 4096 range, mask in [0, 4096]
            hi = (1 << bd) - 1
            wsrc = (rng.integers(0, hi + 1, h * w) * rng.integers(0, 4097, h * w)).astype(np.int32)
            omask = rng.integers(0, 4097, h * w).astype(np.int32)
            if it == 1:
                wsrc[:] = hi * 4096
                omask[:] = 0
            elif it == 2:
                wsrc[:] = 0
                omask[:] = 4096
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], None, 0, wsrc=wsrc, omask=omask)
            sname = ("aom_highbd_obmc_sad%dx%d_c" if highbd else "aom_obmc_sad%dx%d_c") % (w, h)
            want = getattr(R, sname)(bp(pre, highbd), pre.shape[1], ptr(wsrc, i32p), ptr(omask, i32p))
            assert O.av1o_obmc_sad(C.byref(c)) == want & 0xffffffff, (w, h, it)
            if highbd:
                vp = "aom_highbd_obmc_" if bd == 8 else "aom_highbd_%d_obmc_" % bd
            else:
                vp = "aom_obmc_"
            sse_a, sse_b = C.c_uint(0), C.c_uint(0)
            want = getattr(R, vp + "variance%dx%d_c" % (w, h))(bp(pre, highbd), pre.shape[1], ptr(wsrc, i32p),
                                                              ptr(omask, i32p), C.byref(sse_a))
            got = O.av1o_obmc_variance(C.byref(c), C.byref(sse_b))
            assert (got & 0xffffffff, sse_b.value) == (want & 0xffffffff, sse_a.value), ("ovar", w, h, it)
            xo, yo = int(rng.integers(0, 8)), int(rng.integers(0, 8))
            want = getattr(R, vp + "sub_pixel_variance%dx%d_c" % (w, h))(
                bp(pre, highbd), pre.shape[1], xo, yo, ptr(wsrc, i32p), ptr(omask, i32p), C.byref(sse_a))
            c = cstruct(w, h, highbd, bd, pre, pre.shape[1], None, 0, wsrc=wsrc, omask=omask, subpel=1, xoff=xo,
                        yoff=yo)
            got = O.av1o_obmc_variance(C.byref(c), C.byref(sse_b))
            assert (got & 0xffffffff, sse_b.value) == (want & 0xffffffff, sse_a.value), ("osvar", w, h, it)
