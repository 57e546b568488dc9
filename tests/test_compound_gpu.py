"""GPU parity of b200_compound_launch (compound / masked / OBMC SAD and variance + the compound
predictor builders) against oracle/av1o_compound.c, bit-exact: every op,
the 22 block sizes in one batch, uint8 and uint16 pixels at 8 / 10 / 12 bit."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from test_compound_oracle import Compound, DIST_PAIRS

pytestmark = pytest.mark.gpu

NONE, AVG, DIST, MASK = range(4)


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def _batch(rng, highbd, bd, reps=3):
    """planes + one descriptor per (size, rep): pre / tgt planes, w-pitch second_pred / wsrc / obmc_mask
    pools, a mask plane"""
    import b200av1
    dt = np.uint16 if highbd else np.uint8
    hi = (1 << bd) - 1
    pw, ph = 3 * 130 + 16, 22 * 129 + 8
    pre = rng.integers(0, hi + 1, (ph, pw)).astype(dt)
    tgt = rng.integers(0, hi + 1, (ph, pw + 6)).astype(dt)
    maskp = rng.integers(0, 65, (ph, pw + 3)).astype(np.uint8)
    pre[:129] = hi      # extreme rows
    tgt[:129] = 0
    maskp[:129] = 64
    blocks, masks, pool = [], [], 0
    for si, (w, h) in enumerate(H.BLOCK_SIZES):
        for r in range(reps):
            x0, y0 = r * 130 + int(rng.integers(0, 3)), si * 129 + int(rng.integers(0, 2))
            blocks.append((y0 * pw + x0, y0 * (pw + 6) + x0, pool, pw, pw + 6, w, h,
                           int(rng.integers(0, 8)), int(rng.integers(0, 8))))
            masks.append((y0 * (pw + 3) + x0, pw + 3, (si + r) & 1, 0))
            pool += w * h
    blocks = np.array(blocks, b200av1.PIXBLOCK_DTYPE)
    masks = np.array(masks, b200av1.PIXMASK_DTYPE)
    second = rng.integers(0, hi + 1, pool).astype(dt)
    wsrc = (rng.integers(0, hi + 1, pool) * rng.integers(0, 4097, pool)).astype(np.int32)
    omask = rng.integers(0, 4097, pool).astype(np.int32)
    return pre, tgt, maskp, blocks, masks, second, wsrc, omask


def _oracle(op, blocks, masks, pre, tgt, maskp, second, wsrc, omask, highbd, bd, fwd, bck):
    import b200av1
    O = H.oracle()
    isz = pre.itemsize
    traits = {b200av1.CMP_MASKED_SAD: (0, MASK, "sad"), b200av1.CMP_MASKED_SUBPEL_VAR: (1, MASK, "var"),
              b200av1.CMP_SUBPEL_AVG_VAR: (1, AVG, "var"), b200av1.CMP_DIST_WTD_SUBPEL_AVG_VAR: (1, DIST, "var"),
              b200av1.CMP_DIST_WTD_SAD_AVG: (0, DIST, "sad"), b200av1.CMP_OBMC_SAD: (0, NONE, "osad"),
              b200av1.CMP_OBMC_VAR: (0, NONE, "ovar"), b200av1.CMP_OBMC_SUBPEL_VAR: (1, NONE, "ovar"),
              b200av1.CMP_AVG_PRED: (0, AVG, "pred"), b200av1.CMP_DIST_WTD_AVG_PRED: (0, DIST, "pred"),
              b200av1.CMP_MASK_PRED: (0, MASK, "pred")}[op]
    out = np.zeros((len(blocks), 2), np.uint32)
    comp = np.zeros(second.size, second.dtype)
    for i, (b, m) in enumerate(zip(blocks, masks)):
        w, h, aux = int(b["w"]), int(b["h"]), int(b["aux_offset"])
        c = Compound(w, h, highbd, bd, pre.ctypes.data + int(b["a_offset"]) * isz, int(b["a_stride"]),
                     traits[0], int(b["xoff"]), int(b["yoff"]),
                     tgt.ctypes.data + int(b["b_offset"]) * isz, int(b["b_stride"]), traits[1],
                     second.ctypes.data + aux * isz, maskp.ctypes.data + int(m["mask_offset"]),
                     int(m["mask_stride"]), int(m["invert_mask"]), fwd, bck,
                     wsrc.ctypes.data + aux * 4, omask.ctypes.data + aux * 4)
        sse = C.c_uint(0)
        if traits[2] == "sad":
            out[i, 0] = O.av1o_compound_sad(C.byref(c)) & 0xffffffff
        elif traits[2] == "var":
            out[i, 0] = O.av1o_compound_variance(C.byref(c), C.byref(sse)) & 0xffffffff
            out[i, 1] = sse.value
        elif traits[2] == "osad":
            out[i, 0] = O.av1o_obmc_sad(C.byref(c)) & 0xffffffff
        elif traits[2] == "ovar":
            out[i, 0] = O.av1o_obmc_variance(C.byref(c), C.byref(sse)) & 0xffffffff
            out[i, 1] = sse.value
        else:
            O.av1o_compound_pred(C.byref(c), C.c_void_p(comp.ctypes.data + aux * isz))
    return out, comp


@pytest.mark.parametrize("highbd,bd", [(0, 8), (1, 8), (1, 10), (1, 12)])
@pytest.mark.parametrize("op", range(11))
def test_compound_launch_vs_oracle(ctx, op, highbd, bd):
    import b200av1
    rng = np.random.default_rng(9500 + op * 8 + bd + highbd)
    pre, tgt, maskp, blocks, masks, second, wsrc, omask = _batch(rng, highbd, bd)
    fwd, bck = DIST_PAIRS[op % len(DIST_PAIRS)]
    builder = op >= b200av1.CMP_AVG_PRED
    if builder:  # builders write comp_pred + b_offset: point b_offset at the block's pool slot
        blocks = blocks.copy()
        blocks["b_offset"] = blocks["aux_offset"]
    want, want_pred = _oracle(op, blocks, masks, pre.reshape(-1), tgt.reshape(-1), maskp.reshape(-1), second,
                              wsrc, omask, highbd, bd, fwd, bck)
    d = {k: b200av1.to_device(ctx, v) for k, v in dict(blocks=blocks, masks=masks, pre=pre, tgt=tgt, mask=maskp,
                                                       second=second, wsrc=wsrc, omask=omask).items()}
    d_out = b200av1.to_device(ctx, np.full((len(blocks), 2), 0xdeadbeef, np.uint32))
    d_comp = b200av1.to_device(ctx, np.zeros_like(second))
    l = b200av1.CompoundLaunch(op, highbd, bd, len(blocks), d["blocks"].ptr, d["masks"].ptr, d["pre"].ptr,
                               d["tgt"].ptr, d["second"].ptr, d["mask"].ptr, d["wsrc"].ptr, d["omask"].ptr,
                               fwd, bck, d_out.ptr, d_comp.ptr)
    b200av1.check(b200av1.lib().b200_compound_launch(ctx.h, C.byref(l), None))
    ctx.sync()
    if builder:
        got = b200av1.to_host(ctx, d_comp, second.dtype, second.size)
        assert np.array_equal(got, want_pred)
    else:
        got = b200av1.to_host(ctx, d_out, np.uint32, len(blocks) * 2).reshape(-1, 2)
        bad = np.nonzero((got != want).any(axis=1))[0]
        assert bad.size == 0, (op, bad[:5], got[bad[:5]], want[bad[:5]], blocks[bad[:5]])
