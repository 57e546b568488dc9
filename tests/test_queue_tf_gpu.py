"""Temporal-filter block queue on the GPU (a25 / f1): ONE b200_queue_tf_motion_search call per frame
pair -- four chained launches -- against the reference's per-block loop (av1_tf_do_filtering_row,
temporal_filter.c:804-872, around tf_motion_search :88-260) run with the reference's own
av1_full_pixel_search / av1_find_best_sub_pixel_tree (oracle/ref_shim.c)."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu


def frame_pair(width, height, bd, seed):
    """textured frame + the same content displaced by a smooth per-region motion plus noise"""
    rng = np.random.default_rng(seed)
    pad = 48
    yy, xx = np.mgrid[0:height + 2 * pad, 0:width + 2 * pad]
    base = 128 + 50 * np.sin(xx / 7.0) * np.cos(yy / 11.0) + 35 * np.sin((xx + 3 * yy) / 23.0) + rng.normal(0, 10, xx.shape)
    cur = base[pad:pad + height, pad:pad + width]
    ref = np.empty_like(cur)
    for by in range(0, height, 64):
        for bx in range(0, width, 64):
            dy, dx = rng.integers(-9, 10, 2)
            h, w = min(64, height - by), min(64, width - bx)
            ref[by:by + h, bx:bx + w] = base[pad + by + dy:pad + by + dy + h, pad + bx + dx:pad + bx + dx + w]
    ref = ref + rng.normal(0, 2.0, ref.shape)
    scale = 1 << (bd - 8)
    dt = np.uint8 if bd == 8 else np.uint16
    q = lambda a: np.clip(np.rint(a * scale), 0, (1 << bd) - 1).astype(dt)
    return q(cur), q(ref)


def reference_tf(cur_b, ref_b, width, height, border, bd, params, ref_mvs):
    import b200av1
    R = H.ref()
    rows, cols = (height + params.bh - 1) // params.bh, (width + params.bw - 1) // params.bw
    out = np.zeros(rows * cols, b200av1.TF_RESULT_DTYPE)
    stride = cur_b.shape[1]
    px = cur_b.dtype.itemsize
    off = (border * stride + border) * px
    mv = None if ref_mvs is None else np.ascontiguousarray(ref_mvs, np.int16)
    R.refshim_tf_motion_search_frame(C.c_void_p(cur_b.ctypes.data + off), C.c_void_p(ref_b.ctypes.data + off), stride,
                                     width, height, int(bd > 8), bd, C.byref(params),
                                     C.c_void_p(mv.ctypes.data) if mv is not None else None,
                                     C.c_void_p(out.ctypes.data))
    return out.reshape(rows, cols)


def test_frame_upload_replicates_the_border():
    import b200av1
    ctx = b200av1.Context(0)
    rng = np.random.default_rng(1)
    for bd, shape, border in ((8, (37, 53), 32), (10, (64, 48), 160)):
        y = rng.integers(0, 1 << bd, shape).astype(np.uint8 if bd == 8 else np.uint16)
        f = b200av1.Frame(ctx, y, border, bd)
        assert np.array_equal(f.download(), np.pad(y, border, mode="edge"))
        f.free()


@pytest.mark.parametrize("width,height,bd,tree,chain", [(352, 288, 8, 0, False), (176, 144, 10, 0, True),
                                                       (200, 120, 8, 1, True), (160, 96, 12, 2, False)])
def test_tf_motion_search_frame_pair_equals_reference_loop(width, height, bd, tree, chain):
    import b200av1
    if H.ref() is None:
        pytest.skip("oracle/_ref not built")
    ctx = b200av1.Context(0)
    cur, ref = frame_pair(width, height, bd, 11 + width)
    p = b200av1.tf_default_params(width, height)
    p.subpel_tree, p.want_cost_list = tree, int(tree > 0)
    fc, fr = b200av1.Frame(ctx, cur, p.border_in_pixels, bd), b200av1.Frame(ctx, ref, p.border_in_pixels, bd)
    rows, cols = (height + 15) // 16, (width + 15) // 16
    mvs = None
    if chain:  # a ref_mv handed down from the previous frame of the window
        mvs = np.random.default_rng(3).integers(-40, 41, (rows * cols, 2)).astype(np.int16)
    got = b200av1.queue_tf_motion_search(ctx, fc, fr, p, mvs)
    want = reference_tf(fc.download(), fr.download(), width, height, p.border_in_pixels, bd, p, mvs)
    for name in b200av1.TF_RESULT_DTYPE.names:
        bad = np.argwhere(~np.all((got[name] == want[name]).reshape(rows, cols, -1), axis=2))
        assert bad.size == 0, "%s differs at blocks %s: got %s want %s" % (
            name, bad[:4].tolist(), got[name][tuple(bad[0])], want[name][tuple(bad[0])])
    # the motion is real: most blocks found a non-zero vector, and some blocks split
    assert np.count_nonzero(np.any(got["block_mv"] != 0, axis=-1)) > rows * cols // 2
    fc.free()
    fr.free()
