"""TPL block queue on the GPU (a25 / a26 / f1): motion_estimation (tpl_model.c:249-304) for every block
of a frame in two chained launches, and txfm_quant_rdcost (:228-247) for every block of a frame in one
fused launch, against the reference's own functions called per block (oracle/ref_shim.c)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from test_queue_tf_gpu import frame_pair

pytestmark = pytest.mark.gpu


def mv_cost_tables(seed):
    """monotone synthetic nmv cost tables like the ones the encoder derives from its CDFs"""
    rng = np.random.default_rng(seed)
    idx = np.arange(-16383, 16384)
    t0 = (np.log2(1 + np.abs(idx)) * 180 + 300 + rng.integers(0, 40, idx.size)).astype(np.int32)
    t1 = (np.log2(1 + np.abs(idx)) * 170 + 280 + rng.integers(0, 40, idx.size)).astype(np.int32)
    joint = np.array([200, 700, 650, 1100], np.int32)
    return joint, np.ascontiguousarray(t0), np.ascontiguousarray(t1)


@pytest.mark.parametrize("width,height,bd,cost_list,centred", [(352, 288, 8, 0, False), (176, 144, 10, 1, True),
                                                             (200, 120, 8, 1, True)])
def test_tpl_motion_estimation_frame_equals_reference_loop(width, height, bd, cost_list, centred):
    import b200av1
    R = H.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    ctx = b200av1.Context(0)
    cur, ref = frame_pair(width, height, bd, 5 + height)
    p = b200av1.tpl_default_params(width, height)
    p.want_cost_list = cost_list
    p.subpel_tree = 1 if cost_list else 0
    joint, t0, t1 = mv_cost_tables(9)
    p.mvjcost = joint.ctypes.data
    p.mvcost[0] = t0.ctypes.data + 16383 * 4
    p.mvcost[1] = t1.ctypes.data + 16383 * 4
    fc, fr = b200av1.Frame(ctx, cur, p.border_in_pixels, bd), b200av1.Frame(ctx, ref, p.border_in_pixels, bd)
    rows, cols = b200av1.tpl_grid(width, height, p)
    mvs = None
    if centred:
        mvs = np.random.default_rng(4).integers(-60, 61, (rows * cols, 2)).astype(np.int16)
    got = b200av1.queue_tpl_motion_estimation(ctx, fc, fr, p, mvs)
    cb, rb = fc.download(), fr.download()
    off = (p.border_in_pixels * cb.shape[1] + p.border_in_pixels) * cb.dtype.itemsize
    want = np.zeros(rows * cols, b200av1.TPL_ME_DTYPE)
    R.refshim_tpl_motion_estimation_frame(C.c_void_p(cb.ctypes.data + off), C.c_void_p(rb.ctypes.data + off),
                                          cb.shape[1], width, height, int(bd > 8), bd, C.byref(p),
                                          C.c_void_p(mvs.ctypes.data) if mvs is not None else None,
                                          C.c_void_p(want.ctypes.data))
    want = want.reshape(rows, cols)
    assert np.array_equal(got["mv"], want["mv"]), np.argwhere(np.any(got["mv"] != want["mv"], axis=-1))[:5]
    assert np.array_equal(got["sme"], want["sme"])
    assert np.count_nonzero(np.any(got["mv"] != 0, axis=-1)) > rows * cols // 2
    fc.free()
    fr.free()


@pytest.mark.parametrize("bd,tx_size,qindex", [(8, 2, 60), (8, 3, 140), (10, 2, 100), (12, 1, 200), (8, 0, 20)])
def test_tpl_txfm_quant_rdcost_equals_reference(bd, tx_size, qindex):
    import b200av1
    R = H.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    ctx = b200av1.Context(0)
    w = H.TX_W[tx_size]
    rows, cols = 9, 13
    cur, pred = frame_pair(cols * w + 8, rows * w + 4, bd, 31 + tx_size)
    qrow = b200av1.qrow_from_oracle(H.oracle_qrow(bd, qindex))
    got = b200av1.queue_tpl_txfm_quant_rdcost(ctx, cur, pred, rows, cols, bd, tx_size, qrow)
    want = np.zeros(rows * cols, b200av1.TPL_RD_DTYPE)
    q = np.ascontiguousarray(qrow)
    R.refshim_tpl_txfm_quant_rdcost(C.c_void_p(cur.ctypes.data), cur.shape[1], C.c_void_p(pred.ctypes.data),
                                    pred.shape[1], rows, cols, bd, tx_size, C.c_void_p(q.ctypes.data),
                                    C.c_void_p(want.ctypes.data))
    want = want.reshape(rows, cols)
    for name in want.dtype.names:
        assert np.array_equal(got[name], want[name]), name
    assert got["eob"].max() > 0
