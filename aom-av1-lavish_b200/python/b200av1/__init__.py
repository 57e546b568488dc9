"""ctypes view of libb200av1.so (the C ABI declared in include/b200av1.h).

This is harness plumbing for tests and bench.py -- the product is the shared library
itself.  Loading fails loudly when the library has not been built; calls fail loudly
when there is no sm_100 GPU (the library has no CPU fallback)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.dirname(os.path.dirname(_HERE))
LIB_PATH = os.environ.get("B200AV1_LIB", os.path.join(PKG_DIR, "libb200av1.so"))

OK = 0
IN_RESIDUAL, IN_PIXELS8, IN_PIXELS16 = 0, 1, 2
Q_NONE, Q_FP, Q_FP_HBD, Q_B, Q_B_HBD = 0, 1, 2, 3, 4
TXQ_STORE_COEFF, TXQ_BLOCK_ERROR, TXQ_FORCE_EXACT = 1, 2, 4


class QRow(C.Structure):
    _fields_ = [("zbin", C.c_int16 * 2), ("round", C.c_int16 * 2), ("quant", C.c_int16 * 2),
                ("quant_shift", C.c_int16 * 2), ("round_fp", C.c_int16 * 2),
                ("quant_fp", C.c_int16 * 2), ("dequant", C.c_int16 * 2),
                ("reserved", C.c_int16 * 2)]


TXBLOCK_DTYPE = np.dtype([("src_offset", "<u4"), ("out_offset", "<u4"), ("src_stride", "<u2"),
                          ("qrow", "<u2"), ("tx_size", "u1"), ("tx_type", "u1"),
                          ("reserved", "<u2")])
TXPRED_DTYPE = np.dtype([("pred_offset", "<u4"), ("pred_stride", "<u2"), ("reserved", "<u2")])
QROW_DTYPE = np.dtype([("zbin", "<i2", 2), ("round", "<i2", 2), ("quant", "<i2", 2),
                       ("quant_shift", "<i2", 2), ("round_fp", "<i2", 2), ("quant_fp", "<i2", 2),
                       ("dequant", "<i2", 2), ("reserved", "<i2", 2)])
assert TXBLOCK_DTYPE.itemsize == 16 and TXPRED_DTYPE.itemsize == 8 and QROW_DTYPE.itemsize == 32


class TxqLaunch(C.Structure):
    _fields_ = [("tx_size", C.c_int), ("in_kind", C.c_int), ("q_kind", C.c_int), ("bd", C.c_int),
                ("flags", C.c_uint), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("preds", C.c_void_p), ("src", C.c_void_p), ("pred", C.c_void_p),
                ("qrows", C.c_void_p), ("coeff", C.c_void_p), ("qcoeff", C.c_void_p),
                ("dqcoeff", C.c_void_p), ("eob", C.c_void_p), ("dist", C.c_void_p)]


class ItxLaunch(C.Structure):
    _fields_ = [("tx_size", C.c_int), ("highbd", C.c_int), ("bd", C.c_int), ("n_blocks", C.c_int),
                ("blocks", C.c_void_p), ("preds", C.c_void_p), ("dqcoeff", C.c_void_p),
                ("eob", C.c_void_p), ("src", C.c_void_p), ("pred", C.c_void_p),
                ("recon", C.c_void_p), ("dist", C.c_void_p), ("lossless", C.c_int)]


TXSEARCH_DTYPE = np.dtype([("dist_px", "<i8"), ("dist_tx", "<i8"), ("ssz", "<i8"), ("rate_est", "<i4"),
                           ("eob", "<u2"), ("reserved", "<u2")])
assert TXSEARCH_DTYPE.itemsize == 32
TXS_DIST_PX, TXS_DIST_TX, TXS_RATE_EST = 1, 2, 4


def txsearch_batch_host(ctx, in_kind, q_kind, bd, want, blocks, preds, src, pred, qrows):
    """b200_txsearch_batch_host: one TXSEARCH_DTYPE record per descriptor; returns (results, d2h_bytes)"""
    out = np.zeros(len(blocks), TXSEARCH_DTYPE)
    d2h = C.c_size_t(0)
    check(lib().b200_txsearch_batch_host(ctx.h, in_kind, q_kind, bd, C.c_uint(want), vp(blocks), vp(preds),
                                         len(blocks), vp(src), C.c_size_t(src.size), vp(pred),
                                         C.c_size_t(pred.size), vp(qrows), len(qrows), vp(out), C.byref(d2h)))
    return out, d2h.value


class CompactLaunch(C.Structure):
    _fields_ = [("n_blocks", C.c_int), ("blocks", C.c_void_p), ("qcoeff", C.c_void_p),
                ("eob", C.c_void_p), ("wide", C.c_int), ("levels", C.c_void_p),
                ("level_offset", C.c_void_p)]


def expand_compact(blocks, qrows, levels, level_offset, eob, out_elems, wide=False):
    """host-side expansion of a compact batch into dense qcoeff / dqcoeff (b200_txq_expand_block
    per block)"""
    qc = np.zeros(out_elems, np.int32)
    dq = np.zeros(out_elems, np.int32)
    L = lib()
    isz = levels.itemsize
    for i, b in enumerate(blocks):
        oo = int(b["out_offset"])
        row = qrows[int(b["qrow"]):int(b["qrow"]) + 1]
        check(L.b200_txq_expand_block(int(b["tx_size"]), int(b["tx_type"]),
                                      C.c_void_p(levels.ctypes.data + int(level_offset[i]) * isz),
                                      int(wide), int(eob[i]), vp(row),
                                      C.c_void_p(qc.ctypes.data + oo * 4), C.c_void_p(dq.ctypes.data + oo * 4)))
    return qc, dq


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("libb200av1.so is not built (run __graft_entry__.build() or "
                               "`make -C aom-av1-lavish_b200`): " + LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.b200_last_error.restype = C.c_char_p
        L.b200_ctx_stream.restype = C.c_void_p
        L.b200_default_ctx.restype = C.c_void_p
        L.b200_scan.restype = C.POINTER(C.c_int16)
        L.b200_iscan.restype = C.POINTER(C.c_int16)
        _lib = L
    return _lib


class B200Error(RuntimeError):
    pass


def check(rc):
    if rc != OK:
        raise B200Error("libb200av1 error %d: %s" % (rc, lib().b200_last_error().decode()))


def vp(a):
    """void* of a numpy array (or None)"""
    return None if a is None else C.c_void_p(a.ctypes.data)


class Context:
    """b200_ctx wrapper.  Raises B200Error when no sm_100 GPU is present."""

    def __init__(self, device=0):
        self.h = C.c_void_p()
        check(lib().b200_ctx_create(device, C.byref(self.h)))

    def close(self):
        if self.h:
            lib().b200_ctx_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def stream(self):
        return lib().b200_ctx_stream(self.h)

    def sync(self):
        check(lib().b200_sync(self.h))

    def txq_batch_host(self, in_kind, q_kind, bd, flags, blocks, preds, src, pred, qrows,
                       coeff, qcoeff, dqcoeff, eob, dist):
        n = len(blocks)
        out_elems = 0
        for o in (coeff, qcoeff, dqcoeff):
            if o is not None:
                out_elems = o.size
        check(lib().b200_txq_batch_host(
            self.h, in_kind, q_kind, bd, C.c_uint(flags), vp(blocks), vp(preds), n, vp(src),
            C.c_size_t(src.size), vp(pred), C.c_size_t(0 if pred is None else pred.size),
            vp(qrows), 0 if qrows is None else len(qrows), vp(coeff), vp(qcoeff), vp(dqcoeff),
            C.c_size_t(out_elems), vp(eob), vp(dist)))

    def txq_batch_host_compact(self, in_kind, q_kind, bd, flags, blocks, preds, src, pred, qrows,
                               out_elems, levels, level_offset, eob, dist=None, coeff=None):
        """b200_txq_batch_host_compact; returns the bytes that crossed PCIe towards the host"""
        d2h = C.c_size_t(0)
        check(lib().b200_txq_batch_host_compact(
            self.h, in_kind, q_kind, bd, C.c_uint(flags), vp(blocks), vp(preds), len(blocks), vp(src),
            C.c_size_t(src.size), vp(pred), C.c_size_t(0 if pred is None else pred.size),
            vp(qrows), 0 if qrows is None else len(qrows), C.c_size_t(out_elems), vp(levels),
            C.c_size_t(levels.size), vp(level_offset), vp(eob), vp(dist), vp(coeff), C.byref(d2h)))
        return d2h.value

    def txq_launch(self, launch, stream=None):
        check(lib().b200_txq_launch(self.h, C.byref(launch), C.c_void_p(stream)))


def qrow_from_oracle(row):
    """harness.QRow (8-wide reference-style rows) -> one QROW_DTYPE record"""
    r = np.zeros((), QROW_DTYPE)
    for name in ("zbin", "round", "quant", "quant_shift", "round_fp", "quant_fp", "dequant"):
        r[name] = [getattr(row, name)[0], getattr(row, name)[1]]
    return r


# ------------------------------------------------------------------ device buffers ----
class DevBuf:
    """device allocation owned by a Context (b200_malloc / b200_free)"""

    def __init__(self, ctx, nbytes):
        self.ctx, self.nbytes = ctx, nbytes
        self.ptr = C.c_void_p()
        check(lib().b200_malloc(ctx.h, C.c_size_t(max(nbytes, 1)), C.byref(self.ptr)))

    def free(self):
        if self.ptr and self.ctx.h:  # a closed context has already released its device
            lib().b200_free(self.ctx.h, self.ptr)
        self.ptr = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def to_device(ctx, arr):
    arr = np.ascontiguousarray(arr)
    buf = DevBuf(ctx, arr.nbytes)
    check(lib().b200_h2d(ctx.h, buf.ptr, vp(arr), C.c_size_t(arr.nbytes)))
    ctx.sync()
    return buf


def to_host(ctx, buf, dtype, count):
    out = np.zeros(count, dtype)
    check(lib().b200_d2h(ctx.h, vp(out), buf.ptr, C.c_size_t(out.nbytes)))
    ctx.sync()
    return out


PIXBLOCK_DTYPE = np.dtype([("a_offset", "<u4"), ("b_offset", "<u4"), ("aux_offset", "<u4"),
                           ("a_stride", "<u2"), ("b_stride", "<u2"), ("w", "u1"), ("h", "u1"),
                           ("xoff", "u1"), ("yoff", "u1")])
QUANTBLOCK_DTYPE = np.dtype([("coeff_offset", "<u4"), ("out_offset", "<u4"), ("iscan_offset", "<u4"),
                             ("qm_offset", "<u4"), ("n_coeffs", "<u2"), ("qrow", "<u2"),
                             ("log_scale", "u1"), ("reserved", "u1", 3)])
ERRBLOCK_DTYPE = np.dtype([("coeff_offset", "<u4"), ("dqcoeff_offset", "<u4"), ("n_coeffs", "<u4"),
                           ("aux_offset", "<u4")])
assert PIXBLOCK_DTYPE.itemsize == 20 and QUANTBLOCK_DTYPE.itemsize == 24 and ERRBLOCK_DTYPE.itemsize == 16

PIX_SAD, PIX_SAD_SKIP, PIX_SAD_AVG, PIX_VAR, PIX_MSE, PIX_SUBPEL_VAR, PIX_SSE = range(7)
(QUANT_FP, QUANT_FP_HBD, QUANT_LP, QUANT_B, QUANT_B_HBD, QUANT_B_ADAPT, QUANT_B_ADAPT_HBD) = range(7)
ERR_LOWBD, ERR_LP, ERR_HBD, ERR_QM = range(4)


class PixLaunch(C.Structure):
    _fields_ = [("op", C.c_int), ("highbd", C.c_int), ("bd", C.c_int), ("n_blocks", C.c_int),
                ("blocks", C.c_void_p), ("a", C.c_void_p), ("b", C.c_void_p), ("aux", C.c_void_p),
                ("out", C.c_void_p)]


PIXMASK_DTYPE = np.dtype([("mask_offset", "<u4"), ("mask_stride", "<u2"), ("invert_mask", "u1"),
                          ("reserved", "u1")])
assert PIXMASK_DTYPE.itemsize == 8
(CMP_MASKED_SAD, CMP_MASKED_SUBPEL_VAR, CMP_SUBPEL_AVG_VAR, CMP_DIST_WTD_SUBPEL_AVG_VAR, CMP_DIST_WTD_SAD_AVG,
 CMP_OBMC_SAD, CMP_OBMC_VAR, CMP_OBMC_SUBPEL_VAR, CMP_AVG_PRED, CMP_DIST_WTD_AVG_PRED, CMP_MASK_PRED) = range(11)


class CompoundLaunch(C.Structure):
    _fields_ = [("op", C.c_int), ("highbd", C.c_int), ("bd", C.c_int), ("n_blocks", C.c_int),
                ("blocks", C.c_void_p), ("masks", C.c_void_p), ("pre", C.c_void_p), ("tgt", C.c_void_p),
                ("second_pred", C.c_void_p), ("mask", C.c_void_p), ("wsrc", C.c_void_p),
                ("obmc_mask", C.c_void_p), ("fwd_offset", C.c_int), ("bck_offset", C.c_int),
                ("out", C.c_void_p), ("comp_pred", C.c_void_p)]


class QuantLaunch(C.Structure):
    _fields_ = [("kind", C.c_int), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("coeff", C.c_void_p), ("qrows", C.c_void_p), ("iscan", C.c_void_p),
                ("qm", C.c_void_p), ("iqm", C.c_void_p), ("qcoeff", C.c_void_p),
                ("dqcoeff", C.c_void_p), ("eob", C.c_void_p)]


class ErrLaunch(C.Structure):
    _fields_ = [("kind", C.c_int), ("bd", C.c_int), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("coeff", C.c_void_p), ("dqcoeff", C.c_void_p), ("qm", C.c_void_p),
                ("scan", C.c_void_p), ("out", C.c_void_p)]


# ------------------------------------------------------------------ motion search ----
MEBLOCK_DTYPE = np.dtype([("src_offset", "<u4"), ("ref_offset", "<u4"), ("col_min", "<i2"),
                          ("col_max", "<i2"), ("row_min", "<i2"), ("row_max", "<i2"),
                          ("start_row", "<i2"), ("start_col", "<i2"), ("ref_mv_row", "<i2"),
                          ("ref_mv_col", "<i2"), ("bw", "u1"), ("bh", "u1"), ("search_method", "u1"),
                          ("step_param", "u1"), ("flags", "u1"), ("mv_cost_type", "u1"),
                          ("mesh_mv_diff_threshold", "u1"), ("subpel", "u1"),
                          ("force_mesh_thresh", "<i4"), ("sad_per_bit", "<i4"),
                          ("error_per_bit", "<i4")])
MERESULT_DTYPE = np.dtype([("best_row", "<i2"), ("best_col", "<i2"), ("second_row", "<i2"),
                           ("second_col", "<i2"), ("cost_list", "<i4", 5), ("ret", "<i4"),
                           ("distortion", "<i4"), ("sse", "<u4")])
assert MEBLOCK_DTYPE.itemsize == 44 and MERESULT_DTYPE.itemsize == 40
ME_RUN_MESH, ME_PRUNE_MESH, ME_FINE_INTERVAL, ME_SKIP_SAD, ME_ALLOW_HP = 1, 2, 4, 8, 16
ME_COST_LIST = 32
ME_REFINE_8P = 11
SUBPEL_TREE, SUBPEL_TREE_PRUNED, SUBPEL_TREE_PRUNED_MORE = 0, 1, 2


class MeLaunch(C.Structure):
    _fields_ = [("highbd", C.c_int), ("bd", C.c_int), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("cur", C.c_void_p), ("ref", C.c_void_p), ("cur_stride", C.c_int),
                ("ref_stride", C.c_int), ("mesh_range", C.c_int * 4), ("mesh_interval", C.c_int * 4),
                ("mvjcost", C.c_void_p), ("mvcost", C.c_void_p * 2), ("results", C.c_void_p),
                ("window_bytes", C.c_int), ("max_bw", C.c_int), ("max_bh", C.c_int),
                ("cost_lists", C.c_void_p), ("cost_list_stride", C.c_int), ("families", C.c_uint),
                ("cur_rows", C.c_int), ("ref_rows", C.c_int), ("max_col_span", C.c_int),
                ("max_row_span", C.c_int), ("start_from", C.c_void_p), ("stats", C.c_void_p),
                ("start_group", C.c_int), ("compound", C.c_void_p), ("second_pred", C.c_void_p),
                ("mask", C.c_void_p)]


MECOMPOUND_DTYPE = np.dtype([("second_offset", "<u4"), ("mask_offset", "<u4"), ("mask_stride", "<u2"),
                             ("invert_mask", "u1"), ("use_mask", "u1")])
assert MECOMPOUND_DTYPE.itemsize == 12


ME_FAMILY_NSTEP, ME_FAMILY_PATTERN, ME_FAMILY_REFINE = 1, 2, 4


# ---- Hadamard / SATD / averages (include/b200av1.h) ----
HAD_LOWBD, HAD_LP, HAD_HIGHBD, HAD_WHT = 0, 1, 2, 3
AVG_4X4, AVG_8X8, MINMAX_8X8 = 0, 1, 2
HADBLOCK_DTYPE = np.dtype([("src_offset", "<u4"), ("out_offset", "<u4"), ("stride", "<i4")])
AVGBLOCK_DTYPE = np.dtype([("a_offset", "<u4"), ("b_offset", "<u4"), ("a_stride", "<i4"),
                           ("b_stride", "<i4")])
assert HADBLOCK_DTYPE.itemsize == 12 and AVGBLOCK_DTYPE.itemsize == 16


class HadamardLaunch(C.Structure):
    _fields_ = [("n", C.c_int), ("kind", C.c_int), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("src", C.c_void_p), ("coeff", C.c_void_p), ("satd", C.c_void_p)]


class AvgLaunch(C.Structure):
    _fields_ = [("op", C.c_int), ("highbd", C.c_int), ("n_blocks", C.c_int), ("blocks", C.c_void_p),
                ("a", C.c_void_p), ("b", C.c_void_p), ("out", C.c_void_p)]


# ---- frame tier + block queue (csrc/queue.cu) -----------------------------------------
class TfParams(C.Structure):
    """B200TfParams (include/b200av1.h): tf_motion_search's parameter set-up."""
    _fields_ = [("bw", C.c_int), ("bh", C.c_int), ("border_in_pixels", C.c_int), ("search_method", C.c_int),
                ("step_param", C.c_int), ("mv_cost_type", C.c_int), ("mesh_range", C.c_int * 4),
                ("mesh_interval", C.c_int * 4), ("force_mesh_thresh", C.c_int), ("prune_mesh_search", C.c_int),
                ("mesh_search_mv_diff_threshold", C.c_int), ("use_downsampled_sad", C.c_int),
                ("sad_per_bit", C.c_int), ("error_per_bit", C.c_int), ("allow_hp", C.c_int),
                ("iters_per_step", C.c_int), ("subpel_tree", C.c_int), ("want_cost_list", C.c_int)]


TF_RESULT_DTYPE = np.dtype([("block_mv", "<i2", (2,)), ("block_mse", "<i4"), ("sub_mv", "<i2", (4, 2)),
                            ("sub_mse", "<i4", (4,)), ("ref_mv", "<i2", (2,))])
assert TF_RESULT_DTYPE.itemsize == 44


def tf_default_params(width, height):
    p = TfParams()
    lib().b200_queue_tf_default_params(width, height, C.byref(p))
    return p


class Frame:
    """b200_frame: a luma plane on the device with its replicated border."""

    def __init__(self, ctx, y, border, bd=8):
        y = np.ascontiguousarray(y)
        assert y.ndim == 2 and y.dtype == (np.uint8 if bd == 8 else np.uint16)
        self.ctx, self.h = ctx, C.c_void_p()
        self.height, self.width = y.shape
        self.border, self.bd = border, bd
        check(lib().b200_frame_upload(ctx.h, vp(y), y.shape[1], y.shape[1], y.shape[0], border, bd,
                                      C.byref(self.h)))

    def download(self):
        out = np.zeros((self.height + 2 * self.border, self.width + 2 * self.border),
                       np.uint8 if self.bd == 8 else np.uint16)
        check(lib().b200_frame_download(self.ctx.h, self.h, vp(out)))
        return out

    def free(self):
        if self.h:
            lib().b200_frame_free(self.ctx.h, self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def queue_tf_motion_search(ctx, cur, ref, params, ref_mvs=None):
    """b200_queue_tf_motion_search: every TF block of the frame pair; returns a TF_RESULT_DTYPE array."""
    rows = (cur.height + params.bh - 1) // params.bh
    cols = (cur.width + params.bw - 1) // params.bw
    out = np.zeros(rows * cols, TF_RESULT_DTYPE)
    mv = None
    if ref_mvs is not None:
        mv = np.ascontiguousarray(ref_mvs, np.int16)
        assert mv.size == 2 * rows * cols
    check(lib().b200_queue_tf_motion_search(ctx.h, cur.h, ref.h, C.byref(params), vp(mv) if mv is not None else None,
                                            vp(out)))
    return out.reshape(rows, cols)


class TplParams(C.Structure):
    """B200TplParams (include/b200av1.h): TPL motion_estimation's parameter set-up."""
    _fields_ = [("bw", C.c_int), ("bh", C.c_int), ("border_in_pixels", C.c_int), ("search_method", C.c_int),
                ("step_param", C.c_int), ("forced_stop", C.c_int), ("mv_cost_type", C.c_int),
                ("mesh_range", C.c_int * 4), ("mesh_interval", C.c_int * 4), ("force_mesh_thresh", C.c_int),
                ("prune_mesh_search", C.c_int), ("use_downsampled_sad", C.c_int), ("sad_per_bit", C.c_int),
                ("error_per_bit", C.c_int), ("allow_hp", C.c_int), ("iters_per_step", C.c_int),
                ("subpel_tree", C.c_int), ("want_cost_list", C.c_int), ("mvjcost", C.c_void_p),
                ("mvcost", C.c_void_p * 2)]


TPL_ME_DTYPE = np.dtype([("mv", "<i2", (2,)), ("sme", "<u4")])
TPL_RD_DTYPE = np.dtype([("rate_cost", "<i4"), ("eob", "<i4"), ("recon_error", "<i8"), ("sse", "<i8")])
assert TPL_ME_DTYPE.itemsize == 8 and TPL_RD_DTYPE.itemsize == 24


def tpl_default_params(width, height):
    p = TplParams()
    lib().b200_queue_tpl_default_params(width, height, C.byref(p))
    return p


def tpl_grid(width, height, params):
    mi_rows, mi_cols = ((height + 7) & ~7) >> 2, ((width + 7) & ~7) >> 2
    return (mi_rows * 4 + params.bh - 1) // params.bh, (mi_cols * 4 + params.bw - 1) // params.bw


def queue_tpl_motion_estimation(ctx, cur, ref, params, center_mvs=None):
    rows, cols = tpl_grid(cur.width, cur.height, params)
    out = np.zeros(rows * cols, TPL_ME_DTYPE)
    mv = None if center_mvs is None else np.ascontiguousarray(center_mvs, np.int16)
    check(lib().b200_queue_tpl_motion_estimation(ctx.h, cur.h, ref.h, C.byref(params),
                                                 vp(mv) if mv is not None else None, vp(out)))
    return out.reshape(rows, cols)


def queue_tpl_txfm_quant_rdcost(ctx, src, pred, rows, cols, bd, tx_size, qrow):
    """src / pred: 2-D host planes (uint8 or uint16); qrow: one QROW_DTYPE record."""
    src, pred = np.ascontiguousarray(src), np.ascontiguousarray(pred)
    out = np.zeros(rows * cols, TPL_RD_DTYPE)
    q = np.ascontiguousarray(qrow)
    check(lib().b200_queue_tpl_txfm_quant_rdcost(ctx.h, vp(src), src.shape[1], vp(pred), pred.shape[1], rows, cols,
                                                 bd, tx_size, vp(q), vp(out)))
    return out.reshape(rows, cols)
