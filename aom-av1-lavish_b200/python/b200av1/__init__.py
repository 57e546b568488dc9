"""ctypes view of libb200av1.so (the C ABI declared in include/b200av1.h).

This is harness plumbing for tests and bench.py -- the product is the shared library
itself.  Loading fails loudly when the library has not been built; calls fail loudly
when there is no sm_100 GPU (the library has no CPU fallback)."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.dirname(os.path.dirname(_HERE))
LIB_PATH = os.path.join(PKG_DIR, "libb200av1.so")

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

    def txq_launch(self, launch, stream=None):
        check(lib().b200_txq_launch(self.h, C.byref(launch), C.c_void_p(stream)))


def qrow_from_oracle(row):
    """harness.QRow (8-wide reference-style rows) -> one QROW_DTYPE record"""
    r = np.zeros((), QROW_DTYPE)
    for name in ("zbin", "round", "quant", "quant_shift", "round_fp", "quant_fp", "dequant"):
        r[name] = [getattr(row, name)[0], getattr(row, name)[1]]
    return r
