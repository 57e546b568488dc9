"""Shared test plumbing: ctypes loaders for the oracle, the reference build (when
present) and small helpers.  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_LIB = os.path.join(ORACLE_DIR, "_ref", "libaomref.so")

TX_W = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TX_H = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]
TX_NAMES = ["%dx%d" % (w, h) for w, h in zip(TX_W, TX_H)]
N_TX_SIZES = 19
N_TX_TYPES = 16

# SAD / variance block sizes of aom_dsp_rtcd_defs.pl:42-58
BLOCK_SIZES = [(128, 128), (128, 64), (64, 128), (64, 64), (64, 32), (32, 64), (32, 32),
               (32, 16), (16, 32), (16, 16), (16, 8), (8, 16), (8, 8), (8, 4), (4, 8),
               (4, 4), (4, 16), (16, 4), (8, 32), (32, 8), (16, 64), (64, 16)]

i16p = C.POINTER(C.c_int16)
i32p = C.POINTER(C.c_int32)
u8p = C.POINTER(C.c_uint8)
u16p = C.POINTER(C.c_uint16)
u32p = C.POINTER(C.c_uint32)
i64p = C.POINTER(C.c_int64)


def ptr(a, t):
    return a.ctypes.data_as(t)


_oracle = None


def oracle():
    """Build (if needed) and load oracle/liboracle.so."""
    global _oracle
    if _oracle is None:
        subprocess.run(["make", "-s", "-C", ORACLE_DIR], check=True)
        lib = C.CDLL(os.path.join(ORACLE_DIR, "liboracle.so"))
        lib.av1o_scan.restype = i16p
        lib.av1o_iscan.restype = i16p
        lib.av1o_cospi.restype = i32p
        lib.av1o_sinpi.restype = i32p
        lib.av1o_dc_q.restype = C.c_int16
        lib.av1o_ac_q.restype = C.c_int16
        for n in ("av1o_block_error", "av1o_block_error_lp", "av1o_highbd_block_error",
                  "av1o_block_error_qm", "av1o_sse", "av1o_highbd_sse"):
            getattr(lib, n).restype = C.c_int64
        for n in ("av1o_sad", "av1o_sad_skip", "av1o_sad_avg", "av1o_highbd_sad",
                  "av1o_highbd_sad_skip", "av1o_variance", "av1o_mse",
                  "av1o_sub_pixel_variance", "av1o_highbd_variance",
                  "av1o_highbd_sub_pixel_variance"):
            getattr(lib, n).restype = C.c_uint
        _oracle = lib
    return _oracle


_ref = None


def have_ref():
    return os.path.exists(REF_LIB)


def _shim_protos(lib):
    for n in ("refshim_fwd_txfm_fn", "refshim_quantize_fp_fn"):
        getattr(lib, n).restype = C.c_void_p
        getattr(lib, n).argtypes = [C.c_int, C.c_int]
    return lib


def ref():
    """The reference's own C functions (oracle/_ref), or None when not built."""
    global _ref
    if _ref is None and have_ref():
        _ref = _shim_protos(C.CDLL(REF_LIB))
    return _ref


REF_AVX2_LIB = os.path.join(os.path.dirname(REF_LIB), "libaomref_avx2.so")
_ref_avx2 = None


def host_has_avx2():
    try:
        return " avx2 " in open("/proc/cpuinfo").read().replace("\n", " ")
    except OSError:
        return False


def ref_avx2():
    """The reference built with its x86 intrinsics and run-time dispatch (oracle/mk_ref.py, avx2
    flavour), or None when not built / the host has no AVX2."""
    global _ref_avx2
    if _ref_avx2 is None and os.path.exists(REF_AVX2_LIB) and host_has_avx2():
        _ref_avx2 = _shim_protos(C.CDLL(REF_AVX2_LIB))
    return _ref_avx2


class QRow(C.Structure):
    _fields_ = [(n, C.c_int16 * 8) for n in
                ("quant", "quant_shift", "zbin", "round", "quant_fp", "round_fp", "dequant")]


def oracle_qrow(bd, qindex, plane=0, sharpness=0):
    row = QRow()
    oracle().av1o_build_qrow(bd, qindex, plane, sharpness, C.byref(row))
    return row


def oracle_scan(tx_size, tx_type):
    n = oracle().av1o_max_eob(tx_size)
    s = np.ctypeslib.as_array(oracle().av1o_scan(tx_size, tx_type), shape=(n,)).copy()
    i = np.ctypeslib.as_array(oracle().av1o_iscan(tx_size, tx_type), shape=(n,)).copy()
    return s, i


def max_eob(tx_size):
    return min(TX_W[tx_size], 32) * min(TX_H[tx_size], 32)


def tx_valid(tx_size, tx_type):
    return bool(oracle().av1o_tx_valid(tx_size, tx_type))


def log_scale(tx_size):
    """av1_get_tx_scale, av1/common/idct.c:24"""
    pels = TX_W[tx_size] * TX_H[tx_size]
    return int(pels > 256) + int(pels > 1024)


def oracle_fwd_txfm2d(resid, stride, tx_size, tx_type, bd=8):
    out = np.zeros(max_eob(tx_size), np.int32)
    oracle().av1o_fwd_txfm2d(ptr(resid, i16p), ptr(out, i32p), stride, tx_size, tx_type, bd)
    return out


# ---------------------------------------------------------------- batch oracle ----
def oracle_txq_batch(src, blocks, rows, q_kind, bd, out_elems, in_kind=0, pred=None, preds=None,
                     with_err=True):
    """Run the oracle over a descriptor batch exactly like b200_txq_batch_host would.
    `rows` is a list of harness.QRow (oracle rows) indexed by block['qrow'].
    q_kind / in_kind use the B200_* numbering of include/b200av1.h."""
    O = oracle()
    coeff = np.zeros(out_elems, np.int32)
    qc = np.zeros(out_elems, np.int32)
    dq = np.zeros(out_elems, np.int32)
    eob = np.zeros(len(blocks), np.uint16)
    dist = np.zeros((len(blocks), 2), np.int64)
    for i, b in enumerate(blocks):
        ts, tt = int(b["tx_size"]), int(b["tx_type"])
        W, Hh = TX_W[ts], TX_H[ts]
        n = max_eob(ts)
        so, st = int(b["src_offset"]), int(b["src_stride"])
        if in_kind == 0:
            resid = src[so: so + (Hh - 1) * st + W]
            rstride = st
        else:
            po, ps = int(preds[i]["pred_offset"]), int(preds[i]["pred_stride"])
            resid = np.zeros((Hh, W), np.int16)
            s2 = np.lib.stride_tricks.as_strided(src[so:], (Hh, W), (st * src.itemsize, src.itemsize))
            p2 = np.lib.stride_tricks.as_strided(pred[po:], (Hh, W), (ps * pred.itemsize, pred.itemsize))
            resid[:] = (s2.astype(np.int32) - p2.astype(np.int32)).astype(np.int16)
            resid = resid.reshape(-1)
            rstride = W
        resid = np.ascontiguousarray(resid)
        oo = int(b["out_offset"])
        c = np.zeros(n, np.int32)
        O.av1o_fwd_txfm2d(ptr(resid, i16p), ptr(c, i32p), rstride, ts, tt, bd)
        coeff[oo:oo + n] = c
        if q_kind == 0:
            continue
        row = rows[int(b["qrow"])]
        scan, iscan = oracle_scan(ts, tt)
        q = np.zeros(n, np.int32)
        d = np.zeros(n, np.int32)
        e = C.c_uint16(0)
        ls = log_scale(ts)
        fp = q_kind in (1, 2)
        fn = {1: O.av1o_quantize_fp, 2: O.av1o_highbd_quantize_fp, 3: O.av1o_quantize_b,
              4: O.av1o_highbd_quantize_b}[q_kind]
        fn(ptr(c, i32p), C.c_ssize_t(n), row.zbin, row.round_fp if fp else row.round,
           row.quant_fp if fp else row.quant, row.quant_shift, ptr(q, i32p), ptr(d, i32p),
           row.dequant, C.byref(e), ptr(scan, i16p), ptr(iscan, i16p), None, None, ls)
        qc[oo:oo + n] = q
        dq[oo:oo + n] = d
        eob[i] = e.value
        if with_err:
            ssz = C.c_int64(0)
            if q_kind in (2, 4):
                err = O.av1o_highbd_block_error(ptr(c, i32p), ptr(d, i32p), C.c_ssize_t(n),
                                                C.byref(ssz), bd)
            else:
                err = O.av1o_block_error(ptr(c, i32p), ptr(d, i32p), C.c_ssize_t(n), C.byref(ssz))
            dist[i] = (err, ssz.value)
    return coeff, qc, dq, eob, dist


# ------------------------------------------------------------------ motion search ----
class MeParams(C.Structure):
    """oracle/av1o_me.h Av1oMeParams"""
    _fields_ = [("src", C.c_void_p), ("ref", C.c_void_p), ("src_stride", C.c_int),
                ("ref_stride", C.c_int), ("bw", C.c_int), ("bh", C.c_int), ("highbd", C.c_int),
                ("bd", C.c_int), ("col_min", C.c_int), ("col_max", C.c_int), ("row_min", C.c_int),
                ("row_max", C.c_int), ("search_method", C.c_int), ("step_param", C.c_int),
                ("run_mesh_search", C.c_int), ("prune_mesh_search", C.c_int),
                ("mesh_search_mv_diff_threshold", C.c_int), ("force_mesh_thresh", C.c_int),
                ("mesh_range", C.c_int * 4), ("mesh_interval", C.c_int * 4),
                ("fine_search_interval", C.c_int), ("use_skip_sad", C.c_int),
                ("mv_cost_type", C.c_int), ("ref_mv_row", C.c_int), ("ref_mv_col", C.c_int),
                ("sad_per_bit", C.c_int), ("error_per_bit", C.c_int), ("mvjcost", C.c_void_p),
                ("mvcost0", C.c_void_p), ("mvcost1", C.c_void_p), ("start_row", C.c_int),
                ("start_col", C.c_int), ("want_cost_list", C.c_int), ("want_second_best", C.c_int),
                ("allow_hp", C.c_int), ("forced_stop", C.c_int), ("iters_per_step", C.c_int),
                ("subpel_search_type", C.c_int), ("sp_col_min", C.c_int), ("sp_col_max", C.c_int),
                ("sp_row_min", C.c_int), ("sp_row_max", C.c_int), ("sp_start_row", C.c_int),
                ("sp_start_col", C.c_int), ("sp_tree", C.c_int), ("sp_have_cost_list", C.c_int),
                ("sp_cost_list", C.c_int * 5), ("second_pred", C.c_void_p), ("mask", C.c_void_p),
                ("mask_stride", C.c_int), ("inv_mask", C.c_int)]


class MeResult(C.Structure):
    _fields_ = [("best_row", C.c_int), ("best_col", C.c_int), ("second_row", C.c_int),
                ("second_col", C.c_int), ("cost_list", C.c_int * 5), ("ret", C.c_int),
                ("distortion", C.c_int), ("sse", C.c_uint)]

    def key(self, subpel=False):
        if subpel:
            return (self.best_row, self.best_col, self.ret, self.distortion, self.sse)
        return (self.best_row, self.best_col, self.second_row, self.second_col,
                tuple(self.cost_list), self.ret)


MV_TABLE_HALF = 1 << 14


def make_mv_cost_tables(seed=0):
    """synthetic monotone entropy-style cost tables: joint[4], comp[2][2*MV_TABLE_HALF+1]
    (index by centred pointer), in 1/512-bit units like av1_fill_mv_costs produces"""
    rng = np.random.default_rng(seed)
    d = np.abs(np.arange(-MV_TABLE_HALF, MV_TABLE_HALF + 1))
    base = (np.log2(d + 1.0) * 2 * 512 + 300).astype(np.int32)
    t0 = (base + rng.integers(0, 40, base.size)).astype(np.int32)
    t1 = (base + rng.integers(0, 40, base.size)).astype(np.int32)
    joint = np.array([300, 900, 1000, 1500], np.int32)
    return joint, t0, t1


def mv_cost_ptrs(joint, t0, t1):
    return (joint.ctypes.data, t0.ctypes.data + 4 * MV_TABLE_HALF, t1.ctypes.data + 4 * MV_TABLE_HALF)


def bordered(img, border):
    """replicate-pad like the encoder's frame border (aom_scale/yv12config.h:28-32)"""
    return np.ascontiguousarray(np.pad(img, border, mode="edge"))
