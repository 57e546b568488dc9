"""Oracle vs committed golden vectors (tests/golden/hotpath_golden.npz, produced from the
reference build by tests/golden/gen_golden.py).  Runs anywhere -- no /root/reference."""
import ctypes as C
import os

import numpy as np

import harness as H
from harness import i16p, i32p, ptr, u8p

G = np.load(os.path.join(H.ROOT, "tests", "golden", "hotpath_golden.npz"))


def test_txfm_golden():
    for ts in range(19):
        W = H.TX_W[ts]
        inp, exp = G["txfm_in_%d" % ts], G["txfm_out_%d" % ts]
        for tt in range(16):
            if not H.tx_valid(ts, tt):
                continue
            got = H.oracle_fwd_txfm2d(np.ascontiguousarray(inp[tt]), W, ts, tt, 10)
            assert np.array_equal(got, exp[tt]), (ts, tt)


def test_quantize_fp_golden():
    O = H.oracle()
    row = H.oracle_qrow(8, 60)
    for ts in range(19):
        n, ls = H.max_eob(ts), H.log_scale(ts)
        coeff = np.ascontiguousarray(G["q_coeff_%d" % ts])
        scan, iscan = H.oracle_scan(ts, 0)
        qc, dq, eob = np.zeros(n, np.int32), np.zeros(n, np.int32), C.c_uint16(0)
        O.av1o_quantize_fp(ptr(coeff, i32p), C.c_ssize_t(n), row.zbin, row.round_fp, row.quant_fp,
                           row.quant_shift, ptr(qc, i32p), ptr(dq, i32p), row.dequant, C.byref(eob),
                           ptr(scan, i16p), ptr(iscan, i16p), None, None, ls)
        assert np.array_equal(qc, G["q_qc_%d" % ts]) and np.array_equal(dq, G["q_dq_%d" % ts])
        assert eob.value == int(G["q_eob_%d" % ts][0])


def test_pixel_golden():
    O = H.oracle()
    a, b = np.ascontiguousarray(G["pix_a"]), np.ascontiguousarray(G["pix_b"])
    for (w, h), exp in zip(H.BLOCK_SIZES, G["pix_results"]):
        sse, sse2 = C.c_uint(0), C.c_uint(0)
        got = [O.av1o_sad(ptr(a, u8p), 144, ptr(b, u8p), 144, w, h),
               O.av1o_variance(ptr(a, u8p), 144, ptr(b, u8p), 144, w, h, C.byref(sse)), 0,
               O.av1o_sub_pixel_variance(ptr(a, u8p), 144, 3, 5, ptr(b, u8p), 144, w, h, C.byref(sse2)), 0]
        got[2], got[4] = sse.value, sse2.value
        assert got == [int(v) for v in exp], (w, h)


# ---- motion search + Hadamard golden vectors (tests/golden/gen_golden_me.py) ----
GM = np.load(os.path.join(H.ROOT, "tests", "golden", "me_avg_golden.npz"))
ME_FIELDS = [n for n, _ in H.MeParams._fields_ if n not in ("src", "ref", "mvjcost", "mvcost0", "mvcost1",
             "second_pred", "mask", "mask_stride", "inv_mask")]  # compound operands: not part of these vectors


def _unpack(row, cur, refb, tables, border):
    p = H.MeParams()
    it = iter(int(v) for v in row)
    for n in ME_FIELDS:
        cur_v = getattr(p, n)
        if hasattr(cur_v, "__len__"):
            for i in range(len(cur_v)):
                cur_v[i] = next(it)
        else:
            setattr(p, n, next(it))
    bx, by = next(it), next(it)
    esz = cur.itemsize
    p.src = cur.ctypes.data + (by * cur.shape[1] + bx) * esz
    p.ref = refb.ctypes.data + ((by + border) * refb.shape[1] + bx + border) * esz
    p.mvjcost, p.mvcost0, p.mvcost1 = H.mv_cost_ptrs(*tables)
    return p


def test_motion_search_golden():
    """full-pel (11 methods), 8-point refinement, sub-pel tree / pruned / pruned_more: the
    oracle reproduces the results the reference build produced for the committed frames"""
    import me_common as M
    O = H.oracle()
    tables = tuple(np.ascontiguousarray(GM[k]) for k in ("mv_joint", "mv_t0", "mv_t1"))
    fns = [O.av1o_full_pixel_search, O.av1o_refining_search_8p, O.av1o_find_best_sub_pixel_tree,
           O.av1o_find_best_sub_pixel_tree_pruned, O.av1o_find_best_sub_pixel_tree_pruned]
    n = 0
    for bd in (8, 10):
        cur, refb = np.ascontiguousarray(GM["cur_%d" % bd]), np.ascontiguousarray(GM["ref_%d" % bd])
        for row, exp, kind in zip(GM["me_params_%d" % bd], GM["me_results_%d" % bd], GM["me_kinds_%d" % bd]):
            p = _unpack(row, cur, refb, tables, M.BORDER)
            r = H.MeResult()
            fns[int(kind)](C.byref(p), C.byref(r))
            exp = [int(v) for v in exp]
            if kind >= 2:
                got = [r.best_row, r.best_col, r.ret, r.distortion, r.sse]
                assert got == [exp[0], exp[1], exp[9], exp[10], exp[11]], (bd, n, int(kind))
            else:
                got = [r.best_row, r.best_col, r.ret]
                assert got == [exp[0], exp[1], exp[9]], (bd, n, int(kind))
                if kind == 0 and p.want_second_best:
                    assert [r.second_row, r.second_col] == exp[2:4], (bd, n)
                if kind == 0 and p.want_cost_list:
                    assert list(r.cost_list) == exp[4:9], (bd, n)
            n += 1
    assert n == 2 * 6 * (11 + 1 + 12)


def test_hadamard_golden():
    O = H.oracle()
    for (n, kind) in ((4, 0), (8, 0), (16, 0), (32, 0), (8, 1), (16, 1), (8, 2), (16, 2), (32, 2)):
        src, exp = GM["had_in_%d_%d" % (n, kind)], GM["had_out_%d_%d" % (n, kind)]
        for i in range(2):
            s = np.ascontiguousarray(src[i])
            got = np.zeros(n * n, np.int32)
            O.av1o_hadamard(ptr(s, i16p), C.c_ssize_t(n), ptr(got, i32p), n, kind)
            assert np.array_equal(got, exp[i]), (n, kind, i)
