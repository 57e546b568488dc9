"""Oracle vs the committed round-2 golden vectors (tests/golden/r2_golden.npz, produced from the reference
build by tests/golden/gen_golden_r2.py): inverse transforms incl. the lossless WHT, the Laplacian rate
estimate, the compound / masked / OBMC SAD and variance families.  Runs anywhere -- no /root/reference."""
import ctypes as C
import os

import numpy as np

import harness as H
from harness import i32p, ptr, u16p
from test_compound_oracle import Compound

G = np.load(os.path.join(H.ROOT, "tests", "golden", "r2_golden.npz"))


def test_inverse_transform_golden():
    O = H.oracle()
    n_checked = 0
    for ts in range(19):
        w, h = H.TX_W[ts], H.TX_H[ts]
        for bd in (8, 10, 12):
            key = "itx_co_%d_%d" % (ts, bd)
            if key not in G:
                continue
            co, pr, rc = G[key], G["itx_pred_%d_%d" % (ts, bd)], G["itx_recon_%d_%d" % (ts, bd)]
            for tt in (0, 1, 5, 6, 9, 10, 13, 15):
                if not H.tx_valid(ts, tt):
                    continue
                a = np.ascontiguousarray(pr[tt])
                O.av1o_inv_txfm2d_add(ptr(np.ascontiguousarray(co[tt]), i32p), ptr(a, u16p), w, ts, tt, bd)
                assert np.array_equal(a, rc[tt]), (ts, tt, bd)
                n_checked += 1
    assert n_checked > 150
    for eob in (1, 2, 16):
        a = np.ascontiguousarray(G["iwht_pred_%d" % eob])
        O.av1o_iwht4x4_add(ptr(np.ascontiguousarray(G["iwht_co_%d" % eob]), i32p), ptr(a, u16p), 4, eob, 10)
        assert np.array_equal(a, G["iwht_recon_%d" % eob])


def test_rate_estimate_golden():
    O = H.oracle()
    for ts, tt, eob, want in G["est"]:
        q = np.ascontiguousarray(G["est_q_%d_%d" % (ts, tt)])
        assert O.av1o_cost_coeffs_txb_estimate(ptr(q, i32p), int(eob), int(ts), int(tt)) == int(want)


def test_compound_golden():
    O = H.oracle()
    for hb, bd in ((0, 8), (1, 10)):
        tag = "cmp%d_" % bd
        pre, tgt, mask = (np.ascontiguousarray(G[tag + k]) for k in ("pre", "tgt", "mask"))
        second, wsrc, omask = (np.ascontiguousarray(G[tag + k]) for k in ("second", "wsrc", "omask"))

        def cs(w, h, blend=0, subpel=0, inv=0, obmc=False):
            return Compound(w, h, hb, bd, pre.ctypes.data, 80, subpel, 3, 5, tgt.ctypes.data, 72, blend,
                            second.ctypes.data, mask.ctypes.data, 70, inv, 11, 5,
                            wsrc.ctypes.data if obmc else None, omask.ctypes.data if obmc else None)
        for row in G[tag + "expect"]:
            w, h = int(row[0]), int(row[1])
            sse = C.c_uint(0)
            got = [w, h, O.av1o_compound_sad(C.byref(cs(w, h, 3, 0, 1)))]
            got += [O.av1o_compound_variance(C.byref(cs(w, h, 3, 1, 0)), C.byref(sse)), sse.value]
            got += [O.av1o_compound_variance(C.byref(cs(w, h, 1, 1)), C.byref(sse)), sse.value]
            got += [O.av1o_compound_variance(C.byref(cs(w, h, 2, 1)), C.byref(sse)), sse.value]
            got += [O.av1o_compound_sad(C.byref(cs(w, h, 2, 0)))]
            got += [O.av1o_obmc_sad(C.byref(cs(w, h, obmc=True)))]
            got += [O.av1o_obmc_variance(C.byref(cs(w, h, obmc=True)), C.byref(sse)), sse.value]
            got += [O.av1o_obmc_variance(C.byref(cs(w, h, subpel=1, obmc=True)), C.byref(sse)), sse.value]
            assert [int(v) & 0xffffffff for v in got] == [int(v) for v in row], (hb, w, h)
