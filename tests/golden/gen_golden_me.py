#!/usr/bin/env python3
"""Generate tests/golden/me_avg_golden.npz from the REFERENCE build (oracle/_ref): one seeded
frame pair with the results of the reference's av1_full_pixel_search (all methods),
av1_refining_search_8p_c, av1_find_best_sub_pixel_tree / _pruned / _pruned_more, and the
aom_dsp/avg.c Hadamard family, for machines where /root/reference is absent.
Run here: python tests/golden/gen_golden_me.py"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import harness as H  # noqa: E402
import me_common as M  # noqa: E402
from harness import i16p, i32p, ptr  # noqa: E402

R = H.ref()
assert R is not None, "oracle/_ref not built"
FIELDS = [n for n, _ in H.MeParams._fields_ if n not in ("src", "ref", "mvjcost", "mvcost0", "mvcost1",
          "second_pred", "mask", "mask_stride", "inv_mask")]  # compound operands: not part of these vectors


def pack(p):
    row = []
    for n in FIELDS:
        v = getattr(p, n)
        row += list(v) if hasattr(v, "__len__") else [v]
    return row + [p._bx, p._by]


def result(r):
    return [r.best_row, r.best_col, r.second_row, r.second_col] + list(r.cost_list) + \
        [r.ret, r.distortion, r.sse]


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    tables = H.make_mv_cost_tables(7)
    out["mv_joint"], out["mv_t0"], out["mv_t1"] = tables
    for bd in (8, 10):
        cur, refb = M.make_frames(rng, 192, 128, bd, flat=False)
        cur[:32] = cur[0, 0]                       # a flat band: SAD ties
        refb[M.BORDER:M.BORDER + 32] = cur[0, 0]
        out["cur_%d" % bd], out["ref_%d" % bd] = cur, refb
        rows, res, kinds = [], [], []
        for (bw, bh) in ((64, 64), (32, 32), (16, 16), (8, 8), (16, 8), (4, 4)):
            for method in range(11):
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                p.search_method = method
                p.step_param = int(rng.integers(0, 9))
                r = H.MeResult()
                R.refshim_full_pixel_search(C.byref(p), C.byref(r))
                rows.append(pack(p)); res.append(result(r)); kinds.append(0)
            p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
            r = H.MeResult()
            R.refshim_refining_search_8p(C.byref(p), C.byref(r))
            rows.append(pack(p)); res.append(result(r)); kinds.append(1)
            for tree in range(3):
                for st in range(4):
                    p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, st)
                    p.sp_tree = tree
                    M.random_cost_list(rng, p)
                    r = H.MeResult()
                    if tree == 0:
                        R.refshim_find_best_sub_pixel_tree(C.byref(p), C.byref(r))
                    else:
                        R.refshim_find_best_sub_pixel_tree_pruned(C.byref(p), C.byref(r))
                    rows.append(pack(p)); res.append(result(r)); kinds.append(2 + tree)
        out["me_params_%d" % bd] = np.array(rows, np.int64)
        out["me_results_%d" % bd] = np.array(res, np.int64)
        out["me_kinds_%d" % bd] = np.array(kinds, np.int32)
    # Hadamard family: 9-bit (13-bit for highbd) residual plus one full-range int16 block
    names = [("aom_hadamard_4x4_c", 4, 0), ("aom_hadamard_8x8_c", 8, 0), ("aom_hadamard_16x16_c", 16, 0),
             ("aom_hadamard_32x32_c", 32, 0), ("aom_hadamard_lp_8x8_c", 8, 1),
             ("aom_hadamard_lp_16x16_c", 16, 1), ("aom_highbd_hadamard_8x8_c", 8, 2),
             ("aom_highbd_hadamard_16x16_c", 16, 2), ("aom_highbd_hadamard_32x32_c", 32, 2)]
    for name, n, kind in names:
        lim = 4095 if kind == 2 else 255
        src = rng.integers(-lim, lim + 1, (2, n, n)).astype(np.int16)
        src[1] = rng.integers(-32768, 32768, (n, n)).astype(np.int16)
        exp = np.zeros((2, n * n), np.int32)
        for i in range(2):
            s = np.ascontiguousarray(src[i])
            if kind == 1:
                o = np.zeros(n * n, np.int16)
                getattr(R, name)(ptr(s, i16p), C.c_ssize_t(n), ptr(o, i16p))
            else:
                o = np.zeros(n * n, np.int32)
                getattr(R, name)(ptr(s, i16p), C.c_ssize_t(n), ptr(o, i32p))
            exp[i] = o
        out["had_in_%d_%d" % (n, kind)], out["had_out_%d_%d" % (n, kind)] = src, exp
    np.savez_compressed(os.path.join(HERE, "me_avg_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "me_avg_golden.npz"))


if __name__ == "__main__":
    main()
