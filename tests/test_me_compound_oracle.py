"""Compound motion search (second_pred / mask in the MSBuffers, av1_set_ms_compound_refs): the oracle's
av1_refining_search_8p_c and sub-pel trees against the reference's own functions (oracle/_ref) -- the two
searches of one iteration of av1_joint_motion_search (motion_search_facade.c:594-652)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
import me_common as M

pytestmark = [pytest.mark.ref,
              pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]

SIZES = [(64, 64), (32, 32), (16, 16), (8, 8), (16, 8), (32, 64), (128, 128), (4, 4), (8, 32)]


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_compound_refining_search_8p(bd):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(9900 + bd)
    tables = H.make_mv_cost_tables(5)
    cur, refb = M.make_frames(rng, 320, 256, bd, flat=False)
    n = 0
    for (bw, bh) in SIZES:
        for rep in range(12):
            keep = []
            p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
            M.add_compound(rng, p, bd, keep, masked=rep % 2 == 1)
            a, b = H.MeResult(), H.MeResult()
            R.refshim_refining_search_8p(C.byref(p), C.byref(a))
            O.av1o_refining_search_8p(C.byref(p), C.byref(b))
            assert (a.best_row, a.best_col, a.ret) == (b.best_row, b.best_col, b.ret), (bd, bw, bh, rep)
            n += 1
    assert n == 108


@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("tree", [0, 1, 2])
def test_compound_subpel_trees(bd, tree):
    R, O = H.ref(), H.oracle()
    rng = np.random.default_rng(9950 + bd + 7 * tree)
    tables = H.make_mv_cost_tables(6)
    cur, refb = M.make_frames(rng, 320, 256, bd, flat=False)
    fr = [R.refshim_find_best_sub_pixel_tree, R.refshim_find_best_sub_pixel_tree_pruned,
          R.refshim_find_best_sub_pixel_tree_pruned][tree]
    fo = [O.av1o_find_best_sub_pixel_tree, O.av1o_find_best_sub_pixel_tree_pruned,
          O.av1o_find_best_sub_pixel_tree_pruned][tree]
    for (bw, bh) in SIZES[:7]:
        for rep in range(6):
            keep = []
            p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, int(rng.integers(0, 4)))
            p.sp_tree = tree
            if tree:
                M.random_cost_list(rng, p)
            M.add_compound(rng, p, bd, keep, masked=rep % 2 == 1)
            a, b = H.MeResult(), H.MeResult()
            fr(C.byref(p), C.byref(a))
            fo(C.byref(p), C.byref(b))
            assert a.key(True) == b.key(True), (bd, tree, bw, bh, rep, p.subpel_search_type)
