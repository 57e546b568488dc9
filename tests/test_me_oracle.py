"""Pin the oracle's motion search (oracle/av1o_mcomp.c) against the reference's own
av1_full_pixel_search / av1_find_best_sub_pixel_tree called through oracle/ref_shim.c.
The reference has no unit test for these loops, so the inputs are
synthetic frames incl. flat regions that force SAD ties."""
import ctypes as C

import numpy as np
import pytest

import harness as H
import me_common as M

pytestmark = [pytest.mark.ref, pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")]


def run_both(fn_o, fn_r, p, subpel=False):
    a, b = H.MeResult(), H.MeResult()
    fn_o(C.byref(p), C.byref(a))
    fn_r(C.byref(p), C.byref(b))
    return a.key(subpel), b.key(subpel)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_full_pixel_search_vs_reference(bd):
    O, R = H.oracle(), H.ref()
    rng = np.random.default_rng(300 + bd)
    tables = H.make_mv_cost_tables(1)
    n = 0
    for trial in range(10):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        for (bw, bh) in M.FULLPEL_SIZES:
            for rep in range(2):
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                a, b = run_both(O.av1o_full_pixel_search, R.refshim_full_pixel_search, p)
                assert a == b, (trial, bw, bh, p.search_method, p.step_param, p.run_mesh_search,
                                p.use_skip_sad, p.mv_cost_type, a, b)
                n += 1
    assert n == 200


def test_init_search_range():
    O, R = H.oracle(), H.ref()
    for size in (1, 8, 16, 32, 64, 128, 500, 1023, 2000):
        assert O.av1o_init_search_range(size) == R.av1_init_search_range(size)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_sub_pixel_tree_vs_reference(bd):
    O, R = H.oracle(), H.ref()
    rng = np.random.default_rng(400 + bd)
    tables = H.make_mv_cost_tables(2)
    for trial in range(6):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        for (bw, bh) in M.SUBPEL_SIZES:
            for st in range(4):
                p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, st)
                a, b = run_both(O.av1o_find_best_sub_pixel_tree, R.refshim_find_best_sub_pixel_tree,
                                p, subpel=True)
                assert a == b, (trial, bw, bh, st, p.allow_hp, p.forced_stop, p.iters_per_step, a, b)


PATTERN_METHODS = [4, 5, 6, 7, 8, 9, 10]  # HEX, BIGDIA, SQUARE, FAST_HEX, FAST_DIAMOND, FAST_BIGDIA, VFAST_DIAMOND


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_pattern_search_vs_reference(bd):
    """pattern_search family (mcomp.c:1017-1316): with and without a cost list (the cost
    list changes the last scale of the BIGDIA family), limits clipped at the frame edge,
    flat regions for ties, row-skipping SAD with its redo rule, optional mesh follow-up."""
    O, R = H.oracle(), H.ref()
    rng = np.random.default_rng(500 + bd)
    tables = H.make_mv_cost_tables(5)
    n = 0
    for trial in range(6):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        for (bw, bh) in M.FULLPEL_SIZES:
            for method in PATTERN_METHODS:
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                p.search_method = method
                p.step_param = int(rng.integers(0, 12))
                p.run_mesh_search = int(rng.integers(0, 4) == 0)
                a, b = run_both(O.av1o_full_pixel_search, R.refshim_full_pixel_search, p)
                assert a == b, (trial, bw, bh, method, p.step_param, p.want_cost_list, p.use_skip_sad,
                                p.mv_cost_type, a, b)
                n += 1
    assert n == 6 * len(M.FULLPEL_SIZES) * len(PATTERN_METHODS)


@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("tree", [1, 2])
def test_sub_pixel_tree_pruned_vs_reference(bd, tree):
    """av1_find_best_sub_pixel_tree_pruned (mcomp.c:2992) / _pruned_more (:2907): with no
    cost list, an unusable one (INT_MAX entry), a well-behaved bowl and an arbitrary one."""
    O, R = H.oracle(), H.ref()
    rng = np.random.default_rng(600 + bd + 7 * tree)
    tables = H.make_mv_cost_tables(3)
    for trial in range(6):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        for (bw, bh) in M.SUBPEL_SIZES:
            for rep in range(4):
                p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, rep)
                p.sp_tree = tree
                M.random_cost_list(rng, p)
                a, b = run_both(O.av1o_find_best_sub_pixel_tree_pruned,
                                R.refshim_find_best_sub_pixel_tree_pruned, p, subpel=True)
                assert a == b, (trial, bw, bh, tree, p.allow_hp, p.forced_stop, p.iters_per_step,
                                p.sp_have_cost_list, list(p.sp_cost_list), a, b)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_refining_search_8p_vs_reference(bd):
    """av1_refining_search_8p_c (mcomp.c:1683), single reference"""
    O, R = H.oracle(), H.ref()
    rng = np.random.default_rng(700 + bd)
    tables = H.make_mv_cost_tables(4)
    for trial in range(6):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        for (bw, bh) in M.FULLPEL_SIZES:
            for rep in range(3):
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                a, b = run_both(O.av1o_refining_search_8p, R.refshim_refining_search_8p, p)
                assert a == b, (trial, bw, bh, p.use_skip_sad, p.mv_cost_type, a, b)
