"""GPU parity of the batched motion search (b200_me_fullpel_launch / b200_me_subpel_launch)
against the oracle (itself pinned to the reference's av1_full_pixel_search /
av1_find_best_sub_pixel_tree): best mv, second-best mv, cost list, returned variance,
distortion and sse must be identical."""
import ctypes as C

import numpy as np
import pytest

import harness as H
import me_common as M

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def upload_tables(ctx, tables):
    import b200av1
    joint, t0, t1 = tables
    dj, d0, d1 = (b200av1.to_device(ctx, t) for t in (joint, t0, t1))
    return (dj, d0, d1), (dj.ptr.value, d0.ptr.value + 4 * H.MV_TABLE_HALF, d1.ptr.value + 4 * H.MV_TABLE_HALF)


def me_launch(ctx, cur, refb, params, tables_dev, subpel, window_bytes, tma=True, compound=None):
    import b200av1
    n = len(params)
    blocks = np.zeros(n, b200av1.MEBLOCK_DTYPE)
    for i, p in enumerate(params):
        b = blocks[i]
        b["src_offset"] = p._by * cur.shape[1] + p._bx
        b["ref_offset"] = (p._by + M.BORDER) * refb.shape[1] + p._bx + M.BORDER
        if subpel:
            b["col_min"], b["col_max"], b["row_min"], b["row_max"] = (p.sp_col_min, p.sp_col_max,
                                                                      p.sp_row_min, p.sp_row_max)
            b["start_row"], b["start_col"] = p.sp_start_row, p.sp_start_col
            want_cl = p.sp_have_cost_list
        else:
            want_cl = p.want_cost_list
            b["col_min"], b["col_max"], b["row_min"], b["row_max"] = (p.col_min, p.col_max,
                                                                      p.row_min, p.row_max)
            b["start_row"], b["start_col"] = p.start_row, p.start_col
        b["ref_mv_row"], b["ref_mv_col"] = p.ref_mv_row, p.ref_mv_col
        b["bw"], b["bh"] = p.bw, p.bh
        b["search_method"], b["step_param"] = (p.sp_tree if subpel else p.search_method), p.step_param
        b["flags"] = (p.run_mesh_search * b200av1.ME_RUN_MESH | p.prune_mesh_search * b200av1.ME_PRUNE_MESH |
                      p.fine_search_interval * b200av1.ME_FINE_INTERVAL |
                      p.use_skip_sad * b200av1.ME_SKIP_SAD | p.allow_hp * b200av1.ME_ALLOW_HP |
                      want_cl * b200av1.ME_COST_LIST)
        b["mv_cost_type"] = p.mv_cost_type
        b["mesh_mv_diff_threshold"] = p.mesh_search_mv_diff_threshold
        b["subpel"] = p.forced_stop | (p.iters_per_step << 2) | (p.subpel_search_type << 4)
        b["force_mesh_thresh"] = p.force_mesh_thresh
        b["sad_per_bit"], b["error_per_bit"] = p.sad_per_bit, p.error_per_bit
    d_cur, d_ref = b200av1.to_device(ctx, cur), b200av1.to_device(ctx, refb)
    d_blk = b200av1.to_device(ctx, blocks)
    d_res = b200av1.DevBuf(ctx, n * 40)
    l = b200av1.MeLaunch()
    l.highbd, l.bd, l.n_blocks = int(cur.itemsize == 2), params[0].bd, n
    l.blocks, l.cur, l.ref = d_blk.ptr, d_cur.ptr, d_ref.ptr
    l.cur_stride, l.ref_stride = cur.shape[1], refb.shape[1]
    for i in range(4):  # one mesh pattern per launch (MV_SPEED_FEATURES::mesh_patterns)
        l.mesh_range[i], l.mesh_interval[i] = params[0].mesh_range[i], params[0].mesh_interval[i]
    l.mvjcost, l.mvcost[0], l.mvcost[1] = tables_dev
    l.results = d_res.ptr
    l.window_bytes = window_bytes
    if tma and not subpel:
        # plane extents + largest window of the batch: the TMA engine stages block and window
        l.cur_rows, l.ref_rows = cur.shape[0], refb.shape[0]
        l.max_col_span = max(1, max(p.col_max - p.col_min for p in params))
        l.max_row_span = max(1, max(p.row_max - p.row_min for p in params))
        l.max_bw, l.max_bh = max(p.bw for p in params), max(p.bh for p in params)
    if subpel and any(p.sp_have_cost_list for p in params):
        cls = np.array([list(p.sp_cost_list) for p in params], np.int32)
        d_cls = b200av1.to_device(ctx, cls)
        l.cost_lists, l.cost_list_stride = d_cls.ptr, 5
    if compound is not None:
        # compound[i] = (second_pred block, mask array or None): pooled into one buffer each
        cd = np.zeros(n, b200av1.MECOMPOUND_DTYPE)
        sec, msk, so, mo = [], [], 0, 0
        for i, (p, (second, mask)) in enumerate(zip(params, compound)):
            cd[i]["second_offset"] = so
            sec.append(second)
            so += second.size
            if mask is not None:
                cd[i]["mask_offset"], cd[i]["mask_stride"] = mo, p.mask_stride
                cd[i]["invert_mask"], cd[i]["use_mask"] = p.inv_mask, 1
                msk.append(mask)
                mo += mask.size
        d_cd = b200av1.to_device(ctx, cd)
        d_sec = b200av1.to_device(ctx, np.concatenate(sec))
        d_msk = b200av1.to_device(ctx, np.concatenate(msk) if msk else np.zeros(4, np.uint8))
        l.compound, l.second_pred, l.mask = d_cd.ptr, d_sec.ptr, d_msk.ptr
    fn = b200av1.lib().b200_me_subpel_launch if subpel else b200av1.lib().b200_me_fullpel_launch
    b200av1.check(fn(ctx.h, C.byref(l), None))
    return b200av1.to_host(ctx, d_res, b200av1.MERESULT_DTYPE, n)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_fullpel_batch_vs_oracle(ctx, bd):
    O = H.oracle()
    rng = np.random.default_rng(700 + bd)
    tables = H.make_mv_cost_tables(1)
    keep, tdev = upload_tables(ctx, tables)
    for trial in range(4):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        params = [M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                  for (bw, bh) in M.FULLPEL_SIZES for _ in range(6)]
        exp = []
        for p in params:
            r = H.MeResult()
            O.av1o_full_pixel_search(C.byref(p), C.byref(r))
            exp.append(r)
        # window staged in shared memory by the TMA engine / by the threads / forced through
        # global memory: same answers
        for window, tma in ((150 * 1024, True), (150 * 1024, False), (0, False)):
            got = me_launch(ctx, cur, refb, params, tdev, False, window, tma)
            for i, (p, e) in enumerate(zip(params, exp)):
                g = got[i]
                tag = (trial, i, p.bw, p.bh, p.search_method, p.step_param, p.run_mesh_search,
                       p.use_skip_sad, p.mv_cost_type, window)
                assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"])) == \
                    (e.best_row, e.best_col, e.ret), tag
                if p.want_second_best:
                    assert (int(g["second_row"]), int(g["second_col"])) == (e.second_row, e.second_col), tag
                if p.want_cost_list:
                    assert tuple(int(v) for v in g["cost_list"]) == tuple(e.cost_list), tag


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_subpel_batch_vs_oracle(ctx, bd):
    O = H.oracle()
    rng = np.random.default_rng(800 + bd)
    tables = H.make_mv_cost_tables(2)
    keep, tdev = upload_tables(ctx, tables)
    for trial in range(3):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        params = [M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, st)
                  for (bw, bh) in M.SUBPEL_SIZES for st in range(4) for _ in range(2)]
        got = me_launch(ctx, cur, refb, params, tdev, True, 0)
        for i, p in enumerate(params):
            e = H.MeResult()
            O.av1o_find_best_sub_pixel_tree(C.byref(p), C.byref(e))
            g = got[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                    int(g["sse"])) == e.key(True), (trial, i, p.bw, p.bh, p.subpel_search_type,
                                                    p.forced_stop, p.iters_per_step)


def test_config3_shape_64x64_pm64_two_refs(ctx):
    """BASELINE config 3 shape at reduced frame size: 64x64 superblocks, +-64 range, NSTEP +
    forced mesh {64,8},{28,4},{15,1}, two reference frames, L1 (HD) mv cost; full-pel result
    feeds the sub-pel tree (8 taps) like av1_single_motion_search chains them."""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(33)
    tables = H.make_mv_cost_tables(3)
    keep, tdev = upload_tables(ctx, tables)
    w, h = 448, 256
    for ref_id in range(2):
        cur, refb = M.make_frames(rng, w, h, 8, flat=(ref_id == 1))
        params = []
        for by in range(0, h, 64):
            for bx in range(0, w, 64):
                p = M.base_params(cur, refb, bx, by, 64, 64, 8, tables)
                p.col_min, p.col_max = max(-64, -(bx + M.BORDER - 8)), min(64, w - 64 - bx + M.BORDER - 8)
                p.row_min, p.row_max = max(-64, -(by + M.BORDER - 8)), min(64, h - 64 - by + M.BORDER - 8)
                p.search_method, p.step_param = 0, O.av1o_init_search_range(64)
                p.run_mesh_search, p.mv_cost_type = 1, 3
                p.want_cost_list = p.want_second_best = 1
                params.append(p)
        got = me_launch(ctx, cur, refb, params, tdev, False, 150 * 1024)
        sub = []
        for i, p in enumerate(params):
            e = H.MeResult()
            O.av1o_full_pixel_search(C.byref(p), C.byref(e))
            g = got[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["second_row"]), int(g["second_col"]),
                    tuple(int(v) for v in g["cost_list"]), int(g["ret"])) == e.key(), (ref_id, i)
            p.subpel_search_type, p.allow_hp, p.forced_stop, p.iters_per_step = 3, 1, 0, 2
            p.sp_start_row, p.sp_start_col = e.best_row * 8, e.best_col * 8
            p.sp_row_min, p.sp_row_max = p.row_min * 8, p.row_max * 8
            p.sp_col_min, p.sp_col_max = p.col_min * 8, p.col_max * 8
            sub.append(p)
        got2 = me_launch(ctx, cur, refb, sub, tdev, True, 0)
        for i, p in enumerate(sub):
            e = H.MeResult()
            O.av1o_find_best_sub_pixel_tree(C.byref(p), C.byref(e))
            g = got2[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                    int(g["sse"])) == e.key(True), (ref_id, i)


MESH_PATTERNS = [  # speed_features.c:25-44
    ((64, 16), (24, 8), (12, 4), (7, 1)),   # good quality, mesh speed >= 3: 16-pixel stage
    ((64, 8), (14, 2), (7, 1), (7, 1)),     # mesh speed 2: 2-pixel stage (no strip path)
    ((64, 4), (16, 1), (0, 0), (0, 0)),     # intrabc
    ((24, 1), (24, 1), (0, 0), (0, 0)),     # dense first stage: several bands of rows
]


@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("pattern", range(len(MESH_PATTERNS)))
def test_fullpel_mesh_patterns(ctx, pattern, bd):
    """Every mesh stage shape of the reference's tables through the register-blocked SAD
    maps (step 1: column lanes; step 4/8/16: strips) and the generic path (step 2, row
    skipping, 4/8-pixel blocks), forced mesh, limits clipped at the frame edge."""
    O = H.oracle()
    rng = np.random.default_rng(900 + pattern + 17 * (bd - 8))
    tables = H.make_mv_cost_tables(4)
    keep, tdev = upload_tables(ctx, tables)
    mesh = MESH_PATTERNS[pattern]
    for trial in range(2):
        cur, refb = M.make_frames(rng, 320, 256, bd, flat=(trial == 1))
        if bd > 8 and trial == 0:  # extreme pixels: 64 rows of 12-bit extremes fill the packed 16-bit sums
            cur[64:200, 32:224] = (1 << bd) - 1
            refb[M.BORDER + 64:M.BORDER + 200, M.BORDER + 32:M.BORDER + 224] = 0
        params = []
        for (bw, bh) in [(64, 64), (32, 32), (16, 16), (128, 128), (8, 8), (16, 8), (64, 32), (32, 8),
                         (128, 64), (16, 64)]:
            for rep in range(3):
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                for i, (r, it) in enumerate(mesh):
                    p.mesh_range[i], p.mesh_interval[i] = r, it
                p.run_mesh_search, p.prune_mesh_search = 1, 0
                p.search_method = 0
                p.use_skip_sad = int(rep == 2 and bh >= 8)
                p.want_cost_list = p.want_second_best = 1
                params.append(p)
        got = me_launch(ctx, cur, refb, params, tdev, False, 150 * 1024)
        for i, p in enumerate(params):
            e = H.MeResult()
            O.av1o_full_pixel_search(C.byref(p), C.byref(e))
            g = got[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["second_row"]), int(g["second_col"]),
                    tuple(int(v) for v in g["cost_list"]), int(g["ret"])) == e.key(), \
                (pattern, bd, trial, i, p.bw, p.bh, p.use_skip_sad, p.fine_search_interval)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_pattern_search_methods(ctx, bd):
    """HEX / BIGDIA / SQUARE / FAST_HEX / FAST_DIAMOND / FAST_BIGDIA / VFAST_DIAMOND
    (pattern_search, mcomp.c:1017-1316) with and without a cost list, staged window and
    global-memory path."""
    O = H.oracle()
    rng = np.random.default_rng(950 + bd)
    tables = H.make_mv_cost_tables(6)
    keep, tdev = upload_tables(ctx, tables)
    for trial in range(3):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial == 2))
        params = []
        for (bw, bh) in M.FULLPEL_SIZES:
            for method in range(4, 11):
                p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                p.search_method = method
                p.step_param = int(rng.integers(0, 12))
                p.run_mesh_search = int(rng.integers(0, 4) == 0)
                params.append(p)
        exp = []
        for p in params:
            e = H.MeResult()
            O.av1o_full_pixel_search(C.byref(p), C.byref(e))
            exp.append(e)
        for window in (150 * 1024, 0):
            got = me_launch(ctx, cur, refb, params, tdev, False, window)
            for i, (p, e) in enumerate(zip(params, exp)):
                g = got[i]
                tag = (trial, i, p.bw, p.bh, p.search_method, p.step_param, p.want_cost_list,
                       p.use_skip_sad, p.run_mesh_search, window)
                assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"])) == \
                    (e.best_row, e.best_col, e.ret), tag
                if p.want_cost_list:
                    assert tuple(int(v) for v in g["cost_list"]) == tuple(e.cost_list), tag
                if p.want_second_best:
                    assert (int(g["second_row"]), int(g["second_col"])) == (e.second_row, e.second_col), tag


@pytest.mark.parametrize("bd", [8, 10, 12])
@pytest.mark.parametrize("tree", [1, 2])
def test_subpel_pruned_batch_vs_oracle(ctx, bd, tree):
    """av1_find_best_sub_pixel_tree_pruned / _pruned_more through b200_me_subpel_launch
    (search_method = B200_SUBPEL_TREE_PRUNED / _PRUNED_MORE, cost lists in device memory)."""
    O = H.oracle()
    rng = np.random.default_rng(900 + bd + 7 * tree)
    tables = H.make_mv_cost_tables(2)
    keep, tdev = upload_tables(ctx, tables)
    for trial in range(3):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        params = []
        for (bw, bh) in M.SUBPEL_SIZES:
            for rep in range(8):
                p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, rep % 4)
                p.sp_tree = tree
                params.append(M.random_cost_list(rng, p))
        got = me_launch(ctx, cur, refb, params, tdev, True, 0)
        for i, p in enumerate(params):
            e = H.MeResult()
            O.av1o_find_best_sub_pixel_tree_pruned(C.byref(p), C.byref(e))
            g = got[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                    int(g["sse"])) == e.key(True), (trial, i, p.bw, p.bh, tree, p.forced_stop,
                                                    p.iters_per_step, p.sp_have_cost_list,
                                                    list(p.sp_cost_list))


def test_fullpel_cost_list_chains_into_pruned_subpel(ctx):
    """the full-pel launch's results[].cost_list feeds the pruned sub-pel launch in place
    (cost_lists = &results[0].cost_list, stride 10 ints), like av1_single_motion_search does
    with its cost_list array (motion_search_facade.c:120-494)."""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(951)
    tables = H.make_mv_cost_tables(3)
    keep, tdev = upload_tables(ctx, tables)
    cur, refb = M.make_frames(rng, 256, 192, 8)
    params = []
    for (bw, bh) in ((32, 32), (16, 16), (64, 64)):
        for _ in range(10):
            p = M.random_fullpel_params(rng, cur, refb, bw, bh, 8, tables)
            p.search_method, p.run_mesh_search, p.use_skip_sad = 0, 0, 0
            p.want_cost_list = 1
            params.append(p)
    got = me_launch(ctx, cur, refb, params, tdev, False, 150 * 1024)
    sub = []
    for i, p in enumerate(params):
        e = H.MeResult()
        O.av1o_full_pixel_search(C.byref(p), C.byref(e))
        assert tuple(int(v) for v in got[i]["cost_list"]) == tuple(e.cost_list)
        p.sp_tree, p.sp_have_cost_list = 1 + i % 2, 1
        for j in range(5):
            p.sp_cost_list[j] = e.cost_list[j]
        p.subpel_search_type, p.allow_hp, p.forced_stop, p.iters_per_step = 0, 1, 0, 2
        p.sp_start_row, p.sp_start_col = e.best_row * 8, e.best_col * 8
        p.sp_row_min, p.sp_row_max = p.row_min * 8, p.row_max * 8
        p.sp_col_min, p.sp_col_max = p.col_min * 8, p.col_max * 8
        sub.append(p)
    got2 = me_launch(ctx, cur, refb, sub, tdev, True, 0)
    for i, p in enumerate(sub):
        e = H.MeResult()
        O.av1o_find_best_sub_pixel_tree_pruned(C.byref(p), C.byref(e))
        g = got2[i]
        assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                int(g["sse"])) == e.key(True), (i, p.bw, p.sp_tree)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_refining_search_8p_batch_vs_oracle(ctx, bd):
    """av1_refining_search_8p_c through b200_me_fullpel_launch (search_method = B200_ME_REFINE_8P)"""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(970 + bd)
    tables = H.make_mv_cost_tables(4)
    keep, tdev = upload_tables(ctx, tables)
    for trial in range(3):
        cur, refb = M.make_frames(rng, 256, 192, bd, flat=(trial % 3 == 2))
        params = [M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
                  for (bw, bh) in M.FULLPEL_SIZES for _ in range(5)]
        exp = []
        for p in params:
            e = H.MeResult()
            O.av1o_refining_search_8p(C.byref(p), C.byref(e))
            exp.append(e)
            p.search_method = b200av1.ME_REFINE_8P
        for window in (150 * 1024, 0):
            got = me_launch(ctx, cur, refb, params, tdev, False, window)
            for i, (p, e) in enumerate(zip(params, exp)):
                g = got[i]
                assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"])) == \
                    (e.best_row, e.best_col, e.ret), (trial, i, p.bw, p.bh, p.use_skip_sad, window)


@pytest.mark.parametrize("bd", [8, 10, 12])
def test_compound_refine_and_subpel_vs_oracle(ctx, bd):
    """compound search (second_pred / mask, av1_set_ms_compound_refs): av1_refining_search_8p_c and the
    three sub-pel trees -- one iteration of av1_joint_motion_search (motion_search_facade.c:594-652) --
    against the oracle, average and masked blends, all block sizes"""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(1900 + bd)
    tables = H.make_mv_cost_tables(7)
    keep, tdev = upload_tables(ctx, tables)
    cur, refb = M.make_frames(rng, 320, 256, bd, flat=False)
    sizes = [(64, 64), (32, 32), (16, 16), (8, 8), (16, 8), (32, 64), (128, 128), (4, 4), (8, 32), (4, 16)]
    alive, params, comp = [], [], []
    for (bw, bh) in sizes:
        for rep in range(6):
            p = M.random_fullpel_params(rng, cur, refb, bw, bh, bd, tables)
            p.search_method = b200av1.ME_REFINE_8P
            k = []
            M.add_compound(rng, p, bd, k, masked=rep % 2 == 1)
            alive.append(k)
            params.append(p)
            comp.append((k[0], k[1] if len(k) > 1 else None))
    got = me_launch(ctx, cur, refb, params, tdev, False, 0, tma=False, compound=comp)
    for i, p in enumerate(params):
        e = H.MeResult()
        O.av1o_refining_search_8p(C.byref(p), C.byref(e))
        g = got[i]
        assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"])) == (e.best_row, e.best_col, e.ret), \
            (bd, i, p.bw, p.bh)
    for tree in (0, 1, 2):
        params, comp = [], []
        for (bw, bh) in sizes:
            for rep in range(4):
                p = M.random_subpel_params(rng, cur, refb, bw, bh, bd, tables, int(rng.integers(0, 4)))
                p.sp_tree = tree
                if tree:
                    M.random_cost_list(rng, p)
                k = []
                M.add_compound(rng, p, bd, k, masked=rep % 2 == 1)
                alive.append(k)
                params.append(p)
                comp.append((k[0], k[1] if len(k) > 1 else None))
        got = me_launch(ctx, cur, refb, params, tdev, True, 0, compound=comp)
        fo = O.av1o_find_best_sub_pixel_tree if tree == 0 else O.av1o_find_best_sub_pixel_tree_pruned
        for i, p in enumerate(params):
            e = H.MeResult()
            fo(C.byref(p), C.byref(e))
            g = got[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]), int(g["sse"])) == \
                e.key(True), (bd, tree, i, p.bw, p.bh, p.subpel_search_type)


@pytest.mark.parametrize("bd,shuffle", [(8, False), (8, True), (10, False)])
def test_me_batch_host_pipelined_vs_launch_tier_and_oracle(ctx, bd, shuffle):
    """b200_me_batch_host cuts a large queue into chunks whose plane ranges are copied while the
    previous chunks' kernels run: the results must equal those of ONE device-resident launch pair
    over the same descriptors (any descriptor order), and the oracle's on a sample."""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(4100 + bd + shuffle)
    w, h, B = 1024, 576, M.BORDER
    dt = np.uint8 if bd == 8 else np.uint16
    hi = (1 << bd) - 1
    base = rng.integers(0, hi + 1, (h // 8, w // 8)).astype(np.float32)
    cur = np.clip(np.kron(base, np.ones((8, 8), np.float32)) + rng.integers(-6, 7, (h, w)), 0, hi).astype(dt)
    rs, rh = w + 2 * B, h + 2 * B
    refs = np.zeros((2, rh, rs), dt)
    for r in range(2):
        dy, dx = int(rng.integers(-5, 6)), int(rng.integers(-5, 6))
        ref = np.roll(cur, (dy, dx), (0, 1)).astype(np.int32) + rng.integers(-2, 3, (h, w))
        refs[r] = np.pad(np.clip(ref, 0, hi).astype(dt), B, mode="edge")
    sites = [(r, bx, by) for r in range(2) for by in range(0, h, 16) for bx in range(0, w, 16)]
    if shuffle:
        sites = [sites[i] for i in rng.permutation(len(sites))]
    n = len(sites)
    assert n >= 4 * 148 * 6  # several chunks of two waves on a B200
    blocks = np.zeros(n, b200av1.MEBLOCK_DTYPE)
    blocks["src_offset"] = [by * w + bx for (_, bx, by) in sites]
    blocks["ref_offset"] = [(r * rh + by + B) * rs + bx + B for (r, bx, by) in sites]
    blocks["col_min"], blocks["col_max"], blocks["row_min"], blocks["row_max"] = -12, 12, -12, 12
    blocks["bw"], blocks["bh"] = 16, 16
    blocks["search_method"] = 0
    blocks["step_param"] = 6
    blocks["mv_cost_type"] = 3
    blocks["sad_per_bit"], blocks["error_per_bit"] = 40, 90
    sub = blocks.copy()
    sub["col_min"], sub["col_max"], sub["row_min"], sub["row_max"] = -96, 96, -96, 96
    sub["flags"] = b200av1.ME_ALLOW_HP
    sub["subpel"] = 0 | (2 << 2) | (3 << 4)

    def launch_struct(pb, pc, pr, pres):
        l = b200av1.MeLaunch()
        l.highbd, l.bd, l.n_blocks = int(bd > 8), bd, n
        l.blocks, l.cur, l.ref, l.results = pb, pc, pr, pres
        l.cur_stride, l.ref_stride = w, rs
        l.window_bytes, l.max_bw, l.max_bh = 16 * 1024, 16, 16
        l.families = b200av1.ME_FAMILY_NSTEP
        l.cur_rows, l.ref_rows = h, 2 * rh
        l.max_col_span, l.max_row_span = 24, 24
        return l

    lib = b200av1.lib()
    # host tier
    rf = np.zeros(n, b200av1.MERESULT_DTYPE)
    rsub = np.zeros(n, b200av1.MERESULT_DTYPE)
    hl = launch_struct(blocks.ctypes.data, cur.ctypes.data, refs.ctypes.data, None)
    for _ in range(2):  # the second call reuses the scratch of the first
        rf[:] = 0
        rsub[:] = 0
        b200av1.check(lib.b200_me_batch_host(ctx.h, C.byref(hl), C.c_size_t(cur.nbytes), C.c_size_t(refs.nbytes),
                                             C.c_void_p(sub.ctypes.data), C.c_void_p(rf.ctypes.data),
                                             C.c_void_p(rsub.ctypes.data)))
    # launch tier, everything resident
    d_cur, d_ref = b200av1.to_device(ctx, cur), b200av1.to_device(ctx, refs)
    d_blk, d_sub = b200av1.to_device(ctx, blocks), b200av1.to_device(ctx, sub)
    d_rf, d_rs = b200av1.DevBuf(ctx, n * 40), b200av1.DevBuf(ctx, n * 40)
    l = launch_struct(d_blk.ptr, d_cur.ptr, d_ref.ptr, d_rf.ptr)
    b200av1.check(lib.b200_me_fullpel_launch(ctx.h, C.byref(l), None))
    l.blocks, l.results, l.start_from = d_sub.ptr, d_rs.ptr, d_rf.ptr
    b200av1.check(lib.b200_me_subpel_launch(ctx.h, C.byref(l), None))
    ef = b200av1.to_host(ctx, d_rf, b200av1.MERESULT_DTYPE, n)
    es = b200av1.to_host(ctx, d_rs, b200av1.MERESULT_DTYPE, n)
    assert rf.tobytes() == ef.tobytes()
    assert rsub.tobytes() == es.tobytes()
    assert np.any(rf["best_row"] != 0) and np.all(rsub["ret"] < 2 ** 30)
    # oracle on a sample (first, last and random searches)
    esz = cur.itemsize
    for i in [0, n - 1] + [int(v) for v in rng.integers(0, n, 40)]:
        r, bx, by = sites[i]
        p = H.MeParams()
        p.src = cur.ctypes.data + (by * w + bx) * esz
        p.ref = refs.ctypes.data + ((r * rh + by + B) * rs + bx + B) * esz
        p.src_stride, p.ref_stride = w, rs
        p.bw, p.bh, p.highbd, p.bd = 16, 16, int(bd > 8), bd
        p.col_min, p.col_max, p.row_min, p.row_max = -12, 12, -12, 12
        p.search_method, p.step_param = 0, 6
        p.mv_cost_type, p.sad_per_bit, p.error_per_bit = 3, 40, 90
        p.want_cost_list = p.want_second_best = 1
        p.allow_hp, p.forced_stop, p.iters_per_step, p.subpel_search_type = 1, 0, 2, 3
        p.sp_col_min, p.sp_col_max, p.sp_row_min, p.sp_row_max = -96, 96, -96, 96
        e = H.MeResult()
        O.av1o_full_pixel_search(C.byref(p), C.byref(e))
        assert (int(rf[i]["best_row"]), int(rf[i]["best_col"]), int(rf[i]["ret"])) == (e.best_row, e.best_col, e.ret), i
        p.sp_start_row, p.sp_start_col = e.best_row * 8, e.best_col * 8
        e2 = H.MeResult()
        O.av1o_find_best_sub_pixel_tree(C.byref(p), C.byref(e2))
        assert (int(rsub[i]["best_row"]), int(rsub[i]["best_col"]), int(rsub[i]["ret"]),
                int(rsub[i]["distortion"]), int(rsub[i]["sse"])) == \
            (e2.best_row, e2.best_col, e2.ret, e2.distortion, e2.sse), i
