"""Full-size runs of the BASELINE configurations through the C ABI on the GPU.

config[3]: 10-bit txfm + av1_highbd_quantize_fp + av1_highbd_block_error on a synthetic 4K
frame (luma plane, fused subtract from uint16 planes).  The oracle covers the whole plane
for the large transform sizes and a seeded sample of blocks for the small ones; on top of
that, size-independent properties are checked on EVERY block of the frame:
  * block error: dist == sum (coeff - dqcoeff)^2, ssz == sum coeff^2, each rounded
    >> 2*(bd-8) as av1_highbd_block_error_c does (rdopt.c:663-682);
  * eob == 1 + last scan position holding a non-zero qcoeff (av1_quantize.c:60-66);
  * dqcoeff == (qcoeff * dequant) >> log_scale in magnitude, sign kept (av1_quantize.c:156-160);
  * the guarded 32-bit butterflies and the forced 64-bit path agree on every output word.
config[1] at full 1080p size gets the same properties for all 19 sizes."""
import numpy as np
import pytest

import harness as H
import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def run_plane(ctx, src, pred, stride, pstride, w, h, ts, bd, rows, qrows, seed, force_exact=False):
    import b200av1
    W, Hh, n = H.TX_W[ts], H.TX_H[ts], H.max_eob(ts)
    offs = synth.tile_blocks(w, h, stride, W, Hh)
    poffs = synth.tile_blocks(w, h, pstride, W, Hh)
    types = np.array([t for t in range(16) if H.tx_valid(ts, t)], np.uint8)
    rng = np.random.default_rng(seed + ts)
    nb = len(offs)
    blocks = np.zeros(nb, b200av1.TXBLOCK_DTYPE)
    preds = np.zeros(nb, b200av1.TXPRED_DTYPE)
    blocks["src_offset"], blocks["src_stride"] = offs, stride
    blocks["out_offset"] = np.arange(nb, dtype=np.uint32) * n
    blocks["qrow"] = rng.integers(0, len(rows), nb)
    blocks["tx_size"] = ts
    blocks["tx_type"] = types[rng.integers(0, len(types), nb)]
    preds["pred_offset"], preds["pred_stride"] = poffs, pstride
    coeff, qc, dq = (np.zeros(nb * n, np.int32) for _ in range(3))
    eob = np.zeros(nb, np.uint16)
    dist = np.zeros((nb, 2), np.int64)
    q_kind = b200av1.Q_FP if bd == 8 else b200av1.Q_FP_HBD
    in_kind = b200av1.IN_PIXELS8 if bd == 8 else b200av1.IN_PIXELS16
    flags = b200av1.TXQ_STORE_COEFF | b200av1.TXQ_BLOCK_ERROR | (b200av1.TXQ_FORCE_EXACT if force_exact else 0)
    ctx.txq_batch_host(in_kind, q_kind, bd, flags, blocks, preds, src, pred, qrows, coeff, qc, dq, eob, dist)
    return blocks, preds, (coeff, qc, dq, eob, dist), (q_kind, in_kind)


def check_properties(ts, bd, blocks, outs, rows):
    coeff, qc, dq, eob, dist = outs
    n, ls = H.max_eob(ts), H.log_scale(ts)
    nb = len(blocks)
    c2, q2, d2 = (a.reshape(nb, n).astype(np.int64) for a in (coeff, qc, dq))
    # block error with the reference's rounding
    shift = 2 * (bd - 8)
    rnd = (1 << shift) >> 1
    err = ((c2 - d2) ** 2).sum(1)
    ssz = (c2 ** 2).sum(1)
    if shift:
        err, ssz = (err + rnd) >> shift, (ssz + rnd) >> shift
    assert np.array_equal(dist[:, 0], err) and np.array_equal(dist[:, 1], ssz), H.TX_NAMES[ts]
    # eob and de-quantisation, per tx type (the scan depends on it)
    dequant = np.array([[r.dequant[0], r.dequant[1]] for r in rows], np.int64)
    for tt in np.unique(blocks["tx_type"]):
        sel = np.nonzero(blocks["tx_type"] == tt)[0]
        scan, _ = H.oracle_scan(ts, int(tt))
        qs = q2[sel][:, scan]
        nz = qs != 0
        last = np.where(nz.any(1), n - np.argmax(nz[:, ::-1], axis=1), 0)
        assert np.array_equal(eob[sel].astype(np.int64), last), (H.TX_NAMES[ts], int(tt))
        dqv = dequant[blocks["qrow"][sel]]
        dq_tab = np.repeat(dqv[:, 1:2], n, axis=1)
        dq_tab[:, 0] = dqv[:, 0]
        expect = np.sign(q2[sel]) * ((np.abs(q2[sel]) * dq_tab) >> ls)
        assert np.array_equal(d2[sel], expect), (H.TX_NAMES[ts], int(tt))


def compare_with_oracle(ts, bd, blocks, preds, outs, rows, kinds, src, pred, sample, seed):
    n = H.max_eob(ts)
    idx = np.arange(len(blocks))
    if sample and sample < len(blocks):
        idx = np.sort(np.random.default_rng(seed).choice(len(blocks), sample, replace=False))
    sub, subp = blocks[idx].copy(), preds[idx].copy()
    sub["out_offset"] = np.arange(len(idx), dtype=np.uint32) * n
    r = H.oracle_txq_batch(src, sub, rows, kinds[0], bd, len(idx) * n, kinds[1], pred, subp)
    coeff, qc, dq, eob, dist = outs
    pick = (idx[:, None] * n + np.arange(n)[None, :]).reshape(-1)
    tag = (H.TX_NAMES[ts], bd, len(idx))
    assert np.array_equal(coeff[pick], r[0]), tag
    assert np.array_equal(qc[pick], r[1]), tag
    assert np.array_equal(dq[pick], r[2]), tag
    assert np.array_equal(eob[idx], r[3]), tag
    assert np.array_equal(dist[idx], r[4]), tag


_PLANES = {}


def make_planes(rng, w, h, bd, stress_rows):
    key = (w, h, bd, stress_rows)
    if key not in _PLANES:
        _PLANES[key] = _make_planes(rng, w, h, bd, stress_rows)
    return _PLANES[key]


def _make_planes(rng, w, h, bd, stress_rows):
    stride, pstride = w + 64, w + 96
    dt = np.uint8 if bd == 8 else np.uint16
    srcp = np.zeros((h, stride), dt)
    predp = np.zeros((h, pstride), dt)
    srcp[:, :w] = synth.smooth_frame(rng, w, h, bd)
    predp[:, :w] = synth.pred_from(rng, srcp[:, :w], bd)
    hi = (1 << bd) - 1
    # stress band: residual uniform over the whole +-(2^bd - 1) range
    srcp[:stress_rows, :w] = rng.integers(0, hi + 1, (stress_rows, w))
    predp[:stress_rows, :w] = rng.integers(0, hi + 1, (stress_rows, w))
    return srcp.reshape(-1), predp.reshape(-1), stride, pstride


def test_config4_4k_10bit_full_plane(ctx):
    import b200av1
    bd, w, h = 10, 3840, 2160
    rng = np.random.default_rng(4001)
    src, pred, stride, pstride = make_planes(rng, w, h, bd, 128)
    rows = [H.oracle_qrow(bd, q) for q in (0, 48, 128, 255)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    # (tx size, oracle sample: 0 = whole plane)
    plan = [(0, 3000), (1, 3000), (2, 4000), (3, 0), (4, 0), (9, 3000), (12, 0), (13, 3000), (18, 0)]
    for ts, sample in plan:
        blocks, preds, outs, kinds = run_plane(ctx, src, pred, stride, pstride, w, h, ts, bd, rows, qrows, 4100)
        check_properties(ts, bd, blocks, outs, rows)
        compare_with_oracle(ts, bd, blocks, preds, outs, rows, kinds, src, pred, sample, 4200 + ts)
        _, _, outs2, _ = run_plane(ctx, src, pred, stride, pstride, w, h, ts, bd, rows, qrows, 4100,
                                   force_exact=True)
        for a, b in zip(outs, outs2):
            assert np.array_equal(a, b), (H.TX_NAMES[ts], "guarded vs exact")


@pytest.mark.parametrize("ts", range(19), ids=H.TX_NAMES)
def test_config2_1080p_8bit_full_plane_properties(ctx, ts):
    import b200av1
    bd, w, h = 8, 1920, 1080
    rng = np.random.default_rng(2001)
    src, pred, stride, pstride = make_planes(rng, w, h, bd, 64)
    rows = [H.oracle_qrow(bd, q) for q in (0, 32, 96, 160, 255)]
    qrows = np.array([b200av1.qrow_from_oracle(r) for r in rows], b200av1.QROW_DTYPE)
    blocks, preds, outs, kinds = run_plane(ctx, src, pred, stride, pstride, w, h, ts, bd, rows, qrows, 2100)
    check_properties(ts, bd, blocks, outs, rows)
    compare_with_oracle(ts, bd, blocks, preds, outs, rows, kinds, src, pred, 1500, 2200 + ts)


@pytest.mark.parametrize("bd", [8, 10])
def test_config3_1080p_motion_search_full_frame(ctx, bd):
    """BASELINE config[2] at full size: every 64x64 superblock of a 1080p frame (480) against
    two references, +-64, NSTEP + the 4-stage good-quality mesh, L1 (HD) cost, then the 8-tap
    sub-pel tree started from the full-pel winner.  8-bit: all 960 searches against the oracle;
    10-bit (the config[4] variant, highbd SAD >> 2): every third one."""
    import ctypes as C
    import b200av1
    import me_common as M
    from test_me_gpu import me_launch, upload_tables
    O = H.oracle()
    rng = np.random.default_rng(3000 + bd)
    tables = H.make_mv_cost_tables(3)
    keep, tdev = upload_tables(ctx, tables)
    w, h = 1920, 1080
    cur = synth.smooth_frame(rng, w, h, bd)
    step = 1 if bd == 8 else 3
    for ref_id in range(2):
        dy, dx = int(rng.integers(-20, 21)), int(rng.integers(-20, 21))
        ref = np.roll(cur, (dy, dx), (0, 1)).astype(np.int32) + np.rint(rng.normal(0, 2, cur.shape)).astype(np.int32)
        refb = H.bordered(np.clip(ref, 0, (1 << bd) - 1).astype(cur.dtype), M.BORDER)
        params = []
        for by in range(0, h - 63, 64):
            for bx in range(0, w - 63, 64):
                p = M.base_params(cur, refb, bx, by, 64, 64, bd, tables)
                p.col_min, p.col_max, p.row_min, p.row_max = -64, 64, -64, 64
                p.search_method, p.step_param = 0, O.av1o_init_search_range(64)
                p.run_mesh_search, p.mv_cost_type = 1, 3
                p.want_cost_list = p.want_second_best = 1
                params.append(p)
        assert len(params) == 480
        got = me_launch(ctx, cur, refb, params, tdev, False, (48 if bd == 8 else 88) * 1024)
        sub, exp_full = [], {}
        for i, p in enumerate(params):
            if i % step == 0:
                e = H.MeResult()
                O.av1o_full_pixel_search(C.byref(p), C.byref(e))
                g = got[i]
                assert (int(g["best_row"]), int(g["best_col"]), int(g["second_row"]), int(g["second_col"]),
                        tuple(int(v) for v in g["cost_list"]), int(g["ret"])) == e.key(), (bd, ref_id, i)
            p.subpel_search_type, p.allow_hp, p.forced_stop, p.iters_per_step = 3, 1, 0, 2
            p.sp_start_row, p.sp_start_col = int(got[i]["best_row"]) * 8, int(got[i]["best_col"]) * 8
            p.sp_row_min, p.sp_row_max, p.sp_col_min, p.sp_col_max = -512, 512, -512, 512
            sub.append(p)
        got2 = me_launch(ctx, cur, refb, sub, tdev, True, 0)
        for i, p in enumerate(sub):
            if i % step:
                continue
            e = H.MeResult()
            O.av1o_find_best_sub_pixel_tree(C.byref(p), C.byref(e))
            g = got2[i]
            assert (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                    int(g["sse"])) == e.key(True), (bd, ref_id, i)
