"""GPU parity of the stand-alone batched quantiser and block-error launches
(b200_quant_launch / b200_block_error_launch, include/b200av1.h) called DIRECTLY -- the
drop-in tier only reaches them with qm == NULL.  Covers the qmatrix branch of the FP
quantisers (reference av1/encoder/av1_quantize.c:96-120, :150-169), of the B quantisers
(aom_dsp/quantize.c:108, :261) and the fork's av1_block_error_qm (tx_search.c:1049)."""
import ctypes as C

import numpy as np
import pytest

import harness as H
from harness import i16p, i32p, ptr, u8p

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def _coeffs(rng, n, bd, mode, highbd):
    lim = (1 << (7 + bd)) if highbd else 32767
    c = np.zeros(n, np.int32)
    if mode == "rand":
        c[:] = rng.integers(-lim, lim, n)
    elif mode == "small":
        c[:] = rng.integers(-40, 41, n)
    elif mode == "sparse":
        k = int(rng.integers(1, n + 1))
        c[:k] = rng.integers(-lim, lim, k)
    elif mode == "neg":
        c[:] = -8191
    return c


KINDS = [  # (B200_QUANT_*, oracle function, uses fp rows, highbd, takes qm)
    (0, "av1o_quantize_fp", True, 0, True),
    (1, "av1o_highbd_quantize_fp", True, 1, True),
    (3, "av1o_quantize_b", False, 0, True),
    (4, "av1o_highbd_quantize_b", False, 1, True),
    (5, "av1o_quantize_b_adaptive", False, 0, False),
    (6, "av1o_highbd_quantize_b_adaptive", False, 1, False),
]


@pytest.mark.parametrize("with_qm", [0, 1], ids=["flat", "qmatrix"])
@pytest.mark.parametrize("kind", KINDS, ids=[k[1] for k in KINDS])
def test_quant_launch_vs_oracle(ctx, kind, with_qm):
    import b200av1
    qk, oname, fp, highbd, takes_qm = kind
    if with_qm and not takes_qm:
        pytest.skip("the adaptive quantisers take no qmatrix (aom_dsp/quantize.c:16)")
    O = H.oracle()
    ofn = getattr(O, oname)
    rng = np.random.default_rng(4100 + qk * 7 + with_qm)
    bds = (10, 12) if highbd else (8,)
    cases = []
    for ts in (0, 1, 2, 3, 4, 5, 9, 11, 13, 16, 18):
        for qi in (0, 17, 96, 200, 255):
            for mode in ("rand", "small", "sparse", "neg"):
                cases.append((ts, qi, mode, bds[len(cases) % len(bds)]))
    n_tot = sum(H.max_eob(c[0]) for c in cases)
    coeff = np.zeros(n_tot, np.int32)
    iscan_all = np.zeros(n_tot, np.int16)
    qm_all = rng.integers(16, 200, n_tot).astype(np.uint8)   # qm_val_t weights, AOM_QM_BITS = 5
    iqm_all = rng.integers(16, 200, n_tot).astype(np.uint8)
    blocks = np.zeros(len(cases), b200av1.QUANTBLOCK_DTYPE)
    rows, qrows = [], np.zeros(len(cases), b200av1.QROW_DTYPE)
    exp_q, exp_dq, exp_eob = np.zeros(n_tot, np.int32), np.zeros(n_tot, np.int32), []
    o = 0
    for i, (ts, qi, mode, bd) in enumerate(cases):
        n, ls = H.max_eob(ts), H.log_scale(ts)
        tt = [t for t in range(16) if H.tx_valid(ts, t)][i % 3 if ts < 3 else 0]
        scan, iscan = H.oracle_scan(ts, tt)
        c = _coeffs(rng, n, bd, mode, highbd)
        coeff[o:o + n] = c
        iscan_all[o:o + n] = iscan
        row = H.oracle_qrow(bd, qi, plane=i % 3)
        rows.append(row)
        qrows[i] = b200av1.qrow_from_oracle(row)
        blocks[i] = (o, o, o, o, n, i, ls, 0)
        q, dq, e = np.zeros(n, np.int32), np.zeros(n, np.int32), C.c_uint16(0)
        qm = np.ascontiguousarray(qm_all[o:o + n])
        iqm = np.ascontiguousarray(iqm_all[o:o + n])
        args = [ptr(c, i32p), C.c_ssize_t(n), row.zbin, row.round_fp if fp else row.round,
                row.quant_fp if fp else row.quant, row.quant_shift, ptr(q, i32p), ptr(dq, i32p),
                row.dequant, C.byref(e), ptr(scan, i16p), ptr(iscan, i16p)]
        if takes_qm:
            args += [ptr(qm, u8p), ptr(iqm, u8p)] if with_qm else [None, None]
        args.append(ls)
        ofn(*args)
        exp_q[o:o + n], exp_dq[o:o + n] = q, dq
        exp_eob.append(e.value)
        o += n
    d_coeff, d_blk = b200av1.to_device(ctx, coeff), b200av1.to_device(ctx, blocks)
    d_rows, d_iscan = b200av1.to_device(ctx, qrows), b200av1.to_device(ctx, iscan_all)
    d_qm, d_iqm = b200av1.to_device(ctx, qm_all), b200av1.to_device(ctx, iqm_all)
    d_q, d_dq = b200av1.DevBuf(ctx, n_tot * 4), b200av1.DevBuf(ctx, n_tot * 4)
    d_eob = b200av1.DevBuf(ctx, len(cases) * 2)
    l = b200av1.QuantLaunch()
    l.kind, l.n_blocks = qk, len(cases)
    l.blocks, l.coeff, l.qrows, l.iscan = d_blk.ptr, d_coeff.ptr, d_rows.ptr, d_iscan.ptr
    if with_qm:
        l.qm, l.iqm = d_qm.ptr, d_iqm.ptr
    l.qcoeff, l.dqcoeff, l.eob = d_q.ptr, d_dq.ptr, d_eob.ptr
    b200av1.check(b200av1.lib().b200_quant_launch(ctx.h, C.byref(l), None))
    got_q = b200av1.to_host(ctx, d_q, np.int32, n_tot)
    got_dq = b200av1.to_host(ctx, d_dq, np.int32, n_tot)
    got_eob = b200av1.to_host(ctx, d_eob, np.uint16, len(cases))
    o = 0
    for i, (ts, qi, mode, bd) in enumerate(cases):
        n = H.max_eob(ts)
        tag = (oname, H.TX_NAMES[ts], qi, mode, bd, with_qm)
        assert int(got_eob[i]) == exp_eob[i], tag
        assert np.array_equal(got_q[o:o + n], exp_q[o:o + n]), tag
        assert np.array_equal(got_dq[o:o + n], exp_dq[o:o + n]), tag
        o += n


@pytest.mark.parametrize("kind", [0, 1, 2, 3], ids=["lowbd", "lp", "hbd", "qm"])
def test_block_error_launch_vs_oracle(ctx, kind):
    """B200_ERR_LOWBD / LP / HBD / QM through b200_block_error_launch"""
    import b200av1
    O = H.oracle()
    rng = np.random.default_rng(4200 + kind)
    cases = [(ts, bd) for ts in range(19) for bd in ((10, 12) if kind == 2 else (8,))] * 2
    n_tot = sum(H.max_eob(ts) for ts, _ in cases)
    lp = kind == 1
    dt = np.int16 if lp else np.int32
    coeff, dq = np.zeros(n_tot, dt), np.zeros(n_tot, dt)
    qm_all = rng.integers(1, 256, n_tot).astype(np.uint8)
    scan_all = np.zeros(n_tot, np.int16)
    blocks = np.zeros(len(cases), b200av1.ERRBLOCK_DTYPE)
    exp = np.zeros((len(cases), 2), np.int64)
    o = 0
    for i, (ts, bd) in enumerate(cases):
        n = H.max_eob(ts)
        lim = 32767 if lp else (1 << (7 + bd + (3 if i % 5 == 0 else 0)))  # some near-overflow magnitudes
        c = rng.integers(-lim, lim + 1, n).astype(dt)
        d = (c.astype(np.int64) + rng.integers(-300, 301, n)).clip(-lim, lim).astype(dt)
        tt = [t for t in range(16) if H.tx_valid(ts, t)][i % 2 if ts < 3 else 0]
        scan, _ = H.oracle_scan(ts, tt)
        coeff[o:o + n], dq[o:o + n], scan_all[o:o + n] = c, d, scan
        blocks[i] = (o, o, n, o)
        ssz = C.c_int64(0)
        if kind == 0:
            err = O.av1o_block_error(ptr(c, i32p), ptr(d, i32p), C.c_ssize_t(n), C.byref(ssz))
        elif kind == 1:
            err = O.av1o_block_error_lp(ptr(c, i16p), ptr(d, i16p), C.c_ssize_t(n))
        elif kind == 2:
            err = O.av1o_highbd_block_error(ptr(c, i32p), ptr(d, i32p), C.c_ssize_t(n), C.byref(ssz), bd)
        else:
            qm = np.ascontiguousarray(qm_all[o:o + n])
            err = O.av1o_block_error_qm(ptr(c, i32p), ptr(d, i32p), C.c_ssize_t(n), ptr(qm, u8p),
                                        ptr(scan, i16p), C.byref(ssz))
        exp[i] = (err, ssz.value)
        o += n
    # one launch per bit depth (bd is a launch parameter)
    for bd in sorted({b for _, b in cases}):
        sel = np.array([i for i, (_, b) in enumerate(cases) if b == bd])
        d_blk = b200av1.to_device(ctx, blocks[sel])
        d_c, d_d = b200av1.to_device(ctx, coeff), b200av1.to_device(ctx, dq)
        d_qm, d_scan = b200av1.to_device(ctx, qm_all), b200av1.to_device(ctx, scan_all)
        d_out = b200av1.DevBuf(ctx, sel.size * 16)
        l = b200av1.ErrLaunch()
        l.kind, l.bd, l.n_blocks = kind, bd, int(sel.size)
        l.blocks, l.coeff, l.dqcoeff, l.out = d_blk.ptr, d_c.ptr, d_d.ptr, d_out.ptr
        if kind == 3:
            l.qm, l.scan = d_qm.ptr, d_scan.ptr
        b200av1.check(b200av1.lib().b200_block_error_launch(ctx.h, C.byref(l), None))
        got = b200av1.to_host(ctx, d_out, np.int64, sel.size * 2).reshape(-1, 2)
        for j, i in enumerate(sel):
            want = exp[i].copy()
            if kind == 1:
                want[1] = got[j, 1]  # av1_block_error_lp has no ssz output
            assert tuple(got[j]) == tuple(want), (kind, H.TX_NAMES[cases[i][0]], bd)


def test_launch_argument_validation(ctx):
    """bad bd / kind / missing tables are refused with B200_EINVAL, not run"""
    import b200av1
    l = b200av1.ErrLaunch()
    l.kind, l.bd, l.n_blocks = 2, 9, 1
    d = b200av1.DevBuf(ctx, 64)
    l.blocks = l.coeff = l.dqcoeff = l.out = d.ptr
    assert b200av1.lib().b200_block_error_launch(ctx.h, C.byref(l), None) == -2
    l.kind, l.bd = 3, 8  # QM without qm / scan
    assert b200av1.lib().b200_block_error_launch(ctx.h, C.byref(l), None) == -2
