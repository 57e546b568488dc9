"""b200_txq_compact_launch_noalloc (include/b200av1.h): the packing of the compact result format with its
scratch in a file-scope device array instead of cudaMalloc'ed memory.  No counterpart in the reference; it
must produce exactly what b200_txq_compact_launch produces (integer work: bit-exact), on two calls in a
row (the static scratch keeps the first call's contents), and refuse batches beyond its stated maximum.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import harness as H

pytestmark = pytest.mark.gpu

NOALLOC_MAX_BLOCKS = 1 << 23  # B200_COMPACT_NOALLOC_MAX_BLOCKS


@pytest.fixture(scope="module")
def ctx():
    import b200av1
    c = b200av1.Context(0)
    yield c
    c.close()


def make_batch(rng, n, wide):
    """n blocks of mixed tx sizes / types with a dense qcoeff array and eob per block, as b200_txq_launch
    leaves them in device memory"""
    import b200av1
    blocks = np.zeros(n, b200av1.TXBLOCK_DTYPE)
    sizes = rng.integers(0, H.N_TX_SIZES, n)
    off = 0
    eob = np.zeros(n, np.uint16)
    for i, ts in enumerate(sizes):
        ts = int(ts)
        types = [t for t in range(16) if H.tx_valid(ts, t)]
        blocks[i]["tx_size"] = ts
        blocks[i]["tx_type"] = types[int(rng.integers(0, len(types)))]
        blocks[i]["out_offset"] = off
        me = H.max_eob(ts)
        eob[i] = 0 if rng.random() < 0.2 else int(rng.integers(1, me + 1))
        off += me
    lim = (1 << 20) if wide else (1 << 15)
    qcoeff = rng.integers(-lim, lim, off).astype(np.int32)
    return blocks, eob, qcoeff


def run(ctx, fn, blocks, eob, qcoeff, wide):
    import b200av1
    n = len(blocks)
    d_blk, d_eob, d_q = (b200av1.to_device(ctx, a) for a in (blocks, eob, qcoeff))
    lv_dt = np.int32 if wide else np.int16
    total = int(eob.astype(np.int64).sum())
    d_lv = b200av1.to_device(ctx, np.full(max(total, 1), -1, lv_dt))
    d_off = b200av1.to_device(ctx, np.full(n + 1, 0xffffffff, np.uint32))
    l = b200av1.CompactLaunch()
    l.n_blocks, l.blocks, l.qcoeff, l.eob = n, d_blk.ptr, d_q.ptr, d_eob.ptr
    l.wide, l.levels, l.level_offset = int(wide), d_lv.ptr, d_off.ptr
    b200av1.check(fn(ctx.h, C.byref(l), None))
    b200av1.check(b200av1.lib().b200_sync(ctx.h))
    return (b200av1.to_host(ctx, d_off, np.uint32, n + 1), b200av1.to_host(ctx, d_lv, lv_dt, max(total, 1)))


def test_noalloc_compaction_matches_the_allocating_entry_point_twice_in_a_row(ctx):
    import b200av1
    lib = b200av1.lib()
    rng = np.random.default_rng(9100)
    # second batch smaller than the first: the scratch still holds the first call's chunk sums beyond it
    for n, wide in ((9000, False), (2500, True)):
        blocks, eob, qcoeff = make_batch(rng, n, wide)
        want_off, want_lv = run(ctx, lib.b200_txq_compact_launch, blocks, eob, qcoeff, wide)
        got_off, got_lv = run(ctx, lib.b200_txq_compact_launch_noalloc, blocks, eob, qcoeff, wide)
        assert np.array_equal(got_off, want_off)
        assert np.array_equal(got_lv, want_lv)  # bit-exact: integer gather
        # and against the definition: exclusive prefix sum of eob, levels in scan order
        assert np.array_equal(got_off, np.concatenate([[0], np.cumsum(eob.astype(np.uint64))]).astype(np.uint32))
        for i in list(range(5)) + [n - 1]:
            scan, _ = H.oracle_scan(int(blocks[i]["tx_size"]), int(blocks[i]["tx_type"]))
            e, o = int(eob[i]), int(blocks[i]["out_offset"])
            assert np.array_equal(got_lv[got_off[i]:got_off[i] + e].astype(np.int64),
                                  qcoeff[o + np.asarray(scan[:e], np.int64)].astype(got_lv.dtype).astype(np.int64))


def test_noalloc_compaction_rejects_batches_beyond_its_maximum_and_handles_empty_ones(ctx):
    import b200av1
    lib = b200av1.lib()
    d = b200av1.to_device(ctx, np.zeros(64, np.uint32))
    l = b200av1.CompactLaunch()
    l.blocks = l.qcoeff = l.eob = l.levels = l.level_offset = d.ptr
    l.n_blocks = NOALLOC_MAX_BLOCKS + 1
    assert lib.b200_txq_compact_launch_noalloc(ctx.h, C.byref(l), None) != 0
    lib.b200_last_error.restype = C.c_char_p
    assert b"maximum" in lib.b200_last_error()
    l.n_blocks = 0
    b200av1.check(lib.b200_txq_compact_launch_noalloc(ctx.h, C.byref(l), None))
    b200av1.check(lib.b200_sync(ctx.h))
    assert b200av1.to_host(ctx, d, np.uint32, 1)[0] == 0
