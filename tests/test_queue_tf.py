"""Host side of the temporal-filter block queue (csrc/queue.cu): the descriptors
b200_queue_tf_describe builds for a frame carry the limits the reference's own
av1_set_mv_limits / av1_set_mv_search_range / av1_set_subpel_mv_search_range produce inside
tf_motion_search (temporal_filter.c:88-260, loop :804-872).  No device work."""
import ctypes as C

import numpy as np
import pytest

import harness as H

b200av1 = pytest.importorskip("b200av1")

ME_BLOCK = np.dtype([("src_offset", "<u4"), ("ref_offset", "<u4"), ("col_min", "<i2"), ("col_max", "<i2"),
                     ("row_min", "<i2"), ("row_max", "<i2"), ("start_row", "<i2"), ("start_col", "<i2"),
                     ("ref_mv_row", "<i2"), ("ref_mv_col", "<i2"), ("bw", "u1"), ("bh", "u1"),
                     ("search_method", "u1"), ("step_param", "u1"), ("flags", "u1"), ("mv_cost_type", "u1"),
                     ("mesh_mv_diff_threshold", "u1"), ("subpel", "u1"), ("force_mesh_thresh", "<i4"),
                     ("sad_per_bit", "<i4"), ("error_per_bit", "<i4")])
assert ME_BLOCK.itemsize == 44


def describe(width, height, pitch, origin, params, ref_mvs, stage):
    rows, cols = (height + params.bh - 1) // params.bh, (width + params.bw - 1) // params.bw
    n = rows * cols * (4 if stage else 1)
    full, sub = np.zeros(n, ME_BLOCK), np.zeros(n, ME_BLOCK)
    mv = None if ref_mvs is None else np.ascontiguousarray(ref_mvs, np.int16)
    got = b200av1.lib().b200_queue_tf_describe(width, height, pitch, origin, C.byref(params),
                                               b200av1.vp(mv) if mv is not None else None, stage,
                                               b200av1.vp(full), b200av1.vp(sub))
    assert got == n
    return full, sub, rows, cols


@pytest.mark.parametrize("width,height", [(352, 288), (1920, 1080), (100, 70)])
def test_tf_descriptor_limits_match_reference(width, height):
    R = H.ref()
    if R is None:
        pytest.skip("oracle/_ref not built")
    p = b200av1.tf_default_params(width, height)
    pitch = (width + 2 * 160 + 15) & ~15
    origin = 160 * pitch + 160
    for stage in (0, 1):
        full, sub, rows, cols = describe(width, height, pitch, origin, p, None, stage)
        per = 4 if stage else 1
        for r in range(rows):
            for c in range(cols):
                ref = (C.c_int * 8)()
                R.refshim_tf_limits(width, height, p.bw, p.bh, p.border_in_pixels, r, c, ref)
                for k in range(per):
                    f, s = full[(r * cols + c) * per + k], sub[(r * cols + c) * per + k]
                    assert [f["col_min"], f["col_max"], f["row_min"], f["row_max"]] == list(ref[:4])
                    assert [s["col_min"], s["col_max"], s["row_min"], s["row_max"]] == list(ref[4:])


def test_tf_descriptor_geometry_and_start():
    width, height = 352, 288
    p = b200av1.tf_default_params(width, height)
    assert (p.bw, p.bh) == (16, 16) and p.mv_cost_type == 1  # B200_MV_COST_L1_LOWRES: min(352, 288) < 480
    pitch, origin = 672, 160 * 672 + 160
    rows, cols = 18, 22
    rng = np.random.default_rng(5)
    mvs = rng.integers(-200, 200, (rows * cols, 2)).astype(np.int16)
    full, sub, _r, _c = describe(width, height, pitch, origin, p, mvs, 0)
    for i in (0, 7, rows * cols - 1):
        r, c = divmod(i, cols)
        assert full[i]["src_offset"] == origin + r * 16 * pitch + c * 16 == sub[i]["ref_offset"]
        # get_fullmv_from_mv (mv.h:79): GET_MV_RAWPEL rounds half away from zero-ish
        want = [(int(v) + 3 + (int(v) >= 0)) >> 3 for v in mvs[i]]
        assert [full[i]["start_row"], full[i]["start_col"]] == want
        assert full[i]["flags"] & 1 and sub[i]["mv_cost_type"] == 4 and sub[i]["subpel"] == (0 | 2 << 2 | 3 << 4)
    full1, sub1, _r, _c = describe(width, height, pitch, origin, p, mvs, 1)
    assert len(full1) == 4 * rows * cols
    i = 5 * cols + 3
    offs = [origin + (5 * 16 + (k >> 1) * 8) * pitch + 3 * 16 + (k & 1) * 8 for k in range(4)]
    assert list(full1[4 * i:4 * i + 4]["src_offset"]) == offs
    assert set(full1["bw"]) == {8} and set(sub1["bh"]) == {8}
