"""Shared generators for the motion-search tests (oracle pinning on CPU, CUDA parity on
GPU): synthetic frame pairs with a replicated border and random search parameters that
exercise every branch of av1_full_pixel_search / av1_find_best_sub_pixel_tree."""
import numpy as np

import harness as H
import synth

BORDER = 96
FULLPEL_SIZES = [(64, 64), (32, 32), (16, 16), (8, 8), (16, 8), (8, 16), (64, 32), (4, 4),
                 (128, 128), (32, 8)]
SUBPEL_SIZES = [(64, 64), (32, 32), (16, 16), (8, 8), (16, 8), (4, 4), (128, 64), (8, 32)]
MESH = ((64, 8), (28, 4), (15, 1), (7, 1))  # speed_features.c:25-33 (good quality)


def make_frames(rng, w, h, bd, flat=False):
    """current frame + bordered reference = shifted current + noise (ties when flat)"""
    cur = synth.smooth_frame(rng, w, h, bd)
    if flat:
        cur[: h // 2] = cur[0, 0]
    dy, dx = int(rng.integers(-9, 10)), int(rng.integers(-9, 10))
    ref = np.roll(cur, (dy, dx), (0, 1)).astype(np.int32) + \
        np.rint(rng.normal(0, 2, cur.shape)).astype(np.int32)
    ref = np.clip(ref, 0, (1 << bd) - 1).astype(cur.dtype)
    if flat:
        ref[: h // 2] = cur[0, 0]
    return np.ascontiguousarray(cur), H.bordered(ref, BORDER)


def base_params(cur, refb, bx, by, bw, bh, bd, tables):
    p = H.MeParams()
    esz = cur.itemsize
    stride_r = refb.shape[1]
    p.src = cur.ctypes.data + (by * cur.shape[1] + bx) * esz
    p.ref = refb.ctypes.data + ((by + BORDER) * stride_r + bx + BORDER) * esz
    p.src_stride, p.ref_stride = cur.shape[1], stride_r
    p.bw, p.bh, p.highbd, p.bd = bw, bh, int(bd > 8), bd
    p.mvjcost, p.mvcost0, p.mvcost1 = H.mv_cost_ptrs(*tables)
    p.sad_per_bit, p.error_per_bit = 40, 90
    for i, (r, it) in enumerate(MESH):
        p.mesh_range[i], p.mesh_interval[i] = r, it
    # bookkeeping for the CUDA tests (not part of the C struct)
    p._bx, p._by = bx, by
    return p


def random_fullpel_params(rng, cur, refb, bw, bh, bd, tables):
    fh, fw = cur.shape
    bx = int(rng.integers(0, (fw - bw) // 4 + 1)) * 4
    by = int(rng.integers(0, (fh - bh) // 4 + 1)) * 4
    p = base_params(cur, refb, bx, by, bw, bh, bd, tables)
    rngr = int(rng.choice([8, 24, 64]))
    # window around 0, clipped so that every candidate block stays inside the border
    p.col_min = max(-rngr, -(bx + BORDER - 8))
    p.col_max = min(rngr, fw - bw - bx + BORDER - 8)
    p.row_min = max(-rngr, -(by + BORDER - 8))
    p.row_max = min(rngr, fh - bh - by + BORDER - 8)
    p.search_method = int(rng.integers(0, 4))
    p.step_param = int(rng.integers(0, 9)) if p.search_method != 3 else int(rng.integers(0, 6))
    p.run_mesh_search = int(rng.integers(0, 2))
    p.prune_mesh_search = int(rng.integers(0, 2))
    p.mesh_search_mv_diff_threshold = int(rng.integers(0, 5))
    p.force_mesh_thresh = int(rng.choice([0, 1000, 1 << 20]))
    p.fine_search_interval = int(rng.integers(0, 2))
    p.use_skip_sad = int(rng.integers(0, 2)) if bh >= 8 else 0
    p.mv_cost_type = int(rng.integers(0, 5))
    p.ref_mv_row, p.ref_mv_col = int(rng.integers(-60, 61)), int(rng.integers(-60, 61))
    p.start_row = int(rng.integers(p.row_min - 3, p.row_max + 4))
    p.start_col = int(rng.integers(p.col_min - 3, p.col_max + 4))
    p.want_cost_list = int(rng.integers(0, 2))
    p.want_second_best = int(rng.integers(0, 2))
    return p


def random_subpel_params(rng, cur, refb, bw, bh, bd, tables, search_type):
    fh, fw = cur.shape
    bx = int(rng.integers(0, (fw - bw) // 4 + 1)) * 4
    by = int(rng.integers(0, (fh - bh) // 4 + 1)) * 4
    p = base_params(cur, refb, bx, by, bw, bh, bd, tables)
    p.subpel_search_type = search_type
    p.allow_hp = int(rng.integers(0, 2))
    p.forced_stop = int(rng.integers(0, 4))
    p.iters_per_step = int(rng.integers(1, 3))
    p.mv_cost_type = int(rng.integers(0, 5))
    p.ref_mv_row, p.ref_mv_col = int(rng.integers(-60, 61)), int(rng.integers(-60, 61))
    fr, fc = int(rng.integers(-12, 13)), int(rng.integers(-12, 13))
    p.sp_start_row, p.sp_start_col = fr * 8, fc * 8
    lim = int(rng.choice([8, 30, 200]))
    p.sp_row_min, p.sp_row_max = fr * 8 - lim, fr * 8 + lim
    p.sp_col_min, p.sp_col_max = fc * 8 - lim, fc * 8 + lim
    return p


def random_cost_list(rng, p):
    """fill p.sp_cost_list / sp_have_cost_list: absent, holding an INT_MAX entry, a
    well-behaved bowl (centre strictly smallest) or an arbitrary list"""
    mode = int(rng.integers(0, 5))
    p.sp_have_cost_list = int(mode != 0)
    vals = [int(v) for v in rng.integers(200, 6000, 5)]
    if mode == 2:
        vals[int(rng.integers(0, 5))] = 2 ** 31 - 1
    elif mode >= 3:
        vals[0] = min(vals) - int(rng.integers(1, 150))
    for i in range(5):
        p.sp_cost_list[i] = vals[i]
    return p


def add_compound(rng, p, bd, keep, masked):
    """attach a second_pred block (bw-pitch, the other reference's prediction) and, when `masked`, a 6-bit
    blend mask to a search descriptor; `keep` receives the arrays (they must outlive the call)"""
    dt = np.uint16 if bd > 8 else np.uint8
    second = rng.integers(0, 1 << bd, p.bw * p.bh).astype(dt)
    keep.append(second)
    p.second_pred = second.ctypes.data
    if masked:
        ms = p.bw + int(rng.integers(0, 3)) * 4
        mask = rng.integers(0, 65, p.bh * ms).astype(np.uint8)
        keep.append(mask)
        p.mask, p.mask_stride, p.inv_mask = mask.ctypes.data, ms, int(rng.integers(0, 2))
    return p
