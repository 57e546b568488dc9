#!/usr/bin/env python3
"""Dev only: per-phase cycles of the full-pel search on the bench's configs[2] workload, from the
-DB200_ME_TIMING build (tools/me_timing_build.sh).  Run with B200AV1_LIB=<...>/libb200av1_timing.so."""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "aom-av1-lavish_b200", "python"))
import b200av1  # noqa: E402
import bench  # noqa: E402


def main():
    F = 4
    _, _, luma = bench.make_residual_frames(list(range(F)), 5)
    bd = 10 if "--hbd" in sys.argv else 8
    if bd > 8:  # the 10-bit searches of configs[4] (same per-search work on 1080p planes)
        luma = (luma.astype(np.uint16) << 2) | (np.random.default_rng(6).integers(0, 4, luma.shape).astype(np.uint16))
    work = bench.MeWork(luma, list(range(F)), 5, bd=bd)
    if "--tf" in sys.argv:
        # the temporal filter's searches (queue.cu b200_queue_tf_default_params): 16x16 (or 8x8 with --tf8) blocks on
        # a 16 / 8 pixel grid of frame 0, frame-wide mv limits, step_param 0, forced mesh
        n = 8 if "--tf8" in sys.argv else 16
        B = bench.ME_BORDER
        sites = [(bx, by) for by in range(0, work.h - n + 1, n) for bx in range(0, work.w - n + 1, n)]
        blk = np.zeros(len(sites), b200av1.MEBLOCK_DTYPE)
        for i, (bx, by) in enumerate(sites):
            blk[i]["src_offset"] = by * work.w + bx
            blk[i]["ref_offset"] = (by + B) * work.rstride + bx + B
            blk[i]["col_min"], blk[i]["col_max"] = max(-(bx + B - 8), -1000), min(work.w - bx - n + B - 8, 1000)
            blk[i]["row_min"], blk[i]["row_max"] = max(-(by + B - 8), -1000), min(work.h - by - n + B - 8, 1000)
        blk["bw"], blk["bh"] = n, n
        blk["search_method"], blk["step_param"] = 0, 0
        blk["flags"] = b200av1.ME_RUN_MESH
        blk["mv_cost_type"] = 3
        blk["sad_per_bit"], blk["error_per_bit"] = 16, 10
        work.blocks, work.n = blk, len(sites)
        work.tf_block = n
    if "--no-mesh" in sys.argv:
        work.blocks["flags"] = 0
        work.blocks["force_mesh_thresh"] = 2 ** 31 - 1
    ctx = b200av1.Context(0)
    d_cur, d_ref = b200av1.to_device(ctx, work.cur), b200av1.to_device(ctx, work.refs)
    d_blk = b200av1.to_device(ctx, work.blocks)
    d_res = b200av1.DevBuf(ctx, work.n * 40)
    l = work.launch_struct(d_blk.ptr, d_cur.ptr, d_ref.ptr, d_res.ptr, work.n, F * work.h, 2 * F * work.rh)
    if hasattr(work, "tf_block"):
        n = work.tf_block
        words = ((15 + (2 * 64 + n) + 3) // 4 + 4 + 11) & ~3
        l.window_bytes, l.max_bw, l.max_bh = (2 * 64 + n) * (words + 8) * 4, n, n
        l.max_col_span, l.max_row_span = 0, 0  # frame-wide limits: no TMA box
    for _ in range(2):
        b200av1.check(b200av1.lib().b200_me_fullpel_launch(ctx.h, C.byref(l), None))
    r = b200av1.to_host(ctx, d_res, b200av1.MERESULT_DTYPE, work.n)
    cl = r["cost_list"].astype(np.int64) * 16
    whole = r["sse"].astype(np.int64) * 16
    if "--t2" in sys.argv:  # -DB200_ME_TIMING2 build: inside the n-step phase
        rows = [("staging", cl[:, 4]), ("n-step phase", r["distortion"].astype(np.int64) * 16),
                ("  diamond_runs", cl[:, 2]), ("  replay: var_cost of the runs", cl[:, 1]), ("  sad_list", cl[:, 3]),
                ("whole search", whole)]
    elif "--t3" in sys.argv:  # -DB200_ME_TIMING3: eval / control split of diamond_runs, rounds
        rows = [("n-step phase", r["distortion"].astype(np.int64) * 16), ("  rounds: eval (to the barrier)", cl[:, 1]),
                ("  rounds: control (decide + publish)", cl[:, 3]), ("whole search", whole)]
        print("rounds/search %.2f" % (r["cost_list"][:, 2].mean()))
    else:
        rows = [("staging", cl[:, 4]), ("n-step (diamond runs + var + cost list)", r["distortion"].astype(np.int64) * 16),
                ("mesh stage 0", cl[:, 1]), ("mesh stage 1", cl[:, 2]), ("mesh stage 2 (interval 1)", cl[:, 3]),
                ("whole search", whole)]
    for name, v in rows:
        print("%-42s mean %9.0f cycles  %5.1f %%" % (name, v.mean(), 100.0 * v.mean() / whole.mean()))
    steps = r["cost_list"][:, 0]
    print("n-step rounds/search %.1f  candidates/search %.1f" % ((steps & 0xff).mean(), (steps >> 8).mean()))


if __name__ == "__main__":
    main()
