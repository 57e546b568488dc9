#!/usr/bin/env python3
"""bench.py -- headline benchmark of the B200 AV1 hot path (BASELINE.json configs[1]).

Workload ("txq_sweep_1080p_8bit"): a batch of F synthetic 1080p 8-bit 4:2:0 frames
(int16 residual planes = source - prediction, seeded) goes through the fused
forward-transform + quantize_fp kernel once for EACH of the 19 AV1 transform sizes
(every plane tiled with WxH blocks, tx_type drawn per block from the valid set, qindex
drawn per block from {0,32,96,160,255}).  One step = 19 launches = 19*F frame
equivalents (FE, 3 110 400 samples each).  Inputs are resident in HBM for `value`;
`e2e` runs the same sweep through the host-buffer C-ABI call (H2D + D2H inside).

Contract: python bench.py --gpus N --steps K --warmup W   (torchrun for N > 1)
prints ONE JSON line on rank 0.  `--impl reference` times the reference's own C
functions (oracle/_ref) on the host cores instead.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "aom-av1-lavish_b200", "python"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

FE_SAMPLES = 1920 * 1080 * 3 // 2
TX_W = [4, 8, 16, 32, 64, 4, 8, 8, 16, 16, 32, 32, 64, 4, 16, 8, 32, 16, 64]
TX_H = [4, 8, 16, 32, 64, 8, 4, 16, 8, 32, 16, 64, 32, 16, 4, 32, 8, 64, 16]
VTX = [0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3]
HTX = [0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2]
QINDEXES = [0, 32, 96, 160, 255]

TXBLOCK_DTYPE = np.dtype([("src_offset", "<u4"), ("out_offset", "<u4"), ("src_stride", "<u2"),
                          ("qrow", "<u2"), ("tx_size", "u1"), ("tx_type", "u1"),
                          ("reserved", "<u2")])


def valid_types(ts):
    out = []
    for tt in range(16):
        ok = True
        for dim, kind in ((TX_H[ts], VTX[tt]), (TX_W[ts], HTX[tt])):
            if dim == 64 and kind != 0:
                ok = False
            if dim == 32 and kind in (1, 2):
                ok = False
        if ok:
            out.append(tt)
    return out


def max_eob(ts):
    return min(TX_W[ts], 32) * min(TX_H[ts], 32)


def make_residual_frames(frame_ids, seed):
    """int16 residual of 1080p 4:2:0 frames: planes Y 1920x1080, U, V 960x540 stored back to
    back (strides = widths).  Smooth source panned per frame, prediction = shifted source +
    noise.  A frame's content depends only on (seed, frame id), never on the rank owning it."""
    import synth
    if isinstance(frame_ids, int):
        frame_ids = list(range(frame_ids))
    planes = [(1920, 1080), (960, 540), (960, 540)]
    frame_elems = sum(w * h for w, h in planes)
    out = np.zeros((len(frame_ids), frame_elems), np.int16)
    rng0 = np.random.default_rng(seed)
    base = [synth.smooth_frame(rng0, w, h, 8) for (w, h) in planes]
    for i, f in enumerate(frame_ids):
        rng = np.random.default_rng([seed, f])
        o = 0
        for (w, h), s in zip(planes, base):
            src = np.roll(s, (f * 2, f * 3), (0, 1))
            pred = synth.pred_from(rng, src, 8)
            out[i, o:o + w * h] = (src.astype(np.int16) - pred.astype(np.int16)).reshape(-1)
            o += w * h
    return out, planes


def make_descriptors(ts, n_frames, planes, frame_elems, seed):
    """all blocks of size ts tiling every plane of every frame, in batch-queue order"""
    rng = np.random.default_rng(seed * 100 + ts)
    W, Hh = TX_W[ts], TX_H[ts]
    offs, strides, planes_id = [], [], []
    o = 0
    for pi, (w, h) in enumerate(planes):
        xs = np.arange(w // W) * W
        ys = np.arange(h // Hh) * Hh
        po = (o + ys[:, None] * w + xs[None, :]).reshape(-1)
        offs.append(po)
        strides.append(np.full(po.size, w))
        planes_id.append(np.full(po.size, pi))
        o += w * h
    offs = np.concatenate(offs)
    strides = np.concatenate(strides)
    planes_id = np.concatenate(planes_id)
    per_frame = offs.size
    n = per_frame * n_frames
    blocks = np.zeros(n, TXBLOCK_DTYPE)
    blocks["src_offset"] = (offs[None, :] + (np.arange(n_frames) * frame_elems)[:, None]).reshape(-1)
    blocks["src_stride"] = np.tile(strides, n_frames)
    vt = np.array(valid_types(ts), np.uint8)
    blocks["tx_type"] = vt[rng.integers(0, vt.size, n)]
    blocks["tx_size"] = ts
    # quantiser row = plane * 5 + qindex slot
    blocks["qrow"] = np.tile(planes_id, n_frames) * len(QINDEXES) + rng.integers(0, len(QINDEXES), n)
    # batch-queue order: within each frame group the blocks by tx_type (warps stay on one
    # code path) but keep frames apart, so that the 16 type groups of a frame re-touch the
    # same ~6 MB of residual while it is still in L2
    frame_id = np.repeat(np.arange(n_frames), per_frame)
    order = np.lexsort((blocks["tx_type"], frame_id))
    blocks = blocks[order]
    blocks["out_offset"] = np.arange(n, dtype=np.uint64) * max_eob(ts)
    samples = n * W * Hh
    return blocks, samples


def algorithmic_bytes(ts, n_blocks):
    """int16 in + coeff/qcoeff/dqcoeff int32 out (kept region) + eob"""
    return n_blocks * (TX_W[ts] * TX_H[ts] * 2 + max_eob(ts) * 12 + 2)


class ClockSampler:
    """nvidia-smi clocks during the timed region (B200_PROFILING.md clocks line)"""

    def __init__(self, gpu):
        self.gpu = gpu
        self.proc = None
        self.lines = []

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thr = threading.Thread(target=self._read, daemon=True)
            self.thr.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0]))
                smax = float(p[1])
            except ValueError:
                continue
            for nm, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------- CPU arm ----
def load_cpu_libs():
    import harness as H
    O = H.oracle()
    R = H.ref()
    return H, O, R


def cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes=range(19),
                  keep_outputs=False):
    """One frame per size through the CPU path.  Returns (seconds, samples, outputs)."""
    rows = (H.QRow * (3 * len(QINDEXES)))()
    for pl in range(3):
        for qi, q in enumerate(QINDEXES):
            O.av1o_build_qrow(8, q, pl, 0, C.byref(rows[pl * len(QINDEXES) + qi]))
    total_t, total_samples, outs = 0.0, 0, {}
    src = np.ascontiguousarray(frames[:1].reshape(-1))
    for ts in sizes:
        blocks, samples = make_descriptors(ts, 1, planes, frame_elems, seed)
        n = blocks.size
        ne = n * max_eob(ts)
        coeff = np.zeros(ne, np.int32)
        qc = np.zeros(ne, np.int32)
        dq = np.zeros(ne, np.int32)
        eob = np.zeros(n, np.uint16)
        if R is not None:
            fwd = C.cast(getattr(R, "av1_fwd_txfm2d_%dx%d_c" % (TX_W[ts], TX_H[ts])), C.c_void_p)
            ls = H.log_scale(ts)
            qn = ["av1_quantize_fp_c", "av1_quantize_fp_32x32_c", "av1_quantize_fp_64x64_c"][ls]
            quant = C.cast(getattr(R, qn), C.c_void_p)
        else:
            fwd, quant = None, None
        t0 = time.perf_counter()
        O.av1o_drv_txq_fp(fwd, quant, 0, 8, ts, C.c_void_p(blocks.ctypes.data), n,
                          C.c_void_p(src.ctypes.data), rows, C.c_void_p(coeff.ctypes.data),
                          C.c_void_p(qc.ctypes.data), C.c_void_p(dq.ctypes.data),
                          C.c_void_p(eob.ctypes.data), n_threads)
        total_t += time.perf_counter() - t0
        total_samples += samples
        if keep_outputs:
            outs[ts] = (blocks, coeff, qc, dq, eob)
    return total_t, total_samples, outs


# ------------------------------------------------------------------ ME workload ----
ME_BORDER = 160  # AOM_ENC_NO_SCALE_BORDER, aom_scale/yv12config.h:28-32


def bench_me(ctx, dev, stream, steps, warmup, seed, frame_id, n_frames=8):
    """BASELINE configs[2]: full-pel (NSTEP + mesh {64,8},{28,4},{15,1}) and sub-pel (8-tap
    tree) search of every 64x64 superblock of `n_frames` 1080p 8-bit frames against 2
    references each, +-64 range, L1 (HD) mv cost.  All frames and references live in one
    plane pool (stacked vertically), so ONE launch per kernel covers the whole batch.
    Returns device-timed ms per frame for both kernels."""
    import torch
    import b200av1
    import synth
    rng = np.random.default_rng([seed, frame_id, 77])
    w, h = 1920, 1080
    base = synth.smooth_frame(rng, w, h, 8)
    rstride, rh = w + 2 * ME_BORDER, h + 2 * ME_BORDER
    curs = np.zeros((n_frames, h, w), np.uint8)
    refs = np.zeros((n_frames * 2, rh, rstride), np.uint8)
    for f in range(n_frames):
        curs[f] = np.roll(base, (f * 2, f * 3), (0, 1))
        for r in range(2):
            dy, dx = int(rng.integers(-20, 21)), int(rng.integers(-20, 21))
            ref = np.roll(curs[f], (dy, dx), (0, 1)).astype(np.int16) + \
                np.rint(rng.normal(0, 2, (h, w))).astype(np.int16)
            refs[f * 2 + r] = np.pad(np.clip(ref, 0, 255).astype(np.uint8), ME_BORDER, mode="edge")
    sbs = [(bx, by) for by in range(0, h - 63, 64) for bx in range(0, w - 63, 64)]
    n_sb = len(sbs)
    n = n_sb * 2 * n_frames
    blocks = np.zeros(n, b200av1.MEBLOCK_DTYPE)
    i = 0
    for f in range(n_frames):
        for r in range(2):
            for (bx, by) in sbs:
                b = blocks[i]
                b["src_offset"] = (f * h + by) * w + bx
                b["ref_offset"] = ((f * 2 + r) * rh + by + ME_BORDER) * rstride + bx + ME_BORDER
                i += 1
    blocks["col_min"], blocks["col_max"], blocks["row_min"], blocks["row_max"] = -64, 64, -64, 64
    blocks["bw"], blocks["bh"] = 64, 64
    blocks["search_method"] = 0  # B200_ME_NSTEP
    blocks["flags"] = b200av1.ME_RUN_MESH
    blocks["mv_cost_type"] = 3  # MV_COST_L1_HDRES
    blocks["sad_per_bit"], blocks["error_per_bit"] = 40, 90
    blocks["step_param"] = b200av1.lib().b200_me_init_search_range(64)
    d_cur = torch.from_numpy(curs.reshape(-1)).to(dev)
    d_ref = torch.from_numpy(refs.reshape(-1)).to(dev)
    d_blk = torch.from_numpy(blocks.view(np.uint8)).to(dev)
    d_res = torch.zeros(n * 40, dtype=torch.uint8, device=dev)

    def launch(kind, d_blocks):
        l = b200av1.MeLaunch()
        l.highbd, l.bd, l.n_blocks = 0, 8, n
        l.blocks, l.cur, l.ref = d_blocks.data_ptr(), d_cur.data_ptr(), d_ref.data_ptr()
        l.cur_stride, l.ref_stride = w, rstride
        for i, (r, it) in enumerate(((64, 8), (28, 4), (15, 1), (7, 1))):
            l.mesh_range[i], l.mesh_interval[i] = r, it
        l.results = d_res.data_ptr()
        l.window_bytes, l.max_bw, l.max_bh = 48 * 1024, 64, 64
        l.families = b200av1.ME_FAMILY_NSTEP
        l.cur_rows, l.ref_rows = n_frames * h, n_frames * 2 * rh
        l.max_col_span, l.max_row_span = 128, 128
        fn = b200av1.lib().b200_me_fullpel_launch if kind == 0 else b200av1.lib().b200_me_subpel_launch
        b200av1.check(fn(ctx.h, C.byref(l), C.c_void_p(stream.cuda_stream)))

    # full-pel once (untimed) to build the sub-pel descriptors from its results
    launch(0, d_blk)
    torch.cuda.synchronize(dev)
    res = d_res.cpu().numpy().view(b200av1.MERESULT_DTYPE)
    if os.environ.get("B200_ME_DEBUG"):  # library built with -DB200_ME_TIMING (dev only)
        print("ME timing (cycles): staging %.0f diamond %.0f mesh %.0f / %.0f / %.0f total %.0f" % (
            16 * res["cost_list"][:, 4].mean(), 16.0 * res["distortion"].mean(),
            16.0 * res["cost_list"][:, 1].mean(), 16.0 * res["cost_list"][:, 2].mean(),
            16.0 * res["cost_list"][:, 3].mean(), 16.0 * res["sse"].mean()), file=sys.stderr)
        print("ME diamond: steps/search %.1f candidates/search %.1f" % (
            (res["cost_list"][:, 0] & 255).mean(), (res["cost_list"][:, 0] >> 8).mean()), file=sys.stderr)
    sb = blocks.copy()
    sb["start_row"], sb["start_col"] = res["best_row"] * 8, res["best_col"] * 8
    sb["col_min"], sb["col_max"], sb["row_min"], sb["row_max"] = -512, 512, -512, 512
    sb["flags"] = b200av1.ME_ALLOW_HP
    sb["subpel"] = 0 | (2 << 2) | (3 << 4)  # EIGHTH_PEL, 2 iters/step, USE_8_TAPS
    d_sub = torch.from_numpy(sb.view(np.uint8)).to(dev)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_full = t_sub = 0.0
    for it in range(warmup + steps):
        ev[0].record(stream)
        launch(0, d_blk)
        ev[1].record(stream)
        launch(1, d_sub)
        ev[2].record(stream)
        torch.cuda.synchronize(dev)
        if it >= warmup:
            t_full += ev[0].elapsed_time(ev[1])
            t_sub += ev[1].elapsed_time(ev[2])
    t_full /= steps * n_frames  # ms per frame
    t_sub /= steps * n_frames
    # mesh candidates per search (lower bound of the SAD work; diamond steps come on top):
    # 17x17 @8, 15x15 @4, 31 rows x 30 cols @1 (the reference's tail-loop quirk)
    cands = 17 * 17 + 15 * 15 + 31 * 30
    per_frame = 2 * n_sb
    return {"workload": "config[2]: %d frames x 1080p 8-bit, %d 64x64 SBs x 2 refs per frame, +-64, "
                        "NSTEP+mesh, 8-tap sub-pel; one launch per kernel" % (n_frames, n_sb),
            "fullpel_ms_per_frame": t_full, "subpel_ms_per_frame": t_sub,
            "frames_per_s": 1e3 / (t_full + t_sub),
            "mesh_gabsdiff_per_s": per_frame * cands * 4096 / (t_full * 1e-3) / 1e9,
            "unique_bytes_per_frame": per_frame * (4096 + 192 * 192 + 64),
            "unique_GBps": per_frame * (4096 + 192 * 192 + 64) / (t_full * 1e-3) / 1e9,
            "launches_per_step": 2, "frames_per_launch": n_frames}



def bench_hbd4k(ctx, dev, stream, steps, warmup, seed, n_frames=3):
    """BASELINE configs[3]: 10-bit fused subtract + fwd txfm + av1_highbd_quantize_fp +
    av1_highbd_block_error over all 19 TX sizes of `n_frames` synthetic 4K 4:2:0 frames
    (uint16 source and prediction planes resident in HBM, 150 MB > L2)."""
    import torch
    import b200av1
    planes = [(3840, 2160), (1920, 1080), (1920, 1080)]
    frame_elems = sum(w * h for w, h in planes)
    rng = np.random.default_rng([seed, 4001])
    src = np.zeros((n_frames, frame_elems), np.uint16)
    pred = np.zeros((n_frames, frame_elems), np.uint16)
    for f in range(n_frames):
        o = 0
        for (w, h) in planes:
            low = rng.integers(100, 900, (h // 8 + 1, w // 8 + 1))
            img = np.kron(low, np.ones((8, 8), np.int64))[:h, :w] + rng.integers(-8, 9, (h, w))
            pr = np.roll(img, (1, 1), (0, 1)) + rng.integers(-12, 13, (h, w))
            src[f, o:o + w * h] = np.clip(img, 0, 1023).reshape(-1)
            pred[f, o:o + w * h] = np.clip(pr, 0, 1023).reshape(-1)
            o += w * h
    d_src = torch.from_numpy(src.reshape(-1).view(np.int16)).to(dev)
    d_pred = torch.from_numpy(pred.reshape(-1).view(np.int16)).to(dev)
    qrows = build_qrows(10)
    d_qrows = torch.from_numpy(qrows.view(np.uint8)).to(dev)
    n_out = n_frames * frame_elems
    d_coeff, d_qc, d_dq = (torch.empty(n_out, dtype=torch.int32, device=dev) for _ in range(3))
    launches, keep, tot_bytes, tot_samples = [], [], 0, 0
    for ts in range(19):
        blocks, samples = make_descriptors(ts, n_frames, planes, frame_elems, seed + 4)
        preds = np.zeros(blocks.size, b200av1.TXPRED_DTYPE)
        preds["pred_offset"], preds["pred_stride"] = blocks["src_offset"], blocks["src_stride"]
        d_blocks = torch.from_numpy(blocks.view(np.uint8)).to(dev)
        d_preds = torch.from_numpy(preds.view(np.uint8)).to(dev)
        d_eob = torch.zeros(blocks.size, dtype=torch.int16, device=dev)
        d_dist = torch.zeros(blocks.size * 2, dtype=torch.int64, device=dev)
        keep += [d_blocks, d_preds, d_eob, d_dist]
        l = b200av1.TxqLaunch()
        l.tx_size, l.in_kind, l.q_kind, l.bd = ts, b200av1.IN_PIXELS16, b200av1.Q_FP_HBD, 10
        l.flags = b200av1.TXQ_STORE_COEFF | b200av1.TXQ_BLOCK_ERROR
        l.n_blocks = int(blocks.size)
        l.blocks, l.preds = d_blocks.data_ptr(), d_preds.data_ptr()
        l.src, l.pred, l.qrows = d_src.data_ptr(), d_pred.data_ptr(), d_qrows.data_ptr()
        l.coeff, l.qcoeff, l.dqcoeff = d_coeff.data_ptr(), d_qc.data_ptr(), d_dq.data_ptr()
        l.eob, l.dist = d_eob.data_ptr(), d_dist.data_ptr()
        launches.append(l)
        # 2 + 2 B/sample in, 12 B/kept coeff out, 2 B eob + 16 B dist per block
        tot_bytes += blocks.size * (TX_W[ts] * TX_H[ts] * 4 + max_eob(ts) * 12 + 18)
        tot_samples += samples
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ms = 0.0
    for it in range(warmup + steps):
        ev[0].record(stream)
        for l in launches:
            ctx.txq_launch(l, stream.cuda_stream)
        ev[1].record(stream)
        torch.cuda.synchronize(dev)
        if it >= warmup:
            ms += ev[0].elapsed_time(ev[1])
    ms /= steps
    fe4k = 3840 * 2160 * 3 // 2
    return {"workload": "config[3]: %d synthetic 4K 10-bit 4:2:0 frames, fused subtract + txfm + "
                        "av1_highbd_quantize_fp + av1_highbd_block_error, all 19 TX sizes" % n_frames,
            "ms_per_sweep": ms, "fe4k_per_s": tot_samples / fe4k / (ms * 1e-3),
            "fe1080p_per_s": tot_samples / FE_SAMPLES / (ms * 1e-3),
            "GBps": tot_bytes / (ms * 1e-3) / 1e9, "bytes_per_sweep": int(tot_bytes),
            "launches_per_step": 19}


def bench_hadamard(ctx, dev, stream, steps, warmup, d_src, n_frames):
    """Hadamard + SATD over the luma residual of the resident 1080p frames (the TPL /
    tx-prune cost, aom_dsp/avg.c): coefficient + SATD output, and SATD only."""
    import torch
    import b200av1
    w, h = 1920, 1080
    frame_elems = FE_SAMPLES
    out = {}
    for n in (8, 16, 32):
        xs, ys = np.arange(w // n) * n, np.arange(h // n) * n
        offs = (ys[:, None] * w + xs[None, :]).reshape(-1)
        nb = offs.size * n_frames
        blocks = np.zeros(nb, b200av1.HADBLOCK_DTYPE)
        blocks["src_offset"] = (offs[None, :] + (np.arange(n_frames) * frame_elems)[:, None]).reshape(-1)
        blocks["out_offset"] = np.arange(nb, dtype=np.uint64) * n * n
        blocks["stride"] = w
        d_blk = torch.from_numpy(blocks.view(np.uint8)).to(dev)
        d_coeff = torch.empty(nb * n * n, dtype=torch.int32, device=dev)
        d_satd = torch.empty(nb, dtype=torch.int32, device=dev)
        for mode in ("coeff+satd", "satd"):
            l = b200av1.HadamardLaunch()
            l.n, l.kind, l.n_blocks = n, b200av1.HAD_LOWBD, nb
            l.blocks, l.src = d_blk.data_ptr(), d_src.data_ptr()
            l.coeff = d_coeff.data_ptr() if mode != "satd" else None
            l.satd = d_satd.data_ptr()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
            ms = 0.0
            for it in range(warmup + steps):
                ev[0].record(stream)
                b200av1.check(b200av1.lib().b200_hadamard_launch(ctx.h, C.byref(l), C.c_void_p(stream.cuda_stream)))
                ev[1].record(stream)
                torch.cuda.synchronize(dev)
                if it >= warmup:
                    ms += ev[0].elapsed_time(ev[1])
            ms /= steps
            nbytes = nb * (n * n * (6 if mode != "satd" else 2) + 4)
            out["%dx%d %s" % (n, n, mode)] = {"ms": round(ms, 4), "GBps": round(nbytes / (ms * 1e-3) / 1e9, 1),
                                              "luma_frames_per_s": round(n_frames / (ms * 1e-3), 1)}
    return out


def build_qrows_product():
    """3 planes x 5 qindexes, all delta_q = 0 (av1_build_quantizer, av1_quantize.c:602)"""
    return build_qrows(8)


def build_qrows(bd):
    """the product's own b200_build_quantizer_row (no oracle on the b200 arm)"""
    import b200av1
    rows = np.zeros(3 * len(QINDEXES), b200av1.QROW_DTYPE)
    row = b200av1.QRow()
    for pl in range(3):
        for j, q in enumerate(QINDEXES):
            b200av1.check(b200av1.lib().b200_build_quantizer_row(bd, q, 0, 0, 0, C.byref(row)))
            rows[pl * len(QINDEXES) + j] = np.frombuffer(bytes(row), b200av1.QROW_DTYPE)[0]
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--frames", type=int, default=32, help="frames per rank per launch")
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--force-exact", action="store_true", help="always take the 64-bit half_btf path")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-me", action="store_true", help="skip the motion-search section")
    ap.add_argument("--no-extras", action="store_true", help="skip the config[3] 4K 10-bit and Hadamard sections")
    ap.add_argument("--sizes", default="", help="comma list of tx sizes (debug)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    sizes = [int(s) for s in args.sizes.split(",")] if args.sizes else list(range(19))
    seed = 2001
    metric = "1080p frame-equiv/s for txfm+quant+ME SAD; HBM GB/s as % of B200 peak"
    config = {"workload": "txq_sweep_1080p_8bit: fused fwd txfm + quantize_fp over all %d TX sizes, "
                          "%d synthetic 1080p 4:2:0 frames per GPU per launch" % (len(sizes), args.frames),
              "frames_per_gpu": args.frames, "tx_sizes": len(sizes), "quantizer": "av1_quantize_fp",
              "qindex_set": QINDEXES, "l2": "inputs (%.0f MB/GPU) larger than L2; outputs streamed"
                                            % (args.frames * FE_SAMPLES * 2 / 1e6),
              "parallelism": "frames sharded over %d GPU(s), no collective" % world,
              "butterflies": "exact64 (forced)" if args.force_exact else "guarded (32-bit when |residual|<=4094, proved exact; else 64-bit)"}

    # ----------------------------------------------------------- reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        H, O, R = load_cpu_libs()
        frames, planes = make_residual_frames(1, seed)
        frame_elems = frames.shape[1]
        n_threads = os.cpu_count() or 1
        for _ in range(min(args.warmup, 1)):
            cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes)
        tt, ss = 0.0, 0
        for _ in range(args.steps):
            t, s, _o = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes)
            tt += t
            ss += s
        val = ss / FE_SAMPLES / tt
        kind = "reference" if R is not None else "port"
        line = {"impl": "reference", "metric": metric, "value": val, "unit": "1080p frame-equiv/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": tt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config,
                "cpu_baseline": {"value": val, "unit": "1080p frame-equiv/s", "cores": n_threads,
                                 "kind": kind,
                                 "sample": "per step: 1 frame x %d TX sizes (%.1f FE) through "
                                           "av1_fwd_txfm2d_*_c + av1_quantize_fp*_c"
                                           % (len(sizes), ss / args.steps / FE_SAMPLES)},
                "e2e": {"value": val, "unit": "1080p frame-equiv/s", "h2d_bytes_per_step": 0,
                        "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        return 0

    # ---------------------------------------------------------------- b200 arm
    import torch
    import b200av1
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    ctx = b200av1.Context(local_rank)
    stream = torch.cuda.Stream(device=dev)

    from b200av1 import shard
    frame_ids = shard.frames_of_rank(args.frames * world, rank, world)
    frames, planes = make_residual_frames(frame_ids, seed)
    frame_elems = frames.shape[1]
    d_src = torch.from_numpy(frames.reshape(-1)).to(dev)
    qrows = build_qrows_product()
    d_qrows = torch.from_numpy(qrows.view(np.uint8)).to(dev)
    launches, keep = [], []
    max_out = 0
    total_samples, total_bytes = 0, 0
    per_size = []
    for ts in sizes:
        blocks, samples = make_descriptors(ts, args.frames, planes, frame_elems, seed + rank)
        max_out = max(max_out, blocks.size * max_eob(ts))
        d_blocks = torch.from_numpy(blocks.view(np.uint8)).to(dev)
        d_eob = torch.zeros(blocks.size, dtype=torch.int16, device=dev)
        keep += [d_blocks, d_eob]
        per_size.append({"ts": ts, "n_blocks": int(blocks.size), "samples": int(samples),
                         "bytes": int(algorithmic_bytes(ts, blocks.size)), "d_blocks": d_blocks,
                         "d_eob": d_eob})
        total_samples += samples
        total_bytes += algorithmic_bytes(ts, blocks.size)
    d_coeff = torch.empty(max_out, dtype=torch.int32, device=dev)
    d_qc = torch.empty(max_out, dtype=torch.int32, device=dev)
    d_dq = torch.empty(max_out, dtype=torch.int32, device=dev)
    flags = b200av1.TXQ_STORE_COEFF | (b200av1.TXQ_FORCE_EXACT if args.force_exact else 0)
    for ps in per_size:
        l = b200av1.TxqLaunch()
        l.tx_size, l.in_kind, l.q_kind, l.bd, l.flags = ps["ts"], 0, b200av1.Q_FP, 8, flags
        l.n_blocks = ps["n_blocks"]
        l.blocks = ps["d_blocks"].data_ptr()
        l.src = d_src.data_ptr()
        l.qrows = d_qrows.data_ptr()
        l.coeff, l.qcoeff, l.dqcoeff = d_coeff.data_ptr(), d_qc.data_ptr(), d_dq.data_ptr()
        l.eob = ps["d_eob"].data_ptr()
        launches.append(l)
    sptr = stream.cuda_stream

    def step(events=None):
        for i, l in enumerate(launches):
            if events is not None:
                events[i].record(stream)
            ctx.txq_launch(l, sptr)
        if events is not None:
            events[len(launches)].record(stream)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
    barrier()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(len(launches) + 1)]
          for _ in range(args.steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    for k in range(args.steps):
        step(ev[k])
    barrier()
    clocks = sampler.stop()
    step_ms = [ev[k][0].elapsed_time(ev[k][len(launches)]) for k in range(args.steps)]
    total_ms = ev[0][0].elapsed_time(ev[args.steps - 1][len(launches)])
    kern_ms = np.zeros(len(launches))
    for k in range(args.steps):
        for i in range(len(launches)):
            kern_ms[i] += ev[k][i].elapsed_time(ev[k][i + 1])
    kern_ms /= args.steps
    t_local = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_local, op=dist.ReduceOp.MAX)
    total_ms_max = float(t_local.item())
    fe_per_step_rank = total_samples / FE_SAMPLES
    value = fe_per_step_rank * world * args.steps / (total_ms_max / 1e3)

    # ---- e2e: the same sweep through the host-buffer C-ABI call (1 frame / step)
    e2e = None
    if args.e2e_steps > 0:
        eb, e_samples = [], 0
        out_off = 0
        for ts in sizes:
            b, s = make_descriptors(ts, 1, planes, frame_elems, seed + rank)
            b["out_offset"] += out_off
            out_off += b.size * max_eob(ts)
            eb.append(b)
            e_samples += s
        eblocks = np.concatenate(eb)
        lib = b200av1.lib()

        def pinned(nbytes, dtype):
            p = C.c_void_p()
            b200av1.check(lib.b200_malloc_host(C.c_size_t(nbytes), C.byref(p)))
            n = nbytes // np.dtype(dtype).itemsize
            return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,)).view(dtype)[:n], p

        h_src, p1 = pinned(frame_elems * 2, np.int16)
        h_src[:] = frames[0]
        h_coeff, p2 = pinned(out_off * 4, np.int32)
        h_qc, p3 = pinned(out_off * 4, np.int32)
        h_dq, p4 = pinned(out_off * 4, np.int32)
        h_eob = np.zeros(eblocks.size, np.uint16)
        h_blk, p5 = pinned(eblocks.nbytes, np.uint8)  # the descriptor queue is pinned too
        h_blk[:] = eblocks.view(np.uint8)
        eblocks = h_blk.view(TXBLOCK_DTYPE)

        def e2e_step():
            ctx.txq_batch_host(0, b200av1.Q_FP, 8, flags, eblocks, None, h_src, None, qrows,
                               h_coeff, h_qc, h_dq, h_eob, None)

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.e2e_steps):
            e2e_step()
        torch.cuda.synchronize(dev)
        dt = time.perf_counter() - t0
        t_e = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_e, op=dist.ReduceOp.MAX)
        e2e = {"value": e_samples / FE_SAMPLES * world * args.e2e_steps / float(t_e.item()),
               "unit": "1080p frame-equiv/s",
               "h2d_bytes_per_step": int(frame_elems * 2 + eblocks.nbytes + qrows.nbytes),
               "d2h_bytes_per_step": int(out_off * 12 + eblocks.size * 2),
               "note": "b200_txq_batch_host: pinned host buffers, 1 frame x %d sizes per step; "
                       "descriptor bucketing + H2D + %d launches + D2H (per tx-size group, overlapped "
                       "on a second stream) inside the timed region" % (len(sizes), len(sizes))}
        for p in (p1, p2, p3, p4):
            pass  # pinned buffers live until process exit

    me = None
    if not args.no_me:
        me = bench_me(ctx, dev, stream, max(2, args.steps // 2), 2, seed, frame_ids[0])
        t_me = torch.tensor([me["fullpel_ms_per_frame"] + me["subpel_ms_per_frame"]],
                            dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t_me, op=dist.ReduceOp.MAX)
        me["frames_per_s_all_gpus"] = world * 1e3 / float(t_me.item())

    extras = None
    if not args.no_extras and world == 1:
        extras = {"hadamard": bench_hadamard(ctx, dev, stream, max(2, args.steps // 2), 2, d_src, args.frames),
                  "hbd4k": bench_hbd4k(ctx, dev, stream, max(2, args.steps // 2), 2, seed)}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = total_bytes / (kern_ms.sum() / 1e3) / 1e9
    kernels = []
    for ps, ms in zip(per_size, kern_ms):
        kernels.append({"tx": "%dx%d" % (TX_W[ps["ts"]], TX_H[ps["ts"]]), "ms": round(float(ms), 4),
                        "GBps": round(ps["bytes"] / (ms / 1e3) / 1e9, 1),
                        "frac": round(ps["bytes"] / (ms / 1e3) / 1e9 / peak, 3)})
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            tj = json.load(open(tp))
            # ncu capture was taken at tj["frames_per_launch"] frames; DRAM bytes scale with frames
            traffic = tj["txq_kernel_dram_bytes_per_step_at_8_frames"] * args.frames / tj["frames_per_launch"]
        except (OSError, ValueError, KeyError):
            traffic = None
    line = {"metric": metric, "value": value, "unit": "1080p frame-equiv/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": config, "clocks": clocks,
            "gpu_launches": len(launches) * args.steps,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak, "traffic": traffic,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
                         "kernel": "txq_kernel<W,H> (19 instances, one launch each per step); "
                                   "algorithmic bytes = 2 B/sample in + 12 B/kept coeff out + 2 B/block",
                         "bytes_per_step": int(total_bytes)},
            "kernels": kernels, "e2e": e2e, "me": me, "extras": extras,
            "step_ms_min_max": [float(min(step_ms)), float(max(step_ms))]}

    # ---- cpu_baseline (rank 0, N = 1 only) + full-size parity of frame 0
    if world == 1 and not args.no_cpu:
        H, O, R = load_cpu_libs()
        n_threads = os.cpu_count() or 1
        t, s, outs = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed + rank, n_threads,
                                   sizes, keep_outputs=True)  # warm-up pass; its outputs feed the parity check
        reps, t, s = 8, 0.0, 0
        for _ in range(reps):  # a bounded sample, long enough to average the thread start-up noise out
            ti, si, _o = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed + rank, n_threads, sizes)
            t += ti
            s += si
        line["cpu_baseline"] = {"value": s / FE_SAMPLES / t, "unit": "1080p frame-equiv/s",
                                "cores": n_threads, "kind": "reference" if R is not None else "port",
                                "sample": "%d x (1 frame x %d TX sizes) = %.1f FE after one warm-up pass, "
                                          "av1_fwd_txfm2d_*_c + av1_quantize_fp*_c on %d host threads"
                                          % (reps, len(sizes), s / FE_SAMPLES, n_threads)}
        # parity at full size: GPU (host-buffer call, frame 0) vs CPU outputs, bit-exact
        ok = True
        for ts in sizes:
            blocks, coeff, qc, dq, eob = outs[ts]
            g = [np.zeros_like(coeff), np.zeros_like(qc), np.zeros_like(dq), np.zeros_like(eob)]
            ctx.txq_batch_host(0, b200av1.Q_FP, 8, flags, blocks, None,
                               np.ascontiguousarray(frames[0]), None, qrows, g[0], g[1], g[2], g[3],
                               None)
            ok = ok and all(np.array_equal(a, b) for a, b in zip(g, (coeff, qc, dq, eob)))
        line["parity_full_frame"] = "bit-exact" if ok else "MISMATCH"
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
