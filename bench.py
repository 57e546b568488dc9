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


# tx-type sets of INTER blocks, the ones an encoder batches: av1_get_ext_tx_set_type
# (av1/common/blockd.h:1097-1107, is_inter = 1, reduced_tx_set = 0) + av1_ext_tx_used (:1036-1043).
# They are also the only types the reference's SIMD transforms implement for each size.
_SET_DCTONLY = (0,)
_SET_DCT_IDTX = (0, 9)
_SET_DTT9_IDTX_1DDCT = tuple(range(12))
_SET_ALL16 = tuple(range(16))


def valid_types(ts):
    up = max(TX_W[ts], TX_H[ts])
    if up > 32:
        return list(_SET_DCTONLY)
    if up == 32:
        return list(_SET_DCT_IDTX)
    if min(TX_W[ts], TX_H[ts]) == 16:
        return list(_SET_DTT9_IDTX_1DDCT)
    return list(_SET_ALL16)


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
    luma = np.zeros((len(frame_ids), 1080, 1920), np.uint8)
    rng0 = np.random.default_rng(seed)
    base = [synth.smooth_frame(rng0, w, h, 8) for (w, h) in planes]
    for i, f in enumerate(frame_ids):
        rng = np.random.default_rng([seed, f])
        o = 0
        for (w, h), s in zip(planes, base):
            src = np.roll(s, (f * 2, f * 3), (0, 1))
            if o == 0:
                luma[i] = src
            pred = synth.pred_from(rng, src, 8)
            out[i, o:o + w * h] = (src.astype(np.int16) - pred.astype(np.int16)).reshape(-1)
            o += w * h
    return out, planes, luma


def make_descriptors(ts, n_frames, planes, frame_elems, seed, bytes_per_sample=2, band_bytes=1 << 20):
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
    # batch-queue order: the queue walks the frame in bands of whole 64-row units per plane holding about
    # band_bytes of input and groups the blocks of a band by tx_type (warps stay on one code path).  The
    # 16 type groups of a band re-touch the same ~1 MB of input while it is still in L2 -- with whole
    # frames as the unit (round 1) a 4K 10-bit frame pair is 50 MB and the type groups re-read it from
    # DRAM (ncu: 3.3 x the algorithmic input bytes at 4x4, profiles/r2_hbd_v1_ncu_summary.md).  Smaller
    # bands cost warps that straddle two tx types (a 64-row band of 1080p has ~150 4x16 blocks per type);
    # raster order with the types mixed inside a warp is slower still (measured: 4x4 0.67 -> 0.57).
    def band_rows(w):
        return 64 * max(1, int(round(band_bytes / (w * bytes_per_sample * 64.0))))
    frame_id = np.repeat(np.arange(n_frames), per_frame)
    band = np.tile(np.concatenate([(np.arange(h // Hh) * Hh // band_rows(w)).repeat(w // W) + 1000 * pi
                                   for pi, (w, h) in enumerate(planes)]), n_frames)
    order = np.lexsort((blocks["tx_type"], band, frame_id))
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
    R = H.ref_avx2() or H.ref()
    return H, O, R


def run_configs0():
    """BASELINE configs[0]: the reference's own encoder, `aomenc --cpu-used=6 --good`, one thread, on the
    synthetic 352x288 8-bit 4:2:0 10-frame clip of baseline/gen_y4m.py (oracle/_ref/aomenc*, built by
    oracle/mk_ref.py).  Context for the CPU baseline; not part of the metric."""
    import subprocess
    import tempfile
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "baseline"))
    out = {}
    try:
        import gen_y4m
        with tempfile.TemporaryDirectory() as td:
            clip = os.path.join(td, "cif.y4m")
            gen_y4m.write_y4m(clip)
            for name in ("aomenc", "aomenc_avx2"):
                exe = os.path.join(here, "oracle", "_ref", name)
                if not os.path.exists(exe) or (name == "aomenc_avx2" and " avx2 " not in open("/proc/cpuinfo").read().replace("\n", " ")):
                    continue
                t0 = time.perf_counter()
                r = subprocess.run([exe, "--cpu-used=6", "--good", "--threads=1", "--limit=10", "-q", "-o",
                                    os.path.join(td, name + ".ivf"), clip], capture_output=True, timeout=300)
                dt = time.perf_counter() - t0
                if r.returncode == 0:
                    out[name] = {"seconds": round(dt, 3), "fps": round(10 / dt, 2),
                                 "bytes": os.path.getsize(os.path.join(td, name + ".ivf"))}
    except Exception as e:  # reported, never fatal to the bench line
        out["error"] = repr(e)[:200]
    if out:
        out["cmd"] = "aomenc --cpu-used=6 --good --threads=1 --limit=10, 352x288 8-bit 4:2:0 synthetic y4m, 10 frames, 2 passes"
    return out


def cpu_kind(R):
    """cpu_baseline.kind: the reference built with its x86 intrinsics (AVX2 dispatch; assembler-only
    kernels in C, see oracle/mk_ref.py), its generic C build, or the oracle port."""
    if R is None:
        return "port"
    return "reference-avx2" if R.refshim_simd() else "reference"


def cpu_desc(R):
    return {"port": "oracle restatement", "reference": "generic C build of the reference",
            "reference-avx2": "reference built with its SSE2..AVX2 intrinsics, run-time dispatch; the kernels "
                              "that exist only as assembler run as C"}[cpu_kind(R)]


def aligned_zeros(n, dtype, align=64):
    """The reference's SIMD flavours use aligned loads/stores on the encoder's 32-byte aligned buffers."""
    item = np.dtype(dtype).itemsize
    raw = np.zeros(n * item + align, np.uint8)
    off = (-raw.ctypes.data) % align
    return raw[off:off + n * item].view(dtype)


def cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes=range(19),
                  keep_outputs=False):
    """One frame per size through the CPU path.  Returns (seconds, samples, outputs)."""
    rows_mem = aligned_zeros(C.sizeof(H.QRow) * 3 * len(QINDEXES), np.uint8)  # SIMD quantisers load rows aligned
    rows = (H.QRow * (3 * len(QINDEXES))).from_buffer(rows_mem)
    for pl in range(3):
        for qi, q in enumerate(QINDEXES):
            O.av1o_build_qrow(8, q, pl, 0, C.byref(rows[pl * len(QINDEXES) + qi]))
    total_t, total_samples, outs = 0.0, 0, {}
    src = aligned_zeros(frames[:1].size, frames.dtype)
    src[:] = frames[:1].reshape(-1)
    for ts in sizes:
        blocks, samples = make_descriptors(ts, 1, planes, frame_elems, seed)
        n = blocks.size
        ne = n * max_eob(ts)
        coeff = aligned_zeros(ne, np.int32)
        qc = aligned_zeros(ne, np.int32)
        dq = aligned_zeros(ne, np.int32)
        eob = np.zeros(n, np.uint16)
        if R is not None:
            # av1_lowbd_fwd_txfm (hybrid_fwd_txfm.c:236) and av1_quantize_fp* as the encoder's own
            # dispatch binds them: the _c flavours in the generic build, SSE4.1/AVX2 in the avx2 one
            fwd = C.c_void_p(R.refshim_fwd_txfm_fn(ts, 0))
            quant = C.c_void_p(R.refshim_quantize_fp_fn(H.log_scale(ts), 0))
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
ME_MESH = ((64, 8), (28, 4), (15, 1), (7, 1))  # speed_features.c:25-33 (good quality)


class MeWork:
    """BASELINE configs[2] on a set of frames: every 64x64 superblock of each frame against 2
    reference frames, +-64 range clipped to the bordered reference, NSTEP + 4-stage mesh, L1 (HD)
    mv cost; the sub-pel tree (8 taps, 1/8 pel, 2 iterations per step) starts at the full-pel
    result.  All frames and references live in one plane pool (stacked vertically), so ONE launch
    per kernel covers the whole set.  References are the frame translated by a random mv plus
    noise, a function of (seed, frame id, reference) only."""

    def __init__(self, luma, frame_ids, seed, bd=8):
        import b200av1
        self.bd = bd
        F, h, w = luma.shape
        self.F, self.h, self.w = F, h, w
        self.rstride, self.rh = w + 2 * ME_BORDER, h + 2 * ME_BORDER
        self.cur = np.ascontiguousarray(luma)
        self.refs = np.zeros((F * 2, self.rh, self.rstride), luma.dtype)
        hi = (1 << bd) - 1
        for i, f in enumerate(frame_ids):
            rng = np.random.default_rng([seed, int(f), 77])
            for r in range(2):
                dy, dx = int(rng.integers(-20, 21)), int(rng.integers(-20, 21))
                noise = rng.integers(-3 << (bd - 8), (3 << (bd - 8)) + 1, (h, w), dtype=np.int16)
                ref = np.roll(luma[i], (dy, dx), (0, 1)).astype(np.int16) + noise
                self.refs[i * 2 + r] = np.pad(np.clip(ref, 0, hi).astype(luma.dtype), ME_BORDER, mode="edge")
        self.sbs = [(bx, by) for by in range(0, h - 63, 64) for bx in range(0, w - 63, 64)]
        self.n_sb = len(self.sbs)
        self.per_frame = 2 * self.n_sb
        self.n = self.per_frame * F
        self.blocks = self._blocks(range(F), pooled=True)
        self.sub = self._sub(self.blocks)
        self.step_param = int(b200av1.lib().b200_me_init_search_range(64))

    def _blocks(self, frames, pooled):
        """full-pel descriptors of `frames` (pool indices); pooled=False: offsets relative to a
        single frame's own cur / 2-reference buffers (the host-buffer call ships one frame)"""
        import b200av1
        h, w, rh, rs = self.h, self.w, self.rh, self.rstride
        blocks = np.zeros(self.per_frame * len(frames), b200av1.MEBLOCK_DTYPE)
        i = 0
        for j, f in enumerate(frames):
            fc, fr = (f, f * 2) if pooled else (0, 0)
            for r in range(2):
                for (bx, by) in self.sbs:
                    b = blocks[i]
                    b["src_offset"] = (fc * h + by) * w + bx
                    b["ref_offset"] = ((fr + r) * rh + by + ME_BORDER) * rs + bx + ME_BORDER
                    i += 1
        blocks["col_min"], blocks["col_max"], blocks["row_min"], blocks["row_max"] = -64, 64, -64, 64
        blocks["bw"], blocks["bh"] = 64, 64
        blocks["search_method"] = 0  # B200_ME_NSTEP
        blocks["flags"] = b200av1.ME_RUN_MESH
        blocks["mv_cost_type"] = 3  # MV_COST_L1_HDRES
        blocks["sad_per_bit"], blocks["error_per_bit"] = 40, 90
        blocks["step_param"] = b200av1.lib().b200_me_init_search_range(64)
        return blocks

    @staticmethod
    def _sub(blocks):
        import b200av1
        sb = blocks.copy()
        sb["col_min"], sb["col_max"], sb["row_min"], sb["row_max"] = -512, 512, -512, 512
        sb["flags"] = b200av1.ME_ALLOW_HP
        sb["subpel"] = 0 | (2 << 2) | (3 << 4)  # EIGHTH_PEL, 2 iters/step, USE_8_TAPS
        return sb

    def launch_struct(self, blocks_ptr, cur_ptr, ref_ptr, res_ptr, n, cur_rows, ref_rows):
        import b200av1
        l = b200av1.MeLaunch()
        l.highbd, l.bd, l.n_blocks = int(self.bd > 8), self.bd, n
        l.blocks, l.cur, l.ref = blocks_ptr, cur_ptr, ref_ptr
        l.cur_stride, l.ref_stride = self.w, self.rstride
        for i, (r, it) in enumerate(ME_MESH):
            l.mesh_range[i], l.mesh_interval[i] = r, it
        l.results = res_ptr
        l.window_bytes, l.max_bw, l.max_bh = (48 if self.bd == 8 else 88) * 1024, 64, 64
        l.families = b200av1.ME_FAMILY_NSTEP
        l.cur_rows, l.ref_rows = cur_rows, ref_rows
        l.max_col_span, l.max_row_span = 128, 128
        return l

    def cpu_params(self, H, frame):
        """harness.MeParams of every search of pool frame `frame` (host pointers)"""
        out = (H.MeParams * self.per_frame)()
        esz = self.cur.itemsize
        i = 0
        for r in range(2):
            ref = self.refs[frame * 2 + r]
            for (bx, by) in self.sbs:
                p = out[i]
                p.src = self.cur.ctypes.data + ((frame * self.h + by) * self.w + bx) * esz
                p.ref = ref.ctypes.data + ((by + ME_BORDER) * self.rstride + bx + ME_BORDER) * esz
                p.src_stride, p.ref_stride = self.w, self.rstride
                p.bw, p.bh, p.highbd, p.bd = 64, 64, int(self.bd > 8), self.bd
                p.col_min, p.col_max, p.row_min, p.row_max = -64, 64, -64, 64
                p.search_method, p.step_param, p.run_mesh_search = 0, self.step_param, 1
                for k, (rg, it) in enumerate(ME_MESH):
                    p.mesh_range[k], p.mesh_interval[k] = rg, it
                p.mv_cost_type, p.sad_per_bit, p.error_per_bit = 3, 40, 90
                p.want_cost_list = p.want_second_best = 1
                p.allow_hp, p.forced_stop, p.iters_per_step, p.subpel_search_type = 1, 0, 2, 3
                p.sp_col_min, p.sp_col_max, p.sp_row_min, p.sp_row_max = -512, 512, -512, 512
                i += 1
        return out


def cpu_run_me(H, O, R, work, frame, n_threads):
    """one frame's searches (full-pel + sub-pel) through the CPU path: the reference's own
    av1_full_pixel_search / av1_find_best_sub_pixel_tree (oracle/_ref via ref_shim.c) when
    built, else the oracle's restatement.  Returns (seconds, full results, sub results)."""
    params = work.cpu_params(H, frame)
    n = work.per_frame
    rf, rs = (H.MeResult * n)(), (H.MeResult * n)()
    if R is not None:
        full = C.cast(R.refshim_full_pixel_search, C.c_void_p)
        sub = C.cast(R.refshim_find_best_sub_pixel_tree, C.c_void_p)
    else:
        full = C.cast(O.av1o_full_pixel_search, C.c_void_p)
        sub = C.cast(O.av1o_find_best_sub_pixel_tree, C.c_void_p)
    t0 = time.perf_counter()
    O.av1o_drv_me(full, sub, params, n, rf, rs, n_threads)
    return time.perf_counter() - t0, rf, rs


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
    launches, keep, tot_bytes, tot_samples, size_bytes = [], [], 0, 0, []
    for ts in range(19):
        blocks, samples = make_descriptors(ts, n_frames, planes, frame_elems, seed + 4, bytes_per_sample=4)
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
        size_bytes.append(blocks.size * (TX_W[ts] * TX_H[ts] * 4 + max_eob(ts) * 12 + 18))
        tot_bytes += size_bytes[-1]
        tot_samples += samples
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(len(launches) + 1)]
    ms, per = 0.0, np.zeros(len(launches))
    for it in range(warmup + steps):
        ev[0].record(stream)
        for i, l in enumerate(launches):
            ctx.txq_launch(l, stream.cuda_stream)
            ev[i + 1].record(stream)
        torch.cuda.synchronize(dev)
        if it >= warmup:
            ms += ev[0].elapsed_time(ev[-1])
            per += [ev[i].elapsed_time(ev[i + 1]) for i in range(len(launches))]
    ms /= steps
    per /= steps
    peak = hbm_peak_gbs()
    kernels = [{"tx": "%dx%d" % (TX_W[ts], TX_H[ts]), "ms": round(float(per[ts]), 4),
                "GBps": round(size_bytes[ts] / (per[ts] * 1e-3) / 1e9, 1),
                "frac": round(size_bytes[ts] / (per[ts] * 1e-3) / 1e9 / peak, 3)} for ts in range(19)]
    fe4k = 3840 * 2160 * 3 // 2
    return {"workload": "config[3]: %d synthetic 4K 10-bit 4:2:0 frames, fused subtract + txfm + "
                        "av1_highbd_quantize_fp + av1_highbd_block_error, all 19 TX sizes" % n_frames,
            "ms_per_sweep": ms, "fe4k_per_s": tot_samples / fe4k / (ms * 1e-3),
            "fe1080p_per_s": tot_samples / FE_SAMPLES / (ms * 1e-3),
            "GBps": tot_bytes / (ms * 1e-3) / 1e9, "frac": tot_bytes / (ms * 1e-3) / 1e9 / peak,
            "bytes_per_sweep": int(tot_bytes), "launches_per_step": 19, "kernels": kernels}


def hbm_peak_gbs():
    """roofline denominator: the driver-measured copy bandwidth, else the profiling guide's fallback"""
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0))
    except (OSError, ValueError):
        return 6650.0


def bench_itx(ctx, dev, stream, steps, warmup, seed, hbm_peak, n_frames=8):
    """inverse transform + reconstruction + pixel-domain distortion of every block of
    `n_frames` synthetic 1080p 8-bit 4:2:0 frames, all 19 TX sizes, chained on the descriptors, dqcoeff
    and eob a fused forward launch (av1_quantize_fp from uint8 planes) left in HBM; then the transform-type
    search through host buffers (b200_txsearch_batch_host): every valid tx type of every 8x8 and 16x16
    luma block of one frame."""
    import torch
    import b200av1
    planes = [(1920, 1080), (960, 540), (960, 540)]
    frame_elems = sum(w * h for w, h in planes)
    rng = np.random.default_rng([seed, 9001])
    src = np.zeros((n_frames, frame_elems), np.uint8)
    pred = np.zeros((n_frames, frame_elems), np.uint8)
    for f in range(n_frames):
        o = 0
        for (w, h) in planes:
            low = rng.integers(30, 220, (h // 8 + 1, w // 8 + 1))
            img = np.kron(low, np.ones((8, 8), np.int64))[:h, :w] + rng.integers(-6, 7, (h, w))
            pr = np.roll(img, (1, 1), (0, 1)) + rng.integers(-4, 5, (h, w))
            src[f, o:o + w * h] = np.clip(img, 0, 255).reshape(-1)
            pred[f, o:o + w * h] = np.clip(pr, 0, 255).reshape(-1)
            o += w * h
    d_src = torch.from_numpy(src.reshape(-1)).to(dev)
    d_pred = torch.from_numpy(pred.reshape(-1)).to(dev)
    d_recon = torch.empty_like(d_pred)
    qrows = build_qrows(8)
    d_qrows = torch.from_numpy(qrows.view(np.uint8)).to(dev)
    n_out = n_frames * frame_elems
    d_qc, d_dq = (torch.empty(n_out, dtype=torch.int32, device=dev) for _ in range(2))
    fwd, inv, keep, per_size = [], [], [], []
    for ts in range(19):
        blocks, samples = make_descriptors(ts, n_frames, planes, frame_elems, seed + 9)
        preds = np.zeros(blocks.size, b200av1.TXPRED_DTYPE)
        preds["pred_offset"], preds["pred_stride"] = blocks["src_offset"], blocks["src_stride"]
        d_blocks = torch.from_numpy(blocks.view(np.uint8)).to(dev)
        d_preds = torch.from_numpy(preds.view(np.uint8)).to(dev)
        d_eob = torch.zeros(blocks.size, dtype=torch.int16, device=dev)
        d_dist = torch.zeros(blocks.size, dtype=torch.int64, device=dev)
        keep += [d_blocks, d_preds, d_eob, d_dist]
        l = b200av1.TxqLaunch()
        l.tx_size, l.in_kind, l.q_kind, l.bd, l.flags = ts, b200av1.IN_PIXELS8, b200av1.Q_FP, 8, 0
        l.n_blocks = int(blocks.size)
        l.blocks, l.preds = d_blocks.data_ptr(), d_preds.data_ptr()
        l.src, l.pred, l.qrows = d_src.data_ptr(), d_pred.data_ptr(), d_qrows.data_ptr()
        l.qcoeff, l.dqcoeff, l.eob = d_qc.data_ptr(), d_dq.data_ptr(), d_eob.data_ptr()
        fwd.append(l)
        il = b200av1.ItxLaunch(ts, 0, 8, int(blocks.size), d_blocks.data_ptr(), d_preds.data_ptr(),
                               d_dq.data_ptr(), d_eob.data_ptr(), d_src.data_ptr(), d_pred.data_ptr(),
                               d_recon.data_ptr(), d_dist.data_ptr(), 0)
        inv.append(il)
        # algorithmic bytes: kept dqcoeff (4 B) + pred + src in, recon out (1 B each), eob 2 B + dist 8 B per block
        per_size.append((blocks.size * (max_eob(ts) * 4 + TX_W[ts] * TX_H[ts] * 3 + 10), samples, d_eob))
    lib = b200av1.lib()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    kernels, tot_ms, tot_bytes, tot_samples, coded = [], 0.0, 0, 0, []
    for ts in range(19):
        ctx.txq_launch(fwd[ts], stream.cuda_stream)  # dqcoeff / eob of this size (the buffers are shared)
        ms = 0.0
        for it in range(warmup + steps):
            ev[0].record(stream)
            b200av1.check(lib.b200_itx_launch(ctx.h, C.byref(inv[ts]), C.c_void_p(stream.cuda_stream)))
            ev[1].record(stream)
            torch.cuda.synchronize(dev)
            if it >= warmup:
                ms += ev[0].elapsed_time(ev[1])
        ms /= steps
        nbytes, samples, d_eob = per_size[ts]
        coded.append(float((d_eob != 0).float().mean().item()))
        gbps = nbytes / (ms * 1e-3) / 1e9
        kernels.append({"tx": "%dx%d" % (TX_W[ts], TX_H[ts]), "ms": round(ms, 4), "GBps": round(gbps, 1),
                        "frac": round(gbps / hbm_peak, 3), "coded_blocks": round(coded[-1], 3)})
        tot_ms += ms
        tot_bytes += nbytes
        tot_samples += samples
    out = {"workload": "%d synthetic 1080p 8-bit 4:2:0 frames, all 19 TX sizes: b200_itx_launch (inverse transform + "
                       "reconstruction written + dist_block_px_domain) on the dqcoeff / eob of the fused forward launch"
                       % n_frames,
           "ms_per_sweep": tot_ms, "fe1080p_per_s": tot_samples / FE_SAMPLES / (tot_ms * 1e-3),
           "GBps": tot_bytes / (tot_ms * 1e-3) / 1e9, "frac": tot_bytes / (tot_ms * 1e-3) / 1e9 / hbm_peak,
           "bytes_per_sweep": int(tot_bytes), "launches_per_step": 19, "kernels": kernels,
           "bytes_rule": "4 B per kept dqcoeff + 1 B pred + 1 B src in, 1 B recon out per sample, 10 B per block"}
    # ---- transform-type search through host buffers: all valid types of the 8x8 / 16x16 luma blocks of a frame
    recs = []
    w, h = planes[0]
    for ts in (1, 2):
        W, Hh = TX_W[ts], TX_H[ts]
        offs = (np.arange(h // Hh)[:, None] * Hh * w + np.arange(w // W)[None, :] * W).reshape(-1)
        for tt in valid_types(ts):
            b = np.zeros(offs.size, b200av1.TXBLOCK_DTYPE)
            b["src_offset"], b["src_stride"], b["tx_size"], b["tx_type"] = offs, w, ts, tt
            b["qrow"] = 2
            recs.append(b)
    sblocks = np.concatenate(recs)
    spreds = np.zeros(sblocks.size, b200av1.TXPRED_DTYPE)
    spreds["pred_offset"], spreds["pred_stride"] = sblocks["src_offset"], sblocks["src_stride"]
    h_src, h_pred = pinned_copy(src[0, :w * h]), pinned_copy(pred[0, :w * h])
    res, d2h = None, 0
    t_best = None
    for it in range(4):
        t0 = time.perf_counter()
        res, d2h = b200av1.txsearch_batch_host(ctx, b200av1.IN_PIXELS8, b200av1.Q_B, 8, 7, sblocks, spreds,
                                               h_src, h_pred, qrows)
        dt = time.perf_counter() - t0
        if it and (t_best is None or dt < t_best):
            t_best = dt
    cand_samples = int(sum(TX_W[int(t)] * TX_H[int(t)] for t in sblocks["tx_size"]))
    out["txsearch_e2e"] = {
        "workload": "b200_txsearch_batch_host: every valid tx type (16 per block) of the %d 8x8 and 16x16 luma blocks "
                    "of one 1080p frame = %d candidates; host planes in, {eob, tx-domain error, rate estimate, "
                    "pixel-domain distortion} out" % (sblocks.size // 16, sblocks.size),
        "ms": t_best * 1e3, "candidates_per_s": sblocks.size / t_best,
        "fe1080p_per_s": cand_samples / FE_SAMPLES / t_best,
        "h2d_bytes": int(h_src.nbytes + h_pred.nbytes + sblocks.nbytes + spreds.nbytes + qrows.nbytes),
        "d2h_bytes": int(d2h), "d2h_bytes_dense_equivalent": int(cand_samples * 12 + sblocks.size * 2),
        "coded_fraction": float((res["eob"] != 0).mean())}
    return out


def bench_tf_queue(ctx, luma, seed, n_threads, reps=3):
    """The first consumer of the block queue: tf_motion_search over one 1080p
    frame pair -- 8160 16x16 blocks + 32640 8x8 sub-blocks, full-pel NSTEP + mesh and 8-tap sub-pel each
    (temporal_filter.c:88-260) -- as ONE b200_queue_tf_motion_search call (host planes in, per-block
    results out: uploads, border extension, 4 chained launches, D2H all inside the timed region), next
    to the reference's per-block loop (:804-872) on the host threads, results compared."""
    import b200av1
    H, _O, R = load_cpu_libs()
    cur = np.ascontiguousarray(luma[0])
    rng = np.random.default_rng([seed, 77])
    ref = np.roll(cur, (3, -5), (0, 1)).astype(np.int16) + rng.integers(-3, 4, cur.shape)
    ref = np.clip(ref, 0, 255).astype(np.uint8)
    h, w = cur.shape
    p = b200av1.tf_default_params(w, h)
    got, t_e2e, t_res = None, [], []
    for _ in range(reps + 1):
        t0 = time.perf_counter()
        fc, fr = b200av1.Frame(ctx, cur, p.border_in_pixels), b200av1.Frame(ctx, ref, p.border_in_pixels)
        got = b200av1.queue_tf_motion_search(ctx, fc, fr, p)
        t_e2e.append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        b200av1.queue_tf_motion_search(ctx, fc, fr, p)
        t_res.append(time.perf_counter() - t0)
        fc.free()
        fr.free()
    out = {"workload": "tf_motion_search over one 1080p 8-bit frame pair: %d blocks of 16x16 + 4 x 8x8 each = %d searches"
                       % (got.size, got.size * 5),
           "gpu_e2e_ms": min(t_e2e[1:]) * 1e3, "gpu_frames_resident_ms": min(t_res[1:]) * 1e3,
           "launches_per_pair": 5, "h2d_bytes": int(2 * cur.nbytes + got.size * 10 * 44),
           "d2h_bytes": int(got.size * 10 * 40)}
    if R is not None:
        cb, rb = np.pad(cur, p.border_in_pixels, mode="edge"), np.pad(ref, p.border_in_pixels, mode="edge")
        off = p.border_in_pixels * cb.shape[1] + p.border_in_pixels
        want = np.zeros(got.size, b200av1.TF_RESULT_DTYPE)
        t0 = time.perf_counter()
        R.refshim_tf_motion_search_frame_mt(C.c_void_p(cb.ctypes.data + off), C.c_void_p(rb.ctypes.data + off),
                                            cb.shape[1], w, h, 0, 8, C.byref(p), None, C.c_void_p(want.ctypes.data),
                                            n_threads)
        out["cpu_ms"] = (time.perf_counter() - t0) * 1e3
        out["cpu"] = "%s, %d threads over block rows" % (cpu_kind(R), n_threads)
        out["parity"] = "bit-exact" if all(np.array_equal(got.reshape(-1)[n], want[n]) for n in want.dtype.names) \
            else "MISMATCH"
    return out


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


def pinned(nbytes, dtype):
    """numpy view of page-locked host memory (b200_malloc_host); lives until process exit"""
    import b200av1
    p = C.c_void_p()
    b200av1.check(b200av1.lib().b200_malloc_host(C.c_size_t(max(nbytes, 16)), C.byref(p)))
    n = nbytes // np.dtype(dtype).itemsize
    return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(max(nbytes, 16),)).view(dtype)[:n]


def pinned_copy(a):
    out = pinned(a.nbytes, a.dtype)
    out[:] = a.reshape(-1)
    return out.reshape(a.shape)


def combined(r_txq, r_me):
    """frame-equiv/s for txfm+quant+ME = FE / (t_txq + t_me): one
    frame-equivalent passes through the transform+quantise stage once (rate r_txq, FE/s averaged
    over the 19 size passes) and through the motion search once (rate r_me, frames/s)."""
    return 1.0 / (1.0 / r_txq + 1.0 / r_me)


def me_mismatches(H, got_full, got_sub, rf, rs, n):
    bad = 0
    for i in range(n):
        g, e = got_full[i], rf[i]
        if (int(g["best_row"]), int(g["best_col"]), int(g["second_row"]), int(g["second_col"]),
                tuple(int(v) for v in g["cost_list"]), int(g["ret"])) != e.key():
            bad += 1
        g, e = got_sub[i], rs[i]
        if (int(g["best_row"]), int(g["best_col"]), int(g["ret"]), int(g["distortion"]),
                int(g["sse"])) != e.key(True):
            bad += 1
    return bad


def bench_config5(ctx, dev, stream, rank, world, seed, total_frames, batch, e2e_frames):
    """BASELINE configs[4]: `total_frames` synthetic 4K 10-bit 4:2:0 frames, frame i -> rank
    i mod world (strong scaling); every frame goes through the fused subtract + txfm +
    av1_highbd_quantize_fp + av1_highbd_block_error sweep over all 19 TX sizes (uint16 planes)
    and through the high-bit-depth motion search of all its 64x64 superblocks against 2
    references.  Resident: frames uploaded in batches of `batch` (untimed), launches timed with
    CUDA events.  e2e: the first `e2e_frames` frames of the rank through the host-buffer calls
    (b200_txq_batch_host_compact + b200_me_batch_host), copies inside the timed region."""
    import torch
    import b200av1
    from b200av1 import shard
    ids = shard.frames_of_rank(total_frames, rank, world)
    planes = [(3840, 2160), (1920, 1080), (1920, 1080)]
    frame_elems = sum(w * h for w, h in planes)
    rng = np.random.default_rng([seed, 5000])
    base, noise = [], []
    for _ in range(2):  # two base pictures; frame i = base[i % 2] panned by i
        pl = []
        for (w, h) in planes:
            low = rng.integers(100, 900, (h // 8 + 1, w // 8 + 1))
            pl.append((np.kron(low, np.ones((8, 8), np.int64))[:h, :w] + rng.integers(-8, 9, (h, w))).astype(np.int16))
        base.append(pl)
    for (w, h) in planes:  # one noise plane per picture plane, panned per frame (cheap and seeded)
        noise.append(rng.integers(-12, 13, (h, w), dtype=np.int16))

    def frame(i):
        srcs, preds = [], []
        for img, nz in zip(base[i % 2], noise):
            s0 = np.roll(img, (3 * i % 64, 5 * i % 64), (0, 1))
            pr = np.roll(s0, (1, 1), (0, 1)) + np.roll(nz, (7 * i % 97, 11 * i % 89), (0, 1))
            srcs.append(np.clip(s0, 0, 1023).astype(np.uint16).reshape(-1))
            preds.append(np.clip(pr, 0, 1023).astype(np.uint16).reshape(-1))
        return np.concatenate(srcs), np.concatenate(preds)

    qrows = build_qrows(10)
    d_qrows = torch.from_numpy(qrows.view(np.uint8)).to(dev)
    # descriptors of one batch (identical for every batch: offsets are batch-relative)
    n_out = batch * frame_elems
    d_coeff, d_qc, d_dq = (torch.empty(n_out, dtype=torch.int32, device=dev) for _ in range(3))
    d_src = torch.empty(batch * frame_elems, dtype=torch.int16, device=dev)
    d_pred = torch.empty(batch * frame_elems, dtype=torch.int16, device=dev)
    launches, keep, txq_bytes, txq_samples = [], [], 0, 0
    one_frame_blocks, one_frame_preds, out_off = [], [], 0
    for ts in range(19):
        blocks, samples = make_descriptors(ts, batch, planes, frame_elems, seed + 5, bytes_per_sample=4)
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
        txq_bytes += blocks.size * (TX_W[ts] * TX_H[ts] * 4 + max_eob(ts) * 12 + 18)
        txq_samples += samples
        if e2e_frames:
            b1, _s = make_descriptors(ts, 1, planes, frame_elems, seed + 5, bytes_per_sample=4)
            b1["out_offset"] += out_off
            out_off += b1.size * max_eob(ts)
            p1 = np.zeros(b1.size, b200av1.TXPRED_DTYPE)
            p1["pred_offset"], p1["pred_stride"] = b1["src_offset"], b1["src_stride"]
            one_frame_blocks.append(b1)
            one_frame_preds.append(p1)
    # motion search of a batch: cur = luma plane of each frame inside d_src (stride 3840)
    w, h = planes[0]
    rs, rh = w + 2 * ME_BORDER, h + 2 * ME_BORDER
    sbs = [(bx, by) for by in range(0, h - 63, 64) for bx in range(0, w - 63, 64)]
    per_frame = 2 * len(sbs)

    def me_blocks(nf, frame_stride_elems):
        blocks = np.zeros(per_frame * nf, b200av1.MEBLOCK_DTYPE)
        i = 0
        for f in range(nf):
            for r in range(2):
                for (bx, by) in sbs:
                    blocks[i]["src_offset"] = f * frame_stride_elems + by * w + bx
                    blocks[i]["ref_offset"] = ((f * 2 + r) * rh + by + ME_BORDER) * rs + bx + ME_BORDER
                    i += 1
        blocks["col_min"], blocks["col_max"], blocks["row_min"], blocks["row_max"] = -64, 64, -64, 64
        blocks["bw"], blocks["bh"] = 64, 64
        blocks["flags"] = b200av1.ME_RUN_MESH
        blocks["mv_cost_type"] = 3
        blocks["sad_per_bit"], blocks["error_per_bit"] = 40, 90
        blocks["step_param"] = b200av1.lib().b200_me_init_search_range(64)
        return blocks

    mb = me_blocks(batch, frame_elems)
    d_mb = torch.from_numpy(mb.view(np.uint8)).to(dev)
    d_msub = torch.from_numpy(MeWork._sub(mb).view(np.uint8)).to(dev)
    d_refs = torch.empty(batch * 2 * rh * rs, dtype=torch.int16, device=dev)
    d_rf = torch.zeros(mb.size * 40, dtype=torch.uint8, device=dev)
    d_rsub = torch.zeros(mb.size * 40, dtype=torch.uint8, device=dev)

    def me_launch(kind):
        l = b200av1.MeLaunch()
        l.highbd, l.bd, l.n_blocks = 1, 10, int(mb.size)
        l.blocks = (d_mb if kind == 0 else d_msub).data_ptr()
        l.cur, l.ref = d_src.data_ptr(), d_refs.data_ptr()
        l.cur_stride, l.ref_stride = w, rs
        for i, (r, it) in enumerate(ME_MESH):
            l.mesh_range[i], l.mesh_interval[i] = r, it
        l.results = (d_rf if kind == 0 else d_rsub).data_ptr()
        l.window_bytes, l.max_bw, l.max_bh = 88 * 1024, 64, 64
        l.families = b200av1.ME_FAMILY_NSTEP
        # cur rows: the luma planes sit frame_elems apart inside d_src (not a multiple of the
        # stride times rows only when chroma follows): describe the pool as rows of 3840 samples
        l.cur_rows, l.ref_rows = batch * frame_elems // w, batch * 2 * rh
        l.max_col_span, l.max_row_span = 128, 128
        if kind == 1:
            l.start_from = d_rf.data_ptr()
        fn = b200av1.lib().b200_me_fullpel_launch if kind == 0 else b200av1.lib().b200_me_subpel_launch
        b200av1.check(fn(ctx.h, C.byref(l), C.c_void_p(stream.cuda_stream)))

    def make_refs(i, cur_luma):
        frng = np.random.default_rng([seed, 5000 + int(i), 9])
        out = np.zeros((2, rh, rs), np.uint16)
        for r in range(2):
            dy, dx = int(frng.integers(-20, 21)), int(frng.integers(-20, 21))
            ref = np.roll(cur_luma, (dy, dx), (0, 1)).astype(np.int16) + np.roll(noise[0], (13 * i + r, 3 * i + 5 * r), (0, 1))
            out[r] = np.pad(np.clip(ref, 0, 1023).astype(np.uint16), ME_BORDER, mode="edge")
        return out

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    t_txq = t_me = 0.0
    n_done = 0
    e2e_t = e2e_n = 0
    e2e_h2d = e2e_d2h = 0
    lib = b200av1.lib()
    if e2e_frames:
        eblocks = pinned_copy(np.concatenate(one_frame_blocks))
        epreds = pinned_copy(np.concatenate(one_frame_preds))
        h_levels = pinned(out_off * 4, np.int32)
        h_loff = pinned((eblocks.size + 1) * 4, np.uint32)
        h_eob = pinned(eblocks.size * 2, np.uint16)
        h_dist = pinned(eblocks.size * 16, np.int64).reshape(-1, 2)
        h_src, h_pred = pinned(frame_elems * 2, np.uint16), pinned(frame_elems * 2, np.uint16)
        h_refs = pinned(2 * rh * rs * 2, np.uint16)
        emb = pinned_copy(me_blocks(1, frame_elems))
        esub = pinned_copy(MeWork._sub(emb))
        h_rf = pinned(emb.size * 40, b200av1.MERESULT_DTYPE)
        h_rs = pinned(emb.size * 40, b200av1.MERESULT_DTYPE)
    for b0 in range(0, len(ids), batch):
        chunk = ids[b0:b0 + batch]
        hs = np.zeros((batch, frame_elems), np.uint16)
        hp = np.zeros((batch, frame_elems), np.uint16)
        hr = np.zeros((batch * 2, rh, rs), np.uint16)
        for j, i in enumerate(chunk):
            hs[j], hp[j] = frame(i)
            hr[2 * j:2 * j + 2] = make_refs(i, hs[j][:w * h].reshape(h, w))
            if e2e_n < e2e_frames:
                # ---- e2e: this frame through the host-buffer calls, copies inside the timed region
                h_src[:], h_pred[:], h_refs[:] = hs[j], hp[j], hr[2 * j:2 * j + 2].reshape(-1)
                def e2e_frame():
                    d2h = ctx.txq_batch_host_compact(b200av1.IN_PIXELS16, b200av1.Q_FP_HBD, 10,
                                                     b200av1.TXQ_BLOCK_ERROR, eblocks, epreds, h_src, h_pred, qrows,
                                                     out_off, h_levels, h_loff, h_eob, h_dist)
                    hl = b200av1.MeLaunch()
                    hl.highbd, hl.bd, hl.n_blocks = 1, 10, int(emb.size)
                    hl.blocks, hl.cur, hl.ref = emb.ctypes.data, h_src.ctypes.data, h_refs.ctypes.data
                    hl.cur_stride, hl.ref_stride = w, rs
                    for k, (rg, it) in enumerate(ME_MESH):
                        hl.mesh_range[k], hl.mesh_interval[k] = rg, it
                    hl.window_bytes, hl.max_bw, hl.max_bh = 88 * 1024, 64, 64
                    hl.families = b200av1.ME_FAMILY_NSTEP
                    hl.cur_rows, hl.ref_rows = h, 2 * rh
                    hl.max_col_span, hl.max_row_span = 128, 128
                    b200av1.check(lib.b200_me_batch_host(ctx.h, C.byref(hl), C.c_size_t(w * h * 2),
                                                         C.c_size_t(h_refs.nbytes), C.c_void_p(esub.ctypes.data),
                                                         C.c_void_p(h_rf.ctypes.data), C.c_void_p(h_rs.ctypes.data)))
                    return d2h
                if e2e_n == 0:
                    e2e_frame()  # untimed once: scratch growth, lazy kernel loading of the uint16 instances
                torch.cuda.synchronize(dev)
                t0 = time.perf_counter()
                d2h = e2e_frame()
                e2e_t += time.perf_counter() - t0
                e2e_n += 1
                e2e_h2d += 2 * frame_elems * 2 + eblocks.nbytes + epreds.nbytes + w * h * 2 + h_refs.nbytes + 2 * emb.nbytes
                e2e_d2h += d2h + 2 * h_rf.nbytes
        d_src.copy_(torch.from_numpy(hs.reshape(-1).view(np.int16)))
        d_pred.copy_(torch.from_numpy(hp.reshape(-1).view(np.int16)))
        d_refs.copy_(torch.from_numpy(hr.reshape(-1).view(np.int16)))
        torch.cuda.synchronize(dev)
        reps = 2 if b0 == 0 else 1  # the first batch once untimed (warm-up)
        for rep in range(reps):
            ev[0].record(stream)
            for l in launches:
                ctx.txq_launch(l, stream.cuda_stream)
            ev[1].record(stream)
            me_launch(0)
            me_launch(1)
            ev[2].record(stream)
            torch.cuda.synchronize(dev)
        frac = len(chunk) / batch  # a last partial batch still runs `batch` frames of work
        t_txq += ev[0].elapsed_time(ev[1]) * frac
        t_me += ev[1].elapsed_time(ev[2]) * frac
        n_done += len(chunk)
    return {"frames_rank": n_done, "t_txq_ms": t_txq, "t_me_ms": t_me, "txq_samples_per_frame": txq_samples / batch,
            "txq_bytes_per_frame": txq_bytes / batch, "me_searches_per_frame": per_frame,
            "e2e_s": e2e_t, "e2e_frames": e2e_n, "e2e_h2d": e2e_h2d, "e2e_d2h": e2e_d2h}


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
    ap.add_argument("--no-me", action="store_true", help="(debug) time the transform half only")
    ap.add_argument("--no-extras", action="store_true", help="skip the config[3] / config[4] / Hadamard sections")
    ap.add_argument("--config5-frames", type=int, default=120, help="frames of the 4K 10-bit batch (all ranks)")
    ap.add_argument("--sizes", default="", help="comma list of tx sizes (debug)")
    ap.add_argument("--only", default="", choices=["", "hbd4k", "itx", "hadamard", "tf"],
                    help="(profiling) run one extras section alone and print its JSON")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    # one process per GPU on one host: the host-buffer tier's validation threads divide the cores between the ranks
    os.environ.setdefault("B200_HOST_THREADS", str(max(1, min(16, (os.cpu_count() or 1) // max(world, 1)))))
    sizes = [int(s) for s in args.sizes.split(",")] if args.sizes else list(range(19))
    seed = 2001
    metric = "1080p frame-equiv/s for txfm+quant+ME SAD; HBM GB/s as % of B200 peak"
    config = {"workload": "txq_me_1080p_8bit: per GPU %d synthetic 1080p 4:2:0 frames; each step = fused fwd txfm + "
                          "quantize_fp over all %d TX sizes of every frame (configs[1]) + full-pel (NSTEP + 4-stage "
                          "mesh, +-64) and sub-pel (8-tap tree) motion search of every 64x64 superblock against 2 "
                          "references (configs[2]); value = 1 / (1/txq FE/s + 1/ME frames/s)" % (args.frames, len(sizes)),
              "frames_per_gpu": args.frames, "tx_sizes": len(sizes), "quantizer": "av1_quantize_fp",
              "qindex_set": QINDEXES, "me": "960 searches/frame: 480 SBs x 2 refs, MV_COST_L1_HDRES",
              "l2": "inputs (%.0f MB residual + %.0f MB planes per GPU) larger than L2; outputs streamed"
                    % (args.frames * FE_SAMPLES * 2 / 1e6, args.frames * (1920 * 1080 + 2 * 2240 * 1400) / 1e6),
              "parallelism": "frames sharded over %d GPU(s), no collective" % world,
              "butterflies": "exact64 (forced)" if args.force_exact else "guarded (32-bit when |residual|<=4094, proved exact; else 64-bit)"}

    # ----------------------------------------------------------- reference arm
    if args.impl == "reference":
        if rank != 0:
            return 0
        H, O, R = load_cpu_libs()
        frames, planes, luma = make_residual_frames(1, seed)
        frame_elems = frames.shape[1]
        n_threads = os.cpu_count() or 1
        work = MeWork(luma, [0], seed)
        for _ in range(min(args.warmup, 1)):
            cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes)
        tt, ss, tm = 0.0, 0, 0.0
        for _ in range(args.steps):
            t, s_, _o = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed, n_threads, sizes)
            tt += t
            ss += s_
            tm += cpu_run_me(H, O, R, work, 0, n_threads)[0]
        r_txq, r_me = ss / FE_SAMPLES / tt, args.steps / tm
        val = combined(r_txq, r_me)
        kind = cpu_kind(R)
        line = {"impl": "reference", "metric": metric, "value": val, "unit": "1080p frame-equiv/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": (tt + tm) / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "int32", "data": "synthetic", "config": config,
                "txq_fe_per_s": r_txq, "me_frames_per_s": r_me,
                "cpu_baseline": {"value": val, "unit": "1080p frame-equiv/s", "cores": n_threads,
                                 "kind": kind,
                                 "sample": "per step: 1 frame x %d TX sizes (%.1f FE) through av1_lowbd_fwd_txfm + "
                                           "av1_quantize_fp*, and the frame's 960 searches through "
                                           "av1_full_pixel_search + av1_find_best_sub_pixel_tree (%s)"
                                           % (len(sizes), ss / args.steps / FE_SAMPLES, cpu_desc(R))},
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
    lib = b200av1.lib()
    stream = torch.cuda.Stream(device=dev)
    if args.only:
        if args.only == "hbd4k":
            out = bench_hbd4k(ctx, dev, stream, args.steps, args.warmup, seed)
        elif args.only == "itx":
            out = bench_itx(ctx, dev, stream, args.steps, args.warmup, seed, hbm_peak_gbs())
        elif args.only == "tf":
            out = bench_tf_queue(ctx, make_residual_frames([0], seed)[2], seed, os.cpu_count() or 1)
        else:
            d_src = torch.from_numpy(make_residual_frames(range(args.frames), seed)[0].reshape(-1)).to(dev)
            out = bench_hadamard(ctx, dev, stream, args.steps, args.warmup, d_src, args.frames)
        print(json.dumps({"only": args.only, "result": out}))
        return 0

    def allmax(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    from b200av1 import shard
    frame_ids = shard.frames_of_rank(args.frames * world, rank, world)
    frames, planes, luma = make_residual_frames(frame_ids, seed)
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

    # ---- motion search of the same frames (their luma planes), resident in HBM
    work = None
    if not args.no_me:
        work = MeWork(luma, frame_ids, seed)
        d_cur = torch.from_numpy(work.cur.reshape(-1)).to(dev)
        d_ref = torch.from_numpy(work.refs.reshape(-1)).to(dev)
        d_mblk = torch.from_numpy(work.blocks.view(np.uint8)).to(dev)
        d_msub = torch.from_numpy(work.sub.view(np.uint8)).to(dev)
        d_rfull = torch.zeros(work.n * 40, dtype=torch.uint8, device=dev)
        d_rsub = torch.zeros(work.n * 40, dtype=torch.uint8, device=dev)
        d_stats = torch.zeros(2, dtype=torch.int64, device=dev)
        l_full = work.launch_struct(d_mblk.data_ptr(), d_cur.data_ptr(), d_ref.data_ptr(), d_rfull.data_ptr(),
                                    work.n, work.F * work.h, work.F * 2 * work.rh)
        l_full.stats = d_stats.data_ptr()
        l_sub = work.launch_struct(d_msub.data_ptr(), d_cur.data_ptr(), d_ref.data_ptr(), d_rsub.data_ptr(),
                                   work.n, work.F * work.h, work.F * 2 * work.rh)
        l_sub.start_from = d_rfull.data_ptr()  # chained on the device: start = 8 x full-pel best

    n_txq = len(launches)

    def step(events=None):
        for i, l in enumerate(launches):
            if events is not None:
                events[i].record(stream)
            ctx.txq_launch(l, sptr)
        if events is not None:
            events[n_txq].record(stream)
        if work is not None:
            b200av1.check(lib.b200_me_fullpel_launch(ctx.h, C.byref(l_full), C.c_void_p(sptr)))
            if events is not None:
                events[n_txq + 1].record(stream)
            b200av1.check(lib.b200_me_subpel_launch(ctx.h, C.byref(l_sub), C.c_void_p(sptr)))
        elif events is not None:
            events[n_txq + 1].record(stream)
        if events is not None:
            events[n_txq + 2].record(stream)

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    with torch.cuda.stream(stream):
        for _ in range(args.warmup):
            step()
    barrier()
    if work is not None:
        d_stats.zero_()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(n_txq + 3)] for _ in range(args.steps)]
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier()
    for k in range(args.steps):
        step(ev[k])
    barrier()
    clocks = sampler.stop()
    step_ms = [ev[k][0].elapsed_time(ev[k][n_txq + 2]) for k in range(args.steps)]
    total_ms = ev[0][0].elapsed_time(ev[args.steps - 1][n_txq + 2])
    kern_ms = np.zeros(n_txq)
    t_txq = t_full = t_sub = 0.0
    for k in range(args.steps):
        for i in range(n_txq):
            kern_ms[i] += ev[k][i].elapsed_time(ev[k][i + 1])
        t_txq += ev[k][0].elapsed_time(ev[k][n_txq])
        t_full += ev[k][n_txq].elapsed_time(ev[k][n_txq + 1])
        t_sub += ev[k][n_txq + 1].elapsed_time(ev[k][n_txq + 2])
    kern_ms /= args.steps
    total_ms_max = allmax(total_ms)
    t_txq_max, t_me_max = allmax(t_txq), allmax(t_full + t_sub)
    fe_per_step_rank = total_samples / FE_SAMPLES
    r_txq = fe_per_step_rank * world * args.steps / (t_txq_max / 1e3)
    if work is not None:
        r_me = args.frames * world * args.steps / (t_me_max / 1e3)
        value = combined(r_txq, r_me)
        stats = d_stats.cpu().numpy()
        sad_pixels_per_step = float(stats[1]) / args.steps
    else:
        r_me, value = None, r_txq

    # ---- e2e: the same work through the host-buffer C-ABI calls, 1 frame per step:
    # b200_txq_batch_host_compact (19 sizes; levels in scan order up to eob come back) and
    # b200_me_batch_host (planes + descriptors up, 2 x 40 B per search back)
    e2e = None
    if args.e2e_steps > 0:
        eb, e_samples = [], 0
        out_off = 0
        for ts in sizes:
            b, s_ = make_descriptors(ts, 1, planes, frame_elems, seed + rank)
            b["out_offset"] += out_off
            out_off += b.size * max_eob(ts)
            eb.append(b)
            e_samples += s_
        eblocks = pinned_copy(np.concatenate(eb))
        h_src = pinned_copy(frames[0])
        h_levels = pinned(out_off * 2, np.int16)
        h_loff = pinned((eblocks.size + 1) * 4, np.uint32)
        h_eob = pinned(eblocks.size * 2, np.uint16)
        d2h_txq = [0]
        if work is not None:
            # one host-buffer call carries E frames (their planes stacked like the resident pool): a single
            # 1080p frame is 960 searches = 2.2 waves of resident CTAs, so calls of one frame time the tail
            E = min(8, args.frames)
            e_mblk = pinned_copy(work._blocks(range(E), pooled=True))
            e_msub = pinned_copy(MeWork._sub(e_mblk))
            h_cur = pinned_copy(work.cur[0:E])
            h_ref = pinned_copy(work.refs[0:2 * E])
            h_rf = pinned(e_mblk.size * 40, b200av1.MERESULT_DTYPE)
            h_rs = pinned(e_mblk.size * 40, b200av1.MERESULT_DTYPE)
            hl = work.launch_struct(e_mblk.ctypes.data, h_cur.ctypes.data, h_ref.ctypes.data, None,
                                    int(e_mblk.size), E * work.h, 2 * E * work.rh)

        def e2e_txq():
            d2h_txq[0] = ctx.txq_batch_host_compact(0, b200av1.Q_FP, 8, flags & ~b200av1.TXQ_STORE_COEFF, eblocks,
                                                    None, h_src, None, qrows, out_off, h_levels, h_loff, h_eob)

        def e2e_me():
            b200av1.check(lib.b200_me_batch_host(ctx.h, C.byref(hl), C.c_size_t(h_cur.nbytes),
                                                 C.c_size_t(h_ref.nbytes), C.c_void_p(e_msub.ctypes.data),
                                                 C.c_void_p(h_rf.ctypes.data), C.c_void_p(h_rs.ctypes.data)))

        e2e_txq()
        if work is not None:
            e2e_me()
        barrier()
        te_txq = te_me = 0.0
        for _ in range(args.e2e_steps):
            t0 = time.perf_counter()
            e2e_txq()
            t1 = time.perf_counter()
            if work is not None:
                e2e_me()
            t2 = time.perf_counter()
            te_txq += t1 - t0
            te_me += t2 - t1
        torch.cuda.synchronize(dev)
        te_txq, te_me = allmax(te_txq), allmax(te_me)
        re_txq = e_samples / FE_SAMPLES * world * args.e2e_steps / te_txq
        h2d = int(frame_elems * 2 + eblocks.nbytes + qrows.nbytes)
        d2h = int(d2h_txq[0])
        if work is not None:
            re_me = world * E * args.e2e_steps / te_me
            e2e_value = combined(re_txq, re_me)
            h2d += int(h_cur.nbytes + h_ref.nbytes + 2 * e_mblk.nbytes)
            d2h += int(2 * h_rf.nbytes)
        else:
            re_me, e2e_value = None, re_txq
        e2e = {"value": e2e_value, "unit": "1080p frame-equiv/s", "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "txq_fe_per_s": re_txq, "me_frames_per_s": re_me,
               "d2h_dense_bytes_per_step": int(out_off * 12 + eblocks.size * 2),
               "me_frames_per_call": E if work is not None else 0,
               "note": "per step: 1 frame through b200_txq_batch_host_compact (pinned host buffers; descriptor "
                       "bucketing + H2D + %d launches + on-device packing + D2H of eob, offsets and the levels in "
                       "scan order up to eob) and %d frames through ONE b200_me_batch_host call (cur + 2 reference "
                       "planes per frame + descriptors H2D, full-pel + chained sub-pel launch, 2 x 40 B per search "
                       "D2H), all inside the timed region; the byte counts are those of one such step"
                       % (len(sizes), E if work is not None else 0)}

    extras = None
    if not args.no_extras:
        extras = {}
        if world == 1:
            extras["hadamard"] = bench_hadamard(ctx, dev, stream, max(2, args.steps // 2), 2, d_src, args.frames)
            extras["hbd4k"] = bench_hbd4k(ctx, dev, stream, max(2, args.steps // 2), 2, seed)
            extras["itx"] = bench_itx(ctx, dev, stream, max(2, args.steps // 2), 2, seed, hbm_peak_gbs())
            if not args.no_cpu:
                extras["tf_queue"] = bench_tf_queue(ctx, luma, seed, os.cpu_count() or 1)
        # BASELINE configs[4] at every N: 120 frames over the ranks
        del d_coeff, d_qc, d_dq
        torch.cuda.empty_cache()
        c5 = bench_config5(ctx, dev, stream, rank, world, seed, args.config5_frames, 4, 2)
        t5_txq, t5_me = allmax(c5["t_txq_ms"]), allmax(c5["t_me_ms"])
        te5 = allmax(c5["e2e_s"] / max(c5["e2e_frames"], 1))
        fe4k = 3840 * 2160 * 3 // 2
        r5_txq = c5["txq_samples_per_frame"] / fe4k * args.config5_frames / (t5_txq / 1e3)
        r5_me = args.config5_frames / (t5_me / 1e3)
        extras["config5"] = {
            "workload": "configs[4]: %d synthetic 4K 10-bit 4:2:0 frames, frame i -> GPU i mod %d (strong scaling); "
                        "fused subtract + txfm + av1_highbd_quantize_fp + av1_highbd_block_error over all 19 TX sizes "
                        "from uint16 planes + high-bit-depth motion search (1980 64x64 SBs x 2 refs, +-64, NSTEP + "
                        "mesh + 8-tap sub-pel) per frame" % (args.config5_frames, world),
            "frames_total": args.config5_frames, "txq_fe4k_per_s": r5_txq, "me_frames4k_per_s": r5_me,
            "fe4k_per_s": combined(r5_txq, r5_me),
            "txq_GBps_per_gpu": c5["txq_bytes_per_frame"] * c5["frames_rank"] / (c5["t_txq_ms"] / 1e3) / 1e9,
            "ms_txq_max_rank": t5_txq, "ms_me_max_rank": t5_me,
            "e2e_fe4k_per_s": world / te5 if te5 > 0 else None,
            "e2e_sample": "%d frames per rank through b200_txq_batch_host_compact + b200_me_batch_host" % c5["e2e_frames"],
            "e2e_h2d_bytes_per_frame": c5["e2e_h2d"] // max(c5["e2e_frames"], 1),
            "e2e_d2h_bytes_per_frame": c5["e2e_d2h"] // max(c5["e2e_frames"], 1)}

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
    if os.path.exists(tp) and len(sizes) == 19:
        try:
            tj = json.load(open(tp))
            # ncu capture (one --set full pass per kernel) at tj["frames_per_launch"] frames per launch;
            # DRAM bytes scale with the frames
            traffic = tj["txq_kernel_dram_bytes_per_step"] * args.frames / tj["frames_per_launch"]
            assert traffic >= 0.95 * total_bytes, "DRAM traffic below the algorithmic bytes"
        except (OSError, ValueError, KeyError):
            traffic = None
    roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
            "frac": achieved / peak, "traffic": traffic,
            "peak_source": "MEASURED_PEAKS.json hbm_gbs" if peaks else "fallback 6650",
            "kernel": "txq_kernel<W,H> (19 instances, one launch each per step); "
                      "algorithmic bytes = 2 B/sample in + 12 B/kept coeff out + 2 B/block",
            "bytes_per_step": int(total_bytes),
            "share_of_step": float(t_txq / (t_txq + t_full + t_sub))}
    me = None
    if work is not None:
        sad_peak = C.c_double(0)
        b200av1.check(lib.b200_measure_sad_peak(ctx.h, C.byref(sad_peak)))
        full_s = t_full / args.steps / 1e3
        unique = work.n * (4096 + 192 * 192 + 64)
        me = {"fullpel_ms_per_frame": t_full / args.steps / args.frames,
              "subpel_ms_per_frame": t_sub / args.steps / args.frames,
              "frames_per_s_per_gpu": args.frames * args.steps / ((t_full + t_sub) / 1e3),
              "searches_per_frame": work.per_frame, "launches_per_step": 2, "frames_per_launch": args.frames}
        roof["me"] = {"bound": "int_alu", "kernel": "fullpel_kernel<0> (n-step + mesh), %.0f %% of the step"
                                                    % (100 * t_full / (t_txq + t_full + t_sub)),
                      "achieved": sad_pixels_per_step / full_s / 1e9, "peak": sad_peak.value,
                      "unit": "G pixel-absdiff/s", "frac": sad_pixels_per_step / full_s / 1e9 / sad_peak.value,
                      "peak_source": "b200_measure_sad_peak: VABSDIFF4.ACC issued back to back on every SM, this run",
                      "sad_pixels_per_step": sad_pixels_per_step,
                      "unique_bytes_per_step": unique, "unique_GBps": unique / full_s / 1e9,
                      "hbm_frac": unique / full_s / 1e9 / peak}
    line = {"metric": metric, "value": value, "unit": "1080p frame-equiv/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int32",
            "data": "synthetic", "config": config, "clocks": clocks,
            "gpu_launches": (n_txq + (2 if work is not None else 0)) * args.steps,
            "txq_fe_per_s": r_txq, "me_frames_per_s": r_me,
            "roofline": roof, "kernels": kernels, "e2e": e2e, "me": me, "extras": extras,
            "step_ms_min_max": [float(min(step_ms)), float(max(step_ms))]}

    # ---- cpu_baseline (rank 0, N = 1 only) + full-size parity of frame 0
    if world == 1 and not args.no_cpu:
        H, O, R = load_cpu_libs()
        n_threads = os.cpu_count() or 1
        t, s_, outs = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed + rank, n_threads,
                                    sizes, keep_outputs=True)  # warm-up pass; its outputs feed the parity check
        reps, t, s_ = 4, 0.0, 0
        for _ in range(reps):  # a bounded sample, long enough to average the thread start-up noise out
            ti, si, _o = cpu_run_sweep(H, O, R, frames, planes, frame_elems, seed + rank, n_threads, sizes)
            t += ti
            s_ += si
        c_txq = s_ / FE_SAMPLES / t
        kind = cpu_kind(R)
        cb = {"unit": "1080p frame-equiv/s", "cores": n_threads, "kind": kind, "txq_fe_per_s": c_txq}
        sample = "%d x (1 frame x %d TX sizes) = %.1f FE after one warm-up pass, av1_lowbd_fwd_txfm + " \
                 "av1_quantize_fp*" % (reps, len(sizes), s_ / FE_SAMPLES)
        if work is not None:
            tm, rf, rs = cpu_run_me(H, O, R, work, 0, n_threads)
            cb["me_frames_per_s"] = 1.0 / tm
            cb["value"] = combined(c_txq, 1.0 / tm)
            sample += "; 1 frame = %d searches through av1_full_pixel_search + av1_find_best_sub_pixel_tree" % work.per_frame
        else:
            cb["value"] = c_txq
        cb["sample"] = sample + " (%s) on %d host threads" % (cpu_desc(R), n_threads)
        Rc = H.ref()
        if Rc is not None and Rc is not R:  # the generic C build beside it, one pass
            tc, sc_, _o = cpu_run_sweep(H, O, Rc, frames, planes, frame_elems, seed + rank, n_threads, sizes)
            gc = {"kind": "reference", "txq_fe_per_s": sc_ / FE_SAMPLES / tc}
            if work is not None:
                tmc = cpu_run_me(H, O, Rc, work, 0, n_threads)[0]
                gc["me_frames_per_s"] = 1.0 / tmc
                gc["value"] = combined(gc["txq_fe_per_s"], 1.0 / tmc)
            cb["generic_c"] = gc
        cfg0 = run_configs0()
        if cfg0:
            cb["configs0_aomenc"] = cfg0
        line["cpu_baseline"] = cb
        # parity at full size: GPU (host-buffer calls, frame 0) vs CPU outputs, bit-exact
        ok = True
        for ts in sizes:
            blocks, coeff, qc, dq, eob = outs[ts]
            g = [np.zeros_like(coeff), np.zeros_like(qc), np.zeros_like(dq), np.zeros_like(eob)]
            ctx.txq_batch_host(0, b200av1.Q_FP, 8, flags, blocks, None,
                               np.ascontiguousarray(frames[0]), None, qrows, g[0], g[1], g[2], g[3],
                               None)
            ok = ok and all(np.array_equal(a, b) for a, b in zip(g, (coeff, qc, dq, eob)))
        line["parity_full_frame"] = "bit-exact" if ok else "MISMATCH"
        if work is not None and args.e2e_steps > 0:
            bad = me_mismatches(H, h_rf, h_rs, rf, rs, work.per_frame)
            line["parity_me_full_frame"] = "bit-exact (%d searches, full-pel + sub-pel)" % work.per_frame \
                if bad == 0 else "MISMATCH in %d results" % bad
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
