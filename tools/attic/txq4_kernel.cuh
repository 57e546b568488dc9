// NOT BUILT -- experiment record (round 1).  Generalised warp-tile kernel (CPL columns / RPL rows per
// lane, slot-fastest lane mapping).  Bit-exact (107 GPU tests) but slower than txq_kernel.cuh on
// B200: its lane mapping makes every global store touch one L1 wavefront per 16 bytes (8x8: 0.51
// vs 0.67 of the HBM roofline).  Kept for the measurements quoted in DESIGN.md section 6.
// txq4_kernel.cuh -- fused (subtract +) forward 2-D transform (+ quantise) (+ block error)
// kernel, version 4 (mid-size transforms).
//
// Same contract, arguments and bit-exact results as txq_kernel.cuh; the warp-tile mapping is
// generalised so that a lane carries 16..32 samples through each pass instead of 4..8:
//   * a lane owns CPL adjacent columns in the column pass and RPL adjacent rows in the
//     row pass; a block is served by LC = W/CPL lanes (columns) and LR = Hk/RPL lanes
//     (rows) and a warp tile holds TB = 32/min(LC, LR) blocks.  The per-tile bookkeeping
//     (descriptors, addresses, staging, reductions) is thus amortised over 2-4x more
//     samples than in the one-column-per-lane kernel;
//   * lane -> (slot = lane % TB, group = lane / TB): a lane serves the same block in every
//     phase, and with the pitches of tools/bank_search.py the packed column loads
//     (LDS.32/64 of CPL int16), the vector column stores (STS.64/128) and the LDS.128
//     row loads are all free of bank conflicts;
//   * adjacent rows are adjacent in the codec's column-major coefficient storage, so the
//     RPL results of a lane leave as one 8/16-byte store per coefficient column;
//   * the tile's descriptors and quantiser rows travel through the cp.async ring with the
//     samples (the column / row phases never wait on a dependent global load) and the
//     descriptor for the staging step is fetched one iteration ahead;
//   * the inverse scan tables sit in shared memory.
// Used where CI == RI == 1 (every pass is one round); the 64-point sizes stay on
// txq_kernel.cuh.
#pragma once
#include "txq_kernel.cuh"

namespace b200 {

// (CPL, RPL) per size; mirrored by tools/bank_search.py
template <int W, int H>
struct Tx4Map {
  static constexpr int CPL = (W == 8 && H == 8)     ? 2
                             : (W == 16 && H == 4)  ? 4
                             : (W == 8 && H == 16)  ? 2
                             : (W == 16 && H == 8)  ? 4
                             : (W == 16 && H == 16) ? 2
                             : (W == 32 && H == 8)  ? 4
                             : (W == 32 && H == 16) ? 2
                             : (W == 8 && H == 4)   ? 2
                                                    : 1;
  static constexpr int RPL = (W == 8 && H == 8)     ? 2
                             : (W == 4 && H == 16)  ? 4
                             : (W == 8 && H == 16)  ? 4
                             : (W == 16 && H == 8)  ? 2
                             : (W == 16 && H == 16) ? 2
                             : (W == 8 && H == 32)  ? 4
                             : (W == 16 && H == 32) ? 2
                             : (W == 4 && H == 8)   ? 2
                                                    : 1;
};

// shared-memory pitches chosen by tools/bank_search.py: transpose-tile row pitch (words),
// extra words per slot of the transpose tile, extra bytes per slot of the staging area for
// 1- and 2-byte samples.  With these every LDS / STS of the kernel is conflict-free (the
// LDS.128 row loads of 16x16 / 8x32 and the column stores of 16x32 take two wavefronts).
struct Pitch4 {
  int ldx, xs_pad, slot_pad8, slot_pad16;
};
__host__ __device__ constexpr Pitch4 pitch4_of(int w, int h) {
  switch (w * 100 + h) {
    case 808: return Pitch4{ 8, 4, 8, 16 };
    case 416: return Pitch4{ 4, 4, 4, 8 };
    case 1604: return Pitch4{ 16, 4, 16, 16 };
    case 816: return Pitch4{ 8, 4, 8, 16 };
    case 1608: return Pitch4{ 16, 4, 16, 16 };
    case 1616: return Pitch4{ 16, 8, 16, 32 };
    case 832: return Pitch4{ 8, 8, 8, 16 };
    case 3208: return Pitch4{ 36, 8, 32, 32 };
    case 1632: return Pitch4{ 20, 4, 16, 32 };
    case 3216: return Pitch4{ 36, 16, 32, 64 };
    case 3232: return Pitch4{ 36, 0, 0, 0 };
    case 408: return Pitch4{ 4, 4, 4, 8 };
    case 804: return Pitch4{ 8, 4, 8, 16 };
    case 404: return Pitch4{ 4, 4, 0, 8 };
    default: return Pitch4{ w + 4, 0, 0, 0 };
  }
}

template <int W, int H>
struct Tx4Cfg {
  using C = TxCfg<W, H>;
  static constexpr int CPL = Tx4Map<W, H>::CPL, RPL = Tx4Map<W, H>::RPL;
  static constexpr int Wk = C::Wk, Hk = C::Hk;
  static constexpr int LC = W / CPL, LR = Hk / RPL;
  static constexpr int TB = 32 / cx_min(LC, LR);
  static_assert(TB * LC == 32 && TB * LR == 32, "v4 serves sizes whose passes are one round each");
  // transpose tile: row pitch LDX, slot pitch XS (words), see tools/bank_search.py
  static constexpr int LDX = pitch4_of(W, H).ldx;
  static constexpr int XS = Hk * LDX + pitch4_of(W, H).xs_pad;
};

template <int W, int H, int ESZ>
struct Stage4Cfg {
  using C4 = Tx4Cfg<W, H>;
  static constexpr int kRowBytes = W * ESZ;
  static constexpr int kChunk = cx_min(16, kRowBytes);
  static constexpr int kCpr = kRowBytes / kChunk;
  static constexpr int kCpb = H * kCpr;
  static constexpr int kChunksPerTile = C4::TB * kCpb;
  static_assert(kChunksPerTile % 32 == 0, "every lane copies the same number of chunks");
  static constexpr int kCpl = kChunksPerTile / 32;
  static constexpr int kTileBytes = H * kRowBytes;
  static constexpr int kSlotBytes =
      kTileBytes + (ESZ == 1 ? pitch4_of(W, H).slot_pad8 : pitch4_of(W, H).slot_pad16);
  static constexpr int kPlaneBytes = cx_round_up(C4::TB * kSlotBytes, 16);
};

template <int W, int H, int IN>
struct Smem4Cfg {
  using C4 = Tx4Cfg<W, H>;
  static constexpr int ESZ = IN == B200_IN_PIXELS8 ? 1 : 2;
  using S = Stage4Cfg<W, H, ESZ>;
  static constexpr int kOperands = IN == B200_IN_RESIDUAL ? 1 : 2;
  static constexpr int kDescOff = kOperands * S::kPlaneBytes;        // TB x 16 B descriptors
  static constexpr int kQrowOff = kDescOff + C4::TB * 16;             // TB x 32 B quantiser rows
  static constexpr int kStageBytes = kQrowOff + C4::TB * 32;
  static constexpr int kStages = S::kPlaneBytes >= 2048 ? 2 : 3;
  static constexpr int kXposeBytes = C4::TB * C4::XS * 4;
  static constexpr int kWarpBytes = cx_round_up(kStages * kStageBytes + kXposeBytes, 16);
  static constexpr int kIscanBytes = cx_round_up(3 * C4::Wk * C4::Hk * 2, 16);
  static constexpr int kCtaBytes = kIscanBytes + kWarpsPerCta * kWarpBytes;
};

// ------------------------------------------------------------------- staging -----
__device__ __forceinline__ void cp_async_ca16(uint32_t dst, const void *src) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ B200TxBlock decode_desc(const uint4 &raw) {
  B200TxBlock d;
  memcpy(&d, &raw, sizeof(d));
  return d;
}
// raw descriptor of the block this lane serves in warp-tile wt (zeros past the end).  Only
// the words the staging step needs are loaded (src_offset; stride/qrow/size/type): a
// 128-bit load would leave a dead destination register that the compiler recycles at
// once, stalling the warp on the write-after-write hazard.
template <int TB>
__device__ __forceinline__ uint4 fetch_desc4(const TxqArgs &a, int wt, int n_tiles, int lane) {
  const int bi = wt * TB + lane % TB;
  uint4 r = make_uint4(0, 0, 0, 0);
  if (wt < n_tiles && bi < a.n_blocks) {
    const uint32_t *p = reinterpret_cast<const uint32_t *>(a.blocks + bi);
    r.x = __ldg(p);
    const uint2 zw = __ldg(reinterpret_cast<const uint2 *>(p + 2));
    r.z = zw.x, r.w = zw.y;
  }
  return r;
}

// Issue the cp.async copies of warp-tile `wt` into `stage`: the samples of its TB blocks
// (upside-down for FLIPADST columns), their descriptors and their quantiser rows.
template <int W, int H, int IN>
__device__ __forceinline__ void stage_tile4(const TxqArgs &a, int wt, const uint4 &draw,
                                            unsigned char *stage, int lane) {
  using C4 = Tx4Cfg<W, H>;
  using M = Smem4Cfg<W, H, IN>;
  using S = typename M::S;
  constexpr int ESZ = M::ESZ, TB = C4::TB;
  const int slot = lane % TB, bi = wt * TB + slot;
  if (bi >= a.n_blocks) return;
#ifdef B200_DBG_NO_STAGE
  return;
#endif
  const B200TxBlock d = decode_desc(draw);
  const bool flip = vtx_of(d.tx_type & 15) == 2;  // ud_flip (av1_fwd_txfm2d.c:86)
  auto copy_plane = [&](const void *plane, uint32_t offset, int stride, unsigned char *dst0) {
    const unsigned char *g0 = static_cast<const unsigned char *>(plane) + (size_t)offset * ESZ;
    const int pitch = stride * ESZ;
    const bool aligned = ((reinterpret_cast<uintptr_t>(g0) | (uintptr_t)pitch) & (S::kChunk - 1)) == 0;
#pragma unroll
    for (int i = 0; i < S::kCpl; ++i) {
      const int rem = lane / TB + (32 / TB) * i;
      const int row = rem / S::kCpr, part = rem % S::kCpr;
      const int drow = flip ? H - 1 - row : row;
      unsigned char *dst = dst0 + drow * S::kRowBytes + part * S::kChunk;
      const unsigned char *g = g0 + (size_t)row * pitch + part * S::kChunk;
      if (aligned) {
        cp_async<S::kChunk>(smem_u32(dst), g);
      } else {  // unaligned source: element-wise copy (rare)
#pragma unroll
        for (int e = 0; e < S::kChunk; ++e) dst[e] = g[e];
      }
    }
  };
  copy_plane(a.src, d.src_offset, d.src_stride, stage + slot * S::kSlotBytes);
  if constexpr (IN != B200_IN_RESIDUAL) {
    const B200TxPred pd = a.preds[bi];
    copy_plane(a.pred, pd.pred_offset, pd.pred_stride, stage + S::kPlaneBytes + slot * S::kSlotBytes);
  }
  if (lane < TB) {  // one lane per block forwards the descriptor and the quantiser row
    // through L1 (.ca): the few quantiser rows are shared by every block of the launch and
    // an L1-bypassing copy would hammer a handful of L2 sectors from all SMs
    cp_async_ca16(smem_u32(stage + M::kDescOff + slot * 16), a.blocks + bi);
    if (a.q_kind != B200_Q_NONE) {
      const unsigned char *q = reinterpret_cast<const unsigned char *>(a.qrows + d.qrow);
      cp_async_ca16(smem_u32(stage + M::kQrowOff + slot * 32), q);
      cp_async_ca16(smem_u32(stage + M::kQrowOff + slot * 32 + 16), q + 16);
    }
  }
}

// ---------------------------------------------------------------- column pass ----
// Read CPL adjacent columns (c0 ..) of the staged block: x[cc*H + r], up-shifted.  Returns
// whether all residuals read are inside the fast-arithmetic guard.
template <int W, int H, int IN>
__device__ __forceinline__ bool load_columns4(const unsigned char *stg, int c0, int32_t *x) {
  using C = TxCfg<W, H>;
  constexpr int CPL = Tx4Cfg<W, H>::CPL;
  constexpr int PB = Smem4Cfg<W, H, IN>::S::kPlaneBytes;
  auto up = [](int32_t v) { return (int32_t)((uint32_t)v << C::kUp); };
  if constexpr (IN == B200_IN_PIXELS8) {
#pragma unroll
    for (int r = 0; r < H; ++r) {
#pragma unroll
      for (int cc = 0; cc < CPL; ++cc)
        x[cc * H + r] = up((int)stg[r * W + c0 + cc] - (int)stg[PB + r * W + c0 + cc]);
    }
    return true;
  } else if constexpr (CPL == 1) {
    int32_t mx = 0, mn = 0;
#pragma unroll
    for (int r = 0; r < H; ++r) {
      int32_t v;
      if constexpr (IN == B200_IN_RESIDUAL) {
        v = reinterpret_cast<const int16_t *>(stg)[r * W + c0];
      } else {
        v = (int16_t)((int)reinterpret_cast<const uint16_t *>(stg)[r * W + c0] -
                      (int)reinterpret_cast<const uint16_t *>(stg + PB)[r * W + c0]);
      }
      mx = v > mx ? v : mx;
      mn = v < mn ? v : mn;
      x[r] = up(v);
    }
    return mx <= kGuard && mn >= -kGuard;
  } else {
    uint32_t mx = 0, mn = 0;  // packed int16 pairs
#pragma unroll
    for (int r = 0; r < H; ++r) {
      uint32_t w[CPL / 2];
      const unsigned char *p = stg + (r * W + c0) * 2;
      if constexpr (CPL == 2) {
        w[0] = *reinterpret_cast<const uint32_t *>(p);
        if constexpr (IN == B200_IN_PIXELS16) w[0] = __vsub2(w[0], *reinterpret_cast<const uint32_t *>(p + PB));
      } else {
        const uint2 t = *reinterpret_cast<const uint2 *>(p);
        w[0] = t.x, w[1] = t.y;
        if constexpr (IN == B200_IN_PIXELS16) {
          const uint2 q = *reinterpret_cast<const uint2 *>(p + PB);
          w[0] = __vsub2(w[0], q.x), w[1] = __vsub2(w[1], q.y);
        }
      }
#pragma unroll
      for (int k = 0; k < CPL / 2; ++k) {
        mx = __vmaxs2(mx, w[k]);
        mn = __vmins2(mn, w[k]);
        x[(2 * k) * H + r] = up((int32_t)(int16_t)(w[k] & 0xffff));
        x[(2 * k + 1) * H + r] = up((int32_t)w[k] >> 16);
      }
    }
    const uint32_t hi = __vmaxs2(mx, __byte_perm(mx, 0, 0x1032));
    const uint32_t lo = __vmins2(mn, __byte_perm(mn, 0, 0x1032));
    return (int16_t)(hi & 0xffff) <= kGuard && (int16_t)(lo & 0xffff) >= -kGuard;
  }
}

// Column transforms of the lane's CPL columns -> transpose tile of the block (vector stores;
// lr_flip of FLIPADST rows mirrors the column index, av1_fwd_txfm2d.c:99-105).
template <int W, int H, bool EXACT>
__device__ __forceinline__ void column_pass4(int vt, int ht, int32_t *x, int32_t *tile, int c0) {
  using C = TxCfg<W, H>;
  using C4 = Tx4Cfg<W, H>;
  constexpr int CPL = C4::CPL, Hk = C4::Hk;
  int32_t y[CPL * Hk];
#pragma unroll
  for (int cc = 0; cc < CPL; ++cc) txfm1d<H, C::kCosCol, EXACT, Hk>(vt, x + cc * H, y + cc * Hk);
  const bool flip = ht == 2;
  int32_t *dst = tile + (flip ? W - CPL - c0 : c0);
#pragma unroll
  for (int r = 0; r < Hk; ++r) {
    int32_t v[CPL];
#pragma unroll
    for (int e = 0; e < CPL; ++e) v[e] = down_shift<EXACT, C::kDown1>(y[e * Hk + r]);
    if constexpr (CPL == 1) {
      dst[r * C4::LDX] = v[0];
    } else if constexpr (CPL == 2) {
      *reinterpret_cast<int2 *>(dst + r * C4::LDX) = flip ? make_int2(v[1], v[0]) : make_int2(v[0], v[1]);
    } else {
      *reinterpret_cast<int4 *>(dst + r * C4::LDX) =
          flip ? make_int4(v[3], v[2], v[1], v[0]) : make_int4(v[0], v[1], v[2], v[3]);
    }
  }
}

// ------------------------------------------------------------------- epilogue ----
template <int R, bool VEC>
__device__ __forceinline__ void store_rows(int32_t *p, const int32_t *v) {
  if constexpr (R == 1 || !VEC) {
#pragma unroll
    for (int j = 0; j < R; ++j) p[j] = v[j];
  } else if constexpr (R == 2) {
    *reinterpret_cast<int2 *>(p) = make_int2(v[0], v[1]);
  } else {
    *reinterpret_cast<int4 *>(p) = make_int4(v[0], v[1], v[2], v[3]);
  }
}

// Quantise the lane's RPL rows (t0 .. t0+RPL-1): y[j*Wk + c] is coefficient (c, t0 + j).
// qc / dqc point at the lane's first output element; element (c, t0 + j) is at [c*Hk + j].
template <int W, int H, int QK, bool ERR, bool VEC>
__device__ __forceinline__ void quant_epilogue4(const B200QRow &qr, int t0, const int32_t *y,
                                                int32_t *qc, int32_t *dqc, int *last_c,
                                                int64_t &err, int64_t &ssz) {
  using C = TxCfg<W, H>;
  using C4 = Tx4Cfg<W, H>;
  constexpr int RPL = C4::RPL, Wk = C4::Wk, Hk = C4::Hk;
  constexpr int LS = C::kLogScale;
  constexpr bool fp = QK == B200_Q_FP || QK == B200_Q_FP_HBD;
  constexpr bool hbd = QK == B200_Q_FP_HBD || QK == B200_Q_B_HBD;
  auto rpo2 = [](int v) { return LS ? ((v + (1 << (LS - 1))) >> LS) : v; };  // ROUND_POWER_OF_TWO
  const int rnd0 = rpo2(fp ? qr.round_fp[0] : qr.round[0]);
  const int rnd1 = rpo2(fp ? qr.round_fp[1] : qr.round[1]);
  const int qnt0 = fp ? qr.quant_fp[0] : qr.quant[0];
  const int qnt1 = fp ? qr.quant_fp[1] : qr.quant[1];
  const int deq0 = qr.dequant[0], deq1 = qr.dequant[1];
  const int qsh0 = qr.quant_shift[0], qsh1 = qr.quant_shift[1];
  // dead-zone test rewritten as |coeff| >= threshold (see txq_kernel.cuh)
  constexpr int K = 1 + LS;
  const int thr0 = fp ? ((deq0 + (1 << K) - 1) >> K) : rpo2(qr.zbin[0]);
  const int thr1 = fp ? ((deq1 + (1 << K) - 1) >> K) : rpo2(qr.zbin[1]);
#pragma unroll
  for (int c = 0; c < Wk; ++c) {
    int32_t qs[RPL], dqs[RPL];
#pragma unroll
    for (int j = 0; j < RPL; ++j) {
      const bool dc = (c == 0) && (j == 0) && (t0 == 0);  // compile-time false unless c == j == 0
      const int rnd_k = dc ? rnd0 : rnd1, qnt_k = dc ? qnt0 : qnt1, deq_k = dc ? deq0 : deq1;
      const int qsh_k = dc ? qsh0 : qsh1, thr_k = dc ? thr0 : thr1;
      (void)qsh_k;
      const int32_t cf = y[j * Wk + c];
      const int32_t ab = cf < 0 ? -cf : cf;
      int32_t q;
      if constexpr (fp) {
        if constexpr (!hbd) {
          // clamp64(abs + round, INT16_MIN, INT16_MAX): abs, round >= 0 -> one upper clamp
          int32_t tt = ab + rnd_k;
          tt = tt > 32767 ? 32767 : tt;
          tt = ab >= thr_k ? tt : 0;
          q = (tt * qnt_k) >> (16 - LS);
        } else {
          const uint32_t tt = ab >= thr_k ? (uint32_t)(ab + rnd_k) : 0u;
          q = (int32_t)(((uint64_t)tt * (uint32_t)qnt_k) >> (16 - LS));
        }
      } else {
        int64_t tt = (int64_t)ab + rnd_k;
        if constexpr (!hbd) tt = tt > 32767 ? 32767 : tt;
        tt = ab >= thr_k ? tt * 32 : 0;
        q = (int32_t)(((((tt * qnt_k) >> 16) + tt) * qsh_k) >> (16 - LS + 5));
      }
      const int32_t dq = wmul(q, deq_k) >> LS;
      const int32_t sg = cf >> 31;  // 0 or -1
      qs[j] = (q ^ sg) - sg;
      dqs[j] = (dq ^ sg) - sg;
      if (q) last_c[j] = c;  // iscan grows with c along a row: the last non-zero c wins
      if constexpr (ERR) {
        const int32_t df = wsub(cf, dqs[j]);
        if constexpr (hbd) {
          err += (int64_t)df * df;
          ssz += (int64_t)cf * cf;
        } else {  // av1_block_error_c squares in 32-bit int (rdopt.c:642-643)
          err += (int64_t)wmul(df, df);
          ssz += (int64_t)wmul(cf, cf);
        }
      }
    }
    store_rows<RPL, VEC>(qc + c * Hk, qs);
    store_rows<RPL, VEC>(dqc + c * Hk, dqs);
  }
}

// --------------------------------------------------------------------- kernel ----
template <int W, int H, int IN>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 5) txq4_kernel(const TxqArgs a) {
  using C = TxCfg<W, H>;
  using C4 = Tx4Cfg<W, H>;
  using M = Smem4Cfg<W, H, IN>;
  using S = typename M::S;
  constexpr int NS = M::kStages, TB = C4::TB, RPL = C4::RPL, CPL = C4::CPL;
  constexpr int Wk = C4::Wk, Hk = C4::Hk;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int16_t *s_iscan = reinterpret_cast<int16_t *>(smem_raw);
  unsigned char *wbase = smem_raw + M::kIscanBytes + warp * M::kWarpBytes;
  int32_t *xpose = reinterpret_cast<int32_t *>(wbase + NS * M::kStageBytes);
  const int n_tiles = (a.n_blocks + TB - 1) / TB;
  const int wstride = gridDim.x * kWarpsPerCta;
  const int wt0 = blockIdx.x * kWarpsPerCta + warp;
  const bool force_exact = (a.flags & B200_TXQ_FORCE_EXACT) != 0;
  const bool quant = a.q_kind != B200_Q_NONE;
  const bool want_err = (a.flags & B200_TXQ_BLOCK_ERROR) != 0;
  const bool want_coeff = (a.flags & B200_TXQ_STORE_COEFF) != 0;
  const int slot = lane % TB, grp = lane / TB;  // the lane's block and column / row group

  if (quant)
    for (int i = threadIdx.x; i < 3 * Wk * Hk; i += blockDim.x) s_iscan[i] = a.iscan3[i];
  __syncthreads();

  // prologue: fill NS-1 stages; keep the descriptor of the next tile to stage in registers
#pragma unroll
  for (int s = 0; s < NS - 1; ++s) {
    const int wt = wt0 + s * wstride;
    if (wt < n_tiles)
      stage_tile4<W, H, IN>(a, wt, fetch_desc4<TB>(a, wt, n_tiles, lane), wbase + s * M::kStageBytes, lane);
    cp_async_commit();
  }
  uint4 dpre = fetch_desc4<TB>(a, wt0 + (NS - 1) * wstride, n_tiles, lane);

  int it = 0, st = 0;  // st = it % NS
  for (int wt = wt0; wt < n_tiles; wt += wstride, ++it) {
    {  // keep the ring full: tile it+NS-1 goes to the stage consumed in the previous round
      const int wtp = wt + (NS - 1) * wstride;
      const int sp = st == 0 ? NS - 1 : st - 1;
      if (wtp < n_tiles) stage_tile4<W, H, IN>(a, wtp, dpre, wbase + sp * M::kStageBytes, lane);
      cp_async_commit();
      dpre = fetch_desc4<TB>(a, wtp + wstride, n_tiles, lane);
    }
    cp_async_wait<NS - 1>();
    __syncwarp();
    const unsigned char *stage = wbase + st * M::kStageBytes;
    st = st + 1 == NS ? 0 : st + 1;
    const int bi = wt * TB + slot;
#ifdef B200_DBG_NO_COMPUTE
    const bool act = false;
#else
    const bool act = bi < a.n_blocks;
#endif
    B200TxBlock d;
    {
      const uint4 raw = act ? *reinterpret_cast<const uint4 *>(stage + M::kDescOff + slot * 16)
                            : make_uint4(0, 0, 0, 0);
      d = decode_desc(raw);
    }
    const int tx_type = d.tx_type & 15;
    const int vt = vtx_of(tx_type), ht = htx_of(tx_type);
    int32_t *tile = xpose + slot * C4::XS;

    // ---------------------------------------------------------------- column pass
    bool fast;
    {
      int32_t x[CPL * H];
      bool fits = true;
      if (act) fits = load_columns4<W, H, IN>(stage + slot * S::kSlotBytes, grp * CPL, x);
      fast = !force_exact;
      if constexpr (IN != B200_IN_PIXELS8) fast = fast && __all_sync(0xffffffffu, fits);
      if (act) {
        if (fast) column_pass4<W, H, false>(vt, ht, x, tile, grp * CPL);
        else column_pass4<W, H, true>(vt, ht, x, tile, grp * CPL);
      }
    }
    __syncwarp();

    // ------------------------------------------------------------------- row pass
    const int t0 = grp * RPL;
    int last_c[RPL];
#pragma unroll
    for (int j = 0; j < RPL; ++j) last_c[j] = -1;
    int64_t err = 0, ssz = 0;
    if (act) {
      int32_t y[RPL * Wk];
#pragma unroll
      for (int j = 0; j < RPL; ++j) {
        const int32_t *xrow = tile + (t0 + j) * C4::LDX;
        if (fast) row_pass<W, H, false>(ht, xrow, y + j * Wk);
        else row_pass<W, H, true>(ht, xrow, y + j * Wk);
      }
      const size_t obase = (size_t)d.out_offset + t0;
      uintptr_t al = 0;
      if (want_coeff) al |= reinterpret_cast<uintptr_t>(a.coeff + obase);
      if (quant) al |= reinterpret_cast<uintptr_t>(a.qcoeff + obase) | reinterpret_cast<uintptr_t>(a.dqcoeff + obase);
      const bool vec = (al & (RPL * 4 - 1)) == 0;
      if (want_coeff) {
#pragma unroll
        for (int c = 0; c < Wk; ++c) {
          int32_t v[RPL];
#pragma unroll
          for (int j = 0; j < RPL; ++j) v[j] = y[j * Wk + c];
          if (vec) store_rows<RPL, true>(a.coeff + obase + c * Hk, v);
          else store_rows<RPL, false>(a.coeff + obase + c * Hk, v);
        }
      }
      if (quant) {
        B200QRow qr;
        {
          const uint4 q0 = *reinterpret_cast<const uint4 *>(stage + M::kQrowOff + slot * 32);
          const uint4 q1 = *reinterpret_cast<const uint4 *>(stage + M::kQrowOff + slot * 32 + 16);
          memcpy(&qr, &q0, 16);
          memcpy(reinterpret_cast<char *>(&qr) + 16, &q1, 16);
        }
        int32_t *qc = a.qcoeff + obase, *dqc = a.dqcoeff + obase;
        auto run = [&](auto qk, auto we) {
          constexpr int QK = decltype(qk)::value;
          constexpr bool WE = decltype(we)::value;
          if (vec) quant_epilogue4<W, H, QK, WE, true>(qr, t0, y, qc, dqc, last_c, err, ssz);
          else quant_epilogue4<W, H, QK, WE, false>(qr, t0, y, qc, dqc, last_c, err, ssz);
        };
        using T = std::true_type;
        using F = std::false_type;
        if (a.q_kind == B200_Q_FP) {
          if (want_err) run(std::integral_constant<int, B200_Q_FP>{}, T{});
          else run(std::integral_constant<int, B200_Q_FP>{}, F{});
        } else if (a.q_kind == B200_Q_FP_HBD) {
          if (want_err) run(std::integral_constant<int, B200_Q_FP_HBD>{}, T{});
          else run(std::integral_constant<int, B200_Q_FP_HBD>{}, F{});
        } else if (a.q_kind == B200_Q_B) {
          if (want_err) run(std::integral_constant<int, B200_Q_B>{}, T{});
          else run(std::integral_constant<int, B200_Q_B>{}, F{});
        } else {
          if (want_err) run(std::integral_constant<int, B200_Q_B_HBD>{}, T{});
          else run(std::integral_constant<int, B200_Q_B_HBD>{}, F{});
        }
      }
    }
    if (quant) {
      int eob = 0;
      {
        const int cls = d.tx_type < 10 ? 0 : ((d.tx_type & 1) ? 2 : 1);
        const int16_t *isc = s_iscan + cls * (Wk * Hk) + t0;
#pragma unroll
        for (int j = 0; j < RPL; ++j) {
          if (last_c[j] >= 0) {
            const int e = isc[last_c[j] * Hk + j] + 1;
            eob = e > eob ? e : eob;
          }
        }
      }
      // reduce over the lanes that serve the block: lane = slot + TB * group
#pragma unroll
      for (int o = 16; o >= TB; o >>= 1) {
        const int e2 = __shfl_xor_sync(0xffffffffu, eob, o);
        eob = e2 > eob ? e2 : eob;
      }
      if (want_err) {
#pragma unroll
        for (int o = 16; o >= TB; o >>= 1) {
          err += __shfl_xor_sync(0xffffffffu, err, o);
          ssz += __shfl_xor_sync(0xffffffffu, ssz, o);
        }
      }
      if (act && grp == 0) {
        a.eob[bi] = (uint16_t)eob;
        if (want_err && a.dist) {
          if (a.q_kind == B200_Q_FP_HBD || a.q_kind == B200_Q_B_HBD) {
            const int sh = 2 * (a.bd - 8);
            const int64_t r = sh > 0 ? ((int64_t)1 << (sh - 1)) : 0;
            err = (err + r) >> sh;
            ssz = (ssz + r) >> sh;
          }
          a.dist[2 * (size_t)bi] = err;
          a.dist[2 * (size_t)bi + 1] = ssz;
        }
      }
    }
    __syncwarp();  // stage and transpose tile are reused by the next round
  }
  cp_async_wait<0>();
}

template <int W, int H, int IN>
TxqKernelInfo txq4_kernel_info() {
  TxqKernelInfo k;
  k.fn = txq4_kernel<W, H, IN>;
  k.smem_bytes = Smem4Cfg<W, H, IN>::kCtaBytes;
  k.bpw = Tx4Cfg<W, H>::TB;
  return k;
}

}  // namespace b200
