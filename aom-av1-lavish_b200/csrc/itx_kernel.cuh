// itx_kernel.cuh -- inverse 2-D transform + reconstruction + pixel-domain distortion for a whole
// batch of same-size transform blocks in one launch, version 2.
//
// Replaces the reference's per-block chain of the tx-type search
//   av1_inverse_transform_block    av1/common/idct.c:303  (av1_inv_txfm_add_c :283,
//                                  av1_highbd_inv_txfm_add_c :216 -> av1_inv_txfm2d_add_WxH_c,
//                                  av1/common/av1_inv_txfm2d.c:234-326)
//   dist_block_px_domain           av1/encoder/tx_search.c:969-1022 (pixel_dist_visible_only :920)
// so that a speculative evaluation of several tx types can stay on the device and only
// {eob, rate inputs, distortion} cross PCIe.
//
// Mapping: the forward kernel's warp-tile geometry run the other way round.  A warp owns TB =
// 32 / min(W, Hk) blocks and strides over the tiles of the launch.  Per tile:
//   1. the blocks' dequantised coefficients (contiguous in the codec's column-major storage) and, for
//      the sizes below 64 points, their prediction and source rows arrive in shared memory through
//      cp.async, a ring of two stages filled one tile ahead (version 1 loaded them synchronously and
//      was latency bound: 11-28 long-scoreboard cycles per issued instruction, profiles/r2_itx_v1_*);
//   2. row rounds: lane <-> (block, kept row t): W-point inverse row transform in registers, results
//      (left-right flipped for FLIPADST rows) into the padded per-warp int32 tile;
//   3. column rounds: lane <-> (block, column c): H-point inverse column transform, the H rounded
//      residuals written back into the lane's own tile column (upside down for FLIPADST columns);
//   4. pixel rounds: lane <-> (block, row): prediction row + residual row, clip, packed 4..16-byte
//      stores of the reconstruction, squared difference against the source row; the block's sum is a
//      shuffle reduction over its rows.
// __syncwarp only -- no CTA-wide barrier.
#pragma once
#include "itxfm1d.cuh"
#include "txq_kernel.cuh"

namespace b200 {

struct ItxArgs {
  const B200TxBlock *blocks;
  const B200TxPred *preds;
  const int32_t *dqcoeff;
  const uint16_t *eob;  // optional
  const void *src;      // optional (distortion)
  const void *pred;
  void *recon;          // optional
  int64_t *dist;        // optional
  int n_blocks;
  int bd;
};

// av1_inv_txfm_shift_ls[tx_size][0] as a down-shift (av1_inv_txfm2d.c:133-159); [1] is always 4
template <int W, int H>
struct ItxCfg {
  static constexpr int kMax = cx_max(W, H), kMin = cx_min(W, H);
  static constexpr int kRowShift =
      (W == H) ? (W == 4 ? 0 : (W == 8 ? 1 : 2))
      : (kMax == 2 * kMin) ? (kMax == 8 ? 0 : 1)   /* 4x8 8x4: 0; 8x16 .. 64x32: 1 */
                           : (kMax == 16 ? 1 : 2); /* 4x16 16x4: 1; 8x32 32x8 16x64 64x16: 2 */
};

template <int W, int H, int PX>
struct ItxSmem {
  using C = TxCfg<W, H>;
  using S = StageCfg<W, H, PX>;
  static constexpr int kStages = 2;
  // coefficient slot pitch (words): + Hk so that the (slot, row) lanes of a row round hit distinct banks
  static constexpr int kDqPitch = C::Wk * C::Hk + (C::TB > 1 ? C::Hk : 0);
  static constexpr int kDqBytes = C::TB * kDqPitch * 4;
  // pixels are staged below 64 points only (a 64x64 tile pair is 8-16 KB per stage: the ring would halve
  // the resident warps, as it did in the forward kernel)
  static constexpr bool kStagePix = cx_max(W, H) < 64;
  static constexpr int kPixBytes = kStagePix ? cx_round_up(2 * S::kPlaneBytes, 16) : 0;
  static constexpr int kStageBytes = cx_round_up(kDqBytes + kPixBytes, 16);
  // the tile holds the row-pass results (Hk rows) and then the H residual rows of the column pass
  static constexpr int XS = (C::TB > 1 && W < 32) ? cx_round_up(H * C::LDX, 32) + W : H * C::LDX;
  static constexpr int kXposeBytes = cx_round_up(C::TB * XS * 4, 16);
  static constexpr int kWarpBytes = cx_round_up(kStages * kStageBytes + kXposeBytes, 128);
  static constexpr int kCtaBytes = kWarpsPerCta * kWarpBytes;
};

// cp.async copies of one tile into `stage`: coefficients of every coded block, then (staged sizes) the
// prediction and source rows.
template <int W, int H, int PX>
__device__ __forceinline__ void itx_stage_tile(const ItxArgs &a, int wt, unsigned char *stage, int lane) {
  using C = TxCfg<W, H>;
  using M = ItxSmem<W, H, PX>;
  using S = typename M::S;
  constexpr int kDqChunks = C::Wk * C::Hk / 4;  // 16-byte chunks per block
  constexpr int kDqTotal = C::TB * kDqChunks;
#pragma unroll
  for (int i = 0; i < (kDqTotal + 31) / 32; ++i) {
    const int k = lane + 32 * i;
    if (kDqTotal % 32 != 0 && k >= kDqTotal) break;
    const int slot = k / kDqChunks, part = k % kDqChunks;
    const int bi = wt * C::TB + slot;
    if (bi >= a.n_blocks) continue;
    if (a.eob && a.eob[bi] == 0) continue;
    const int32_t *g = a.dqcoeff + (size_t)a.blocks[bi].out_offset + 4 * part;
    unsigned char *dst = stage + (slot * M::kDqPitch + 4 * part) * 4;
    if ((reinterpret_cast<uintptr_t>(g) & 15) == 0) {
      cp_async<16>(smem_u32(dst), g);
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) reinterpret_cast<int32_t *>(dst)[e] = g[e];
    }
  }
  if constexpr (M::kStagePix) {
    unsigned char *pix = stage + M::kDqBytes;
#pragma unroll
    for (int i = 0; i < S::kCpl; ++i) {
      const int k = lane + 32 * i;
      if (S::kChunksPerTile % 32 != 0 && k >= S::kChunksPerTile) break;
      const int slot = k / S::kCpb, rem = k % S::kCpb;
      const int row = rem / S::kCpr, part = rem % S::kCpr;
      const int bi = wt * C::TB + slot;
      if (bi >= a.n_blocks) continue;
      const B200TxPred pd = a.preds[bi];
      unsigned char *dst = pix + slot * S::kSlotBytes + row * S::kRowBytes + part * S::kChunk;
      const unsigned char *gp = static_cast<const unsigned char *>(a.pred) +
                                ((size_t)pd.pred_offset + (size_t)row * pd.pred_stride) * PX + part * S::kChunk;
      if ((reinterpret_cast<uintptr_t>(gp) & (S::kChunk - 1)) == 0) {
        cp_async<S::kChunk>(smem_u32(dst), gp);
      } else {
#pragma unroll
        for (int e = 0; e < S::kChunk; ++e) dst[e] = gp[e];
      }
      if (a.src) {
        const B200TxBlock d = load_desc(a.blocks + bi);
        unsigned char *dsts = dst + S::kPlaneBytes;
        const unsigned char *gs = static_cast<const unsigned char *>(a.src) +
                                  ((size_t)d.src_offset + (size_t)row * d.src_stride) * PX + part * S::kChunk;
        if ((reinterpret_cast<uintptr_t>(gs) & (S::kChunk - 1)) == 0) {
          cp_async<S::kChunk>(smem_u32(dsts), gs);
        } else {
#pragma unroll
          for (int e = 0; e < S::kChunk; ++e) dsts[e] = gs[e];
        }
      }
    }
  }
}

// inverse row transform of kept row t of one block: coefficients from the staged block, results into the tile
template <int W, int H, bool WIDE>
__device__ __forceinline__ void itx_row(const int32_t *dq, int t, int ht, int bd, int32_t *tile_row) {
  using C = TxCfg<W, H>;
  int32_t x[C::Wk], y[W];
  const int32_t hi = (1 << (bd + 7)) - 1, lo = -hi - 1;  // clamp_buf(temp_in, W, bd + 8) = the row range
#pragma unroll
  for (int c = 0; c < C::Wk; ++c) {
    int32_t v = dq[c * C::Hk + t];
    if constexpr (C::kRect)  // round_shift((int64_t)input * NewInvSqrt2, NewSqrt2Bits)
      v = (int32_t)(((int64_t)v * 2896 + 2048) >> 12);
    x[c] = clamp_range(v, lo, hi);
  }
  itxfm1d<W, WIDE, C::Wk>(ht, x, y, lo, hi);
#pragma unroll
  for (int c = 0; c < W; ++c) y[c] = rshift_round<ItxCfg<W, H>::kRowShift>(y[c]);
  if (ht == 2) {  // lr_flip (av1_inv_txfm2d.c:298-302): the column pass reads column W-1-c as column c
#pragma unroll
    for (int c = 0; c < W; c += 4)
      *reinterpret_cast<int4 *>(tile_row + c) = make_int4(y[W - 1 - c], y[W - 2 - c], y[W - 3 - c], y[W - 4 - c]);
  } else {
#pragma unroll
    for (int c = 0; c < W; c += 4)
      *reinterpret_cast<int4 *>(tile_row + c) = make_int4(y[c], y[c + 1], y[c + 2], y[c + 3]);
  }
}

// NW consecutive 32-bit words of one pixel row (shared or global), widest loads the alignment allows
template <int NW>
__device__ __forceinline__ void load_row_words(const unsigned char *p, uint32_t *w) {
  if constexpr (NW >= 4) {
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
      for (int i = 0; i < NW; i += 4) {
        const uint4 t = *reinterpret_cast<const uint4 *>(p + 4 * i);
        w[i] = t.x, w[i + 1] = t.y, w[i + 2] = t.z, w[i + 3] = t.w;
      }
      return;
    }
  } else if constexpr (NW == 2) {
    if ((reinterpret_cast<uintptr_t>(p) & 7) == 0) {
      const uint2 t = *reinterpret_cast<const uint2 *>(p);
      w[0] = t.x, w[1] = t.y;
      return;
    }
  }
  if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
#pragma unroll
    for (int i = 0; i < NW; ++i) w[i] = reinterpret_cast<const uint32_t *>(p)[i];
  } else {
#pragma unroll
    for (int i = 0; i < NW; ++i)
      w[i] = (uint32_t)p[4 * i] | ((uint32_t)p[4 * i + 1] << 8) | ((uint32_t)p[4 * i + 2] << 16) |
             ((uint32_t)p[4 * i + 3] << 24);
  }
}
template <int NW>
__device__ __forceinline__ void store_row_words(unsigned char *p, const uint32_t *w) {
  if constexpr (NW >= 4) {
    if ((reinterpret_cast<uintptr_t>(p) & 15) == 0) {
#pragma unroll
      for (int i = 0; i < NW; i += 4) *reinterpret_cast<uint4 *>(p + 4 * i) = make_uint4(w[i], w[i + 1], w[i + 2], w[i + 3]);
      return;
    }
  } else if constexpr (NW == 2) {
    if ((reinterpret_cast<uintptr_t>(p) & 7) == 0) {
      *reinterpret_cast<uint2 *>(p) = make_uint2(w[0], w[1]);
      return;
    }
  }
  if ((reinterpret_cast<uintptr_t>(p) & 3) == 0) {
#pragma unroll
    for (int i = 0; i < NW; ++i) reinterpret_cast<uint32_t *>(p)[i] = w[i];
  } else {
#pragma unroll
    for (int i = 0; i < 4 * NW; ++i) p[i] = (unsigned char)(w[i >> 2] >> (8 * (i & 3)));
  }
}

template <int W, int H, int PX>
__global__ void __launch_bounds__(kWarpsPerCta * 32) itx_kernel(const ItxArgs a) {
  using C = TxCfg<W, H>;
  using M = ItxSmem<W, H, PX>;
  using S = typename M::S;
  constexpr int NS = M::kStages;
  constexpr int NW = W * PX / 4;               // 32-bit words per pixel row
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char *wbase = smem_raw + warp * M::kWarpBytes;
  int32_t *xpose = reinterpret_cast<int32_t *>(wbase + NS * M::kStageBytes);
  const int n_tiles = (a.n_blocks + C::TB - 1) / C::TB;
  const int wstride = gridDim.x * kWarpsPerCta;
  const int wt0 = blockIdx.x * kWarpsPerCta + warp;
  const int pmax = (1 << a.bd) - 1;
  // column pass: clamp_buf(temp_in, H, max(bd + 6, 16)) = the column stage range
  const int32_t chi = (1 << (a.bd == 12 ? 17 : 15)) - 1, clo = -chi - 1;

  if (wt0 < n_tiles) itx_stage_tile<W, H, PX>(a, wt0, wbase, lane);
  cp_async_commit();

  int it = 0;
  for (int wt = wt0; wt < n_tiles; wt += wstride, ++it) {
    {
      const int wtp = wt + wstride;
      if (wtp < n_tiles) itx_stage_tile<W, H, PX>(a, wtp, wbase + ((it + 1) & 1) * M::kStageBytes, lane);
      cp_async_commit();
    }
    cp_async_wait<1>();
    __syncwarp();
    const unsigned char *stage = wbase + (it & 1) * M::kStageBytes;
    const int32_t *dqs = reinterpret_cast<const int32_t *>(stage);

    // ------------------------------------------------------------------- row rounds
#pragma unroll 1
    for (int ri = 0; ri < C::RI; ++ri) {
      const int item = lane + 32 * ri;
      const int slot = item / C::Hk, t = item % C::Hk;
      const int bi = wt * C::TB + slot;
      if (bi < a.n_blocks && (!a.eob || a.eob[bi])) {
        const int ht = htx_of(a.blocks[bi].tx_type & 15);
        int32_t *row = xpose + slot * M::XS + t * C::LDX;
        if (a.bd == 12) itx_row<W, H, true>(dqs + slot * M::kDqPitch, t, ht, a.bd, row);
        else itx_row<W, H, false>(dqs + slot * M::kDqPitch, t, ht, a.bd, row);
      }
    }
    __syncwarp();

    // ---------------------------------------------------------------- column rounds
    // a lane reads its column's Hk row-pass results, then writes the column's H residuals back
#pragma unroll 1
    for (int ci = 0; ci < C::CI; ++ci) {
      const int item = lane + 32 * ci;
      const int slot = item / W, c = item % W;
      const int bi = wt * C::TB + slot;
      if (bi < a.n_blocks && (!a.eob || a.eob[bi])) {
        const int vt = vtx_of(a.blocks[bi].tx_type & 15);
        int32_t x[C::Hk], y[H];
        int32_t *col = xpose + slot * M::XS + c;
#pragma unroll
        for (int r = 0; r < C::Hk; ++r) x[r] = clamp_range(col[r * C::LDX], clo, chi);
        itxfm1d<H, false, C::Hk>(vt, x, y, clo, chi);
        if (vt == 2) {  // ud_flip (:313-318): output r of the transform lands on pixel row H-1-r
#pragma unroll
          for (int r = 0; r < H; ++r) col[(H - 1 - r) * C::LDX] = rshift_round<4>(y[r]);
        } else {
#pragma unroll
          for (int r = 0; r < H; ++r) col[r * C::LDX] = rshift_round<4>(y[r]);
        }
      }
    }
    __syncwarp();
    // ----------------------------------------------------------------- pixel rounds
    constexpr int PI = C::TB * H / 32;           // rounds of lane <-> (block, pixel row)
    constexpr int RPB = H > 32 ? H / 32 : 1;     // rounds per block
    unsigned sse_blk = 0;
#pragma unroll
    for (int pi = 0; pi < PI; ++pi) {
      const int item = lane + 32 * pi;
      const int slot = item / H, g = item % H;
      const int bi = wt * C::TB + slot;
      unsigned sse = 0;
      if (bi < a.n_blocks) {
        const B200TxBlock d = load_desc(a.blocks + bi);
        const B200TxPred pd = a.preds[bi];
        const bool coded = !a.eob || a.eob[bi];
        uint32_t pw[NW], sw[NW], ow[NW];
        const size_t poff = ((size_t)pd.pred_offset + (size_t)g * pd.pred_stride) * PX;
        if constexpr (M::kStagePix) {
          const unsigned char *prow = stage + M::kDqBytes + slot * S::kSlotBytes + g * S::kRowBytes;
          load_row_words<NW>(prow, pw);
          if (a.src) load_row_words<NW>(prow + S::kPlaneBytes, sw);
        } else {
          load_row_words<NW>(static_cast<const unsigned char *>(a.pred) + poff, pw);
          if (a.src)
            load_row_words<NW>(static_cast<const unsigned char *>(a.src) +
                                   ((size_t)d.src_offset + (size_t)g * d.src_stride) * PX, sw);
        }
        const int32_t *rrow = xpose + slot * M::XS + g * C::LDX;
        const int vis_w = d.reserved ? (d.reserved & 255) : W;
        const int vis_h = d.reserved ? (d.reserved >> 8) : H;
#pragma unroll
        for (int c4 = 0; c4 < W; c4 += 4) {
          int4 rv = make_int4(0, 0, 0, 0);
          if (coded) rv = *reinterpret_cast<const int4 *>(rrow + c4);
          const int res[4] = { rv.x, rv.y, rv.z, rv.w };
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int c = c4 + e;
            int p, sv = 0;
            if constexpr (PX == 1) {
              p = (pw[c >> 2] >> (8 * (c & 3))) & 255;
              if (a.src) sv = (sw[c >> 2] >> (8 * (c & 3))) & 255;
            } else {
              p = (pw[c >> 1] >> (16 * (c & 1))) & 0xffff;
              if (a.src) sv = (sw[c >> 1] >> (16 * (c & 1))) & 0xffff;
            }
            const int v = coded ? min(max(p + res[e], 0), pmax) : p;
            if constexpr (PX == 1) {
              if ((c & 3) == 0) ow[c >> 2] = (uint32_t)v;
              else ow[c >> 2] |= (uint32_t)v << (8 * (c & 3));
            } else {
              if ((c & 1) == 0) ow[c >> 1] = (uint32_t)v;
              else ow[c >> 1] |= (uint32_t)v << 16;
            }
            if (a.src && c < vis_w && g < vis_h) {
              const int df = sv - v;
              sse += (unsigned)(df * df);
            }
          }
        }
        if (a.recon) store_row_words<NW>(static_cast<unsigned char *>(a.recon) + poff, ow);
      }
      if (a.dist) {
        // a block's rows: min(H, 32) lanes of this round, over RPB consecutive rounds
        sse_blk = (pi % RPB == 0) ? sse : sse_blk + sse;
        if (pi % RPB == RPB - 1) {
          unsigned long long t = sse_blk;
#pragma unroll
          for (int o = cx_min(H, 32) / 2; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
          if (bi < a.n_blocks && (g % cx_min(H, 32)) == 0 && g / 32 == RPB - 1) {
            if (PX == 2 && a.bd > 8) {  // ROUND_POWER_OF_TWO(sse, 2 * (bd - 8)) of the highbd variance / sse
              const int sh = 2 * (a.bd - 8);
              t = (t + (1ull << (sh - 1))) >> sh;
            }
            a.dist[bi] = 16 * (long long)(unsigned)t;
          }
        }
      }
    }
    __syncwarp();  // stage and tile are reused by the next round
  }
  cp_async_wait<0>();
}

struct ItxKernelInfo {
  void (*fn)(const ItxArgs);
  int smem_bytes;
  int bpw;
};
template <int W, int H, int PX>
__global__ void itx_direct_kernel(const ItxArgs a);  // itx_direct.cuh

template <int W, int H, int PX>
ItxKernelInfo itx_kernel_info() {
  ItxKernelInfo k;
  constexpr bool direct = W * H <= 128 || (W == 8 && H == 32);  // see itx_direct.cuh
  if constexpr (direct) {
    k.fn = itx_direct_kernel<W, H, PX>;
    k.smem_bytes = kWarpsPerCta * TxCfg<W, H>::TB * TxCfg<W, H>::XS * 4;
  } else {
    k.fn = itx_kernel<W, H, PX>;
    k.smem_bytes = ItxSmem<W, H, PX>::kCtaBytes;
  }
  k.bpw = TxCfg<W, H>::TB;
  return k;
}

}  // namespace b200
