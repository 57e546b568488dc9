// txq_kernel.cuh -- fused (subtract +) forward 2-D transform (+ quantise) (+ block error)
// kernel, version 2.
//
// Replaces, for a whole batch of same-size transform blocks in one launch, the
// reference's per-block chain
//   aom_subtract_block        aom_dsp/subtract.c:20
//   av1_fwd_txfm2d_WxH        av1/encoder/av1_fwd_txfm2d.c:56-312
//   av1_quantize_fp* / aom_quantize_b* (+highbd)   av1/encoder/av1_quantize.c:36-264,
//                                                  aom_dsp/quantize.c:108-316
//   av1_block_error / av1_highbd_block_error       av1/encoder/rdopt.c:635-682
//
// Mapping: a warp owns a tile of TB = 32/min(W, Hk) transform
// blocks at a time and strides over the tiles of the launch (persistent grid); the
// tile's columns, then its rows, are spread over the 32 lanes so that no lane idles in
// either pass whatever the block shape.  Per tile:
//   1. input rows arrive in shared memory through cp.async (16/8/4-byte chunks, one
//      address computation per chunk, ring of S stages filled S-1 tiles ahead; the
//      upside-down flip of FLIPADST columns is applied by the copy);
//   2. a lane reads its column with immediate-offset LDS and runs the whole H-point
//      column transform in registers;
//   3. the column results cross to row ownership through a padded per-warp int32 tile
//      (__syncwarp only -- there is no CTA-wide barrier in this kernel);
//   4. a lane reads its row with LDS.128, runs the W-point row transform, quantises its
//      W coefficients in registers and stores coeff / qcoeff / dqcoeff -- which the
//      codec keeps column-major, i.e. lane-contiguous -- with coalesced 32-bit stores;
//   5. eob and the error sums are shuffle-reduced over the block's lanes.
// Arithmetic: butterflies use the 3-instruction all-32-bit form when every residual of
// the warp's tile satisfies |x| <= kGuard (proved exact by tools/range_proof.py), and
// the literal 64-bit-sum form of half_btf otherwise; results are bit-identical to the
// reference either way.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "internal.h"
#include "txfm1d.cuh"

namespace b200 {

__host__ __device__ constexpr int cx_min(int a, int b) { return a < b ? a : b; }
__host__ __device__ constexpr int cx_max(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int cx_round_up(int v, int m) { return (v + m - 1) / m * m; }

constexpr int kGuard = 4094;  // tools/range_proof.py: fast butterflies exact for |x| <= 4094
constexpr int kWarpsPerCta = 4;

// ---------------------------------------------------------------- configuration --
// av1_fwd_txfm_shift_ls (av1_fwd_txfm2d.c:314-340): shift[0] is an up-shift, shift[1]
// and shift[2] are rounding down-shifts; stored here as non-negative amounts.
template <int W, int H>
struct TxCfg {
  static constexpr int kPels = W * H;
  static constexpr int kMaxDim = cx_max(W, H), kMinDim = cx_min(W, H);
  static constexpr int kUp =
      (kMaxDim == 64 && !(W == 64 && H == 32) && !(W == 64 && H == 16)) ? 0 : 2;
  static constexpr int kDown1 =
      (W == 4 && H == 4)                                   ? 0
      : (kMaxDim == 64 && kUp == 0)                        ? 2 /* 64x64 32x64 16x64 */
      : (W == 64)                                          ? 4 /* 64x32 64x16 */
      : (kMaxDim == 8 || (kMaxDim == 16 && kMinDim == 4))  ? 1
      : (kMaxDim == 16 || (kMaxDim == 32 && kMinDim == 8)) ? 2
                                                           : 4; /* 32x32 16x32 32x16 */
  static constexpr int kDown2 =
      ((W == 64 && H == 64) || (W == 32 && H == 64) || (W == 64 && H == 32)) ? 2 : 0;
  static constexpr bool kRect = (W == 2 * H) || (H == 2 * W);
  static constexpr int kLogScale = (kPels > 256) + (kPels > 1024);
  // av1_fwd_cos_bit_col/_row[log2(W)-2][log2(H)-2] (av1_fwd_txfm2d.c:342-358)
  static constexpr int cosbit_col() {
    constexpr int8_t t[5][5] = { { 13, 13, 13, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 13, 12, 13 },
                                 { 0, 13, 13, 12, 13 }, { 0, 0, 13, 12, 13 } };
    return t[cx_log2(W) - 2][cx_log2(H) - 2];
  }
  static constexpr int cosbit_row() {
    constexpr int8_t t[5][5] = { { 13, 13, 12, 0, 0 }, { 13, 13, 13, 12, 0 }, { 13, 13, 12, 13, 12 },
                                 { 0, 12, 13, 12, 11 }, { 0, 0, 12, 11, 10 } };
    return t[cx_log2(W) - 2][cx_log2(H) - 2];
  }
  static constexpr int kCosCol = cosbit_col(), kCosRow = cosbit_row();

  // kept region and lane mapping.  A warp works on a tile of TB blocks.  Column pass:
  // the TB*W columns of the tile are spread over the 32 lanes (CI rounds); row pass: the
  // TB*Hk kept rows likewise (RI rounds).  TB is chosen so that the narrower of the two
  // passes fills the warp exactly -- no lane idles in either pass for any block shape.
  static constexpr int Wk = cx_min(W, 32), Hk = cx_min(H, 32);
  static constexpr int TB = 32 / cx_min(W, Hk);  // blocks per tile
  static constexpr int BPW = TB;
  static constexpr int CI = TB * W / 32;         // column rounds
  static constexpr int RI = TB * Hk / 32;        // row rounds
  // transpose tile (int32): pitch = 4 (mod 8) words -> LDS.128 row reads and the
  // column-wise STS are both bank-conflict free
  static constexpr int LDX = W == 4 ? 4 : W + 4;
  static constexpr int XS = (TB > 1 && W < 32) ? cx_round_up(Hk * LDX, 32) + W : Hk * LDX;  // words / slot
};

// Input staging geometry for element size ESZ (2: int16 residual / uint16 pixels, 1: uint8)
template <int W, int H, int ESZ>
struct StageCfg {
  using C = TxCfg<W, H>;
  static constexpr int kRowBytes = W * ESZ;
  static constexpr int kChunk = cx_min(16, kRowBytes);        // bytes per cp.async
  static constexpr int kCpr = kRowBytes / kChunk;             // chunks per row
  static constexpr int kCpb = H * kCpr;                       // chunks per block
  static constexpr int kChunksPerTile = C::BPW * kCpb;
  static constexpr int kCpl = (kChunksPerTile + 31) / 32;     // chunks per lane
  static constexpr int kTileBytes = H * kRowBytes;
  // slot pitch: distinct banks for the column reads of the BPW slots of a warp
  static constexpr int kSlotBytes = (C::TB > 1 && kRowBytes < 128)
                                        ? cx_round_up(kTileBytes, 128) + kRowBytes
                                        : cx_round_up(kTileBytes, 16);
  static constexpr int kPlaneBytes = C::BPW * kSlotBytes;     // one operand, all slots
};

template <int W, int H, int IN>
struct SmemCfg {
  using C = TxCfg<W, H>;
  static constexpr int ESZ = IN == B200_IN_PIXELS8 ? 1 : 2;
  using S = StageCfg<W, H, ESZ>;
  static constexpr int kOperands = IN == B200_IN_RESIDUAL ? 1 : 2;  // src (+ pred)
  static constexpr int kStageBytes = cx_round_up(kOperands * S::kPlaneBytes, 16);
  // deeper prefetch for small tiles (bytes in flight per warp)
  // bytes in flight per warp count BOTH operand planes of the pixel-input kinds: fewer stages
  // there (same prefetch depth in bytes, less shared memory, more resident warps).  Measured
  // on B200, 4K 10-bit: 8x8 / 16x16 / 8x16 / 16x8 4-9 % faster, other sizes unchanged
  // (-DB200_TXQ_STAGES_PER_OPERAND restores the per-operand rule).
#ifndef B200_TXQ_STAGES_PER_OPERAND
  static constexpr int kFlight = kOperands * S::kPlaneBytes;
#else
  static constexpr int kFlight = S::kPlaneBytes;
#endif
  // 64-point sizes: no staging ring at all.  Their tiles are 2-16 KB per operand, so two stages of a
  // 4-warp CTA took 84-166 KB of shared memory and left 4-8 warps per SM resident (ncu round 2: 6-12 %
  // warps active, the kernels wait instead of streaming); the column pass reads its samples straight
  // from global memory instead (a lane's H loads are independent, consecutive lanes = consecutive
  // columns of a row, so every request is a few full sectors) and occupancy doubles.
  // (residual input: measured better only when the rows are 64 wide; 32x64 / 16x64 keep the ring)
  static constexpr bool kDirect = W == 64 || (cx_max(W, H) == 64 && IN != B200_IN_RESIDUAL);
  static constexpr int kStages = kDirect ? 0 : (kFlight >= 2048 ? 2 : (kFlight >= 1024 ? 3 : 4));
  static constexpr int kXposeBytes = cx_round_up(C::BPW * C::XS * 4, 16);
  // Output tiles (optional bulk-store path, -DB200_TXQ_BULK_OUT): the tile's coeff / qcoeff /
  // dqcoeff blocks are assembled in shared memory in the codec's storage order and leave
  // through the bulk-copy engine instead of costing one L1 store wavefront per 32..64 bytes
  // (short storage columns, Hk <= 16).  Measured on B200 (round 1): one 256..2048-byte bulk
  // copy per block and array is slower than the direct stores for every size (8x8 0.67 ->
  // 0.52 of the HBM roofline, 16x16 0.75 -> 0.61), so the path is off by default.
#ifdef B200_TXQ_BULK_OUT
  static constexpr bool kBulkOut = C::Hk <= 16;
#else
  static constexpr bool kBulkOut = false;
#endif
  static constexpr int kOutBlockBytes = C::Wk * C::Hk * 4;
  static constexpr int kOutPitchWords = C::Wk * C::Hk + C::Hk;  // + Hk: the slots of a round hit distinct banks
  static constexpr int kOutArrayBytes = C::BPW * kOutPitchWords * 4;
  static constexpr int kOutBytes = kBulkOut ? 3 * kOutArrayBytes : 0;
  static constexpr int kOutOff = cx_round_up(kStages * kStageBytes + kXposeBytes, 128);
  static constexpr int kWarpBytes = cx_round_up(kOutOff + kOutBytes, 128);
  static constexpr int kCtaBytes = kWarpsPerCta * kWarpBytes;
};

struct TxqArgs {
  const B200TxBlock *blocks;
  const B200TxPred *preds;
  const void *src;
  const void *pred;
  const B200QRow *qrows;
  const int16_t *iscan3;  // default, mrow, mcol inverse scans of this tx size
  const unsigned long long *pmask3;  // small sizes: scan prefix masks, 3 x (N + 1), see txq.cu
  int32_t *coeff, *qcoeff, *dqcoeff;
  uint16_t *eob;
  int64_t *dist;
  int n_blocks;
  int q_kind, bd;
  unsigned flags;
};

// 1-D kind per 2-D type: 0 DCT, 1 ADST, 2 FLIPADST, 3 IDENTITY
// (vtx_tab / htx_tab, av1/common/common_data.h:149-159), packed 2 bits per type
constexpr uint32_t kVtxPacked = 0u | (1u << 2) | (0u << 4) | (1u << 6) | (2u << 8) | (0u << 10) |
                                (2u << 12) | (1u << 14) | (2u << 16) | (3u << 18) | (0u << 20) |
                                (3u << 22) | (1u << 24) | (3u << 26) | (2u << 28) | (3u << 30);
constexpr uint32_t kHtxPacked = 0u | (0u << 2) | (1u << 4) | (1u << 6) | (0u << 8) | (2u << 10) |
                                (2u << 12) | (2u << 14) | (1u << 16) | (3u << 18) | (3u << 20) |
                                (0u << 22) | (3u << 24) | (1u << 26) | (3u << 28) | (2u << 30);
__host__ __device__ __forceinline__ int vtx_of(int tx_type) { return (kVtxPacked >> (2 * tx_type)) & 3; }
__host__ __device__ __forceinline__ int htx_of(int tx_type) { return (kHtxPacked >> (2 * tx_type)) & 3; }

// --------------------------------------------------------------- small helpers --
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
template <int BYTES>
__device__ __forceinline__ void cp_async(uint32_t dst, const void *src) {
  if constexpr (BYTES == 16)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
  else if constexpr (BYTES == 8)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// shared -> global bulk copy (TMA engine, no L1 store wavefronts); 16-byte aligned, n % 16 == 0
__device__ __forceinline__ void bulk_store(void *gdst, const void *ssrc, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst),
               "r"(smem_u32(ssrc)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ B200TxBlock load_desc(const B200TxBlock *p) {
  const uint4 raw = __ldg(reinterpret_cast<const uint4 *>(p));
  B200TxBlock d;
  memcpy(&d, &raw, sizeof(d));
  return d;
}

// Arithmetic flavour of one pass: EXACT = literal reference widths everywhere.
template <bool EXACT, int K>
__device__ __forceinline__ int32_t down_shift(int32_t x) {
  if constexpr (K == 0) return x;
  else if constexpr (EXACT) return rshift_round<K>(x);
  else return (x + (1 << (K - 1))) >> K;  // cannot overflow under the guard
}
// round_shift((int64)v * 5793 * MUL, 12)
template <bool EXACT, int MUL>
__device__ __forceinline__ int32_t sqrt2_scale(int32_t v) {
  if constexpr (EXACT) {
    return scale_sqrt2<MUL>(v);
  } else {
    // 5793*MUL = MUL*4096 + R  ->  MUL*v + ((v*R + 2048) >> 12), all in 32 bits
    constexpr int R = 5793 * MUL - 4096 * MUL;
    return MUL * v + ((v * R + 2048) >> 12);
  }
}

template <int N, int BIT, bool EXACT, int KEEP>
__device__ __forceinline__ void txfm1d(int kind, int32_t *x, int32_t *y) {
  using B = Btf<BIT, EXACT>;
  if constexpr (N == 64) {
    fdct<64, B, KEEP>(x, y);  // only DCT exists at 64 (av1_txfm.c:89-100)
  } else if constexpr (N == 32) {
    if (kind == 3) {
      fidentity<32>(x, y);
    } else {
      fdct<32, B, KEEP>(x, y);
    }
  } else {
    if (kind == 0) {
      fdct<N, B, KEEP>(x, y);
    } else if (kind == 3) {
      if constexpr (EXACT || N == 8) {
        fidentity<N>(x, y);
      } else {
        static_for<0, N>([&](auto ic) {
          constexpr int i = decltype(ic)::value;
          y[i] = sqrt2_scale<false, N == 4 ? 1 : 2>(x[i]);
        });
      }
    } else {
      fadst<N, B>(x, y);
    }
  }
}

// ------------------------------------------------------------------- staging -----
// Issue the cp.async copies of one tile (all BPW blocks of warp-tile `wt`) into `stage`.
template <int W, int H, int IN>
__device__ __forceinline__ void stage_tile(const TxqArgs &a, int wt, unsigned char *stage, int lane) {
  using C = TxCfg<W, H>;
  using M = SmemCfg<W, H, IN>;
  using S = typename M::S;
  constexpr int ESZ = M::ESZ;
#pragma unroll
  for (int i = 0; i < S::kCpl; ++i) {
    const int k = lane + 32 * i;
    if (S::kChunksPerTile % 32 != 0 && k >= S::kChunksPerTile) break;
    const int slot = k / S::kCpb, rem = k % S::kCpb;
    const int row = rem / S::kCpr, part = rem % S::kCpr;
    const int bi = wt * C::BPW + slot;
    if (bi >= a.n_blocks) continue;
    const B200TxBlock d = load_desc(a.blocks + bi);
    const int drow = vtx_of(d.tx_type & 15) == 2 ? H - 1 - row : row;  // ud_flip
    unsigned char *dst = stage + slot * S::kSlotBytes + drow * S::kRowBytes + part * S::kChunk;
    const unsigned char *g = static_cast<const unsigned char *>(a.src) +
                             ((size_t)d.src_offset + (size_t)row * d.src_stride) * ESZ + part * S::kChunk;
    if ((reinterpret_cast<uintptr_t>(g) & (S::kChunk - 1)) == 0) {
      cp_async<S::kChunk>(smem_u32(dst), g);
    } else {  // unaligned source: element-wise copy (rare)
#pragma unroll
      for (int e = 0; e < S::kChunk; ++e) dst[e] = g[e];
    }
    if constexpr (IN != B200_IN_RESIDUAL) {
      const B200TxPred pd = a.preds[bi];
      unsigned char *dstp = dst + S::kPlaneBytes;
      const unsigned char *gp = static_cast<const unsigned char *>(a.pred) +
                                ((size_t)pd.pred_offset + (size_t)row * pd.pred_stride) * ESZ +
                                part * S::kChunk;
      if ((reinterpret_cast<uintptr_t>(gp) & (S::kChunk - 1)) == 0) {
        cp_async<S::kChunk>(smem_u32(dstp), gp);
      } else {
#pragma unroll
        for (int e = 0; e < S::kChunk; ++e) dstp[e] = gp[e];
      }
    }
  }
}

// ---------------------------------------------------------------- the passes -----
// Row pass + final scalings: y[c] = coefficient (c, t) of the block, c < Wk.
template <int W, int H, bool EXACT>
__device__ __forceinline__ void row_pass(int ht, const int32_t *xrow, int32_t *y) {
  using C = TxCfg<W, H>;
  int32_t x[W];
  if constexpr (W == 4) {
    const int4 v = *reinterpret_cast<const int4 *>(xrow);
    x[0] = v.x, x[1] = v.y, x[2] = v.z, x[3] = v.w;
  } else {
#pragma unroll
    for (int c = 0; c < W; c += 4) {
      const int4 v = *reinterpret_cast<const int4 *>(xrow + c);
      x[c] = v.x, x[c + 1] = v.y, x[c + 2] = v.z, x[c + 3] = v.w;
    }
  }
  txfm1d<W, C::kCosRow, EXACT, C::Wk>(ht, x, y);
#pragma unroll
  for (int c = 0; c < C::Wk; ++c) {
    int32_t v = down_shift<EXACT, C::kDown2>(y[c]);
    if constexpr (C::kRect) v = sqrt2_scale<EXACT, 1>(v);
    y[c] = v;
  }
}

// Read column c of the staged block (residual formed here for pixel inputs), track the
// value range for the guard, apply the up-shift (av1_round_shift_array(temp_in, H,
// -shift[0]); the input is int16 so the reference's clamp can never trigger).
template <int W, int H, int IN>
__device__ __forceinline__ void load_column(const unsigned char *stg, int c, int32_t *x,
                                            int32_t &mx, int32_t &mn) {
  using C = TxCfg<W, H>;
  constexpr int PB = SmemCfg<W, H, IN>::S::kPlaneBytes;
#pragma unroll
  for (int r = 0; r < H; ++r) {
    int32_t v;
    if constexpr (IN == B200_IN_RESIDUAL) {
      v = reinterpret_cast<const int16_t *>(stg)[r * W + c];
    } else if constexpr (IN == B200_IN_PIXELS8) {
      v = (int)stg[r * W + c] - (int)stg[PB + r * W + c];
    } else {
      v = (int16_t)((int)reinterpret_cast<const uint16_t *>(stg)[r * W + c] -
                    (int)reinterpret_cast<const uint16_t *>(stg + PB)[r * W + c]);
    }
    if constexpr (IN != B200_IN_PIXELS8) {
      mx = v > mx ? v : mx;
      mn = v < mn ? v : mn;
    }
    x[r] = (int32_t)((uint32_t)v << C::kUp);
  }
}

// Same, straight from global memory (SmemCfg::kDirect): column c of block d.
template <int W, int H, int IN>
__device__ __forceinline__ void load_column_direct(const TxqArgs &a, const B200TxBlock &d, int bi, int c,
                                                   bool ud, int32_t *x, int32_t &mx, int32_t &mn) {
  using C = TxCfg<W, H>;
  if constexpr (IN == B200_IN_RESIDUAL) {
    const int16_t *p = static_cast<const int16_t *>(a.src) + d.src_offset + c;
#pragma unroll
    for (int r = 0; r < H; ++r) x[r] = __ldg(p + (size_t)(ud ? H - 1 - r : r) * d.src_stride);
  } else {
    const B200TxPred pd = a.preds[bi];
    if constexpr (IN == B200_IN_PIXELS8) {
      const uint8_t *p = static_cast<const uint8_t *>(a.src) + d.src_offset + c;
      const uint8_t *q = static_cast<const uint8_t *>(a.pred) + pd.pred_offset + c;
#pragma unroll
      for (int r = 0; r < H; ++r) {
        const int rr = ud ? H - 1 - r : r;
        x[r] = (int)__ldg(p + (size_t)rr * d.src_stride) - (int)__ldg(q + (size_t)rr * pd.pred_stride);
      }
    } else {
      const uint16_t *p = static_cast<const uint16_t *>(a.src) + d.src_offset + c;
      const uint16_t *q = static_cast<const uint16_t *>(a.pred) + pd.pred_offset + c;
#pragma unroll
      for (int r = 0; r < H; ++r) {
        const int rr = ud ? H - 1 - r : r;
        x[r] = (int16_t)((int)__ldg(p + (size_t)rr * d.src_stride) - (int)__ldg(q + (size_t)rr * pd.pred_stride));
      }
    }
  }
#pragma unroll
  for (int r = 0; r < H; ++r) {
    if constexpr (IN != B200_IN_PIXELS8) {
      mx = x[r] > mx ? x[r] : mx;
      mn = x[r] < mn ? x[r] : mn;
    }
    x[r] = (int32_t)((uint32_t)x[r] << C::kUp);
  }
}

// Column transform of one column (already in x) -> transpose tile of its block.
template <int W, int H, bool EXACT>
__device__ __forceinline__ void column_pass(int vt, int ht, int32_t *x, int32_t *tile, int c) {
  using C = TxCfg<W, H>;
  int32_t y[C::Hk];
  txfm1d<H, C::kCosCol, EXACT, C::Hk>(vt, x, y);
  const int cc = (ht == 2) ? (W - 1 - c) : c;  // lr_flip (av1_fwd_txfm2d.c:99-105)
#pragma unroll
  for (int r = 0; r < C::Hk; ++r) tile[r * C::LDX + cc] = down_shift<EXACT, C::kDown1>(y[r]);
}

// ------------------------------------------------------------------- epilogue ----
template <int W, int H, int QK, bool ERR>
__device__ __forceinline__ void quant_epilogue(const TxqArgs &a, const B200TxBlock &d, int t,
                                               const int32_t *y, int32_t *qc, int32_t *dqc,
                                               int &last_c, int64_t &err, int64_t &ssz) {
  // qc / dqc: this lane's first output element (global); element c is at [c * Hk]
  using C = TxCfg<W, H>;
  constexpr int LS = C::kLogScale;
  constexpr bool fp = QK == B200_Q_FP || QK == B200_Q_FP_HBD;
  constexpr bool hbd = QK == B200_Q_FP_HBD || QK == B200_Q_B_HBD;
  // one 32-byte quantiser row per block (broadcast load)
  const uint4 q0 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + d.qrow));
  const uint4 q1 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + d.qrow) + 1);
  B200QRow qr;
  memcpy(&qr, &q0, 16);
  memcpy(reinterpret_cast<char *>(&qr) + 16, &q1, 16);
  auto rpo2 = [](int v) { return LS ? ((v + (1 << (LS - 1))) >> LS) : v; };  // ROUND_POWER_OF_TWO
  const int rnd0 = rpo2(fp ? qr.round_fp[0] : qr.round[0]);
  const int rnd1 = rpo2(fp ? qr.round_fp[1] : qr.round[1]);
  const int qnt0 = fp ? qr.quant_fp[0] : qr.quant[0];
  const int qnt1 = fp ? qr.quant_fp[1] : qr.quant[1];
  const int deq0 = qr.dequant[0], deq1 = qr.dequant[1];
  const int qsh0 = qr.quant_shift[0], qsh1 = qr.quant_shift[1];
  // dead-zone test rewritten as |coeff| >= threshold:
  //   FP: (abs << (1+LS)) >= dequant  <=>  abs >= ceil(dequant / 2^(1+LS))
  //   B : abs*32 >= zbin*32           <=>  abs >= zbin
  constexpr int K = 1 + LS;
  const int thr0 = fp ? ((deq0 + (1 << K) - 1) >> K) : rpo2(qr.zbin[0]);
  const int thr1 = fp ? ((deq1 + (1 << K) - 1) >> K) : rpo2(qr.zbin[1]);
#pragma unroll
  for (int c = 0; c < C::Wk; ++c) {
    const bool dc = (c == 0) && (t == 0);  // compile-time false for c > 0
    const int rnd_k = dc ? rnd0 : rnd1, qnt_k = dc ? qnt0 : qnt1, deq_k = dc ? deq0 : deq1;
    const int qsh_k = dc ? qsh0 : qsh1, thr_k = dc ? thr0 : thr1;
    (void)qsh_k;
    const int32_t cf = y[c];
    const int32_t ab = cf < 0 ? -cf : cf;
    const int32_t sgn = cf < 0 ? -1 : 1;
    int32_t q, dq;
    if constexpr (fp) {
      if constexpr (!hbd) {
        // clamp64(abs + round, INT16_MIN, INT16_MAX): abs >= 0 so only the upper bound acts
        int32_t tt = (ab > 32767 ? 32767 : ab) + rnd_k;
        tt = tt > 32767 ? 32767 : tt;
        tt = ab >= thr_k ? tt : 0;
        q = (tt * qnt_k) >> (16 - LS);
      } else {
        const uint32_t tt = ab >= thr_k ? (uint32_t)(ab + rnd_k) : 0u;
        q = (int32_t)(((uint64_t)tt * (uint32_t)qnt_k) >> (16 - LS));
      }
      dq = wmul(q, deq_k) >> LS;
    } else {
      int64_t tt = (int64_t)ab + rnd_k;
      if constexpr (!hbd) tt = tt > 32767 ? 32767 : tt;
      tt = ab >= thr_k ? tt * 32 : 0;
      q = (int32_t)(((((tt * qnt_k) >> 16) + tt) * qsh_k) >> (16 - LS + 5));
      dq = wmul(q, deq_k) >> LS;
    }
    const int32_t dqs = dq * sgn;
    qc[c * C::Hk] = q * sgn;
    dqc[c * C::Hk] = dqs;
    if (q) last_c = c;  // iscan grows with c along a row: the last non-zero c wins
    if constexpr (ERR) {
      const int32_t df = wsub(cf, dqs);
      if constexpr (hbd) {
        err += (int64_t)df * df;
        ssz += (int64_t)cf * cf;
      } else {  // av1_block_error_c squares in 32-bit int (rdopt.c:642-643)
        err += (int64_t)wmul(df, df);
        ssz += (int64_t)wmul(cf, cf);
      }
    }
  }
}

// --------------------------------------------------------------------- kernel ----
template <int W, int H, int IN>
__global__ void __launch_bounds__(kWarpsPerCta * 32) txq_kernel(const TxqArgs a) {
  using C = TxCfg<W, H>;
  using M = SmemCfg<W, H, IN>;
  using S = typename M::S;
  constexpr int NS = M::kStages;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned char *wbase = smem_raw + warp * M::kWarpBytes;
  int32_t *xpose = reinterpret_cast<int32_t *>(wbase + NS * M::kStageBytes);
  const int n_tiles = (a.n_blocks + C::TB - 1) / C::TB;
  const int wstride = gridDim.x * kWarpsPerCta;
  const int wt0 = blockIdx.x * kWarpsPerCta + warp;
  const bool force_exact = (a.flags & B200_TXQ_FORCE_EXACT) != 0;

  // prologue: fill NS-1 stages
  if constexpr (!M::kDirect) {
#pragma unroll
    for (int s = 0; s < NS - 1; ++s) {
      const int wt = wt0 + s * wstride;
      if (wt < n_tiles) stage_tile<W, H, IN>(a, wt, wbase + s * M::kStageBytes, lane);
      cp_async_commit();
    }
  }

  int it = 0;
  for (int wt = wt0; wt < n_tiles; wt += wstride, ++it) {
    const unsigned char *stage = nullptr;
    if constexpr (!M::kDirect) {
      {  // keep the ring full: tile it+NS-1 goes to the stage consumed in the previous round
        const int wtp = wt + (NS - 1) * wstride;
        if (wtp < n_tiles)
          stage_tile<W, H, IN>(a, wtp, wbase + ((it + NS - 1) % (NS > 0 ? NS : 1)) * M::kStageBytes, lane);
        cp_async_commit();
      }
      cp_async_wait<(NS > 0 ? NS - 1 : 0)>();
      __syncwarp();
      stage = wbase + (it % (NS > 0 ? NS : 1)) * M::kStageBytes;
    }

    // ---------------------------------------------------------------- column rounds
    bool tile_fast = !force_exact;
#pragma unroll 1
    for (int ci = 0; ci < C::CI; ++ci) {
      const int item = lane + 32 * ci;
      const int slot = item / W, c = item % W;
      const int bi = wt * C::TB + slot;
      const bool act = bi < a.n_blocks;
      int tx_type = 0;
      B200TxBlock dcol;
      if (act) {
        if constexpr (M::kDirect) {
          dcol = load_desc(a.blocks + bi);
          tx_type = dcol.tx_type & 15;
        } else {
          tx_type = a.blocks[bi].tx_type & 15;
        }
      }
      const int vt = vtx_of(tx_type), ht = htx_of(tx_type);
      int32_t x[H], mx = 0, mn = 0;
      if (act) {
        if constexpr (M::kDirect) load_column_direct<W, H, IN>(a, dcol, bi, c, vt == 2, x, mx, mn);
        else load_column<W, H, IN>(stage + slot * S::kSlotBytes, c, x, mx, mn);
      }
      bool ok = !force_exact;
      if constexpr (IN != B200_IN_PIXELS8)
        ok = ok && __all_sync(0xffffffffu, mx <= kGuard && mn >= -kGuard);
      tile_fast = tile_fast && ok;
      if (act) {
        if (ok) column_pass<W, H, false>(vt, ht, x, xpose + slot * C::XS, c);
        else column_pass<W, H, true>(vt, ht, x, xpose + slot * C::XS, c);
      }
    }
    if constexpr (M::kBulkOut) {
      if (lane < C::TB) bulk_wait_read();  // the previous tile's output blocks have left shared memory
    }
    __syncwarp();

    // ------------------------------------------------------------------- row rounds
#pragma unroll 1
    for (int ri = 0; ri < C::RI; ++ri) {
      const int item = lane + 32 * ri;
      const int slot = item / C::Hk, t = item % C::Hk;
      const int bi = wt * C::TB + slot;
      const bool act = bi < a.n_blocks;
      B200TxBlock d;
      if (act) d = load_desc(a.blocks + bi);
      else memset(&d, 0, sizeof(d));
      const int ht = htx_of(d.tx_type & 15);
      int last_c = -1;
      int64_t err = 0, ssz = 0;
      if (act) {
        int32_t y[C::Wk];
        const int32_t *xrow = xpose + slot * C::XS + t * C::LDX;
        if (tile_fast) row_pass<W, H, false>(ht, xrow, y);
        else row_pass<W, H, true>(ht, xrow, y);
        const size_t obase = (size_t)d.out_offset + t;
        int32_t *oc, *qc, *dqc;
        if constexpr (M::kBulkOut) {  // assemble the blocks in shared memory, stored below
          int32_t *ot = reinterpret_cast<int32_t *>(wbase + M::kOutOff) + slot * M::kOutPitchWords + t;
          oc = ot, qc = ot + M::kOutArrayBytes / 4, dqc = ot + 2 * (M::kOutArrayBytes / 4);
        } else {
          oc = a.coeff + obase, qc = a.qcoeff + obase, dqc = a.dqcoeff + obase;
        }
        if (a.flags & B200_TXQ_STORE_COEFF) {
#pragma unroll
          for (int c = 0; c < C::Wk; ++c) oc[c * C::Hk] = y[c];
        }
        if (a.q_kind != B200_Q_NONE) {
          const bool we = (a.flags & B200_TXQ_BLOCK_ERROR) != 0;
          switch (a.q_kind) {
            case B200_Q_FP:
              if (we) quant_epilogue<W, H, B200_Q_FP, true>(a, d, t, y, qc, dqc, last_c, err, ssz);
              else quant_epilogue<W, H, B200_Q_FP, false>(a, d, t, y, qc, dqc, last_c, err, ssz);
              break;
            case B200_Q_FP_HBD:
              if (we) quant_epilogue<W, H, B200_Q_FP_HBD, true>(a, d, t, y, qc, dqc, last_c, err, ssz);
              else quant_epilogue<W, H, B200_Q_FP_HBD, false>(a, d, t, y, qc, dqc, last_c, err, ssz);
              break;
            case B200_Q_B:
              if (we) quant_epilogue<W, H, B200_Q_B, true>(a, d, t, y, qc, dqc, last_c, err, ssz);
              else quant_epilogue<W, H, B200_Q_B, false>(a, d, t, y, qc, dqc, last_c, err, ssz);
              break;
            default:
              if (we) quant_epilogue<W, H, B200_Q_B_HBD, true>(a, d, t, y, qc, dqc, last_c, err, ssz);
              else quant_epilogue<W, H, B200_Q_B_HBD, false>(a, d, t, y, qc, dqc, last_c, err, ssz);
              break;
          }
        }
      }
      if (a.q_kind != B200_Q_NONE) {
        int eob = 0;
        if (last_c >= 0) {
          const int cls = d.tx_type < 10 ? 0 : ((d.tx_type & 1) ? 2 : 1);
          eob = a.iscan3[cls * (C::Wk * C::Hk) + last_c * C::Hk + t] + 1;
        }
        // reduce over the Hk lanes that own the block's rows (xor stays inside the group)
#pragma unroll
        for (int o = C::Hk / 2; o > 0; o >>= 1) {
          const int e2 = __shfl_xor_sync(0xffffffffu, eob, o);
          eob = e2 > eob ? e2 : eob;
        }
        if (a.flags & B200_TXQ_BLOCK_ERROR) {
#pragma unroll
          for (int o = C::Hk / 2; o > 0; o >>= 1) {
            err += __shfl_xor_sync(0xffffffffu, err, o);
            ssz += __shfl_xor_sync(0xffffffffu, ssz, o);
          }
        }
        if (act && t == 0) {
          a.eob[bi] = (uint16_t)eob;
          if ((a.flags & B200_TXQ_BLOCK_ERROR) && a.dist) {
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
    }
    if constexpr (M::kBulkOut) {
      fence_async_smem();  // generic-proxy writes above -> visible to the bulk-copy engine
      __syncwarp();
      const int bi = wt * C::TB + lane;
      const bool mine = lane < C::TB && bi < a.n_blocks;
      const bool want_c = (a.flags & B200_TXQ_STORE_COEFF) != 0, want_q = a.q_kind != B200_Q_NONE;
      uint32_t ob = 0;
      bool aligned = true;
      if (mine) {
        ob = a.blocks[bi].out_offset;
        uintptr_t al = 0;
        if (want_c) al |= reinterpret_cast<uintptr_t>(a.coeff + ob);
        if (want_q) al |= reinterpret_cast<uintptr_t>(a.qcoeff + ob) | reinterpret_cast<uintptr_t>(a.dqcoeff + ob);
        aligned = (al & 15) == 0;
      }
      const int32_t *ot = reinterpret_cast<const int32_t *>(wbase + M::kOutOff);
      if (mine && aligned) {
        const int32_t *src = ot + lane * M::kOutPitchWords;
        if (want_c) bulk_store(a.coeff + ob, src, M::kOutBlockBytes);
        if (want_q) {
          bulk_store(a.qcoeff + ob, src + M::kOutArrayBytes / 4, M::kOutBlockBytes);
          bulk_store(a.dqcoeff + ob, src + 2 * (M::kOutArrayBytes / 4), M::kOutBlockBytes);
        }
        bulk_commit();
      }
      // output blocks that are not 16-byte aligned: plain cooperative copy (rare)
      unsigned bad = __ballot_sync(0xffffffffu, mine && !aligned);
      while (bad) {
        const int b = __ffs(bad) - 1;
        bad &= bad - 1;
        const uint32_t obb = __shfl_sync(0xffffffffu, ob, b);
        const int32_t *src = ot + b * M::kOutPitchWords;
        for (int i = lane; i < C::Wk * C::Hk; i += 32) {
          if (want_c) a.coeff[(size_t)obb + i] = src[i];
          if (want_q) {
            a.qcoeff[(size_t)obb + i] = src[M::kOutArrayBytes / 4 + i];
            a.dqcoeff[(size_t)obb + i] = src[2 * (M::kOutArrayBytes / 4) + i];
          }
        }
      }
    }
    __syncwarp();  // stage and transpose tile are reused by the next round
  }
  if constexpr (!M::kDirect) cp_async_wait<0>();
  if constexpr (M::kBulkOut) bulk_wait_all();  // shared memory must outlive the last bulk reads
}

// one launcher per (size, input kind); defined in txq_inst_*.cu
struct TxqKernelInfo {
  void (*fn)(const TxqArgs);
  int smem_bytes;
  int bpw;
};
template <int W, int H, int IN>
TxqKernelInfo txq_kernel_info() {
  TxqKernelInfo k;
  k.fn = txq_kernel<W, H, IN>;
  k.smem_bytes = SmemCfg<W, H, IN>::kCtaBytes;
  k.bpw = TxCfg<W, H>::TB;
  return k;
}

}  // namespace b200
