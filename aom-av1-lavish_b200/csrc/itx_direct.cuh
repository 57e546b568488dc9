// itx_direct.cuh -- version 1 of the inverse kernel, kept for the small transform sizes (at most 128
// samples, and 8x32): no staging ring -- a lane loads its row's coefficients straight from global memory,
// the column lanes read prediction / source and write the reconstruction one pixel per row.  For these
// sizes the per-tile bookkeeping of the staged kernel (itx_kernel.cuh) costs more than the latency it
// hides (measured on B200, 8 x 1080p: 4x4 0.30 vs 0.18, 4x8 0.28 vs 0.16, 8x8 0.25 vs 0.22 of the HBM
// roofline; from 16x16 up the staged kernel wins, 32x32 0.27 vs 0.10).  Same results bit for bit.
#pragma once
#include "itx_kernel.cuh"

namespace b200 {

template <int W, int H, bool WIDE>
__device__ __forceinline__ void itx_direct_row(const ItxArgs &a, const B200TxBlock &d, int t, int ht,
                                        int32_t *tile_row) {
  using C = TxCfg<W, H>;
  int32_t x[C::Wk], y[W];
  const int32_t *g = a.dqcoeff + (size_t)d.out_offset + t;
#pragma unroll
  for (int c = 0; c < C::Wk; ++c) x[c] = __ldg(g + c * C::Hk);
  const int32_t hi = (1 << (a.bd + 7)) - 1, lo = -hi - 1;  // clamp_buf(temp_in, W, bd + 8) = the row range
#pragma unroll
  for (int c = 0; c < C::Wk; ++c) {
    int32_t v = x[c];
    if constexpr (C::kRect)  // round_shift((int64_t)input * NewInvSqrt2, NewSqrt2Bits)
      v = (int32_t)(((int64_t)v * 2896 + 2048) >> 12);
    x[c] = clamp_range(v, lo, hi);
  }
  itxfm1d<W, WIDE, C::Wk>(ht, x, y, lo, hi);
#pragma unroll
  for (int c = 0; c < W; c += 4) {
    int4 v;
    v.x = rshift_round<ItxCfg<W, H>::kRowShift>(y[c]);
    v.y = rshift_round<ItxCfg<W, H>::kRowShift>(y[c + 1]);
    v.z = rshift_round<ItxCfg<W, H>::kRowShift>(y[c + 2]);
    v.w = rshift_round<ItxCfg<W, H>::kRowShift>(y[c + 3]);
    *reinterpret_cast<int4 *>(tile_row + c) = v;
  }
}

template <int W, int H, int PX>
__global__ void __launch_bounds__(kWarpsPerCta * 32) itx_direct_kernel(const ItxArgs a) {
  using C = TxCfg<W, H>;
  using pix_t = typename std::conditional<PX == 1, uint8_t, uint16_t>::type;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int32_t *xpose = reinterpret_cast<int32_t *>(smem_raw) + warp * (C::TB * C::XS);
  const int n_tiles = (a.n_blocks + C::TB - 1) / C::TB;
  const int wstride = gridDim.x * kWarpsPerCta;
  const int pmax = (1 << a.bd) - 1;
  // column pass: clamp_buf(temp_in, H, max(bd + 6, 16)) = the column stage range
  const int32_t chi = (1 << (a.bd == 12 ? 17 : 15)) - 1, clo = -chi - 1;

  for (int wt = blockIdx.x * kWarpsPerCta + warp; wt < n_tiles; wt += wstride) {
    // ------------------------------------------------------------------- row rounds
#pragma unroll 1
    for (int ri = 0; ri < C::RI; ++ri) {
      const int item = lane + 32 * ri;
      const int slot = item / C::Hk, t = item % C::Hk;
      const int bi = wt * C::TB + slot;
      if (bi < a.n_blocks && (!a.eob || a.eob[bi])) {
        const B200TxBlock d = load_desc(a.blocks + bi);
        const int ht = htx_of(d.tx_type & 15);
        int32_t *row = xpose + slot * C::XS + t * C::LDX;
        if (a.bd == 12) itx_direct_row<W, H, true>(a, d, t, ht, row);
        else itx_direct_row<W, H, false>(a, d, t, ht, row);
      }
    }
    __syncwarp();
    // ---------------------------------------------------------------- column rounds
    unsigned long long carry = 0;
#pragma unroll 1
    for (int ci = 0; ci < C::CI; ++ci) {
      const int item = lane + 32 * ci;
      const int slot = item / W, c = item % W;
      const int bi = wt * C::TB + slot;
      const bool act = bi < a.n_blocks;
      unsigned sse = 0;
      if (act) {
        const B200TxBlock d = load_desc(a.blocks + bi);
        const B200TxPred pd = a.preds[bi];
        const int vt = vtx_of(d.tx_type & 15), ht = htx_of(d.tx_type & 15);
        const bool coded = !a.eob || a.eob[bi];
        int32_t y[H];
        if (coded) {
          int32_t x[C::Hk];
          const int cc = (ht == 2) ? (W - 1 - c) : c;  // lr_flip (av1_inv_txfm2d.c:298-302)
          const int32_t *col = xpose + slot * C::XS + cc;
#pragma unroll
          for (int r = 0; r < C::Hk; ++r) x[r] = clamp_range(col[r * C::LDX], clo, chi);
          itxfm1d<H, false, C::Hk>(vt, x, y, clo, chi);
          if (vt == 2) {  // ud_flip (:313-318)
#pragma unroll
            for (int r = 0; r < H / 2; ++r) {
              const int32_t p = y[r], q = y[H - 1 - r];
              y[r] = q;
              y[H - 1 - r] = p;
            }
          }
        }
        const int vis_w = d.reserved ? (d.reserved & 255) : W;
        const int vis_h = d.reserved ? (d.reserved >> 8) : H;
        const pix_t *pp = static_cast<const pix_t *>(a.pred) + pd.pred_offset + c;
        const pix_t *sp = a.src ? static_cast<const pix_t *>(a.src) + d.src_offset + c : nullptr;
        pix_t *rp = a.recon ? static_cast<pix_t *>(a.recon) + pd.pred_offset + c : nullptr;
#pragma unroll
        for (int r = 0; r < H; ++r) {
          int v = pp[(size_t)r * pd.pred_stride];
          if (coded) v = min(max(v + rshift_round<4>(y[r]), 0), pmax);
          if (rp) rp[(size_t)r * pd.pred_stride] = (pix_t)v;
          if (sp && c < vis_w && r < vis_h) {
            const int df = (int)sp[(size_t)r * d.src_stride] - v;
            sse += (unsigned)(df * df);
          }
        }
      }
      if (a.dist) {
        // a block's columns: min(W, 32) lanes of this round (and the next round when W = 64)
        unsigned long long s = sse;
#pragma unroll
        for (int o = cx_min(W, 32) / 2; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if constexpr (W == 64) {
          s += carry;
          carry = (ci & 1) ? 0ull : s;
        }
        if (act && c == (W == 64 ? 32 : 0) ) {
          if (PX == 2 && a.bd > 8) {  // ROUND_POWER_OF_TWO(sse, 2 * (bd - 8)) of the highbd variance / sse
            const int sh = 2 * (a.bd - 8);
            s = (s + (1ull << (sh - 1))) >> sh;
          }
          a.dist[bi] = 16 * (long long)(unsigned)s;
        }
      }
    }
    __syncwarp();  // the tile is reused by the next round
  }
}

}  // namespace b200
