// txq.cu -- fused (subtract +) forward 2-D transform (+ quantise) (+ block error).
//
// Replaces, for a whole batch of same-size transform blocks in one launch, the
// reference's per-block chain
//   aom_subtract_block        aom_dsp/subtract.c:20
//   av1_fwd_txfm2d_WxH        av1/encoder/av1_fwd_txfm2d.c:56-312
//   av1_quantize_fp* / aom_quantize_b* (+highbd)   av1/encoder/av1_quantize.c:36-264,
//                                                  aom_dsp/quantize.c:108-316
//   av1_block_error / av1_highbd_block_error       av1/encoder/rdopt.c:635-682
//
// Mapping: a transform block is owned by NT <= 32 lanes of ONE warp;
// lane c runs the whole H-point column transform of column c in registers, the
// column results cross to row ownership through a padded per-warp shared-memory tile
// (only __syncwarp, no CTA barrier anywhere), lane r then runs the W-point row
// transform of row r, quantises its W coefficients in registers and writes
// coeff / qcoeff / dqcoeff -- which the codec stores column-major, i.e. exactly
// lane-contiguous -- with fully coalesced stores.  eob and the error sums are
// shuffle-reduced over the block's lanes.
#include <math.h>
#include <stdarg.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "internal.h"
#include "txfm1d.cuh"

namespace b200 {

// 1-D kind per 2-D type: 0 DCT, 1 ADST, 2 FLIPADST, 3 IDENTITY
// (vtx_tab / htx_tab, av1/common/common_data.h:149-159)
__constant__ uint8_t c_vtx[16] = { 0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3 };
__constant__ uint8_t c_htx[16] = { 0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2 };

static const uint8_t kVtxHost[16] = { 0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3 };
static const uint8_t kHtxHost[16] = { 0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2 };

// ------------------------------------------------------------ static geometry --
__host__ __device__ constexpr int cx_min(int a, int b) { return a < b ? a : b; }
__host__ __device__ constexpr int cx_max(int a, int b) { return a > b ? a : b; }

// av1_fwd_txfm_shift_ls (av1_fwd_txfm2d.c:314-340): shift[0] is an up-shift, shift[1]
// and shift[2] are rounding down-shifts; stored here as non-negative amounts.
template <int W, int H>
struct TxCfg {
  static constexpr int kPels = W * H;
  static constexpr int kMaxDim = cx_max(W, H), kMinDim = cx_min(W, H);
  // up-shift before the column pass
  static constexpr int kUp = (kMaxDim == 64 && !(W == 64 && H == 32) && !(W == 64 && H == 16)) ? 0 : 2;
  // down-shift after the column pass
  static constexpr int kDown1 =
      (W == 4 && H == 4)                                  ? 0
      : (kMaxDim == 64 && kUp == 0)                       ? 2  /* 64x64 32x64 16x64 */
      : (W == 64)                                         ? 4  /* 64x32 64x16 */
      : (kMaxDim == 8 || (kMaxDim == 16 && kMinDim == 4)) ? 1
      : (kMaxDim == 16 || (kMaxDim == 32 && kMinDim == 8)) ? 2
                                                          : 4; /* 32x32 16x32 32x16 */
  // down-shift after the row pass
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

  // kept region and lane mapping
  static constexpr int Wk = cx_min(W, 32), Hk = cx_min(H, 32);
  static constexpr int CW = cx_min(W, 32);  // columns handled per column iteration
  static constexpr int CI = W / CW;         // column iterations (2 for 64-wide)
  static constexpr int NT = cx_max(CW, Hk); // lanes per transform block
  static constexpr int BPW = 32 / NT;       // transform blocks per warp
  static constexpr int LD = W + 1;          // odd smem pitch: conflict-free both ways
  static constexpr int kTile = Hk * LD;     // int32 words of smem per block
};

static const TxInfo kTxInfo[B200_TX_SIZES_ALL] = {
#define TXI(W, H)                                                                          \
  { W, H, TxCfg<W, H>::Wk, TxCfg<W, H>::Hk, { TxCfg<W, H>::kUp, -TxCfg<W, H>::kDown1,      \
    -TxCfg<W, H>::kDown2 }, TxCfg<W, H>::kCosCol, TxCfg<W, H>::kCosRow, TxCfg<W, H>::kRect, \
    TxCfg<W, H>::kLogScale }
  TXI(4, 4),   TXI(8, 8),   TXI(16, 16), TXI(32, 32), TXI(64, 64), TXI(4, 8),  TXI(8, 4),
  TXI(8, 16),  TXI(16, 8),  TXI(16, 32), TXI(32, 16), TXI(32, 64), TXI(64, 32), TXI(4, 16),
  TXI(16, 4),  TXI(8, 32),  TXI(32, 8),  TXI(16, 64), TXI(64, 16),
#undef TXI
};
const TxInfo &tx_info(int tx_size) { return kTxInfo[tx_size]; }

// ------------------------------------------------------------- kernel args -----
struct TxqArgs {
  const B200TxBlock *blocks;
  const B200TxPred *preds;
  const void *src;
  const void *pred;
  const B200QRow *qrows;
  const int16_t *iscan3;  // default, mrow, mcol inverse scans of this tx size
  int32_t *coeff, *qcoeff, *dqcoeff;
  uint16_t *eob;
  int64_t *dist;
  int n_blocks;
  int in_kind, q_kind, bd;
  unsigned flags;
};

constexpr int kWarpsPerCta = 4;

template <int N, int BIT, bool EXACT, int KEEP>
__device__ __forceinline__ void txfm1d(int kind, int32_t *x, int32_t *y) {
  using B = Btf<BIT, EXACT>;
  if constexpr (N == 64) {
    fdct<64, B, KEEP>(x, y);  // only DCT exists at 64 (av1_txfm.c:89-100)
  } else if constexpr (N == 32) {
    if (kind == 3) fidentity<32>(x, y);
    else fdct<32, B, KEEP>(x, y);
  } else {
    if (kind == 0) fdct<N, B, KEEP>(x, y);
    else if (kind == 3) fidentity<N>(x, y);
    else fadst<N, B>(x, y);
  }
}

template <int W, int H, bool EXACT>
__global__ void __launch_bounds__(kWarpsPerCta * 32) txq_kernel(const TxqArgs a) {
  using C = TxCfg<W, H>;
  extern __shared__ int32_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int slot = lane / C::NT, t = lane % C::NT;
  int32_t *tile = smem + (warp * C::BPW + slot) * C::kTile;
  const int n_tiles = (a.n_blocks + C::BPW - 1) / C::BPW;
  const int warp_stride = gridDim.x * kWarpsPerCta;
  const unsigned group_mask =
      C::NT == 32 ? 0xffffffffu : (((1u << C::NT) - 1u) << (slot * C::NT));
  (void)group_mask;

  for (int wt = blockIdx.x * kWarpsPerCta + warp; wt < n_tiles; wt += warp_stride) {
    const int bi = wt * C::BPW + slot;
    const bool active = bi < a.n_blocks;
    B200TxBlock d;
    if (active) {
      const uint4 raw = *reinterpret_cast<const uint4 *>(a.blocks + bi);
      memcpy(&d, &raw, sizeof(d));
    } else {
      memset(&d, 0, sizeof(d));
    }
    const int vt = c_vtx[d.tx_type & 15], ht = c_htx[d.tx_type & 15];

    // ------------------------------------------------------------ column pass
    if (active && t < C::CW) {
      const int ud = vt == 2;
      const int stride = d.src_stride;
      const int row0 = ud ? (H - 1) : 0;
      const int rstep = ud ? -stride : stride;
#pragma unroll
      for (int ci = 0; ci < C::CI; ++ci) {
        const int c = t + ci * C::CW;
        int32_t x[H], y[C::Hk];
        if (a.in_kind == B200_IN_RESIDUAL) {
          const int16_t *p = static_cast<const int16_t *>(a.src) + d.src_offset + row0 * stride + c;
#pragma unroll
          for (int r = 0; r < H; ++r) x[r] = p[r * rstep];
        } else {
          const B200TxPred pd = a.preds[bi];
          const int pstride = pd.pred_stride;
          const int prow0 = ud ? (H - 1) * pstride : 0;
          const int pstep = ud ? -pstride : pstride;
          if (a.in_kind == B200_IN_PIXELS8) {
            const uint8_t *s = static_cast<const uint8_t *>(a.src) + d.src_offset + row0 * stride + c;
            const uint8_t *q = static_cast<const uint8_t *>(a.pred) + pd.pred_offset + prow0 + c;
#pragma unroll
            for (int r = 0; r < H; ++r) x[r] = (int)s[r * rstep] - (int)q[r * pstep];
          } else {
            const uint16_t *s = static_cast<const uint16_t *>(a.src) + d.src_offset + row0 * stride + c;
            const uint16_t *q = static_cast<const uint16_t *>(a.pred) + pd.pred_offset + prow0 + c;
#pragma unroll
            for (int r = 0; r < H; ++r) x[r] = (int16_t)((int)s[r * rstep] - (int)q[r * pstep]);
          }
        }
        // av1_round_shift_array(temp_in, H, -shift[0]): int16 input, no clamp can trigger
#pragma unroll
        for (int r = 0; r < H; ++r) x[r] = (int32_t)((uint32_t)x[r] << C::kUp);
        txfm1d<H, C::kCosCol, EXACT, C::Hk>(vt, x, y);
        const int cc = (ht == 2) ? (W - 1 - c) : c;  // lr_flip (av1_fwd_txfm2d.c:99-105)
#pragma unroll
        for (int r = 0; r < C::Hk; ++r) tile[r * C::LD + cc] = rshift_round<C::kDown1>(y[r]);
      }
    }
    __syncwarp();

    // --------------------------------------------------------------- row pass
    int eob = 0;
    int64_t err = 0, ssz = 0;
    if (active && t < C::Hk) {
      int32_t x[W], y[C::Wk];
#pragma unroll
      for (int c = 0; c < W; ++c) x[c] = tile[t * C::LD + c];
      txfm1d<W, C::kCosRow, EXACT, C::Wk>(ht, x, y);
#pragma unroll
      for (int c = 0; c < C::Wk; ++c) {
        int32_t v = rshift_round<C::kDown2>(y[c]);
        if (C::kRect) v = scale_sqrt2<1>(v);
        y[c] = v;
      }
      const size_t obase = (size_t)d.out_offset + t;
      if (a.flags & B200_TXQ_STORE_COEFF) {
#pragma unroll
        for (int c = 0; c < C::Wk; ++c) a.coeff[obase + c * C::Hk] = y[c];
      }
      if (a.q_kind != B200_Q_NONE) {
        const B200QRow qr = a.qrows[d.qrow];
        const int cls = d.tx_type < 10 ? 0 : ((d.tx_type & 1) ? 2 : 1);
        const int16_t *iscan = a.iscan3 + cls * (C::Wk * C::Hk) + t;
        constexpr int LS = C::kLogScale;
        const bool fp = a.q_kind == B200_Q_FP || a.q_kind == B200_Q_FP_HBD;
        const bool hbd = a.q_kind == B200_Q_FP_HBD || a.q_kind == B200_Q_B_HBD;
        // per-block constants for DC (suffix 0) and AC (suffix 1)
        auto rpo2 = [](int v) { return LS ? ((v + (1 << (LS - 1))) >> LS) : v; };
        const int rnd0 = rpo2(fp ? qr.round_fp[0] : qr.round[0]);
        const int rnd1 = rpo2(fp ? qr.round_fp[1] : qr.round[1]);
        const int qnt0 = fp ? qr.quant_fp[0] : qr.quant[0];
        const int qnt1 = fp ? qr.quant_fp[1] : qr.quant[1];
        const int deq0 = qr.dequant[0], deq1 = qr.dequant[1];
        const int qsh0 = qr.quant_shift[0], qsh1 = qr.quant_shift[1];
        const int thr0 = fp ? deq0 : rpo2(qr.zbin[0]);
        const int thr1 = fp ? deq1 : rpo2(qr.zbin[1]);
        const bool with_err = (a.flags & B200_TXQ_BLOCK_ERROR) != 0;
        const int err_shift = hbd ? 2 * (a.bd - 8) : 0;
        (void)err_shift;
#pragma unroll
        for (int c = 0; c < C::Wk; ++c) {
          const bool dc = (c == 0) && (t == 0);  // compile-time false for c > 0
          const int rnd_k = dc ? rnd0 : rnd1, qnt_k = dc ? qnt0 : qnt1, deq_k = dc ? deq0 : deq1;
          const int qsh_k = dc ? qsh0 : qsh1, thr_k = dc ? thr0 : thr1;
          const int32_t cf = y[c];
          const int32_t sign = cf >> 31;
          const int32_t ab = (cf ^ sign) - sign;
          int32_t q = 0, dq = 0;
          if (fp) {
            // av1_quantize_fp_no_qmatrix :36-68 / highbd_quantize_fp_helper_c :170-196
            if ((int32_t)((uint32_t)ab << (1 + LS)) >= thr_k || (!hbd && ab >= (1 << 29))) {
              if (!hbd) {
                // clamp64(abs + round, INT16_MIN, INT16_MAX); min() first keeps it in 32 bits
                int32_t tt = (ab > 32767 ? 32767 : ab) + rnd_k;
                tt = tt > 32767 ? 32767 : (tt < -32768 ? -32768 : tt);
                q = (tt * qnt_k) >> (16 - LS);
              } else {
                const int64_t tt = (int64_t)ab + rnd_k;
                q = (int32_t)((tt * qnt_k) >> (16 - LS));
              }
              dq = wmul(q, deq_k) >> LS;
            }
          } else {
            // aom_quantize_b_helper_c :108-169 / aom_highbd_quantize_b_helper_c :261-316
            if (wmul(ab, 32) >= (thr_k << 5)) {
              int64_t tt = (int64_t)ab + rnd_k;
              if (!hbd) tt = tt > 32767 ? 32767 : tt;
              tt *= 32;
              q = (int32_t)(((((tt * qnt_k) >> 16) + tt) * qsh_k) >> (16 - LS + 5));
              dq = wmul(q, deq_k) >> LS;
            }
          }
          const int32_t qs = (q ^ sign) - sign, dqs = (dq ^ sign) - sign;
          a.qcoeff[obase + c * C::Hk] = qs;
          a.dqcoeff[obase + c * C::Hk] = dqs;
          if (q) {
            const int pos = iscan[c * C::Hk] + 1;
            eob = pos > eob ? pos : eob;
          }
          if (with_err) {
            const int32_t df = wsub(cf, dqs);
            if (hbd) {
              err += (int64_t)df * df;
              ssz += (int64_t)cf * cf;
            } else {  // av1_block_error_c squares in 32-bit int (rdopt.c:642-643)
              err += (int64_t)wmul(df, df);
              ssz += (int64_t)wmul(cf, cf);
            }
          }
        }
      }
    }
    if (a.q_kind != B200_Q_NONE) {
      // reduce over the NT lanes of the block (xor shuffles stay inside the group)
#pragma unroll
      for (int o = C::NT / 2; o > 0; o >>= 1) {
        const int e2 = __shfl_xor_sync(0xffffffffu, eob, o);
        eob = e2 > eob ? e2 : eob;
      }
      if (a.flags & B200_TXQ_BLOCK_ERROR) {
#pragma unroll
        for (int o = C::NT / 2; o > 0; o >>= 1) {
          err += __shfl_xor_sync(0xffffffffu, err, o);
          ssz += __shfl_xor_sync(0xffffffffu, ssz, o);
        }
      }
      if (active && t == 0) {
        a.eob[bi] = (uint16_t)eob;
        if ((a.flags & B200_TXQ_BLOCK_ERROR) && a.dist) {
          const bool hbd = a.q_kind == B200_Q_FP_HBD || a.q_kind == B200_Q_B_HBD;
          if (hbd) {
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
    __syncwarp();
  }
}

// ------------------------------------------------------------------ host side --

typedef void (*TxqKernel)(const TxqArgs);
struct KernelEntry {
  TxqKernel fn[2];  // [exact, fast32]
  int smem_bytes;
  int bpw;
  int max_ctas_per_sm[2];
};

template <int W, int H>
static KernelEntry make_entry() {
  KernelEntry e;
  e.fn[0] = txq_kernel<W, H, true>;
  e.fn[1] = txq_kernel<W, H, false>;
  e.smem_bytes = kWarpsPerCta * TxCfg<W, H>::BPW * TxCfg<W, H>::kTile * (int)sizeof(int32_t);
  e.bpw = TxCfg<W, H>::BPW;
  e.max_ctas_per_sm[0] = e.max_ctas_per_sm[1] = 0;
  return e;
}

static KernelEntry g_entries[B200_TX_SIZES_ALL];
static std::once_flag g_entries_once;
static void init_entries() {
  const KernelEntry e[B200_TX_SIZES_ALL] = {
    make_entry<4, 4>(),   make_entry<8, 8>(),   make_entry<16, 16>(), make_entry<32, 32>(),
    make_entry<64, 64>(), make_entry<4, 8>(),   make_entry<8, 4>(),   make_entry<8, 16>(),
    make_entry<16, 8>(),  make_entry<16, 32>(), make_entry<32, 16>(), make_entry<32, 64>(),
    make_entry<64, 32>(), make_entry<4, 16>(),  make_entry<16, 4>(),  make_entry<8, 32>(),
    make_entry<32, 8>(),  make_entry<16, 64>(), make_entry<64, 16>(),
  };
  memcpy(g_entries, e, sizeof(e));
}

// Scan orders.  The codec stores coefficients column-major (position c*Hk + r).  Three
// classes exist (av1_scan_orders, scan.c:1666): "default" walks the anti-diagonals
// (tall blocks with r increasing, wide blocks with c increasing, square blocks
// alternating, odd diagonals r-increasing), "mrow" (V_* types) walks row by row,
// "mcol" (H_* types) is storage order.  64-point sizes use their kept 32-point region.
static std::vector<int16_t> g_scan[B200_TX_SIZES_ALL][3], g_iscan[B200_TX_SIZES_ALL][3];
static size_t g_iscan_dev_offset[B200_TX_SIZES_ALL];
static size_t g_iscan_total;
static std::once_flag g_scan_once;

static void build_scans() {
  size_t off = 0;
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    const int wk = kTxInfo[ts].wk, hk = kTxInfo[ts].hk, n = wk * hk;
    for (int cls = 0; cls < 3; ++cls) {
      std::vector<int16_t> s;
      s.reserve(n);
      if (cls == 2) {
        for (int i = 0; i < n; ++i) s.push_back((int16_t)i);
      } else if (cls == 1) {
        for (int r = 0; r < hk; ++r)
          for (int c = 0; c < wk; ++c) s.push_back((int16_t)(c * hk + r));
      } else {
        for (int dgn = 0; dgn < wk + hk - 1; ++dgn) {
          const bool r_up = hk > wk ? true : (wk > hk ? false : (dgn & 1) != 0);
          for (int k = 0; k <= dgn; ++k) {
            const int r = r_up ? k : dgn - k, c = dgn - r;
            if (r < hk && c < wk) s.push_back((int16_t)(c * hk + r));
          }
        }
      }
      std::vector<int16_t> inv(n);
      for (int i = 0; i < n; ++i) inv[s[i]] = (int16_t)i;
      g_scan[ts][cls] = s;
      g_iscan[ts][cls] = inv;
    }
    g_iscan_dev_offset[ts] = off;
    off += 3 * (size_t)n;
  }
  g_iscan_total = off;
}

static int scan_class(int tx_type) { return tx_type < 10 ? 0 : ((tx_type & 1) ? 2 : 1); }

int txq_init_device(b200_ctx *ctx) {
  std::call_once(g_entries_once, init_entries);
  std::call_once(g_scan_once, build_scans);
  int32_t cosv[4][64], ncosv[4][64];
  const double pi = 3.14159265358979323846;
  for (int b = 0; b < 4; ++b)
    for (int j = 0; j < 64; ++j) {
      cosv[b][j] = (int32_t)floor(cos(pi * j / 128.0) * (double)(1 << (10 + b)) + 0.5);
      ncosv[b][j] = -cosv[b][j];
    }
  // av1_sinpi_arr_data (av1_txfm.c:62-69) for bits 10..13: round(sqrt2*sin(j*pi/9)*2/3 * 2^bit),
  // entries adjusted by the codec so that s1 + s2 == s4.
  const int32_t sinv[4][5] = { { 0, 330, 621, 836, 951 },
                               { 0, 660, 1241, 1672, 1901 },
                               { 0, 1321, 2482, 3344, 3803 },
                               { 0, 2642, 4964, 6689, 7606 } };
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_cospi, cosv, sizeof(cosv)));
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_ncospi, ncosv, sizeof(ncosv)));
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_sinpi, sinv, sizeof(sinv)));
  std::vector<int16_t> all(g_iscan_total);
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int cls = 0; cls < 3; ++cls)
      memcpy(all.data() + g_iscan_dev_offset[ts] + cls * g_iscan[ts][cls].size(),
             g_iscan[ts][cls].data(), g_iscan[ts][cls].size() * sizeof(int16_t));
  B200_CUDA_TRY(cudaMalloc(&ctx->d_iscan, all.size() * sizeof(int16_t)));
  B200_CUDA_TRY(cudaMemcpy(ctx->d_iscan, all.data(), all.size() * sizeof(int16_t),
                           cudaMemcpyHostToDevice));
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int v = 0; v < 2; ++v) {
      KernelEntry &e = g_entries[ts];
      B200_CUDA_TRY(cudaFuncSetAttribute(e.fn[v], cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         e.smem_bytes));
      int nb = 0;
      B200_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, e.fn[v], kWarpsPerCta * 32,
                                                                  e.smem_bytes));
      e.max_ctas_per_sm[v] = nb > 0 ? nb : 1;
    }
  return B200_OK;
}

int txq_launch(b200_ctx *ctx, const B200TxqLaunch &l, cudaStream_t s) {
  if (l.tx_size < 0 || l.tx_size >= B200_TX_SIZES_ALL || l.n_blocks < 0 || !l.blocks || !l.src) {
    set_error("b200_txq_launch: bad tx_size / n_blocks / null blocks or src");
    return B200_EINVAL;
  }
  if (l.in_kind != B200_IN_RESIDUAL && (!l.preds || !l.pred)) {
    set_error("b200_txq_launch: pixel input needs preds and pred");
    return B200_EINVAL;
  }
  if (l.q_kind != B200_Q_NONE && (!l.qrows || !l.qcoeff || !l.dqcoeff || !l.eob)) {
    set_error("b200_txq_launch: quantiser selected but qrows/qcoeff/dqcoeff/eob missing");
    return B200_EINVAL;
  }
  if ((l.flags & B200_TXQ_STORE_COEFF) && !l.coeff) {
    set_error("b200_txq_launch: STORE_COEFF without coeff buffer");
    return B200_EINVAL;
  }
  if (l.n_blocks == 0) return B200_OK;
  const int v = (l.flags & B200_TXQ_FAST32) ? 1 : 0;
  const KernelEntry &e = g_entries[l.tx_size];
  TxqArgs a;
  a.blocks = l.blocks;
  a.preds = l.preds;
  a.src = l.src;
  a.pred = l.pred;
  a.qrows = l.qrows;
  a.iscan3 = ctx->d_iscan + g_iscan_dev_offset[l.tx_size];
  a.coeff = l.coeff;
  a.qcoeff = l.qcoeff;
  a.dqcoeff = l.dqcoeff;
  a.eob = l.eob;
  a.dist = l.dist;
  a.n_blocks = l.n_blocks;
  a.in_kind = l.in_kind;
  a.q_kind = l.q_kind;
  a.bd = l.bd;
  a.flags = l.flags;
  const int n_tiles = (l.n_blocks + e.bpw - 1) / e.bpw;
  int grid = (n_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
  const int resident = ctx->sm_count * e.max_ctas_per_sm[v];
  if (grid > resident) grid = resident;  // persistent: warps stride over the tiles
  e.fn[v]<<<grid, kWarpsPerCta * 32, e.smem_bytes, s>>>(a);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

// host-visible table accessors (tests, host mirrors)
extern "C" int b200_tx_wide(int tx_size) { return b200::kTxInfo[tx_size].w; }
extern "C" int b200_tx_high(int tx_size) { return b200::kTxInfo[tx_size].h; }
extern "C" int b200_tx_max_eob(int tx_size) {
  return b200::kTxInfo[tx_size].wk * b200::kTxInfo[tx_size].hk;
}
extern "C" int b200_tx_valid(int tx_size, int tx_type) {
  if (tx_size < 0 || tx_size >= B200_TX_SIZES_ALL || tx_type < 0 || tx_type >= 16) return 0;
  const int dims[2] = { b200::kTxInfo[tx_size].h, b200::kTxInfo[tx_size].w };
  const int kinds[2] = { b200::kVtxHost[tx_type], b200::kHtxHost[tx_type] };
  for (int i = 0; i < 2; ++i) {
    if (dims[i] == 64 && kinds[i] != 0) return 0;
    if (dims[i] == 32 && (kinds[i] == 1 || kinds[i] == 2)) return 0;
  }
  return 1;
}
extern "C" const int16_t *b200_scan(int tx_size, int tx_type) {
  std::call_once(b200::g_scan_once, b200::build_scans);
  return b200::g_scan[tx_size][b200::scan_class(tx_type)].data();
}
extern "C" const int16_t *b200_iscan(int tx_size, int tx_type) {
  std::call_once(b200::g_scan_once, b200::build_scans);
  return b200::g_iscan[tx_size][b200::scan_class(tx_type)].data();
}
