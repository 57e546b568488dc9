// txq_small.cuh -- fused (subtract +) forward 2-D transform (+ quantise) (+ block error)
// for the transform sizes of at most 64 samples (4x4, 4x8, 8x4, 8x8, 4x16, 16x4).
//
// Same contract as txq_kernel.cuh (same reference chain, same bit-exact results, same
// launch arguments); different mapping.  For these sizes the tile machinery of the big
// kernel costs more instructions than the arithmetic, so here ONE LANE OWNS ONE WHOLE
// BLOCK: the W*H residuals sit in that lane's registers from the global load to the
// coefficient stores.  There is no shared memory, no transpose and no barrier:
//   1. the lane reads its block's rows with 4..16-byte global loads (adjacent lanes ->
//      adjacent blocks -> adjacent addresses), forms the residual for pixel inputs and
//      checks the fast-arithmetic guard with packed 16x2 min/max;
//   2. column transforms, then row transforms, run in registers (flips are static
//      re-indexings of the register file, selected by a switch on the 1-D kind);
//   3. the W*H coefficients -- stored column-major by the codec, i.e. contiguous per
//      lane -- are quantised in register groups of eight and leave through 256-bit
//      stores (one full 32-byte sector per lane per instruction);
//   4. eob: a bitmask of the non-zero positions in storage order is searched against
//      prefix masks of the scan order (log2(N)+1 steps) instead of a per-coefficient
//      table walk.
// The descriptor of the warp's next group of 32 blocks is fetched one iteration ahead.
#pragma once
#include "txq_kernel.cuh"

namespace b200 {

// ----------------------------------------------------------------- memory helpers --
__device__ __forceinline__ void stg256(int32_t *p, const int32_t *x) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(x[0]), "r"(x[1]),
               "r"(x[2]), "r"(x[3]), "r"(x[4]), "r"(x[5]), "r"(x[6]), "r"(x[7])
               : "memory");
}
static __device__ __noinline__ void store8_scalar(int32_t *p, int32_t x0, int32_t x1, int32_t x2,
                                                 int32_t x3, int32_t x4, int32_t x5, int32_t x6,
                                                 int32_t x7) {
  p[0] = x0, p[1] = x1, p[2] = x2, p[3] = x3, p[4] = x4, p[5] = x5, p[6] = x6, p[7] = x7;
}
// eight consecutive int32 results: one 256-bit store when the address allows it (the
// fallback is an out-of-line call so that the hot path carries no predicated-off stores)
__device__ __forceinline__ void store8(int32_t *p, const int32_t *x, bool vec) {
  if (vec) stg256(p, x);
  else store8_scalar(p, x[0], x[1], x[2], x[3], x[4], x[5], x[6], x[7]);
}
// Output staging.  A lane's results are contiguous in global memory (N int32 per block)
// but consecutive lanes are N*4 bytes apart, so direct stores would touch 32 cache lines
// per instruction.  Instead every lane parks its values in a padded row of a per-warp
// shared-memory buffer and the warp then writes the 32 blocks out with 16-byte stores in
// which consecutive lanes cover consecutive addresses of one block.
#ifndef B200_SMALL_STAGED_MIN_N
#define B200_SMALL_STAGED_MIN_N 32  // stage outputs through shared memory for N >= this
#endif
template <int N>
struct StageBuf {
  static constexpr bool kOn = N >= B200_SMALL_STAGED_MIN_N;
  static constexpr int kPitch = N + 4;               // words; 16-byte rows, conflict-free
  static constexpr int kWords = 32 * kPitch;         // one array, one warp
};
__device__ __forceinline__ void stage8(int32_t *row, const int32_t *x) {
  *reinterpret_cast<int4 *>(row) = make_int4(x[0], x[1], x[2], x[3]);
  *reinterpret_cast<int4 *>(row + 4) = make_int4(x[4], x[5], x[6], x[7]);
}
// buf: the warp's staging buffer; out: array base; ob: this lane's block offset (elements);
// nact: number of valid blocks among the warp's 32 (lanes 0..nact-1)
template <int N>
static __device__ __noinline__ void flush_staged(const int32_t *buf, int32_t *out, uint32_t ob, int nact,
                                                 bool vec, int lane) {
  constexpr int CPB = N / 4;        // 16-byte chunks per block
  constexpr int BPI = 32 / CPB;     // blocks covered by one store instruction
  const int sub = lane / CPB, e4 = lane % CPB;
  __syncwarp();
#pragma unroll
  for (int j = 0; j < CPB; ++j) {
    const int b = j * BPI + sub;
    const uint32_t obb = __shfl_sync(0xffffffffu, ob, b);
    const int4 t = *reinterpret_cast<const int4 *>(buf + b * StageBuf<N>::kPitch + 4 * e4);
    if (b < nact) {
      int32_t *dst = out + (size_t)obb + 4 * e4;
      if (vec) {
        *reinterpret_cast<int4 *>(dst) = t;
      } else {
        dst[0] = t.x, dst[1] = t.y, dst[2] = t.z, dst[3] = t.w;
      }
    }
  }
  __syncwarp();
}

// The common case: the warp's 32 blocks are consecutive in the output arrays (block b at
// run + b*N) and the tile is full -- the whole tile is one contiguous run of 32*N int32
// and chunk j of lane l is simply run[(j*32 + l)*4 ..]; every address is an immediate
// offset from one per-lane pointer on both the shared and the global side.
template <int N>
__device__ __forceinline__ void flush_staged_run(const int32_t *buf, int32_t *run, int lane) {
  constexpr int CPB = N / 4, BPI = 32 / CPB;
  const int32_t *src = buf + (lane / CPB) * StageBuf<N>::kPitch + 4 * (lane % CPB);
  int32_t *dst = run + 4 * lane;
  __syncwarp();
#pragma unroll
  for (int j = 0; j < CPB; ++j)
    *reinterpret_cast<int4 *>(dst + j * 128) =
        *reinterpret_cast<const int4 *>(src + j * BPI * StageBuf<N>::kPitch);
  __syncwarp();
}

// BYTES (4, 8, 16 or 32) from an address aligned to min(BYTES, 16)
template <int BYTES>
__device__ __forceinline__ void ldg_row(const unsigned char *p, uint32_t *w) {
  if constexpr (BYTES == 4) {
    w[0] = __ldg(reinterpret_cast<const uint32_t *>(p));
  } else if constexpr (BYTES == 8) {
    const uint2 t = __ldg(reinterpret_cast<const uint2 *>(p));
    w[0] = t.x, w[1] = t.y;
  } else {
#pragma unroll
    for (int i = 0; i < BYTES / 16; ++i) {
      const uint4 t = __ldg(reinterpret_cast<const uint4 *>(p) + i);
      w[4 * i] = t.x, w[4 * i + 1] = t.y, w[4 * i + 2] = t.z, w[4 * i + 3] = t.w;
    }
  }
}

// One operand plane of the block -> raw[H][W*ESZ/4] words (row-major, little endian).
template <int W, int H, int ESZ>
__device__ __forceinline__ void load_plane(const void *plane, uint32_t offset, int stride,
                                           uint32_t *raw) {
  constexpr int RB = W * ESZ, WPR = RB / 4, VB = RB < 16 ? RB : 16;
  const unsigned char *p0 = static_cast<const unsigned char *>(plane) + (size_t)offset * ESZ;
  const int pitch = stride * ESZ;
  if (((reinterpret_cast<uintptr_t>(p0) | (uintptr_t)pitch) & (VB - 1)) == 0) {
#pragma unroll
    for (int r = 0; r < H; ++r) ldg_row<RB>(p0 + (size_t)r * pitch, raw + r * WPR);
  } else {  // unaligned block: element loads, packed the same way
#pragma unroll
    for (int r = 0; r < H; ++r) {
      const unsigned char *p = p0 + (size_t)r * pitch;
#pragma unroll
      for (int k = 0; k < WPR; ++k) {
        if constexpr (ESZ == 2) {
          const unsigned short *q = reinterpret_cast<const unsigned short *>(p) + 2 * k;
          raw[r * WPR + k] = (uint32_t)q[0] | ((uint32_t)q[1] << 16);
        } else {
          raw[r * WPR + k] = (uint32_t)p[4 * k] | ((uint32_t)p[4 * k + 1] << 8) |
                             ((uint32_t)p[4 * k + 2] << 16) | ((uint32_t)p[4 * k + 3] << 24);
        }
      }
    }
  }
}

// Raw operand words of one block, as loaded (kept in registers one iteration ahead).
template <int W, int H, int IN>
struct RawBlock {
  static constexpr int ESZ = IN == B200_IN_PIXELS8 ? 1 : 2;
  static constexpr int NW = W * H * ESZ / 4;
  uint32_t s[NW];
  uint32_t p[IN == B200_IN_RESIDUAL ? 1 : NW];
};

template <int W, int H, int IN>
__device__ __forceinline__ void fetch_block(const TxqArgs &a, const B200TxBlock &d, int bi,
                                            RawBlock<W, H, IN> &raw) {
  load_plane<W, H, RawBlock<W, H, IN>::ESZ>(a.src, d.src_offset, d.src_stride, raw.s);
  if constexpr (IN != B200_IN_RESIDUAL) {
    const B200TxPred pd = a.preds[bi];
    load_plane<W, H, RawBlock<W, H, IN>::ESZ>(a.pred, pd.pred_offset, pd.pred_stride, raw.p);
  }
}

// The block's residuals as v[r*W + c] (no up-shift yet).  Returns whether every residual
// is inside the fast-arithmetic guard.
template <int W, int H, int IN>
__device__ __forceinline__ bool unpack_block(const RawBlock<W, H, IN> &raw, int32_t *v) {
  constexpr int N = W * H;
  uint32_t res[N / 2];  // packed int16 residual pairs (c even | c odd << 16)
  if constexpr (IN == B200_IN_RESIDUAL) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) res[i] = raw.s[i];
  } else if constexpr (IN == B200_IN_PIXELS16) {
#pragma unroll
    for (int i = 0; i < N / 2; ++i) res[i] = __vsub2(raw.s[i], raw.p[i]);  // (int16)(s - p)
  } else {
#pragma unroll
    for (int i = 0; i < N / 4; ++i) {
      res[2 * i] = __vsub2(__byte_perm(raw.s[i], 0, 0x4140), __byte_perm(raw.p[i], 0, 0x4140));
      res[2 * i + 1] = __vsub2(__byte_perm(raw.s[i], 0, 0x4342), __byte_perm(raw.p[i], 0, 0x4342));
    }
  }
  bool fits = true;
  if constexpr (IN != B200_IN_PIXELS8) {  // 8-bit differences are always conformant
    uint32_t mx = res[0], mn = res[0];
#pragma unroll
    for (int i = 1; i < N / 2; ++i) {
      mx = __vmaxs2(mx, res[i]);
      mn = __vmins2(mn, res[i]);
    }
    const uint32_t hi = __vmaxs2(mx, __byte_perm(mx, 0, 0x1032));  // max of both halves, twice
    const uint32_t lo = __vmins2(mn, __byte_perm(mn, 0, 0x1032));
    fits = (int16_t)(hi & 0xffff) <= kGuard && (int16_t)(lo & 0xffff) >= -kGuard;
  }
#pragma unroll
  for (int i = 0; i < N / 2; ++i) {
    v[2 * i] = (int32_t)(int16_t)(res[i] & 0xffff);
    v[2 * i + 1] = (int32_t)res[i] >> 16;
  }
  return fits;
}

// ------------------------------------------------------------------- transforms ----
// one N-point line; KIND: 0 DCT, 1/2 ADST (the flip is applied by the caller's indexing), 3 IDTX
template <int N, int BIT, bool EXACT, int KIND>
__device__ __forceinline__ void line1d(int32_t *x, int32_t *y) {
  using B = Btf<BIT, EXACT>;
  if constexpr (KIND == 0) {
    fdct<N, B, N>(x, y);
  } else if constexpr (KIND == 3) {
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

// v[r*W + c], in place: all W columns (COLS) or all H rows of the block
template <int W, int H, bool EXACT, bool COLS, int KIND>
__device__ __forceinline__ void pass_kind(int32_t *v) {
  using C = TxCfg<W, H>;
  constexpr int LEN = COLS ? H : W, LINES = COLS ? W : H;
  static_for<0, LINES>([&](auto lc) {
    constexpr int l = decltype(lc)::value;
    int32_t x[LEN], y[LEN];
    static_for<0, LEN>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      constexpr int s = KIND == 2 ? LEN - 1 - i : i;  // ud_flip / lr_flip (av1_fwd_txfm2d.c:86,99)
      if constexpr (COLS) x[i] = (int32_t)((uint32_t)v[s * W + l] << C::kUp);
      else x[i] = v[l * W + s];
    });
    line1d<LEN, COLS ? C::kCosCol : C::kCosRow, EXACT, KIND>(x, y);
    static_for<0, LEN>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      if constexpr (COLS) {
        v[i * W + l] = down_shift<EXACT, C::kDown1>(y[i]);
      } else {
        int32_t t = down_shift<EXACT, C::kDown2>(y[i]);
        if constexpr (C::kRect) t = sqrt2_scale<EXACT, 1>(t);
        v[l * W + i] = t;
      }
    });
  });
}

template <int W, int H, bool EXACT, bool COLS>
__device__ __forceinline__ void pass(int kind, int32_t *v) {
  // an if-chain: a switch becomes an indexed constant load + BRX, a visible stall here
  if (kind == 0) pass_kind<W, H, EXACT, COLS, 0>(v);
  else if (kind == 1) pass_kind<W, H, EXACT, COLS, 1>(v);
  else if (kind == 2) pass_kind<W, H, EXACT, COLS, 2>(v);
  else pass_kind<W, H, EXACT, COLS, 3>(v);
}

// --------------------------------------------------------------------- epilogue ----
// Quantise the lane's block.  v[r*W + c] holds coefficient (c, r); storage index is
// i = c*H + r.  `nz` collects bit i for every non-zero quantised coefficient.  Results go
// to the lane's rows of the warp's staging buffers (sq, sdq), see flush_staged.
template <int W, int H, int QK, bool ERR>
__device__ __forceinline__ void small_quant(const B200QRow &qr, const int32_t *v, int32_t *sq,
                                            int32_t *sdq, bool vec, unsigned long long &nz,
                                            int64_t &err, int64_t &ssz) {
  using C = TxCfg<W, H>;
  static_assert(C::kLogScale == 0, "small sizes never use the 32x32 / 64x64 quantiser flavours");
  constexpr int N = W * H;
  constexpr bool fp = QK == B200_Q_FP || QK == B200_Q_FP_HBD;
  constexpr bool hbd = QK == B200_Q_FP_HBD || QK == B200_Q_B_HBD;
  const int rnd0 = fp ? qr.round_fp[0] : qr.round[0], rnd1 = fp ? qr.round_fp[1] : qr.round[1];
  const int qnt0 = fp ? qr.quant_fp[0] : qr.quant[0], qnt1 = fp ? qr.quant_fp[1] : qr.quant[1];
  const int deq0 = qr.dequant[0], deq1 = qr.dequant[1];
  const int qsh0 = qr.quant_shift[0], qsh1 = qr.quant_shift[1];
  // dead zone as |coeff| >= threshold: FP (abs << 1) >= dequant; B abs >= zbin
  const int thr0 = fp ? ((deq0 + 1) >> 1) : qr.zbin[0];
  const int thr1 = fp ? ((deq1 + 1) >> 1) : qr.zbin[1];
  uint32_t nz_lo = 0, nz_hi = 0;
  static_for<0, N / 8>([&](auto gc) {
    constexpr int g = decltype(gc)::value;
    int32_t qs[8], dqs[8];
    static_for<0, 8>([&](auto jc) {
      constexpr int j = decltype(jc)::value;
      constexpr int i = 8 * g + j, c = i / H, r = i % H;
      constexpr bool dc = i == 0;
      const int rnd_k = dc ? rnd0 : rnd1, qnt_k = dc ? qnt0 : qnt1, deq_k = dc ? deq0 : deq1;
      const int qsh_k = dc ? qsh0 : qsh1, thr_k = dc ? thr0 : thr1;
      (void)qsh_k;
      const int32_t cf = v[r * W + c];
      const int32_t ab = cf < 0 ? -cf : cf;
      int32_t q;
      if constexpr (fp) {
        if constexpr (!hbd) {
          // clamp64(abs + round, INT16_MIN, INT16_MAX): abs, round >= 0 -> one upper clamp
          int32_t tt = ab + rnd_k;
          tt = tt > 32767 ? 32767 : tt;
          tt = ab >= thr_k ? tt : 0;
          q = (tt * qnt_k) >> 16;
        } else {
          const uint32_t tt = ab >= thr_k ? (uint32_t)(ab + rnd_k) : 0u;
          q = (int32_t)(((uint64_t)tt * (uint32_t)qnt_k) >> 16);
        }
      } else {
        int64_t tt = (int64_t)ab + rnd_k;
        if constexpr (!hbd) tt = tt > 32767 ? 32767 : tt;
        tt = ab >= thr_k ? tt * 32 : 0;
        q = (int32_t)(((((tt * qnt_k) >> 16) + tt) * qsh_k) >> (16 + 5));
      }
      const int32_t sg = cf >> 31;              // 0 or -1
      const int32_t q_s = (q ^ sg) - sg;        // q * sign
      const int32_t dq_s = wmul(q_s, deq_k);    // (q * dequant) * sign, log_scale 0
      qs[j] = q_s;
      dqs[j] = dq_s;
      if (q) {
        if constexpr (i < 32) nz_lo |= 1u << i;
        else nz_hi |= 1u << (i - 32);
      }
      if constexpr (ERR) {
        const int32_t df = wsub(cf, dq_s);
        if constexpr (hbd) {
          err += (int64_t)df * df;
          ssz += (int64_t)cf * cf;
        } else {  // av1_block_error_c squares in 32-bit int (rdopt.c:642-643)
          err += (int64_t)wmul(df, df);
          ssz += (int64_t)wmul(cf, cf);
        }
      }
    });
    if constexpr (StageBuf<N>::kOn) {
      stage8(sq + 8 * g, qs);
      stage8(sdq + 8 * g, dqs);
    } else {  // sq / sdq are the block's global output rows
      store8(sq + 8 * g, qs, vec);
      store8(sdq + 8 * g, dqs, vec);
    }
  });
  nz = ((unsigned long long)nz_hi << 32) | nz_lo;
}

// --------------------------------------------------------------------- kernel ------
template <int N>
struct SmallSmem {
  static constexpr int kMaskBytes = cx_round_up(3 * (N + 1) * 8, 16);
  static constexpr int kCtaBytes =
      kMaskBytes + (StageBuf<N>::kOn ? kWarpsPerCta * 2 * StageBuf<N>::kWords * 4 : 0);
};

template <int W, int H, int IN>
#ifndef B200_SMALL_MINB64
#define B200_SMALL_MINB64 2
#endif
__global__ void __launch_bounds__(kWarpsPerCta * 32, (W * H == 64) ? B200_SMALL_MINB64 : (W * H == 32 ? 3 : 4))
    txq_small_kernel(const TxqArgs a) {
  constexpr int N = W * H;
  static_assert(N <= 64 && N % 8 == 0, "lane-owns-block mapping is for blocks of <= 64 samples");
  const int lane = threadIdx.x & 31;
  const int n_tiles = (a.n_blocks + 31) >> 5;  // a warp takes 32 blocks at a time
  const int wstride = gridDim.x * kWarpsPerCta;
  const bool force_exact = (a.flags & B200_TXQ_FORCE_EXACT) != 0;
  int wt = blockIdx.x * kWarpsPerCta + (threadIdx.x >> 5);

  // dynamic shared memory: scan prefix masks of the three scan classes (eob search
  // below), then two output staging buffers per warp
  extern __shared__ __align__(16) unsigned char smem_small[];
  unsigned long long *s_pm = reinterpret_cast<unsigned long long *>(smem_small);
  int32_t *sbuf0 = reinterpret_cast<int32_t *>(smem_small + SmallSmem<N>::kMaskBytes) +
                   (threadIdx.x >> 5) * 2 * StageBuf<N>::kWords;
  int32_t *sbuf1 = sbuf0 + StageBuf<N>::kWords;
  int32_t *row0 = sbuf0 + lane * StageBuf<N>::kPitch, *row1 = sbuf1 + lane * StageBuf<N>::kPitch;
  for (int i = threadIdx.x; i < 3 * (N + 1); i += blockDim.x) s_pm[i] = a.pmask3[i];
  __syncthreads();

  // software pipeline: raw descriptors two rounds ahead (decoded only when used), block
  // data and quantiser row one round ahead
  auto desc_raw = [&](int t) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (t < n_tiles && t * 32 + lane < a.n_blocks)
      r = __ldg(reinterpret_cast<const uint4 *>(a.blocks + t * 32 + lane));
    return r;
  };
  auto decode = [](const uint4 &r) {
    B200TxBlock d;
    memcpy(&d, &r, sizeof(d));
    return d;
  };
  const bool quant = a.q_kind != B200_Q_NONE;
  uint4 r0 = desc_raw(wt), r1 = desc_raw(wt + wstride);
  RawBlock<W, H, IN> rn;
  uint4 qn0 = make_uint4(0, 0, 0, 0), qn1 = qn0;
  if (wt * 32 + lane < a.n_blocks) {
    const B200TxBlock d = decode(r0);
    fetch_block<W, H, IN>(a, d, wt * 32 + lane, rn);
    if (quant) {
      qn0 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + d.qrow));
      qn1 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + d.qrow) + 1);
    }
  }
  for (; wt < n_tiles; wt += wstride) {
    const int bi = wt * 32 + lane;
    const bool act = bi < a.n_blocks;
    const B200TxBlock d = decode(r0);
    const RawBlock<W, H, IN> raw = rn;
    B200QRow qr;
    memcpy(&qr, &qn0, 16);
    memcpy(reinterpret_cast<char *>(&qr) + 16, &qn1, 16);
    r0 = r1;
    r1 = desc_raw(wt + 2 * wstride);
    if (wt + wstride < n_tiles && bi + wstride * 32 < a.n_blocks) {
      const B200TxBlock dn = decode(r0);
      fetch_block<W, H, IN>(a, dn, bi + wstride * 32, rn);
      if (quant) {
        qn0 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + dn.qrow));
        qn1 = __ldg(reinterpret_cast<const uint4 *>(a.qrows + dn.qrow) + 1);
      }
    }
    int32_t v[N];
    bool fits = true;
    if (act) fits = unpack_block<W, H, IN>(raw, v);
    const bool fast = !force_exact && __all_sync(0xffffffffu, fits);
    const int nact = a.n_blocks - wt * 32 < 32 ? a.n_blocks - wt * 32 : 32;
    const int tx_type = d.tx_type & 15;
    const int vt = vtx_of(tx_type), ht = htx_of(tx_type);
    if (act) {
      if (fast) {
        pass<W, H, false, true>(vt, v);
        pass<W, H, false, false>(ht, v);
      } else {
        pass<W, H, true, true>(vt, v);
        pass<W, H, true, false>(ht, v);
      }
    }
    const uint32_t ob = d.out_offset;
    const bool want_coeff = (a.flags & B200_TXQ_STORE_COEFF) != 0;
    uintptr_t al = 0;
    if (want_coeff) al |= reinterpret_cast<uintptr_t>(a.coeff + ob);
    if (quant) al |= reinterpret_cast<uintptr_t>(a.qcoeff + ob) | reinterpret_cast<uintptr_t>(a.dqcoeff + ob);
    const bool vec = __all_sync(0xffffffffu, !act || (al & (StageBuf<N>::kOn ? 15 : 31)) == 0);
    // staged output: are the warp's 32 output blocks one contiguous, 16-byte aligned run?
    const uint32_t ob0 = __shfl_sync(0xffffffffu, ob, 0);
    const bool run_ok = StageBuf<N>::kOn && vec &&
                        __all_sync(0xffffffffu, act && ob == ob0 + (uint32_t)lane * N);
    int32_t *oq = StageBuf<N>::kOn ? row0 : a.qcoeff + ob;
    int32_t *odq = StageBuf<N>::kOn ? row1 : a.dqcoeff + ob;
    if (want_coeff) {
      if (act) {
        static_for<0, N / 8>([&](auto gc) {
          constexpr int g = decltype(gc)::value;
          int32_t t[8];
          static_for<0, 8>([&](auto jc) {
            constexpr int i = 8 * g + decltype(jc)::value;
            t[decltype(jc)::value] = v[(i % H) * W + i / H];
          });
          if constexpr (StageBuf<N>::kOn) stage8(row0 + 8 * g, t);
          else store8(a.coeff + ob + 8 * g, t, vec);
        });
      }
      if constexpr (StageBuf<N>::kOn) {
        if (run_ok) flush_staged_run<N>(sbuf0, a.coeff + ob0, lane);
        else flush_staged<N>(sbuf0, a.coeff, ob, nact, vec, lane);
      }
    }
    if (!quant) continue;
    unsigned long long nz = 0;
    int64_t err = 0, ssz = 0;
    const bool we = (a.flags & B200_TXQ_BLOCK_ERROR) != 0;
    if (act) {
      if (a.q_kind == B200_Q_FP) {
        if (we) small_quant<W, H, B200_Q_FP, true>(qr, v, oq, odq, vec, nz, err, ssz);
        else small_quant<W, H, B200_Q_FP, false>(qr, v, oq, odq, vec, nz, err, ssz);
      } else if (a.q_kind == B200_Q_FP_HBD) {
        if (we) small_quant<W, H, B200_Q_FP_HBD, true>(qr, v, oq, odq, vec, nz, err, ssz);
        else small_quant<W, H, B200_Q_FP_HBD, false>(qr, v, oq, odq, vec, nz, err, ssz);
      } else if (a.q_kind == B200_Q_B) {
        if (we) small_quant<W, H, B200_Q_B, true>(qr, v, oq, odq, vec, nz, err, ssz);
        else small_quant<W, H, B200_Q_B, false>(qr, v, oq, odq, vec, nz, err, ssz);
      } else {
        if (we) small_quant<W, H, B200_Q_B_HBD, true>(qr, v, oq, odq, vec, nz, err, ssz);
        else small_quant<W, H, B200_Q_B_HBD, false>(qr, v, oq, odq, vec, nz, err, ssz);
      }
    }
    if constexpr (StageBuf<N>::kOn) {
      if (run_ok) {
        flush_staged_run<N>(sbuf0, a.qcoeff + ob0, lane);
        flush_staged_run<N>(sbuf1, a.dqcoeff + ob0, lane);
      } else {
        flush_staged<N>(sbuf0, a.qcoeff, ob, nact, vec, lane);
        flush_staged<N>(sbuf1, a.dqcoeff, ob, nact, vec, lane);
      }
    }
    if (!act) continue;
    // eob = smallest k with no non-zero position outside the first k scan entries
    const int cls = d.tx_type < 10 ? 0 : ((d.tx_type & 1) ? 2 : 1);
    const unsigned long long *pm = s_pm + cls * (N + 1);
    int lo = 0, hi = N;
#pragma unroll
    for (int s = 0; s <= cx_log2(N); ++s) {
      const int mid = (lo + hi) >> 1;
      const bool inside = (nz & ~pm[mid]) == 0;  // eob <= mid
      if (lo < hi) {
        if (inside) hi = mid;
        else lo = mid + 1;
      }
    }
    a.eob[bi] = (uint16_t)lo;
    if (we && a.dist) {
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

template <int W, int H, int IN>
TxqKernelInfo txq_small_kernel_info() {
  TxqKernelInfo k;
  k.fn = txq_small_kernel<W, H, IN>;
  k.smem_bytes = SmallSmem<W * H>::kCtaBytes;
  k.bpw = 32;
  return k;
}

}  // namespace b200
