// itxfm1d.cuh -- AV1 inverse 1-D integer transforms as fully unrolled register code.
//
// Product code (sm_100a).  Behavioural contract: bit-exact with the reference's
// av1_idct{4,8,16,32,64}, av1_iadst{4,8,16}, av1_iidentity{4,8,16,32}_c
// (/root/reference/av1/common/av1_inv_txfm1d.c:16,59,142,308,1062,656,713,821,1029-1060) incl.
// clamp_value after every add/sub stage (av1_inv_txfm1d.h:21).
//
// Design: the inverse networks are not written down again.  Every forward network of
// txfm1d.cuh is a product of 2x2 rotations and add/sub butterflies; the inverse the codec
// defines is that same flow graph traversed backwards with each 2x2 block transposed (the
// butterflies and most rotations are symmetric) and a clamp to the pass's stage range after
// every add/sub.  So the generators below mirror dct_core / dct_odd_half / adst_level in reverse
// order, still with compile-time indices and constant-bank weights only.
//
// Arithmetic width: inputs of every rotation are either pass inputs (clamped to the stage range
// r by the caller, av1_inv_txfm2d.c:277,286,304) or clamped add/sub results, so |x| <= 2^(r-1),
// r <= 20: a single product |w x| < 2^31 never wraps, and the sum of the two products fits int32
// whenever r <= 18 (5793 * 2^17 < 2^31).  Only the 12-bit row pass (r = 20) needs the 64-bit sum
// of half_btf (Btf<12, true>); everything else runs the 3-instruction 32-bit form.
#pragma once
#include "txfm1d.cuh"

namespace b200 {

__device__ __forceinline__ int32_t clamp_range(int32_t v, int32_t lo, int32_t hi) {
  return min(max(v, lo), hi);
}

// ------------------------------------------------------------------ DCT --------
template <int M, int P, class B>
__device__ __forceinline__ void idct_odd_levels(int32_t *u, int32_t lo, int32_t hi) {
  if constexpr (P <= M) {
    constexpr int G = P / 2;
    static_for<0, M / G>([&](auto gc) {
      constexpr int g = decltype(gc)::value;
      static_for<0, G / 2>([&](auto ic) {
        constexpr int i = decltype(ic)::value;
        constexpr int ia = g * G + i, ib = g * G + G - 1 - i;
        const int32_t a = u[ia], b = u[ib];
        if constexpr ((g & 1) == 0) {
          u[ia] = clamp_range(a + b, lo, hi);
          u[ib] = clamp_range(a - b, lo, hi);
        } else {
          u[ia] = clamp_range(b - a, lo, hi);
          u[ib] = clamp_range(b + a, lo, hi);
        }
      });
    });
    static_for<0, M / 2>([&](auto jc) {
      constexpr int j = decltype(jc)::value;
      constexpr int t = j % P, k = j / P;
      constexpr int n = (P == M) ? 1 : M / (2 * P);
      constexpr int A = (P == M) ? 32 : 16 / n + (64 / n) * cx_bitrev(k, cx_log2(n));
      if constexpr (t >= P / 4 && t < P / 2) {
        const int32_t l = u[j], h = u[M - 1 - j];
        u[j] = B::template rot<-1, A, +1, 64 - A>(l, h);
        u[M - 1 - j] = B::template rot<+1, 64 - A, +1, A>(l, h);
      } else if constexpr (t >= P / 2 && t < 3 * P / 4) {
        const int32_t l = u[j], h = u[M - 1 - j];
        u[j] = B::template rot<-1, 64 - A, -1, A>(l, h);
        u[M - 1 - j] = B::template rot<-1, A, +1, 64 - A>(l, h);
      }
    });
    idct_odd_levels<M, P * 2, B>(u, lo, hi);
  }
}

template <int M, class B>
__device__ __forceinline__ void idct_odd_half(int32_t *u, int32_t lo, int32_t hi) {
  static_for<0, M / 2>([&](auto jc) {
    constexpr int j = decltype(jc)::value;
    constexpr int Y = 32 / M + (128 / M) * cx_bitrev(j, cx_log2(M) - 1);
    const int32_t l = u[j], h = u[M - 1 - j];
    u[j] = B::template rot<+1, 64 - Y, -1, Y>(l, h);
    u[M - 1 - j] = B::template rot<+1, Y, +1, 64 - Y>(l, h);
  });
  idct_odd_levels<M, 4, B>(u, lo, hi);
}

// in place: slot i holds coefficient bitrev(i) on entry and sample i on exit
template <int N, class B>
__device__ __forceinline__ void idct_core(int32_t *v, int32_t lo, int32_t hi) {
  if constexpr (N == 2) {
    const int32_t a = v[0], b = v[1];
    v[0] = B::template rot<+1, 32, +1, 32>(a, b);
    v[1] = B::template rot<+1, 32, -1, 32>(a, b);
  } else {
    constexpr int M = N / 2;
    idct_core<M, B>(v, lo, hi);
    idct_odd_half<M, B>(v + M, lo, hi);
    static_for<0, M>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      const int32_t e = v[i], d = v[N - 1 - i];
      v[i] = clamp_range(e + d, lo, hi);
      v[N - 1 - i] = clamp_range(e - d, lo, hi);
    });
  }
}

// N-point inverse DCT of x[0..NZ) (coefficients NZ..N-1 are zero: the 64-point sizes keep 32)
template <int N, class B, int NZ = N>
__device__ __forceinline__ void idct(const int32_t *x, int32_t *y, int32_t lo, int32_t hi) {
  static_for<0, N>([&](auto ic) {
    constexpr int i = decltype(ic)::value;
    constexpr int k = cx_bitrev(i, cx_log2(N));
    if constexpr (k < NZ) y[i] = x[k];
    else y[i] = 0;
  });
  idct_core<N, B>(y, lo, hi);
}

// ------------------------------------------------------------------ ADST -------
// av1_iadst4 (av1_inv_txfm1d.c:656): wrapping 32-bit products and sums, one rounding
template <class B>
__device__ __forceinline__ void iadst4(const int32_t *x, int32_t *y) {
  const int32_t s1 = c_sinpi[B::kRow][1], s2 = c_sinpi[B::kRow][2];
  const int32_t s3 = c_sinpi[B::kRow][3], s4 = c_sinpi[B::kRow][4];
  const int32_t x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
  const int32_t t0 = wadd(wadd(wmul(s1, x0), wmul(s4, x2)), wmul(s2, x3));
  const int32_t t1 = wsub(wsub(wmul(s2, x0), wmul(s1, x2)), wmul(s4, x3));
  const int32_t t3 = wmul(s3, x1);
  const int32_t t2 = wmul(s3, wadd(wsub(x0, x2), x3));
  y[0] = rshift_round_rt(wadd(t0, t3), B::kBit);
  y[1] = rshift_round_rt(wadd(t1, t3), B::kBit);
  y[2] = rshift_round_rt(t2, B::kBit);
  y[3] = rshift_round_rt(wsub(wadd(t0, t1), t3), B::kBit);
}

template <int N, int G, class B>
__device__ __forceinline__ void iadst_level(int32_t *v, int32_t lo, int32_t hi) {
  if constexpr (G >= 4) {
    constexpr int pairs = G / 4;
    static_for<0, N / G>([&](auto bc) {
      constexpr int base = decltype(bc)::value * G;
      static_for<0, G / 2>([&](auto ic) {
        constexpr int i = base + decltype(ic)::value;
        const int32_t a = v[i], b = v[i + G / 2];
        v[i] = clamp_range(a + b, lo, hi);
        v[i + G / 2] = clamp_range(a - b, lo, hi);
      });
      static_for<0, pairs>([&](auto qc) {
        constexpr int q = decltype(qc)::value;
        constexpr int i0 = base + G / 2 + 2 * q;
        const int32_t a = v[i0], b = v[i0 + 1];
        if constexpr (pairs == 1) {
          v[i0] = B::template rot<+1, 32, +1, 32>(a, b);
          v[i0 + 1] = B::template rot<+1, 32, -1, 32>(a, b);
        } else {
          constexpr int half = pairs / 2, qi = q % half;
          constexpr int A = 16 / half + (64 / half) * qi;
          if constexpr (q < half) {
            v[i0] = B::template rot<+1, A, +1, 64 - A>(a, b);
            v[i0 + 1] = B::template rot<+1, 64 - A, -1, A>(a, b);
          } else {
            v[i0] = B::template rot<-1, 64 - A, +1, A>(a, b);
            v[i0 + 1] = B::template rot<+1, A, +1, 64 - A>(a, b);
          }
        }
      });
    });
    iadst_level<N, G / 2, B>(v, lo, hi);
  }
}

template <int N, class B>
__device__ __forceinline__ void iadst(const int32_t *x, int32_t *y, int32_t lo, int32_t hi) {
  if constexpr (N == 4) {
    iadst4<B>(x, y);
  } else {
    int32_t v[N];
    static_for<0, N / 2>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      v[2 * i + 1] = x[2 * i];
      v[N - 2 - 2 * i] = x[2 * i + 1];
    });
    static_for<0, N / 2>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      constexpr int A = 32 / N + (128 / N) * i;
      const int32_t a = v[2 * i], b = v[2 * i + 1];
      v[2 * i] = B::template rot<+1, A, +1, 64 - A>(a, b);
      v[2 * i + 1] = B::template rot<+1, 64 - A, -1, A>(a, b);
    });
    iadst_level<N, N, B>(v, lo, hi);
    static_for<0, N>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      y[AdstPerm<N>::idx(i)] = AdstPerm<N>::neg(i) ? (int32_t)(0u - (uint32_t)v[i]) : v[i];
    });
  }
}

// ---------------------------------------------------------------- identity -----
template <int N>
__device__ __forceinline__ void iidentity(const int32_t *x, int32_t *y) {
  static_for<0, N>([&](auto ic) {
    constexpr int i = decltype(ic)::value;
    if constexpr (N == 4) y[i] = scale_sqrt2<1>(x[i]);
    else if constexpr (N == 8) y[i] = (int32_t)((uint32_t)x[i] * 2u);
    else if constexpr (N == 16) y[i] = scale_sqrt2<2>(x[i]);
    else y[i] = (int32_t)((uint32_t)x[i] * 4u);
  });
}

// kind: 0 DCT, 1 / 2 ADST (the flips are applied by the caller), 3 identity
template <int N, bool WIDE, int NZ>
__device__ __forceinline__ void itxfm1d(int kind, const int32_t *x, int32_t *y, int32_t lo, int32_t hi) {
  using B = Btf<12, WIDE>;  // INV_COS_BIT (av1_txfm.h)
  if constexpr (N == 64) {
    idct<64, B, NZ>(x, y, lo, hi);
  } else if constexpr (N == 32) {
    if (kind == 3) iidentity<32>(x, y);
    else idct<32, B, NZ>(x, y, lo, hi);
  } else {
    if (kind == 0) idct<N, B, NZ>(x, y, lo, hi);
    else if (kind == 3) iidentity<N>(x, y);
    else iadst<N, B>(x, y, lo, hi);
  }
}

}  // namespace b200
