// txfm1d.cuh -- AV1 forward 1-D integer transforms as fully unrolled register code.
//
// Product code (sm_100a).  Behavioural contract: bit-exact with the reference's
// av1_fdct{4,8,16,32,64}, av1_fadst{4,8,16}, av1_fidentity{4,8,16,32}_c
// (/root/reference/av1/encoder/av1_fwd_txfm1d.c:16,59,144,315,1096,676,735,849,
// 1064-1094) and half_btf/round_shift (av1/common/av1_txfm.h:75-102).
//
// Design: one thread owns one whole N-point transform in registers.  The butterfly
// networks are not spelled out stage by stage; they are generated at compile time
// from the recursive structure of the AV1 DCT (even half = N/2-point DCT of the
// mirrored sums, odd half = alternating rotation / add-sub levels whose angles follow
// a bit-reversal rule) with constexpr index arithmetic, so every array index and
// every cosine is a compile-time constant: values live in registers and the weights
// are constant-bank operands of IMAD.
#pragma once
#include <stdint.h>

#include <type_traits>

namespace b200 {

// cos(j*pi/128) * 2^bit for bit = 10..13 and the negated values (so that a negative
// weight is also a plain constant-bank operand), plus the ADST4 sine table.
// (one private copy per translation unit; each TU uploads its own, see txq_inst.cuh)
static __constant__ int32_t c_cospi[4][64];
static __constant__ int32_t c_ncospi[4][64];
static __constant__ int32_t c_sinpi[4][5];

__host__ __device__ constexpr int cx_log2(int n) { return n <= 1 ? 0 : 1 + cx_log2(n >> 1); }
__host__ __device__ constexpr int cx_bitrev(int v, int bits) {
  int r = 0;
  for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1) << (bits - 1 - i);
  return r;
}

// compile-time loop: f(std::integral_constant<int, I>) for I in [Begin, End)
template <int Begin, int End, class F>
__device__ __forceinline__ void static_for(F &&f) {
  if constexpr (Begin < End) {
    f(std::integral_constant<int, Begin>{});
    static_for<Begin + 1, End>(f);
  }
}

__device__ __forceinline__ int32_t wadd(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a + (uint32_t)b);
}
__device__ __forceinline__ int32_t wsub(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a - (uint32_t)b);
}
__device__ __forceinline__ int32_t wmul(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a * (uint32_t)b);
}

// Arithmetic policy.  EXACT reproduces half_btf literally: two wrapping 32-bit
// products, 64-bit sum, round, shift (av1_txfm.h:82-101).  !EXACT is the all-32-bit
// form which the reference documents as equivalent for conformant input (same comment
// block) and which its own SSE4/AVX2 kernels use.
template <int BIT, bool EXACT>
struct Btf {
  static constexpr int kBit = BIT;
  static constexpr int kRow = BIT - 10;
  // S0*cos[A0]*x0 + S1*cos[A1]*x1, rounded and shifted
  template <int S0, int A0, int S1, int A1>
  static __device__ __forceinline__ int32_t rot(int32_t x0, int32_t x1) {
    const int32_t w0 = S0 > 0 ? c_cospi[kRow][A0] : c_ncospi[kRow][A0];
    const int32_t w1 = S1 > 0 ? c_cospi[kRow][A1] : c_ncospi[kRow][A1];
    if constexpr (EXACT) {
      const int64_t s = (int64_t)wmul(w0, x0) + (int64_t)wmul(w1, x1) + (int64_t)(1 << (BIT - 1));
      return (int32_t)(s >> BIT);
    } else {
      const uint32_t s = (uint32_t)w0 * (uint32_t)x0 + (uint32_t)w1 * (uint32_t)x1 +
                         (uint32_t)(1 << (BIT - 1));
      return (int32_t)s >> BIT;
    }
  }
};

// round_shift(int64 value, bit) of av1_txfm.h:75 for an int32 value, without a
// 64-bit add: floor((x + 2^(k-1)) / 2^k) == ((x >> (k-1)) + 1) >> 1 for k >= 2.
template <int K>
__device__ __forceinline__ int32_t rshift_round(int32_t x) {
  if constexpr (K == 0) return x;
  else if constexpr (K == 1) return (x >> 1) + (x & 1);
  else return ((x >> (K - 1)) + 1) >> 1;
}
__device__ __forceinline__ int32_t rshift_round_rt(int32_t x, int k) {  // k >= 1
  return (int32_t)(((int64_t)x + ((int64_t)1 << (k - 1))) >> k);
}

// round_shift((int64)x * 5793 * MUL, 12): sqrt(2) scalings of identity4/16 and of the
// 2:1 rectangular transforms (av1_fwd_txfm1d.c:1068,1084; av1_fwd_txfm2d.c:115-122)
template <int MUL>
__device__ __forceinline__ int32_t scale_sqrt2(int32_t x) {
  const int64_t v = (int64_t)x * (int64_t)(5793 * MUL) + (int64_t)(1 << 11);
  return (int32_t)(v >> 12);
}

// ------------------------------------------------------------------ DCT --------
// One rotation level of the odd half (period P) followed by its add/sub level.
template <int M, int P, class B>
__device__ __forceinline__ void dct_odd_level(int32_t *u) {
  if constexpr (P >= 4) {
    static_for<0, M / 2>([&](auto jc) {
      constexpr int j = decltype(jc)::value;
      constexpr int t = j % P, k = j / P;
      constexpr int n = (P == M) ? 1 : M / (2 * P);
      constexpr int A = (P == M) ? 32 : 16 / n + (64 / n) * cx_bitrev(k, cx_log2(n));
      if constexpr (t >= P / 4 && t < P / 2) {
        const int32_t lo = u[j], hi = u[M - 1 - j];
        u[j] = B::template rot<-1, A, +1, 64 - A>(lo, hi);
        u[M - 1 - j] = B::template rot<+1, A, +1, 64 - A>(hi, lo);
      } else if constexpr (t >= P / 2 && t < 3 * P / 4) {
        const int32_t lo = u[j], hi = u[M - 1 - j];
        u[j] = B::template rot<-1, 64 - A, -1, A>(lo, hi);
        u[M - 1 - j] = B::template rot<+1, 64 - A, -1, A>(hi, lo);
      }
    });
    constexpr int G = P / 2;
    static_for<0, M / G>([&](auto gc) {
      constexpr int g = decltype(gc)::value;
      static_for<0, G / 2>([&](auto ic) {
        constexpr int i = decltype(ic)::value;
        constexpr int ia = g * G + i, ib = g * G + G - 1 - i;
        const int32_t a = u[ia], b = u[ib];
        if constexpr ((g & 1) == 0) {
          u[ia] = wadd(a, b);
          u[ib] = wsub(a, b);
        } else {
          u[ia] = wsub(b, a);
          u[ib] = wadd(b, a);
        }
      });
    });
    dct_odd_level<M, P / 2, B>(u);
  }
}

// Odd half acting on the M differences u[j] = x[M-1-j] - x[M+j].
template <int M, class B>
__device__ __forceinline__ void dct_odd_half(int32_t *u) {
  dct_odd_level<M, M, B>(u);
  static_for<0, M / 2>([&](auto jc) {
    constexpr int j = decltype(jc)::value;
    constexpr int Y = 32 / M + (128 / M) * cx_bitrev(j, cx_log2(M) - 1);
    const int32_t lo = u[j], hi = u[M - 1 - j];
    u[j] = B::template rot<+1, 64 - Y, +1, Y>(lo, hi);
    u[M - 1 - j] = B::template rot<+1, 64 - Y, -1, Y>(hi, lo);
  });
}

// in place; slot i ends up holding output coefficient bitrev(i)
template <int N, class B>
__device__ __forceinline__ void dct_core(int32_t *v) {
  if constexpr (N == 2) {
    const int32_t a = v[0], b = v[1];
    v[0] = B::template rot<+1, 32, +1, 32>(a, b);
    v[1] = B::template rot<-1, 32, +1, 32>(b, a);
  } else {
    constexpr int M = N / 2;
    int32_t d[M];
    static_for<0, M>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      const int32_t a = v[i], b = v[N - 1 - i];
      v[i] = wadd(a, b);
      d[M - 1 - i] = wsub(a, b);  // d[j] = x[M-1-j] - x[M+j]
    });
    static_for<0, M>([&](auto ic) { v[M + decltype(ic)::value] = d[decltype(ic)::value]; });
    dct_core<M, B>(v);
    dct_odd_half<M, B>(v + M);
  }
}

// N-point forward DCT.  x is overwritten; y[k] receives coefficient k for k < KEEP
// (KEEP = 32 for the 64-point transforms whose upper half the codec discards).
template <int N, class B, int KEEP = N>
__device__ __forceinline__ void fdct(int32_t *x, int32_t *y) {
  dct_core<N, B>(x);
  static_for<0, N>([&](auto ic) {
    constexpr int i = decltype(ic)::value;
    constexpr int k = cx_bitrev(i, cx_log2(N));
    if constexpr (k < KEEP) y[k] = x[i];
  });
}

// ------------------------------------------------------------------ ADST -------
// av1_fadst4 (av1_fwd_txfm1d.c:676): sinpi products, one rounding at the end.  The
// reference's all-zero early-out produces the same zeros as the general path.
template <class B>
__device__ __forceinline__ void fadst4(const int32_t *x, int32_t *y) {
  const int32_t s1 = c_sinpi[B::kRow][1], s2 = c_sinpi[B::kRow][2];
  const int32_t s3 = c_sinpi[B::kRow][3], s4 = c_sinpi[B::kRow][4];
  const int32_t x0 = x[0], x1 = x[1], x2 = x[2], x3 = x[3];
  const int32_t a7 = wsub(wadd(x0, x1), x3);
  const int32_t y0 = wadd(wadd(wmul(s1, x0), wmul(s2, x1)), wmul(s4, x3));
  const int32_t y1 = wmul(s3, a7);
  const int32_t y2 = wadd(wsub(wmul(s4, x0), wmul(s1, x1)), wmul(s2, x3));
  const int32_t y3 = wmul(s3, x2);
  y[0] = rshift_round_rt(wadd(y0, y3), B::kBit);
  y[1] = rshift_round_rt(y1, B::kBit);
  y[2] = rshift_round_rt(wsub(y2, y3), B::kBit);
  y[3] = rshift_round_rt(wadd(wsub(y2, y0), y3), B::kBit);
}

// av1_fadst8 (:735) / av1_fadst16 (:849).
template <int N>
struct AdstPerm;
template <>
struct AdstPerm<8> {
  // source index and sign of stage-1 slot i
  static constexpr int idx(int i) { constexpr int t[8] = { 0, 7, 3, 4, 1, 6, 2, 5 }; return t[i]; }
  static constexpr int neg(int i) { constexpr int t[8] = { 0, 1, 1, 0, 1, 0, 0, 1 }; return t[i]; }
};
template <>
struct AdstPerm<16> {
  static constexpr int idx(int i) {
    constexpr int t[16] = { 0, 15, 7, 8, 3, 12, 4, 11, 1, 14, 6, 9, 2, 13, 5, 10 };
    return t[i];
  }
  static constexpr int neg(int i) {
    constexpr int t[16] = { 0, 1, 1, 0, 1, 0, 0, 1, 1, 0, 0, 1, 0, 1, 1, 0 };
    return t[i];
  }
};

template <int N, int G, class B>
__device__ __forceinline__ void adst_level(int32_t *v) {
  if constexpr (G <= N) {
    constexpr int pairs = G / 4;
    static_for<0, N / G>([&](auto bc) {
      constexpr int base = decltype(bc)::value * G;
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
      static_for<0, G / 2>([&](auto ic) {
        constexpr int i = base + decltype(ic)::value;
        const int32_t a = v[i], b = v[i + G / 2];
        v[i] = wadd(a, b);
        v[i + G / 2] = wsub(a, b);
      });
    });
    adst_level<N, G * 2, B>(v);
  }
}

template <int N, class B>
__device__ __forceinline__ void fadst(const int32_t *x, int32_t *y) {
  if constexpr (N == 4) {
    fadst4<B>(x, y);
  } else {
    int32_t v[N];
    static_for<0, N>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      const int32_t s = x[AdstPerm<N>::idx(i)];
      v[i] = AdstPerm<N>::neg(i) ? (int32_t)(0u - (uint32_t)s) : s;
    });
    adst_level<N, 4, B>(v);
    static_for<0, N / 2>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      constexpr int A = 32 / N + (128 / N) * i;
      const int32_t a = v[2 * i], b = v[2 * i + 1];
      v[2 * i] = B::template rot<+1, A, +1, 64 - A>(a, b);
      v[2 * i + 1] = B::template rot<+1, 64 - A, -1, A>(a, b);
    });
    static_for<0, N / 2>([&](auto ic) {
      constexpr int i = decltype(ic)::value;
      y[2 * i] = v[2 * i + 1];
      y[2 * i + 1] = v[N - 2 - 2 * i];
    });
  }
}

// ---------------------------------------------------------------- identity -----
template <int N>
__device__ __forceinline__ void fidentity(const int32_t *x, int32_t *y) {
  static_for<0, N>([&](auto ic) {
    constexpr int i = decltype(ic)::value;
    if constexpr (N == 4) y[i] = scale_sqrt2<1>(x[i]);
    else if constexpr (N == 8) y[i] = (int32_t)((uint32_t)x[i] * 2u);
    else if constexpr (N == 16) y[i] = scale_sqrt2<2>(x[i]);
    else y[i] = (int32_t)((uint32_t)x[i] * 4u);
  });
}

}  // namespace b200
