/*
 * av1o_avg.c -- CPU ORACLE (test infrastructure only, see av1o.h): Hadamard / SATD /
 * block averages / integer projections used by tx-search pruning, TPL and the
 * variance-based partitioner.  Restates aom_dsp/avg.c (line numbers below).
 */
#include <stdlib.h>
#include <string.h>

#include "av1o.h"

/* hadamard_col4 (avg.c:102): 4-point butterfly with a halving first stage, int16 stores */
static void col4(const int16_t *s, ptrdiff_t st, int16_t *o) {
  const int16_t b0 = (int16_t)((s[0] + s[st]) >> 1), b1 = (int16_t)((s[0] - s[st]) >> 1);
  const int16_t b2 = (int16_t)((s[2 * st] + s[3 * st]) >> 1), b3 = (int16_t)((s[2 * st] - s[3 * st]) >> 1);
  o[0] = (int16_t)(b0 + b2);
  o[1] = (int16_t)(b1 + b3);
  o[2] = (int16_t)(b0 - b2);
  o[3] = (int16_t)(b1 - b3);
}

/* aom_hadamard_4x4_c (avg.c:115) */
void av1o_hadamard_4x4(const int16_t *src, ptrdiff_t stride, int32_t *coeff) {
  int16_t t[16], u[16];
  for (int c = 0; c < 4; ++c) col4(src + c, stride, t + 4 * c);
  for (int c = 0; c < 4; ++c) col4(t + c, 4, u + 4 * c);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) coeff[i * 4 + j] = u[j * 4 + i];
}

/* the 8-point butterfly network of hadamard_col8 (avg.c:147) / the highbd passes (:350,:383)
 * on 32-bit values; the caller narrows to int16 where the reference stores int16 */
static void net8(const int32_t *x, int32_t *o) {
  const int32_t b0 = x[0] + x[1], b1 = x[0] - x[1], b2 = x[2] + x[3], b3 = x[2] - x[3];
  const int32_t b4 = x[4] + x[5], b5 = x[4] - x[5], b6 = x[6] + x[7], b7 = x[6] - x[7];
  const int32_t c0 = b0 + b2, c1 = b1 + b3, c2 = b0 - b2, c3 = b1 - b3;
  const int32_t c4 = b4 + b6, c5 = b5 + b7, c6 = b4 - b6, c7 = b5 - b7;
  o[0] = c0 + c4; o[7] = c1 + c5; o[3] = c2 + c6; o[4] = c3 + c7;
  o[2] = c0 - c4; o[6] = c1 - c5; o[1] = c2 - c6; o[5] = c3 - c7;
}
/* int16 flavour: every intermediate is an int16_t variable in the reference */
static void net8_i16(const int32_t *x, int32_t *o) {
#define W(v) ((int32_t)(int16_t)(v))
  const int32_t b0 = W(x[0] + x[1]), b1 = W(x[0] - x[1]), b2 = W(x[2] + x[3]), b3 = W(x[2] - x[3]);
  const int32_t b4 = W(x[4] + x[5]), b5 = W(x[4] - x[5]), b6 = W(x[6] + x[7]), b7 = W(x[6] - x[7]);
  const int32_t c0 = W(b0 + b2), c1 = W(b1 + b3), c2 = W(b0 - b2), c3 = W(b1 - b3);
  const int32_t c4 = W(b4 + b6), c5 = W(b5 + b7), c6 = W(b4 - b6), c7 = W(b5 - b7);
  o[0] = W(c0 + c4); o[7] = W(c1 + c5); o[3] = W(c2 + c6); o[4] = W(c3 + c7);
  o[2] = W(c0 - c4); o[6] = W(c1 - c5); o[1] = W(c2 - c6); o[5] = W(c3 - c7);
#undef W
}

/* 8x8 core.  kind 0: aom_hadamard_8x8_c (:177) and 1: aom_hadamard_lp_8x8_c (:207) --
 * both passes int16, output transposed; kind 2: aom_highbd_hadamard_8x8_c (:416) -- first
 * pass int16, second pass int32, no transpose. */
static void had8(const int16_t *src, ptrdiff_t stride, int32_t *out, int kind) {
  int32_t t[64], u[64], x[8];
  for (int c = 0; c < 8; ++c) {
    for (int r = 0; r < 8; ++r) x[r] = src[r * stride + c];
    net8_i16(x, t + 8 * c);
  }
  for (int c = 0; c < 8; ++c) {
    for (int r = 0; r < 8; ++r) x[r] = t[r * 8 + c];
    if (kind == 2) net8(x, u + 8 * c);
    else net8_i16(x, u + 8 * c);
  }
  if (kind == 2) {
    memcpy(out, u, sizeof(u));
  } else {
    for (int i = 0; i < 8; ++i)
      for (int j = 0; j < 8; ++j) out[i * 8 + j] = u[j * 8 + i];
  }
}

/* combine four quarter transforms (avg.c:258-276, :298-315, :329-347, :455-472, :485-503) */
static void combine(int32_t *c, int q, int shift, int lp) {
  for (int i = 0; i < q; ++i) {
    const int32_t a0 = c[i], a1 = c[i + q], a2 = c[i + 2 * q], a3 = c[i + 3 * q];
    int32_t b0 = (a0 + a1) >> shift, b1 = (a0 - a1) >> shift;
    int32_t b2 = (a2 + a3) >> shift, b3 = (a2 - a3) >> shift;
    if (lp) { b0 = (int16_t)b0; b1 = (int16_t)b1; b2 = (int16_t)b2; b3 = (int16_t)b3; }
    c[i] = b0 + b2; c[i + q] = b1 + b3; c[i + 2 * q] = b0 - b2; c[i + 3 * q] = b1 - b3;
    if (lp) {
      c[i] = (int16_t)c[i]; c[i + q] = (int16_t)c[i + q];
      c[i + 2 * q] = (int16_t)c[i + 2 * q]; c[i + 3 * q] = (int16_t)c[i + 3 * q];
    }
  }
}

static void had16(const int16_t *src, ptrdiff_t stride, int32_t *out, int kind) {
  for (int i = 0; i < 4; ++i) had8(src + (i >> 1) * 8 * stride + (i & 1) * 8, stride, out + 64 * i, kind);
  combine(out, 64, 1, kind == 1);
  if (kind == 0) { /* the swap that matches the AVX2 output order (avg.c:279-286) */
    for (int i = 0; i < 16; ++i)
      for (int j = 0; j < 4; ++j) {
        const int32_t t = out[i * 16 + 4 + j];
        out[i * 16 + 4 + j] = out[i * 16 + 8 + j];
        out[i * 16 + 8 + j] = t;
      }
  }
}

/* Hadamard of an n x n residual block, n in {4 (kind 0 only), 8, 16, 32 (kinds 0, 2)}.
 * kind 0: aom_hadamard_NxN_c; 1: aom_hadamard_lp_NxN_c (values are int16, returned
 * widened); 2: aom_highbd_hadamard_NxN_c. */
void av1o_hadamard(const int16_t *src, ptrdiff_t stride, int32_t *coeff, int n, int kind) {
  if (n == 4) {
    av1o_hadamard_4x4(src, stride, coeff);
  } else if (n == 8) {
    had8(src, stride, coeff, kind);
  } else if (n == 16) {
    had16(src, stride, coeff, kind);
  } else {
    for (int i = 0; i < 4; ++i)
      had16(src + (i >> 1) * 16 * stride + (i & 1) * 16, stride, coeff + 256 * i, kind);
    combine(coeff, 256, 2, 0);
  }
}

/* aom_satd_c (avg.c:509) / aom_satd_lp_c (:518) */
int av1o_satd(const int32_t *coeff, int length) {
  int s = 0;
  for (int i = 0; i < length; ++i) s += abs(coeff[i]);
  return s;
}

/* aom_avg_4x4_c / aom_avg_8x8_c (avg.c:32,42) and the highbd forms (:63,:74); n = 4 or 8 */
unsigned av1o_avg(const void *s, int stride, int n, int highbd) {
  int sum = 0;
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j)
      sum += highbd ? ((const uint16_t *)s)[i * stride + j] : ((const uint8_t *)s)[i * stride + j];
  return (unsigned)(sum + n * n / 2) >> (n == 4 ? 4 : 6);
}

/* aom_minmax_8x8_c (avg.c:18) / aom_highbd_minmax_8x8_c (:85) */
void av1o_minmax_8x8(const void *s, int p, const void *d, int dp, int highbd, int *mn, int *mx) {
  *mn = 255;
  *mx = 0;
  for (int i = 0; i < 8; ++i)
    for (int j = 0; j < 8; ++j) {
      const int a = highbd ? ((const uint16_t *)s)[i * p + j] : ((const uint8_t *)s)[i * p + j];
      const int b = highbd ? ((const uint16_t *)d)[i * dp + j] : ((const uint8_t *)d)[i * dp + j];
      const int df = abs(a - b);
      if (df < *mn) *mn = df;
      if (df > *mx) *mx = df;
    }
}

/* aom_int_pro_row_c (avg.c:528): column sums >> norm_factor, int16 accumulator */
void av1o_int_pro_row(int16_t *hbuf, const uint8_t *ref, int ref_stride, int width, int height,
                      int norm_factor) {
  for (int x = 0; x < width; ++x) {
    int16_t acc = 0;
    for (int y = 0; y < height; ++y) acc = (int16_t)(acc + ref[y * ref_stride + x]);
    hbuf[x] = (int16_t)(acc >> norm_factor);
  }
}

/* aom_int_pro_col_c (avg.c:542): row sums >> norm_factor */
void av1o_int_pro_col(int16_t *vbuf, const uint8_t *ref, int ref_stride, int width, int height,
                      int norm_factor) {
  for (int y = 0; y < height; ++y) {
    int16_t acc = 0;
    for (int x = 0; x < width; ++x) acc = (int16_t)(acc + ref[y * ref_stride + x]);
    vbuf[y] = (int16_t)(acc >> norm_factor);
  }
}

/* aom_vector_var_c (avg.c:556) */
int av1o_vector_var(const int16_t *ref, const int16_t *src, int bwl) {
  const int width = 4 << bwl;
  int sse = 0, mean = 0;
  for (int i = 0; i < width; ++i) {
    const int d = ref[i] - src[i];
    mean += d;
    sse += d * d;
  }
  const unsigned m = (unsigned)abs(mean);
  return sse - (int)((m * m) >> (bwl + 2));
}
