/*
 * av1o_itxfm.c -- ORACLE (test infrastructure only): inverse transforms, reconstruction and
 * pixel-domain distortion.
 *
 * Restates av1/common/av1_inv_txfm1d.c (av1_idct4 :16, av1_idct8 :59, av1_idct16 :142,
 * av1_idct32 :308, av1_iadst4 :656, av1_iadst8 :713, av1_iadst16 :821, av1_iidentity* :1029-1060,
 * av1_idct64 :1062), av1/common/av1_inv_txfm2d.c (inv_txfm2d_add_c :234-326, the 64-point
 * wrappers :385-460, av1_highbd_iwht4x4_16_add_c / _1_add_c :20-109), av1/common/idct.c
 * (av1_inv_txfm_add_c :283, av1_highbd_inv_txfm_add_4x4_c's lossless switch) and
 * dist_block_px_domain / pixel_dist_visible_only (av1/encoder/tx_search.c:920-1022).
 *
 * The reference spells every stage out; here each inverse network is the forward network of
 * av1o_txfm.c run backwards: the forward 2x2 rotations and add/sub butterflies are visited in
 * reverse order with transposed weights, and every add/sub result is clamped to the pass's stage
 * range (clamp_value, av1_inv_txfm1d.h:21) as the reference does.  tests/test_itx_oracle.py pins
 * that against the reference's own functions for all sizes, types and bit depths.
 */
#include "av1o.h"

#include <stdlib.h>
#include <string.h>

static const uint8_t kW[AV1O_TX_SIZES_ALL] = { 4, 8, 16, 32, 64, 4, 8, 8, 16, 16,
                                                32, 32, 64, 4, 16, 8, 32, 16, 64 };
static const uint8_t kH[AV1O_TX_SIZES_ALL] = { 4, 8, 16, 32, 64, 8, 4, 16, 8, 32,
                                                16, 64, 32, 16, 4, 32, 8, 64, 16 };
/* av1_inv_txfm_shift_ls[tx_size][0] (av1_inv_txfm2d.c:133-159), as a down-shift; [1] is 4 */
static const uint8_t kRowShift[AV1O_TX_SIZES_ALL] = { 0, 1, 2, 2, 2, 0, 0, 1, 1, 1,
                                                       1, 1, 1, 1, 1, 2, 2, 2, 2 };
static const uint8_t kVtx[16] = { 0, 1, 0, 1, 2, 0, 2, 1, 2, 3, 0, 3, 1, 3, 2, 3 };
static const uint8_t kHtx[16] = { 0, 0, 1, 1, 0, 2, 2, 2, 1, 3, 3, 0, 3, 1, 3, 2 };
#define INV_COS_BIT 12

static int32_t rshift_round(int64_t v, int bit) {
  return (int32_t)((v + ((int64_t)1 << (bit - 1))) >> bit);
}
/* half_btf, av1_txfm.h:80-102 */
static int32_t hbtf(int32_t w0, int32_t x0, int32_t w1, int32_t x1) {
  const int32_t p0 = (int32_t)((uint32_t)w0 * (uint32_t)x0);
  const int32_t p1 = (int32_t)((uint32_t)w1 * (uint32_t)x1);
  return (int32_t)(((int64_t)p0 + (int64_t)p1 + ((int64_t)1 << (INV_COS_BIT - 1))) >> INV_COS_BIT);
}
/* clamp_value, av1_inv_txfm1d.h:21 */
static int32_t clampb(int64_t v, int bit) {
  if (bit <= 0) return (int32_t)v;
  const int64_t hi = ((int64_t)1 << (bit - 1)) - 1, lo = -((int64_t)1 << (bit - 1));
  return (int32_t)(v < lo ? lo : (v > hi ? hi : v));
}
static int32_t add_c(int32_t a, int32_t b, int bit) { return clampb((int64_t)a + b, bit); }
static int32_t sub_c(int32_t a, int32_t b, int bit) { return clampb((int64_t)a - b, bit); }
static unsigned bitrev(unsigned v, int bits) {
  unsigned r = 0;
  for (int i = 0; i < bits; ++i) r |= ((v >> i) & 1u) << (bits - 1 - i);
  return r;
}
static int ilog2(int n) {
  int l = 0;
  while ((1 << l) < n) ++l;
  return l;
}

/* Inverse of the forward odd half (av1o_txfm.c dct_odd_half): last rotation level first
 * (transposed), then for p = 4 .. M the add/sub level followed by its rotation level (both
 * symmetric, so their own transposes). */
static void idct_odd_half(int32_t *u, int M, const int32_t *c, int rb) {
  const int lb = ilog2(M) - 1;
  for (int j = 0; j < M / 2; ++j) {
    const int Y = 32 / M + (128 / M) * (int)bitrev((unsigned)j, lb);
    const int32_t lo = u[j], hi = u[M - 1 - j];
    u[j] = hbtf(c[64 - Y], lo, -c[Y], hi);
    u[M - 1 - j] = hbtf(c[Y], lo, c[64 - Y], hi);
  }
  for (int p = 4; p <= M; p <<= 1) {
    const int G = p / 2;
    for (int g = 0; g < M / G; ++g) {
      int32_t *v = u + g * G;
      for (int i = 0; i < G / 2; ++i) {
        const int32_t a = v[i], b = v[G - 1 - i];
        if ((g & 1) == 0) {
          v[i] = add_c(a, b, rb);
          v[G - 1 - i] = sub_c(a, b, rb);
        } else {
          v[i] = sub_c(b, a, rb);
          v[G - 1 - i] = add_c(b, a, rb);
        }
      }
    }
    for (int j = 0; j < M / 2; ++j) {
      const int t = j % p, k = j / p;
      int A = 32;
      if (p != M) {
        const int n = M / (2 * p);
        A = 16 / n + (64 / n) * (int)bitrev((unsigned)k, ilog2(n));
      }
      const int32_t lo = u[j], hi = u[M - 1 - j];
      if (t >= p / 4 && t < p / 2) {
        u[j] = hbtf(-c[A], lo, c[64 - A], hi);
        u[M - 1 - j] = hbtf(c[64 - A], lo, c[A], hi);
      } else if (t >= p / 2 && t < 3 * p / 4) {
        u[j] = hbtf(-c[64 - A], lo, -c[A], hi);
        u[M - 1 - j] = hbtf(-c[A], lo, c[64 - A], hi);
      }
    }
  }
}

/* slot i holds coefficient bitrev(i) on entry, sample i on exit */
static void idct_core(int32_t *v, int n, const int32_t *c, int rb) {
  if (n == 2) {
    const int32_t a = v[0], b = v[1];
    v[0] = hbtf(c[32], a, c[32], b);
    v[1] = hbtf(c[32], a, -c[32], b);
    return;
  }
  const int m = n / 2;
  idct_core(v, m, c, rb);
  idct_odd_half(v + m, m, c, rb);
  int32_t tmp[64];
  for (int i = 0; i < m; ++i) {
    const int32_t e = v[i], d = v[m + (m - 1 - i)];
    tmp[i] = add_c(e, d, rb);
    tmp[n - 1 - i] = sub_c(e, d, rb);
  }
  memcpy(v, tmp, sizeof(int32_t) * (size_t)n);
}

static void idct(const int32_t *in, int32_t *out, int n, int rb) {
  int32_t v[64];
  const int lb = ilog2(n);
  for (int i = 0; i < n; ++i) v[i] = in[bitrev((unsigned)i, lb)];
  idct_core(v, n, av1o_cospi(INV_COS_BIT), rb);
  memcpy(out, v, sizeof(int32_t) * (size_t)n);
}

/* av1_iadst4, av1_inv_txfm1d.c:656-711 (wrapping 32-bit products and sums, one rounding) */
static void iadst4(const int32_t *in, int32_t *out) {
  const int32_t *s = av1o_sinpi(INV_COS_BIT);
  const uint32_t x0 = (uint32_t)in[0], x1 = (uint32_t)in[1], x2 = (uint32_t)in[2], x3 = (uint32_t)in[3];
  if (!(x0 | x1 | x2 | x3)) {
    out[0] = out[1] = out[2] = out[3] = 0;
    return;
  }
  const uint32_t s1 = (uint32_t)s[1], s2 = (uint32_t)s[2], s3 = (uint32_t)s[3], s4 = (uint32_t)s[4];
  const uint32_t t0 = s1 * x0 + s4 * x2 + s2 * x3;
  const uint32_t t1 = s2 * x0 - s1 * x2 - s4 * x3;
  const uint32_t t3 = s3 * x1;
  const uint32_t t2 = s3 * (x0 - x2 + x3);
  out[0] = rshift_round((int32_t)(t0 + t3), INV_COS_BIT);
  out[1] = rshift_round((int32_t)(t1 + t3), INV_COS_BIT);
  out[2] = rshift_round((int32_t)t2, INV_COS_BIT);
  out[3] = rshift_round((int32_t)(t0 + t1 - t3), INV_COS_BIT);
}

/* av1_iadst8 :713 / av1_iadst16 :821 = fadst_n of av1o_txfm.c backwards */
static void iadst_n(const int32_t *in, int32_t *out, int n, int rb) {
  static const int8_t perm8[8] = { 0, -7, -3, 4, -1, 6, 2, -5 };
  static const int8_t perm16[16] = { 0,  -15, -7, 8,  -3, 12,  4,  -11,
                                     -1, 14,  6,  -9, 2,  -13, -5, 10 };
  const int8_t *perm = n == 8 ? perm8 : perm16;
  const int32_t *c = av1o_cospi(INV_COS_BIT);
  int32_t v[16];
  for (int i = 0; i < n / 2; ++i) {
    v[2 * i + 1] = in[2 * i];
    v[n - 2 - 2 * i] = in[2 * i + 1];
  }
  for (int i = 0; i < n / 2; ++i) {
    const int A = 32 / n + (128 / n) * i;
    const int32_t a = v[2 * i], b = v[2 * i + 1];
    v[2 * i] = hbtf(c[A], a, c[64 - A], b);
    v[2 * i + 1] = hbtf(c[64 - A], a, -c[A], b);
  }
  for (int g = n; g >= 4; g >>= 1) {
    const int pairs = g / 4;
    for (int base = 0; base < n; base += g) {
      for (int i = 0; i < g / 2; ++i) {
        const int32_t a = v[base + i], b = v[base + g / 2 + i];
        v[base + i] = add_c(a, b, rb);
        v[base + g / 2 + i] = sub_c(a, b, rb);
      }
      for (int q = 0; q < pairs; ++q) {
        int32_t *x = v + base + g / 2 + 2 * q;
        const int32_t a = x[0], b = x[1];
        if (pairs == 1) {
          x[0] = hbtf(c[32], a, c[32], b);
          x[1] = hbtf(c[32], a, -c[32], b);
        } else {
          const int half = pairs / 2, qi = q % half;
          const int A = 16 / half + (64 / half) * qi;
          if (q < half) {
            x[0] = hbtf(c[A], a, c[64 - A], b);
            x[1] = hbtf(c[64 - A], a, -c[A], b);
          } else {
            x[0] = hbtf(-c[64 - A], a, c[A], b);
            x[1] = hbtf(c[A], a, c[64 - A], b);
          }
        }
      }
    }
  }
  for (int i = 0; i < n; ++i) {
    if (perm[i] < 0) out[-perm[i]] = (int32_t)(0u - (uint32_t)v[i]);
    else out[perm[i]] = v[i];
  }
}

static void iidentity(const int32_t *in, int32_t *out, int n) {
  for (int i = 0; i < n; ++i) {
    switch (n) {
      case 4: out[i] = rshift_round((int64_t)5793 * in[i], 12); break;
      case 8: out[i] = (int32_t)((int64_t)in[i] * 2); break;
      case 16: out[i] = rshift_round((int64_t)5793 * 2 * in[i], 12); break;
      default: out[i] = (int32_t)((int64_t)in[i] * 4); break;
    }
  }
}

/* kind: 0 DCT, 1 ADST, 2 identity; range_bits = the pass's stage range (av1_gen_inv_stage_range,
 * av1_inv_txfm2d.c:189-232: rows 16/18/20, columns 16/16/18 for bd 8/10/12) */
void av1o_inv_txfm1d(int kind, int n, const int32_t *in, int32_t *out, int range_bits) {
  if (kind == 0) idct(in, out, n, range_bits);
  else if (kind == 1) {
    if (n == 4) iadst4(in, out);
    else iadst_n(in, out, n, range_bits);
  } else iidentity(in, out, n);
}

static uint16_t clip_add(uint16_t d, int32_t t, int bd) {
  const int v = (int)d + (int)t, hi = (1 << bd) - 1;
  return (uint16_t)(v < 0 ? 0 : (v > hi ? hi : v));
}

/* av1_inv_txfm2d_add_WxH_c: input = the av1o_max_eob(tx_size) kept coefficients, column-major with
 * pitch min(H, 32) as the forward transform writes them; dest += residual, clipped to bd bits. */
void av1o_inv_txfm2d_add(const int32_t *input, uint16_t *dest, int stride, int tx_size, int tx_type,
                         int bd) {
  const int W = kW[tx_size], H = kH[tx_size];
  const int Wk = W > 32 ? 32 : W, Hk = H > 32 ? 32 : H;
  const int vt = kVtx[tx_type], ht = kHtx[tx_type];
  const int ud = vt == 2, lr = ht == 2;
  const int vkind = vt == 0 ? 0 : (vt == 3 ? 2 : 1), hkind = ht == 0 ? 0 : (ht == 3 ? 2 : 1);
  const int rect = (W == 2 * H) || (H == 2 * W);
  const int row_range = bd == 8 ? 16 : (bd == 10 ? 18 : 20);
  const int col_range = bd == 12 ? 18 : 16;
  const int col_clamp = bd + 6 > 16 ? bd + 6 : 16;
  int32_t *buf = (int32_t *)calloc((size_t)(W * H), sizeof(int32_t));
  int32_t tin[64], tout[64];
  for (int r = 0; r < H; ++r) {
    for (int c = 0; c < W; ++c) {
      int32_t v = (c < Wk && r < Hk) ? input[c * Hk + r] : 0;
      if (rect) v = rshift_round((int64_t)v * 2896, 12);
      tin[c] = clampb(v, bd + 8);
    }
    av1o_inv_txfm1d(hkind, W, tin, tout, row_range);
    for (int c = 0; c < W; ++c)
      buf[r * W + c] = kRowShift[tx_size] ? rshift_round(tout[c], kRowShift[tx_size]) : tout[c];
  }
  for (int c = 0; c < W; ++c) {
    for (int r = 0; r < H; ++r) tin[r] = clampb(buf[r * W + (lr ? W - 1 - c : c)], col_clamp);
    av1o_inv_txfm1d(vkind, H, tin, tout, col_range);
    for (int r = 0; r < H; ++r) {
      const int32_t t = rshift_round(tout[ud ? H - 1 - r : r], 4);
      dest[r * stride + c] = clip_add(dest[r * stride + c], t, bd);
    }
  }
  free(buf);
}

/* av1_highbd_iwht4x4_16_add_c / av1_highbd_iwht4x4_1_add_c (av1_inv_txfm2d.c:20-109) behind the
 * eob switch of av1_highbd_inv_txfm_add_4x4_c (idct.c): the lossless inverse */
void av1o_iwht4x4_add(const int32_t *input, uint16_t *dest, int stride, int eob, int bd) {
  if (eob > 1) {
    int32_t o[16];
    for (int i = 0; i < 4; ++i) {
      int32_t a = input[i] >> 2, c = input[4 + i] >> 2, d = input[8 + i] >> 2, b = input[12 + i] >> 2, e;
      a += c, d -= b, e = (a - d) >> 1, b = e - b, c = e - c, a -= b, d += c;
      o[i] = a, o[4 + i] = b, o[8 + i] = c, o[12 + i] = d;
    }
    for (int i = 0; i < 4; ++i) {
      int32_t a = o[4 * i], c = o[4 * i + 1], d = o[4 * i + 2], b = o[4 * i + 3], e;
      a += c, d -= b, e = (a - d) >> 1, b = e - b, c = e - c, a -= b, d += c;
      dest[i] = clip_add(dest[i], a, bd);
      dest[stride + i] = clip_add(dest[stride + i], b, bd);
      dest[2 * stride + i] = clip_add(dest[2 * stride + i], c, bd);
      dest[3 * stride + i] = clip_add(dest[3 * stride + i], d, bd);
    }
  } else {
    int32_t a = input[0] >> 2, e = a >> 1, t[4];
    a -= e;
    t[0] = a, t[1] = t[2] = t[3] = e;
    for (int i = 0; i < 4; ++i) {
      const int32_t e1 = t[i] >> 1, a1 = t[i] - e1;
      dest[i] = clip_add(dest[i], a1, bd);
      dest[stride + i] = clip_add(dest[stride + i], e1, bd);
      dest[2 * stride + i] = clip_add(dest[2 * stride + i], e1, bd);
      dest[3 * stride + i] = clip_add(dest[3 * stride + i], e1, bd);
    }
  }
}

/* av1_inverse_transform_block (idct.c:303) for one block over pixel buffers of uint8 (highbd == 0,
 * av1_inv_txfm_add_c idct.c:283: widened, transformed at bd, narrowed) or uint16 samples:
 * recon = pred + inverse(dqcoeff), or pred itself when eob == 0. */
void av1o_recon_block(const int32_t *dqcoeff, const void *pred, int pred_stride, void *recon,
                      int recon_stride, int tx_size, int tx_type, int eob, int lossless, int highbd,
                      int bd) {
  const int W = kW[tx_size], H = kH[tx_size];
  uint16_t tmp[64 * 64];
  for (int r = 0; r < H; ++r)
    for (int c = 0; c < W; ++c)
      tmp[r * 64 + c] = highbd ? ((const uint16_t *)pred)[r * pred_stride + c]
                               : ((const uint8_t *)pred)[r * pred_stride + c];
  if (eob) {
    if (lossless) av1o_iwht4x4_add(dqcoeff, tmp, 64, eob, bd);
    else av1o_inv_txfm2d_add(dqcoeff, tmp, 64, tx_size, tx_type, bd);
  }
  for (int r = 0; r < H; ++r)
    for (int c = 0; c < W; ++c) {
      if (highbd) ((uint16_t *)recon)[r * recon_stride + c] = tmp[r * 64 + c];
      else ((uint8_t *)recon)[r * recon_stride + c] = (uint8_t)tmp[r * 64 + c];
    }
}

/* dist_block_px_domain (tx_search.c:969-1022): 16 x the SSE between source and reconstruction over
 * the visible part of the block (pixel_dist_visible_only :920; the vf / sse_odd_size variants agree:
 * plain sum for 8 bit, ROUND_POWER_OF_TWO(sum, 2 (bd - 8)) as unsigned for high bit depth). */
int64_t av1o_dist_block_px(const void *src, int src_stride, const void *recon, int recon_stride,
                           int vis_w, int vis_h, int highbd, int bd) {
  uint64_t sse = 0;
  for (int r = 0; r < vis_h; ++r)
    for (int c = 0; c < vis_w; ++c) {
      const int d = highbd ? (int)((const uint16_t *)src)[r * src_stride + c] -
                                 (int)((const uint16_t *)recon)[r * recon_stride + c]
                           : (int)((const uint8_t *)src)[r * src_stride + c] -
                                 (int)((const uint8_t *)recon)[r * recon_stride + c];
      sse += (uint64_t)(d * d);
    }
  if (highbd && bd > 8) sse = (sse + ((uint64_t)1 << (2 * (bd - 8) - 1))) >> (2 * (bd - 8));
  return 16 * (int64_t)(unsigned)sse;
}

/* av1_cost_coeffs_txb_estimate (av1/encoder/txb_rdopt.c:547-582; costLUT / const_term / loge_par
 * txb_rdopt_utils.h:31-37): the Laplacian rate estimate prune_txk_type / prune_txk_type_separ use
 * (tx_search.c:1179,1215,1302).  scan = av1o_scan(tx_size, tx_type); eob >= 1. */
int av1o_cost_coeffs_txb_estimate(const int32_t *qcoeff, int eob, int tx_size, int tx_type) {
  static const int lut[15] = { -1143, 53, 545, 825, 1031, 1209, 1393, 1577, 1763, 1947, 2132, 2317, 2501, 2686, 2871 };
  const int16_t *scan = av1o_scan(tx_size, tx_type);
  const int const_term = 1 << 9, loge_par = ((14427 << 9) + 5000) / 10000;
  int c = eob - 1;
  int cost = (abs(qcoeff[scan[c]]) - 1) << (9 + 2);
  for (c = eob - 2; c >= 0; --c) {
    const int v = abs(qcoeff[scan[c]]);
    cost += lut[v < 14 ? v : 14];
  }
  return cost + (const_term + loge_par) * (eob - 1);
}
