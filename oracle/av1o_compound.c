/*
 * av1o_compound.c -- ORACLE (test infrastructure only): compound / masked / OBMC SAD and variance and
 * the compound predictor builders (the remaining aom_variance_fn_ptr_t
 * members: sdaf with dist-wtd, svaf, jsvaf, msdf, msvf, osdf, ovf, osvf).
 *
 * Restates aom_dsp/sad_av1.c (masked_sad :21, highbd_masked_sad :95, obmc_sad :163, highbd_obmc_sad
 * :213), aom_dsp/sad.c (aom_dist_wtd_sadWxH_avg_c :57, highbd :290), aom_dsp/variance.c
 * (SUBPIX_AVG_VAR :147 incl. the dist_wtd flavour :165, aom_comp_avg_pred_c :285,
 * aom_dist_wtd_comp_avg_pred_c :300, HIGHBD_SUBPIX_AVG_VAR :530-665, aom_highbd_comp_avg_pred_c :708,
 * aom_highbd_dist_wtd_comp_avg_pred_c :725, aom_comp_mask_pred_c :749, MASK_SUBPIX_VAR :769,
 * aom_highbd_comp_mask_pred_c :816, HIGHBD_MASK_SUBPIX_VAR :838, obmc_variance :933, OBMC_VAR :954,
 * OBMC_SUBPIX_VAR :962, highbd_obmc_variance64 :1040 .. HIGHBD_OBMC_SUBPIX_VAR :1126) and the blend
 * macros of aom_dsp/blend.h:23-40.  One evaluator: predictor pixel (optionally bilinear 1/8-pel),
 * optional blend with second_pred, then the measure.  tests/test_compound_oracle.py pins every entry
 * against the reference's per-size functions.
 */
#include "av1o.h"

#include <stdlib.h>

static const uint8_t kBil[8][2] = { { 128, 0 }, { 112, 16 }, { 96, 32 }, { 80, 48 },
                                    { 64, 64 }, { 48, 80 },  { 32, 96 }, { 16, 112 } };

static int px(const void *p, int highbd, long i) {
  return highbd ? ((const uint16_t *)p)[i] : ((const uint8_t *)p)[i];
}

/* aom_var_filter_block2d_bil_first_pass / second_pass (variance.c:64-116): the interpolated pixel */
static int bil(const void *a, int highbd, int stride, int x, int y, int xoff, int yoff) {
  const long o = (long)y * stride + x;
  const int h0 = (px(a, highbd, o) * kBil[xoff][0] + px(a, highbd, o + 1) * kBil[xoff][1] + 64) >> 7;
  const int h1 =
      (px(a, highbd, o + stride) * kBil[xoff][0] + px(a, highbd, o + stride + 1) * kBil[xoff][1] + 64) >> 7;
  const int v = (h0 * kBil[yoff][0] + h1 * kBil[yoff][1] + 64) >> 7;
  return highbd ? (uint16_t)v : (uint8_t)v;
}

/* the predictor sample of every compound flavour */
static int cmp_pred(const Av1oCompound *c, int x, int y) {
  int p = c->subpel ? bil(c->pre, c->highbd, c->pre_stride, x, y, c->xoff, c->yoff)
                    : px(c->pre, c->highbd, (long)y * c->pre_stride + x);
  if (c->blend == AV1O_BLEND_NONE) return p;
  const int s = px(c->second_pred, c->highbd, (long)y * c->w + x);
  if (c->blend == AV1O_BLEND_AVG) return (s + p + 1) >> 1;                    /* AOM_BLEND_AVG */
  if (c->blend == AV1O_BLEND_DIST_WTD) return (s * c->bck_offset + p * c->fwd_offset + 8) >> 4;
  const int m = c->mask[y * c->mask_stride + x];                              /* AOM_BLEND_A64 */
  return c->invert_mask ? (m * s + (64 - m) * p + 32) >> 6 : (m * p + (64 - m) * s + 32) >> 6;
}

void av1o_compound_pred(const Av1oCompound *c, void *comp_pred) {
  for (int y = 0; y < c->h; ++y)
    for (int x = 0; x < c->w; ++x) {
      const int v = cmp_pred(c, x, y);
      if (c->highbd) ((uint16_t *)comp_pred)[y * c->w + x] = (uint16_t)v;
      else ((uint8_t *)comp_pred)[y * c->w + x] = (uint8_t)v;
    }
}

unsigned av1o_compound_sad(const Av1oCompound *c) {
  unsigned sad = 0;
  for (int y = 0; y < c->h; ++y)
    for (int x = 0; x < c->w; ++x)
      sad += (unsigned)abs(cmp_pred(c, x, y) - px(c->tgt, c->highbd, (long)y * c->tgt_stride + x));
  return sad;
}

/* aom_varianceWxH / aom_highbd_{8,10,12}_varianceWxH of the predictor against tgt */
unsigned av1o_compound_variance(const Av1oCompound *c, unsigned *sse_out) {
  int64_t sum = 0;
  uint64_t sse = 0;
  for (int y = 0; y < c->h; ++y)
    for (int x = 0; x < c->w; ++x) {
      const int d = cmp_pred(c, x, y) - px(c->tgt, c->highbd, (long)y * c->tgt_stride + x);
      sum += d;
      sse += (uint32_t)(d * d);
    }
  const int n = c->w * c->h;
  if (c->highbd && c->bd > 8) {
    const int s2 = c->bd - 8, s1 = 2 * s2;
    const unsigned sse32 = (unsigned)((sse + ((uint64_t)1 << (s1 - 1))) >> s1);
    const int isum = (int)((sum + ((int64_t)1 << (s2 - 1))) >> s2);
    const int64_t var = (int64_t)sse32 - ((int64_t)isum * isum) / n;
    *sse_out = sse32;
    return var >= 0 ? (unsigned)var : 0;
  }
  *sse_out = (unsigned)sse;
  return (unsigned)sse - (unsigned)(((int64_t)(int)sum * (int)sum) / n);
}

static int rpo2_signed(int v, int n) {
  return v < 0 ? -((-v + (1 << (n - 1))) >> n) : (v + (1 << (n - 1))) >> n;
}

unsigned av1o_obmc_sad(const Av1oCompound *c) {
  unsigned sad = 0;
  for (int y = 0; y < c->h; ++y)
    for (int x = 0; x < c->w; ++x) {
      const int i = y * c->w + x;
      sad += (unsigned)((abs(c->wsrc[i] - cmp_pred(c, x, y) * c->obmc_mask[i]) + 2048) >> 12);
    }
  return sad;
}

unsigned av1o_obmc_variance(const Av1oCompound *c, unsigned *sse_out) {
  int64_t sum64 = 0;
  uint64_t sse64 = 0;
  unsigned sse_lo = 0; /* the 8-bit function accumulates in unsigned int / int */
  int sum_lo = 0;
  for (int y = 0; y < c->h; ++y)
    for (int x = 0; x < c->w; ++x) {
      const int i = y * c->w + x;
      const int d = rpo2_signed(c->wsrc[i] - cmp_pred(c, x, y) * c->obmc_mask[i], 12);
      sum64 += d;
      sse64 += (uint64_t)(int64_t)(d * d);
      sum_lo += d;
      sse_lo += (unsigned)(d * d);
    }
  const int n = c->w * c->h;
  unsigned sse;
  int sum;
  if (!c->highbd) {
    sse = sse_lo, sum = sum_lo;
  } else if (c->bd == 8) {
    sse = (unsigned)sse64, sum = (int)sum64;
  } else {
    const int s2 = c->bd - 8, s1 = 2 * s2;
    sum = (int)((sum64 + ((int64_t)1 << (s2 - 1))) >> s2);
    sse = (unsigned)((sse64 + ((uint64_t)1 << (s1 - 1))) >> s1);
  }
  *sse_out = sse;
  if (c->highbd && c->bd > 8) {
    const int64_t var = (int64_t)sse - ((int64_t)sum * sum) / n;
    return var >= 0 ? (unsigned)var : 0;
  }
  return sse - (unsigned)(((int64_t)sum * sum) / n);
}
