/*
 * av1o_sadvar.c -- ORACLE (test infrastructure only): SAD, variance, sub-pel
 * variance and the sub-pel predictor used by motion search.
 * Restates aom_dsp/sad.c:22-140,240-330, aom_dsp/variance.c:38-145,321-530,
 * av1/encoder/reconinter_enc.c:424-496,562-640 (unscaled branch) with
 * aom_dsp/aom_convolve.c:22-110.  Size-generic (w, h arguments) instead of the
 * reference's per-size macro instantiations.
 */
#include "av1o.h"

#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------ SAD ------ */

unsigned av1o_sad(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride,
                  int w, int h) {
  unsigned acc = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      acc += (unsigned)abs(src[y * src_stride + x] - ref[y * ref_stride + x]);
  return acc;
}

/* aom_sad_skip_WxH_c sad.c:65: even rows only, doubled */
unsigned av1o_sad_skip(const uint8_t *src, int src_stride, const uint8_t *ref,
                       int ref_stride, int w, int h) {
  return 2 * av1o_sad(src, 2 * src_stride, ref, 2 * ref_stride, w, h / 2);
}

/* aom_sadWxH_avg_c sad.c:50 with aom_comp_avg_pred_c variance.c:285 */
unsigned av1o_sad_avg(const uint8_t *src, int src_stride, const uint8_t *ref,
                      int ref_stride, const uint8_t *second_pred, int w, int h) {
  unsigned acc = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int p = (second_pred[y * w + x] + ref[y * ref_stride + x] + 1) >> 1;
      acc += (unsigned)abs(src[y * src_stride + x] - p);
    }
  return acc;
}

/* aom_sadWxHx4d_c :95 / x3d :125 (the C x3d calls x4d, i.e. it also fills [3]) /
 * aom_sad_skip_WxHx4d_c :113 */
void av1o_sad_x4d(const uint8_t *src, int src_stride, const uint8_t *const ref[4],
                  int ref_stride, uint32_t sad[4], int w, int h, int n_refs, int skip) {
  for (int i = 0; i < n_refs; ++i)
    sad[i] = skip ? av1o_sad_skip(src, src_stride, ref[i], ref_stride, w, h)
                  : av1o_sad(src, src_stride, ref[i], ref_stride, w, h);
}

unsigned av1o_highbd_sad(const uint16_t *src, int src_stride, const uint16_t *ref,
                         int ref_stride, int w, int h) {
  unsigned acc = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      acc += (unsigned)abs(src[y * src_stride + x] - ref[y * ref_stride + x]);
  return acc;
}

unsigned av1o_highbd_sad_skip(const uint16_t *src, int src_stride, const uint16_t *ref,
                              int ref_stride, int w, int h) {
  return 2 * av1o_highbd_sad(src, 2 * src_stride, ref, 2 * ref_stride, w, h / 2);
}

/* -------------------------------------------------------------- variance ----- */

unsigned av1o_variance(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride,
                       int w, int h, unsigned *sse) {
  int sum = 0;
  uint32_t sq = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int d = a[y * a_stride + x] - b[y * b_stride + x];
      sum += d;
      sq += (uint32_t)(d * d);
    }
  *sse = sq;
  return sq - (uint32_t)(((int64_t)sum * sum) / (w * h));
}

unsigned av1o_mse(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int w,
                  int h, unsigned *sse) {
  av1o_variance(a, a_stride, b, b_stride, w, h, sse);
  return *sse;
}

static const uint8_t kBil[8][2] = { { 128, 0 }, { 112, 16 }, { 96, 32 }, { 80, 48 },
                                    { 64, 64 }, { 48, 80 },  { 32, 96 }, { 16, 112 } };

/* SUBPIX_VAR variance.c:132: horizontal 2-tap on (h+1) rows, then vertical 2-tap */
unsigned av1o_sub_pixel_variance(const uint8_t *a, int a_stride, int xoff, int yoff,
                                 const uint8_t *b, int b_stride, int w, int h,
                                 unsigned *sse) {
  uint16_t *mid = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)((h + 1) * w));
  uint8_t *pred = (uint8_t *)malloc((size_t)(h * w));
  for (int y = 0; y < h + 1; ++y)
    for (int x = 0; x < w; ++x)
      mid[y * w + x] = (uint16_t)((a[y * a_stride + x] * kBil[xoff][0] +
                                   a[y * a_stride + x + 1] * kBil[xoff][1] + 64) >> 7);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      pred[y * w + x] = (uint8_t)((mid[y * w + x] * kBil[yoff][0] +
                                   mid[(y + 1) * w + x] * kBil[yoff][1] + 64) >> 7);
  const unsigned v = av1o_variance(pred, w, b, b_stride, w, h, sse);
  free(mid);
  free(pred);
  return v;
}

/* highbd_variance64 :321, highbd_{8,10,12}_variance :351-379, HIGHBD_VAR :381-408 */
unsigned av1o_highbd_variance(const uint16_t *a, int a_stride, const uint16_t *b,
                              int b_stride, int w, int h, int bd, unsigned *sse) {
  int64_t sum = 0;
  uint64_t sq = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int d = a[y * a_stride + x] - b[y * b_stride + x];
      sum += d;
      sq += (uint32_t)(d * d);
    }
  const int s2 = bd - 8, s1 = 2 * s2; /* sum shift, sse shift */
  if (s2 > 0) {
    sq = (sq + ((uint64_t)1 << (s1 - 1))) >> s1;
    sum = (sum + ((int64_t)1 << (s2 - 1))) >> s2;
  }
  *sse = (uint32_t)sq;
  const int isum = (int)sum;
  if (bd == 8) return *sse - (uint32_t)(((int64_t)isum * isum) / (w * h));
  const int64_t var = (int64_t)(*sse) - (((int64_t)isum * isum) / (w * h));
  return var >= 0 ? (uint32_t)var : 0;
}

unsigned av1o_highbd_sub_pixel_variance(const uint16_t *a, int a_stride, int xoff,
                                        int yoff, const uint16_t *b, int b_stride, int w,
                                        int h, int bd, unsigned *sse) {
  uint16_t *mid = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)((h + 1) * w));
  uint16_t *pred = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(h * w));
  for (int y = 0; y < h + 1; ++y)
    for (int x = 0; x < w; ++x)
      mid[y * w + x] = (uint16_t)((a[y * a_stride + x] * kBil[xoff][0] +
                                   a[y * a_stride + x + 1] * kBil[xoff][1] + 64) >> 7);
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x)
      pred[y * w + x] = (uint16_t)((mid[y * w + x] * kBil[yoff][0] +
                                    mid[(y + 1) * w + x] * kBil[yoff][1] + 64) >> 7);
  const unsigned v = av1o_highbd_variance(pred, w, b, b_stride, w, h, bd, sse);
  free(mid);
  free(pred);
  return v;
}

/* ------------------------------------------------------- sub-pel predictor --- */

/* Even rows (subpel_q3 << 1) of av1_bilinear_filters, av1_sub_pel_filters_4 and
 * av1_sub_pel_filters_8 (av1/common/filter.h:110-135,210-222): AV1 interpolation
 * kernels at 1/8-pel positions, taps -3..+4. */
static const int16_t kTaps2[8][8] = {
  { 0, 0, 0, 128, 0, 0, 0, 0 },  { 0, 0, 0, 112, 16, 0, 0, 0 }, { 0, 0, 0, 96, 32, 0, 0, 0 },
  { 0, 0, 0, 80, 48, 0, 0, 0 },  { 0, 0, 0, 64, 64, 0, 0, 0 },  { 0, 0, 0, 48, 80, 0, 0, 0 },
  { 0, 0, 0, 32, 96, 0, 0, 0 },  { 0, 0, 0, 16, 112, 0, 0, 0 },
};
static const int16_t kTaps4[8][8] = {
  { 0, 0, 0, 128, 0, 0, 0, 0 },     { 0, 0, -8, 122, 18, -4, 0, 0 },
  { 0, 0, -12, 110, 38, -8, 0, 0 }, { 0, 0, -14, 94, 58, -10, 0, 0 },
  { 0, 0, -12, 76, 76, -12, 0, 0 }, { 0, 0, -10, 58, 94, -14, 0, 0 },
  { 0, 0, -8, 38, 110, -12, 0, 0 }, { 0, 0, -4, 18, 122, -8, 0, 0 },
};
static const int16_t kTaps8[8][8] = {
  { 0, 0, 0, 128, 0, 0, 0, 0 },      { 0, 2, -10, 122, 18, -4, 0, 0 },
  { 0, 2, -14, 110, 38, -10, 2, 0 }, { 0, 2, -16, 94, 58, -12, 2, 0 },
  { 0, 2, -14, 76, 76, -14, 2, 0 },  { 0, 2, -12, 58, 94, -16, 2, 0 },
  { 0, 2, -10, 38, 110, -14, 2, 0 }, { 0, 0, -4, 18, 122, -10, 2, 0 },
};

static const int16_t *taps_for(int taps, int q3) {
  return taps == 2 ? kTaps2[q3] : (taps == 4 ? kTaps4[q3] : kTaps8[q3]);
}

static int clip_px(int v, int maxv) { return v < 0 ? 0 : (v > maxv ? maxv : v); }

/* generic over pixel type through a small macro; the intermediate between the
 * horizontal and the vertical pass is clipped to the pixel range (the reference
 * goes through aom_convolve8_horiz_c into a pixel-typed temp). */
#define UPSAMPLED_BODY(PIX, MAXV)                                                        \
  const int16_t *kx = taps_for(taps, subpel_x_q3), *ky = taps_for(taps, subpel_y_q3);    \
  if (!subpel_x_q3 && !subpel_y_q3) {                                                    \
    for (int y = 0; y < height; ++y)                                                     \
      memcpy(comp_pred + y * width, ref + y * ref_stride, sizeof(PIX) * (size_t)width);  \
    return;                                                                              \
  }                                                                                      \
  if (!subpel_y_q3) {                                                                    \
    for (int y = 0; y < height; ++y)                                                     \
      for (int x = 0; x < width; ++x) {                                                  \
        int s = 0;                                                                       \
        for (int k = 0; k < 8; ++k) s += ref[y * ref_stride + x - 3 + k] * kx[k];        \
        comp_pred[y * width + x] = (PIX)clip_px((s + 64) >> 7, MAXV);                    \
      }                                                                                  \
    return;                                                                              \
  }                                                                                      \
  if (!subpel_x_q3) {                                                                    \
    for (int y = 0; y < height; ++y)                                                     \
      for (int x = 0; x < width; ++x) {                                                  \
        int s = 0;                                                                       \
        for (int k = 0; k < 8; ++k) s += ref[(y - 3 + k) * ref_stride + x] * ky[k];      \
        comp_pred[y * width + x] = (PIX)clip_px((s + 64) >> 7, MAXV);                    \
      }                                                                                  \
    return;                                                                              \
  }                                                                                      \
  {                                                                                      \
    const int ih = height + 7;                                                           \
    PIX *tmp = (PIX *)malloc(sizeof(PIX) * (size_t)(ih * width));                        \
    for (int y = 0; y < ih; ++y)                                                         \
      for (int x = 0; x < width; ++x) {                                                  \
        int s = 0;                                                                       \
        for (int k = 0; k < 8; ++k) s += ref[(y - 3) * ref_stride + x - 3 + k] * kx[k];  \
        tmp[y * width + x] = (PIX)clip_px((s + 64) >> 7, MAXV);                          \
      }                                                                                  \
    for (int y = 0; y < height; ++y)                                                     \
      for (int x = 0; x < width; ++x) {                                                  \
        int s = 0;                                                                       \
        for (int k = 0; k < 8; ++k) s += tmp[(y + k) * width + x] * ky[k];               \
        comp_pred[y * width + x] = (PIX)clip_px((s + 64) >> 7, MAXV);                    \
      }                                                                                  \
    free(tmp);                                                                           \
  }

void av1o_upsampled_pred(uint8_t *comp_pred, int width, int height, int subpel_x_q3,
                         int subpel_y_q3, const uint8_t *ref, int ref_stride, int taps) {
  UPSAMPLED_BODY(uint8_t, 255)
}

void av1o_highbd_upsampled_pred(uint16_t *comp_pred, int width, int height,
                                int subpel_x_q3, int subpel_y_q3, const uint16_t *ref,
                                int ref_stride, int bd, int taps) {
  UPSAMPLED_BODY(uint16_t, (1 << bd) - 1)
}
