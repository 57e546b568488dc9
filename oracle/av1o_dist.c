/*
 * av1o_dist.c -- ORACLE (test infrastructure only): transform-domain distortion,
 * residual and pixel SSE.  Restates av1/encoder/rdopt.c:635-682,
 * av1/encoder/tx_search.c:1049-1075, aom_dsp/subtract.c:20-50, aom_dsp/sse.c:19-53.
 */
#include "av1o.h"

static int32_t mul32(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }

/* av1_block_error_c: the squares are formed in 32-bit int (wrapping) before they
 * are widened -- rdopt.c:642-643 */
int64_t av1o_block_error(const int32_t *coeff, const int32_t *dqcoeff, intptr_t n,
                         int64_t *ssz) {
  int64_t err = 0, sq = 0;
  for (intptr_t i = 0; i < n; ++i) {
    const int32_t d = (int32_t)((uint32_t)coeff[i] - (uint32_t)dqcoeff[i]);
    err += mul32(d, d);
    sq += mul32(coeff[i], coeff[i]);
  }
  *ssz = sq;
  return err;
}

/* av1_block_error_lp_c rdopt.c:650 */
int64_t av1o_block_error_lp(const int16_t *coeff, const int16_t *dqcoeff, intptr_t n) {
  int64_t err = 0;
  for (intptr_t i = 0; i < n; ++i) {
    const int d = coeff[i] - dqcoeff[i];
    err += mul32(d, d);
  }
  return err;
}

/* av1_highbd_block_error_c rdopt.c:663: 64-bit squares, rounded down-shift 2(bd-8) */
int64_t av1o_highbd_block_error(const int32_t *coeff, const int32_t *dqcoeff,
                                intptr_t n, int64_t *ssz, int bd) {
  int64_t err = 0, sq = 0;
  const int shift = 2 * (bd - 8);
  const int64_t rnd = shift > 0 ? (int64_t)1 << (shift - 1) : 0;
  for (intptr_t i = 0; i < n; ++i) {
    const int64_t d = (int32_t)((uint32_t)coeff[i] - (uint32_t)dqcoeff[i]);
    err += d * d;
    sq += (int64_t)coeff[i] * coeff[i];
  }
  *ssz = (sq + rnd) >> shift;
  return (err + rnd) >> shift;
}

/* av1_block_error_qm tx_search.c:1049 (fork addition).  Note the reference indexes
 * the weight as qmatrix[scan[i]] while coeff is indexed by i; kept as is. */
int64_t av1o_block_error_qm(const int32_t *coeff, const int32_t *dqcoeff, intptr_t n,
                            const uint8_t *qm, const int16_t *scan, int64_t *ssz) {
  int64_t err = 0, sq = 0;
  for (intptr_t i = 0; i < n; ++i) {
    const int64_t w = qm[scan[i]];
    const int64_t dd = (int64_t)(int32_t)((uint32_t)coeff[i] - (uint32_t)dqcoeff[i]) * w;
    const int64_t cc = (int64_t)coeff[i] * w;
    err += (dd * dd + (1 << 9)) >> 10;
    sq += (cc * cc + (1 << 9)) >> 10;
  }
  *ssz = sq;
  return err;
}

void av1o_subtract_block(int rows, int cols, int16_t *diff, ptrdiff_t diff_stride,
                         const uint8_t *src, ptrdiff_t src_stride, const uint8_t *pred,
                         ptrdiff_t pred_stride) {
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c)
      diff[r * diff_stride + c] = (int16_t)(src[r * src_stride + c] - pred[r * pred_stride + c]);
}

void av1o_highbd_subtract_block(int rows, int cols, int16_t *diff, ptrdiff_t diff_stride,
                                const uint16_t *src, ptrdiff_t src_stride,
                                const uint16_t *pred, ptrdiff_t pred_stride) {
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c)
      diff[r * diff_stride + c] = (int16_t)(src[r * src_stride + c] - pred[r * pred_stride + c]);
}

int64_t av1o_sse(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int w,
                 int h) {
  int64_t sse = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int d = a[y * a_stride + x] - b[y * b_stride + x];
      sse += d * d;
    }
  return sse;
}

int64_t av1o_highbd_sse(const uint16_t *a, int a_stride, const uint16_t *b, int b_stride,
                        int w, int h) {
  int64_t sse = 0;
  for (int y = 0; y < h; ++y)
    for (int x = 0; x < w; ++x) {
      const int32_t d = (int32_t)a[y * a_stride + x] - (int32_t)b[y * b_stride + x];
      sse += mul32(d, d);
    }
  return sse;
}
