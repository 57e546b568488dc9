/*
 * av1o_scan.c -- ORACLE (test infrastructure only): coefficient scan orders.
 *
 * The reference ships av1_scan_orders[TX_SIZES_ALL][TX_TYPES] as ~12.9k lines of
 * literal tables (av1/common/scan.c:17-1664, table of pointers :1666).  They follow
 * three rules, generated here instead (coefficients are stored column-major,
 * position = c * H + r):
 *   "mcol"    (tx types H_DCT, H_ADST, H_FLIPADST): storage order, 0,1,2,...
 *   "mrow"    (tx types V_DCT, V_ADST, V_FLIPADST): row by row, c*H + r for r, c
 *   "default" (all other types): anti-diagonals d = r + c in increasing d; tall
 *             blocks (H > W) walk each diagonal with r increasing, wide blocks with
 *             c increasing, square blocks alternate (odd d: r increasing).
 * 64-point dimensions use the 32-point table of the kept region (scan.c:1743).
 * tests/test_oracle_vs_ref.py compares all 19 x 16 tables with the reference's.
 */
#include "av1o.h"

#include <stdlib.h>

static int16_t *g_scan[AV1O_TX_SIZES_ALL][3];
static int16_t *g_iscan[AV1O_TX_SIZES_ALL][3];

static int scan_class(int tx_type) {
  if (tx_type < 10) return 0;         /* default */
  return (tx_type & 1) ? 2 : 1;       /* 11,13,15 -> mcol ; 10,12,14 -> mrow */
}

static void build(int tx_size, int cls) {
  int W = av1o_tx_wide(tx_size), H = av1o_tx_high(tx_size);
  if (W > 32) W = 32;
  if (H > 32) H = 32;
  const int n = W * H;
  int16_t *s = (int16_t *)malloc(sizeof(int16_t) * (size_t)n);
  int16_t *is = (int16_t *)malloc(sizeof(int16_t) * (size_t)n);
  int k = 0;
  if (cls == 2) {
    for (int i = 0; i < n; ++i) s[k++] = (int16_t)i;
  } else if (cls == 1) {
    for (int r = 0; r < H; ++r)
      for (int c = 0; c < W; ++c) s[k++] = (int16_t)(c * H + r);
  } else {
    for (int d = 0; d < W + H - 1; ++d) {
      int r_up; /* walk the diagonal with r increasing? */
      if (H > W) r_up = 1;
      else if (W > H) r_up = 0;
      else r_up = d & 1;
      for (int t = 0; t <= d; ++t) {
        const int r = r_up ? t : d - t, c = d - r;
        if (r < H && c < W) s[k++] = (int16_t)(c * H + r);
      }
    }
  }
  for (int i = 0; i < n; ++i) is[s[i]] = (int16_t)i;
  g_scan[tx_size][cls] = s;
  g_iscan[tx_size][cls] = is;
}

const int16_t *av1o_scan(int tx_size, int tx_type) {
  const int cls = scan_class(tx_type);
  if (!g_scan[tx_size][cls]) build(tx_size, cls);
  return g_scan[tx_size][cls];
}

const int16_t *av1o_iscan(int tx_size, int tx_type) {
  const int cls = scan_class(tx_type);
  if (!g_iscan[tx_size][cls]) build(tx_size, cls);
  return g_iscan[tx_size][cls];
}
