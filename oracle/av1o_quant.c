/*
 * av1o_quant.c -- ORACLE (test infrastructure only): quantisers.
 *
 * Restates av1/encoder/av1_quantize.c:36-264,565-686 ("FP" quantisers, quantiser
 * table build) and aom_dsp/quantize.c:16-316 ("B" dead-zone quantisers, adaptive
 * variants).  The reference walks coefficients in scan order and tracks the last
 * non-zero position; this restatement is written per coefficient position with
 * eob = 1 + max(iscan[rc]) over non-zero outputs, which is the form the CUDA
 * kernels use -- tests/test_oracle_vs_ref.py shows both give identical outputs.
 */
#include "av1o.h"

#include <string.h>

#include "av1o_qlookup.inc"

#define QM_BITS 5 /* AOM_QM_BITS */

static int rpo2(int v, int n) { return n == 0 ? v : (v + ((1 << n) >> 1)) >> n; }
static int32_t apply_sign(int32_t mag, int32_t sign) { return (mag ^ sign) - sign; }
static int32_t mul32(int32_t a, int32_t b) { return (int32_t)((uint32_t)a * (uint32_t)b); }
static int64_t clamp16(int64_t v) { return v > 32767 ? 32767 : (v < -32768 ? -32768 : v); }

int16_t av1o_dc_q(int qindex, int delta, int bd) {
  int q = qindex + delta;
  q = q < 0 ? 0 : (q > 255 ? 255 : q);
  return k_dc_q[(bd - 8) / 2][q];
}
int16_t av1o_ac_q(int qindex, int delta, int bd) {
  int q = qindex + delta;
  q = q < 0 ? 0 : (q > 255 ? 255 : q);
  return k_ac_q[(bd - 8) / 2][q];
}

/* av1_build_quantizer, av1_quantize.c:580-686, one (qindex, plane) row, deltas 0.
 * `sharpness` is the fork's quant_sharpness argument (:602-625). */
void av1o_build_qrow(int bd, int qindex, int plane, int sharpness, Av1oQRow *row) {
  (void)plane; /* all delta_q are 0 here, so Y/U/V rows coincide */
  const int q = qindex;
  const int dcq = av1o_dc_q(q, 0, bd);
  const int lim = bd == 8 ? 148 : (bd == 10 ? 592 : 2368);
  int zbin_factor = q == 0 ? 64 : (dcq < lim ? 84 : 80);
  int round_factor = q == 0 ? 64 : 48;
  int adj = 16 * (7 - sharpness) / 7;
  if (sharpness > 0 && q > 0) {
    zbin_factor = 64 + adj;
    round_factor = 64 - adj;
  } else if (sharpness < 0 && q > 0) {
    adj = 16 * (7 + sharpness) / 7;
    zbin_factor = 64 + adj;   /* 64 - -adj */
    round_factor = 64 - adj;  /* 64 + -adj */
  }
  int round_factor_fp = 64;
  if (sharpness > 0) round_factor_fp = 64 - adj;
  if (sharpness < 0) round_factor_fp = 64 - adj;
  for (int i = 0; i < 8; ++i) {
    const int d = i == 0 ? dcq : av1o_ac_q(q, 0, bd);
    /* invert_quant :580 */
    int l = 0;
    for (uint32_t t = (uint32_t)d; t > 1; t >>= 1) ++l;
    const int m = 1 + (1 << (16 + l)) / d;
    row->quant[i] = (int16_t)(m - (1 << 16));
    row->quant_shift[i] = (int16_t)(1 << (16 - l));
    row->quant_fp[i] = (int16_t)((1 << 16) / d);
    row->round_fp[i] = (int16_t)((round_factor_fp * d) >> 7);
    row->zbin[i] = (int16_t)rpo2(zbin_factor * d, 7);
    row->round[i] = (int16_t)((round_factor * d) >> 7);
    row->dequant[i] = (int16_t)d;
  }
}

static uint16_t eob_from(const int32_t *q, const int16_t *iscan, intptr_t n) {
  int eob = 0;
  for (intptr_t rc = 0; rc < n; ++rc)
    if (q[rc] && iscan[rc] + 1 > eob) eob = iscan[rc] + 1;
  return (uint16_t)eob;
}

/* quantize_fp_helper_c (:70-123) incl. av1_quantize_fp_no_qmatrix (:36-68):
 * low bit depth: abs+round is clamped to int16 */
void av1o_quantize_fp(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                      const int16_t *round, const int16_t *quant,
                      const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                      const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                      const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                      int log_scale) {
  (void)zbin, (void)quant_shift, (void)scan;
  const int rnd[2] = { rpo2(round[0], log_scale), rpo2(round[1], log_scale) };
  for (intptr_t rc = 0; rc < n; ++rc) {
    const int ac = rc != 0;
    const int32_t c = coeff[rc], sign = c < 0 ? -1 : 0;
    int64_t a = apply_sign(c, sign);
    int32_t q = 0, dq = 0;
    if (!qm && !iqm) {
      if ((a << (1 + log_scale)) >= (int32_t)dequant[ac]) {
        a = clamp16(a + rnd[ac]);
        q = (int32_t)((a * quant[ac]) >> (16 - log_scale));
        if (q) dq = mul32(q, dequant[ac]) >> log_scale;
      }
    } else {
      const int wt = qm ? qm[rc] : 32, iwt = iqm ? iqm[rc] : 32;
      const int dqv = (dequant[ac] * iwt + 16) >> QM_BITS;
      if (a * wt >= (dequant[ac] << (QM_BITS - (1 + log_scale)))) {
        a = clamp16(a + rnd[ac]);
        q = (int32_t)((a * wt * quant[ac]) >> (16 - log_scale + QM_BITS));
        dq = mul32(q, dqv) >> log_scale;
      }
    }
    qcoeff[rc] = apply_sign(q, sign);
    dqcoeff[rc] = apply_sign(dq, sign);
  }
  *eob = eob_from(qcoeff, iscan, n);
}

/* highbd_quantize_fp_helper_c (:125-198): no int16 clamp */
void av1o_highbd_quantize_fp(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                             const int16_t *round, const int16_t *quant,
                             const int16_t *quant_shift, int32_t *qcoeff,
                             int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                             const int16_t *scan, const int16_t *iscan,
                             const uint8_t *qm, const uint8_t *iqm, int log_scale) {
  (void)zbin, (void)quant_shift, (void)scan;
  const int shift = 16 - log_scale;
  const int rnd[2] = { rpo2(round[0], log_scale), rpo2(round[1], log_scale) };
  for (intptr_t rc = 0; rc < n; ++rc) {
    const int ac = rc != 0;
    const int32_t c = coeff[rc], sign = c < 0 ? -1 : 0;
    const int32_t a = apply_sign(c, sign);
    int32_t q = 0, dq = 0;
    if (qm || iqm) {
      const int wt = qm ? qm[rc] : 32, iwt = iqm ? iqm[rc] : 32;
      const int dqv = (dequant[ac] * iwt + 16) >> QM_BITS;
      if ((int64_t)a * wt >= (dequant[ac] << (QM_BITS - (1 + log_scale)))) {
        const int64_t t = (int64_t)a + rnd[ac];
        q = (int32_t)((t * quant[ac] * wt) >> (shift + QM_BITS));
        dq = mul32(q, dqv) >> log_scale;
      }
    } else {
      if ((int32_t)((uint32_t)a << (1 + log_scale)) >= dequant[ac]) {
        const int64_t t = (int64_t)a + rnd[ac];
        q = (int32_t)((t * quant[ac]) >> shift);
        dq = mul32(q, dequant[ac]) >> log_scale;
      }
    }
    qcoeff[rc] = apply_sign(q, sign);
    dqcoeff[rc] = apply_sign(dq, sign);
  }
  *eob = eob_from(qcoeff, iscan, n);
}

/* av1_quantize_lp_c (:212-240): int16 in/out, no dead zone */
void av1o_quantize_lp(const int16_t *coeff, intptr_t n, const int16_t *round,
                      const int16_t *quant, int16_t *qcoeff, int16_t *dqcoeff,
                      const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                      const int16_t *iscan) {
  (void)scan;
  int e = 0;
  for (intptr_t rc = 0; rc < n; ++rc) {
    const int ac = rc != 0;
    const int c = coeff[rc], sign = c < 0 ? -1 : 0;
    const int a = (c ^ sign) - sign;
    int t = (int)clamp16(a + round[ac]);
    t = (t * quant[ac]) >> 16;
    qcoeff[rc] = (int16_t)((t ^ sign) - sign);
    dqcoeff[rc] = (int16_t)(qcoeff[rc] * dequant[ac]);
    if (t && iscan[rc] + 1 > e) e = iscan[rc] + 1;
  }
  *eob = (uint16_t)e;
}

/* aom_quantize_b_helper_c (aom_dsp/quantize.c:108-169) and
 * aom_highbd_quantize_b_helper_c (:261-316).  The reference's backwards pre-scan
 * only skips coefficients that the per-coefficient zbin test rejects anyway. */
static void quantize_b_any(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                           const int16_t *round, const int16_t *quant,
                           const int16_t *quant_shift, int32_t *qcoeff,
                           int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                           const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                           int log_scale, int highbd) {
  const int zb[2] = { rpo2(zbin[0], log_scale), rpo2(zbin[1], log_scale) };
  for (intptr_t rc = 0; rc < n; ++rc) {
    const int ac = rc != 0;
    const int32_t c = coeff[rc], sign = c < 0 ? -1 : 0;
    const int32_t a = apply_sign(c, sign);
    const int wt = qm ? qm[rc] : 32, iwt = iqm ? iqm[rc] : 32;
    int32_t q = 0, dq = 0;
    if (mul32(a, wt) >= (zb[ac] << QM_BITS)) {
      int64_t t = (int64_t)a + rpo2(round[ac], log_scale);
      if (!highbd) t = clamp16(t);
      t *= wt;
      q = (int32_t)(((((t * quant[ac]) >> 16) + t) * quant_shift[ac]) >>
                    (16 - log_scale + QM_BITS));
      const int dqv = (dequant[ac] * iwt + 16) >> QM_BITS;
      dq = mul32(q, dqv) >> log_scale;
    }
    qcoeff[rc] = apply_sign(q, sign);
    dqcoeff[rc] = apply_sign(dq, sign);
  }
  *eob = eob_from(qcoeff, iscan, n);
}

void av1o_quantize_b(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                     const int16_t *round, const int16_t *quant,
                     const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                     const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                     const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                     int log_scale) {
  (void)scan;
  quantize_b_any(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff, dequant, eob,
                 iscan, qm, iqm, log_scale, 0);
}

void av1o_highbd_quantize_b(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                            const int16_t *round, const int16_t *quant,
                            const int16_t *quant_shift, int32_t *qcoeff,
                            int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                            const int16_t *scan, const int16_t *iscan,
                            const uint8_t *qm, const uint8_t *iqm, int log_scale) {
  (void)scan;
  quantize_b_any(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff, dequant, eob,
                 iscan, qm, iqm, log_scale, 1);
}

/* aom_quantize_b_adaptive_helper_c (:16-105) / highbd (:173-258), no qmatrix.
 * Step 1: cut the trailing run (scan order) whose |coeff*32| stays below
 *         zbin*32 + round(dequant*325/128).
 * Step 2: plain B quantisation of what is left.
 * Step 3: if exactly one coefficient survived and it is +-1, drop it when its
 *         |coeff*32| is below zbin*32 + round(dequant*525/128). */
static void quantize_b_adaptive_any(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                                    const int16_t *round, const int16_t *quant,
                                    const int16_t *quant_shift, int32_t *qcoeff,
                                    int32_t *dqcoeff, const int16_t *dequant,
                                    uint16_t *eob, const int16_t *scan, int log_scale,
                                    int highbd) {
  const int zb[2] = { rpo2(zbin[0], log_scale), rpo2(zbin[1], log_scale) };
  const int add[2] = { rpo2(dequant[0] * 325, 7), rpo2(dequant[1] * 325, 7) };
  intptr_t keep = n;
  while (keep > 0) {
    const int rc = scan[keep - 1], ac = rc != 0;
    const int32_t cw = mul32(coeff[rc], 32);
    if (cw < zb[ac] * 32 + add[ac] && cw > -zb[ac] * 32 - add[ac]) --keep;
    else break;
  }
  memset(qcoeff, 0, sizeof(int32_t) * (size_t)n);
  memset(dqcoeff, 0, sizeof(int32_t) * (size_t)n);
  int last = -1, first = -1;
  for (intptr_t i = 0; i < keep; ++i) {
    const int rc = scan[i], ac = rc != 0;
    const int32_t c = coeff[rc], sign = c < 0 ? -1 : 0;
    const int32_t a = apply_sign(c, sign);
    if (mul32(a, 32) >= (zb[ac] << QM_BITS)) {
      int64_t t = (int64_t)a + rpo2(round[ac], log_scale);
      if (!highbd) t = clamp16(t);
      t *= 32;
      const int32_t q = (int32_t)(((((t * quant[ac]) >> 16) + t) * quant_shift[ac]) >>
                                  (16 - log_scale + QM_BITS));
      qcoeff[rc] = apply_sign(q, sign);
      dqcoeff[rc] = apply_sign(mul32(q, dequant[ac]) >> log_scale, sign);
      if (q) {
        last = (int)i;
        if (first < 0) first = (int)i;
      }
    }
  }
  if (last >= 0 && first == last) {
    const int rc = scan[last], ac = rc != 0;
    if (qcoeff[rc] == 1 || qcoeff[rc] == -1) {
      const int32_t cw = mul32(coeff[rc], 32);
      const int add2 = rpo2(dequant[ac] * (325 + 200), 7);
      if (cw < zb[ac] * 32 + add2 && cw > -zb[ac] * 32 - add2) {
        qcoeff[rc] = 0;
        dqcoeff[rc] = 0;
        last = -1;
      }
    }
  }
  *eob = (uint16_t)(last + 1);
}

void av1o_quantize_b_adaptive(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                              const int16_t *round, const int16_t *quant,
                              const int16_t *quant_shift, int32_t *qcoeff,
                              int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                              const int16_t *scan, const int16_t *iscan, int log_scale) {
  (void)iscan;
  quantize_b_adaptive_any(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff,
                          dequant, eob, scan, log_scale, 0);
}

void av1o_highbd_quantize_b_adaptive(const int32_t *coeff, intptr_t n,
                                     const int16_t *zbin, const int16_t *round,
                                     const int16_t *quant, const int16_t *quant_shift,
                                     int32_t *qcoeff, int32_t *dqcoeff,
                                     const int16_t *dequant, uint16_t *eob,
                                     const int16_t *scan, const int16_t *iscan,
                                     int log_scale) {
  (void)iscan;
  quantize_b_adaptive_any(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff,
                          dequant, eob, scan, log_scale, 1);
}
