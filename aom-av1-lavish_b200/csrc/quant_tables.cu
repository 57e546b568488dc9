// quant_tables.cu -- host side of the quantiser rows: one (qindex, plane) row of the tables
// av1_build_quantizer fills (av1/encoder/av1_quantize.c:580-686, incl. the fork's
// quant_sharpness argument) from the AV1 step-size tables (quant_common.c:193-215).
#include "internal.h"

namespace {
#include "qlookup.inc"

inline int clamp_q(int v) { return v < 0 ? 0 : (v > 255 ? 255 : v); }
}  // namespace

extern "C" int b200_dc_quant(int qindex, int delta, int bd) {
  const int b = bd == 8 ? 0 : (bd == 10 ? 1 : (bd == 12 ? 2 : -1));
  return b < 0 ? -1 : kDcQ[b][clamp_q(qindex + delta)];
}
extern "C" int b200_ac_quant(int qindex, int delta, int bd) {
  const int b = bd == 8 ? 0 : (bd == 10 ? 1 : (bd == 12 ? 2 : -1));
  return b < 0 ? -1 : kAcQ[b][clamp_q(qindex + delta)];
}

extern "C" int b200_build_quantizer_row(int bd, int qindex, int dc_delta_q, int ac_delta_q,
                                        int quant_sharpness, B200QRow *row) {
  if (!row || qindex < 0 || qindex > 255 || (bd != 8 && bd != 10 && bd != 12)) {
    b200::set_error("b200_build_quantizer_row: qindex 0..255 and bd 8/10/12 expected");
    return B200_EINVAL;
  }
  // get_qzbin_factor (:585-598) looks at the DC step with delta 0
  const int dc0 = b200_dc_quant(qindex, 0, bd);
  const int knee = 148 << (bd - 8);
  int zbin_factor = qindex == 0 ? 64 : (dc0 < knee ? 84 : 80);
  int round_factor = qindex == 0 ? 64 : 48;
  // quant_sharpness of either sign trades dead zone against rounding by the same amount.  The
  // reference recomputes the adjustment from |sharpness| only inside its q > 0 branch, so at
  // qindex 0 a negative sharpness leaves the first value (7 - s, i.e. 7 + |s|) in place for
  // the FP rounding factor (:606-625) -- kept.
  int adj = 16 * (7 - quant_sharpness) / 7;
  if (quant_sharpness < 0 && qindex > 0) adj = 16 * (7 + quant_sharpness) / 7;
  if (quant_sharpness != 0 && qindex > 0) {
    zbin_factor = 64 + adj;
    round_factor = 64 - adj;
  }
  const int round_factor_fp = quant_sharpness != 0 ? 64 - adj : 64;
  for (int i = 0; i < 2; ++i) {
    const int step = i == 0 ? b200_dc_quant(qindex, dc_delta_q, bd) : b200_ac_quant(qindex, ac_delta_q, bd);
    int msb = 0;  // invert_quant (:580): 16-bit reciprocal split into multiplier and shift
    while ((step >> (msb + 1)) != 0) ++msb;
    row->quant[i] = (int16_t)(1 + (1 << (16 + msb)) / step - (1 << 16));
    row->quant_shift[i] = (int16_t)(1 << (16 - msb));
    row->quant_fp[i] = (int16_t)((1 << 16) / step);
    row->round_fp[i] = (int16_t)((round_factor_fp * step) >> 7);
    row->zbin[i] = (int16_t)((zbin_factor * step + 64) >> 7);
    row->round[i] = (int16_t)((round_factor * step) >> 7);
    row->dequant[i] = (int16_t)step;
  }
  row->reserved[0] = row->reserved[1] = 0;
  return B200_OK;
}
