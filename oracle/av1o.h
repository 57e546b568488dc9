/*
 * av1o.h -- CPU ORACLE for the B200 AV1 hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is a from-scratch plain-C restatement of the reference's (aom-av1-lavish,
 * libaom 3.6 based) C path for: forward 2-D transforms, quantisers, block error,
 * subtract, SAD / variance / sub-pel variance and the full-pel / sub-pel motion
 * search loops.  Every function cites the reference file:line it restates.
 *
 * Parity status: PINNED.  tests/test_oracle_vs_ref.py checks every entry point
 * bit-exactly against the reference's own `_c` functions compiled from
 * /root/reference into oracle/_ref/libaomref.so (recipe: oracle/mk_ref.py), on the
 * input distributions of the reference's unit tests, and
 * tests/golden/ holds vectors produced by that reference build for machines where
 * /root/reference is absent.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library.  The product (libb200av1.so) never links or calls it.
 */
#ifndef AV1O_H_
#define AV1O_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* TX_SIZE / TX_TYPE numbering follows aom_dsp/txfm_common.h:25-71 */
enum {
  AV1O_TX_4X4, AV1O_TX_8X8, AV1O_TX_16X16, AV1O_TX_32X32, AV1O_TX_64X64,
  AV1O_TX_4X8, AV1O_TX_8X4, AV1O_TX_8X16, AV1O_TX_16X8, AV1O_TX_16X32,
  AV1O_TX_32X16, AV1O_TX_32X64, AV1O_TX_64X32, AV1O_TX_4X16, AV1O_TX_16X4,
  AV1O_TX_8X32, AV1O_TX_32X8, AV1O_TX_16X64, AV1O_TX_64X16, AV1O_TX_SIZES_ALL
};
enum { AV1O_TX_TYPES = 16 };

/* ---- transforms (av1o_txfm.c) ------------------------------------------------ */
int av1o_tx_wide(int tx_size);
int av1o_tx_high(int tx_size);
/* number of coefficients the codec keeps: av1_get_max_eob, blockd.h:1596 */
int av1o_max_eob(int tx_size);
/* 1 when (tx_size, tx_type) has a forward transform (test/av1_txfm_test.h:90) */
int av1o_tx_valid(int tx_size, int tx_type);
const int32_t *av1o_cospi(int cos_bit); /* 64 entries, cos_bit 10..16 */
const int32_t *av1o_sinpi(int cos_bit); /* 5 entries */
/* 1-D kernels: kind 0=DCT 1=ADST 2=IDENTITY, n in {4,8,16,32,64} */
void av1o_fwd_txfm1d(int kind, int n, const int32_t *in, int32_t *out, int cos_bit);
/* av1_fwd_txfm2d_WxH_c: writes av1o_max_eob(tx_size) coefficients (column-major,
 * 64-point sizes already re-packed); anything beyond that is not part of parity. */
void av1o_fwd_txfm2d(const int16_t *input, int32_t *output, int stride, int tx_size,
                     int tx_type, int bd);
void av1o_fwht4x4(const int16_t *input, int32_t *output, int stride);
void av1o_round_shift_array(int32_t *arr, int size, int bit);

/* ---- inverse transforms, reconstruction, pixel-domain distortion (av1o_itxfm.c) -------- */
/* 1-D inverse kernels (av1_inv_txfm1d.c): kind 0=DCT 1=ADST 2=IDENTITY, range_bits = stage range */
void av1o_inv_txfm1d(int kind, int n, const int32_t *in, int32_t *out, int range_bits);
/* av1_inv_txfm2d_add_WxH_c (av1_inv_txfm2d.c:234): input = kept coefficients, column-major */
void av1o_inv_txfm2d_add(const int32_t *input, uint16_t *dest, int stride, int tx_size, int tx_type,
                         int bd);
void av1o_iwht4x4_add(const int32_t *input, uint16_t *dest, int stride, int eob, int bd);
/* av1_inverse_transform_block (idct.c:303) on uint8 / uint16 pixel buffers */
void av1o_recon_block(const int32_t *dqcoeff, const void *pred, int pred_stride, void *recon,
                      int recon_stride, int tx_size, int tx_type, int eob, int lossless, int highbd,
                      int bd);
/* dist_block_px_domain (tx_search.c:969): 16 * sse over the visible region */
/* av1_cost_coeffs_txb_estimate (txb_rdopt.c:547) */
int av1o_cost_coeffs_txb_estimate(const int32_t *qcoeff, int eob, int tx_size, int tx_type);
int64_t av1o_dist_block_px(const void *src, int src_stride, const void *recon, int recon_stride,
                           int vis_w, int vis_h, int highbd, int bd);

/* ---- Hadamard / SATD / averages (av1o_avg.c; aom_dsp/avg.c) -------------------- */
/* n x n Hadamard; kind 0 aom_hadamard_NxN_c (n 4..32), 1 aom_hadamard_lp_NxN_c (n 8, 16; int16
 * values widened), 2 aom_highbd_hadamard_NxN_c (n 8..32) */
void av1o_hadamard(const int16_t *src, ptrdiff_t stride, int32_t *coeff, int n, int kind);
void av1o_hadamard_4x4(const int16_t *src, ptrdiff_t stride, int32_t *coeff);
int av1o_satd(const int32_t *coeff, int length);
unsigned av1o_avg(const void *s, int stride, int n, int highbd);
void av1o_minmax_8x8(const void *s, int p, const void *d, int dp, int highbd, int *mn, int *mx);
void av1o_int_pro_row(int16_t *hbuf, const uint8_t *ref, int ref_stride, int width, int height,
                      int norm_factor);
void av1o_int_pro_col(int16_t *vbuf, const uint8_t *ref, int ref_stride, int width, int height,
                      int norm_factor);
int av1o_vector_var(const int16_t *ref, const int16_t *src, int bwl);

/* ---- compound / masked / OBMC SAD and variance (av1o_compound.c) ------------------ */
enum { AV1O_BLEND_NONE = 0, AV1O_BLEND_AVG, AV1O_BLEND_DIST_WTD, AV1O_BLEND_MASK };
typedef struct {
  int w, h, highbd, bd;
  const void *pre;          /* predictor plane (the reference functions' `ref` / interpolated `a` / `pre`) */
  int pre_stride;
  int subpel, xoff, yoff;   /* bilinear 1/8-pel interpolation of pre (the *_sub_pixel_* functions) */
  const void *tgt;          /* the plane compared against (`src` / `b`) */
  int tgt_stride;
  int blend;                /* AV1O_BLEND_* */
  const void *second_pred;  /* w-pitch block */
  const uint8_t *mask;      /* AV1O_BLEND_MASK */
  int mask_stride, invert_mask;
  int fwd_offset, bck_offset; /* DIST_WTD_COMP_PARAMS */
  const int32_t *wsrc, *obmc_mask; /* OBMC: w-pitch blocks */
} Av1oCompound;
void av1o_compound_pred(const Av1oCompound *c, void *comp_pred);
unsigned av1o_compound_sad(const Av1oCompound *c);
unsigned av1o_compound_variance(const Av1oCompound *c, unsigned *sse);
unsigned av1o_obmc_sad(const Av1oCompound *c);
unsigned av1o_obmc_variance(const Av1oCompound *c, unsigned *sse);

/* ---- scan tables (av1o_scan.c) ----------------------------------------------- */
/* scan / iscan of av1_scan_orders[tx_size][tx_type] (scan.c:1666); n = max_eob */
const int16_t *av1o_scan(int tx_size, int tx_type);
const int16_t *av1o_iscan(int tx_size, int tx_type);

/* ---- quantisers (av1o_quant.c) ------------------------------------------------ */
/* One qindex row of the QUANTS/Dequants tables (av1_quantize.h:45-84), plane 0..2.
 * Each array has 8 int16 entries like the reference rows ([0]=DC, [1..7]=AC). */
typedef struct {
  int16_t quant[8], quant_shift[8], zbin[8], round[8];
  int16_t quant_fp[8], round_fp[8], dequant[8];
} Av1oQRow;
int16_t av1o_dc_q(int qindex, int delta, int bd);
int16_t av1o_ac_q(int qindex, int delta, int bd);
/* av1_build_quantizer (av1_quantize.c:602) for one qindex / plane, all deltas 0 */
void av1o_build_qrow(int bd, int qindex, int plane, int sharpness, Av1oQRow *row);

void av1o_quantize_fp(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                      const int16_t *round, const int16_t *quant,
                      const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                      const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                      const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                      int log_scale);
void av1o_highbd_quantize_fp(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                             const int16_t *round, const int16_t *quant,
                             const int16_t *quant_shift, int32_t *qcoeff,
                             int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                             const int16_t *scan, const int16_t *iscan,
                             const uint8_t *qm, const uint8_t *iqm, int log_scale);
void av1o_quantize_lp(const int16_t *coeff, intptr_t n, const int16_t *round,
                      const int16_t *quant, int16_t *qcoeff, int16_t *dqcoeff,
                      const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                      const int16_t *iscan);
void av1o_quantize_b(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                     const int16_t *round, const int16_t *quant,
                     const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                     const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                     const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                     int log_scale);
void av1o_highbd_quantize_b(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                            const int16_t *round, const int16_t *quant,
                            const int16_t *quant_shift, int32_t *qcoeff,
                            int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                            const int16_t *scan, const int16_t *iscan,
                            const uint8_t *qm, const uint8_t *iqm, int log_scale);
void av1o_quantize_b_adaptive(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                              const int16_t *round, const int16_t *quant,
                              const int16_t *quant_shift, int32_t *qcoeff,
                              int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                              const int16_t *scan, const int16_t *iscan, int log_scale);
void av1o_highbd_quantize_b_adaptive(const int32_t *coeff, intptr_t n,
                                     const int16_t *zbin, const int16_t *round,
                                     const int16_t *quant, const int16_t *quant_shift,
                                     int32_t *qcoeff, int32_t *dqcoeff,
                                     const int16_t *dequant, uint16_t *eob,
                                     const int16_t *scan, const int16_t *iscan,
                                     int log_scale);

/* ---- distortion / residual (av1o_dist.c) --------------------------------------- */
int64_t av1o_block_error(const int32_t *coeff, const int32_t *dqcoeff, intptr_t n,
                         int64_t *ssz);
int64_t av1o_block_error_lp(const int16_t *coeff, const int16_t *dqcoeff, intptr_t n);
int64_t av1o_highbd_block_error(const int32_t *coeff, const int32_t *dqcoeff,
                                intptr_t n, int64_t *ssz, int bd);
int64_t av1o_block_error_qm(const int32_t *coeff, const int32_t *dqcoeff, intptr_t n,
                            const uint8_t *qm, const int16_t *scan, int64_t *ssz);
void av1o_subtract_block(int rows, int cols, int16_t *diff, ptrdiff_t diff_stride,
                         const uint8_t *src, ptrdiff_t src_stride, const uint8_t *pred,
                         ptrdiff_t pred_stride);
void av1o_highbd_subtract_block(int rows, int cols, int16_t *diff, ptrdiff_t diff_stride,
                                const uint16_t *src, ptrdiff_t src_stride,
                                const uint16_t *pred, ptrdiff_t pred_stride);
int64_t av1o_sse(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int w,
                 int h);
int64_t av1o_highbd_sse(const uint16_t *a, int a_stride, const uint16_t *b, int b_stride,
                        int w, int h);

/* ---- SAD / variance (av1o_sadvar.c) -------------------------------------------- */
unsigned av1o_sad(const uint8_t *src, int src_stride, const uint8_t *ref, int ref_stride,
                  int w, int h);
unsigned av1o_sad_skip(const uint8_t *src, int src_stride, const uint8_t *ref,
                       int ref_stride, int w, int h);
unsigned av1o_sad_avg(const uint8_t *src, int src_stride, const uint8_t *ref,
                      int ref_stride, const uint8_t *second_pred, int w, int h);
void av1o_sad_x4d(const uint8_t *src, int src_stride, const uint8_t *const ref[4],
                  int ref_stride, uint32_t sad[4], int w, int h, int n_refs, int skip);
unsigned av1o_highbd_sad(const uint16_t *src, int src_stride, const uint16_t *ref,
                         int ref_stride, int w, int h);
unsigned av1o_highbd_sad_skip(const uint16_t *src, int src_stride, const uint16_t *ref,
                              int ref_stride, int w, int h);
/* variance: a = "src_ptr" argument of aom_varianceWxH, b = "ref_ptr" */
unsigned av1o_variance(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride,
                       int w, int h, unsigned *sse);
unsigned av1o_mse(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int w,
                  int h, unsigned *sse);
unsigned av1o_sub_pixel_variance(const uint8_t *a, int a_stride, int xoff, int yoff,
                                 const uint8_t *b, int b_stride, int w, int h,
                                 unsigned *sse);
unsigned av1o_highbd_variance(const uint16_t *a, int a_stride, const uint16_t *b,
                              int b_stride, int w, int h, int bd, unsigned *sse);
unsigned av1o_highbd_sub_pixel_variance(const uint16_t *a, int a_stride, int xoff,
                                        int yoff, const uint16_t *b, int b_stride, int w,
                                        int h, int bd, unsigned *sse);
/* aom_upsampled_pred_c (reconinter_enc.c:424), unscaled references only.
 * taps: 2 (bilinear), 4 or 8 == USE_2_TAPS / USE_4_TAPS / USE_8_TAPS */
void av1o_upsampled_pred(uint8_t *comp_pred, int width, int height, int subpel_x_q3,
                         int subpel_y_q3, const uint8_t *ref, int ref_stride, int taps);
void av1o_highbd_upsampled_pred(uint16_t *comp_pred, int width, int height,
                                int subpel_x_q3, int subpel_y_q3, const uint16_t *ref,
                                int ref_stride, int bd, int taps);

#ifdef __cplusplus
}
#endif
#endif /* AV1O_H_ */
