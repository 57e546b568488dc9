/*
 * b200av1_rtcd.h -- drop-in tier: the reference's RTCD prototypes with a `_b200` suffix.
 *
 * libaom dispatches its DSP kernels through run-time CPU dispatch ("RTCD"): for every
 * `add_proto` in av1/common/av1_rtcd_defs.pl and aom_dsp/aom_dsp_rtcd_defs.pl the
 * generated header declares one prototype per flavour (`_c`, `_sse2`, `_avx2`, ...) and
 * either a `#define name name_c` or a function pointer `name` set in
 * setup_rtcd_internal() (build/cmake/rtcd.pl:168-290).  This header adds the flavour
 * `_b200` for the hot-path members, with EXACTLY the reference argument lists, so that
 * a maintainer can `specialize` them.  Each call moves one block to
 * the GPU, runs the same kernels as the batched tier on a batch of one, and copies the
 * result back: correct and complete, but latency-bound -- throughput comes from the
 * batched tier (b200av1.h).  Calls are serialised on the default context
 * (b200_default_ctx()); a CUDA failure is fatal (abort) because these signatures have
 * no error channel.
 *
 * High-bit-depth pixel pointers follow the reference convention: uint16_t data passed
 * as CONVERT_TO_BYTEPTR(ptr) (aom_ports/mem.h:79-80), i.e. (uint8_t *)((uintptr_t)p >> 1).
 */
#ifndef B200AV1_RTCD_H_
#define B200AV1_RTCD_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int32_t b200_tran_low_t; /* tran_low_t, aom_dsp/aom_dsp_common.h:57 */
typedef uint8_t b200_qm_val_t;   /* qm_val_t */

/* TxfmParam, aom_dsp/txfm_common.h:84-101 (layout-compatible).  Inside libaom (generated rtcd
 * headers, tools/gen_rtcd_b200.py) define B200AV1_RTCD_USE_AOM_TYPES: the prototypes then use the
 * encoder's own TxfmParam / DIST_WTD_COMP_PARAMS so that they are the very types of the add_proto
 * lines they replace. */
#ifdef B200AV1_RTCD_USE_AOM_TYPES
typedef struct txfm_param B200TxfmParam;
#else
typedef struct {
  uint8_t tx_type;     /* TX_TYPE */
  uint8_t tx_size;     /* TX_SIZE */
  int lossless;
  int bd;
  int is_hbd;
  uint8_t tx_set_type; /* TxSetType */
  int eob;
} B200TxfmParam;
#endif

/* ---- forward transforms: av1_rtcd_defs.pl:351-399 (av1_fwd_txfm2d_WxH), :269 ---- */
#define B200_DECL_FWD_TXFM(W, H)                                                           \
  void av1_fwd_txfm2d_##W##x##H##_b200(const int16_t *input, int32_t *output, int stride, \
                                       uint8_t tx_type, int bd);
B200_DECL_FWD_TXFM(4, 4)
B200_DECL_FWD_TXFM(8, 8)
B200_DECL_FWD_TXFM(16, 16)
B200_DECL_FWD_TXFM(32, 32)
B200_DECL_FWD_TXFM(64, 64)
B200_DECL_FWD_TXFM(4, 8)
B200_DECL_FWD_TXFM(8, 4)
B200_DECL_FWD_TXFM(8, 16)
B200_DECL_FWD_TXFM(16, 8)
B200_DECL_FWD_TXFM(16, 32)
B200_DECL_FWD_TXFM(32, 16)
B200_DECL_FWD_TXFM(32, 64)
B200_DECL_FWD_TXFM(64, 32)
B200_DECL_FWD_TXFM(4, 16)
B200_DECL_FWD_TXFM(16, 4)
B200_DECL_FWD_TXFM(8, 32)
B200_DECL_FWD_TXFM(32, 8)
B200_DECL_FWD_TXFM(16, 64)
B200_DECL_FWD_TXFM(64, 16)
/* av1_lowbd_fwd_txfm (av1_rtcd_defs.pl:349; hybrid_fwd_txfm.c:241): size switch;
 * txfm_param->lossless takes the WHT (av1_fwht4x4_b200). */
void av1_lowbd_fwd_txfm_b200(const int16_t *src_diff, b200_tran_low_t *coeff, int diff_stride,
                             B200TxfmParam *txfm_param);

/* ---- inverse transforms: av1_rtcd_defs.pl:137-141 (av1_inv_txfm_add, av1_highbd_inv_txfm_add),
 * :182-219 (av1_inv_txfm2d_add_WxH); output / dst holds the prediction and receives the
 * reconstruction; lossless blocks take av1_highbd_iwht4x4_{16,1}_add by txfm_param->eob ---- */
#define B200_DECL_INV_TXFM(W, H)                                                                 \
  void av1_inv_txfm2d_add_##W##x##H##_b200(const int32_t *input, uint16_t *output, int stride,  \
                                           uint8_t tx_type, int bd);
B200_DECL_INV_TXFM(4, 4)
B200_DECL_INV_TXFM(8, 8)
B200_DECL_INV_TXFM(16, 16)
B200_DECL_INV_TXFM(32, 32)
B200_DECL_INV_TXFM(64, 64)
B200_DECL_INV_TXFM(4, 8)
B200_DECL_INV_TXFM(8, 4)
B200_DECL_INV_TXFM(8, 16)
B200_DECL_INV_TXFM(16, 8)
B200_DECL_INV_TXFM(16, 32)
B200_DECL_INV_TXFM(32, 16)
B200_DECL_INV_TXFM(32, 64)
B200_DECL_INV_TXFM(64, 32)
B200_DECL_INV_TXFM(4, 16)
B200_DECL_INV_TXFM(16, 4)
B200_DECL_INV_TXFM(8, 32)
B200_DECL_INV_TXFM(32, 8)
B200_DECL_INV_TXFM(16, 64)
B200_DECL_INV_TXFM(64, 16)
void av1_inv_txfm_add_b200(const b200_tran_low_t *dqcoeff, uint8_t *dst, int stride,
                           const B200TxfmParam *txfm_param);
void av1_highbd_inv_txfm_add_b200(const b200_tran_low_t *input, uint8_t *dest, int stride,
                                  const B200TxfmParam *txfm_param);

/* ---- quantisers: av1_rtcd_defs.pl:334-347,428; aom_dsp_rtcd_defs.pl:655-694 ---- */
#define B200_QUANT_ARGS                                                                       \
  const b200_tran_low_t *coeff_ptr, intptr_t n_coeffs, const int16_t *zbin_ptr,               \
      const int16_t *round_ptr, const int16_t *quant_ptr, const int16_t *quant_shift_ptr,     \
      b200_tran_low_t *qcoeff_ptr, b200_tran_low_t *dqcoeff_ptr, const int16_t *dequant_ptr,  \
      uint16_t *eob_ptr, const int16_t *scan, const int16_t *iscan
void av1_quantize_fp_b200(B200_QUANT_ARGS);
void av1_quantize_fp_32x32_b200(B200_QUANT_ARGS);
void av1_quantize_fp_64x64_b200(B200_QUANT_ARGS);
void av1_highbd_quantize_fp_b200(B200_QUANT_ARGS, int log_scale);
void av1_quantize_lp_b200(const int16_t *coeff_ptr, intptr_t n_coeffs, const int16_t *round_ptr,
                          const int16_t *quant_ptr, int16_t *qcoeff_ptr, int16_t *dqcoeff_ptr,
                          const int16_t *dequant_ptr, uint16_t *eob_ptr, const int16_t *scan,
                          const int16_t *iscan);
void aom_quantize_b_b200(B200_QUANT_ARGS);
void aom_quantize_b_32x32_b200(B200_QUANT_ARGS);
void aom_quantize_b_64x64_b200(B200_QUANT_ARGS);
void aom_quantize_b_adaptive_b200(B200_QUANT_ARGS);
void aom_quantize_b_32x32_adaptive_b200(B200_QUANT_ARGS);
void aom_quantize_b_64x64_adaptive_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_32x32_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_64x64_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_adaptive_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_32x32_adaptive_b200(B200_QUANT_ARGS);
void aom_highbd_quantize_b_64x64_adaptive_b200(B200_QUANT_ARGS);
void aom_quantize_b_helper_b200(B200_QUANT_ARGS, const b200_qm_val_t *qm_ptr,
                                const b200_qm_val_t *iqm_ptr, const int log_scale);
void aom_highbd_quantize_b_helper_b200(B200_QUANT_ARGS, const b200_qm_val_t *qm_ptr,
                                       const b200_qm_val_t *iqm_ptr, const int log_scale);

/* ---- transform-domain distortion: av1_rtcd_defs.pl:328-332,423 ---- */
int64_t av1_block_error_b200(const b200_tran_low_t *coeff, const b200_tran_low_t *dqcoeff,
                             intptr_t block_size, int64_t *ssz);
int64_t av1_block_error_lp_b200(const int16_t *coeff, const int16_t *dqcoeff, intptr_t block_size);
int64_t av1_highbd_block_error_b200(const b200_tran_low_t *coeff, const b200_tran_low_t *dqcoeff,
                                    intptr_t block_size, int64_t *ssz, int bd);

/* ---- residual / SSE: aom_dsp_rtcd_defs.pl:725-739 ---- */
void aom_subtract_block_b200(int rows, int cols, int16_t *diff_ptr, ptrdiff_t diff_stride,
                             const uint8_t *src_ptr, ptrdiff_t src_stride,
                             const uint8_t *pred_ptr, ptrdiff_t pred_stride);
void aom_highbd_subtract_block_b200(int rows, int cols, int16_t *diff_ptr, ptrdiff_t diff_stride,
                                    const uint8_t *src_ptr, ptrdiff_t src_stride,
                                    const uint8_t *pred_ptr, ptrdiff_t pred_stride);
int64_t aom_sse_b200(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int width,
                     int height);
int64_t aom_highbd_sse_b200(const uint8_t *a8, int a_stride, const uint8_t *b8, int b_stride,
                            int width, int height);

/* ---- SAD / variance: aom_dsp_rtcd_defs.pl:762-768,879-885,1001-1008,1136-1145,
 *      1303-1323,1349-1367,1466-1516; 22 block sizes (:42-58) ---- */
#define B200_DECL_BLOCK_FNS(W, H)                                                                 \
  unsigned int aom_sad##W##x##H##_b200(const uint8_t *src_ptr, int src_stride,                     \
                                       const uint8_t *ref_ptr, int ref_stride);                    \
  unsigned int aom_sad_skip_##W##x##H##_b200(const uint8_t *src_ptr, int src_stride,               \
                                             const uint8_t *ref_ptr, int ref_stride);              \
  unsigned int aom_sad##W##x##H##_avg_b200(const uint8_t *src_ptr, int src_stride,                 \
                                           const uint8_t *ref_ptr, int ref_stride,                 \
                                           const uint8_t *second_pred);                            \
  void aom_sad##W##x##H##x4d_b200(const uint8_t *src_ptr, int src_stride,                          \
                                  const uint8_t *const ref_ptr[4], int ref_stride,                 \
                                  uint32_t sad_array[4]);                                          \
  void aom_sad##W##x##H##x3d_b200(const uint8_t *src_ptr, int src_stride,                          \
                                  const uint8_t *const ref_ptr[4], int ref_stride,                 \
                                  uint32_t sad_array[4]);                                          \
  void aom_sad_skip_##W##x##H##x4d_b200(const uint8_t *src_ptr, int src_stride,                    \
                                        const uint8_t *const ref_ptr[4], int ref_stride,           \
                                        uint32_t sad_array[4]);                                    \
  unsigned int aom_highbd_sad##W##x##H##_b200(const uint8_t *src_ptr, int src_stride,              \
                                              const uint8_t *ref_ptr, int ref_stride);             \
  unsigned int aom_highbd_sad_skip_##W##x##H##_b200(const uint8_t *src_ptr, int src_stride,        \
                                                    const uint8_t *ref_ptr, int ref_stride);       \
  void aom_highbd_sad##W##x##H##x4d_b200(const uint8_t *src_ptr, int src_stride,                   \
                                         const uint8_t *const ref_ptr[4], int ref_stride,          \
                                         uint32_t sad_array[4]);                                   \
  unsigned int aom_variance##W##x##H##_b200(const uint8_t *src_ptr, int source_stride,             \
                                            const uint8_t *ref_ptr, int ref_stride,                \
                                            unsigned int *sse);                                    \
  uint32_t aom_sub_pixel_variance##W##x##H##_b200(const uint8_t *src_ptr, int source_stride,       \
                                                  int xoffset, int yoffset,                        \
                                                  const uint8_t *ref_ptr, int ref_stride,          \
                                                  uint32_t *sse);                                  \
  unsigned int aom_highbd_8_variance##W##x##H##_b200(const uint8_t *src_ptr, int source_stride,    \
                                                     const uint8_t *ref_ptr, int ref_stride,       \
                                                     unsigned int *sse);                           \
  unsigned int aom_highbd_10_variance##W##x##H##_b200(const uint8_t *src_ptr, int source_stride,   \
                                                      const uint8_t *ref_ptr, int ref_stride,      \
                                                      unsigned int *sse);                          \
  unsigned int aom_highbd_12_variance##W##x##H##_b200(const uint8_t *src_ptr, int source_stride,   \
                                                      const uint8_t *ref_ptr, int ref_stride,      \
                                                      unsigned int *sse);                          \
  uint32_t aom_highbd_8_sub_pixel_variance##W##x##H##_b200(                                        \
      const uint8_t *src_ptr, int source_stride, int xoffset, int yoffset,                         \
      const uint8_t *ref_ptr, int ref_stride, uint32_t *sse);                                      \
  uint32_t aom_highbd_10_sub_pixel_variance##W##x##H##_b200(                                       \
      const uint8_t *src_ptr, int source_stride, int xoffset, int yoffset,                         \
      const uint8_t *ref_ptr, int ref_stride, uint32_t *sse);                                      \
  uint32_t aom_highbd_12_sub_pixel_variance##W##x##H##_b200(                                       \
      const uint8_t *src_ptr, int source_stride, int xoffset, int yoffset,                         \
      const uint8_t *ref_ptr, int ref_stride, uint32_t *sse);
B200_DECL_BLOCK_FNS(128, 128)
B200_DECL_BLOCK_FNS(128, 64)
B200_DECL_BLOCK_FNS(64, 128)
B200_DECL_BLOCK_FNS(64, 64)
B200_DECL_BLOCK_FNS(64, 32)
B200_DECL_BLOCK_FNS(32, 64)
B200_DECL_BLOCK_FNS(32, 32)
B200_DECL_BLOCK_FNS(32, 16)
B200_DECL_BLOCK_FNS(16, 32)
B200_DECL_BLOCK_FNS(16, 16)
B200_DECL_BLOCK_FNS(16, 8)
B200_DECL_BLOCK_FNS(8, 16)
B200_DECL_BLOCK_FNS(8, 8)
B200_DECL_BLOCK_FNS(8, 4)
B200_DECL_BLOCK_FNS(4, 8)
B200_DECL_BLOCK_FNS(4, 4)
B200_DECL_BLOCK_FNS(4, 16)
B200_DECL_BLOCK_FNS(16, 4)
B200_DECL_BLOCK_FNS(8, 32)
B200_DECL_BLOCK_FNS(32, 8)
B200_DECL_BLOCK_FNS(16, 64)
B200_DECL_BLOCK_FNS(64, 16)
/* aom_mse16x16 / 16x8 / 8x16 / 8x8 (aom_dsp_rtcd_defs.pl:1349-1352) */
unsigned int aom_mse16x16_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr,
                               int recon_stride, unsigned int *sse);
unsigned int aom_mse16x8_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr,
                              int recon_stride, unsigned int *sse);
unsigned int aom_mse8x16_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr,
                              int recon_stride, unsigned int *sse);
unsigned int aom_mse8x8_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr,
                             int recon_stride, unsigned int *sse);

/* ---- sub-pel predictor: av1_rtcd_defs.pl:282-285,309-312.  xd / cm / mv are only used by
 * the reference for scaled references; this flavour implements the unscaled branch
 * (reconinter_enc.c:461-495) and ignores them. ---- */
#ifdef B200AV1_RTCD_USE_AOM_TYPES
#define B200_XD_T MACROBLOCKD
#define B200_CM_T struct AV1Common
#define B200_MV_T MV
#else
#define B200_XD_T void
#define B200_CM_T void
#define B200_MV_T void
#endif
void aom_upsampled_pred_b200(B200_XD_T *xd, const B200_CM_T *const cm, int mi_row, int mi_col,
                             const B200_MV_T *const mv, uint8_t *comp_pred, int width, int height,
                             int subpel_x_q3, int subpel_y_q3, const uint8_t *ref, int ref_stride,
                             int subpel_search);
void aom_highbd_upsampled_pred_b200(B200_XD_T *xd, const B200_CM_T *const cm, int mi_row, int mi_col,
                                    const B200_MV_T *const mv, uint8_t *comp_pred8, int width, int height,
                                    int subpel_x_q3, int subpel_y_q3, const uint8_t *ref8,
                                    int ref_stride, int bd, int subpel_search);
/* ---- Hadamard / SATD / averages: aom_dsp_rtcd_defs.pl:1209-1279 (aom_dsp/avg.c), and the
 * lossless WHT av1_rtcd_defs.pl:351 (hybrid_fwd_txfm.c:24) ---- */
void aom_hadamard_4x4_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_hadamard_8x8_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_hadamard_16x16_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_hadamard_32x32_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_hadamard_lp_8x8_b200(const int16_t *src_diff, ptrdiff_t src_stride, int16_t *coeff);
void aom_hadamard_lp_16x16_b200(const int16_t *src_diff, ptrdiff_t src_stride, int16_t *coeff);
void aom_hadamard_lp_8x8_dual_b200(const int16_t *src_diff, ptrdiff_t src_stride, int16_t *coeff);
void aom_highbd_hadamard_8x8_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_highbd_hadamard_16x16_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
void aom_highbd_hadamard_32x32_b200(const int16_t *src_diff, ptrdiff_t src_stride, b200_tran_low_t *coeff);
int aom_satd_b200(const b200_tran_low_t *coeff, int length);
int aom_satd_lp_b200(const int16_t *coeff, int length);
void av1_fwht4x4_b200(const int16_t *input, b200_tran_low_t *output, int stride);
unsigned int aom_avg_4x4_b200(const uint8_t *s, int p);
unsigned int aom_avg_8x8_b200(const uint8_t *s, int p);
void aom_avg_8x8_quad_b200(const uint8_t *s, int p, int x16_idx, int y16_idx, int *avg);
void aom_minmax_8x8_b200(const uint8_t *s, int p, const uint8_t *d, int dp, int *min, int *max);
unsigned int aom_highbd_avg_4x4_b200(const uint8_t *s8, int p);
unsigned int aom_highbd_avg_8x8_b200(const uint8_t *s8, int p);
void aom_highbd_minmax_8x8_b200(const uint8_t *s8, int p, const uint8_t *d8, int dp, int *min,
                                int *max);
void aom_int_pro_row_b200(int16_t *hbuf, const uint8_t *ref, const int ref_stride, const int width,
                          const int height, int norm_factor);
void aom_int_pro_col_b200(int16_t *vbuf, const uint8_t *ref, const int ref_stride, const int width,
                          const int height, int norm_factor);
int aom_vector_var_b200(const int16_t *ref, const int16_t *src, int bwl);

/* ---- compound / masked / OBMC members: aom_dsp_rtcd_defs.pl:884-889 (highbd
 *      _avg, dist_wtd), :1016-1032 (masked SAD), :1040-1060 (OBMC SAD), :1139-1141 (highbd x3d / skip
 *      x4d), :1303-1304 (aom_get8x8var / 16x16var), :1370-1420 (OBMC variance), :1440-1465 (avg /
 *      dist_wtd sub-pel variance), :1530-1560 (masked sub-pel variance, comp preds) ---- */
#ifdef B200AV1_RTCD_USE_AOM_TYPES
typedef DIST_WTD_COMP_PARAMS B200DistWtdCompParams;
#else
typedef struct { /* DIST_WTD_COMP_PARAMS, aom_dsp/aom_dsp_common.h (layout-compatible) */
  int use_dist_wtd_comp_avg;
  int fwd_offset;
  int bck_offset;
} B200DistWtdCompParams;
#endif

#define B200_DECL_HBD3(NAME, ARGS)               \
  unsigned int aom_highbd_8_##NAME##_b200 ARGS;  \
  unsigned int aom_highbd_10_##NAME##_b200 ARGS; \
  unsigned int aom_highbd_12_##NAME##_b200 ARGS;
#define B200_DECL_HBD_OBMC(NAME, ARGS)           \
  unsigned int aom_highbd_##NAME##_b200 ARGS;    \
  unsigned int aom_highbd_10_##NAME##_b200 ARGS; \
  unsigned int aom_highbd_12_##NAME##_b200 ARGS;
#define B200_MSVAR_ARGS                                                                                \
  (const uint8_t *src, int src_stride, int xoffset, int yoffset, const uint8_t *ref, int ref_stride,   \
   const uint8_t *second_pred, const uint8_t *msk, int msk_stride, int invert_mask, unsigned int *sse)
#define B200_SVAVAR_ARGS                                                                               \
  (const uint8_t *src_ptr, int source_stride, int xoffset, int yoffset, const uint8_t *ref_ptr,        \
   int ref_stride, uint32_t *sse, const uint8_t *second_pred)
#define B200_JSVAVAR_ARGS                                                                              \
  (const uint8_t *src_ptr, int source_stride, int xoffset, int yoffset, const uint8_t *ref_ptr,        \
   int ref_stride, uint32_t *sse, const uint8_t *second_pred, const B200DistWtdCompParams *jcp_param)
#define B200_OVAR_ARGS \
  (const uint8_t *pre, int pre_stride, const int32_t *wsrc, const int32_t *mask, unsigned int *sse)
#define B200_OSVAR_ARGS                                                                                \
  (const uint8_t *pre, int pre_stride, int xoffset, int yoffset, const int32_t *wsrc, const int32_t *mask, \
   unsigned int *sse)
#define B200_DECL_COMPOUND_FNS(W, H)                                                                          \
  unsigned int aom_highbd_sad##W##x##H##_avg_b200(const uint8_t *src_ptr, int src_stride,                      \
                                                  const uint8_t *ref_ptr, int ref_stride,                      \
                                                  const uint8_t *second_pred);                                 \
  void aom_highbd_sad##W##x##H##x3d_b200(const uint8_t *src_ptr, int src_stride,                               \
                                         const uint8_t *const ref_ptr[], int ref_stride, uint32_t *sad_array); \
  void aom_highbd_sad_skip_##W##x##H##x4d_b200(const uint8_t *src_ptr, int src_stride,                         \
                                               const uint8_t *const ref_ptr[], int ref_stride,                 \
                                               uint32_t *sad_array);                                           \
  unsigned int aom_masked_sad##W##x##H##_b200(const uint8_t *src, int src_stride, const uint8_t *ref,          \
                                              int ref_stride, const uint8_t *second_pred, const uint8_t *msk,  \
                                              int msk_stride, int invert_mask);                                \
  unsigned int aom_highbd_masked_sad##W##x##H##_b200(const uint8_t *src8, int src_stride, const uint8_t *ref8, \
                                                     int ref_stride, const uint8_t *second_pred8,              \
                                                     const uint8_t *msk, int msk_stride, int invert_mask);     \
  unsigned int aom_masked_sub_pixel_variance##W##x##H##_b200 B200_MSVAR_ARGS;                                  \
  B200_DECL_HBD3(masked_sub_pixel_variance##W##x##H, B200_MSVAR_ARGS)                                          \
  uint32_t aom_sub_pixel_avg_variance##W##x##H##_b200 B200_SVAVAR_ARGS;                                        \
  B200_DECL_HBD3(sub_pixel_avg_variance##W##x##H, B200_SVAVAR_ARGS)                                            \
  uint32_t aom_dist_wtd_sub_pixel_avg_variance##W##x##H##_b200 B200_JSVAVAR_ARGS;                              \
  B200_DECL_HBD3(dist_wtd_sub_pixel_avg_variance##W##x##H, B200_JSVAVAR_ARGS)                                  \
  unsigned int aom_dist_wtd_sad##W##x##H##_avg_b200(const uint8_t *src_ptr, int src_stride,                    \
                                                    const uint8_t *ref_ptr, int ref_stride,                    \
                                                    const uint8_t *second_pred,                                \
                                                    const B200DistWtdCompParams *jcp_param);                   \
  unsigned int aom_highbd_dist_wtd_sad##W##x##H##_avg_b200(const uint8_t *src_ptr, int src_stride,             \
                                                           const uint8_t *ref_ptr, int ref_stride,             \
                                                           const uint8_t *second_pred,                         \
                                                           const B200DistWtdCompParams *jcp_param);            \
  unsigned int aom_obmc_sad##W##x##H##_b200(const uint8_t *pre, int pre_stride, const int32_t *wsrc,           \
                                            const int32_t *mask);                                              \
  unsigned int aom_highbd_obmc_sad##W##x##H##_b200(const uint8_t *pre, int pre_stride, const int32_t *wsrc,    \
                                                   const int32_t *mask);                                       \
  unsigned int aom_obmc_variance##W##x##H##_b200 B200_OVAR_ARGS;                                               \
  unsigned int aom_obmc_sub_pixel_variance##W##x##H##_b200 B200_OSVAR_ARGS;                                    \
  B200_DECL_HBD_OBMC(obmc_variance##W##x##H, B200_OVAR_ARGS)                                                   \
  B200_DECL_HBD_OBMC(obmc_sub_pixel_variance##W##x##H, B200_OSVAR_ARGS)
B200_DECL_COMPOUND_FNS(128, 128)
B200_DECL_COMPOUND_FNS(128, 64)
B200_DECL_COMPOUND_FNS(64, 128)
B200_DECL_COMPOUND_FNS(64, 64)
B200_DECL_COMPOUND_FNS(64, 32)
B200_DECL_COMPOUND_FNS(32, 64)
B200_DECL_COMPOUND_FNS(32, 32)
B200_DECL_COMPOUND_FNS(32, 16)
B200_DECL_COMPOUND_FNS(16, 32)
B200_DECL_COMPOUND_FNS(16, 16)
B200_DECL_COMPOUND_FNS(16, 8)
B200_DECL_COMPOUND_FNS(8, 16)
B200_DECL_COMPOUND_FNS(8, 8)
B200_DECL_COMPOUND_FNS(8, 4)
B200_DECL_COMPOUND_FNS(4, 8)
B200_DECL_COMPOUND_FNS(4, 4)
B200_DECL_COMPOUND_FNS(4, 16)
B200_DECL_COMPOUND_FNS(16, 4)
B200_DECL_COMPOUND_FNS(8, 32)
B200_DECL_COMPOUND_FNS(32, 8)
B200_DECL_COMPOUND_FNS(16, 64)
B200_DECL_COMPOUND_FNS(64, 16)
void aom_comp_avg_pred_b200(uint8_t *comp_pred, const uint8_t *pred, int width, int height,
                            const uint8_t *ref, int ref_stride);
void aom_highbd_comp_avg_pred_b200(uint8_t *comp_pred8, const uint8_t *pred8, int width, int height,
                                   const uint8_t *ref8, int ref_stride);
void aom_dist_wtd_comp_avg_pred_b200(uint8_t *comp_pred, const uint8_t *pred, int width, int height,
                                     const uint8_t *ref, int ref_stride,
                                     const B200DistWtdCompParams *jcp_param);
void aom_highbd_dist_wtd_comp_avg_pred_b200(uint8_t *comp_pred8, const uint8_t *pred8, int width,
                                            int height, const uint8_t *ref8, int ref_stride,
                                            const B200DistWtdCompParams *jcp_param);
void aom_comp_mask_pred_b200(uint8_t *comp_pred, const uint8_t *pred, int width, int height,
                             const uint8_t *ref, int ref_stride, const uint8_t *mask, int mask_stride,
                             int invert_mask);
void aom_highbd_comp_mask_pred_b200(uint8_t *comp_pred8, const uint8_t *pred8, int width, int height,
                                    const uint8_t *ref8, int ref_stride, const uint8_t *mask,
                                    int mask_stride, int invert_mask);
void aom_get8x8var_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr, int ref_stride,
                        unsigned int *sse, int *sum);
void aom_get16x16var_b200(const uint8_t *src_ptr, int source_stride, const uint8_t *ref_ptr,
                          int ref_stride, unsigned int *sse, int *sum);
/* av1_round_shift_array (av1_rtcd_defs.pl:269) */
void av1_round_shift_array_b200(int32_t *arr, int size, int bit);

/* aom_variance_fn_ptr_t (aom_dsp/variance.h:92-113), member for member, and the filler that binds
 * every member of one block size to its `_b200` flavour -- what av1/encoder/encoder.c:1336-1600
 * (BFP / HIGHBD_BFP / MBFP / OBFP) does for the `_c` and SIMD flavours. */
typedef struct {
  unsigned int (*sdf)(const uint8_t *, int, const uint8_t *, int);
  unsigned int (*sdsf)(const uint8_t *, int, const uint8_t *, int);
  unsigned int (*sdaf)(const uint8_t *, int, const uint8_t *, int, const uint8_t *);
  unsigned int (*vf)(const uint8_t *, int, const uint8_t *, int, unsigned int *);
  unsigned int (*svf)(const uint8_t *, int, int, int, const uint8_t *, int, unsigned int *);
  unsigned int (*svaf)(const uint8_t *, int, int, int, const uint8_t *, int, unsigned int *, const uint8_t *);
  void (*sdx4df)(const uint8_t *, int, const uint8_t *const[], int, unsigned int *);
  void (*sdx3df)(const uint8_t *, int, const uint8_t *const[], int, unsigned int *);
  void (*sdsx4df)(const uint8_t *, int, const uint8_t *const[], int, unsigned int *);
  unsigned int (*msdf)(const uint8_t *, int, const uint8_t *, int, const uint8_t *, const uint8_t *, int, int);
  unsigned int (*msvf)(const uint8_t *, int, int, int, const uint8_t *, int, const uint8_t *, const uint8_t *,
                       int, int, unsigned int *);
  unsigned int (*osdf)(const uint8_t *, int, const int32_t *, const int32_t *);
  unsigned int (*ovf)(const uint8_t *, int, const int32_t *, const int32_t *, unsigned int *);
  unsigned int (*osvf)(const uint8_t *, int, int, int, const int32_t *, const int32_t *, unsigned int *);
  unsigned int (*jsdaf)(const uint8_t *, int, const uint8_t *, int, const uint8_t *,
                        const B200DistWtdCompParams *);
  unsigned int (*jsvaf)(const uint8_t *, int, int, int, const uint8_t *, int, unsigned int *, const uint8_t *,
                        const B200DistWtdCompParams *);
} B200VarianceFnPtr;
int b200_variance_fn_ptr(int block_width, int block_height, int highbd, int bd, B200VarianceFnPtr *vt);

#ifdef __cplusplus
}
#endif
#endif /* B200AV1_RTCD_H_ */
