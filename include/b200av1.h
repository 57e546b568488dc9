/*
 * b200av1.h -- C ABI of libb200av1.so: the B200 (sm_100a) implementation of the
 * aom-av1-lavish encoder's data-parallel hot path (forward transform + quantisation,
 * block error, subtract, SAD / variance, motion search).
 *
 * Plain C, plain pointers and sizes; no torch or CUDA types in any signature
 * (streams are passed as void*, device buffers as void* / typed device pointers).
 *
 * Three tiers:
 *   1. frame tier    -- context, device buffers, table upload            (this file)
 *   2. batched tier  -- descriptor arrays -> one launch per kernel class  (this file)
 *   3. drop-in tier  -- the reference's RTCD prototypes, one block per call:
 *                       include/b200av1_rtcd.h
 *
 * Every entry point cites the reference interface it replaces as
 * (path:line relative to the aom-av1-lavish tree).
 *
 * Error model: RTCD-shaped functions cannot return errors (reference DSP kernels are
 * void / value functions), so a CUDA failure there is fatal:
 * the message is stored for b200_last_error(), printed to stderr, and abort() is
 * called.  Tier 1/2 functions return 0 on success and a negative B200_E* code on
 * failure.  There is no CPU fallback anywhere in this library.
 */
#ifndef B200AV1_H_
#define B200AV1_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ECUDA (-1)   /* a CUDA runtime call failed; see b200_last_error() */
#define B200_EINVAL (-2)  /* bad argument (size/type combination, null pointer) */
#define B200_ENODEV (-3)  /* no CUDA device / device is not sm_100 */

/* TX_SIZE and TX_TYPE numbering: aom_dsp/txfm_common.h:25-71 */
enum {
  B200_TX_4X4, B200_TX_8X8, B200_TX_16X16, B200_TX_32X32, B200_TX_64X64,
  B200_TX_4X8, B200_TX_8X4, B200_TX_8X16, B200_TX_16X8, B200_TX_16X32,
  B200_TX_32X16, B200_TX_32X64, B200_TX_64X32, B200_TX_4X16, B200_TX_16X4,
  B200_TX_8X32, B200_TX_32X8, B200_TX_16X64, B200_TX_64X16, B200_TX_SIZES_ALL
};
enum {
  B200_DCT_DCT, B200_ADST_DCT, B200_DCT_ADST, B200_ADST_ADST, B200_FLIPADST_DCT,
  B200_DCT_FLIPADST, B200_FLIPADST_FLIPADST, B200_ADST_FLIPADST, B200_FLIPADST_ADST,
  B200_IDTX, B200_V_DCT, B200_H_DCT, B200_V_ADST, B200_H_ADST, B200_V_FLIPADST,
  B200_H_FLIPADST, B200_TX_TYPES
};

/* ------------------------------------------------------------------ frame tier */

typedef struct b200_ctx b200_ctx;

int b200_device_count(void);
/* Create a context on `device` (one per GPU; one process per GPU is the intended
 * deployment, several contexts per process also work).  Uploads the constant tables
 * (cospi/sinpi av1_txfm.c:18-69, scan orders scan.c:1666, tx configuration
 * av1_fwd_txfm2d.c:314-381).  Fails with B200_ENODEV when no sm_100 device exists. */
int b200_ctx_create(int device, b200_ctx **ctx);
void b200_ctx_destroy(b200_ctx *ctx);
/* Context used by the drop-in tier (created lazily on device $B200AV1_DEVICE or 0). */
b200_ctx *b200_default_ctx(void);
const char *b200_last_error(void);
/* The context's own stream as a cudaStream_t cast to void*. */
void *b200_ctx_stream(b200_ctx *ctx);
int b200_sync(b200_ctx *ctx);

/* Device memory helpers (thin wrappers so that C callers need no CUDA headers). */
int b200_malloc(b200_ctx *ctx, size_t bytes, void **dptr);
int b200_free(b200_ctx *ctx, void *dptr);
int b200_malloc_host(size_t bytes, void **hptr); /* pinned */
int b200_free_host(void *hptr);
int b200_h2d(b200_ctx *ctx, void *dst, const void *src, size_t bytes); /* async on ctx stream */
int b200_d2h(b200_ctx *ctx, void *dst, const void *src, size_t bytes); /* async on ctx stream */

/* Introspection used by tests: geometry helpers mirroring tx_size_wide/high
 * (av1/common/common_data.h) and av1_get_max_eob (av1/common/blockd.h:1596). */
int b200_tx_wide(int tx_size);
int b200_tx_high(int tx_size);
int b200_tx_max_eob(int tx_size);
int b200_tx_valid(int tx_size, int tx_type);
/* Host copy of the cosine table the kernels use, av1_cospi_arr(bit) (av1/common/av1_txfm.c:18),
 * bit = 10..13; NULL when the table failed its self-check against the reference's literal. */
const int32_t *b200_cospi(int bit);
/* Host copy of scan / iscan of av1_scan_orders[tx_size][tx_type] (scan.c:1666). */
const int16_t *b200_scan(int tx_size, int tx_type);
const int16_t *b200_iscan(int tx_size, int tx_type);

/* ---------------------------------------------------------------- batched tier */

/* One quantiser row: the [0]=DC / [1]=AC entries of the reference's per-qindex
 * tables (QUANTS / Dequants, av1/encoder/av1_quantize.h:45-84), as the encoder
 * keeps them in MACROBLOCK_PLANE (av1/encoder/block.h:129-167).  The encoder may pass its
 * own (av1_build_quantizer, av1_quantize.c:602) or have b200_build_quantizer_row make them. */
typedef struct {
  int16_t zbin[2], round[2], quant[2], quant_shift[2]; /* "B" quantiser */
  int16_t round_fp[2], quant_fp[2];                    /* "FP" quantiser */
  int16_t dequant[2];
  int16_t reserved[2];
} B200QRow; /* 32 bytes */
/* One row of av1_build_quantizer (av1_quantize.c:602-686): plane deltas as in
 * av1_dc_quant_QTX(qindex, dc_delta_q, bd) / av1_ac_quant_QTX(qindex, ac_delta_q, bd)
 * (quant_common.c:193-215; y: (y_dc_delta_q, 0), u: (u_dc_delta_q, u_ac_delta_q), ...);
 * quant_sharpness is this fork's extra argument.  Host only, no GPU needed. */
int b200_build_quantizer_row(int bd, int qindex, int dc_delta_q, int ac_delta_q,
                             int quant_sharpness, B200QRow *row);
int b200_dc_quant(int qindex, int delta, int bd); /* av1_dc_quant_QTX */
int b200_ac_quant(int qindex, int delta, int bd); /* av1_ac_quant_QTX */

/* One transform block of a batch. */
typedef struct {
  uint32_t src_offset; /* elements from the input base to the block's top-left sample */
  uint32_t out_offset; /* int32 elements into coeff / qcoeff / dqcoeff (BLOCK_OFFSET,
                          av1/common/blockd.h:489) */
  uint16_t src_stride; /* elements per input row (diff_stride / plane stride) */
  uint16_t qrow;       /* index into the B200QRow table */
  uint8_t tx_size;     /* B200_TX_* ; all blocks of one launch share it */
  uint8_t tx_type;     /* B200_*_* */
  uint16_t reserved;   /* b200_txq_launch: ignored.  b200_itx_launch: visible part of the block for
                          the distortion (get_txb_dimensions, av1/encoder/tx_search.c:940): 0 = whole
                          block, else visible_cols | visible_rows << 8 */
} B200TxBlock; /* 16 bytes */

/* Companion entry when the kernel forms the residual itself (fused
 * aom_subtract_block, aom_dsp/subtract.c:20): location of the prediction block. */
typedef struct {
  uint32_t pred_offset;
  uint16_t pred_stride;
  uint16_t reserved;
} B200TxPred; /* 8 bytes */

/* what the fused kernel reads */
enum {
  B200_IN_RESIDUAL = 0, /* int16 src_diff, as av1_fwd_txfm (hybrid_fwd_txfm.c:233) */
  B200_IN_PIXELS8 = 1,  /* uint8 source + prediction planes (fused subtract) */
  B200_IN_PIXELS16 = 2  /* uint16 source + prediction planes (high bit depth) */
};
/* which quantiser runs on the coefficients while they are still in registers */
enum {
  B200_Q_NONE = 0,      /* transform only */
  B200_Q_FP = 1,        /* av1_quantize_fp / _32x32 / _64x64 (av1_quantize.c:201-264) */
  B200_Q_FP_HBD = 2,    /* av1_highbd_quantize_fp (av1_quantize.c:565) */
  B200_Q_B = 3,         /* aom_quantize_b / _32x32 / _64x64 (aom_dsp/quantize.c:399-433) */
  B200_Q_B_HBD = 4      /* aom_highbd_quantize_b* (aom_dsp/quantize.c:435-470) */
};
#define B200_TXQ_STORE_COEFF 1u /* write the transform coefficients (p->coeff) */
#define B200_TXQ_BLOCK_ERROR 2u /* also av1_block_error / av1_highbd_block_error */
#define B200_TXQ_FORCE_EXACT 4u /* always use the literal 64-bit half_btf (arithmetic) */

typedef struct {
  int tx_size;             /* all n_blocks blocks have this size */
  int in_kind;             /* B200_IN_* */
  int q_kind;              /* B200_Q_* */
  int bd;                  /* 8, 10 or 12 (block-error rounding, rdopt.c:663) */
  unsigned flags;          /* B200_TXQ_* */
  int n_blocks;
  const B200TxBlock *blocks; /* device */
  const B200TxPred *preds;   /* device, or NULL for B200_IN_RESIDUAL */
  const void *src;           /* device: int16 residual or pixel plane base */
  const void *pred;          /* device: prediction plane base, or NULL */
  const B200QRow *qrows;     /* device */
  int32_t *coeff;            /* device, may be NULL without B200_TXQ_STORE_COEFF */
  int32_t *qcoeff;           /* device */
  int32_t *dqcoeff;          /* device */
  uint16_t *eob;             /* device, n_blocks entries */
  int64_t *dist;             /* device, 2*n_blocks entries {error, ssz}, or NULL */
} B200TxqLaunch;

/* Enqueue one fused (subtract +) transform (+ quantise) (+ block error) launch on
 * `stream` (void* cudaStream_t; NULL = the context's stream).  All pointers are
 * DEVICE pointers.  Replaces the per-block sequence av1_subtract_block ->
 * av1_xform -> av1_quant -> dist_block_tx_domain (encodemb.c:38,295,308;
 * tx_search.c:1077) for a whole batch of same-size blocks. */
int b200_txq_launch(b200_ctx *ctx, const B200TxqLaunch *launch, void *stream);

/* Host-buffer convenience wrapper over b200_txq_launch: copies descriptors, input and
 * tables to the device, launches once per tx_size present in `blocks`, copies the
 * results back and synchronises.  `src_elems` / `pred_elems` are the input buffer
 * lengths in elements, `out_elems` the length of each coefficient buffer. */
int b200_txq_batch_host(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                        const B200TxBlock *blocks, const B200TxPred *preds, int n_blocks,
                        const void *src, size_t src_elems, const void *pred,
                        size_t pred_elems, const B200QRow *qrows, int n_qrows,
                        int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff,
                        size_t out_elems, uint16_t *eob, int64_t *dist);


/* ------------------------------ inverse transform + reconstruction + distortion ---- */
/* av1_inverse_transform_block (av1/common/idct.c:303; av1_inv_txfm_add_c
 * :283, av1_highbd_inv_txfm_add_c :216, av1_inv_txfm2d_add_WxH_c av1/common/av1_inv_txfm2d.c:234-460)
 * followed by dist_block_px_domain (av1/encoder/tx_search.c:969) for a batch of same-size blocks, so
 * that a tx-type search evaluated on the device returns distortions instead of coefficients.
 * The launch reads the SAME descriptor arrays as the b200_txq_launch that produced dqcoeff / eob:
 * blocks[i].out_offset locates the dequantised coefficients, blocks[i].src_offset / src_stride the
 * source block, preds[i] the prediction block; recon (optional) is a plane with the prediction's
 * geometry (it may be the prediction plane itself: in-place, like the reference's dst buffer);
 * dist[i] (optional) = 16 * SSE(source, reconstruction) over the visible part of the block, rounded
 * by 2 (bd - 8) bits for uint16 pixels as the high-bit-depth variance functions do.  A block whose
 * eob is 0 reconstructs to its prediction (idct.c:307).  uint8 pixels (highbd = 0) imply bd = 8. */
typedef struct {
  int tx_size;               /* all n_blocks blocks have this size */
  int highbd;                /* 0: uint8 pixel planes, 1: uint16 */
  int bd;                    /* 8, 10 or 12 */
  int n_blocks;
  const B200TxBlock *blocks; /* device */
  const B200TxPred *preds;   /* device */
  const int32_t *dqcoeff;    /* device: as b200_txq_launch wrote it (column-major kept region) */
  const uint16_t *eob;       /* device, n_blocks entries, or NULL (every block is inverse transformed) */
  const void *src;           /* device source plane (needed for dist) */
  const void *pred;          /* device prediction plane */
  void *recon;               /* device, or NULL */
  int64_t *dist;             /* device, n_blocks entries, or NULL */
  int lossless;              /* 1: tx_size is B200_TX_4X4 and the blocks take the lossless inverse WHT,
                                av1_highbd_iwht4x4_16_add / _1_add by eob (av1_inv_txfm2d.c:20-109;
                                dqcoeff in the WHT's own row-major order, hybrid_fwd_txfm.c:24) */
} B200ItxLaunch;
int b200_itx_launch(b200_ctx *ctx, const B200ItxLaunch *launch, void *stream);

/* ------------------------------------- transform-type search on the device ------- */
/* av1_cost_coeffs_txb_estimate (av1/encoder/txb_rdopt.c:547-582): the coefficient term of the
 * Laplacian rate estimate av1_cost_coeffs_txb_laplacian (:626) that prune_txk_type /
 * prune_txk_type_separ rank tx types with (tx_search.c:1179,1215,1302), for every block of a batch
 * from the dense qcoeff a b200_txq_launch left on the device; rate[i] = 0 when eob[i] = 0 (the
 * reference then takes txb_skip_cost instead, txb_rdopt.c:646). */
typedef struct {
  int n_blocks;
  const B200TxBlock *blocks; /* device: tx_size, tx_type, out_offset are read */
  const int32_t *qcoeff;     /* device */
  const uint16_t *eob;       /* device */
  int32_t *rate;             /* device out, n_blocks */
} B200RateEstLaunch;
int b200_rate_estimate_launch(b200_ctx *ctx, const B200RateEstLaunch *launch, void *stream);

/* One descriptor per (block, candidate tx type) of search_tx_type (av1/encoder/tx_search.c:2047-2330):
 * av1_xform_quant (encodemb.c:295) + dist_block_tx_domain (tx_search.c:1077) + the rate estimate
 * above + dist_block_px_domain (tx_search.c:969) run chained on the device; 32 bytes per descriptor
 * come back instead of the coefficient arrays.  src / pred are HOST pixel planes (uint8 for
 * B200_IN_PIXELS8, uint16 for B200_IN_PIXELS16); blocks[i].out_offset is ignored; the encoder adds
 * the context-dependent rate terms (txb_skip / tx-type / eob costs) from eob and tx_type. */
#define B200_TXS_DIST_PX 1u  /* fill dist_px  (inverse transform + reconstruction + SSE) */
#define B200_TXS_DIST_TX 2u  /* fill dist_tx / ssz (block error in the transform domain) */
#define B200_TXS_RATE_EST 4u /* fill rate_est */
typedef struct {
  int64_t dist_px;  /* 16 * SSE(source, reconstruction), rounded for bd > 8 (dist_block_px_domain) */
  int64_t dist_tx;  /* av1_block_error / av1_highbd_block_error return value (before the caller's shift) */
  int64_t ssz;      /* its *ssz */
  int32_t rate_est; /* av1_cost_coeffs_txb_estimate */
  uint16_t eob;
  uint16_t reserved;
} B200TxSearchResult; /* 32 bytes */
int b200_txsearch_batch_host(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned want,
                             const B200TxBlock *blocks, const B200TxPred *preds, int n_blocks,
                             const void *src, size_t src_elems, const void *pred, size_t pred_elems,
                             const B200QRow *qrows, int n_qrows, B200TxSearchResult *out,
                             size_t *d2h_bytes);

/* ------------------------------------------------ compact result format ---------- */
/* What the encoder consumes after the quantiser is the levels in SCAN order up to eob
 * (av1_write_coeffs_txb / av1_cost_coeffs_txb, av1/encoder/encodetxb.c) and dqcoeff is a
 * function of the level ((|q| * dequant[rc != 0]) >> log_scale, av1_quantize.c:61-63), so the
 * dense int32 coeff + qcoeff + dqcoeff arrays (12 bytes per sample) need not cross PCIe.
 * Compact form of a batch: block i owns eob[i] consecutive levels at level_offset[i]
 * (exclusive prefix sum of eob in batch order; level_offset[n_blocks] = total), int16 for
 * the low-bit-depth quantisers (levels are clamped to int16 there, av1_quantize.c:48-52),
 * int32 (`wide`) for the high-bit-depth ones. */
typedef struct {
  int n_blocks;
  const B200TxBlock *blocks; /* device: tx_size, tx_type and out_offset are read */
  const int32_t *qcoeff;     /* device: dense qcoeff as b200_txq_launch wrote it */
  const uint16_t *eob;       /* device, n_blocks */
  int wide;                  /* 0: int16 levels, 1: int32 levels */
  void *levels;              /* device out: packed levels (capacity >= sum of eob) */
  uint32_t *level_offset;    /* device out: n_blocks + 1 entries */
} B200CompactLaunch;
int b200_txq_compact_launch(b200_ctx *ctx, const B200CompactLaunch *launch, void *stream);
/* The same packing for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Its
 * scratch (the per-chunk sums of the prefix sum) is a file-scope device array of
 * B200_COMPACT_NOALLOC_MAX_BLOCKS / 2048 * 4 bytes = 16 KB, resident on every GPU that loads the library;
 * the call itself performs no cudaMalloc / cudaFree and no allocating library call (the context must exist
 * already).  n_blocks beyond B200_COMPACT_NOALLOC_MAX_BLOCKS is refused with B200_EINVAL and a message in
 * b200_last_error().  Every call on a device shares the array (it keeps the last call's sums; nothing in it
 * needs to start at zero): calls whose kernels overlap on the GPU race -- order them on one stream or with
 * events.  Results are identical to b200_txq_compact_launch. */
#define B200_COMPACT_NOALLOC_MAX_BLOCKS (1 << 23) /* 8 388 608 transform blocks per call */
int b200_txq_compact_launch_noalloc(b200_ctx *ctx, const B200CompactLaunch *launch, void *stream);

/* Host-buffer call returning the compact form: like b200_txq_batch_host, but the dense
 * qcoeff / dqcoeff stay on the device; the host receives eob, level_offset (n_blocks + 1),
 * the packed levels (at most levels_cap elements of 2 or 4 bytes: `wide` = q_kind is a
 * high-bit-depth quantiser), optionally {error, ssz} per block (B200_TXQ_BLOCK_ERROR) and
 * optionally the dense coeff (B200_TXQ_STORE_COEFF; trellis wants it, TPL's
 * txfm_quant_rdcost, tpl_model.c:228, does not).  *d2h_bytes (optional) receives the bytes
 * that crossed PCIe towards the host.  Fails with B200_EINVAL when levels_cap is too small. */
int b200_txq_batch_host_compact(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                                const B200TxBlock *blocks, const B200TxPred *preds, int n_blocks,
                                const void *src, size_t src_elems, const void *pred,
                                size_t pred_elems, const B200QRow *qrows, int n_qrows,
                                size_t out_elems, void *levels, size_t levels_cap,
                                uint32_t *level_offset, uint16_t *eob, int64_t *dist,
                                int32_t *coeff, size_t *d2h_bytes);

/* Host-side expander of one block of the compact form into the dense arrays the quantisers
 * (without qmatrix) write: qcoeff[scan[i]] = level i, dqcoeff from the row's dequant.  Either
 * output may be NULL. */
int b200_txq_expand_block(int tx_size, int tx_type, const void *levels, int wide, int eob,
                          const B200QRow *row, int32_t *qcoeff, int32_t *dqcoeff);

/* ------------------------------------------------ stand-alone quantiser batch ---- */
/* The quantisers as separate launches over coefficient buffers that already sit in
 * device memory (e.g. re-quantising one transform at several qindexes).  One warp per
 * block.  Semantics per kind: the reference function named beside it. */
enum {
  B200_QUANT_FP = 0,          /* av1_quantize_fp* (av1_quantize.c:70-123,201-264), qm optional */
  B200_QUANT_FP_HBD = 1,      /* av1_highbd_quantize_fp (av1_quantize.c:125-198,565) */
  B200_QUANT_LP = 2,          /* av1_quantize_lp (av1_quantize.c:212), int16 in/out */
  B200_QUANT_B = 3,           /* aom_quantize_b_helper (aom_dsp/quantize.c:108) */
  B200_QUANT_B_HBD = 4,       /* aom_highbd_quantize_b_helper (aom_dsp/quantize.c:261) */
  B200_QUANT_B_ADAPT = 5,     /* aom_quantize_b_adaptive_helper (aom_dsp/quantize.c:16) */
  B200_QUANT_B_ADAPT_HBD = 6  /* aom_highbd_quantize_b_adaptive_helper (quantize.c:173) */
};
typedef struct {
  uint32_t coeff_offset; /* elements into coeff (int32; int16 for B200_QUANT_LP) */
  uint32_t out_offset;   /* elements into qcoeff / dqcoeff */
  uint32_t iscan_offset; /* int16 elements into iscan */
  uint32_t qm_offset;    /* elements into qm / iqm (ignored when those are NULL) */
  uint16_t n_coeffs;     /* 16 .. 1024 */
  uint16_t qrow;
  uint8_t log_scale;     /* 0, 1, 2 (av1_get_tx_scale, av1/common/idct.c:24) */
  uint8_t reserved[3];
} B200QuantBlock; /* 24 bytes */
typedef struct {
  int kind; /* B200_QUANT_* */
  int n_blocks;
  const B200QuantBlock *blocks; /* device */
  const void *coeff;            /* device */
  const B200QRow *qrows;        /* device */
  const int16_t *iscan;         /* device */
  const uint8_t *qm, *iqm;      /* device or NULL (qm_val_t weights, AOM_QM_BITS = 5) */
  void *qcoeff, *dqcoeff;       /* device */
  uint16_t *eob;                /* device */
} B200QuantLaunch;
int b200_quant_launch(b200_ctx *ctx, const B200QuantLaunch *launch, void *stream);

/* ------------------------------------------------------ block error batch -------- */
enum {
  B200_ERR_LOWBD = 0, /* av1_block_error (rdopt.c:635): squares formed in 32-bit int */
  B200_ERR_LP = 1,    /* av1_block_error_lp (rdopt.c:650): int16 inputs */
  B200_ERR_HBD = 2,   /* av1_highbd_block_error (rdopt.c:663): rounded >> 2(bd-8) */
  B200_ERR_QM = 3     /* av1_block_error_qm (tx_search.c:1049, fork): qmatrix weighted */
};
typedef struct {
  uint32_t coeff_offset, dqcoeff_offset;
  uint32_t n_coeffs;
  uint32_t aux_offset; /* B200_ERR_QM: element offset into qm and scan */
} B200ErrBlock; /* 16 bytes */
typedef struct {
  int kind, bd, n_blocks;
  const B200ErrBlock *blocks; /* device */
  const void *coeff, *dqcoeff; /* device: int32 (int16 for B200_ERR_LP) */
  const uint8_t *qm;           /* device, B200_ERR_QM only */
  const int16_t *scan;         /* device, B200_ERR_QM only */
  int64_t *out;                /* device: {error, ssz} per block */
} B200ErrLaunch;
int b200_block_error_launch(b200_ctx *ctx, const B200ErrLaunch *launch, void *stream);

/* ------------------------------------------- pixel block batch (SAD / variance) -- */
/* One warp per (a, b) block pair.  `a` follows the first pointer argument of the
 * reference function, `b` the second (SAD: a = src, b = ref; variance: a = src_ptr,
 * b = ref_ptr; sub-pel variance: a is the interpolated operand). */
enum {
  B200_PIX_SAD = 0,      /* aom_sadWxH (aom_dsp/sad.c:45)                    out0 = sad */
  B200_PIX_SAD_SKIP = 1, /* aom_sad_skip_WxH (sad.c:65)                      out0 = sad */
  B200_PIX_SAD_AVG = 2,  /* aom_sadWxH_avg (sad.c:50), aux = second_pred     out0 = sad */
  B200_PIX_VAR = 3,      /* aom_varianceWxH (variance.c:123)          out0 = var, out1 = sse */
  B200_PIX_MSE = 4,      /* aom_mseWxH (variance.c:235)               out0 = sse, out1 = sse */
  B200_PIX_SUBPEL_VAR = 5, /* aom_sub_pixel_varianceWxH (variance.c:132)   out0 = var, out1 = sse */
  B200_PIX_SSE = 6,      /* aom_sse (aom_dsp/sse.c:19)         out0 = low 32, out1 = high 32 */
  B200_PIX_VAR_SUM = 7   /* aom_get8x8var / aom_get16x16var (variance.c GET_VAR)  out0 = (int) sum, out1 = sse */
};
typedef struct {
  uint32_t a_offset, b_offset; /* pixels from the plane bases */
  uint32_t aux_offset;         /* second_pred offset (B200_PIX_SAD_AVG) */
  uint16_t a_stride, b_stride;
  uint8_t w, h;                /* 4 .. 128 */
  uint8_t xoff, yoff;          /* 1/8-pel bilinear offsets 0..7 (B200_PIX_SUBPEL_VAR) */
} B200PixBlock; /* 20 bytes */
typedef struct {
  int op;      /* B200_PIX_* */
  int highbd;  /* 0: uint8 pixels, 1: uint16 pixels */
  int bd;      /* 8 / 10 / 12: variance rounding of aom_highbd_{8,10,12}_variance */
  int n_blocks;
  const B200PixBlock *blocks; /* device */
  const void *a, *b, *aux;    /* device pixel planes */
  uint32_t *out;              /* device: 2 entries per block */
} B200PixLaunch;
int b200_pix_launch(b200_ctx *ctx, const B200PixLaunch *launch, void *stream);

/* ------------------ compound / masked / OBMC SAD and variance ---- */
/* The aom_variance_fn_ptr_t members (aom_dsp/variance.h:92-113) beyond sdf / vf / svf, one warp per
 * block.  Operands: `pre` = the predictor plane at a_offset / a_stride (the reference functions' `ref`
 * in the SAD family, the interpolated first argument `a` / `src` / `pre` in the variance family,
 * `ref` in the builders); `tgt` = the plane compared against at b_offset / b_stride (`src` in the SAD
 * family, `b` / `ref` in the variance family); second_pred, wsrc and obmc_mask are w-pitch blocks at
 * aux_offset; the builders write a w-pitch block to comp_pred + b_offset.  out0 = SAD or variance,
 * out1 = sse (variance ops). */
enum {
  B200_CMP_MASKED_SAD = 0,          /* aom_masked_sadWxH / aom_highbd_masked_sadWxH (sad_av1.c:42,122) */
  B200_CMP_MASKED_SUBPEL_VAR,       /* aom_masked_sub_pixel_varianceWxH, highbd 8/10/12 (variance.c:769,838) */
  B200_CMP_SUBPEL_AVG_VAR,          /* aom_sub_pixel_avg_varianceWxH (variance.c:147), highbd (:551-600) */
  B200_CMP_DIST_WTD_SUBPEL_AVG_VAR, /* aom_dist_wtd_sub_pixel_avg_varianceWxH (variance.c:165), highbd (:603-665) */
  B200_CMP_DIST_WTD_SAD_AVG,        /* aom_dist_wtd_sadWxH_avg (sad.c:57), highbd (:290) */
  B200_CMP_OBMC_SAD,                /* aom_obmc_sadWxH / highbd (sad_av1.c:163-262) */
  B200_CMP_OBMC_VAR,                /* aom_obmc_varianceWxH, highbd 8/10/12 (variance.c:933-960,1040-1125) */
  B200_CMP_OBMC_SUBPEL_VAR,         /* aom_obmc_sub_pixel_varianceWxH, highbd (variance.c:962,1126) */
  B200_CMP_AVG_PRED,                /* aom_comp_avg_pred / aom_highbd_comp_avg_pred (variance.c:285,708) */
  B200_CMP_DIST_WTD_AVG_PRED,       /* aom_dist_wtd_comp_avg_pred / highbd (variance.c:300,725) */
  B200_CMP_MASK_PRED                /* aom_comp_mask_pred / aom_highbd_comp_mask_pred (variance.c:749,816) */
};
typedef struct {
  uint32_t mask_offset; /* bytes from `mask` to the block's first weight (0..64, AOM_BLEND_A64) */
  uint16_t mask_stride;
  uint8_t invert_mask;
  uint8_t reserved;
} B200PixMask; /* 8 bytes */
typedef struct {
  int op;      /* B200_CMP_* */
  int highbd;  /* 0: uint8 pixels, 1: uint16 */
  int bd;      /* 8 / 10 / 12 */
  int n_blocks;
  const B200PixBlock *blocks;   /* device; xoff / yoff = 1/8-pel bilinear offsets of the *_SUBPEL_* ops */
  const B200PixMask *masks;     /* device, masked ops */
  const void *pre, *tgt;        /* device pixel planes */
  const void *second_pred;      /* device */
  const uint8_t *mask;          /* device */
  const int32_t *wsrc, *obmc_mask; /* device, OBMC ops (calc_target_weighted_pred, av1/encoder/rdopt.c) */
  int fwd_offset, bck_offset;   /* DIST_WTD_COMP_PARAMS (av1_dist_wtd_comp_weight_assign) */
  uint32_t *out;                /* device, 2 per block */
  void *comp_pred;              /* device, builders */
} B200CompoundLaunch;
int b200_compound_launch(b200_ctx *ctx, const B200CompoundLaunch *launch, void *stream);

/* aom_subtract_block / aom_highbd_subtract_block (aom_dsp/subtract.c:20,35) over a batch:
 * diff = a - b for every block (a = src, b = pred); diff rows are `w` apart. */
typedef struct {
  int highbd, n_blocks;
  const B200PixBlock *blocks; /* aux_offset = element offset of the block in diff */
  const void *a, *b;
  int16_t *diff;
} B200SubtractLaunch;
int b200_subtract_launch(b200_ctx *ctx, const B200SubtractLaunch *launch, void *stream);

/* aom_upsampled_pred / aom_highbd_upsampled_pred (av1/encoder/reconinter_enc.c:424,562),
 * unscaled references: separable sub-pel interpolation with 2 / 4 / 8 taps; one block
 * per descriptor: a = reference plane, aux_offset = element offset in comp_pred
 * (w*h contiguous), xoff / yoff = subpel_x_q3 / subpel_y_q3. */
typedef struct {
  int highbd, bd, taps, n_blocks;
  const B200PixBlock *blocks;
  const void *ref;
  void *comp_pred;
} B200UpsampleLaunch;
int b200_upsampled_pred_launch(b200_ctx *ctx, const B200UpsampleLaunch *launch, void *stream);


/* ------------------------------------------- Hadamard / SATD / block averages ------ */
/* aom_dsp/avg.c: the Hadamard transforms tx-search pruning and TPL use for their SATD cost
 * (tpl_model.c:97-130, tx_search.c prune paths), plus the lossless WHT (hybrid_fwd_txfm.c:24).
 * One warp per block (a lane per block at 4x4); coeff and satd are both optional so that the
 * SATD-only use (aom_hadamard_NxN followed by aom_satd, the common case) writes 4 bytes per
 * block instead of 4 bytes per sample. */
enum {
  B200_HAD_LOWBD = 0,  /* aom_hadamard_{4x4,8x8,16x16,32x32}_c (avg.c:115,177,247,318): int32 out */
  B200_HAD_LP = 1,     /* aom_hadamard_lp_{8x8,16x16}_c (avg.c:207,289): int16 out */
  B200_HAD_HIGHBD = 2, /* aom_highbd_hadamard_{8x8,16x16,32x32}_c (avg.c:416,442,474): int32 out */
  B200_HAD_WHT = 3     /* av1_fwht4x4_c (hybrid_fwd_txfm.c:24), n = 4: int32 out */
};
typedef struct {
  uint32_t src_offset; /* int16 elements from `src` to the block's first residual */
  uint32_t out_offset; /* elements from `coeff` to the block's n*n coefficients */
  int32_t stride;      /* residual row stride, elements */
} B200HadBlock; /* 12 bytes */
typedef struct {
  int n;    /* 4, 8, 16 or 32 */
  int kind; /* B200_HAD_* */
  int n_blocks;
  const B200HadBlock *blocks; /* device */
  const int16_t *src;         /* device */
  void *coeff;                /* device, int32 (int16 for B200_HAD_LP), or NULL */
  int32_t *satd;              /* device, aom_satd / aom_satd_lp of block i, or NULL */
} B200HadamardLaunch;
int b200_hadamard_launch(b200_ctx *ctx, const B200HadamardLaunch *launch, void *stream);

enum {
  B200_AVG_4X4 = 0,    /* aom_avg_4x4 / aom_highbd_avg_4x4 (avg.c:32,74)        out0 = average */
  B200_AVG_8X8 = 1,    /* aom_avg_8x8 / aom_highbd_avg_8x8 (avg.c:42,63)        out0 = average */
  B200_MINMAX_8X8 = 2  /* aom_minmax_8x8 / highbd (avg.c:18,85), b = second plane   out0 = min, out1 = max */
};
typedef struct {
  uint32_t a_offset, b_offset; /* pixels from the plane bases */
  int32_t a_stride, b_stride;  /* pixels */
} B200AvgBlock; /* 16 bytes */
typedef struct {
  int op, highbd, n_blocks;
  const B200AvgBlock *blocks; /* device */
  const void *a, *b;          /* device planes (b only for B200_MINMAX_8X8) */
  int32_t *out;               /* device, 2 ints per block */
} B200AvgLaunch;
int b200_avg_launch(b200_ctx *ctx, const B200AvgLaunch *launch, void *stream);

/* ------------------------------------------------------ motion search batch ------ */
/* One descriptor = one single-reference translational search of one block: the fields
 * of FULLPEL_MOTION_SEARCH_PARAMS / SUBPEL_MOTION_SEARCH_PARAMS (av1/encoder/mcomp.h:
 * 101-141, 301-324) that av1_full_pixel_search (mcomp.c:1755) and
 * av1_find_best_sub_pixel_tree (mcomp.c:3128) read.  One CTA per descriptor: the
 * candidates of every search step are evaluated in parallel (SAD / variance per warp),
 * the decision rule is the reference's "first candidate in scan order with the strictly
 * smallest sad + mv cost", second_best_mv and the 5-point cost list included. */
enum { B200_ME_NSTEP = 0, B200_ME_NSTEP_8PT = 1, B200_ME_DIAMOND = 2, B200_ME_CLAMPED_DIAMOND = 3,
       /* pattern_search family (mcomp.c:1017-1316); values of the reference's SEARCH_METHODS */
       B200_ME_HEX = 4, B200_ME_BIGDIA = 5, B200_ME_SQUARE = 6, B200_ME_FAST_HEX = 7,
       B200_ME_FAST_DIAMOND = 8, B200_ME_FAST_BIGDIA = 9, B200_ME_VFAST_DIAMOND = 10,
       /* not a SEARCH_METHODS value: av1_refining_search_8p_c (mcomp.c:1683, single reference)
        * instead of av1_full_pixel_search; result = best mv + returned sad-with-cost */
       B200_ME_REFINE_8P = 11 };
/* b200_me_subpel_launch reads search_method as the sub-pel function: av1_find_best_sub_pixel_tree
 * (mcomp.c:3128), _pruned (:2992), _pruned_more (:2907) */
enum { B200_SUBPEL_TREE = 0, B200_SUBPEL_TREE_PRUNED = 1, B200_SUBPEL_TREE_PRUNED_MORE = 2 };
enum { B200_MV_COST_ENTROPY = 0, B200_MV_COST_L1_LOWRES, B200_MV_COST_L1_MIDRES,
       B200_MV_COST_L1_HDRES, B200_MV_COST_NONE };           /* MV_COST_TYPE, mcomp.h:32-38 */
enum { B200_USE_2_TAPS_ORIG = 0, B200_USE_2_TAPS, B200_USE_4_TAPS, B200_USE_8_TAPS };
enum { B200_EIGHTH_PEL = 0, B200_QUARTER_PEL, B200_HALF_PEL, B200_FULL_PEL };
#define B200_ME_RUN_MESH 1u      /* run_mesh_search */
#define B200_ME_PRUNE_MESH 2u    /* prune_mesh_search */
#define B200_ME_FINE_INTERVAL 4u /* fine_search_interval */
#define B200_ME_SKIP_SAD 8u      /* ms_params->sdf = vfp->sdsf (row-skipping SAD) */
#define B200_ME_ALLOW_HP 16u     /* sub-pel: allow_hp */
#define B200_ME_COST_LIST 32u    /* the caller passes cost_list != NULL (pattern_search takes a
                                    different path through its last scale when it does; the pruned
                                    sub-pel trees read it from B200MeLaunch.cost_lists) */
typedef struct {
  uint32_t src_offset; /* pixels from the current-frame base to the block */
  uint32_t ref_offset; /* pixels from the reference-frame base to the co-located block */
  int16_t col_min, col_max, row_min, row_max; /* FullMvLimits (full-pel search) or
                                                 SubpelMvLimits in 1/8 pel (sub-pel search) */
  int16_t start_row, start_col;               /* full-pel (1/8 pel for the sub-pel search) */
  int16_t ref_mv_row, ref_mv_col;             /* 1/8 pel */
  uint8_t bw, bh;                             /* one of the 22 SAD block sizes */
  uint8_t search_method, step_param;
  uint8_t flags;                              /* B200_ME_* */
  uint8_t mv_cost_type;
  uint8_t mesh_mv_diff_threshold;
  uint8_t subpel;                             /* forced_stop | iters_per_step << 2 | search_type << 4 */
  int32_t force_mesh_thresh;
  int32_t sad_per_bit, error_per_bit;
} B200MeBlock; /* 44 bytes */
typedef struct {
  int16_t best_row, best_col;
  int16_t second_row, second_col; /* -32768 (INVALID_MV) when the search left it unset */
  int32_t cost_list[5];           /* full-pel: calc_int_sad_list (mcomp.c:791) */
  int32_t ret;                    /* the reference function's return value */
  int32_t distortion;             /* sub-pel: *distortion */
  uint32_t sse;                   /* sub-pel: *sse1 */
} B200MeResult; /* 40 bytes */
typedef struct {
  int highbd, bd;                /* 0: uint8 planes; 1: uint16 planes of bd bits */
  int n_blocks;
  const B200MeBlock *blocks;     /* device */
  const void *cur, *ref;         /* device planes (the reference plane includes its border) */
  int cur_stride, ref_stride;    /* pixels */
  int mesh_range[4], mesh_interval[4]; /* MESH_PATTERN[MAX_MESH_STEP], speed_features.h:27-32 */
  const int *mvjcost;            /* device, 4 entries (MV_COST_ENTROPY), or NULL */
  const int *mvcost[2];          /* device, centred pointers: index -MV_MAX..MV_MAX */
  B200MeResult *results;         /* device */
  int window_bytes;              /* shared memory reserved per search for the reference window
                                    (limits box + block, e.g. 192*192 = 36864 for a 64x64 block
                                    with +-64); searches whose window does not fit read the
                                    reference through L1/L2 instead.  0 = never stage. */
  int max_bw, max_bh;            /* largest block of the batch (sizes the per-search shared
                                    memory); 0 = assume 128 */
  const int32_t *cost_lists;     /* device; sub-pel pruned trees: ms_params->cost_list of block i at
                                    cost_lists[i * cost_list_stride .. +5], read when the block has
                                    B200_ME_COST_LIST.  Pointing it at the full-pel launch's
                                    results[0].cost_list with stride 10 chains the two launches. */
  int cost_list_stride;          /* in int32 units; 0 = 5 */
  /* ---- optional hints (0 = unknown); a zero-initialised struct behaves as before ---- */
  unsigned families;             /* full-pel: B200_ME_FAMILY_* bits of the search methods present in
                                    the batch; one kernel instance runs per family (0: all three
                                    are launched and the CTAs of absent families exit at once) */
  int cur_rows, ref_rows;        /* rows of the cur / ref plane buffers.  With both set, 16-byte
                                    aligned bases and strides that are multiples of 16 bytes, the
                                    source block and the search window are brought into shared
                                    memory by the TMA engine (cp.async.bulk.tensor.2d) instead of
                                    by the threads */
  int max_col_span, max_row_span; /* largest col_max - col_min / row_max - row_min of the batch
                                    (full-pel units): sizes the TMA box of the search window */
  const B200MeResult *start_from; /* sub-pel launch, device: when set, block i starts at 8 x the
                                    full-pel result start_from[i].best_{row,col} instead of its
                                    descriptor's start mv (chains b200_me_fullpel_launch into
                                    b200_me_subpel_launch without a host round trip, as
                                    av1_single_motion_search does, motion_search_facade.c:120-494) */
  unsigned long long *stats;     /* full-pel launch, device, optional: [0] += SAD candidates evaluated,
                                    [1] += pixels compared (measurement only) */
  int start_group;               /* with start_from: block i reads start_from[i / start_group] (0 = 1).
                                    In a FULL-PEL launch start_from holds SUB-pel results (1/8 pel) and
                                    the block starts at get_fullmv_from_mv() of it (mv.h:79) -- the
                                    4 sub-blocks of tf_motion_search start at their parent block's
                                    refined mv (temporal_filter.c:198), start_group = 4 */
  /* ---- compound search (av1_set_ms_compound_refs, av1/encoder/mcomp.h:86): with `compound` set, block i
   * is searched against blend(reference, second_pred) -- the average (vfp->sdaf / svaf,
   * aom_comp_avg_upsampled_pred) or, with a mask, AOM_BLEND_A64 (vfp->msdf / msvf,
   * aom_comp_mask_upsampled_pred) -- as get_mvpred_compound_sad (mcomp.c:729), setup_center_error (:2781),
   * estimated_pref_error (:2371) and upsampled_pref_error (:2402) do.  Served by B200_ME_REFINE_8P
   * (b200_me_fullpel_launch) and the three sub-pel trees (b200_me_subpel_launch): the two searches of one
   * iteration of av1_joint_motion_search (motion_search_facade.c:594-652). ---- */
  const struct B200MeCompound *compound; /* device, n_blocks entries, or NULL: single reference */
  const void *second_pred;               /* device: bw-pitch prediction blocks from the other reference */
  const uint8_t *mask;                   /* device: 6-bit blend weights (0..64), or NULL */
} B200MeLaunch;
typedef struct B200MeCompound {
  uint32_t second_offset; /* pixels from second_pred to the block's bw x bh prediction */
  uint32_t mask_offset;   /* bytes from mask to the block's first weight */
  uint16_t mask_stride;
  uint8_t invert_mask;    /* inv_mask: the weights apply to second_pred instead of the reference */
  uint8_t use_mask;       /* 0: average, 1: masked blend */
} B200MeCompound; /* 12 bytes */
#define B200_ME_FAMILY_NSTEP 1u   /* NSTEP, NSTEP_8PT, DIAMOND, CLAMPED_DIAMOND */
#define B200_ME_FAMILY_PATTERN 2u /* HEX .. VFAST_DIAMOND (pattern_search) */
#define B200_ME_FAMILY_REFINE 4u  /* B200_ME_REFINE_8P */
/* av1_init_search_range (mcomp.c:239): step_param for a search of +-size pixels */
int b200_me_init_search_range(int size);
int b200_me_fullpel_launch(b200_ctx *ctx, const B200MeLaunch *launch, void *stream);
int b200_me_subpel_launch(b200_ctx *ctx, const B200MeLaunch *launch, void *stream);
/* Host-buffer motion search of one batch (what `e2e` times for the search half): `launch`
 * describes the batch with HOST pointers in blocks / cur / ref / mvjcost / mvcost (cost tables
 * optional; mvcost pointers are centred, each table spans -MV_MAX..MV_MAX = 2 * 16383 + 1 ints);
 * cur_bytes / ref_bytes are the plane buffer sizes.  Copies everything to the device, runs the
 * full-pel launch and, when subpel_blocks is given, the sub-pel launch chained on the full-pel
 * result (start_from), copies fullpel_out / subpel_out (n_blocks each, either may be NULL)
 * back and synchronises. */
int b200_me_batch_host(b200_ctx *ctx, const B200MeLaunch *launch, size_t cur_bytes, size_t ref_bytes,
                       const B200MeBlock *subpel_blocks, B200MeResult *fullpel_out,
                       B200MeResult *subpel_out);
/* ---------------------------------------------- frame tier + encoder block queue ------ */
/* SURVEY.md section 8b tier 3 / rows a25, a26, f1: what an encoder's frame loop binds instead of
 * its per-block loops.  A b200_frame is one luma plane resident on the device with its replicated
 * border (aom_extend_frame_borders, aom_scale/generic/yv12extend.c:223: the motion search reads up
 * to `border` pixels outside the picture); the queue calls build every descriptor of a frame on
 * the host, run ONE chained launch sequence, and scatter the results back. */
typedef struct b200_frame b200_frame;
/* y: host pointer to pixel (0,0) of a width x height plane of uint8 (bd == 8) or uint16 samples,
 * `stride` in pixels; the border is replicated on the device, the host buffer needs none. */
int b200_frame_upload(b200_ctx *ctx, const void *y, int stride, int width, int height, int border, int bd,
                      b200_frame **frame);
int b200_frame_free(b200_ctx *ctx, b200_frame *frame);
/* reads the bordered device plane back (tests): out holds (height + 2*border) rows of
 * (width + 2*border) pixels */
int b200_frame_download(b200_ctx *ctx, const b200_frame *frame, void *out);

/* tf_motion_search (av1/encoder/temporal_filter.c:88-260) for every TF_BLOCK_SIZE block of a frame
 * pair, the loop of av1_tf_do_filtering_row (:804-872): full-pel NSTEP + mesh and the 8-tap
 * sub-pel refinement of the whole block, then of its 4 sub-blocks started at the block's refined
 * mv, the split decision (tf_determine_block_partition, :275) and the ref_mv hand-down rule
 * (:255-259).  Four chained launches per frame pair, no host round trip in between. */
typedef struct {
  int bw, bh;                 /* TF_BLOCK_SIZE (temporal_filter.h:31): 16x16; sub-blocks are bw/2 x bh/2 */
  int border_in_pixels;       /* cpi->oxcf.border_in_pixels, av1_set_mv_{row,col}_limits (mcomp.h:227-251) */
  int search_method;          /* B200_ME_NSTEP (temporal_filter.c:134) */
  int step_param;             /* av1_init_search_range(max(width, height)) (:112) */
  int mv_cost_type;           /* B200_MV_COST_L1_* by min(width, height) (:116-119) */
  int mesh_range[4], mesh_interval[4]; /* sf.mv_sf.mesh_patterns */
  int force_mesh_thresh;      /* sf.mv_sf.exhaustive_searches_thresh */
  int prune_mesh_search;      /* (:152-156) */
  int mesh_search_mv_diff_threshold;
  int use_downsampled_sad;    /* sf.mv_sf.use_downsampled_sad (applies when bh >= 16, mcomp.c:135) */
  int sad_per_bit, error_per_bit; /* x->sadperbit, x->errorperbit */
  int allow_hp;               /* cm->features.allow_high_precision_mv */
  int iters_per_step;         /* sf.mv_sf.subpel_iters_per_step */
  int subpel_tree;            /* B200_SUBPEL_TREE*: cpi->mv_search_params.find_fractional_mv_step */
  int want_cost_list;         /* cond_cost_list(cpi, cost_list) != NULL */
} B200TfParams;
typedef struct {
  int16_t block_mv[2];  /* row, col, 1/8 pel: the whole block's refined mv */
  int32_t block_mse;
  int16_t sub_mv[4][2]; /* subblock_mvs after the split decision */
  int32_t sub_mse[4];   /* subblock_mses after the split decision */
  int16_t ref_mv[2];    /* *ref_mv handed to the next frame of the window */
} B200TfResult; /* 44 bytes */
/* temporal_filter.c:112-119 defaults for a width x height frame + the good-quality speed features */
void b200_queue_tf_default_params(int width, int height, B200TfParams *params);
/* ref_mvs: 2 x int16 (row, col; 1/8 pel) per block in raster order, or NULL for kZeroMv;
 * out: ceil(height / bh) * ceil(width / bw) results in raster order (host). */
int b200_queue_tf_motion_search(b200_ctx *ctx, const b200_frame *to_filter, const b200_frame *ref,
                                const B200TfParams *params, const int16_t *ref_mvs, B200TfResult *out);
/* The descriptors b200_queue_tf_motion_search builds, without touching the device (tests, and
 * callers that want to launch themselves): stage 0 = whole blocks, stage 1 = sub-blocks; full[]
 * and sub[] take n (stage 0) or 4n (stage 1) descriptors.  Returns the count written. */
int b200_queue_tf_describe(int width, int height, int frame_pitch, int frame_origin, const B200TfParams *params,
                           const int16_t *ref_mvs, int stage, B200MeBlock *full, B200MeBlock *sub);

/* TPL model (av1/encoder/tpl_model.c): motion_estimation (:249-304) for every block of a frame
 * against one reference, each from a caller-supplied centre mv (mode_estimation takes them from the
 * neighbouring blocks' results, :640-690 -- a caller runs one call per candidate / wavefront), and
 * txfm_quant_rdcost (:228-247) up to its rate / distortion outputs for every block of a frame. */
typedef struct {
  int bw, bh;                 /* tpl block size (tpl_data->tpl_bsize_1d): 16x16 */
  int border_in_pixels;       /* tpl_data->border_in_pixels (:1353, :1381) */
  int search_method;          /* tpl_sf.search_method */
  int step_param;             /* min(tpl_sf.reduce_first_step_size, MAX_MVSEARCH_STEPS - 2) (:267-268) */
  int forced_stop;            /* tpl_sf.subpel_force_stop (B200_EIGHTH_PEL ..) */
  int mv_cost_type;           /* B200_MV_COST_ENTROPY is what init_mv_cost_params leaves (mcomp.c:42) */
  int mesh_range[4], mesh_interval[4];
  int force_mesh_thresh;      /* sf.mv_sf.exhaustive_searches_thresh (NSTEP may escalate to the mesh) */
  int prune_mesh_search;
  int use_downsampled_sad;
  int sad_per_bit, error_per_bit;
  int allow_hp, iters_per_step, subpel_tree, want_cost_list;
  const int *mvjcost;         /* host, 4 entries, or NULL (then mv cost terms are 0) */
  const int *mvcost[2];       /* host, centred: index -MV_MAX..MV_MAX */
} B200TplParams;
typedef struct {
  int16_t mv[2];   /* row, col, 1/8 pel: best_mv->as_mv */
  uint32_t sme;    /* motion_estimation's return value (bestsme) */
} B200TplMeResult; /* 8 bytes */
void b200_queue_tpl_default_params(int width, int height, B200TplParams *params);
/* out: ceil(mi_rows * 4 / bh) * ceil(mi_cols * 4 / bw) results, raster order */
int b200_queue_tpl_motion_estimation(b200_ctx *ctx, const b200_frame *cur, const b200_frame *ref,
                                     const B200TplParams *params, const int16_t *center_mvs,
                                     B200TplMeResult *out);
int b200_queue_tpl_describe(int width, int height, int frame_pitch, int frame_origin, const B200TplParams *params,
                            const int16_t *center_mvs, B200MeBlock *full, B200MeBlock *sub);
typedef struct {
  int32_t rate_cost;   /* rate_estimator(qcoeff, eob, tx_size) (:207-226) */
  int32_t eob;
  int64_t recon_error; /* max(av1_block_error(coeff, dqcoeff) >> shift, 1) (:97-135) */
  int64_t sse;         /* max(ssz >> shift, 1) */
} B200TplRdResult; /* 24 bytes */
/* src / pred: HOST planes of uint8 (bd 8) or uint16 samples holding rows x cols blocks of the square
 * tx_size (B200_TX_4X4 .. B200_TX_32X32) from their first sample; qrow: the plane-0 quantiser row of
 * the frame's qindex.  Subtract, DCT_DCT, av1_quantize_fp / av1_highbd_quantize_fp and the block
 * error run fused on the device; the levels come back in the compact form and the rate is
 * accumulated from them on the host. */
int b200_queue_tpl_txfm_quant_rdcost(b200_ctx *ctx, const void *src, int src_stride, const void *pred,
                                     int pred_stride, int rows, int cols, int bd, int tx_size,
                                     const B200QRow *qrow, B200TplRdResult *out);

/* Measured rate of the integer pipe on this device: VABSDIFF4 with accumulate (the SAD
 * instruction) issued back to back from every SM, in G pixel-absdiffs per second.  The
 * denominator of the motion search's int-ALU roofline. */
int b200_measure_sad_peak(b200_ctx *ctx, double *gabsdiff_per_s);

#ifdef __cplusplus
}
#endif
#endif /* B200AV1_H_ */
