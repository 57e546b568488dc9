/*
 * ref_shim.c -- compiled ONLY into oracle/_ref/libaomref.so (by oracle/mk_ref.py, when
 * /root/reference is present).  TEST INFRASTRUCTURE.
 *
 * Builds the reference's own FULLPEL_MOTION_SEARCH_PARAMS / SUBPEL_MOTION_SEARCH_PARAMS
 * (real struct definitions from the reference headers) out of the flat Av1oMeParams and
 * calls the UNMODIFIED av1_full_pixel_search / av1_find_best_sub_pixel_tree, so that the
 * oracle's motion-search restatement can be pinned against the reference itself.
 */
#include <stdlib.h>
#include <string.h>

#include "config/aom_config.h"
#include "config/aom_dsp_rtcd.h"
#include "config/av1_rtcd.h"
#include "config/aom_scale_rtcd.h"

#include "av1/common/av1_common_int.h"
#include "av1/common/blockd.h"
#include "av1/encoder/encoder.h"
#include "av1/encoder/mcomp.h"

#include "av1o_me.h"

/* high-bit-depth SADs as the encoder sees them: >> 2 at 10 bit, >> 4 at 12 bit
 * (av1/encoder/encoder_utils.h:153-208) */
/* REFSHIM_DISPATCH (the avx2 flavour, oracle/mk_ref.py): every kernel goes through the
 * reference's own run-time dispatch names instead of the `_c` flavour. */
#ifdef REFSHIM_DISPATCH
#define RF(name) name
#else
#define RF(name) name##_c
#endif

#define SHIM_FNS(W, H)                                                                         \
  static unsigned sad##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *b, int bs) {     \
    return RF(aom_highbd_sad##W##x##H)(a, as, b, bs) >> 2;                                     \
  }                                                                                            \
  static unsigned sad##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *b, int bs) {     \
    return RF(aom_highbd_sad##W##x##H)(a, as, b, bs) >> 4;                                     \
  }                                                                                            \
  static unsigned sads##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *b, int bs) {    \
    return RF(aom_highbd_sad_skip_##W##x##H)(a, as, b, bs) >> 2;                               \
  }                                                                                            \
  static unsigned sads##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *b, int bs) {    \
    return RF(aom_highbd_sad_skip_##W##x##H)(a, as, b, bs) >> 4;                               \
  }                                                                                            \
  static void sad4##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *const b[4], int bs, \
                                  uint32_t *o) {                                               \
    RF(aom_highbd_sad##W##x##H##x4d)(a, as, b, bs, o);                                         \
    for (int i = 0; i < 4; ++i) o[i] >>= 2;                                                    \
  }                                                                                            \
  static void sad4##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *const b[4], int bs, \
                                  uint32_t *o) {                                               \
    RF(aom_highbd_sad##W##x##H##x4d)(a, as, b, bs, o);                                         \
    for (int i = 0; i < 4; ++i) o[i] >>= 4;                                                    \
  }                                                                                            \
  static void sads4##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *const b[4],        \
                                   int bs, uint32_t *o) {                                      \
    RF(aom_highbd_sad_skip_##W##x##H##x4d)(a, as, b, bs, o);                                   \
    for (int i = 0; i < 4; ++i) o[i] >>= 2;                                                    \
  }                                                                                            \
  static void sads4##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *const b[4],        \
                                   int bs, uint32_t *o) {                                      \
    RF(aom_highbd_sad_skip_##W##x##H##x4d)(a, as, b, bs, o);                                   \
    for (int i = 0; i < 4; ++i) o[i] >>= 4;                                                    \
  }                                                                                            \
  static int fill##W##x##H(aom_variance_fn_ptr_t *f, int highbd, int bd) {                     \
    memset(f, 0, sizeof(*f));                                                                  \
    if (!highbd) {                                                                             \
      f->sdf = RF(aom_sad##W##x##H);                                                           \
      f->sdsf = RF(aom_sad_skip_##W##x##H);                                                    \
      f->vf = RF(aom_variance##W##x##H);                                                       \
      f->svf = RF(aom_sub_pixel_variance##W##x##H);                                            \
      f->sdx4df = RF(aom_sad##W##x##H##x4d);                                                   \
      f->sdx3df = RF(aom_sad##W##x##H##x3d);                                                   \
      f->sdsx4df = RF(aom_sad_skip_##W##x##H##x4d);                                            \
    } else if (bd == 10) {                                                                     \
      f->sdf = sad##W##x##H##_b10;                                                             \
      f->sdsf = sads##W##x##H##_b10;                                                           \
      f->vf = RF(aom_highbd_10_variance##W##x##H);                                             \
      f->svf = RF(aom_highbd_10_sub_pixel_variance##W##x##H);                                  \
      f->sdx4df = sad4##W##x##H##_b10;                                                         \
      f->sdx3df = sad4##W##x##H##_b10;                                                         \
      f->sdsx4df = sads4##W##x##H##_b10;                                                       \
    } else {                                                                                   \
      f->sdf = sad##W##x##H##_b12;                                                             \
      f->sdsf = sads##W##x##H##_b12;                                                           \
      f->vf = RF(aom_highbd_12_variance##W##x##H);                                             \
      f->svf = RF(aom_highbd_12_sub_pixel_variance##W##x##H);                                  \
      f->sdx4df = sad4##W##x##H##_b12;                                                         \
      f->sdx3df = sad4##W##x##H##_b12;                                                         \
      f->sdsx4df = sads4##W##x##H##_b12;                                                       \
    }                                                                                          \
    return BLOCK_##W##X##H;                                                                    \
  }

SHIM_FNS(128, 128) SHIM_FNS(128, 64) SHIM_FNS(64, 128) SHIM_FNS(64, 64) SHIM_FNS(64, 32)
SHIM_FNS(32, 64) SHIM_FNS(32, 32) SHIM_FNS(32, 16) SHIM_FNS(16, 32) SHIM_FNS(16, 16)
SHIM_FNS(16, 8) SHIM_FNS(8, 16) SHIM_FNS(8, 8) SHIM_FNS(8, 4) SHIM_FNS(4, 8) SHIM_FNS(4, 4)
SHIM_FNS(4, 16) SHIM_FNS(16, 4) SHIM_FNS(8, 32) SHIM_FNS(32, 8) SHIM_FNS(16, 64) SHIM_FNS(64, 16)

static int fill_vfp(aom_variance_fn_ptr_t *f, int w, int h, int highbd, int bd) {
#define TRY(W, H) \
  if (w == W && h == H) return fill##W##x##H(f, highbd, bd);
  TRY(128, 128) TRY(128, 64) TRY(64, 128) TRY(64, 64) TRY(64, 32) TRY(32, 64) TRY(32, 32)
  TRY(32, 16) TRY(16, 32) TRY(16, 16) TRY(16, 8) TRY(8, 16) TRY(8, 8) TRY(8, 4) TRY(4, 8)
  TRY(4, 4) TRY(4, 16) TRY(16, 4) TRY(8, 32) TRY(32, 8) TRY(16, 64) TRY(64, 16)
#undef TRY
  return -1;
}

static const uint8_t *as_ref_ptr(const void *p, int highbd) {
  return highbd ? CONVERT_TO_BYTEPTR(p) : (const uint8_t *)p;
}

static void fill_cost(MV_COST_PARAMS *c, const Av1oMeParams *p, MV *ref_mv_store) {
  ref_mv_store->row = (int16_t)p->ref_mv_row;
  ref_mv_store->col = (int16_t)p->ref_mv_col;
  c->ref_mv = ref_mv_store;
  c->full_ref_mv = get_fullmv_from_mv(ref_mv_store);
  c->mv_cost_type = (MV_COST_TYPE)p->mv_cost_type;
  c->mvjcost = p->mvjcost;
  c->mvcost[0] = p->mvcost0;
  c->mvcost[1] = p->mvcost1;
  c->error_per_bit = p->error_per_bit;
  c->sad_per_bit = p->sad_per_bit;
}

static void fullpel_common(const Av1oMeParams *p, Av1oMeResult *r, int refine8p) {
  FULLPEL_MOTION_SEARCH_PARAMS ms;
  memset(&ms, 0, sizeof(ms));
  aom_variance_fn_ptr_t vfp;
  const int bsize = fill_vfp(&vfp, p->bw, p->bh, p->highbd, p->bd);
  if (bsize < 0) abort();
  struct buf_2d src, ref;
  memset(&src, 0, sizeof(src));
  memset(&ref, 0, sizeof(ref));
  src.buf = (uint8_t *)as_ref_ptr(p->src, p->highbd);
  src.stride = p->src_stride;
  ref.buf = (uint8_t *)as_ref_ptr(p->ref, p->highbd);
  ref.stride = p->ref_stride;
  ms.bsize = (BLOCK_SIZE)bsize;
  ms.vfp = &vfp;
  ms.ms_buffers.src = &src;
  ms.ms_buffers.ref = &ref;
  search_site_config *cfg = (search_site_config *)calloc(1, sizeof(*cfg));
  switch (p->search_method) {
    case AV1O_NSTEP:
      ms.search_method = NSTEP;
      av1_init_motion_compensation_nstep(cfg, p->ref_stride, 0);
      break;
    case AV1O_NSTEP_8PT:
      ms.search_method = NSTEP_8PT;
      av1_init_motion_compensation_nstep(cfg, p->ref_stride, 1);
      break;
    case AV1O_DIAMOND:
      ms.search_method = DIAMOND;
      av1_init_dsmotion_compensation(cfg, p->ref_stride, 0);
      break;
    case AV1O_CLAMPED_DIAMOND:
      ms.search_method = CLAMPED_DIAMOND;
      av1_init_dsmotion_compensation(cfg, p->ref_stride, 1);
      break;
    case AV1O_HEX:
    case AV1O_FAST_HEX:
      ms.search_method = p->search_method == AV1O_HEX ? HEX : FAST_HEX;
      av1_init_motion_compensation_hex(cfg, p->ref_stride, 0);
      break;
    case AV1O_SQUARE:
      ms.search_method = SQUARE;
      av1_init_motion_compensation_square(cfg, p->ref_stride, 0);
      break;
    default:
      ms.search_method = p->search_method == AV1O_BIGDIA        ? BIGDIA
                         : p->search_method == AV1O_FAST_DIAMOND ? FAST_DIAMOND
                         : p->search_method == AV1O_FAST_BIGDIA  ? FAST_BIGDIA
                                                                 : VFAST_DIAMOND;
      av1_init_motion_compensation_bigdia(cfg, p->ref_stride, 0);
      break;
  }
  ms.search_sites = cfg;
  ms.mv_limits.col_min = p->col_min;
  ms.mv_limits.col_max = p->col_max;
  ms.mv_limits.row_min = p->row_min;
  ms.mv_limits.row_max = p->row_max;
  ms.run_mesh_search = p->run_mesh_search;
  ms.prune_mesh_search = p->prune_mesh_search;
  ms.mesh_search_mv_diff_threshold = p->mesh_search_mv_diff_threshold;
  ms.force_mesh_thresh = p->force_mesh_thresh;
  MESH_PATTERN mesh[MAX_MESH_STEP];
  for (int i = 0; i < MAX_MESH_STEP; ++i) {
    mesh[i].range = p->mesh_range[i];
    mesh[i].interval = p->mesh_interval[i];
  }
  ms.mesh_patterns[0] = mesh;
  ms.mesh_patterns[1] = mesh;
  ms.fine_search_interval = p->fine_search_interval;
  ms.is_intra_mode = 0;
  MV ref_mv;
  fill_cost(&ms.mv_cost_params, p, &ref_mv);
  if (p->use_skip_sad) {
    ms.sdf = vfp.sdsf;
    ms.sdx4df = vfp.sdsx4df;
    ms.sdx3df = vfp.sdsx4df;
  } else {
    ms.sdf = vfp.sdf;
    ms.sdx4df = vfp.sdx4df;
    ms.sdx3df = vfp.sdx3df;
  }
  const FULLPEL_MV start = { (int16_t)p->start_row, (int16_t)p->start_col };
  FULLPEL_MV best, second;
  int cost_list[5];
  if (refine8p) {
    memset(r, 0, sizeof(*r));
    r->ret = av1_refining_search_8p_c(&ms, start, &best);
    r->best_row = best.row;
    r->best_col = best.col;
    r->second_row = r->second_col = -32768;
    free(cfg);
    return;
  }
  r->ret = av1_full_pixel_search(start, &ms, p->step_param, p->want_cost_list ? cost_list : NULL,
                                 &best, p->want_second_best ? &second : NULL);
  r->best_row = best.row;
  r->best_col = best.col;
  r->second_row = p->want_second_best ? second.row : -32768;
  r->second_col = p->want_second_best ? second.col : -32768;
  for (int i = 0; i < 5; ++i) r->cost_list[i] = p->want_cost_list ? cost_list[i] : 0;
  free(cfg);
}

static void subpel_common(const Av1oMeParams *p, Av1oMeResult *r, int tree) {
  SUBPEL_MOTION_SEARCH_PARAMS ms;
  memset(&ms, 0, sizeof(ms));
  aom_variance_fn_ptr_t vfp;
  if (fill_vfp(&vfp, p->bw, p->bh, p->highbd, p->bd) < 0) abort();
  struct buf_2d src, ref;
  memset(&src, 0, sizeof(src));
  memset(&ref, 0, sizeof(ref));
  src.buf = (uint8_t *)as_ref_ptr(p->src, p->highbd);
  src.stride = p->src_stride;
  ref.buf = (uint8_t *)as_ref_ptr(p->ref, p->highbd);
  ref.stride = p->ref_stride;
  ms.allow_hp = p->allow_hp;
  int cl[5];
  for (int i = 0; i < 5; ++i) cl[i] = p->sp_cost_list[i];
  ms.cost_list = (tree && p->sp_have_cost_list) ? cl : NULL;
  ms.forced_stop = (SUBPEL_FORCE_STOP)p->forced_stop;
  ms.iters_per_step = p->iters_per_step;
  ms.mv_limits.col_min = p->sp_col_min;
  ms.mv_limits.col_max = p->sp_col_max;
  ms.mv_limits.row_min = p->sp_row_min;
  ms.mv_limits.row_max = p->sp_row_max;
  MV ref_mv;
  fill_cost(&ms.mv_cost_params, p, &ref_mv);
  ms.var_params.vfp = &vfp;
  ms.var_params.subpel_search_type = (SUBPEL_SEARCH_TYPE)p->subpel_search_type;
  ms.var_params.ms_buffers.src = &src;
  ms.var_params.ms_buffers.ref = &ref;
  ms.var_params.w = p->bw;
  ms.var_params.h = p->bh;

  /* the few MACROBLOCKD / AV1_COMMON fields the search touches (mcomp.c:3151-3154,
   * 2421-2428; reconinter_enc.c:430-436) */
  MACROBLOCKD *xd = (MACROBLOCKD *)calloc(1, sizeof(*xd));
  AV1_COMMON *cm = (AV1_COMMON *)calloc(1, sizeof(*cm));
  MB_MODE_INFO *mbmi = (MB_MODE_INFO *)calloc(1, sizeof(*mbmi));
  MB_MODE_INFO *mi_arr[1] = { mbmi };
  YV12_BUFFER_CONFIG *cur = (YV12_BUFFER_CONFIG *)calloc(1, sizeof(*cur));
  struct scale_factors sf;
  memset(&sf, 0, sizeof(sf));
  sf.x_scale_fp = REF_NO_SCALE;
  sf.y_scale_fp = REF_NO_SCALE;
  cm->sf_identity = sf;
  cur->flags = p->highbd ? YV12_FLAG_HIGHBITDEPTH : 0;
  xd->mi = mi_arr;
  xd->cur_buf = cur;
  xd->bd = p->bd;
  xd->block_ref_scale_factors[0] = &sf;
  xd->block_ref_scale_factors[1] = &sf;

  const MV start = { (int16_t)p->sp_start_row, (int16_t)p->sp_start_col };
  MV best;
  int distortion = 0;
  unsigned sse = 0;
  if (tree == 0)
    r->ret = av1_find_best_sub_pixel_tree(xd, cm, &ms, start, &best, &distortion, &sse, NULL);
  else if (tree == 1)
    r->ret = av1_find_best_sub_pixel_tree_pruned(xd, cm, &ms, start, &best, &distortion, &sse, NULL);
  else
    r->ret = av1_find_best_sub_pixel_tree_pruned_more(xd, cm, &ms, start, &best, &distortion, &sse,
                                                      NULL);
  r->best_row = best.row;
  r->best_col = best.col;
  r->distortion = distortion;
  r->sse = sse;
  free(xd);
  free(cm);
  free(mbmi);
  free(cur);
}

void refshim_find_best_sub_pixel_tree(const Av1oMeParams *p, Av1oMeResult *r) {
  subpel_common(p, r, 0);
}

/* av1_find_best_sub_pixel_tree_pruned (sp_tree 1) / _pruned_more (sp_tree 2), unmodified */
void refshim_find_best_sub_pixel_tree_pruned(const Av1oMeParams *p, Av1oMeResult *r) {
  subpel_common(p, r, p->sp_tree == 2 ? 2 : 1);
}

void refshim_full_pixel_search(const Av1oMeParams *p, Av1oMeResult *r) { fullpel_common(p, r, 0); }

/* av1_refining_search_8p_c (mcomp.c:1683), single reference */
void refshim_refining_search_8p(const Av1oMeParams *p, Av1oMeResult *r) { fullpel_common(p, r, 1); }

/* quantize_fp_helper_c / highbd_quantize_fp_helper_c WITH a qmatrix are static
 * (av1_quantize.c:70,125); the only way in is the facade with a hand-filled
 * MACROBLOCK_PLANE + QUANT_PARAM (av1_quantize.c:266, :423). */
void refshim_quantize_fp_qm(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                            const int16_t *round_fp, const int16_t *quant_fp,
                            const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                            const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                            const int16_t *iscan, const uint8_t *qm, const uint8_t *iqm,
                            int log_scale, int highbd) {
  MACROBLOCK_PLANE p;
  memset(&p, 0, sizeof(p));
  p.zbin_QTX = zbin;
  p.round_fp_QTX = round_fp;
  p.quant_fp_QTX = quant_fp;
  p.quant_shift_QTX = quant_shift;
  p.dequant_QTX = dequant;
  SCAN_ORDER sc;
  sc.scan = scan;
  sc.iscan = iscan;
  QUANT_PARAM qp;
  memset(&qp, 0, sizeof(qp));
  qp.log_scale = log_scale;
  qp.qmatrix = qm;
  qp.iqmatrix = iqm;
  if (highbd)
    av1_highbd_quantize_fp_facade(coeff, n, &p, qcoeff, dqcoeff, eob, &sc, &qp);
  else
    av1_quantize_fp_facade(coeff, n, &p, qcoeff, dqcoeff, eob, &sc, &qp);
}

/* ---- function pointers for the timed CPU baseline (oracle/av1o_drv.c) -------------------------
 * The transform as the encoder reaches it: av1_fwd_txfm (hybrid_fwd_txfm.c:231) ->
 * av1_lowbd_fwd_txfm / av1_highbd_fwd_txfm, i.e. the SIMD flavours in the avx2 build. */
#include "av1/encoder/hybrid_fwd_txfm.h"

#define SHIM_TX(TS)                                                                                   \
  static void fwd_lb_##TS(const int16_t *in, int32_t *out, int stride, TX_TYPE t, int bd) {           \
    TxfmParam p;                                                                                      \
    memset(&p, 0, sizeof(p));                                                                         \
    p.tx_type = t, p.tx_size = (TX_SIZE)TS, p.bd = bd, p.tx_set_type = EXT_TX_SET_ALL16;              \
    av1_lowbd_fwd_txfm(in, out, stride, &p);                                                          \
  }                                                                                                   \
  static void fwd_hb_##TS(const int16_t *in, int32_t *out, int stride, TX_TYPE t, int bd) {           \
    TxfmParam p;                                                                                      \
    memset(&p, 0, sizeof(p));                                                                         \
    p.tx_type = t, p.tx_size = (TX_SIZE)TS, p.bd = bd, p.is_hbd = 1, p.tx_set_type = EXT_TX_SET_ALL16; \
    av1_highbd_fwd_txfm(in, out, stride, &p);                                                         \
  }
SHIM_TX(0) SHIM_TX(1) SHIM_TX(2) SHIM_TX(3) SHIM_TX(4) SHIM_TX(5) SHIM_TX(6) SHIM_TX(7) SHIM_TX(8) SHIM_TX(9)
SHIM_TX(10) SHIM_TX(11) SHIM_TX(12) SHIM_TX(13) SHIM_TX(14) SHIM_TX(15) SHIM_TX(16) SHIM_TX(17) SHIM_TX(18)

typedef void (*shim_fwd_fn)(const int16_t *, int32_t *, int, TX_TYPE, int);
void *refshim_fwd_txfm_fn(int tx_size, int highbd) {
#define E(TS) { fwd_lb_##TS, fwd_hb_##TS }
  static const shim_fwd_fn tab[19][2] = { E(0), E(1), E(2), E(3), E(4), E(5), E(6), E(7), E(8), E(9), E(10),
                                          E(11), E(12), E(13), E(14), E(15), E(16), E(17), E(18) };
#undef E
  if (tx_size < 0 || tx_size >= 19) return NULL;
  return (void *)tab[tx_size][highbd != 0];
}

void *refshim_quantize_fp_fn(int log_scale, int highbd) {
  if (highbd) return (void *)av1_highbd_quantize_fp;
  return log_scale == 0 ? (void *)av1_quantize_fp
                        : log_scale == 1 ? (void *)av1_quantize_fp_32x32 : (void *)av1_quantize_fp_64x64;
}

/* 1 when this library was built with the x86 intrinsics and run-time dispatch */
int refshim_simd(void) {
#ifdef REFSHIM_DISPATCH
  return 1;
#else
  return 0;
#endif
}

#ifdef REFSHIM_DISPATCH
__attribute__((constructor)) static void shim_rtcd_init(void) {
  aom_dsp_rtcd();
  av1_rtcd();
  aom_scale_rtcd();
}
#endif
