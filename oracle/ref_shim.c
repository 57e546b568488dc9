/*
 * ref_shim.c -- compiled ONLY into oracle/_ref/libaomref.so (by oracle/mk_ref.py, when
 * /root/reference is present).  TEST INFRASTRUCTURE.
 *
 * Builds the reference's own FULLPEL_MOTION_SEARCH_PARAMS / SUBPEL_MOTION_SEARCH_PARAMS
 * (real struct definitions from the reference headers) out of the flat Av1oMeParams and
 * calls the UNMODIFIED av1_full_pixel_search / av1_find_best_sub_pixel_tree, so that the
 * oracle's motion-search restatement can be pinned against the reference itself.
 */
#include <pthread.h>
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
  static unsigned sada##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *b, int bs,      \
                                      const uint8_t *sp) {                                     \
    return RF(aom_highbd_sad##W##x##H##_avg)(a, as, b, bs, sp) >> 2;                           \
  }                                                                                            \
  static unsigned sada##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *b, int bs,      \
                                      const uint8_t *sp) {                                     \
    return RF(aom_highbd_sad##W##x##H##_avg)(a, as, b, bs, sp) >> 4;                           \
  }                                                                                            \
  static unsigned msad##W##x##H##_b10(const uint8_t *a, int as, const uint8_t *b, int bs,      \
                                      const uint8_t *sp, const uint8_t *m, int ms, int inv) {  \
    return RF(aom_highbd_masked_sad##W##x##H)(a, as, b, bs, sp, m, ms, inv) >> 2;              \
  }                                                                                            \
  static unsigned msad##W##x##H##_b12(const uint8_t *a, int as, const uint8_t *b, int bs,      \
                                      const uint8_t *sp, const uint8_t *m, int ms, int inv) {  \
    return RF(aom_highbd_masked_sad##W##x##H)(a, as, b, bs, sp, m, ms, inv) >> 4;              \
  }                                                                                            \
  static int fill##W##x##H(aom_variance_fn_ptr_t *f, int highbd, int bd) {                     \
    memset(f, 0, sizeof(*f));                                                                  \
    if (!highbd) {                                                                             \
      f->sdaf = RF(aom_sad##W##x##H##_avg);                                                    \
      f->svaf = RF(aom_sub_pixel_avg_variance##W##x##H);                                       \
      f->msdf = RF(aom_masked_sad##W##x##H);                                                   \
      f->msvf = RF(aom_masked_sub_pixel_variance##W##x##H);                                    \
      f->sdf = RF(aom_sad##W##x##H);                                                           \
      f->sdsf = RF(aom_sad_skip_##W##x##H);                                                    \
      f->vf = RF(aom_variance##W##x##H);                                                       \
      f->svf = RF(aom_sub_pixel_variance##W##x##H);                                            \
      f->sdx4df = RF(aom_sad##W##x##H##x4d);                                                   \
      f->sdx3df = RF(aom_sad##W##x##H##x3d);                                                   \
      f->sdsx4df = RF(aom_sad_skip_##W##x##H##x4d);                                            \
    } else if (bd == 10) {                                                                     \
      f->sdaf = sada##W##x##H##_b10;                                                           \
      f->svaf = RF(aom_highbd_10_sub_pixel_avg_variance##W##x##H);                             \
      f->msdf = msad##W##x##H##_b10;                                                           \
      f->msvf = RF(aom_highbd_10_masked_sub_pixel_variance##W##x##H);                          \
      f->sdf = sad##W##x##H##_b10;                                                             \
      f->sdsf = sads##W##x##H##_b10;                                                           \
      f->vf = RF(aom_highbd_10_variance##W##x##H);                                             \
      f->svf = RF(aom_highbd_10_sub_pixel_variance##W##x##H);                                  \
      f->sdx4df = sad4##W##x##H##_b10;                                                         \
      f->sdx3df = sad4##W##x##H##_b10;                                                         \
      f->sdsx4df = sads4##W##x##H##_b10;                                                       \
    } else {                                                                                   \
      f->sdaf = sada##W##x##H##_b12;                                                           \
      f->svaf = RF(aom_highbd_12_sub_pixel_avg_variance##W##x##H);                             \
      f->msdf = msad##W##x##H##_b12;                                                           \
      f->msvf = RF(aom_highbd_12_masked_sub_pixel_variance##W##x##H);                          \
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
  if (p->second_pred) /* av1_set_ms_compound_refs (mcomp.c:80) */
    av1_set_ms_compound_refs(&ms.ms_buffers, as_ref_ptr(p->second_pred, p->highbd), p->mask, p->mask_stride,
                             p->inv_mask);
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
  if (p->second_pred)
    av1_set_ms_compound_refs(&ms.var_params.ms_buffers, as_ref_ptr(p->second_pred, p->highbd), p->mask,
                             p->mask_stride, p->inv_mask);
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

/* ---- temporal filter: tf_motion_search over a frame pair ---------------------------------
 * cur / ref point at picture pixel (0,0) of planes that carry at least border_in_pixels of
 * replicated border around them (as YV12 buffers do), same stride. */
static void tf_fill(Av1oMeParams *P, const Av1oTfParams *tp, const void *cur, const void *ref, int stride,
                    int highbd, int bd, int x, int y, int bw, int bh, const FullMvLimits *mb_limits,
                    FULLPEL_MV start) {
  const MV baseline = kZeroMv; /* temporal_filter.c:124 */
  memset(P, 0, sizeof(*P));
  const int px = highbd ? 2 : 1;
  P->src = (const uint8_t *)cur + ((size_t)y * stride + x) * px;
  P->ref = (const uint8_t *)ref + ((size_t)y * stride + x) * px;
  P->src_stride = P->ref_stride = stride;
  P->bw = bw, P->bh = bh, P->highbd = highbd, P->bd = bd;
  FullMvLimits fl = *mb_limits;
  av1_set_mv_search_range(&fl, &baseline); /* mcomp.c:150 via av1_make_default_fullpel_ms_params */
  P->col_min = fl.col_min, P->col_max = fl.col_max, P->row_min = fl.row_min, P->row_max = fl.row_max;
  P->search_method = tp->search_method;
  P->step_param = tp->step_param;
  P->run_mesh_search = 1; /* :149 */
  P->prune_mesh_search = tp->prune_mesh_search;
  P->mesh_search_mv_diff_threshold = tp->mesh_search_mv_diff_threshold;
  P->force_mesh_thresh = tp->force_mesh_thresh;
  for (int i = 0; i < 4; ++i) P->mesh_range[i] = tp->mesh_range[i], P->mesh_interval[i] = tp->mesh_interval[i];
  P->use_skip_sad = tp->use_downsampled_sad && bh >= 16; /* mcomp.c:134-135 */
  P->mv_cost_type = tp->mv_cost_type;
  P->sad_per_bit = tp->sad_per_bit, P->error_per_bit = tp->error_per_bit;
  P->start_row = start.row, P->start_col = start.col;
  P->want_cost_list = tp->want_cost_list;
  SubpelMvLimits sl;
  av1_set_subpel_mv_search_range(&sl, mb_limits, &baseline); /* mcomp.c:181 */
  P->sp_col_min = sl.col_min, P->sp_col_max = sl.col_max, P->sp_row_min = sl.row_min, P->sp_row_max = sl.row_max;
  P->allow_hp = tp->allow_hp;
  P->forced_stop = EIGHTH_PEL;          /* :161 */
  P->subpel_search_type = USE_8_TAPS;   /* :113 */
  P->iters_per_step = tp->iters_per_step;
  P->sp_tree = tp->subpel_tree;
}

/* one full-pel + sub-pel search as tf_motion_search runs it; returns `error`, mv in *best */
static unsigned tf_search(Av1oMeParams *P, MV *best) {
  Av1oMeResult rf, rs;
  fullpel_common(P, &rf, 0);
  Av1oMeParams S = *P;
  S.mv_cost_type = AV1O_COST_NONE; /* :166 */
  S.sp_start_row = rf.best_row * 8, S.sp_start_col = rf.best_col * 8;
  S.sp_have_cost_list = P->want_cost_list;
  memcpy(S.sp_cost_list, rf.cost_list, sizeof(S.sp_cost_list));
  subpel_common(&S, &rs, S.sp_tree);
  best->row = (int16_t)rs.best_row, best->col = (int16_t)rs.best_col;
  return (unsigned)rs.ret;
}

typedef struct {
  const void *cur, *ref;
  int stride, width, height, highbd, bd;
  const Av1oTfParams *tp;
  const int16_t *ref_mvs;
  Av1oTfResult *out;
  volatile int *next_row;
} TfJob;

/* rows of blocks are independent (av1_tf_do_filtering_row is the encoder's own unit of
 * multi-threading, ethread.c tf_worker_hook): threads take rows dynamically */
static void *tf_rows_worker(void *arg) {
  const TfJob *j = (const TfJob *)arg;
  const void *cur = j->cur, *ref = j->ref;
  const int stride = j->stride, width = j->width, height = j->height, highbd = j->highbd, bd = j->bd;
  const Av1oTfParams *tp = j->tp;
  const int16_t *ref_mvs = j->ref_mvs;
  Av1oTfResult *out = j->out;
  CommonModeInfoParams mi;
  memset(&mi, 0, sizeof(mi));
  mi.mi_rows = ((height + 7) & ~7) >> MI_SIZE_LOG2; /* av1_set_mb_mi */
  mi.mi_cols = ((width + 7) & ~7) >> MI_SIZE_LOG2;
  const int mb_rows = (height + tp->bh - 1) / tp->bh, mb_cols = (width + tp->bw - 1) / tp->bw;
  const int min_frame_size = AOMMIN(width, height);
  const int mb_pels = tp->bw * tp->bh, sw = tp->bw >> 1, sh = tp->bh >> 1, sub_pels = sw * sh;
  for (;;) {
    const int mb_row = __sync_fetch_and_add(j->next_row, 1);
    if (mb_row >= mb_rows) break;
    FullMvLimits lim;
    av1_set_mv_row_limits(&mi, &lim, mb_row * (tp->bh >> MI_SIZE_LOG2), tp->bh >> MI_SIZE_LOG2,
                          tp->border_in_pixels); /* :846-848 */
    for (int mb_col = 0; mb_col < mb_cols; ++mb_col) {
      av1_set_mv_col_limits(&mi, &lim, mb_col * (tp->bw >> MI_SIZE_LOG2), tp->bw >> MI_SIZE_LOG2,
                            tp->border_in_pixels); /* :850-852 */
      const int bi = mb_row * mb_cols + mb_col;
      Av1oTfResult *o = &out[bi];
      MV ref_mv = kZeroMv;
      if (ref_mvs) ref_mv.row = ref_mvs[2 * bi], ref_mv.col = ref_mvs[2 * bi + 1];
      Av1oMeParams P;
      MV block_mv;
      tf_fill(&P, tp, cur, ref, stride, highbd, bd, mb_col * tp->bw, mb_row * tp->bh, tp->bw, tp->bh, &lim,
              get_fullmv_from_mv(&ref_mv));
      const unsigned error = tf_search(&P, &block_mv);
      const int block_mse = (int)DIVIDE_AND_ROUND(error, mb_pels); /* :172 */
      ref_mv = block_mv;
      MV sub_mv[4];
      int sub_mse[4];
      for (int k = 0; k < 4; ++k) {
        tf_fill(&P, tp, cur, ref, stride, highbd, bd, mb_col * tp->bw + (k & 1) * sw, mb_row * tp->bh + (k >> 1) * sh,
                sw, sh, &lim, get_fullmv_from_mv(&ref_mv)); /* :198 */
        const unsigned e = tf_search(&P, &sub_mv[k]);
        sub_mse[k] = (int)DIVIDE_AND_ROUND(e, sub_pels);
      }
      /* tf_determine_block_partition (:275-297, static) */
      int mn = INT_MAX, mx = INT_MIN;
      int64_t sum = 0;
      for (int k = 0; k < 4; ++k) {
        sum += sub_mse[k];
        mn = AOMMIN(mn, sub_mse[k]);
        mx = AOMMAX(mx, sub_mse[k]);
      }
      if (((block_mse * 15 < sum * 4) && mx - mn < 48) || ((block_mse * 14 < sum * 4) && mx - mn < 24))
        for (int k = 0; k < 4; ++k) sub_mv[k] = block_mv, sub_mse[k] = block_mse;
      const int thresh = (min_frame_size >= 720) ? 12 : 3; /* :255-259 */
      if (block_mse > (thresh << (bd - 8))) ref_mv = kZeroMv;
      o->block_mv[0] = block_mv.row, o->block_mv[1] = block_mv.col;
      o->block_mse = block_mse;
      for (int k = 0; k < 4; ++k) o->sub_mv[k][0] = sub_mv[k].row, o->sub_mv[k][1] = sub_mv[k].col, o->sub_mse[k] = sub_mse[k];
      o->ref_mv[0] = ref_mv.row, o->ref_mv[1] = ref_mv.col;
    }
  }
  return NULL;
}

void refshim_tf_motion_search_frame_mt(const void *cur, const void *ref, int stride, int width, int height,
                                       int highbd, int bd, const Av1oTfParams *tp, const int16_t *ref_mvs,
                                       Av1oTfResult *out, int n_threads) {
  volatile int next_row = 0;
  TfJob job = { cur, ref, stride, width, height, highbd, bd, tp, ref_mvs, out, &next_row };
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  pthread_t th[256];
  for (int t = 1; t < n_threads; ++t) pthread_create(&th[t], NULL, tf_rows_worker, &job);
  tf_rows_worker(&job);
  for (int t = 1; t < n_threads; ++t) pthread_join(th[t], NULL);
}

void refshim_tf_motion_search_frame(const void *cur, const void *ref, int stride, int width, int height,
                                    int highbd, int bd, const Av1oTfParams *tp, const int16_t *ref_mvs,
                                    Av1oTfResult *out) {
  refshim_tf_motion_search_frame_mt(cur, ref, stride, width, height, highbd, bd, tp, ref_mvs, out, 1);
}

/* the limits tf_motion_search hands to its searches for block (mb_row, mb_col):
 * out = full-pel {col_min, col_max, row_min, row_max}, then the sub-pel four */
void refshim_tf_limits(int width, int height, int bw, int bh, int border, int mb_row, int mb_col, int *out) {
  CommonModeInfoParams mi;
  memset(&mi, 0, sizeof(mi));
  mi.mi_rows = ((height + 7) & ~7) >> MI_SIZE_LOG2;
  mi.mi_cols = ((width + 7) & ~7) >> MI_SIZE_LOG2;
  FullMvLimits lim;
  av1_set_mv_limits(&mi, &lim, mb_row * (bh >> MI_SIZE_LOG2), mb_col * (bw >> MI_SIZE_LOG2), bh >> MI_SIZE_LOG2,
                    bw >> MI_SIZE_LOG2, border);
  const MV zero = kZeroMv;
  SubpelMvLimits sl;
  av1_set_subpel_mv_search_range(&sl, &lim, &zero);
  av1_set_mv_search_range(&lim, &zero);
  out[0] = lim.col_min, out[1] = lim.col_max, out[2] = lim.row_min, out[3] = lim.row_max;
  out[4] = sl.col_min, out[5] = sl.col_max, out[6] = sl.row_min, out[7] = sl.row_max;
}

/* ---- TPL: motion_estimation (tpl_model.c:249-304) for every block of a frame, each from its
 * centre mv, with the reference's own limit helpers and searches ------------------------- */
void refshim_tpl_motion_estimation_frame(const void *cur, const void *ref, int stride, int width, int height,
                                         int highbd, int bd, const Av1oTplParams *tp, const int16_t *center_mvs,
                                         Av1oTplMeResult *out) {
  CommonModeInfoParams mi;
  memset(&mi, 0, sizeof(mi));
  mi.mi_rows = ((height + 7) & ~7) >> MI_SIZE_LOG2;
  mi.mi_cols = ((width + 7) & ~7) >> MI_SIZE_LOG2;
  const int mi_h = tp->bh >> MI_SIZE_LOG2, mi_w = tp->bw >> MI_SIZE_LOG2;
  const int px = highbd ? 2 : 1;
  int n = 0;
  for (int mi_row = 0; mi_row < mi.mi_rows; mi_row += mi_h)      /* :1376 */
    for (int mi_col = 0; mi_col < mi.mi_cols; mi_col += mi_w, ++n) { /* :1349 */
      FullMvLimits lim;
      av1_set_mv_limits(&mi, &lim, mi_row, mi_col, mi_h, mi_w, tp->border_in_pixels); /* :1353, :1381 */
      MV center = kZeroMv;
      if (center_mvs) center.row = center_mvs[2 * n], center.col = center_mvs[2 * n + 1];
      Av1oMeParams P;
      memset(&P, 0, sizeof(P));
      const size_t off = ((size_t)mi_row * MI_SIZE * stride + (size_t)mi_col * MI_SIZE) * px;
      P.src = (const uint8_t *)cur + off, P.ref = (const uint8_t *)ref + off;
      P.src_stride = P.ref_stride = stride;
      P.bw = tp->bw, P.bh = tp->bh, P.highbd = highbd, P.bd = bd;
      FullMvLimits fl = lim;
      av1_set_mv_search_range(&fl, &center); /* av1_make_default_fullpel_ms_params, mcomp.c:150 */
      P.col_min = fl.col_min, P.col_max = fl.col_max, P.row_min = fl.row_min, P.row_max = fl.row_max;
      P.search_method = tp->search_method;
      P.step_param = tp->step_param;
      P.prune_mesh_search = tp->prune_mesh_search;
      P.mesh_search_mv_diff_threshold = 4;
      P.force_mesh_thresh = tp->force_mesh_thresh;
      for (int i = 0; i < 4; ++i) P.mesh_range[i] = tp->mesh_range[i], P.mesh_interval[i] = tp->mesh_interval[i];
      P.use_skip_sad = tp->use_downsampled_sad && tp->bh >= 16;
      P.mv_cost_type = tp->mv_cost_type;
      P.ref_mv_row = center.row, P.ref_mv_col = center.col;
      P.sad_per_bit = tp->sad_per_bit, P.error_per_bit = tp->error_per_bit;
      P.mvjcost = tp->mvjcost, P.mvcost0 = tp->mvcost[0], P.mvcost1 = tp->mvcost[1];
      const FULLPEL_MV start = get_fullmv_from_mv(&center); /* :259 */
      P.start_row = start.row, P.start_col = start.col;
      P.want_cost_list = tp->want_cost_list;
      Av1oMeResult rf, rs;
      fullpel_common(&P, &rf, 0);
      SubpelMvLimits sl;
      av1_set_subpel_mv_search_range(&sl, &lim, &center);
      P.sp_col_min = sl.col_min, P.sp_col_max = sl.col_max, P.sp_row_min = sl.row_min, P.sp_row_max = sl.row_max;
      P.allow_hp = tp->allow_hp;
      P.forced_stop = tp->forced_stop;       /* :292 */
      P.subpel_search_type = USE_2_TAPS;     /* :293 */
      P.iters_per_step = tp->iters_per_step;
      P.mv_cost_type = AV1O_COST_NONE;       /* :294 */
      P.sp_start_row = rf.best_row * 8, P.sp_start_col = rf.best_col * 8;
      P.sp_tree = tp->subpel_tree;
      P.sp_have_cost_list = tp->want_cost_list;
      memcpy(P.sp_cost_list, rf.cost_list, sizeof(P.sp_cost_list));
      subpel_common(&P, &rs, P.sp_tree);
      out[n].mv[0] = (int16_t)rs.best_row, out[n].mv[1] = (int16_t)rs.best_col;
      out[n].sme = (uint32_t)rs.ret;
    }
}

/* txfm_quant_rdcost (tpl_model.c:228-247) up to its outputs, for rows x cols blocks of a plane pair:
 * av1_subtract_block, av1_quick_txfm, get_quantize_error's av1_(highbd_)quantize_fp_facade +
 * av1_(highbd_)block_error -- all the reference's own -- and rate_estimator (:207-226, static) */
#include "av1/encoder/encodemb.h"
#include "av1/encoder/rdopt.h"
void refshim_tpl_txfm_quant_rdcost(const void *src, int src_stride, const void *pred, int pred_stride, int rows,
                                   int cols, int bd, int tx_size, const int16_t *qrow /* 7 x 8 int16, B200QRow */,
                                   Av1oTplRdResult *out) {
  const int highbd = bd > 8, w = tx_size_wide[tx_size];
  const int pix_num = 1 << num_pels_log2_lookup[txsize_to_bsize[tx_size]];
  BitDepthInfo bd_info;
  bd_info.bit_depth = bd, bd_info.use_highbitdepth_buf = highbd;
  int16_t *src_diff = (int16_t *)aom_memalign(32, sizeof(int16_t) * pix_num);
  tran_low_t *coeff = (tran_low_t *)aom_memalign(32, sizeof(tran_low_t) * pix_num * 3);
  tran_low_t *qcoeff = coeff + pix_num, *dqcoeff = qcoeff + pix_num;
  /* B200QRow: zbin, round, quant, quant_shift, round_fp, quant_fp, dequant (DC, AC entry each) -> the
   * 8-wide rows the quantisers read: [0] = DC, [1..7] = AC */
  int16_t tab[7][8] __attribute__((aligned(16)));
  for (int f = 0; f < 7; ++f)
    for (int k = 0; k < 8; ++k) tab[f][k] = qrow[2 * f + (k != 0)];
  MACROBLOCK_PLANE p;
  memset(&p, 0, sizeof(p));
  p.zbin_QTX = tab[0], p.round_QTX = tab[1], p.quant_QTX = tab[2], p.quant_shift_QTX = tab[3];
  p.round_fp_QTX = tab[4], p.quant_fp_QTX = tab[5], p.dequant_QTX = tab[6];
  const SCAN_ORDER *const scan_order = &av1_scan_orders[tx_size][DCT_DCT];
  const int shift = tx_size == TX_32X32 ? 0 : 2;
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c) {
      Av1oTplRdResult *o = &out[r * cols + c];
      const size_t so = (size_t)r * w * src_stride + (size_t)c * w, po = (size_t)r * w * pred_stride + (size_t)c * w;
      uint8_t *s8 = highbd ? CONVERT_TO_BYTEPTR((uint16_t *)src + so) : (uint8_t *)src + so;
      uint8_t *p8 = highbd ? CONVERT_TO_BYTEPTR((uint16_t *)pred + po) : (uint8_t *)pred + po;
      av1_subtract_block(bd_info, w, w, src_diff, w, s8, src_stride, p8, pred_stride);
      av1_quick_txfm(0, (TX_SIZE)tx_size, bd_info, src_diff, w, coeff);
      QUANT_PARAM qp;
      av1_setup_quant((TX_SIZE)tx_size, 0, AV1_XFORM_QUANT_FP, 0, &qp);
      uint16_t eob;
      int64_t sse, err;
      if (highbd) {
        av1_highbd_quantize_fp_facade(coeff, pix_num, &p, qcoeff, dqcoeff, &eob, scan_order, &qp);
        err = av1_highbd_block_error(coeff, dqcoeff, pix_num, &sse, bd) >> shift;
      } else {
        av1_quantize_fp_facade(coeff, pix_num, &p, qcoeff, dqcoeff, &eob, scan_order, &qp);
        err = av1_block_error(coeff, dqcoeff, pix_num, &sse) >> shift;
      }
      o->recon_error = AOMMAX(err, 1);
      sse >>= shift;
      o->sse = AOMMAX(sse, 1);
      int rate_cost = 1;
      for (int idx = 0; idx < eob; ++idx) {
        const unsigned abs_level = abs(qcoeff[scan_order->scan[idx]]);
        rate_cost += get_msb(abs_level + 1) + 1 + (abs_level > 0);
      }
      o->rate_cost = rate_cost << AV1_PROB_COST_SHIFT;
      o->eob = eob;
    }
  aom_free(src_diff);
  aom_free(coeff);
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

/* ---- av1_cost_coeffs_txb_estimate (txb_rdopt.c:547) on a bare coefficient block: the function reads
 * only x->plane[0].qcoeff and .eobs */
#include "av1/encoder/txb_rdopt.h"
int refshim_cost_coeffs_txb_estimate(const int32_t *qcoeff, int eob, int tx_size, int tx_type) {
  MACROBLOCK *x = (MACROBLOCK *)aom_calloc(1, sizeof(*x));
  uint16_t eobs[1] = { (uint16_t)eob };
  x->plane[0].qcoeff = (tran_low_t *)qcoeff;
  x->plane[0].eobs = eobs;
  const int r = av1_cost_coeffs_txb_estimate(x, 0, 0, (TX_SIZE)tx_size, (TX_TYPE)tx_type);
  aom_free(x);
  return r;
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
