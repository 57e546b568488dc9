/*
 * av1o_me.h -- ORACLE (test infrastructure only): motion search interface.
 *
 * Flat, self-contained description of ONE single-reference translational motion search,
 * carrying exactly the fields of the reference's FULLPEL_MOTION_SEARCH_PARAMS /
 * SUBPEL_MOTION_SEARCH_PARAMS (av1/encoder/mcomp.h:101-141,301-324) that
 * av1_full_pixel_search (mcomp.c:1755) and av1_find_best_sub_pixel_tree (mcomp.c:3128)
 * read.  The same struct drives
 *   - av1o_full_pixel_search / av1o_find_best_sub_pixel_tree  (oracle/av1o_mcomp.c), and
 *   - refshim_full_pixel_search / refshim_find_best_sub_pixel_tree (oracle/ref_shim.c,
 *     compiled only into oracle/_ref: builds the reference's real structs and calls the
 *     reference functions), which is how the oracle is pinned.
 */
#ifndef AV1O_ME_H_
#define AV1O_ME_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* search methods served (SEARCH_METHODS, av1/encoder/speed_features.h) */
enum { AV1O_NSTEP = 0, AV1O_NSTEP_8PT = 1, AV1O_DIAMOND = 2, AV1O_CLAMPED_DIAMOND = 3,
       /* pattern_search family; values as in the reference's enum (mcomp_structs.h:56-85) */
       AV1O_HEX = 4, AV1O_BIGDIA = 5, AV1O_SQUARE = 6, AV1O_FAST_HEX = 7, AV1O_FAST_DIAMOND = 8,
       AV1O_FAST_BIGDIA = 9, AV1O_VFAST_DIAMOND = 10 };
/* MV_COST_TYPE, mcomp.h:32-38 */
enum { AV1O_COST_ENTROPY = 0, AV1O_COST_L1_LOWRES, AV1O_COST_L1_MIDRES, AV1O_COST_L1_HDRES,
       AV1O_COST_NONE };
/* SUBPEL_SEARCH_TYPE (mcomp.h) */
enum { AV1O_USE_2_TAPS_ORIG = 0, AV1O_USE_2_TAPS, AV1O_USE_4_TAPS, AV1O_USE_8_TAPS };
/* SUBPEL_FORCE_STOP (mcomp.h:294-299) */
enum { AV1O_EIGHTH_PEL = 0, AV1O_QUARTER_PEL, AV1O_HALF_PEL, AV1O_FULL_PEL };

typedef struct {
  /* buffers: src = block in the current frame, ref = co-located position in the
   * reference frame (MV 0,0).  highbd: both point at uint16_t samples. */
  const void *src;
  const void *ref;
  int src_stride, ref_stride;
  int bw, bh; /* block size in pixels: one of the 22 SAD sizes */
  int highbd, bd;
  /* FullMvLimits (av1/common/mv.h:64-69), full-pel units */
  int col_min, col_max, row_min, row_max;
  /* full-pel search control */
  int search_method; /* AV1O_NSTEP ... */
  int step_param;
  int run_mesh_search, prune_mesh_search, mesh_search_mv_diff_threshold, force_mesh_thresh;
  int mesh_range[4], mesh_interval[4]; /* MESH_PATTERN[MAX_MESH_STEP] */
  int fine_search_interval;
  int use_skip_sad; /* ms_params->sdf = vfp->sdsf (row-skipping SAD) */
  /* MV_COST_PARAMS (mcomp.h:40-50) */
  int mv_cost_type;
  int ref_mv_row, ref_mv_col; /* 1/8 pel; full_ref_mv = get_fullmv_from_mv(ref_mv) */
  int sad_per_bit, error_per_bit;
  const int *mvjcost;  /* 4 entries, or NULL */
  const int *mvcost0;  /* centred: valid for index -MV_MAX..MV_MAX (rows) */
  const int *mvcost1;  /* centred (cols) */
  /* call arguments */
  int start_row, start_col; /* full-pel start mv */
  int want_cost_list, want_second_best;
  /* sub-pel control (SUBPEL_MOTION_SEARCH_PARAMS) */
  int allow_hp, forced_stop, iters_per_step, subpel_search_type;
  int sp_col_min, sp_col_max, sp_row_min, sp_row_max; /* SubpelMvLimits, 1/8 pel */
  int sp_start_row, sp_start_col;                     /* 1/8 pel start mv */
  /* which sub-pel function: 0 av1_find_best_sub_pixel_tree (mcomp.c:3128), 1 _pruned (:2992),
   * 2 _pruned_more (:2907); sp_have_cost_list = (ms_params->cost_list != NULL) */
  int sp_tree, sp_have_cost_list;
  int sp_cost_list[5];
  /* compound search (av1_set_ms_compound_refs, mcomp.c:80-93; MSBuffers mcomp.h:64-90): second_pred =
   * bw-pitch prediction from the other reference, mask = optional 6-bit blend weights.  Served by
   * av1_refining_search_8p_c and the three sub-pel trees -- the two searches of one iteration of
   * av1_joint_motion_search (motion_search_facade.c:594-652). */
  const void *second_pred;
  const uint8_t *mask;
  int mask_stride, inv_mask;
} Av1oMeParams;

typedef struct {
  int best_row, best_col;     /* full-pel result (or 1/8-pel for the sub-pel search) */
  int second_row, second_col; /* second_best_mv (INVALID_MV halves = -32768 when unset) */
  int cost_list[5];
  int ret;                    /* function return value (var / besterr) */
  int distortion;             /* sub-pel: *distortion */
  unsigned sse;               /* sub-pel: *sse1 */
} Av1oMeResult;

void av1o_full_pixel_search(const Av1oMeParams *p, Av1oMeResult *r);
void av1o_find_best_sub_pixel_tree(const Av1oMeParams *p, Av1oMeResult *r);
/* av1_find_best_sub_pixel_tree_pruned / _pruned_more, selected by p->sp_tree (1 / 2) */
void av1o_find_best_sub_pixel_tree_pruned(const Av1oMeParams *p, Av1oMeResult *r);
/* av1_refining_search_8p_c (mcomp.c:1683): full-pel 8-point refinement around start */
void av1o_refining_search_8p(const Av1oMeParams *p, Av1oMeResult *r);
/* av1_init_search_range (mcomp.c:239) */
int av1o_init_search_range(int size);

/* ---- temporal-filter motion search of a whole frame pair (tests/test_queue_tf_gpu.py) ----
 * Same layout as B200TfParams / B200TfResult (include/b200av1.h).  The reference side is
 * refshim_tf_motion_search_frame (oracle/ref_shim.c): the block loop of av1_tf_do_filtering_row
 * (temporal_filter.c:804-872) around a restatement of tf_motion_search's parameter set-up
 * (:88-260, the function is static and wants a whole AV1_COMP) that calls the reference's own
 * av1_set_mv_{row,col}_limits, av1_set_mv_search_range, av1_set_subpel_mv_search_range,
 * av1_full_pixel_search and av1_find_best_sub_pixel_tree*. */
typedef struct {
  int bw, bh, border_in_pixels, search_method, step_param, mv_cost_type;
  int mesh_range[4], mesh_interval[4];
  int force_mesh_thresh, prune_mesh_search, mesh_search_mv_diff_threshold, use_downsampled_sad;
  int sad_per_bit, error_per_bit, allow_hp, iters_per_step, subpel_tree, want_cost_list;
} Av1oTfParams;
typedef struct {
  int16_t block_mv[2];
  int32_t block_mse;
  int16_t sub_mv[4][2];
  int32_t sub_mse[4];
  int16_t ref_mv[2];
} Av1oTfResult;

/* ---- TPL: motion_estimation / txfm_quant_rdcost over a frame (tests/test_queue_tpl_gpu.py) ----
 * Layouts of B200TplParams / B200TplMeResult / B200TplRdResult (include/b200av1.h). */
typedef struct {
  int bw, bh, border_in_pixels, search_method, step_param, forced_stop, mv_cost_type;
  int mesh_range[4], mesh_interval[4];
  int force_mesh_thresh, prune_mesh_search, use_downsampled_sad;
  int sad_per_bit, error_per_bit, allow_hp, iters_per_step, subpel_tree, want_cost_list;
  const int *mvjcost;
  const int *mvcost[2];
} Av1oTplParams;
typedef struct {
  int16_t mv[2];
  uint32_t sme;
} Av1oTplMeResult;
typedef struct {
  int32_t rate_cost, eob;
  int64_t recon_error, sse;
} Av1oTplRdResult;

#ifdef __cplusplus
}
#endif
#endif /* AV1O_ME_H_ */
