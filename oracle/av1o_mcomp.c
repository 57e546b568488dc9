/*
 * av1o_mcomp.c -- ORACLE (test infrastructure only): single-reference translational
 * motion search.  Restates, from av1/encoder/mcomp.c of the reference:
 *   search-site tables           :369-404 (diamond), :452-496 (n-step)
 *   mv cost                      :269-350
 *   diamond_search_sad           :1318-1477     full_pixel_diamond      :1479-1526
 *   exhaustive_mesh_search       :1530-1599     full_pixel_exhaustive   :1603-1679
 *   calc_int_sad_list            :791-841       av1_full_pixel_search   :1755-1895
 *   pattern_search (hex / bigdia / square and their fast_* variants) :498-660,:883-1316
 *   sub-pel tree                 :2371-2551,2566-2779,2781-2868,3128-3196
 * The reference walks candidates one at a time updating a running best with strict '<';
 * this restatement evaluates the candidates of a step as a list and takes the first
 * minimum, which is the same decision and is the form the CUDA kernels use.
 * Pinned against the reference through oracle/ref_shim.c (tests/test_me_oracle.py).
 */
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#include "av1o.h"
#include "av1o_me.h"

#define MAX_STEPS 11            /* MAX_MVSEARCH_STEPS */
#define MAX_FULL_PEL_VAL 1023
#define MAX_FIRST_STEP 1024

typedef struct { int row, col; } Mv;
typedef struct {
  Mv site[2 * MAX_STEPS][17];
  int n_steps, per_step[2 * MAX_STEPS], radius[2 * MAX_STEPS];
} Sites;

int av1o_init_search_range(int size) {
  int sr = 0;
  if (size < 16) size = 16;
  while ((size << sr) < MAX_FULL_PEL_VAL) sr++;
  return sr < MAX_STEPS - 2 ? sr : MAX_STEPS - 2;
}

/* ---------------------------------------------------------------- site tables --- */
static void put_sites(Sites *s, int stage, int radius, int tan_radius, int npts) {
  const Mv m[13] = { { 0, 0 }, { -radius, 0 }, { radius, 0 }, { 0, -radius }, { 0, radius },
                     { -radius, -tan_radius }, { radius, tan_radius }, { -tan_radius, radius },
                     { tan_radius, -radius }, { -radius, tan_radius }, { radius, -tan_radius },
                     { tan_radius, radius }, { -tan_radius, -radius } };
  for (int i = 0; i <= npts; ++i) s->site[stage][i] = m[i];
  s->per_step[stage] = npts;
  s->radius[stage] = radius;
}

static void init_sites(Sites *s, int method) {
  memset(s, 0, sizeof(*s));
  if (method == AV1O_NSTEP || method == AV1O_NSTEP_8PT) {
    /* av1_init_motion_compensation_nstep: radius 1,2,3,5,8,12,... x1.5; 12 points beyond 5 */
    const int level = method == AV1O_NSTEP_8PT;
    const int stages = level ? 16 : 15;
    int radius = 1;
    for (int st = 0; st < stages; ++st) {
      int tanr = (int)(0.41 * radius);
      if (tanr < 1) tanr = 1;
      int npts = 12;
      if (radius <= 5 || level) {
        tanr = radius;
        npts = 8;
      }
      put_sites(s, st, radius, tanr, npts);
      if (st < 12) {
        const int grown = (int)(radius * 1.5 + 0.5);
        radius = grown > radius + 1 ? grown : radius + 1;
      }
    }
    s->n_steps = stages;
  } else {
    /* av1_init_dsmotion_compensation: 8 points, radius halves from 1024 (or 256) */
    const int level = method == AV1O_CLAMPED_DIAMOND;
    int stage = MAX_STEPS - 1, n = 0;
    for (int radius = level ? MAX_FIRST_STEP / 4 : MAX_FIRST_STEP; radius > 0;) {
      put_sites(s, stage, radius, radius, 8);
      if (!level || stage < 9) radius /= 2;
      --stage;
      ++n;
    }
    s->n_steps = n;
  }
}

/* ------------------------------------------------------------------ pixel ops --- */
static unsigned me_sad(const Av1oMeParams *p, int row, int col, int skip) {
  unsigned v;
  if (p->highbd) {
    const uint16_t *ref = (const uint16_t *)p->ref + row * p->ref_stride + col;
    v = skip ? av1o_highbd_sad_skip((const uint16_t *)p->src, p->src_stride, ref, p->ref_stride,
                                    p->bw, p->bh)
             : av1o_highbd_sad((const uint16_t *)p->src, p->src_stride, ref, p->ref_stride, p->bw,
                               p->bh);
    v >>= p->bd == 10 ? 2 : (p->bd == 12 ? 4 : 0); /* encoder_utils.h:153-208 */
  } else {
    const uint8_t *ref = (const uint8_t *)p->ref + row * p->ref_stride + col;
    v = skip ? av1o_sad_skip((const uint8_t *)p->src, p->src_stride, ref, p->ref_stride, p->bw,
                             p->bh)
             : av1o_sad((const uint8_t *)p->src, p->src_stride, ref, p->ref_stride, p->bw, p->bh);
  }
  return v;
}

/* the predictor of a compound search at mv (row, col) in 1/8 pel: reference sample (optionally
 * interpolated), blended with second_pred by average or mask */
static void compound_view(const Av1oMeParams *p, int fr, int fc, Av1oCompound *c) {
  memset(c, 0, sizeof(*c));
  c->w = p->bw, c->h = p->bh, c->highbd = p->highbd, c->bd = p->bd;
  c->pre = (const char *)p->ref + ((long)fr * p->ref_stride + fc) * (p->highbd ? 2 : 1);
  c->pre_stride = p->ref_stride;
  c->tgt = p->src, c->tgt_stride = p->src_stride;
  c->blend = p->mask ? AV1O_BLEND_MASK : AV1O_BLEND_AVG;
  c->second_pred = p->second_pred;
  c->mask = p->mask, c->mask_stride = p->mask_stride, c->invert_mask = p->inv_mask;
}

/* get_mvpred_compound_sad (mcomp.c:729): msdf / sdaf, with the encoder's >> 2 / >> 4 high-bit-depth
 * wrappers (encoder_utils.h MAKE_BFP_SADAVG_WRAPPER / MAKE_MBFP_COMPOUND_SAD_WRAPPER) */
static unsigned me_sad_compound(const Av1oMeParams *p, int row, int col) {
  Av1oCompound c;
  compound_view(p, row, col, &c);
  unsigned v = av1o_compound_sad(&c);
  if (p->highbd) v >>= p->bd == 10 ? 2 : (p->bd == 12 ? 4 : 0);
  return v;
}

static unsigned me_var(const Av1oMeParams *p, int row, int col, unsigned *sse) {
  if (p->highbd) {
    const uint16_t *ref = (const uint16_t *)p->ref + row * p->ref_stride + col;
    return av1o_highbd_variance((const uint16_t *)p->src, p->src_stride, ref, p->ref_stride, p->bw,
                                p->bh, p->bd, sse);
  }
  const uint8_t *ref = (const uint8_t *)p->ref + row * p->ref_stride + col;
  return av1o_variance((const uint8_t *)p->src, p->src_stride, ref, p->ref_stride, p->bw, p->bh,
                       sse);
}

/* ------------------------------------------------------------------- mv costs --- */
static int mv_joint(int row, int col) { return (row != 0) * 2 + (col != 0); }

static int mv_cost_tab(const Av1oMeParams *p, int drow, int dcol) {
  return p->mvjcost[mv_joint(drow, dcol)] + p->mvcost0[drow] + p->mvcost1[dcol];
}

/* mvsad_err_cost (:329): full-pel candidates, SAD domain */
static int sad_cost(const Av1oMeParams *p, int row, int col) {
  /* full_ref_mv = get_fullmv_from_mv(ref_mv): GET_MV_RAWPEL */
  const int fr = (p->ref_mv_row + 3 + (p->ref_mv_row >= 0)) >> 3;
  const int fc = (p->ref_mv_col + 3 + (p->ref_mv_col >= 0)) >> 3;
  const int dr = (int16_t)((row - fr) * 8), dc = (int16_t)((col - fc) * 8);
  switch (p->mv_cost_type) {
    case AV1O_COST_ENTROPY: {
      const unsigned c = (unsigned)mv_cost_tab(p, dr, dc) * (unsigned)p->sad_per_bit;
      return (int)((c + 256) >> 9);
    }
    case AV1O_COST_L1_LOWRES: return (32 * (abs(dr) + abs(dc))) >> 3;
    case AV1O_COST_L1_MIDRES: return (15 * (abs(dr) + abs(dc))) >> 3;
    case AV1O_COST_L1_HDRES: return (8 * (abs(dr) + abs(dc))) >> 3;
    default: return 0;
  }
}

/* mv_err_cost (:290): 1/8-pel mv, variance domain */
static int err_cost(const Av1oMeParams *p, int row, int col) {
  const int dr = (int16_t)(row - p->ref_mv_row), dc = (int16_t)(col - p->ref_mv_col);
  switch (p->mv_cost_type) {
    case AV1O_COST_ENTROPY:
      if (!p->mvcost0) return 0;
      /* RDDIV_BITS + AV1_PROB_COST_SHIFT - RD_EPB_SHIFT + PIXEL_TRANSFORM_ERROR_SCALE = 14 */
      return (int)(((int64_t)mv_cost_tab(p, dr, dc) * p->error_per_bit + 8192) >> 14);
    case AV1O_COST_L1_LOWRES: return (2 * (abs(dr) + abs(dc))) >> 3;
    case AV1O_COST_L1_MIDRES: return 0;
    case AV1O_COST_L1_HDRES: return (1 * (abs(dr) + abs(dc))) >> 3;
    default: return 0;
  }
}

static int in_range(const Av1oMeParams *p, int row, int col) {
  return col >= p->col_min && col <= p->col_max && row >= p->row_min && row <= p->row_max;
}
static Mv clamp_mv(const Av1oMeParams *p, Mv m) {
  if (m.col < p->col_min) m.col = p->col_min;
  if (m.col > p->col_max) m.col = p->col_max;
  if (m.row < p->row_min) m.row = p->row_min;
  if (m.row > p->row_max) m.row = p->row_max;
  return m;
}

/* get_mvpred_var_cost (:665): variance at a full-pel mv + sub-pel-domain mv cost */
static int var_cost(const Av1oMeParams *p, Mv m) {
  unsigned sse;
  return (int)me_var(p, m.row, m.col, &sse) + err_cost(p, m.row * 8, m.col * 8);
}

/* ------------------------------------------------------------- diamond search --- */
static int diamond_search(const Av1oMeParams *p, const Sites *cfg, int skip, Mv start,
                          int search_step, int *num00, Mv *best, Mv *second) {
  int off_center = 0, center_steps = 0;
  start = clamp_mv(p, start);
  const int tot = cfg->n_steps - search_step;
  Mv tmp_second = second ? *second : start;
  *best = start;
  unsigned bestsad = (unsigned)sad_cost(p, best->row, best->col) + me_sad(p, start.row, start.col, skip);
  for (int step = tot - 1; step >= 0; --step) {
    const Mv *site = cfg->site[step];
    const int n = cfg->per_step[step];
    int best_site = 0;
    for (int i = 1; i <= n; ++i) {
      const int r = best->row + site[i].row, c = best->col + site[i].col;
      if (!in_range(p, r, c)) continue;
      const unsigned s = me_sad(p, r, c, skip);
      if (s < bestsad) {
        const unsigned t = s + (unsigned)sad_cost(p, r, c);
        if (t < bestsad) {
          bestsad = t;
          best_site = i;
        }
      }
    }
    if (best_site) {
      tmp_second = *best;
      best->row += site[best_site].row;
      best->col += site[best_site].col;
      off_center = 1;
    }
    if (!off_center) center_steps++;
    if (best_site == 0 && step > 2) {
      int next = cfg->radius[step - 1];
      while (next == cfg->radius[step] && step > 2) {
        center_steps++;
        --step;
        next = cfg->radius[step - 1];
      }
    }
  }
  *num00 = center_steps;
  if (second) *second = tmp_second;
  return (int)bestsad;
}

/* calc_int_sad_list (:791) with costlist_has_sad == 0 */
static void sad_list(const Av1oMeParams *p, int skip, Mv best, int *cl) {
  static const Mv nb[4] = { { 0, -1 }, { 1, 0 }, { 0, 1 }, { -1, 0 } };
  cl[0] = (int)me_sad(p, best.row, best.col, skip) + sad_cost(p, best.row, best.col);
  for (int i = 0; i < 4; ++i) {
    const int r = best.row + nb[i].row, c = best.col + nb[i].col;
    if (!in_range(p, r, c)) cl[i + 1] = INT_MAX;
    else cl[i + 1] = (int)me_sad(p, r, c, skip) + sad_cost(p, r, c);
  }
}

static int full_pixel_diamond(const Av1oMeParams *p, const Sites *cfg, int skip, Mv start,
                              int step_param, int *cl, Mv *best, Mv *second) {
  int n, num00 = 0;
  diamond_search(p, cfg, skip, start, step_param, &n, best, second);
  int bestsme = var_cost(p, *best);
  const int further = cfg->n_steps - 1 - step_param;
  while (n < further) {
    ++n;
    Mv tmp;
    diamond_search(p, cfg, skip, start, step_param + n, &num00, &tmp, second);
    const int sme = var_cost(p, tmp);
    if (sme < bestsme) {
      bestsme = sme;
      *best = tmp;
    }
    if (num00) {
      n += num00;
      num00 = 0;
    }
  }
  if (cl) sad_list(p, skip, *best, cl);
  return bestsme;
}

/* ------------------------------------------------------------- pattern search --- */
/* Site tables of av1_init_motion_compensation_{hex,bigdia,square} (mcomp.c:498-660), built
 * by rule: scale i reaches 2^i pixels.  hex: 8-point ring at scale 0, then 6-point hexagons;
 * bigdia: 4 nearest points at scale 0, then 8-point diamonds; square: 8-point squares. */
#define PAT_SCALES 11 /* MAX_PATTERN_SCALES */
typedef struct {
  Mv site[PAT_SCALES][8];
  int n[PAT_SCALES];
} PatSites;

static void init_pattern_sites(PatSites *s, int method) {
  static const Mv ring[8] = { { -1, -1 }, { 0, -1 }, { 1, -1 }, { 1, 0 }, { 1, 1 }, { 0, 1 },
                              { -1, 1 }, { -1, 0 } };
  memset(s, 0, sizeof(*s));
  const int hex = method == AV1O_HEX || method == AV1O_FAST_HEX;
  const int square = method == AV1O_SQUARE;
  for (int i = 0; i < PAT_SCALES; ++i) {
    const int d = 1 << i, h = d >> 1;
    if (square || (hex && i == 0)) {
      s->n[i] = 8;
      for (int j = 0; j < 8; ++j) s->site[i][j] = (Mv){ ring[j].row * d, ring[j].col * d };
    } else if (hex) {
      const Mv m[6] = { { -h, -d }, { h, -d }, { d, 0 }, { h, d }, { -h, d }, { -d, 0 } };
      s->n[i] = 6;
      for (int j = 0; j < 6; ++j) s->site[i][j] = m[j];
    } else if (i == 0) {
      const Mv m[4] = { { 0, -1 }, { 1, 0 }, { 0, 1 }, { -1, 0 } };
      s->n[i] = 4;
      for (int j = 0; j < 4; ++j) s->site[i][j] = m[j];
    } else {
      const Mv m[8] = { { -h, -h }, { 0, -d }, { h, -h }, { d, 0 }, { h, h }, { 0, d }, { -h, h },
                        { -d, 0 } };
      s->n[i] = 8;
      for (int j = 0; j < 8; ++j) s->site[i][j] = m[j];
    }
  }
}

typedef struct {
  const Av1oMeParams *p;
  const PatSites *cfg;
  int skip, br, bc;
  unsigned bestsad, raw_bestsad;
} PatState;

/* One round of candidates idx[0..n) of scale s around the centre (calc_sad4/3/_update_bestmv
 * and their bounds-checking twins, :883-1011): candidates outside the limits are skipped,
 * the others update the running best in order.  Returns the position in idx[] of the last
 * improvement or -1; cl (optional) receives the raw SADs at cl[idx + 1]. */
static int pat_round(PatState *st, int s, const int *idx, int n, int *cl, int cl_oob_marks) {
  int best = -1;
  for (int i = 0; i < n; ++i) {
    const Mv m = { st->br + st->cfg->site[s][idx[i]].row, st->bc + st->cfg->site[s][idx[i]].col };
    if (!in_range(st->p, m.row, m.col)) {
      if (cl && cl_oob_marks) cl[idx[i] + 1] = INT_MAX;
      continue;
    }
    const unsigned sad = me_sad(st->p, m.row, m.col, st->skip);
    if (cl) cl[idx[i] + 1] = (int)sad;
    if (sad >= st->bestsad) continue; /* update_mvs_and_sad :858 */
    const unsigned t = sad + (unsigned)sad_cost(st->p, m.row, m.col);
    if (t < st->bestsad) {
      st->raw_bestsad = sad;
      st->bestsad = t;
      best = i;
    }
  }
  return best;
}

/* All candidates of a scale.  Quirk kept from the reference (:1065-1078): when the whole
 * scale is inside the limits (check_bounds) it evaluates groups of four and then calls the
 * scalar routine with (num_candidates = n % 4, cand_start = 4 * (n / 4)), whose loop runs
 * from cand_start up to n % 4 -- i.e. never: the last two points of a 6-point hexagon are
 * only looked at near the limits. */
static int pat_round_all(PatState *st, int s) {
  static const int all[8] = { 0, 1, 2, 3, 4, 5, 6, 7 };
  const Av1oMeParams *p = st->p;
  const int range = 1 << s;
  const int inside = st->br - range >= p->row_min && st->br + range <= p->row_max &&
                     st->bc - range >= p->col_min && st->bc + range <= p->col_max;
  return pat_round(st, s, all, inside ? st->cfg->n[s] & ~3 : st->cfg->n[s], NULL, 0);
}

static void pat_move(PatState *st, int s, int k) {
  st->br += st->cfg->site[s][k].row;
  st->bc += st->cfg->site[s][k].col;
}

/* pattern_search (mcomp.c:1017-1243) */
static int pattern_search(const Av1oMeParams *p, int skip, Mv start, int search_step, int do_init,
                          int *cl, Mv *best_mv) {
  static const int all[8] = { 0, 1, 2, 3, 4, 5, 6, 7 };
  PatSites cfg;
  init_pattern_sites(&cfg, p->search_method);
  const int last_is_4 = cfg.n[0] == 4;
  if (search_step > MAX_STEPS - 1) search_step = MAX_STEPS - 1;
  int best_init_s = MAX_STEPS - 1 - search_step, s, k = -1;
  start = clamp_mv(p, start);
  PatState st = { p, &cfg, skip, start.row, start.col, 0, 0 };
  if (cl) for (int i = 0; i < 5; ++i) cl[i] = INT_MAX;
  int has_sad = 0;
  st.raw_bestsad = me_sad(p, start.row, start.col, skip);
  st.bestsad = st.raw_bestsad + (unsigned)sad_cost(p, start.row, start.col);
  if (do_init) {
    s = best_init_s;
    best_init_s = -1;
    for (int t = 0; t <= s; ++t) {
      const int bs = pat_round_all(&st, t);
      if (bs != -1) {
        best_init_s = t;
        k = bs;
      }
    }
    if (best_init_s != -1) pat_move(&st, best_init_s, k);
  }
  if (best_init_s != -1) {
    const int last_s = last_is_4 && cl != NULL;
    int best_site = -1;
    for (s = best_init_s; s >= last_s; s--) {
      if (!do_init || s != best_init_s) {
        best_site = pat_round_all(&st, s);
        if (best_site == -1) continue;
        pat_move(&st, s, best_site);
        k = best_site;
      }
      do {
        const int chk[3] = { k == 0 ? cfg.n[s] - 1 : k - 1, k, k == cfg.n[s] - 1 ? 0 : k + 1 };
        best_site = pat_round(&st, s, chk, 3, NULL, 0);
        if (best_site != -1) {
          k = chk[best_site];
          pat_move(&st, s, k);
        }
      } while (best_site != -1);
    }
    if (s == 0) { /* only reachable with a cost list: the 4-point scale fills it on the way */
      cl[0] = (int)st.raw_bestsad;
      has_sad = 1;
      if (!do_init || s != best_init_s) {
        best_site = pat_round(&st, 0, all, 4, cl, 0);
        if (best_site != -1) {
          pat_move(&st, 0, best_site);
          k = best_site;
        }
      }
      while (best_site != -1) {
        const int chk[3] = { k == 0 ? 3 : k - 1, k, k == 3 ? 0 : k + 1 };
        cl[1] = cl[2] = cl[3] = cl[4] = INT_MAX;
        cl[((k + 2) % 4) + 1] = cl[0];
        cl[0] = (int)st.raw_bestsad;
        best_site = pat_round(&st, 0, chk, 3, cl, 1);
        if (best_site != -1) {
          k = chk[best_site];
          pat_move(&st, 0, k);
        }
      }
    }
  }
  best_mv->row = st.br;
  best_mv->col = st.bc;
  if (cl) {
    if (!has_sad) {
      sad_list(p, skip, *best_mv, cl);
    } else { /* calc_int_sad_list :791 with costlist_has_sad: only the mv costs are added */
      static const Mv nb[4] = { { 0, -1 }, { 1, 0 }, { 0, 1 }, { -1, 0 } };
      cl[0] += sad_cost(p, st.br, st.bc);
      for (int i = 0; i < 4; ++i)
        if (cl[i + 1] != INT_MAX) cl[i + 1] += sad_cost(p, st.br + nb[i].row, st.bc + nb[i].col);
    }
  }
  return var_cost(p, *best_mv);
}

/* ---------------------------------------------------------------- mesh search --- */
static void try_mv(const Av1oMeParams *p, unsigned sad, Mv m, unsigned *best_sad, Mv *best,
                   Mv *second) {
  if (sad >= *best_sad) return; /* update_mvs_and_sad :858 */
  const unsigned t = sad + (unsigned)sad_cost(p, m.row, m.col);
  if (t < *best_sad) {
    *best_sad = t;
    if (second) *second = *best;
    *best = m;
  }
}

static int mesh_search(const Av1oMeParams *p, int skip, Mv start, int range, int step, Mv *best,
                       Mv *second) {
  start = clamp_mv(p, start);
  *best = start;
  unsigned best_sad = me_sad(p, start.row, start.col, skip) + (unsigned)sad_cost(p, start.row, start.col);
  const int r0 = -range > p->row_min - start.row ? -range : p->row_min - start.row;
  const int c0 = -range > p->col_min - start.col ? -range : p->col_min - start.col;
  const int r1 = range < p->row_max - start.row ? range : p->row_max - start.row;
  const int c1 = range < p->col_max - start.col ? range : p->col_max - start.col;
  for (int r = r0; r <= r1; r += step) {
    if (step > 1) {
      for (int c = c0; c <= c1; c += step) {
        const Mv m = { start.row + r, start.col + c };
        try_mv(p, me_sad(p, m.row, m.col, skip), m, &best_sad, best, second);
      }
    } else {
      /* groups of four; the tail loop stops one short of end_col (mcomp.c:1586) */
      for (int c = c0; c <= c1; c += 4) {
        const int cnt = c + 3 <= c1 ? 4 : c1 - c;
        for (int i = 0; i < cnt; ++i) {
          const Mv m = { start.row + r, start.col + c + i };
          try_mv(p, me_sad(p, m.row, m.col, skip), m, &best_sad, best, second);
        }
      }
    }
  }
  return (int)best_sad;
}

static int full_pixel_exhaustive(const Av1oMeParams *p, int skip, Mv start, int *cl, Mv *best,
                                 Mv *second) {
  int interval = p->mesh_interval[0], range = p->mesh_range[0];
  *best = start;
  if (range < 7 || range > 256 || interval < 1 || interval > range) return INT_MAX;
  const int divisor = range / interval;
  const int mag = abs(best->row) > abs(best->col) ? abs(best->row) : abs(best->col);
  if ((5 * mag) / 4 > range) range = (5 * mag) / 4;
  if (range > 256) range = 256;
  if (range / divisor > interval) interval = range / divisor;
  if (p->fine_search_interval && interval > 4) interval = 4;
  int bestsme = mesh_search(p, skip, *best, range, interval, best, second);
  if (interval > 1 && range > 7) {
    for (int i = 1; i < 4; ++i) {
      bestsme = mesh_search(p, skip, *best, p->mesh_range[i], p->mesh_interval[i], best, second);
      if (p->mesh_interval[i] == 1) break;
    }
  }
  if (bestsme < INT_MAX) bestsme = var_cost(p, *best);
  if (cl) sad_list(p, skip, *best, cl);
  return bestsme;
}

static int ilog2i(int v) {
  int l = 0;
  while ((1 << (l + 1)) <= v) ++l;
  return l;
}

static int full_pixel_search_impl(const Av1oMeParams *p, const Sites *cfg, int skip,
                                  Av1oMeResult *r, int *cl, Mv *best, Mv *second) {
  const Mv start = { p->start_row, p->start_col };
  int run_mesh = p->run_mesh_search;
  best->row = best->col = -32768; /* MARK_MV_INVALID */
  if (second) second->row = second->col = -32768;
  if (cl) for (int i = 0; i < 5; ++i) cl[i] = INT_MAX;
  int var;
  const int big = MAX_STEPS; /* fast_* wrappers clamp the step from below, :1278-1316 */
  switch (p->search_method) {
    case AV1O_HEX: case AV1O_BIGDIA: case AV1O_SQUARE:
      var = pattern_search(p, skip, start, p->step_param, 1, cl, best);
      break;
    case AV1O_FAST_HEX: case AV1O_FAST_DIAMOND:
      var = pattern_search(p, skip, start, p->step_param > big - 2 ? p->step_param : big - 2, 0, cl, best);
      break;
    case AV1O_FAST_BIGDIA:
      var = pattern_search(p, skip, start, p->step_param > big - 3 ? p->step_param : big - 3, 0, cl, best);
      break;
    case AV1O_VFAST_DIAMOND:
      var = pattern_search(p, skip, start, p->step_param > big - 1 ? p->step_param : big - 1, 0, cl, best);
      break;
    default:
      var = full_pixel_diamond(p, cfg, skip, start, p->step_param, cl, best, second);
  }
  const int mi_log2 = ilog2i(p->bw / 4) + ilog2i(p->bh / 4);
  if (!run_mesh && (p->search_method == AV1O_NSTEP || p->search_method == AV1O_NSTEP_8PT)) {
    const int thr = p->force_mesh_thresh >> (10 - mi_log2);
    if (var > thr) run_mesh = 1;
  }
  if (p->prune_mesh_search) {
    const int dr = abs(start.row - best->row), dc = abs(start.col - best->col);
    if ((dr > dc ? dr : dc) <= p->mesh_search_mv_diff_threshold) run_mesh = 0;
  }
  if (skip) {
    /* row-skipping SAD was used: verify against the full SAD, redo without skipping when
     * they disagree by >= 90 % (mcomp.c:1840-1873) */
    const int sad = (int)me_sad(p, best->row, best->col, 0);
    const int ssad = (int)me_sad(p, best->row, best->col, 1);
    if (sad > (1 << mi_log2) && abs(ssad - sad) * 10 >= (sad > 1 ? sad : 1) * 9)
      return full_pixel_search_impl(p, cfg, 0, r, cl, best, second);
  }
  if (run_mesh) {
    Mv tmp;
    const int var_ex = full_pixel_exhaustive(p, skip, *best, cl, &tmp, second);
    if (var_ex < var) {
      var = var_ex;
      *best = tmp;
    }
  }
  return var;
}

void av1o_full_pixel_search(const Av1oMeParams *p, Av1oMeResult *r) {
  Sites *cfg = (Sites *)malloc(sizeof(Sites));
  init_sites(cfg, p->search_method);
  Mv best, second;
  int cl[5];
  memset(r, 0, sizeof(*r));
  r->ret = full_pixel_search_impl(p, cfg, p->use_skip_sad, r, p->want_cost_list ? cl : NULL, &best,
                                  p->want_second_best ? &second : NULL);
  r->best_row = best.row;
  r->best_col = best.col;
  r->second_row = p->want_second_best ? second.row : -32768;
  r->second_col = p->want_second_best ? second.col : -32768;
  if (p->want_cost_list) memcpy(r->cost_list, cl, sizeof(cl));
  free(cfg);
}

/* ================================================================== sub-pel ===== */
typedef struct {
  const Av1oMeParams *p;
  Mv best;
  unsigned besterr, sse;
  int distortion;
} SubpelState;

/* estimated_pref_error (:2371, bilinear svf) / upsampled_pref_error (:2402) */
static unsigned pref_error(const Av1oMeParams *p, Mv mv, int upsampled, unsigned *sse) {
  const int fr = mv.row >> 3, fc = mv.col >> 3, sx = mv.col & 7, sy = mv.row & 7;
  const int w = p->bw, h = p->bh;
  unsigned v;
  if (p->second_pred) {
    /* compound: svaf / msvf (bilinear estimate), or aom_comp_avg_upsampled_pred /
     * aom_comp_mask_upsampled_pred (reconinter_enc.c:498-560) followed by vf */
    Av1oCompound c;
    if (!upsampled) {
      compound_view(p, fr, fc, &c);
      c.subpel = 1, c.xoff = sx, c.yoff = sy;
      return av1o_compound_variance(&c, sse);
    }
    const int taps = p->subpel_search_type == AV1O_USE_2_TAPS ? 2
                     : (p->subpel_search_type == AV1O_USE_4_TAPS ? 4 : 8);
    void *up = malloc((size_t)(w * h) * 2);
    if (p->highbd)
      av1o_highbd_upsampled_pred((uint16_t *)up, w, h, sx, sy, (const uint16_t *)p->ref + fr * p->ref_stride + fc,
                                 p->ref_stride, p->bd, taps);
    else
      av1o_upsampled_pred((uint8_t *)up, w, h, sx, sy, (const uint8_t *)p->ref + fr * p->ref_stride + fc,
                          p->ref_stride, taps);
    compound_view(p, 0, 0, &c);
    c.pre = up, c.pre_stride = w;
    v = av1o_compound_variance(&c, sse);
    free(up);
    return v;
  }
  if (p->highbd) {
    const uint16_t *ref = (const uint16_t *)p->ref + fr * p->ref_stride + fc;
    const uint16_t *src = (const uint16_t *)p->src;
    if (!upsampled)
      return av1o_highbd_sub_pixel_variance(ref, p->ref_stride, sx, sy, src, p->src_stride, w, h,
                                            p->bd, sse);
    uint16_t *pred = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(w * h));
    const int taps = p->subpel_search_type == AV1O_USE_2_TAPS ? 2
                     : (p->subpel_search_type == AV1O_USE_4_TAPS ? 4 : 8);
    av1o_highbd_upsampled_pred(pred, w, h, sx, sy, ref, p->ref_stride, p->bd, taps);
    v = av1o_highbd_variance(pred, w, src, p->src_stride, w, h, p->bd, sse);
    free(pred);
  } else {
    const uint8_t *ref = (const uint8_t *)p->ref + fr * p->ref_stride + fc;
    const uint8_t *src = (const uint8_t *)p->src;
    if (!upsampled)
      return av1o_sub_pixel_variance(ref, p->ref_stride, sx, sy, src, p->src_stride, w, h, sse);
    uint8_t *pred = (uint8_t *)malloc((size_t)(w * h));
    const int taps = p->subpel_search_type == AV1O_USE_2_TAPS ? 2
                     : (p->subpel_search_type == AV1O_USE_4_TAPS ? 4 : 8);
    av1o_upsampled_pred(pred, w, h, sx, sy, ref, p->ref_stride, taps);
    v = av1o_variance(pred, w, src, p->src_stride, w, h, sse);
    free(pred);
  }
  return v;
}

/* check_better (:2526) / check_better_fast (:2496) for unscaled references */
static unsigned check_better(SubpelState *s, Mv mv, int upsampled, int *better) {
  const Av1oMeParams *p = s->p;
  if (mv.col < p->sp_col_min || mv.col > p->sp_col_max || mv.row < p->sp_row_min ||
      mv.row > p->sp_row_max)
    return INT_MAX;
  unsigned sse;
  const int mse = (int)pref_error(p, mv, upsampled, &sse);
  const unsigned cost = (unsigned)err_cost(p, mv.row, mv.col) + (unsigned)mse;
  if (cost < s->besterr) {
    s->besterr = cost;
    s->best = mv;
    s->distortion = mse;
    s->sse = sse;
    if (better) *better |= 1;
  }
  return cost;
}

/* first_level_check (:2689) / first_level_check_fast (:2566): 4 cardinals + best diagonal */
static Mv first_level(SubpelState *s, Mv c, int hstep, int upsampled) {
  const Mv l = { c.row, c.col - hstep }, r = { c.row, c.col + hstep };
  const Mv t = { c.row - hstep, c.col }, b = { c.row + hstep, c.col };
  const unsigned left = check_better(s, l, upsampled, NULL);
  const unsigned right = check_better(s, r, upsampled, NULL);
  const unsigned up = check_better(s, t, upsampled, NULL);
  const unsigned down = check_better(s, b, upsampled, NULL);
  const Mv diag = { up <= down ? -hstep : hstep, left <= right ? -hstep : hstep };
  const Mv d = { c.row + diag.row, c.col + diag.col };
  check_better(s, d, upsampled, NULL);
  return diag;
}

/* second_level_check_v2 (:2728) */
static void second_level(SubpelState *s, Mv c, Mv diag, int upsampled) {
  if (c.row == s->best.row && c.col == s->best.col) return;
  if (c.row == s->best.row) diag.row = -diag.row;
  else if (c.col == s->best.col) diag.col = -diag.col;
  const Mv rb = { s->best.row + diag.row, s->best.col };
  const Mv cb = { s->best.row, s->best.col + diag.col };
  const Mv db = { s->best.row + diag.row, s->best.col + diag.col };
  int better = 0;
  check_better(s, rb, upsampled, &better);
  check_better(s, cb, upsampled, &better);
  if (better) check_better(s, db, upsampled, &better);
}

/* av1_find_best_sub_pixel_tree (:3128), last_mv_search_list == NULL, unscaled */
void av1o_find_best_sub_pixel_tree(const Av1oMeParams *p, Av1oMeResult *r) {
  SubpelState s;
  memset(r, 0, sizeof(*r));
  s.p = p;
  s.best.row = p->sp_start_row;
  s.best.col = p->sp_start_col;
  const int upsampled = p->subpel_search_type != AV1O_USE_2_TAPS_ORIG;
  const int a = AV1O_FULL_PEL - p->forced_stop, b = 3 - !p->allow_hp;
  const int round = a < b ? a : b;
  int hstep = 4; /* INIT_SUBPEL_STEP_SIZE */
  /* (upsampled_)setup_center_error :2781,:2843 */
  unsigned sse;
  unsigned e;
  if (upsampled) {
    e = pref_error(p, s.best, 1, &sse);
  } else {
    /* vf(y, y_stride, src, src_stride): reference first -- the order matters at 10/12 bit
     * because the rounding of the (signed) sum is not symmetric (variance.c:361-379) */
    const int fr = s.best.row >> 3, fc = s.best.col >> 3;
    if (p->second_pred) { /* setup_center_error :2781: aom_comp_avg_pred / aom_comp_mask_pred, then vf */
      Av1oCompound c;
      compound_view(p, fr, fc, &c);
      e = av1o_compound_variance(&c, &sse);
    } else if (p->highbd)
      e = av1o_highbd_variance((const uint16_t *)p->ref + fr * p->ref_stride + fc, p->ref_stride,
                               (const uint16_t *)p->src, p->src_stride, p->bw, p->bh, p->bd, &sse);
    else
      e = av1o_variance((const uint8_t *)p->ref + fr * p->ref_stride + fc, p->ref_stride,
                        (const uint8_t *)p->src, p->src_stride, p->bw, p->bh, &sse);
  }
  s.distortion = (int)e;
  s.sse = sse;
  s.besterr = e + (unsigned)err_cost(p, s.best.row, s.best.col);
  for (int iter = 0; iter < round; ++iter) {
    const Mv center = s.best;
    const Mv diag = first_level(&s, center, hstep, upsampled);
    if (!(center.row == s.best.row && center.col == s.best.col) && p->iters_per_step > 1)
      second_level(&s, center, diag, upsampled);
    hstep >>= 1;
  }
  r->best_row = s.best.row;
  r->best_col = s.best.col;
  r->ret = (int)s.besterr;
  r->distortion = s.distortion;
  r->sse = s.sse;
}

/* ============================================== pruned sub-pel trees (a24) ===== */
/* second_level_check_fast (:2610): two or three more bilinear probes in the best quadrant */
static void second_level_fast(SubpelState *s, Mv c, Mv diag, int hstep) {
  const int tr = c.row, tc = c.col, br = s->best.row, bc = s->best.col;
  if (tr != br && tc != bc) {
    const Mv c1 = { br, bc + diag.col }, c2 = { br + diag.row, bc };
    check_better(s, c1, 0, NULL);
    check_better(s, c2, 0, NULL);
  } else if (tr == br && tc != bc) {
    const Mv bl = { br + hstep, bc + diag.col }, tl = { br - hstep, bc + diag.col };
    const Mv rev = { br - diag.row, bc };
    check_better(s, bl, 0, NULL);
    check_better(s, tl, 0, NULL);
    check_better(s, rev, 0, NULL);
  } else if (tr != br && tc == bc) {
    const Mv rl = { br + diag.row, bc + hstep }, ll = { br + diag.row, bc - hstep };
    const Mv rev = { br, bc - diag.col };
    check_better(s, rl, 0, NULL);
    check_better(s, ll, 0, NULL);
    check_better(s, rev, 0, NULL);
  }
}

/* two_level_checks_fast (:2675) */
static void two_level_fast(SubpelState *s, Mv c, int hstep, int iters) {
  const Mv diag = first_level(s, c, hstep, 0);
  if (iters > 1) second_level_fast(s, c, diag, hstep);
}

/* divide_and_round (:2853) */
static int div_round(int n, int d) { return ((n < 0) ^ (d < 0)) ? ((n - d / 2) / d) : ((n + d / 2) / d); }

/* av1_find_best_sub_pixel_tree_pruned (:2992) for sp_tree == 1, _pruned_more (:2907) for
 * sp_tree == 2; last_mv_search_list == NULL, unscaled references: the centre error is
 * setup_center_error (vf on the full-pel block) and every probe is the bilinear svf. */
void av1o_find_best_sub_pixel_tree_pruned(const Av1oMeParams *p, Av1oMeResult *r) {
  SubpelState s;
  memset(r, 0, sizeof(*r));
  s.p = p;
  s.best.row = p->sp_start_row;
  s.best.col = p->sp_start_col;
  const Mv start = s.best;
  int hstep = 4;
  unsigned sse, e;
  {
    const int fr = s.best.row >> 3, fc = s.best.col >> 3;
    if (p->second_pred) { /* setup_center_error :2781: aom_comp_avg_pred / aom_comp_mask_pred, then vf */
      Av1oCompound c;
      compound_view(p, fr, fc, &c);
      e = av1o_compound_variance(&c, &sse);
    } else if (p->highbd)
      e = av1o_highbd_variance((const uint16_t *)p->ref + fr * p->ref_stride + fc, p->ref_stride,
                               (const uint16_t *)p->src, p->src_stride, p->bw, p->bh, p->bd, &sse);
    else
      e = av1o_variance((const uint8_t *)p->ref + fr * p->ref_stride + fc, p->ref_stride,
                        (const uint8_t *)p->src, p->src_stride, p->bw, p->bh, &sse);
  }
  s.distortion = (int)e;
  s.sse = sse;
  s.besterr = e + (unsigned)err_cost(p, s.best.row, s.best.col);
  if (p->forced_stop != AV1O_FULL_PEL) {
    const int *cl = p->sp_cost_list;
    const int usable = p->sp_have_cost_list && cl[0] != INT_MAX && cl[1] != INT_MAX &&
                       cl[2] != INT_MAX && cl[3] != INT_MAX && cl[4] != INT_MAX;
    if (p->sp_tree == 2 && usable && cl[0] < cl[1] && cl[0] < cl[2] && cl[0] < cl[3] &&
        cl[0] < cl[4]) {
      /* get_cost_surf_min (:2870) with bits = 1 */
      const int ic = div_round((cl[1] - cl[3]) * 1, cl[1] - 2 * cl[0] + cl[3]);
      const int ir = div_round((cl[4] - cl[2]) * 1, cl[4] - 2 * cl[0] + cl[2]);
      if (ir != 0 || ic != 0) {
        const Mv m = { start.row + ir * hstep, start.col + ic * hstep };
        check_better(&s, m, 0, NULL);
      }
    } else if (p->sp_tree == 1 && usable) {
      const int right = !(cl[1] < cl[3]), top = !(cl[2] < cl[4]);
      const Mv h = { start.row, start.col + (right ? hstep : -hstep) };
      const Mv v = { start.row + (top ? -hstep : hstep), start.col };
      const Mv d = { v.row, h.col };
      check_better(&s, h, 0, NULL);
      check_better(&s, v, 0, NULL);
      check_better(&s, d, 0, NULL);
    } else {
      two_level_fast(&s, start, hstep, p->iters_per_step);
    }
    if (p->forced_stop < AV1O_HALF_PEL) {
      hstep >>= 1;
      two_level_fast(&s, s.best, hstep, p->iters_per_step);
    }
    if (p->allow_hp && p->forced_stop == AV1O_EIGHTH_PEL) {
      hstep >>= 1;
      two_level_fast(&s, s.best, hstep, p->iters_per_step);
    }
  }
  r->best_row = s.best.row;
  r->best_col = s.best.col;
  r->ret = (int)s.besterr;
  r->distortion = s.distortion;
  r->sse = s.sse;
}

/* ===================================== av1_refining_search_8p_c (:1683) ===== */
/* get_mvpred_compound_sad (:729): sdf for a single reference, sdaf / msdf with second_pred / mask */
void av1o_refining_search_8p(const Av1oMeParams *p, Av1oMeResult *r) {
  static const int nb[8][2] = { { -1, 0 }, { 0, -1 }, { 0, 1 }, { 1, 0 },
                                { -1, -1 }, { 1, -1 }, { -1, 1 }, { 1, 1 } };
  enum { RANGE = 3, STRIDE = 7 };
  unsigned char grid[STRIDE * STRIDE];
  memset(grid, 0, sizeof(grid));
  memset(r, 0, sizeof(*r));
  int centre = RANGE * STRIDE + RANGE;
  Mv start = { p->start_row, p->start_col };
  Mv best = clamp_mv(p, start);
  unsigned best_sad = p->second_pred ? me_sad_compound(p, best.row, best.col)
                                     : me_sad(p, best.row, best.col, p->use_skip_sad);
  best_sad += (unsigned)sad_cost(p, best.row, best.col);
  grid[centre] = 1;
  for (int i = 0; i < RANGE; ++i) {
    int best_site = -1;
    for (int j = 0; j < 8; ++j) {
      const int g = centre + nb[j][0] * STRIDE + nb[j][1];
      if (grid[g]) continue;
      grid[g] = 1;
      const Mv m = { best.row + nb[j][0], best.col + nb[j][1] };
      if (m.col < p->col_min || m.col > p->col_max || m.row < p->row_min || m.row > p->row_max)
        continue;
      unsigned sad = p->second_pred ? me_sad_compound(p, m.row, m.col) : me_sad(p, m.row, m.col, p->use_skip_sad);
      if (sad < best_sad) {
        sad += (unsigned)sad_cost(p, m.row, m.col);
        if (sad < best_sad) {
          best_sad = sad;
          best_site = j;
        }
      }
    }
    if (best_site < 0) break;
    best.row += nb[best_site][0];
    best.col += nb[best_site][1];
    centre += nb[best_site][0] * STRIDE + nb[best_site][1];
  }
  r->best_row = best.row;
  r->best_col = best.col;
  r->second_row = r->second_col = -32768;
  r->ret = (int)best_sad;
}
