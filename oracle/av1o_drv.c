/*
 * av1o_drv.c -- ORACLE / CPU-BASELINE DRIVER (test infrastructure only).
 *
 * Runs a batch of same-size transform blocks through "forward transform, then
 * quantize_fp" on the host CPU with a pthread pool, calling through FUNCTION POINTERS
 * so that the same driver times either the reference's own C functions
 * (oracle/_ref/libaomref.so: av1_fwd_txfm2d_WxH_c + av1_quantize_fp*_c /
 * av1_highbd_quantize_fp_c) or this oracle's restatement.  It is what bench.py's
 * `cpu_baseline` leg and `--impl reference` arm execute; the product never links it.
 */
#include <pthread.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "av1o.h"

typedef void (*fwd_txfm_fn)(const int16_t *input, int32_t *output, int stride,
                            uint8_t tx_type, int bd);
/* av1_quantize_fp_c family: 12 arguments; av1_highbd_quantize_fp_c: + log_scale */
typedef void (*quant_fp_fn)(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                            const int16_t *round, const int16_t *quant,
                            const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                            const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                            const int16_t *iscan);
typedef void (*quant_fp_hbd_fn)(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                                const int16_t *round, const int16_t *quant,
                                const int16_t *quant_shift, int32_t *qcoeff,
                                int32_t *dqcoeff, const int16_t *dequant, uint16_t *eob,
                                const int16_t *scan, const int16_t *iscan, int log_scale);

typedef struct {
  uint32_t src_offset, out_offset;
  uint16_t src_stride, qrow;
  uint8_t tx_size, tx_type;
  uint16_t reserved;
} DrvBlock; /* same layout as B200TxBlock */

typedef struct {
  fwd_txfm_fn fwd;
  void *quant;
  int highbd, log_scale, bd, n_coeffs, full_coeffs;
  const DrvBlock *blocks;
  int n_blocks;
  const int16_t *src;
  const Av1oQRow *rows;
  const int16_t *scan[3], *iscan[3];
  int32_t *coeff, *qcoeff, *dqcoeff;
  uint16_t *eob;
  int tid, n_threads;
} DrvJob;

static void *drv_worker(void *arg) {
  DrvJob *j = (DrvJob *)arg;
  /* the reference transform uses its output as scratch for the full W*H block */
  /* 64-byte aligned like the encoder's own coefficient buffers (the SIMD flavours use aligned stores) */
  void *tmp_raw = malloc(sizeof(int32_t) * (size_t)(j->full_coeffs + 64) + 64);
  int32_t *tmp = (int32_t *)(((uintptr_t)tmp_raw + 63) & ~(uintptr_t)63);
  const int per = (j->n_blocks + j->n_threads - 1) / j->n_threads;
  const int lo = j->tid * per, hi = lo + per > j->n_blocks ? j->n_blocks : lo + per;
  for (int i = lo; i < hi; ++i) {
    const DrvBlock *b = &j->blocks[i];
    const int cls = b->tx_type < 10 ? 0 : ((b->tx_type & 1) ? 2 : 1);
    const Av1oQRow *r = &j->rows[b->qrow];
    int32_t *c = j->coeff + b->out_offset;
    j->fwd(j->src + b->src_offset, tmp, b->src_stride, b->tx_type, j->bd);
    memcpy(c, tmp, sizeof(int32_t) * (size_t)j->n_coeffs);
    if (j->highbd)
      ((quant_fp_hbd_fn)j->quant)(c, j->n_coeffs, r->zbin, r->round_fp, r->quant_fp,
                                  r->quant_shift, j->qcoeff + b->out_offset,
                                  j->dqcoeff + b->out_offset, r->dequant, &j->eob[i],
                                  j->scan[cls], j->iscan[cls], j->log_scale);
    else
      ((quant_fp_fn)j->quant)(c, j->n_coeffs, r->zbin, r->round_fp, r->quant_fp,
                              r->quant_shift, j->qcoeff + b->out_offset,
                              j->dqcoeff + b->out_offset, r->dequant, &j->eob[i],
                              j->scan[cls], j->iscan[cls]);
  }
  free(tmp_raw);
  return NULL;
}

/* adapters so that the oracle's own functions fit the reference prototypes */
static int g_oracle_tx_size;
static void oracle_fwd(const int16_t *input, int32_t *output, int stride, uint8_t tx_type,
                       int bd) {
  av1o_fwd_txfm2d(input, output, stride, g_oracle_tx_size, tx_type, bd);
}
static void oracle_qfp_hbd(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                           const int16_t *round, const int16_t *quant,
                           const int16_t *quant_shift, int32_t *qcoeff, int32_t *dqcoeff,
                           const int16_t *dequant, uint16_t *eob, const int16_t *scan,
                           const int16_t *iscan, int log_scale) {
  av1o_highbd_quantize_fp(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff, dequant,
                          eob, scan, iscan, NULL, NULL, log_scale);
}
static int g_oracle_ls;
static void oracle_qfp(const int32_t *coeff, intptr_t n, const int16_t *zbin,
                       const int16_t *round, const int16_t *quant, const int16_t *quant_shift,
                       int32_t *qcoeff, int32_t *dqcoeff, const int16_t *dequant,
                       uint16_t *eob, const int16_t *scan, const int16_t *iscan) {
  av1o_quantize_fp(coeff, n, zbin, round, quant, quant_shift, qcoeff, dqcoeff, dequant, eob,
                   scan, iscan, NULL, NULL, g_oracle_ls);
}

/* fwd / quant == NULL selects the oracle's own restatement. */
int av1o_drv_txq_fp(void *fwd, void *quant, int highbd, int bd, int tx_size,
                    const DrvBlock *blocks, int n_blocks, const int16_t *src,
                    const Av1oQRow *rows, int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff,
                    uint16_t *eob, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  const int pels = av1o_tx_wide(tx_size) * av1o_tx_high(tx_size);
  const int ls = (pels > 256) + (pels > 1024);
  g_oracle_tx_size = tx_size;
  g_oracle_ls = ls;
  DrvJob proto;
  memset(&proto, 0, sizeof(proto));
  proto.fwd = fwd ? (fwd_txfm_fn)fwd : oracle_fwd;
  proto.quant = quant ? quant : (highbd ? (void *)oracle_qfp_hbd : (void *)oracle_qfp);
  proto.highbd = highbd;
  proto.log_scale = ls;
  proto.bd = bd;
  proto.n_coeffs = av1o_max_eob(tx_size);
  proto.full_coeffs = pels;
  proto.blocks = blocks;
  proto.n_blocks = n_blocks;
  proto.src = src;
  proto.rows = rows;
  for (int cls = 0; cls < 3; ++cls) {
    const int tt = cls == 0 ? 0 : (cls == 1 ? 10 : 11);
    proto.scan[cls] = av1o_scan(tx_size, tt);
    proto.iscan[cls] = av1o_iscan(tx_size, tt);
  }
  proto.coeff = coeff;
  proto.qcoeff = qcoeff;
  proto.dqcoeff = dqcoeff;
  proto.eob = eob;
  proto.n_threads = n_threads;
  DrvJob *jobs = (DrvJob *)malloc(sizeof(DrvJob) * (size_t)n_threads);
  pthread_t *th = (pthread_t *)malloc(sizeof(pthread_t) * (size_t)n_threads);
  for (int t = 0; t < n_threads; ++t) {
    jobs[t] = proto;
    jobs[t].tid = t;
    if (t > 0) pthread_create(&th[t], NULL, drv_worker, &jobs[t]);
  }
  drv_worker(&jobs[0]);
  for (int t = 1; t < n_threads; ++t) pthread_join(th[t], NULL);
  free(jobs);
  free(th);
  return 0;
}

/* ---- motion search: full-pel search, then the sub-pel tree started at its result, for a
 * list of searches on a pthread pool.  `full` / `sub` are either the reference's own
 * av1_full_pixel_search / av1_find_best_sub_pixel_tree behind oracle/ref_shim.c
 * (refshim_full_pixel_search / refshim_find_best_sub_pixel_tree) or this oracle's
 * restatement (av1o_full_pixel_search / av1o_find_best_sub_pixel_tree): same signature. */
#include "av1o_me.h"
typedef void (*me_fn)(const Av1oMeParams *p, Av1oMeResult *r);
typedef struct {
  me_fn full, sub;
  Av1oMeParams *params;
  Av1oMeResult *rfull, *rsub;
  int n, tid, n_threads;
  volatile int *next;
} MeJob;

static void *me_worker(void *arg) {
  MeJob *j = (MeJob *)arg;
  for (;;) {
    const int i = __sync_fetch_and_add(j->next, 1); /* searches differ a lot in cost: dynamic split */
    if (i >= j->n) break;
    Av1oMeParams p = j->params[i];
    j->full(&p, &j->rfull[i]);
    if (j->sub) {
      p.sp_start_row = j->rfull[i].best_row * 8;
      p.sp_start_col = j->rfull[i].best_col * 8;
      j->sub(&p, &j->rsub[i]);
    }
  }
  return NULL;
}

void av1o_drv_me(void *full, void *sub, Av1oMeParams *params, int n, Av1oMeResult *rfull,
                 Av1oMeResult *rsub, int n_threads) {
  if (n_threads < 1) n_threads = 1;
  if (n_threads > 256) n_threads = 256;
  pthread_t th[256];
  MeJob jobs[256];
  volatile int next = 0;
  for (int t = 0; t < n_threads; ++t) {
    jobs[t] = (MeJob){ (me_fn)full, (me_fn)sub, params, rfull, rsub, n, t, n_threads, &next };
    pthread_create(&th[t], NULL, me_worker, &jobs[t]);
  }
  for (int t = 0; t < n_threads; ++t) pthread_join(th[t], NULL);
}
