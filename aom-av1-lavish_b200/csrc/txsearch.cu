// txsearch.cu -- the transform-type search evaluated on the device.
//
// search_tx_type (av1/encoder/tx_search.c:2047-2330) and its pruning passes prune_txk_type /
// prune_txk_type_separ (:1100-1330) evaluate, per candidate tx type of a block,
//   av1_xform_quant                       -> eob                          (encodemb.c:295-381)
//   dist_block_tx_domain                  -> block error, ssz             (tx_search.c:1077)
//   av1_cost_coeffs_txb_laplacian         -> rate; its coefficient term is av1_cost_coeffs_txb_estimate
//                                            (txb_rdopt.c:547-582), a reduction over the levels
//   dist_block_px_domain                  -> 16 * SSE(src, recon)         (tx_search.c:969)
// One descriptor per (block, candidate); per descriptor 32 bytes come back instead of 12 bytes per
// sample.  The context-dependent rate terms (txb_skip_cost, tx type cost, eob cost) are table look-ups
// by eob / tx_type the encoder adds on the host.
#include <string.h>

#include <vector>

#include "internal.h"

namespace b200 {

struct RateArgs {
  const B200TxBlock *blocks;
  const int32_t *qcoeff;
  const uint16_t *eob;
  const int16_t *scan_all;   // ctx->d_scan
  int32_t *rate;
  int n_blocks;
  uint32_t scan_off[B200_TX_SIZES_ALL];  // element offset of each tx size's three scan classes
  uint16_t max_eob[B200_TX_SIZES_ALL];
};

// costLUT / const_term / loge_par of av1/encoder/txb_rdopt_utils.h:31-37
__constant__ int c_cost_lut[15] = { -1143, 53, 545, 825, 1031, 1209, 1393, 1577, 1763, 1947, 2132, 2317, 2501, 2686, 2871 };

// One warp per block.  The scan-ordered sum of txb_rdopt.c:566-575 is order free:
//   sum_{c < eob-1} LUT[min(|q_c|, 14)] = S - LUT[min(|q_last|, 14)] + LUT[0] * (eob - nnz)
// with S the sum over the non-zero levels, nnz their count and q_last = qcoeff[scan[eob - 1]].
__global__ void __launch_bounds__(128) rate_est_kernel(const RateArgs a) {
  const int lane = threadIdx.x & 31;
  const int bi = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (bi >= a.n_blocks) return;
  const B200TxBlock d = a.blocks[bi];
  const int eob = a.eob[bi];
  if (eob == 0) {
    if (lane == 0) a.rate[bi] = 0;
    return;
  }
  const int n = a.max_eob[d.tx_size];
  const int32_t *q = a.qcoeff + d.out_offset;
  int S = 0, nnz = 0;
  for (int i = lane; i < n; i += 32) {
    const int v = abs(q[i]);
    if (v) S += c_cost_lut[min(v, 14)], ++nnz;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    S += __shfl_xor_sync(0xffffffffu, S, o);
    nnz += __shfl_xor_sync(0xffffffffu, nnz, o);
  }
  if (lane == 0) {
    const int16_t *scan = a.scan_all + a.scan_off[d.tx_size] + scan_class_of(d.tx_type) * n;
    const int last = abs(q[scan[eob - 1]]);
    const int const_term = 1 << 9, loge_par = ((14427 << 9) + 5000) / 10000;
    a.rate[bi] = (int)((unsigned)(last - 1) << 11) + S - c_cost_lut[min(last, 14)] + c_cost_lut[0] * (eob - nnz) +
                 (const_term + loge_par) * (eob - 1);
  }
}

int rate_estimate_launch(b200_ctx *ctx, const B200RateEstLaunch &l, cudaStream_t s) {
  if (l.n_blocks < 0 || !l.blocks || !l.qcoeff || !l.eob || !l.rate) {
    set_error("b200_rate_estimate_launch: null argument");
    return B200_EINVAL;
  }
  if (l.n_blocks == 0) return B200_OK;
  RateArgs a;
  a.blocks = l.blocks;
  a.qcoeff = l.qcoeff;
  a.eob = l.eob;
  a.scan_all = ctx->d_scan;
  a.rate = l.rate;
  a.n_blocks = l.n_blocks;
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    a.scan_off[ts] = (uint32_t)scan_dev_offset(ts);
    a.max_eob[ts] = (uint16_t)b200_tx_max_eob(ts);
  }
  rate_est_kernel<<<(l.n_blocks + 3) / 4, 128, 0, s>>>(a);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rate_estimate_launch(b200_ctx *ctx, const B200RateEstLaunch *launch, void *stream) {
  if (!ctx || !launch) return B200_EINVAL;
  return rate_estimate_launch(ctx, *launch, stream ? (cudaStream_t)stream : ctx->stream);
}

extern "C" int b200_txsearch_batch_host(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned want,
                                        const B200TxBlock *blocks, const B200TxPred *preds, int n_blocks,
                                        const void *src, size_t src_elems, const void *pred,
                                        size_t pred_elems, const B200QRow *qrows, int n_qrows,
                                        B200TxSearchResult *out, size_t *d2h_bytes) {
  if (d2h_bytes) *d2h_bytes = 0;
  if (!ctx || !blocks || !preds || n_blocks < 0 || !src || !pred || !qrows || n_qrows <= 0 || !out) {
    set_error("b200_txsearch_batch_host: null argument");
    return B200_EINVAL;
  }
  if (in_kind != B200_IN_PIXELS8 && in_kind != B200_IN_PIXELS16) {
    set_error("b200_txsearch_batch_host: in_kind must be B200_IN_PIXELS8 or B200_IN_PIXELS16");
    return B200_EINVAL;
  }
  if (q_kind <= B200_Q_NONE || q_kind > B200_Q_B_HBD) {
    set_error("b200_txsearch_batch_host: a quantiser is required");
    return B200_EINVAL;
  }
  if (in_kind == B200_IN_PIXELS8 && bd != 8) {
    set_error("b200_txsearch_batch_host: uint8 planes imply bd 8");
    return B200_EINVAL;
  }
  if (n_blocks == 0) return B200_OK;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t n = (size_t)n_blocks, px = in_kind == B200_IN_PIXELS8 ? 1 : 2;
  const bool want_px = (want & B200_TXS_DIST_PX) != 0, want_tx = (want & B200_TXS_DIST_TX) != 0;
  const bool want_rate = (want & B200_TXS_RATE_EST) != 0;

  // ---- stable counting sort by tx size; every block gets its own slice of the coefficient scratch
  size_t count[B200_TX_SIZES_ALL + 1] = { 0 };
  for (size_t i = 0; i < n; ++i) {
    const unsigned ts = blocks[i].tx_size, tt = blocks[i].tx_type;
    if (ts >= B200_TX_SIZES_ALL || tt >= 16 || !b200_tx_valid((int)ts, (int)tt)) {
      set_error("b200_txsearch_batch_host: block %zu has invalid tx_size/tx_type %u/%u", i, ts, tt);
      return B200_EINVAL;
    }
    const size_t w = (size_t)b200_tx_wide((int)ts), h = (size_t)b200_tx_high((int)ts);
    if ((size_t)blocks[i].src_offset + (h - 1) * blocks[i].src_stride + w > src_elems ||
        (size_t)preds[i].pred_offset + (h - 1) * preds[i].pred_stride + w > pred_elems ||
        blocks[i].qrow >= (unsigned)n_qrows) {
      set_error("b200_txsearch_batch_host: block %zu reads outside src / pred / qrows", i);
      return B200_EINVAL;
    }
    ++count[ts + 1];
  }
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) count[ts + 1] += count[ts];
  size_t hoff = 0;
  auto hcarve = [&](size_t bytes) {
    const size_t o = hoff;
    hoff += (bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t ho_blocks = hcarve(sizeof(B200TxBlock) * n), ho_preds = hcarve(sizeof(B200TxPred) * n);
  const size_t ho_eob = hcarve(n * 2), ho_tx = hcarve(want_tx ? n * 16 : 0);
  const size_t ho_px = hcarve(want_px ? n * 8 : 0), ho_rate = hcarve(want_rate ? n * 4 : 0);
  int rc = host_scratch_reserve(ctx, hoff);
  if (rc != B200_OK) return rc;
  char *hb = (char *)ctx->h_scratch;
  B200TxBlock *sb = (B200TxBlock *)(hb + ho_blocks);
  B200TxPred *sp = (B200TxPred *)(hb + ho_preds);
  std::vector<uint32_t> order(n);
  {
    size_t cursor[B200_TX_SIZES_ALL];
    for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) cursor[ts] = count[ts];
    for (size_t i = 0; i < n; ++i) {
      const size_t pos = cursor[blocks[i].tx_size]++;
      order[pos] = (uint32_t)i;
      sb[pos] = blocks[i];
      sp[pos] = preds[i];
    }
  }
  size_t out_elems = 0;
  for (size_t i = 0; i < n; ++i) {
    if (out_elems > 0xffffffffull - 1024) {
      set_error("b200_txsearch_batch_host: batch needs more than 2^32 coefficients");
      return B200_EINVAL;
    }
    sb[i].out_offset = (uint32_t)out_elems;
    out_elems += (size_t)b200_tx_max_eob(sb[i].tx_size);
  }

  size_t off = 0;
  auto carve = [&](size_t bytes) {
    const size_t o = off;
    off += (bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t o_blocks = carve(sizeof(B200TxBlock) * n), o_preds = carve(sizeof(B200TxPred) * n);
  const size_t o_src = carve(src_elems * px), o_pred = carve(pred_elems * px);
  const size_t o_q = carve(sizeof(B200QRow) * n_qrows);
  const size_t o_qc = carve(out_elems * 4), o_dq = carve(out_elems * 4);
  const size_t o_eob = carve(n * 2), o_tx = carve(want_tx ? n * 16 : 0);
  const size_t o_px = carve(want_px ? n * 8 : 0), o_rate = carve(want_rate ? n * 4 : 0);
  rc = scratch_reserve(ctx, off);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  cudaStream_t s = ctx->stream;
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_blocks, sb, sizeof(B200TxBlock) * n, cudaMemcpyHostToDevice, s));
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_preds, sp, sizeof(B200TxPred) * n, cudaMemcpyHostToDevice, s));
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_src, src, src_elems * px, cudaMemcpyHostToDevice, s));
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_pred, pred, pred_elems * px, cudaMemcpyHostToDevice, s));
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_q, qrows, sizeof(B200QRow) * n_qrows, cudaMemcpyHostToDevice, s));
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    const size_t first = count[ts], cnt = count[ts + 1] - count[ts];
    if (!cnt) continue;
    B200TxqLaunch l;
    memset(&l, 0, sizeof(l));
    l.tx_size = ts;
    l.in_kind = in_kind;
    l.q_kind = q_kind;
    l.bd = bd;
    l.flags = want_tx ? B200_TXQ_BLOCK_ERROR : 0u;
    l.n_blocks = (int)cnt;
    l.blocks = (const B200TxBlock *)(base + o_blocks) + first;
    l.preds = (const B200TxPred *)(base + o_preds) + first;
    l.src = base + o_src;
    l.pred = base + o_pred;
    l.qrows = (const B200QRow *)(base + o_q);
    l.qcoeff = (int32_t *)(base + o_qc);
    l.dqcoeff = (int32_t *)(base + o_dq);
    l.eob = (uint16_t *)(base + o_eob) + first;
    l.dist = want_tx ? (int64_t *)(base + o_tx) + 2 * first : nullptr;
    rc = txq_launch(ctx, l, s);
    if (rc != B200_OK) return rc;
    if (want_px) {
      B200ItxLaunch il;
      memset(&il, 0, sizeof(il));
      il.tx_size = ts;
      il.highbd = in_kind == B200_IN_PIXELS16;
      il.bd = bd;
      il.n_blocks = (int)cnt;
      il.blocks = l.blocks;
      il.preds = l.preds;
      il.dqcoeff = l.dqcoeff;
      il.eob = l.eob;
      il.src = l.src;
      il.pred = l.pred;
      il.dist = (int64_t *)(base + o_px) + first;
      rc = itx_launch(ctx, il, s);
      if (rc != B200_OK) return rc;
    }
  }
  if (want_rate) {
    B200RateEstLaunch rl;
    rl.n_blocks = n_blocks;
    rl.blocks = (const B200TxBlock *)(base + o_blocks);
    rl.qcoeff = (const int32_t *)(base + o_qc);
    rl.eob = (const uint16_t *)(base + o_eob);
    rl.rate = (int32_t *)(base + o_rate);
    rc = rate_estimate_launch(ctx, rl, s);
    if (rc != B200_OK) return rc;
  }
  size_t d2h = n * 2;
  B200_CUDA_TRY(cudaMemcpyAsync(hb + ho_eob, base + o_eob, n * 2, cudaMemcpyDeviceToHost, s));
  if (want_tx) {
    B200_CUDA_TRY(cudaMemcpyAsync(hb + ho_tx, base + o_tx, n * 16, cudaMemcpyDeviceToHost, s));
    d2h += n * 16;
  }
  if (want_px) {
    B200_CUDA_TRY(cudaMemcpyAsync(hb + ho_px, base + o_px, n * 8, cudaMemcpyDeviceToHost, s));
    d2h += n * 8;
  }
  if (want_rate) {
    B200_CUDA_TRY(cudaMemcpyAsync(hb + ho_rate, base + o_rate, n * 4, cudaMemcpyDeviceToHost, s));
    d2h += n * 4;
  }
  B200_CUDA_TRY(cudaStreamSynchronize(s));
  const uint16_t *h_eob = (const uint16_t *)(hb + ho_eob);
  const int64_t *h_tx = (const int64_t *)(hb + ho_tx), *h_px = (const int64_t *)(hb + ho_px);
  const int32_t *h_rate = (const int32_t *)(hb + ho_rate);
  for (size_t i = 0; i < n; ++i) {
    B200TxSearchResult &r = out[order[i]];
    r.eob = h_eob[i];
    r.reserved = 0;
    r.dist_tx = want_tx ? h_tx[2 * i] : 0;
    r.ssz = want_tx ? h_tx[2 * i + 1] : 0;
    r.dist_px = want_px ? h_px[i] : 0;
    r.rate_est = want_rate ? h_rate[i] : 0;
  }
  if (d2h_bytes) *d2h_bytes = d2h;
  return B200_OK;
}
