// capi.cu -- C ABI glue of libb200av1.so: context, memory helpers, batched tier.
// See include/b200av1.h for the contract.  No CPU fallback: every entry point either
// runs on the GPU or fails loudly.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <mutex>
#include <vector>

#include "internal.h"

namespace b200 {

static thread_local char g_err[512] = "";
static char g_err_global[512] = "";

void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  memcpy(g_err_global, g_err, sizeof(g_err));
}

}  // namespace b200

using namespace b200;

extern "C" const char *b200_last_error(void) { return g_err[0] ? g_err : g_err_global; }

extern "C" int b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

extern "C" int b200_ctx_create(int device, b200_ctx **out) {
  if (!out) return B200_EINVAL;
  *out = nullptr;
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0 || device < 0 || device >= n) {
    cudaGetLastError();
    set_error("b200_ctx_create: no CUDA device %d (count %d); this library has no CPU fallback",
              device, n);
    return B200_ENODEV;
  }
  cudaDeviceProp prop;
  B200_CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10) {
    set_error("b200_ctx_create: device %d is sm_%d%d; libb200av1 is built for sm_100a only",
              device, prop.major, prop.minor);
    return B200_ENODEV;
  }
  B200_CUDA_TRY(cudaSetDevice(device));
  b200_ctx *ctx = (b200_ctx *)calloc(1, sizeof(b200_ctx));
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  B200_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  int rc = txq_init_device(ctx);
  if (rc == B200_OK) rc = blockops_init_device(ctx);
  if (rc == B200_OK) rc = me_init_device(ctx);
  if (rc != B200_OK) {
    free(ctx);
    return rc;
  }
  *out = ctx;
  return B200_OK;
}

extern "C" void b200_ctx_destroy(b200_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  if (ctx->d_iscan) cudaFree(ctx->d_iscan);
  if (ctx->d_pmask) cudaFree(ctx->d_pmask);
  if (ctx->d_scratch) cudaFree(ctx->d_scratch);
  if (ctx->d_me_sites) cudaFree(ctx->d_me_sites);
  if (ctx->h_scratch) cudaFreeHost(ctx->h_scratch);
  cudaStreamDestroy(ctx->stream);
  free(ctx);
}

static b200_ctx *g_default_ctx = nullptr;
static std::once_flag g_default_once;

extern "C" b200_ctx *b200_default_ctx(void) {
  std::call_once(g_default_once, [] {
    const char *env = getenv("B200AV1_DEVICE");
    const int dev = env ? atoi(env) : 0;
    if (b200_ctx_create(dev, &g_default_ctx) != B200_OK) {
      fprintf(stderr, "libb200av1: fatal: %s\n", b200_last_error());
      abort();  // RTCD-shaped callers cannot receive an error
    }
  });
  return g_default_ctx;
}

extern "C" void *b200_ctx_stream(b200_ctx *ctx) { return (void *)ctx->stream; }

extern "C" int b200_sync(b200_ctx *ctx) {
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
  return B200_OK;
}

extern "C" int b200_malloc(b200_ctx *ctx, size_t bytes, void **dptr) {
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  B200_CUDA_TRY(cudaMalloc(dptr, bytes ? bytes : 1));
  return B200_OK;
}
extern "C" int b200_free(b200_ctx *ctx, void *dptr) {
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  B200_CUDA_TRY(cudaFree(dptr));
  return B200_OK;
}
extern "C" int b200_malloc_host(size_t bytes, void **hptr) {
  B200_CUDA_TRY(cudaMallocHost(hptr, bytes ? bytes : 1));
  return B200_OK;
}
extern "C" int b200_free_host(void *hptr) {
  B200_CUDA_TRY(cudaFreeHost(hptr));
  return B200_OK;
}
extern "C" int b200_h2d(b200_ctx *ctx, void *dst, const void *src, size_t bytes) {
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
  return B200_OK;
}
extern "C" int b200_d2h(b200_ctx *ctx, void *dst, const void *src, size_t bytes) {
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  B200_CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  return B200_OK;
}

namespace b200 {
// Grow-only device scratch used by the host-buffer tiers.  Returns a 256-byte aligned
// carve-out; `reset` starts a new carve sequence.
int scratch_reserve(b200_ctx *ctx, size_t bytes) {
  if (bytes <= ctx->scratch_bytes) return B200_OK;
  B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
  if (ctx->d_scratch) B200_CUDA_TRY(cudaFree(ctx->d_scratch));
  ctx->d_scratch = nullptr;
  ctx->scratch_bytes = 0;
  size_t want = bytes + bytes / 2 + (1 << 20);
  B200_CUDA_TRY(cudaMalloc(&ctx->d_scratch, want));
  ctx->scratch_bytes = want;
  return B200_OK;
}
}  // namespace b200

extern "C" int b200_txq_launch(b200_ctx *ctx, const B200TxqLaunch *launch, void *stream) {
  if (!ctx || !launch) return B200_EINVAL;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  return txq_launch(ctx, *launch, stream ? (cudaStream_t)stream : ctx->stream);
}

static size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }

extern "C" int b200_txq_batch_host(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                                   const B200TxBlock *blocks, const B200TxPred *preds,
                                   int n_blocks, const void *src, size_t src_elems,
                                   const void *pred, size_t pred_elems, const B200QRow *qrows,
                                   int n_qrows, int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff,
                                   size_t out_elems, uint16_t *eob, int64_t *dist) {
  if (!ctx || !blocks || n_blocks < 0 || !src) return B200_EINVAL;
  if (n_blocks == 0) return B200_OK;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t px = in_kind == B200_IN_PIXELS8 ? 1 : 2;
  const bool has_pred = in_kind != B200_IN_RESIDUAL;
  const bool has_q = q_kind != B200_Q_NONE;
  const bool want_coeff = (flags & B200_TXQ_STORE_COEFF) != 0;
  const bool want_err = (flags & B200_TXQ_BLOCK_ERROR) != 0 && dist;
  if (has_pred && (!preds || !pred)) return B200_EINVAL;

  // group the blocks by tx_size (stable), remembering each block's original index
  std::vector<int> order(n_blocks);
  for (int i = 0; i < n_blocks; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int x, int y) {
    if (blocks[x].tx_size != blocks[y].tx_size) return blocks[x].tx_size < blocks[y].tx_size;
    return blocks[x].tx_type < blocks[y].tx_type;  // keep warps on one code path
  });
  std::vector<B200TxBlock> sorted(n_blocks);
  std::vector<B200TxPred> sorted_p(has_pred ? n_blocks : 0);
  for (int i = 0; i < n_blocks; ++i) {
    sorted[i] = blocks[order[i]];
    if (sorted[i].tx_size >= B200_TX_SIZES_ALL || !b200_tx_valid(sorted[i].tx_size, sorted[i].tx_type)) {
      set_error("b200_txq_batch_host: block %d has invalid tx_size/tx_type %d/%d", order[i],
                sorted[i].tx_size, sorted[i].tx_type);
      return B200_EINVAL;
    }
    if (has_pred) sorted_p[i] = preds[order[i]];
  }

  size_t off = 0;
  auto carve = [&](size_t bytes) {
    size_t o = off;
    off += align_up(bytes);
    return o;
  };
  const size_t o_blocks = carve(sizeof(B200TxBlock) * n_blocks);
  const size_t o_preds = carve(has_pred ? sizeof(B200TxPred) * n_blocks : 0);
  const size_t o_src = carve(src_elems * px);
  const size_t o_pred = carve(has_pred ? pred_elems * px : 0);
  const size_t o_q = carve(has_q ? sizeof(B200QRow) * n_qrows : 0);
  const size_t o_coeff = carve(want_coeff ? out_elems * 4 : 0);
  const size_t o_qc = carve(has_q ? out_elems * 4 : 0);
  const size_t o_dq = carve(has_q ? out_elems * 4 : 0);
  const size_t o_eob = carve(has_q ? n_blocks * 2 : 0);
  const size_t o_dist = carve(want_err ? n_blocks * 16 : 0);
  int rc = scratch_reserve(ctx, off);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  cudaStream_t s = ctx->stream;
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_blocks, sorted.data(), sizeof(B200TxBlock) * n_blocks,
                                cudaMemcpyHostToDevice, s));
  if (has_pred) {
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_preds, sorted_p.data(), sizeof(B200TxPred) * n_blocks,
                                  cudaMemcpyHostToDevice, s));
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_pred, pred, pred_elems * px, cudaMemcpyHostToDevice, s));
  }
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_src, src, src_elems * px, cudaMemcpyHostToDevice, s));
  if (has_q)
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_q, qrows, sizeof(B200QRow) * n_qrows,
                                  cudaMemcpyHostToDevice, s));
  // regions of the output buffers that no block covers read back as zero
  if (want_coeff) B200_CUDA_TRY(cudaMemsetAsync(base + o_coeff, 0, out_elems * 4, s));
  if (has_q) {
    B200_CUDA_TRY(cudaMemsetAsync(base + o_qc, 0, out_elems * 4, s));
    B200_CUDA_TRY(cudaMemsetAsync(base + o_dq, 0, out_elems * 4, s));
  }
  for (int i = 0; i < n_blocks;) {
    int j = i;
    while (j < n_blocks && sorted[j].tx_size == sorted[i].tx_size) ++j;
    B200TxqLaunch l;
    memset(&l, 0, sizeof(l));
    l.tx_size = sorted[i].tx_size;
    l.in_kind = in_kind;
    l.q_kind = q_kind;
    l.bd = bd;
    l.flags = flags & ~(want_err ? 0u : B200_TXQ_BLOCK_ERROR);
    l.n_blocks = j - i;
    l.blocks = (const B200TxBlock *)(base + o_blocks) + i;
    l.preds = has_pred ? (const B200TxPred *)(base + o_preds) + i : nullptr;
    l.src = base + o_src;
    l.pred = has_pred ? base + o_pred : nullptr;
    l.qrows = has_q ? (const B200QRow *)(base + o_q) : nullptr;
    l.coeff = want_coeff ? (int32_t *)(base + o_coeff) : nullptr;
    l.qcoeff = has_q ? (int32_t *)(base + o_qc) : nullptr;
    l.dqcoeff = has_q ? (int32_t *)(base + o_dq) : nullptr;
    l.eob = has_q ? (uint16_t *)(base + o_eob) + i : nullptr;
    l.dist = want_err ? (int64_t *)(base + o_dist) + 2 * (size_t)i : nullptr;
    rc = txq_launch(ctx, l, s);
    if (rc != B200_OK) return rc;
    i = j;
  }
  std::vector<uint16_t> eob_sorted(has_q ? n_blocks : 0);
  std::vector<int64_t> dist_sorted(want_err ? 2 * (size_t)n_blocks : 0);
  if (want_coeff)
    B200_CUDA_TRY(cudaMemcpyAsync(coeff, base + o_coeff, out_elems * 4, cudaMemcpyDeviceToHost, s));
  if (has_q) {
    B200_CUDA_TRY(cudaMemcpyAsync(qcoeff, base + o_qc, out_elems * 4, cudaMemcpyDeviceToHost, s));
    B200_CUDA_TRY(cudaMemcpyAsync(dqcoeff, base + o_dq, out_elems * 4, cudaMemcpyDeviceToHost, s));
    B200_CUDA_TRY(cudaMemcpyAsync(eob_sorted.data(), base + o_eob, n_blocks * 2,
                                  cudaMemcpyDeviceToHost, s));
  }
  if (want_err)
    B200_CUDA_TRY(cudaMemcpyAsync(dist_sorted.data(), base + o_dist, n_blocks * 16,
                                  cudaMemcpyDeviceToHost, s));
  B200_CUDA_TRY(cudaStreamSynchronize(s));
  for (int i = 0; i < n_blocks; ++i) {
    if (has_q && eob) eob[order[i]] = eob_sorted[i];
    if (want_err) {
      dist[2 * (size_t)order[i]] = dist_sorted[2 * (size_t)i];
      dist[2 * (size_t)order[i] + 1] = dist_sorted[2 * (size_t)i + 1];
    }
  }
  return B200_OK;
}
