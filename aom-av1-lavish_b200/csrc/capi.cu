// capi.cu -- C ABI glue of libb200av1.so: context, memory helpers, batched tier.
// See include/b200av1.h for the contract.  No CPU fallback: every entry point either
// runs on the GPU or fails loudly.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>

#include <stdio.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
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

static int ctx_init(b200_ctx *ctx) {
  B200_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  B200_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  B200_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking));
  for (int i = 0; i < B200_TX_SIZES_ALL; ++i)
    B200_CUDA_TRY(cudaEventCreateWithFlags(&ctx->group_done[i], cudaEventDisableTiming));
  B200_CUDA_TRY(cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
  int rc = txq_init_device(ctx);
  if (rc == B200_OK) rc = blockops_init_device(ctx);
  if (rc == B200_OK) rc = me_init_device(ctx);
  if (rc == B200_OK) rc = compact_init_device(ctx);
  if (rc == B200_OK) rc = itx_init_device(ctx);
  return rc;
}

extern "C" void b200_ctx_destroy(b200_ctx *ctx);

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
  if (!ctx) {
    set_error("b200_ctx_create: out of host memory");
    return B200_EINVAL;
  }
  ctx->device = device;
  ctx->sm_count = prop.multiProcessorCount;
  int rc = ctx_init(ctx);
  if (rc != B200_OK) {
    b200_ctx_destroy(ctx);  // releases whatever was created so far
    return rc;
  }
  *out = ctx;
  return B200_OK;
}

extern "C" void b200_ctx_destroy(b200_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream) cudaStreamSynchronize(ctx->stream);
  if (ctx->copy_stream) cudaStreamSynchronize(ctx->copy_stream);
  if (ctx->aux_stream) cudaStreamSynchronize(ctx->aux_stream);
  if (ctx->d_iscan) cudaFree(ctx->d_iscan);
  if (ctx->d_scan) cudaFree(ctx->d_scan);
  if (ctx->d_pmask) cudaFree(ctx->d_pmask);
  if (ctx->d_scratch) cudaFree(ctx->d_scratch);
  if (ctx->d_me_sites) cudaFree(ctx->d_me_sites);
  if (ctx->d_partials) cudaFree(ctx->d_partials);
  if (ctx->h_scratch) cudaFreeHost(ctx->h_scratch);
  for (int i = 0; i < B200_TX_SIZES_ALL; ++i)
    if (ctx->group_done[i]) cudaEventDestroy(ctx->group_done[i]);
  if (ctx->copy_done) cudaEventDestroy(ctx->copy_done);
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  cudaGetLastError();
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

namespace b200 {
// Grow-only pinned host staging.
int host_scratch_reserve(b200_ctx *ctx, size_t bytes) {
  if (bytes <= ctx->h_scratch_bytes) return B200_OK;
  B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
  B200_CUDA_TRY(cudaStreamSynchronize(ctx->copy_stream));
  if (ctx->h_scratch) B200_CUDA_TRY(cudaFreeHost(ctx->h_scratch));
  ctx->h_scratch = nullptr;
  ctx->h_scratch_bytes = 0;
  const size_t want = bytes + bytes / 2 + (1 << 16);
  B200_CUDA_TRY(cudaMallocHost(&ctx->h_scratch, want));
  ctx->h_scratch_bytes = want;
  return B200_OK;
}
}  // namespace b200

extern "C" int b200_txq_launch(b200_ctx *ctx, const B200TxqLaunch *launch, void *stream) {
  if (!ctx || !launch) return B200_EINVAL;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  return txq_launch(ctx, *launch, stream ? (cudaStream_t)stream : ctx->stream);
}

static size_t align_up(size_t v) { return (v + 255) & ~(size_t)255; }
// B200_HOST_TIMING=1: per-phase wall clock of the host-buffer tier on stderr (measurement aid)
struct HostPhases {
  bool on;
  std::chrono::steady_clock::time_point t0;
  char line[512];
  int len;
  HostPhases() : on(getenv("B200_HOST_TIMING") != nullptr), t0(std::chrono::steady_clock::now()), len(0) { line[0] = 0; }
  void mark(const char *name) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    len += snprintf(line + len, sizeof(line) - len, " %s %.3f", name,
                    std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
  ~HostPhases() {
    if (on) fprintf(stderr, "[b200 host ms]%s\n", line);
  }
};

// Host-buffer wrapper of the fused kernel -- the reference-facing call of the frame tier
// (what `e2e` in bench.py times).  Pipeline per call:
//   host   counting sort of the descriptors by (tx_size, tx_type) -- skipped when the queue
//          already arrives in that order -- with bounds checks of every descriptor, and a pass
//          that finds, per tx-size group, the range of the output arrays its blocks write
//   stream   H2D of planes / tables / descriptors, one kernel launch per tx-size group
//   copy_stream   dense mode: as each group's kernel finishes, D2H of exactly that group's
//          output range, so the PCIe read-back (12 of the 14 bytes per sample) overlaps the
//          remaining uploads and kernels instead of following them;
//          compact mode: the dense qcoeff / dqcoeff never leave the device -- one compaction
//          pass (compact.cu) packs the levels, the host reads eob + offsets + total, then
//          exactly `total` levels
// When the groups' output ranges are not dense and disjoint (a caller that interleaves
// sizes inside the arrays, or blocks that overlap) the dense mode falls back to zero-filled
// device arrays and one whole-array read-back at the end.
namespace {
struct CompactOut {
  void *levels;
  size_t levels_cap;
  uint32_t *level_offset;
  size_t *d2h_bytes;
};
}  // namespace

// A few persistent host threads for the per-descriptor passes of the host-buffer tier: a frame's queue is
// ~10^6 descriptors, and spawning std::threads per call cost more (~0.5 ms) than the pass itself.
namespace {
class HostPool {
 public:
  static HostPool &get() {
    static HostPool p;
    return p;
  }
  int width() const { return (int)workers_.size() + 1; }
  // f(0 .. n_parts) on the pool and the calling thread; returns when all parts are done
  void run(int n_parts, const std::function<void(int)> &f) {
    if (n_parts <= 1 || workers_.empty()) {
      for (int i = 0; i < n_parts; ++i) f(i);
      return;
    }
    std::lock_guard<std::mutex> one_job(job_mu_);
    {
      std::lock_guard<std::mutex> g(mu_);
      fn_ = &f;
      n_.store(n_parts);
      pending_.store(n_parts);
      next_.store(0);
      ++gen_;
    }
    cv_.notify_all();
    work();
    std::unique_lock<std::mutex> g(mu_);
    done_cv_.wait(g, [&] { return pending_.load() == 0; });
  }

 private:
  HostPool() {
    // B200_HOST_THREADS: width of the pool (a process per GPU on one host should divide the cores between them)
    const unsigned hw = std::thread::hardware_concurrency();
    int n = (int)std::min<unsigned>(16, hw ? hw : 1);
    if (const char *e = getenv("B200_HOST_THREADS")) n = std::max(1, std::min(64, atoi(e)));
    for (int i = 0; i < n - 1; ++i) workers_.emplace_back([this] { loop(); });
  }
  ~HostPool() {
    {
      std::lock_guard<std::mutex> g(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto &t : workers_) t.join();
  }
  void work() {
    for (;;) {
      const int i = next_.fetch_add(1);
      if (i >= n_.load()) return;
      (*fn_)(i);
      if (pending_.fetch_sub(1) == 1) {
        std::lock_guard<std::mutex> g(mu_);
        done_cv_.notify_all();
      }
    }
  }
  void loop() {
    unsigned long long seen = 0;
    for (;;) {
      {
        std::unique_lock<std::mutex> g(mu_);
        cv_.wait(g, [&] { return stop_ || gen_ != seen; });
        if (stop_) return;
        seen = gen_;
      }
      work();
    }
  }
  std::vector<std::thread> workers_;
  std::mutex mu_, job_mu_;
  std::condition_variable cv_, done_cv_;
  const std::function<void(int)> *fn_ = nullptr;
  std::atomic<int> n_{ 0 }, next_{ 0 }, pending_{ 0 };
  unsigned long long gen_ = 0;
  bool stop_ = false;
};
}  // namespace

static int txq_batch_host_impl(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                               const B200TxBlock *blocks, const B200TxPred *preds, int n_blocks,
                               const void *src, size_t src_elems, const void *pred, size_t pred_elems,
                               const B200QRow *qrows, int n_qrows, int32_t *coeff, int32_t *qcoeff,
                               int32_t *dqcoeff, size_t out_elems, uint16_t *eob, int64_t *dist,
                               const CompactOut *cmp) {
  if (!ctx || !blocks || n_blocks < 0 || !src) return B200_EINVAL;
  if (cmp && cmp->d2h_bytes) *cmp->d2h_bytes = 0;
  if (n_blocks == 0) {
    if (cmp) cmp->level_offset[0] = 0;
    return B200_OK;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t px = in_kind == B200_IN_PIXELS8 ? 1 : 2;
  const bool has_pred = in_kind != B200_IN_RESIDUAL;
  const bool has_q = q_kind != B200_Q_NONE;
  const bool want_coeff = (flags & B200_TXQ_STORE_COEFF) != 0 && coeff;
  const bool want_err = (flags & B200_TXQ_BLOCK_ERROR) != 0 && dist;
  if (has_pred && (!preds || !pred)) return B200_EINVAL;
  if (has_q && (!qrows || n_qrows <= 0)) return B200_EINVAL;
  if (cmp && (!has_q || !cmp->levels || !cmp->level_offset || !eob)) {
    set_error("b200_txq_batch_host_compact: needs a quantiser, levels, level_offset and eob");
    return B200_EINVAL;
  }
  if (!cmp && ((has_q && (!qcoeff || !dqcoeff)) || ((flags & B200_TXQ_STORE_COEFF) && !coeff))) {
    set_error("b200_txq_batch_host: missing output buffer");
    return B200_EINVAL;
  }
  const size_t n = (size_t)n_blocks;
  HostPhases ph;

  // ---- device carve-out
  const bool wide = q_kind == B200_Q_FP_HBD || q_kind == B200_Q_B_HBD;
  const size_t lv_bytes = wide ? 4 : 2;
  size_t off = 0;
  auto carve = [&](size_t bytes) {
    size_t o = off;
    off += align_up(bytes);
    return o;
  };
  const size_t o_blocks = carve(sizeof(B200TxBlock) * n);
  const size_t o_preds = carve(has_pred ? sizeof(B200TxPred) * n : 0);
  const size_t o_src = carve(src_elems * px);
  const size_t o_pred = carve(has_pred ? pred_elems * px : 0);
  const size_t o_q = carve(has_q ? sizeof(B200QRow) * n_qrows : 0);
  const size_t o_coeff = carve(want_coeff ? out_elems * 4 : 0);
  const size_t o_qc = carve(has_q ? out_elems * 4 : 0);
  const size_t o_dq = carve(has_q ? out_elems * 4 : 0);
  const size_t o_eob = carve(has_q ? n * 2 : 0);
  const size_t o_dist = carve(want_err ? n * 16 : 0);
  const size_t o_loff = carve(cmp ? (n + 1) * 4 : 0);
  const size_t o_lv = carve(cmp ? out_elems * lv_bytes : 0);
  int rc = scratch_reserve(ctx, off);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  cudaStream_t s = ctx->stream, cs = ctx->copy_stream;

  // the inputs go up while the host validates and buckets the queue (nothing below depends on the
  // descriptors' contents); the descriptors too, as they are: a queue that turns out not to be grouped by
  // tx_size is overwritten by its sorted copy later.  Every early return from here on drains the stream.
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_src, src, src_elems * px, cudaMemcpyHostToDevice, s));
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_blocks, blocks, sizeof(B200TxBlock) * n, cudaMemcpyHostToDevice, s));
  if (has_pred)
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_preds, preds, sizeof(B200TxPred) * n, cudaMemcpyHostToDevice, s));
  if (has_pred)
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_pred, pred, pred_elems * px, cudaMemcpyHostToDevice, s));
  if (has_q)
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_q, qrows, sizeof(B200QRow) * n_qrows,
                                  cudaMemcpyHostToDevice, s));
  auto bail = [&](int code) {
    cudaStreamSynchronize(s);
    return code;
  };

  // ---- one pass over the queue: validate, bucket by tx_size (one launch per size; the caller's order
  // inside a size is kept -- it is the caller that knows how to keep blocks of one tx type and of one
  // neighbourhood together, e.g. type groups inside ~1 MB bands, see bench.py make_descriptors), and find
  // the range of the output arrays each tx size writes
  constexpr int kKeys = B200_TX_SIZES_ALL * 16;
  static bool s_valid[kKeys];
  static size_t s_max_eob[B200_TX_SIZES_ALL];
  static int s_w[B200_TX_SIZES_ALL], s_h[B200_TX_SIZES_ALL];
  static std::once_flag s_once;
  std::call_once(s_once, [] {
    for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
      s_max_eob[ts] = (size_t)b200_tx_max_eob(ts);
      s_w[ts] = b200_tx_wide(ts);
      s_h[ts] = b200_tx_high(ts);
      for (int tt = 0; tt < 16; ++tt) s_valid[ts * 16 + tt] = b200_tx_valid(ts, tt) != 0;
    }
  });
  std::vector<uint32_t> bucket(kKeys + 1, 0);
  size_t g_lo[B200_TX_SIZES_ALL], g_hi[B200_TX_SIZES_ALL];
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) g_lo[ts] = (size_t)-1, g_hi[ts] = 0;
  bool in_order = true;
  {
    // a frame's queue is ~10^6 descriptors: the validation pass runs on a few host threads
    struct Part {
      uint32_t count[B200_TX_SIZES_ALL];
      size_t lo[B200_TX_SIZES_ALL], hi[B200_TX_SIZES_ALL];
      long long bad_type, bad_range;  // first offending block, or -1
      bool in_order;
    };
    const int n_parts = (int)std::max<size_t>(1, std::min<size_t>(n >> 14, (size_t)HostPool::get().width()));
    std::vector<Part> parts(n_parts);
    auto scan = [&](int pi) {
      Part &pt = parts[pi];
      for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) pt.count[ts] = 0, pt.lo[ts] = (size_t)-1, pt.hi[ts] = 0;
      pt.bad_type = pt.bad_range = -1;
      pt.in_order = true;
      const size_t i0 = n * pi / n_parts, i1 = n * (pi + 1) / n_parts;
      unsigned prev = i0 ? blocks[i0 - 1].tx_size : 0;
      for (size_t i = i0; i < i1; ++i) {
        const unsigned ts = blocks[i].tx_size;
        if (ts >= B200_TX_SIZES_ALL || blocks[i].tx_type >= 16 || !s_valid[ts * 16u + blocks[i].tx_type]) {
          pt.bad_type = (long long)i;
          return;
        }
        // the last sample a block reads must lie inside the input buffers
        const size_t span = (size_t)(s_h[ts] - 1) * blocks[i].src_stride + s_w[ts];
        if (((size_t)blocks[i].src_offset + span > src_elems ||
             (has_pred && (size_t)preds[i].pred_offset + (size_t)(s_h[ts] - 1) * preds[i].pred_stride + s_w[ts] >
                              pred_elems) ||
             (has_q && blocks[i].qrow >= (unsigned)n_qrows)) && pt.bad_range < 0)
          pt.bad_range = (long long)i;
        pt.in_order = pt.in_order && ts >= prev;
        prev = ts;
        ++pt.count[ts];
        const size_t o = blocks[i].out_offset;
        pt.lo[ts] = o < pt.lo[ts] ? o : pt.lo[ts];
        pt.hi[ts] = o > pt.hi[ts] ? o : pt.hi[ts];
      }
    };
    HostPool::get().run(n_parts, scan);
    for (const Part &pt : parts) {
      if (pt.bad_type >= 0) {
        const size_t i = (size_t)pt.bad_type;
        set_error("b200_txq_batch_host: block %zu has invalid tx_size/tx_type %d/%d", i, (int)blocks[i].tx_size,
                  (int)blocks[i].tx_type);
        return bail(B200_EINVAL);
      }
      if (pt.bad_range >= 0) {
        set_error("b200_txq_batch_host: block %zu reads outside src / pred / qrows", (size_t)pt.bad_range);
        return bail(B200_EINVAL);
      }
      in_order = in_order && pt.in_order;
      for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
        bucket[ts * 16 + 1] += pt.count[ts];
        g_lo[ts] = pt.lo[ts] < g_lo[ts] ? pt.lo[ts] : g_lo[ts];
        g_hi[ts] = pt.hi[ts] > g_hi[ts] ? pt.hi[ts] : g_hi[ts];
      }
    }
  }
  for (int k = 0; k < kKeys; ++k) bucket[k + 1] += bucket[k];
  ph.mark("validate");

  // pinned staging: [sorted blocks][sorted preds][eob][dist][level offsets]
  size_t hoff = 0;
  auto hcarve = [&](size_t bytes) {
    const size_t o = hoff;
    hoff += (bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t ho_blocks = hcarve(in_order ? 0 : sizeof(B200TxBlock) * n);
  const size_t ho_preds = hcarve(!in_order && has_pred ? sizeof(B200TxPred) * n : 0);
  const size_t ho_eob = hcarve(has_q ? n * 2 : 0);
  const size_t ho_dist = hcarve(want_err ? n * 16 : 0);
  const size_t ho_loff = hcarve(cmp ? (n + 1) * 4 : 0);
  rc = host_scratch_reserve(ctx, hoff);
  if (rc != B200_OK) return bail(rc);
  char *hbase = (char *)ctx->h_scratch;
  // a queue already grouped by tx_size needs no scatter: eob, distortion and the level offsets are read
  // back straight into the caller's arrays
  char *eob_dst = in_order && eob ? (char *)eob : hbase + ho_eob;
  char *dist_dst = in_order && dist ? (char *)dist : hbase + ho_dist;
  char *loff_dst = in_order && cmp ? (char *)cmp->level_offset : hbase + ho_loff;
  std::vector<uint32_t> order;  // sorted position -> caller's index
  const B200TxBlock *sblocks = blocks;
  const B200TxPred *spreds = preds;
  if (!in_order) {
    order.resize(n);
    std::vector<uint32_t> cursor(bucket.begin(), bucket.end() - 1);
    B200TxBlock *sb = (B200TxBlock *)(hbase + ho_blocks);
    B200TxPred *sp = (B200TxPred *)(hbase + ho_preds);
    for (size_t i = 0; i < n; ++i) {
      const uint32_t pos = cursor[blocks[i].tx_size * 16]++;
      order[pos] = (uint32_t)i;
      sb[pos] = blocks[i];
      if (has_pred) sp[pos] = preds[i];
    }
    sblocks = sb;
    spreds = sp;
  }

  // ---- tx-size groups and the output range each one writes
  struct Group {
    int ts;
    size_t first, count, lo, hi, elems;
  };
  std::vector<Group> groups;
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    const size_t first = bucket[ts * 16], last = bucket[ts * 16 + 16];
    if (last == first) continue;
    const size_t me = s_max_eob[ts];
    Group g = { ts, first, last - first, g_lo[ts], g_hi[ts] + me, (last - first) * me };
    if (g.hi > out_elems) {
      set_error("b200_txq_batch_host: a %s block writes past out_elems", "queued");
      return bail(B200_EINVAL);
    }
    groups.push_back(g);
  }
  // dense = every group's blocks tile its [lo, hi) range exactly once and the groups' ranges
  // are disjoint; checked per block (slot bitmap), not only by the range length
  bool dense = !cmp;
  std::vector<const Group *> by_lo;
  if (!cmp) {
    std::vector<uint8_t> seen;
    for (const Group &g : groups) {
      dense = dense && g.hi - g.lo == g.elems;
      by_lo.push_back(&g);
      if (!dense) break;
      const size_t me = s_max_eob[g.ts];
      seen.assign(g.count, 0);
      for (size_t i = g.first; i < g.first + g.count && dense; ++i) {
        const size_t rel = sblocks[i].out_offset - g.lo;
        if (rel % me != 0 || seen[rel / me]) dense = false;
        else seen[rel / me] = 1;
      }
    }
    std::sort(by_lo.begin(), by_lo.end(), [](const Group *a, const Group *b) { return a->lo < b->lo; });
    for (size_t i = 0; i + 1 < by_lo.size(); ++i) dense = dense && by_lo[i]->hi <= by_lo[i + 1]->lo;
  }

  ph.mark("sort+groups");
  if (!dense && !cmp) {
    // regions of the output arrays that no block covers read back as zero
    if (want_coeff) B200_CUDA_TRY(cudaMemsetAsync(base + o_coeff, 0, out_elems * 4, s));
    if (has_q) {
      B200_CUDA_TRY(cudaMemsetAsync(base + o_qc, 0, out_elems * 4, s));
      B200_CUDA_TRY(cudaMemsetAsync(base + o_dq, 0, out_elems * 4, s));
    }
  }
  if (cmp && want_coeff) B200_CUDA_TRY(cudaMemsetAsync(base + o_coeff, 0, out_elems * 4, s));
  if (!in_order) {  // the sorted queue replaces the caller's order sent up ahead of the validation
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_blocks, sblocks, n * sizeof(B200TxBlock), cudaMemcpyHostToDevice, s));
    if (has_pred)
      B200_CUDA_TRY(cudaMemcpyAsync(base + o_preds, spreds, n * sizeof(B200TxPred), cudaMemcpyHostToDevice, s));
  }
  for (size_t gi = 0; gi < groups.size(); ++gi) {
    const Group &g = groups[gi];
    B200TxqLaunch l;
    memset(&l, 0, sizeof(l));
    l.tx_size = g.ts;
    l.in_kind = in_kind;
    l.q_kind = q_kind;
    l.bd = bd;
    l.flags = flags & ~(want_err ? 0u : B200_TXQ_BLOCK_ERROR) & ~(want_coeff ? 0u : B200_TXQ_STORE_COEFF);
    l.n_blocks = (int)g.count;
    l.blocks = (const B200TxBlock *)(base + o_blocks) + g.first;
    l.preds = has_pred ? (const B200TxPred *)(base + o_preds) + g.first : nullptr;
    l.src = base + o_src;
    l.pred = has_pred ? base + o_pred : nullptr;
    l.qrows = has_q ? (const B200QRow *)(base + o_q) : nullptr;
    l.coeff = want_coeff ? (int32_t *)(base + o_coeff) : nullptr;
    l.qcoeff = has_q ? (int32_t *)(base + o_qc) : nullptr;
    l.dqcoeff = has_q ? (int32_t *)(base + o_dq) : nullptr;
    l.eob = has_q ? (uint16_t *)(base + o_eob) + g.first : nullptr;
    l.dist = want_err ? (int64_t *)(base + o_dist) + 2 * g.first : nullptr;
    rc = txq_launch(ctx, l, s);
    if (rc != B200_OK) return bail(rc);
    if (cmp) continue;
    B200_CUDA_TRY(cudaEventRecord(ctx->group_done[gi], s));
    B200_CUDA_TRY(cudaStreamWaitEvent(cs, ctx->group_done[gi], 0));
    if (dense) {
      const size_t b0 = g.lo * 4, nb = g.elems * 4;
      if (want_coeff)
        B200_CUDA_TRY(cudaMemcpyAsync((char *)coeff + b0, base + o_coeff + b0, nb, cudaMemcpyDeviceToHost, cs));
      if (has_q) {
        B200_CUDA_TRY(cudaMemcpyAsync((char *)qcoeff + b0, base + o_qc + b0, nb, cudaMemcpyDeviceToHost, cs));
        B200_CUDA_TRY(cudaMemcpyAsync((char *)dqcoeff + b0, base + o_dq + b0, nb, cudaMemcpyDeviceToHost, cs));
      }
    }
  }
  ph.mark("enqueue");
  size_t d2h = 0;
  if (cmp) {
    // pack the levels on the device; the host learns the total before it sizes the read-back
    B200CompactLaunch cl;
    cl.n_blocks = n_blocks;
    cl.blocks = (const B200TxBlock *)(base + o_blocks);
    cl.qcoeff = (const int32_t *)(base + o_qc);
    cl.eob = (const uint16_t *)(base + o_eob);
    cl.wide = wide;
    cl.levels = base + o_lv;
    cl.level_offset = (uint32_t *)(base + o_loff);
    rc = b200_txq_compact_launch(ctx, &cl, s);
    if (rc != B200_OK) return bail(rc);
    uint32_t *h_loff = (uint32_t *)loff_dst;
    B200_CUDA_TRY(cudaMemcpyAsync(h_loff, base + o_loff, (n + 1) * 4, cudaMemcpyDeviceToHost, s));
    B200_CUDA_TRY(cudaMemcpyAsync(eob_dst, base + o_eob, n * 2, cudaMemcpyDeviceToHost, s));
    B200_CUDA_TRY(cudaEventRecord(ctx->group_done[0], s));
    B200_CUDA_TRY(cudaStreamWaitEvent(cs, ctx->group_done[0], 0));
    // meanwhile the second stream brings the per-block distortion and the dense coeff back
    if (want_err)
      B200_CUDA_TRY(cudaMemcpyAsync(dist_dst, base + o_dist, n * 16, cudaMemcpyDeviceToHost, cs));
    if (want_coeff)
      B200_CUDA_TRY(cudaMemcpyAsync(coeff, base + o_coeff, out_elems * 4, cudaMemcpyDeviceToHost, cs));
    B200_CUDA_TRY(cudaStreamSynchronize(s));
    ph.mark("kernels+offsets");
    const size_t total = h_loff[n];
    if (total > cmp->levels_cap) {
      cudaStreamSynchronize(cs);
      set_error("b200_txq_batch_host_compact: %zu levels do not fit levels_cap %zu", total, cmp->levels_cap);
      return B200_EINVAL;
    }
    if (total)
      B200_CUDA_TRY(cudaMemcpyAsync(cmp->levels, base + o_lv, total * lv_bytes, cudaMemcpyDeviceToHost, s));
    d2h = (n + 1) * 4 + n * 2 + total * lv_bytes + (want_err ? n * 16 : 0) + (want_coeff ? out_elems * 4 : 0);
  } else {
    if (!dense) {
      if (want_coeff)
        B200_CUDA_TRY(cudaMemcpyAsync(coeff, base + o_coeff, out_elems * 4, cudaMemcpyDeviceToHost, cs));
      if (has_q) {
        B200_CUDA_TRY(cudaMemcpyAsync(qcoeff, base + o_qc, out_elems * 4, cudaMemcpyDeviceToHost, cs));
        B200_CUDA_TRY(cudaMemcpyAsync(dqcoeff, base + o_dq, out_elems * 4, cudaMemcpyDeviceToHost, cs));
      }
    }
    if (has_q)
      B200_CUDA_TRY(cudaMemcpyAsync(eob_dst, base + o_eob, n * 2, cudaMemcpyDeviceToHost, cs));
    if (want_err)
      B200_CUDA_TRY(cudaMemcpyAsync(dist_dst, base + o_dist, n * 16, cudaMemcpyDeviceToHost, cs));
    if (dense) {
      // zero what no group covers, on the host, while the copies run
      size_t at = 0;
      for (const Group *g : by_lo) {
        if (g->lo > at) {
          if (want_coeff) memset(coeff + at, 0, (g->lo - at) * 4);
          if (has_q) memset(qcoeff + at, 0, (g->lo - at) * 4), memset(dqcoeff + at, 0, (g->lo - at) * 4);
        }
        at = g->hi;
      }
      if (out_elems > at) {
        if (want_coeff) memset(coeff + at, 0, (out_elems - at) * 4);
        if (has_q) memset(qcoeff + at, 0, (out_elems - at) * 4), memset(dqcoeff + at, 0, (out_elems - at) * 4);
      }
    }
  }
  B200_CUDA_TRY(cudaStreamSynchronize(cs));
  B200_CUDA_TRY(cudaStreamSynchronize(s));
  ph.mark("readback");
  const uint16_t *eob_s = (const uint16_t *)(hbase + ho_eob);
  const int64_t *dist_s = (const int64_t *)(hbase + ho_dist);
  if (!in_order) {
    const uint32_t *loff_s = (const uint32_t *)(hbase + ho_loff);
    for (size_t i = 0; i < n; ++i) {
      if (has_q && eob) eob[order[i]] = eob_s[i];
      if (want_err) {
        dist[2 * (size_t)order[i]] = dist_s[2 * i];
        dist[2 * (size_t)order[i] + 1] = dist_s[2 * i + 1];
      }
      // the levels stay packed in device-queue order; each block keeps its own offset
      if (cmp) cmp->level_offset[order[i]] = loff_s[i];
    }
    if (cmp) cmp->level_offset[n] = loff_s[n];
  }
  ph.mark("scatter");
  if (cmp && cmp->d2h_bytes) *cmp->d2h_bytes = d2h;
  return B200_OK;
}

extern "C" int b200_txq_batch_host(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                                   const B200TxBlock *blocks, const B200TxPred *preds,
                                   int n_blocks, const void *src, size_t src_elems,
                                   const void *pred, size_t pred_elems, const B200QRow *qrows,
                                   int n_qrows, int32_t *coeff, int32_t *qcoeff, int32_t *dqcoeff,
                                   size_t out_elems, uint16_t *eob, int64_t *dist) {
  return txq_batch_host_impl(ctx, in_kind, q_kind, bd, flags, blocks, preds, n_blocks, src, src_elems,
                             pred, pred_elems, qrows, n_qrows, coeff, qcoeff, dqcoeff, out_elems, eob,
                             dist, nullptr);
}

extern "C" int b200_txq_batch_host_compact(b200_ctx *ctx, int in_kind, int q_kind, int bd, unsigned flags,
                                           const B200TxBlock *blocks, const B200TxPred *preds,
                                           int n_blocks, const void *src, size_t src_elems,
                                           const void *pred, size_t pred_elems, const B200QRow *qrows,
                                           int n_qrows, size_t out_elems, void *levels,
                                           size_t levels_cap, uint32_t *level_offset, uint16_t *eob,
                                           int64_t *dist, int32_t *coeff, size_t *d2h_bytes) {
  CompactOut c = { levels, levels_cap, level_offset, d2h_bytes };
  return txq_batch_host_impl(ctx, in_kind, q_kind, bd, flags, blocks, preds, n_blocks, src, src_elems,
                             pred, pred_elems, qrows, n_qrows, coeff, nullptr, nullptr, out_elems, eob,
                             dist, &c);
}
