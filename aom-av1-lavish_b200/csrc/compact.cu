// compact.cu -- compact result format of the transform+quantise path.
//
// The dense result of one transform block is coeff + qcoeff + dqcoeff as int32 = 12 bytes
// per sample; what the encoder consumes after the quantiser is the levels in SCAN order up
// to eob (av1_write_coeffs_txb / av1_cost_coeffs_txb walk scan[0 .. eob), av1/encoder/
// encodetxb.c), and dqcoeff is a function of the level: (|q| * dequant[rc != 0]) >> log_scale
// with the sign put back (av1_quantize.c:61-63).  This file packs the dense qcoeff the fused
// kernel left in device memory into "levels": block i owns eob[i] consecutive entries at
// level_offset[i] (exclusive prefix sum of eob over the batch, entry n = total), int16 for
// the low-bit-depth quantisers (their levels are clamped to int16 by construction,
// av1_quantize.c:48-52, aom_dsp/quantize.c:142) and int32 for the high-bit-depth ones.
//
//   compact_sums      per 2048-block chunk: sum of eob
//   compact_scan      one CTA: exclusive scan of the chunk sums (+ total)
//   compact_offsets   per chunk: exclusive scan of eob -> level_offset
//   compact_gather    one warp per block: levels[off + i] = qcoeff[out_offset + scan[i]]
#include <string.h>

#include "internal.h"

namespace b200 {

constexpr int kChunk = 2048;   // blocks per CTA of the prefix-sum kernels
constexpr int kScanThreads = 256;
constexpr int kMaxChunks = 1 << 16;

// Scratch of b200_txq_compact_launch_noalloc: the per-chunk sums, in a file-scope device array instead of the
// context's cudaMalloc'ed d_partials.  kStaticMaxChunks * 4 B = 16 KB, held by every GPU that loads the library
// until the process exits.  In HOST code this name is only the host-side stand-in of the symbol, never a
// device address: compact_init_device asks the runtime for the address (cudaGetSymbolAddress, once per context
// with its device current) and the launches below pass that pointer.
constexpr int kStaticMaxChunks = B200_COMPACT_NOALLOC_MAX_BLOCKS / kChunk;
static_assert(kStaticMaxChunks * kChunk == B200_COMPACT_NOALLOC_MAX_BLOCKS, "maximum is a whole number of chunks");
static __device__ __align__(256) uint32_t g_static_partials[kStaticMaxChunks];

__constant__ uint32_t c_scan_off[B200_TX_SIZES_ALL];  // element offset of the size's tables in d_scan
__constant__ uint16_t c_scan_n[B200_TX_SIZES_ALL];    // max_eob of the size

__device__ __forceinline__ uint32_t cta_sum(uint32_t v, uint32_t *sh) {
  v = __reduce_add_sync(0xffffffffu, v);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  uint32_t t = 0;
  for (int w = 0; w < kScanThreads / 32; ++w) t += sh[w];
  __syncthreads();
  return t;
}

__global__ void __launch_bounds__(kScanThreads) compact_sums(const uint16_t *eob, int n, uint32_t *partials) {
  __shared__ uint32_t sh[kScanThreads / 32];
  const int base = blockIdx.x * kChunk;
  uint32_t v = 0;
  for (int i = threadIdx.x; i < kChunk; i += kScanThreads)
    if (base + i < n) v += eob[base + i];
  v = cta_sum(v, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = v;
}

// exclusive scan of `partials` in place; total -> *total_out
__global__ void __launch_bounds__(1024) compact_scan(uint32_t *partials, int n_chunks, uint32_t *total_out) {
  __shared__ uint32_t sh[32];
  __shared__ uint32_t carry;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b = 0; b < n_chunks; b += 1024) {
    const int i = b + threadIdx.x;
    const uint32_t v = i < n_chunks ? partials[i] : 0;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += t;
    }
    if (lane == 31) sh[warp] = inc;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = sh[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      sh[lane] = w;
    }
    __syncthreads();
    const uint32_t before = carry + (warp ? sh[warp - 1] : 0) + inc - v;
    if (i < n_chunks) partials[i] = before;
    __syncthreads();
    if (threadIdx.x == 0) carry += sh[31];
    __syncthreads();
  }
  if (threadIdx.x == 0) *total_out = carry;
}

__global__ void __launch_bounds__(kScanThreads) compact_offsets(const uint16_t *eob, int n,
                                                                 const uint32_t *partials,
                                                                 uint32_t *level_offset) {
  __shared__ uint32_t sh[kScanThreads / 32];
  constexpr int PER = kChunk / kScanThreads;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int first = blockIdx.x * kChunk + threadIdx.x * PER;
  uint32_t e[PER], mine = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    e[j] = first + j < n ? eob[first + j] : 0;
    mine += e[j];
  }
  uint32_t inc = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += t;
  }
  if (lane == 31) sh[warp] = inc;
  __syncthreads();
  uint32_t before = partials[blockIdx.x] + inc - mine;
  for (int w = 0; w < warp; ++w) before += sh[w];
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    if (first + j < n) level_offset[first + j] = before;
    before += e[j];
  }
}

template <class T>
__global__ void __launch_bounds__(256) compact_gather(const B200TxBlock *blocks, const int32_t *qcoeff,
                                                      const uint16_t *eob, const uint32_t *level_offset,
                                                      const int16_t *scan, T *levels, int n) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (bi >= n) return;
  const B200TxBlock d = blocks[bi];
  const int e = eob[bi];
  if (e == 0) return;
  const int16_t *sc = scan + c_scan_off[d.tx_size] + scan_class_of(d.tx_type) * c_scan_n[d.tx_size];
  const int32_t *q = qcoeff + d.out_offset;
  T *out = levels + level_offset[bi];
  for (int i = lane; i < e; i += 32) out[i] = (T)q[sc[i]];
}

int compact_init_device(b200_ctx *ctx) {
  uint32_t off[B200_TX_SIZES_ALL];
  uint16_t cnt[B200_TX_SIZES_ALL];
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    off[ts] = (uint32_t)scan_dev_offset(ts);
    cnt[ts] = (uint16_t)b200_tx_max_eob(ts);
  }
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_scan_off, off, sizeof(off)));
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_scan_n, cnt, sizeof(cnt)));
  B200_CUDA_TRY(cudaMalloc(&ctx->d_partials, sizeof(uint32_t) * kMaxChunks));
  // device address of the static scratch on THIS context's device (ctx_init runs with it current)
  void *p = nullptr;
  B200_CUDA_TRY(cudaGetSymbolAddress(&p, g_static_partials));
  ctx->d_static_partials = static_cast<uint32_t *>(p);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_txq_compact_launch(b200_ctx *ctx, const B200CompactLaunch *l, void *stream) {
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (!l->blocks || !l->qcoeff || !l->eob || !l->levels || !l->level_offset) {
    set_error("b200_txq_compact_launch: missing buffer");
    return B200_EINVAL;
  }
  const int n = l->n_blocks;
  const int n_chunks = (n + kChunk - 1) / kChunk;
  if (n_chunks > kMaxChunks) {
    set_error("b200_txq_compact_launch: more than %d blocks in one batch", kMaxChunks * kChunk);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  if (n == 0) {
    B200_CUDA_TRY(cudaMemsetAsync(l->level_offset, 0, sizeof(uint32_t), s));
    return B200_OK;
  }
  compact_sums<<<n_chunks, kScanThreads, 0, s>>>(l->eob, n, ctx->d_partials);
  compact_scan<<<1, 1024, 0, s>>>(ctx->d_partials, n_chunks, l->level_offset + n);
  compact_offsets<<<n_chunks, kScanThreads, 0, s>>>(l->eob, n, ctx->d_partials, l->level_offset);
  if (l->wide)
    compact_gather<int32_t><<<(n + 7) / 8, 256, 0, s>>>(l->blocks, l->qcoeff, l->eob, l->level_offset,
                                                        ctx->d_scan, static_cast<int32_t *>(l->levels), n);
  else
    compact_gather<int16_t><<<(n + 7) / 8, 256, 0, s>>>(l->blocks, l->qcoeff, l->eob, l->level_offset,
                                                        ctx->d_scan, static_cast<int16_t *>(l->levels), n);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

// The same packing without any allocation (no counterpart in the reference; see include/b200av1.h): the
// per-chunk sums live in g_static_partials.  Nothing in the scratch has to start at zero -- compact_sums writes
// every entry [0, n_chunks) the two later kernels read -- so nothing is cleared; the array keeps the previous
// call's sums.  All calls on a device share it: two calls whose kernels overlap on the GPU (different streams
// without an ordering between them) race.
extern "C" int b200_txq_compact_launch_noalloc(b200_ctx *ctx, const B200CompactLaunch *l, void *stream) {
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (!l->blocks || !l->qcoeff || !l->eob || !l->levels || !l->level_offset) {
    set_error("b200_txq_compact_launch_noalloc: missing buffer");
    return B200_EINVAL;
  }
  const int n = l->n_blocks;
  if (n > B200_COMPACT_NOALLOC_MAX_BLOCKS) {
    set_error("b200_txq_compact_launch_noalloc: %d blocks exceed the static scratch's maximum of %d "
              "(B200_COMPACT_NOALLOC_MAX_BLOCKS); use b200_txq_compact_launch or split the batch",
              n, B200_COMPACT_NOALLOC_MAX_BLOCKS);
    return B200_EINVAL;
  }
  if (!ctx->d_static_partials) {
    set_error("b200_txq_compact_launch_noalloc: static scratch not resolved on device %d", ctx->device);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  if (n == 0) {
    B200_CUDA_TRY(cudaMemsetAsync(l->level_offset, 0, sizeof(uint32_t), s));
    return B200_OK;
  }
  const int n_chunks = (n + kChunk - 1) / kChunk;
  uint32_t *partials = ctx->d_static_partials;
  compact_sums<<<n_chunks, kScanThreads, 0, s>>>(l->eob, n, partials);
  compact_scan<<<1, 1024, 0, s>>>(partials, n_chunks, l->level_offset + n);
  compact_offsets<<<n_chunks, kScanThreads, 0, s>>>(l->eob, n, partials, l->level_offset);
  if (l->wide)
    compact_gather<int32_t><<<(n + 7) / 8, 256, 0, s>>>(l->blocks, l->qcoeff, l->eob, l->level_offset,
                                                        ctx->d_scan, static_cast<int32_t *>(l->levels), n);
  else
    compact_gather<int16_t><<<(n + 7) / 8, 256, 0, s>>>(l->blocks, l->qcoeff, l->eob, l->level_offset,
                                                        ctx->d_scan, static_cast<int16_t *>(l->levels), n);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

// Host-side expander of one block of the compact format: dense qcoeff / dqcoeff as the
// quantisers without qmatrix write them (av1_quantize.c:36-68, aom_dsp/quantize.c:142-160).
extern "C" int b200_txq_expand_block(int tx_size, int tx_type, const void *levels, int wide, int eob,
                                     const B200QRow *row, int32_t *qcoeff, int32_t *dqcoeff) {
  if (tx_size < 0 || tx_size >= B200_TX_SIZES_ALL || tx_type < 0 || tx_type >= 16 || !levels || !row)
    return B200_EINVAL;
  const int n = b200_tx_max_eob(tx_size), ls = tx_info(tx_size).log_scale;
  if (eob < 0 || eob > n) return B200_EINVAL;
  const int16_t *scan = b200_scan(tx_size, tx_type);
  if (qcoeff) memset(qcoeff, 0, sizeof(int32_t) * n);
  if (dqcoeff) memset(dqcoeff, 0, sizeof(int32_t) * n);
  for (int i = 0; i < eob; ++i) {
    const int rc = scan[i];
    const int32_t q = wide ? static_cast<const int32_t *>(levels)[i]
                           : (int32_t) static_cast<const int16_t *>(levels)[i];
    if (qcoeff) qcoeff[rc] = q;
    if (dqcoeff) {
      const int32_t sign = q < 0 ? -1 : 0, a = (q ^ sign) - sign;
      const int32_t dq = (int32_t)((uint32_t)a * (uint32_t)(int32_t)row->dequant[rc != 0]) >> ls;
      dqcoeff[rc] = (dq ^ sign) - sign;
    }
  }
  return B200_OK;
}
