// internal.h -- shared between the translation units of libb200av1.so (not installed).
#pragma once
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/b200av1.h"

namespace b200 {

// Static transform configuration of one TX_SIZE (av1_fwd_txfm2d.c:314-381).
struct TxInfo {
  int w, h, wk, hk;        // size and kept (<= 32) size
  int shift[3];            // fwd_shift_* (positive = up-shift)
  int cos_bit_col, cos_bit_row;
  int rect;                // |log2(w/h)| == 1 -> sqrt(2) rescale
  int log_scale;           // av1_get_tx_scale, av1/common/idct.c:24
};
const TxInfo &tx_info(int tx_size);

void set_error(const char *fmt, ...);
#define B200_CUDA_TRY(expr)                                                        \
  do {                                                                             \
    cudaError_t e_ = (expr);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      b200::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,         \
                      cudaGetErrorString(e_));                                     \
      return B200_ECUDA;                                                           \
    }                                                                              \
  } while (0)

// av1_cospi_arr_data (av1_txfm.c:18-60), bits 10..13: round(cos(j*pi/128) * 2^bit).  Returns
// false when the computed table is not the reference's literal one (FNV-1a over the 256
// entries, constant taken from the reference table), whatever the host libm does.
inline bool fill_cospi_tables(int32_t (*cosv)[64]) {
  const double pi = 3.14159265358979323846;
  unsigned long long fnv = 0xcbf29ce484222325ull;
  for (int b = 0; b < 4; ++b)
    for (int j = 0; j < 64; ++j) {
      cosv[b][j] = (int32_t)floor(cos(pi * j / 128.0) * (double)(1 << (10 + b)) + 0.5);
      fnv = (fnv ^ (unsigned)cosv[b][j]) * 0x100000001b3ull;
    }
  return fnv == 0x66b250b93dc05948ull;
}

}  // namespace b200

struct b200_ctx {
  int device;
  int sm_count;
  cudaStream_t stream;
  int16_t *d_iscan;   // all inverse scan tables, see txq.cu
  int16_t *d_scan;    // all forward scan tables, same layout (compact.cu)
  uint32_t *d_partials;  // per-chunk sums of the compaction prefix sum (compact.cu)
  uint32_t *d_static_partials;  // device address of compact.cu's file-scope scratch (no-alloc entry point)
  unsigned long long *d_pmask;  // scan prefix masks of the <= 64-sample sizes, see txq.cu
  // scratch for the host-buffer (drop-in) tier
  void *d_scratch;
  size_t scratch_bytes;
  void *h_scratch;      // pinned staging of the host-buffer tier (sorted descriptors, eob, dist)
  size_t h_scratch_bytes;
  cudaStream_t copy_stream;            // device->host leg of the host-buffer tier
  cudaEvent_t group_done[B200_TX_SIZES_ALL];  // one per tx-size group of a batch
  cudaEvent_t copy_done;
  cudaStream_t aux_stream;             // second compute stream of b200_me_batch_host (alternating chunks)
  void *d_me_sites;   // search-site tables of the four served methods, see me_host.cuh
};

namespace b200 {
// per-TU table upload / launchers
int txq_init_device(b200_ctx *ctx);
int txq_launch(b200_ctx *ctx, const B200TxqLaunch &l, cudaStream_t s);
int blockops_init_device(b200_ctx *ctx);
int me_init_device(b200_ctx *ctx);
int compact_init_device(b200_ctx *ctx);
int itx_init_device(b200_ctx *ctx);
int itx_launch(b200_ctx *ctx, const B200ItxLaunch &l, cudaStream_t s);
// element offset of the scan tables of `tx_size` inside d_scan / d_iscan (3 classes of
// max_eob entries each, see txq.cu) and the scan class of a tx_type
size_t scan_dev_offset(int tx_size);
__host__ __device__ inline int scan_class_of(int tx_type) { return tx_type < 10 ? 0 : ((tx_type & 1) ? 2 : 1); }
int round_shift_array_launch(b200_ctx *ctx, int32_t *arr, int size, int bit, cudaStream_t s);
int scratch_reserve(b200_ctx *ctx, size_t bytes);
int host_scratch_reserve(b200_ctx *ctx, size_t bytes);  // pinned staging (capi.cu)
// hadamard.cu: single-call helpers behind the drop-in tier
int int_pro_launch(b200_ctx *ctx, short *out, const unsigned char *ref, int ref_stride, int width,
                   int height, int norm, int col, cudaStream_t s);
int vector_var_launch(b200_ctx *ctx, const short *ref, const short *src, int bwl, int *out, cudaStream_t s);
int satd_launch(b200_ctx *ctx, const void *coeff, int lp, int length, int *out, cudaStream_t s);
}  // namespace b200
