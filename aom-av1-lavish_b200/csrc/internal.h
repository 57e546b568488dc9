// internal.h -- shared between the translation units of libb200av1.so (not installed).
#pragma once
#include <cuda_runtime.h>
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

}  // namespace b200

struct b200_ctx {
  int device;
  int sm_count;
  cudaStream_t stream;
  int16_t *d_iscan;   // all inverse scan tables, see txq.cu
  unsigned long long *d_pmask;  // scan prefix masks of the <= 64-sample sizes, see txq.cu
  // scratch for the host-buffer (drop-in) tier
  void *d_scratch;
  size_t scratch_bytes;
  void *h_scratch;      // pinned staging of the host-buffer tier (sorted descriptors, eob, dist)
  size_t h_scratch_bytes;
  cudaStream_t copy_stream;            // device->host leg of the host-buffer tier
  cudaEvent_t group_done[B200_TX_SIZES_ALL];  // one per tx-size group of a batch
  cudaEvent_t copy_done;
  void *d_me_sites;   // search-site tables of the four served methods, see me_host.cuh
};

namespace b200 {
// per-TU table upload / launchers
int txq_init_device(b200_ctx *ctx);
int txq_launch(b200_ctx *ctx, const B200TxqLaunch &l, cudaStream_t s);
int blockops_init_device(b200_ctx *ctx);
int me_init_device(b200_ctx *ctx);
int scratch_reserve(b200_ctx *ctx, size_t bytes);
// hadamard.cu: single-call helpers behind the drop-in tier
int int_pro_launch(b200_ctx *ctx, short *out, const unsigned char *ref, int ref_stride, int width,
                   int height, int norm, int col, cudaStream_t s);
int vector_var_launch(b200_ctx *ctx, const short *ref, const short *src, int bwl, int *out, cudaStream_t s);
int satd_launch(b200_ctx *ctx, const void *coeff, int lp, int length, int *out, cudaStream_t s);
}  // namespace b200
