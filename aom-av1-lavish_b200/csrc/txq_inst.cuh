// txq_inst.cuh -- helper for the txq_inst_*.cu translation units: each one instantiates
// the kernel for a few transform sizes (parallel compilation) and owns a private copy of
// the cosine tables in constant memory, which it uploads in its init function.
#pragma once
#include "txq_kernel.cuh"
#include "txq_small.cuh"

namespace b200 {

static int upload_txfm_constants() {
  int32_t cosv[4][64], ncosv[4][64];
  if (!fill_cospi_tables(cosv)) {
    set_error("cospi table differs from av1_cospi_arr_data (host libm rounding?)");
    return B200_EINVAL;
  }
  for (int b = 0; b < 4; ++b)
    for (int j = 0; j < 64; ++j) ncosv[b][j] = -cosv[b][j];
  // av1_sinpi_arr_data (av1_txfm.c:62-69), bits 10..13: round(sqrt2*sin(j*pi/9)*2/3 * 2^bit)
  // with entries 1 and 2 adjusted by the codec so that s1 + s2 == s4
  const int32_t sinv[4][5] = { { 0, 330, 621, 836, 951 },
                               { 0, 660, 1241, 1672, 1901 },
                               { 0, 1321, 2482, 3344, 3803 },
                               { 0, 2642, 4964, 6689, 7606 } };
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_cospi, cosv, sizeof(cosv)));
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_ncospi, ncosv, sizeof(ncosv)));
  B200_CUDA_TRY(cudaMemcpyToSymbol(c_sinpi, sinv, sizeof(sinv)));
  return B200_OK;
}

template <int W, int H>
static void fill_entries(TxqKernelInfo *e) {
  e[B200_IN_RESIDUAL] = txq_kernel_info<W, H, B200_IN_RESIDUAL>();
  e[B200_IN_PIXELS8] = txq_kernel_info<W, H, B200_IN_PIXELS8>();
  e[B200_IN_PIXELS16] = txq_kernel_info<W, H, B200_IN_PIXELS16>();
}

template <int W, int H>
static void fill_entries_small(TxqKernelInfo *e) {
  e[B200_IN_RESIDUAL] = txq_small_kernel_info<W, H, B200_IN_RESIDUAL>();
  e[B200_IN_PIXELS8] = txq_small_kernel_info<W, H, B200_IN_PIXELS8>();
  e[B200_IN_PIXELS16] = txq_small_kernel_info<W, H, B200_IN_PIXELS16>();
}

}  // namespace b200

#define B200_TXQ_GROUP_BEGIN(NAME)                                   \
  namespace b200 {                                                   \
  int txq_group_##NAME(TxqKernelInfo (*table)[3], bool upload) {     \
    if (upload) return upload_txfm_constants();
#define B200_TXQ_SIZE(ENUM, W, H) fill_entries<W, H>(table[ENUM]);
#define B200_TXQ_SIZE_SMALL(ENUM, W, H) fill_entries_small<W, H>(table[ENUM]);
// lane-owns-block kernel for the residual input, tiled kernel for the two-operand pixel inputs: used for
// 16x4, whose lane-owned pixel kernel sits at 255 registers with spills (configs[3]: 0.42 -> 0.46 of the
// roofline; 4x16 is the other way round, 0.39 lane-owned vs 0.31 tiled, and keeps the lane-owned one)
#define B200_TXQ_SIZE_SMALL_RESID(ENUM, W, H) \
  fill_entries<W, H>(table[ENUM]);            \
  table[ENUM][B200_IN_RESIDUAL] = txq_small_kernel_info<W, H, B200_IN_RESIDUAL>();
#define B200_TXQ_GROUP_END \
  return B200_OK;          \
  }                        \
  }
