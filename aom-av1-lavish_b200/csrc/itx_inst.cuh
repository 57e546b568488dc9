// itx_inst.cuh -- helper for the itx_inst_*.cu translation units: each instantiates the inverse
// kernel for a few transform sizes (parallel compilation) and uploads its private copy of the
// cosine / sine tables (static __constant__ in txfm1d.cuh) in its group function.
#pragma once
#include "itx_kernel.cuh"
#include "itx_direct.cuh"
#include "txq_inst.cuh"  // upload_txfm_constants()

namespace b200 {
template <int W, int H>
static void fill_itx_entries(ItxKernelInfo *e) {
  e[0] = itx_kernel_info<W, H, 1>();
  e[1] = itx_kernel_info<W, H, 2>();
}
}  // namespace b200

#define B200_ITX_GROUP_BEGIN(NAME)                                  \
  namespace b200 {                                                  \
  int itx_group_##NAME(ItxKernelInfo (*table)[2], bool upload) {    \
    if (upload) return upload_txfm_constants();
#define B200_ITX_SIZE(ENUM, W, H) fill_itx_entries<W, H>(table[ENUM]);
#define B200_ITX_GROUP_END \
  return B200_OK;          \
  }                        \
  }
