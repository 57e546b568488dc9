// itx.cu -- host side of the inverse transform + reconstruction + pixel-domain distortion launch
// (b200_itx_launch, include/b200av1.h).  Kernels: itx_kernel.cuh, instantiated in itx_inst_*.cu.
#include <mutex>

#include "itx_kernel.cuh"

namespace b200 {

int itx_group_a(ItxKernelInfo (*table)[2], bool upload);
int itx_group_b(ItxKernelInfo (*table)[2], bool upload);
int itx_group_c(ItxKernelInfo (*table)[2], bool upload);
int itx_group_d(ItxKernelInfo (*table)[2], bool upload);
int itx_group_e(ItxKernelInfo (*table)[2], bool upload);
typedef int (*ItxGroupFn)(ItxKernelInfo (*)[2], bool);
static const ItxGroupFn kItxGroups[] = { itx_group_a, itx_group_b, itx_group_c, itx_group_d, itx_group_e };

static ItxKernelInfo g_itx[B200_TX_SIZES_ALL][2];
static int g_itx_ctas[B200_TX_SIZES_ALL][2];
static std::once_flag g_itx_once;

int itx_init_device(b200_ctx *ctx) {
  (void)ctx;
  std::call_once(g_itx_once, [] {
    for (ItxGroupFn g : kItxGroups) g(g_itx, false);
  });
  for (ItxGroupFn g : kItxGroups) {
    const int rc = g(g_itx, true);  // per-TU constant tables on the current device
    if (rc != B200_OK) return rc;
  }
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int k = 0; k < 2; ++k) {
      const ItxKernelInfo &e = g_itx[ts][k];
      B200_CUDA_TRY(cudaFuncSetAttribute(e.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, e.smem_bytes));
      int nb = 0;
      B200_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, e.fn, kWarpsPerCta * 32, e.smem_bytes));
      g_itx_ctas[ts][k] = nb > 0 ? nb : 1;
    }
  return B200_OK;
}

int itx_launch(b200_ctx *ctx, const B200ItxLaunch &l, cudaStream_t s) {
  if (l.tx_size < 0 || l.tx_size >= B200_TX_SIZES_ALL || l.n_blocks < 0 || !l.blocks || !l.preds ||
      !l.dqcoeff || !l.pred) {
    set_error("b200_itx_launch: bad tx_size / n_blocks / null blocks, preds, dqcoeff or pred");
    return B200_EINVAL;
  }
  if (l.bd != 8 && l.bd != 10 && l.bd != 12) {
    set_error("b200_itx_launch: bd must be 8, 10 or 12 (got %d)", l.bd);
    return B200_EINVAL;
  }
  if (!l.highbd && l.bd != 8) {
    set_error("b200_itx_launch: uint8 pixels imply bd 8 (got %d)", l.bd);
    return B200_EINVAL;
  }
  if (!l.recon && !l.dist) {
    set_error("b200_itx_launch: neither recon nor dist requested");
    return B200_EINVAL;
  }
  if (l.dist && !l.src) {
    set_error("b200_itx_launch: dist needs the source plane");
    return B200_EINVAL;
  }
  if (l.n_blocks == 0) return B200_OK;
  const int k = l.highbd ? 1 : 0;
  const ItxKernelInfo &e = g_itx[l.tx_size][k];
  ItxArgs a;
  a.blocks = l.blocks;
  a.preds = l.preds;
  a.dqcoeff = l.dqcoeff;
  a.eob = l.eob;
  a.src = l.dist ? l.src : nullptr;
  a.pred = l.pred;
  a.recon = l.recon;
  a.dist = l.dist;
  a.n_blocks = l.n_blocks;
  a.bd = l.bd;
  const int n_tiles = (l.n_blocks + e.bpw - 1) / e.bpw;
  int grid = (n_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
  const int resident = ctx->sm_count * g_itx_ctas[l.tx_size][k];
  if (grid > resident) grid = resident;  // persistent: warps stride over the tiles
  e.fn<<<grid, kWarpsPerCta * 32, e.smem_bytes, s>>>(a);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_itx_launch(b200_ctx *ctx, const B200ItxLaunch *launch, void *stream) {
  if (!ctx || !launch) return B200_EINVAL;
  return b200::itx_launch(ctx, *launch, stream ? (cudaStream_t)stream : ctx->stream);
}
