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


// av1_highbd_iwht4x4_16_add_c / av1_highbd_iwht4x4_1_add_c (av1_inv_txfm2d.c:20-109) behind the eob
// switch of av1_highbd_iwht4x4_add (idct.c): the lossless 4x4 inverse, one lane per block.
template <int PX>
__global__ void iwht_kernel(const ItxArgs a) {
  using pix_t = typename std::conditional<PX == 1, uint8_t, uint16_t>::type;
  const int bi = blockIdx.x * blockDim.x + threadIdx.x;
  if (bi >= a.n_blocks) return;
  const B200TxBlock d = load_desc(a.blocks + bi);
  const B200TxPred pd = a.preds[bi];
  const int eob = a.eob ? a.eob[bi] : 16;
  const int32_t *in = a.dqcoeff + d.out_offset;
  int32_t res[16];
  if (eob > 1) {
    int32_t o[16];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int32_t aa = in[i] >> 2, c = in[4 + i] >> 2, dd = in[8 + i] >> 2, b = in[12 + i] >> 2;
      aa += c, dd -= b;
      const int32_t e = (aa - dd) >> 1;
      b = e - b, c = e - c, aa -= b, dd += c;
      o[i] = aa, o[4 + i] = b, o[8 + i] = c, o[12 + i] = dd;
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      int32_t aa = o[4 * i], c = o[4 * i + 1], dd = o[4 * i + 2], b = o[4 * i + 3];
      aa += c, dd -= b;
      const int32_t e = (aa - dd) >> 1;
      b = e - b, c = e - c, aa -= b, dd += c;
      res[i] = aa, res[4 + i] = b, res[8 + i] = c, res[12 + i] = dd;
    }
  } else if (eob == 1) {
    int32_t a1 = in[0] >> 2;
    const int32_t e1 = a1 >> 1;
    a1 -= e1;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int32_t t = i == 0 ? a1 : e1, e = t >> 1;
      res[i] = t - e, res[4 + i] = e, res[8 + i] = e, res[12 + i] = e;
    }
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) res[i] = 0;
  }
  const int pmax = (1 << a.bd) - 1;
  const int vis_w = d.reserved ? (d.reserved & 255) : 4, vis_h = d.reserved ? (d.reserved >> 8) : 4;
  unsigned long long sse = 0;
#pragma unroll
  for (int r = 0; r < 4; ++r)
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const size_t po = (size_t)pd.pred_offset + (size_t)r * pd.pred_stride + c;
      const int v = min(max((int)static_cast<const pix_t *>(a.pred)[po] + res[4 * r + c], 0), pmax);
      if (a.recon) static_cast<pix_t *>(a.recon)[po] = (pix_t)v;
      if (a.src && c < vis_w && r < vis_h) {
        const int df = (int)static_cast<const pix_t *>(a.src)[(size_t)d.src_offset + (size_t)r * d.src_stride + c] - v;
        sse += (unsigned)(df * df);
      }
    }
  if (a.dist) {
    if (PX == 2 && a.bd > 8) {
      const int sh = 2 * (a.bd - 8);
      sse = (sse + (1ull << (sh - 1))) >> sh;
    }
    a.dist[bi] = 16 * (long long)(unsigned)sse;
  }
}

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
  if (l.lossless) {
    if (l.tx_size != B200_TX_4X4) {
      set_error("b200_itx_launch: lossless blocks are 4x4 (WHT)");
      return B200_EINVAL;
    }
    if (l.highbd) iwht_kernel<2><<<(l.n_blocks + 127) / 128, 128, 0, s>>>(a);
    else iwht_kernel<1><<<(l.n_blocks + 127) / 128, 128, 0, s>>>(a);
    B200_CUDA_TRY(cudaGetLastError());
    return B200_OK;
  }
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
