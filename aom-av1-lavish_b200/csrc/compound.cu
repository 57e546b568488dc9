// compound.cu -- compound / masked / OBMC SAD and variance and the compound predictor builders
// (SURVEY.md section 8f rank 4): the members of aom_variance_fn_ptr_t (aom_dsp/variance.h:92-113)
// beyond sdf / vf / svf -- sdaf (dist-wtd flavour), svaf, jsvaf / jsdaf, msdf, msvf, osdf, ovf, osvf --
// for a batch of blocks in one launch.
//
// Reference semantics: aom_dsp/sad_av1.c:21-262, aom_dsp/sad.c:57,290, aom_dsp/variance.c:147-181,
// 285-318,530-760,749-1180, aom_dsp/blend.h:23-40.  One warp per block; every flavour is
//   predictor sample (optionally the 2-tap 1/8-pel bilinear interpolation of variance.c:64-116)
//   -> optional blend with second_pred (average / distance weighted / 6-bit mask)
//   -> measure (SAD, variance, OBMC SAD, OBMC variance) or store (the *_pred builders),
// evaluated per pixel with the reference's intermediate roundings (each interpolation pass rounds to a
// pixel, the blend rounds once) so that every flavour is bit-exact.
#include <string.h>

#include "internal.h"

namespace b200 {
namespace {

constexpr int kWarps = 4;

__constant__ uint8_t c_bil2[8][2] = { { 128, 0 }, { 112, 16 }, { 96, 32 }, { 80, 48 },
                                      { 64, 64 }, { 48, 80 },  { 32, 96 }, { 16, 112 } };

template <class T>
__device__ __forceinline__ T wsum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

struct OpTraits {
  bool subpel, to_pred, obmc, sad;
  int blend;  // 0 none, 1 avg, 2 dist-wtd, 3 mask
};
__device__ __forceinline__ OpTraits traits_of(int op) {
  switch (op) {
    case B200_CMP_MASKED_SAD: return { false, false, false, true, 3 };
    case B200_CMP_MASKED_SUBPEL_VAR: return { true, false, false, false, 3 };
    case B200_CMP_SUBPEL_AVG_VAR: return { true, false, false, false, 1 };
    case B200_CMP_DIST_WTD_SUBPEL_AVG_VAR: return { true, false, false, false, 2 };
    case B200_CMP_DIST_WTD_SAD_AVG: return { false, false, false, true, 2 };
    case B200_CMP_OBMC_SAD: return { false, false, true, true, 0 };
    case B200_CMP_OBMC_VAR: return { false, false, true, false, 0 };
    case B200_CMP_OBMC_SUBPEL_VAR: return { true, false, true, false, 0 };
    case B200_CMP_AVG_PRED: return { false, true, false, false, 1 };
    case B200_CMP_DIST_WTD_AVG_PRED: return { false, true, false, false, 2 };
    default: return { false, true, false, false, 3 };  // B200_CMP_MASK_PRED
  }
}

template <class PIX>
__device__ __forceinline__ void compound_block(const B200CompoundLaunch &l, int bi, int lane) {
  const B200PixBlock d = l.blocks[bi];
  const OpTraits t = traits_of(l.op);
  const PIX *pre = static_cast<const PIX *>(l.pre) + d.a_offset;
  const PIX *tgt = static_cast<const PIX *>(l.tgt) + d.b_offset;  // unused by OBMC / builders
  const PIX *second = static_cast<const PIX *>(l.second_pred) + d.aux_offset;
  const int w = d.w, h = d.h, n = w * h, ps = d.a_stride, ts = d.b_stride;
  const uint8_t *mask = nullptr;
  int ms = 0, inv = 0;
  if (t.blend == 3) {
    const B200PixMask m = l.masks[bi];
    mask = l.mask + m.mask_offset;
    ms = m.mask_stride;
    inv = m.invert_mask;
  }
  const int f0 = c_bil2[d.xoff][0], f1 = c_bil2[d.xoff][1];
  const int g0 = c_bil2[d.yoff][0], g1 = c_bil2[d.yoff][1];
  const int32_t *wsrc = t.obmc ? l.wsrc + d.aux_offset : nullptr;
  const int32_t *omask = t.obmc ? l.obmc_mask + d.aux_offset : nullptr;
  PIX *outp = t.to_pred ? static_cast<PIX *>(l.comp_pred) + d.b_offset : nullptr;

  uint32_t sad = 0, sse_lo = 0;
  int64_t sum = 0;
  uint64_t sse = 0;
  for (int i = lane; i < n; i += 32) {
    const int y = i / w, x = i - y * w;
    const PIX *p = pre + y * ps + x;
    int pv;
    if (t.subpel) {
      const int h0 = ((int)p[0] * f0 + (int)p[1] * f1 + 64) >> 7;
      const int h1 = ((int)p[ps] * f0 + (int)p[ps + 1] * f1 + 64) >> 7;
      pv = (PIX)((h0 * g0 + h1 * g1 + 64) >> 7);
    } else {
      pv = p[0];
    }
    if (t.blend == 1) {
      pv = ((int)second[i] + pv + 1) >> 1;
    } else if (t.blend == 2) {
      pv = ((int)second[i] * l.bck_offset + pv * l.fwd_offset + 8) >> 4;
    } else if (t.blend == 3) {
      const int m = mask[y * ms + x], s = second[i];
      pv = inv ? (m * s + (64 - m) * pv + 32) >> 6 : (m * pv + (64 - m) * s + 32) >> 6;
    }
    if (t.to_pred) {
      outp[i] = (PIX)pv;
    } else if (t.obmc) {
      const int df = wsrc[i] - pv * omask[i];
      if (t.sad) {
        sad += (uint32_t)((abs(df) + 2048) >> 12);
      } else {
        const int r = df < 0 ? -((-df + 2048) >> 12) : (df + 2048) >> 12;
        sum += r;
        sse += (uint64_t)(int64_t)(r * r);
        sse_lo += (uint32_t)(r * r);
      }
    } else {
      const int df = pv - (int)tgt[y * ts + x];
      if (t.sad) {
        sad += (uint32_t)abs(df);
      } else {
        sum += df;
        sse += (uint32_t)(df * df);
      }
    }
  }
  if (t.to_pred) return;
  uint32_t out0, out1 = 0;
  if (t.sad) {
    out0 = wsum(sad);
  } else {
    sum = wsum(sum);
    sse = wsum(sse);
    sse_lo = wsum(sse_lo);
    const bool hb = sizeof(PIX) == 2;
    uint32_t sse32;
    int isum;
    if (hb && l.bd > 8) {
      const int s2 = l.bd - 8, s1 = 2 * s2;
      sse32 = (uint32_t)((sse + ((uint64_t)1 << (s1 - 1))) >> s1);
      isum = (int)((sum + ((int64_t)1 << (s2 - 1))) >> s2);
    } else {
      // the 8-bit OBMC variance accumulates its squares in unsigned int (variance.c:933-951)
      sse32 = (t.obmc && !hb) ? sse_lo : (uint32_t)sse;
      isum = (int)sum;
    }
    const int64_t mean2 = ((int64_t)isum * isum) / n;
    if (hb && l.bd > 8) {
      const int64_t var = (int64_t)sse32 - mean2;
      out0 = var >= 0 ? (uint32_t)var : 0u;
    } else {
      out0 = sse32 - (uint32_t)mean2;
    }
    out1 = sse32;
  }
  if (lane == 0) {
    l.out[2 * (size_t)bi] = out0;
    l.out[2 * (size_t)bi + 1] = out1;
  }
}

__global__ void __launch_bounds__(kWarps * 32) compound_kernel(const B200CompoundLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  if (l.highbd) compound_block<uint16_t>(l, bi, lane);
  else compound_block<uint8_t>(l, bi, lane);
}

}  // namespace
}  // namespace b200

extern "C" int b200_compound_launch(b200_ctx *ctx, const B200CompoundLaunch *l, void *stream) {
  using namespace b200;
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (l->n_blocks == 0) return B200_OK;
  const int op = l->op;
  const bool builder = op == B200_CMP_AVG_PRED || op == B200_CMP_DIST_WTD_AVG_PRED || op == B200_CMP_MASK_PRED;
  const bool obmc = op == B200_CMP_OBMC_SAD || op == B200_CMP_OBMC_VAR || op == B200_CMP_OBMC_SUBPEL_VAR;
  const bool masked = op == B200_CMP_MASKED_SAD || op == B200_CMP_MASKED_SUBPEL_VAR || op == B200_CMP_MASK_PRED;
  if (op < 0 || op > B200_CMP_MASK_PRED || !l->blocks || !l->pre || (builder ? !l->comp_pred : !l->out) ||
      (obmc ? (!l->wsrc || !l->obmc_mask) : !l->second_pred) || (!obmc && !builder && !l->tgt) ||
      (masked && (!l->masks || !l->mask))) {
    set_error("b200_compound_launch: missing buffer or bad op %d", op);
    return B200_EINVAL;
  }
  if (l->highbd && l->bd != 8 && l->bd != 10 && l->bd != 12) {
    set_error("b200_compound_launch: bd must be 8, 10 or 12");
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  compound_kernel<<<(l->n_blocks + kWarps - 1) / kWarps, kWarps * 32, 0, s>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}
