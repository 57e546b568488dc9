// blockops.cu -- stand-alone batched members of the hot path: quantisers, block error,
// SAD / variance / sub-pel variance / SSE, subtract, sub-pel predictor.
//
// One warp per block everywhere: lanes stride over the block's elements, results are
// shuffle-reduced.  These serve the batched tier of include/b200av1.h and, with a batch
// of one, the RTCD drop-in tier (rtcd.cu).  Reference semantics:
//   av1/encoder/av1_quantize.c:36-264,565   aom_dsp/quantize.c:16-316
//   av1/encoder/rdopt.c:635-682             av1/encoder/tx_search.c:1049
//   aom_dsp/sad.c:22-140,240-330            aom_dsp/variance.c:38-145,321-530
//   aom_dsp/subtract.c:20-50                aom_dsp/sse.c:19-53
//   av1/encoder/reconinter_enc.c:424-496,562-640 + aom_dsp/aom_convolve.c:22-110
#include <string.h>

#include "internal.h"

namespace b200 {

constexpr int kWarps = 4;  // warps per CTA

__device__ __forceinline__ int32_t wmul32(int32_t a, int32_t b) {
  return (int32_t)((uint32_t)a * (uint32_t)b);
}
__device__ __forceinline__ int warp_max(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}
__device__ __forceinline__ int warp_min(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const int w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w < v ? w : v;
  }
  return v;
}
template <class T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ int rpo2(int v, int n) { return n ? ((v + ((1 << n) >> 1)) >> n) : v; }

// ------------------------------------------------------------------ quantisers ----
struct QuantArgs {
  B200QuantLaunch l;
};

__global__ void __launch_bounds__(kWarps * 32) quant_kernel(const B200QuantLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  const B200QuantBlock d = l.blocks[bi];
  const B200QRow qr = l.qrows[d.qrow];
  const int n = d.n_coeffs, ls = d.log_scale;
  const int16_t *iscan = l.iscan + d.iscan_offset;
  const uint8_t *qm = l.qm ? l.qm + d.qm_offset : nullptr;
  const uint8_t *iqm = l.iqm ? l.iqm + d.qm_offset : nullptr;
  const int kind = l.kind;
  const bool fp = kind == B200_QUANT_FP || kind == B200_QUANT_FP_HBD || kind == B200_QUANT_LP;
  const bool hbd = kind == B200_QUANT_FP_HBD || kind == B200_QUANT_B_HBD || kind == B200_QUANT_B_ADAPT_HBD;
  const bool adapt = kind == B200_QUANT_B_ADAPT || kind == B200_QUANT_B_ADAPT_HBD;
  int eob = 0;

  if (kind == B200_QUANT_LP) {  // av1_quantize_lp_c: int16, no dead zone, no log_scale
    const int16_t *c16 = static_cast<const int16_t *>(l.coeff) + d.coeff_offset;
    int16_t *q16 = static_cast<int16_t *>(l.qcoeff) + d.out_offset;
    int16_t *dq16 = static_cast<int16_t *>(l.dqcoeff) + d.out_offset;
    for (int rc = lane; rc < n; rc += 32) {
      const int ac = rc != 0;
      const int c = c16[rc], sign = c < 0 ? -1 : 0, a = (c ^ sign) - sign;
      int t = a + qr.round_fp[ac];
      t = t > 32767 ? 32767 : (t < -32768 ? -32768 : t);
      t = (t * qr.quant_fp[ac]) >> 16;
      const int16_t q = (int16_t)((t ^ sign) - sign);
      q16[rc] = q;
      dq16[rc] = (int16_t)(q * qr.dequant[ac]);
      if (t) eob = max(eob, iscan[rc] + 1);
    }
    eob = warp_max(eob);
    if (lane == 0) l.eob[bi] = (uint16_t)eob;
    return;
  }

  const int32_t *coeff = static_cast<const int32_t *>(l.coeff) + d.coeff_offset;
  int32_t *qc = static_cast<int32_t *>(l.qcoeff) + d.out_offset;
  int32_t *dqc = static_cast<int32_t *>(l.dqcoeff) + d.out_offset;
  const int zb[2] = { rpo2(qr.zbin[0], ls), rpo2(qr.zbin[1], ls) };

  // adaptive: cut the trailing run (scan order) that stays inside zbin + dequant*325/128
  int keep = n;
  if (adapt) {
    int last_out = 0;  // 1 + largest scan index of a coefficient outside the widened bin
    for (int rc = lane; rc < n; rc += 32) {
      const int ac = rc != 0;
      const int add = rpo2(qr.dequant[ac] * 325, 7);
      const int32_t cw = wmul32(coeff[rc], 32);
      const bool inside = cw < zb[ac] * 32 + add && cw > -zb[ac] * 32 - add;
      if (!inside) last_out = max(last_out, iscan[rc] + 1);
    }
    keep = warp_max(last_out);
  }

  int first = 0x7fffffff, last = -1;  // scan positions of non-zero outputs (adaptive)
  for (int rc = lane; rc < n; rc += 32) {
    const int ac = rc != 0;
    const int32_t c = coeff[rc], sign = c < 0 ? -1 : 0;
    const int32_t a = (c ^ sign) - sign;
    const int wt = qm ? qm[rc] : 32, iwt = iqm ? iqm[rc] : 32;
    int32_t q = 0, dq = 0;
    if (fp) {
      const int rnd = rpo2(qr.round_fp[ac], ls);
      if (!qm && !iqm) {
        const bool pass = hbd ? ((int32_t)((uint32_t)a << (1 + ls)) >= qr.dequant[ac])
                              : (((int64_t)a << (1 + ls)) >= (int64_t)qr.dequant[ac]);
        if (pass) {
          int64_t t = (int64_t)a + rnd;
          if (!hbd) t = t > 32767 ? 32767 : (t < -32768 ? -32768 : t);
          q = (int32_t)((t * qr.quant_fp[ac]) >> (16 - ls));
          dq = wmul32(q, qr.dequant[ac]) >> ls;
        }
      } else {
        const int dqv = (qr.dequant[ac] * iwt + 16) >> 5;
        if ((int64_t)a * wt >= (qr.dequant[ac] << (5 - (1 + ls)))) {
          int64_t t = (int64_t)a + rnd;
          if (!hbd) t = t > 32767 ? 32767 : (t < -32768 ? -32768 : t);
          q = (int32_t)((t * wt * qr.quant_fp[ac]) >> (16 - ls + 5));
          dq = wmul32(q, dqv) >> ls;
        }
      }
    } else {
      const bool in_range = !adapt || iscan[rc] < keep;
      if (in_range && wmul32(a, wt) >= (zb[ac] << 5)) {
        int64_t t = (int64_t)a + rpo2(qr.round[ac], ls);
        if (!hbd) t = t > 32767 ? 32767 : (t < -32768 ? -32768 : t);
        t *= wt;
        q = (int32_t)(((((t * qr.quant[ac]) >> 16) + t) * qr.quant_shift[ac]) >> (16 - ls + 5));
        const int dqv = (qr.dequant[ac] * iwt + 16) >> 5;
        dq = wmul32(q, dqv) >> ls;
      }
    }
    const int32_t qs = (q ^ sign) - sign, dqs = (dq ^ sign) - sign;
    qc[rc] = qs;
    dqc[rc] = dqs;
    if (qs) {
      const int pos = iscan[rc];
      eob = max(eob, pos + 1);
      first = min(first, pos);
      last = max(last, pos);
    }
  }
  eob = warp_max(eob);
  if (adapt) {
    first = warp_min(first);
    last = warp_max(last);
    if (last >= 0 && first == last) {
      // exactly one coefficient survived: drop it when it is +-1 and inside the wider bin
      __syncwarp();
      for (int rc = lane; rc < n; rc += 32) {
        if (iscan[rc] != last) continue;
        const int32_t qv = qc[rc];
        if (qv == 1 || qv == -1) {
          const int ac = rc != 0;
          const int add2 = rpo2(qr.dequant[ac] * 525, 7);
          const int32_t cw = wmul32(coeff[rc], 32);
          if (cw < zb[ac] * 32 + add2 && cw > -zb[ac] * 32 - add2) {
            qc[rc] = 0;
            dqc[rc] = 0;
            last = -2;
          }
        }
      }
      last = warp_min(last);
      if (last == -2) eob = 0;
    }
  }
  if (lane == 0) l.eob[bi] = (uint16_t)eob;
}

// ----------------------------------------------------------------- block error ----
__global__ void __launch_bounds__(kWarps * 32) block_error_kernel(const B200ErrLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  const B200ErrBlock d = l.blocks[bi];
  int64_t err = 0, ssz = 0;
  if (l.kind == B200_ERR_LP) {
    const int16_t *c = static_cast<const int16_t *>(l.coeff) + d.coeff_offset;
    const int16_t *q = static_cast<const int16_t *>(l.dqcoeff) + d.dqcoeff_offset;
    for (uint32_t i = lane; i < d.n_coeffs; i += 32) {
      const int df = c[i] - q[i];
      err += wmul32(df, df);
    }
  } else {
    const int32_t *c = static_cast<const int32_t *>(l.coeff) + d.coeff_offset;
    const int32_t *q = static_cast<const int32_t *>(l.dqcoeff) + d.dqcoeff_offset;
    for (uint32_t i = lane; i < d.n_coeffs; i += 32) {
      const int32_t cv = c[i];
      const int32_t df = (int32_t)((uint32_t)cv - (uint32_t)q[i]);
      if (l.kind == B200_ERR_LOWBD) {
        err += wmul32(df, df);
        ssz += wmul32(cv, cv);
      } else if (l.kind == B200_ERR_HBD) {
        err += (int64_t)df * df;
        ssz += (int64_t)cv * cv;
      } else {  // QM: weight = qmatrix[scan[i]] (indexing as in the reference)
        const int64_t w = l.qm[d.aux_offset + l.scan[d.aux_offset + i]];
        const int64_t dd = (int64_t)df * w, cc = (int64_t)cv * w;
        err += (dd * dd + 512) >> 10;
        ssz += (cc * cc + 512) >> 10;
      }
    }
  }
  err = warp_sum(err);
  ssz = warp_sum(ssz);
  if (lane == 0) {
    if (l.kind == B200_ERR_HBD) {
      const int sh = 2 * (l.bd - 8);
      const int64_t r = sh > 0 ? ((int64_t)1 << (sh - 1)) : 0;
      err = (err + r) >> sh;
      ssz = (ssz + r) >> sh;
    }
    l.out[2 * (size_t)bi] = err;
    l.out[2 * (size_t)bi + 1] = ssz;
  }
}

// ---------------------------------------------------------- SAD / variance family --
__constant__ uint8_t c_bil[8][2] = { { 128, 0 }, { 112, 16 }, { 96, 32 }, { 80, 48 },
                                     { 64, 64 }, { 48, 80 },  { 32, 96 }, { 16, 112 } };

template <class PIX>
__device__ __forceinline__ void pix_block(const B200PixLaunch &l, const B200PixBlock &d, int lane,
                                          uint32_t &out0, uint32_t &out1) {
  const PIX *a = static_cast<const PIX *>(l.a) + d.a_offset;
  const PIX *b = static_cast<const PIX *>(l.b) + d.b_offset;
  const int w = d.w;
  int h = d.h, as = d.a_stride, bs = d.b_stride;
  const int op = l.op;
  if (op == B200_PIX_SAD_SKIP) {
    h >>= 1;
    as *= 2;
    bs *= 2;
  }
  const int n = w * h;
  uint32_t sad = 0;
  int64_t sum = 0;
  uint64_t sse = 0;
  if (op == B200_PIX_SAD || op == B200_PIX_SAD_SKIP) {
    for (int i = lane; i < n; i += 32) {
      const int y = i / w, x = i - y * w;
      sad += (uint32_t)abs((int)a[y * as + x] - (int)b[y * bs + x]);
    }
  } else if (op == B200_PIX_SAD_AVG) {
    const PIX *sp = static_cast<const PIX *>(l.aux) + d.aux_offset;
    for (int i = lane; i < n; i += 32) {
      const int y = i / w, x = i - y * w;
      const int p = ((int)sp[i] + (int)b[y * bs + x] + 1) >> 1;  // aom_comp_avg_pred
      sad += (uint32_t)abs((int)a[y * as + x] - p);
    }
  } else if (op == B200_PIX_SUBPEL_VAR) {
    const int f0 = c_bil[d.xoff][0], f1 = c_bil[d.xoff][1];
    const int g0 = c_bil[d.yoff][0], g1 = c_bil[d.yoff][1];
    for (int i = lane; i < n; i += 32) {
      const int y = i / w, x = i - y * w;
      const PIX *p = a + y * as + x;
      const int h0 = ((int)p[0] * f0 + (int)p[1] * f1 + 64) >> 7;
      const int h1 = ((int)p[as] * f0 + (int)p[as + 1] * f1 + 64) >> 7;
      const int pv = (PIX)((h0 * g0 + h1 * g1 + 64) >> 7);
      const int df = pv - (int)b[y * bs + x];
      sum += df;
      sse += (uint32_t)(df * df);
    }
  } else {  // VAR, MSE, SSE
    for (int i = lane; i < n; i += 32) {
      const int y = i / w, x = i - y * w;
      const int df = (int)a[y * as + x] - (int)b[y * bs + x];
      sum += df;
      sse += (uint32_t)(df * df);
    }
  }
  sad = warp_sum(sad);
  sum = warp_sum(sum);
  sse = warp_sum(sse);
  if (op == B200_PIX_SAD || op == B200_PIX_SAD_AVG) {
    out0 = sad, out1 = 0;
  } else if (op == B200_PIX_SAD_SKIP) {
    out0 = 2 * sad, out1 = 0;
  } else if (op == B200_PIX_SSE) {
    out0 = (uint32_t)sse, out1 = (uint32_t)(sse >> 32);
  } else {
    uint32_t sse32;
    int isum;
    const bool hb = sizeof(PIX) == 2;
    if (hb && l.bd > 8) {  // highbd_10_variance / highbd_12_variance (variance.c:361-379)
      const int s2 = l.bd - 8, s1 = 2 * s2;
      sse32 = (uint32_t)((sse + ((uint64_t)1 << (s1 - 1))) >> s1);
      isum = (int)((sum + ((int64_t)1 << (s2 - 1))) >> s2);
    } else {
      sse32 = (uint32_t)sse;
      isum = (int)sum;
    }
    out1 = sse32;
    if (op == B200_PIX_MSE) {
      out0 = sse32;
    } else if (op == B200_PIX_VAR_SUM) {
      out0 = (uint32_t)isum;
    } else {
      const int64_t mean2 = ((int64_t)isum * isum) / (d.w * d.h);
      if (hb && l.bd > 8) {
        const int64_t var = (int64_t)sse32 - mean2;
        out0 = var >= 0 ? (uint32_t)var : 0u;
      } else {
        out0 = sse32 - (uint32_t)mean2;
      }
    }
  }
}

__global__ void __launch_bounds__(kWarps * 32) pix_kernel(const B200PixLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  const B200PixBlock d = l.blocks[bi];
  uint32_t o0, o1;
  if (l.highbd) pix_block<uint16_t>(l, d, lane, o0, o1);
  else pix_block<uint8_t>(l, d, lane, o0, o1);
  if (lane == 0) {
    l.out[2 * (size_t)bi] = o0;
    l.out[2 * (size_t)bi + 1] = o1;
  }
}

__global__ void __launch_bounds__(kWarps * 32) subtract_kernel(const B200SubtractLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  const B200PixBlock d = l.blocks[bi];
  const int w = d.w, n = d.w * d.h;
  int16_t *diff = l.diff + d.aux_offset;
  for (int i = lane; i < n; i += 32) {
    const int y = i / w, x = i - y * w;
    int v;
    if (l.highbd)
      v = (int)static_cast<const uint16_t *>(l.a)[d.a_offset + y * d.a_stride + x] -
          (int)static_cast<const uint16_t *>(l.b)[d.b_offset + y * d.b_stride + x];
    else
      v = (int)static_cast<const uint8_t *>(l.a)[d.a_offset + y * d.a_stride + x] -
          (int)static_cast<const uint8_t *>(l.b)[d.b_offset + y * d.b_stride + x];
    diff[i] = (int16_t)v;
  }
}

// Interpolation kernels at 1/8-pel positions: even rows of av1_bilinear_filters,
// av1_sub_pel_filters_4 and av1_sub_pel_filters_8 (av1/common/filter.h:110-135,210-222);
// taps -3..+4 around the sample.
__constant__ int16_t c_taps[3][8][8] = {
  { { 0, 0, 0, 128, 0, 0, 0, 0 }, { 0, 0, 0, 112, 16, 0, 0, 0 }, { 0, 0, 0, 96, 32, 0, 0, 0 },
    { 0, 0, 0, 80, 48, 0, 0, 0 }, { 0, 0, 0, 64, 64, 0, 0, 0 }, { 0, 0, 0, 48, 80, 0, 0, 0 },
    { 0, 0, 0, 32, 96, 0, 0, 0 }, { 0, 0, 0, 16, 112, 0, 0, 0 } },
  { { 0, 0, 0, 128, 0, 0, 0, 0 }, { 0, 0, -8, 122, 18, -4, 0, 0 }, { 0, 0, -12, 110, 38, -8, 0, 0 },
    { 0, 0, -14, 94, 58, -10, 0, 0 }, { 0, 0, -12, 76, 76, -12, 0, 0 }, { 0, 0, -10, 58, 94, -14, 0, 0 },
    { 0, 0, -8, 38, 110, -12, 0, 0 }, { 0, 0, -4, 18, 122, -8, 0, 0 } },
  { { 0, 0, 0, 128, 0, 0, 0, 0 }, { 0, 2, -10, 122, 18, -4, 0, 0 }, { 0, 2, -14, 110, 38, -10, 2, 0 },
    { 0, 2, -16, 94, 58, -12, 2, 0 }, { 0, 2, -14, 76, 76, -14, 2, 0 }, { 0, 2, -12, 58, 94, -16, 2, 0 },
    { 0, 2, -10, 38, 110, -14, 2, 0 }, { 0, 0, -4, 18, 122, -10, 2, 0 } },
};

template <class PIX>
__device__ __forceinline__ int hfilt(const PIX *p, const int16_t *k) {
  int s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += (int)p[i - 3] * k[i];
  return s;
}

template <class PIX>
__device__ __forceinline__ void upsample_block(const B200UpsampleLaunch &l, const B200PixBlock &d,
                                               int lane) {
  const PIX *ref = static_cast<const PIX *>(l.ref) + d.a_offset;
  PIX *out = static_cast<PIX *>(l.comp_pred) + d.aux_offset;
  const int w = d.w, h = d.h, rs = d.a_stride, n = w * h;
  const int set = l.taps == 2 ? 0 : (l.taps == 4 ? 1 : 2);
  const int16_t *kx = c_taps[set][d.xoff], *ky = c_taps[set][d.yoff];
  const int maxv = sizeof(PIX) == 1 ? 255 : (1 << l.bd) - 1;
  for (int i = lane; i < n; i += 32) {
    const int y = i / w, x = i - y * w;
    const PIX *p = ref + y * rs + x;
    int v;
    if (!d.xoff && !d.yoff) {
      v = p[0];
    } else if (!d.yoff) {
      v = min(max((hfilt(p, kx) + 64) >> 7, 0), maxv);
    } else if (!d.xoff) {
      int s = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) s += (int)p[(k - 3) * rs] * ky[k];
      v = min(max((s + 64) >> 7, 0), maxv);
    } else {
      // horizontal pass clipped to the pixel range before the vertical pass, exactly
      // as the reference's pixel-typed intermediate buffer does
      int s = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int hv = min(max((hfilt(p + (k - 3) * rs, kx) + 64) >> 7, 0), maxv);
        s += hv * ky[k];
      }
      v = min(max((s + 64) >> 7, 0), maxv);
    }
    out[i] = (PIX)v;
  }
}

__global__ void __launch_bounds__(kWarps * 32) upsample_kernel(const B200UpsampleLaunch l) {
  const int lane = threadIdx.x & 31;
  const int bi = blockIdx.x * kWarps + (threadIdx.x >> 5);
  if (bi >= l.n_blocks) return;
  const B200PixBlock d = l.blocks[bi];
  if (l.highbd) upsample_block<uint16_t>(l, d, lane);
  else upsample_block<uint8_t>(l, d, lane);
}

// av1_round_shift_array_c (av1/common/av1_txfm.c:71-87)
__global__ void round_shift_array_kernel(int32_t *arr, int size, int bit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= size) return;
  if (bit > 0) {
    arr[i] = (int32_t)(((int64_t)arr[i] + ((int64_t)1 << (bit - 1))) >> bit);
  } else {
    const int64_t v = (int64_t)arr[i] * ((int64_t)1 << (-bit));
    arr[i] = (int32_t)(v > INT32_MAX ? INT32_MAX : (v < INT32_MIN ? INT32_MIN : v));
  }
}
int round_shift_array_launch(b200_ctx *, int32_t *arr, int size, int bit, cudaStream_t s) {
  if (size <= 0 || bit == 0) return B200_OK;
  round_shift_array_kernel<<<(size + 255) / 256, 256, 0, s>>>(arr, size, bit);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

int blockops_init_device(b200_ctx *) { return B200_OK; }

static int grid_for(int n_blocks) { return (n_blocks + kWarps - 1) / kWarps; }

}  // namespace b200

using namespace b200;

#define B200_LAUNCH_PROLOGUE(L)                                                    \
  if (!ctx || !(L)) return B200_EINVAL;                                            \
  if ((L)->n_blocks < 0) return B200_EINVAL;                                       \
  if ((L)->n_blocks == 0) return B200_OK;                                          \
  B200_CUDA_TRY(cudaSetDevice(ctx->device));                                       \
  cudaStream_t s_ = stream ? (cudaStream_t)stream : ctx->stream;

extern "C" int b200_quant_launch(b200_ctx *ctx, const B200QuantLaunch *l, void *stream) {
  B200_LAUNCH_PROLOGUE(l)
  if (!l->blocks || !l->coeff || !l->qrows || !l->iscan || !l->qcoeff || !l->dqcoeff || !l->eob ||
      l->kind < 0 || l->kind > B200_QUANT_B_ADAPT_HBD) {
    set_error("b200_quant_launch: missing buffer or bad kind");
    return B200_EINVAL;
  }
  quant_kernel<<<grid_for(l->n_blocks), kWarps * 32, 0, s_>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_block_error_launch(b200_ctx *ctx, const B200ErrLaunch *l, void *stream) {
  B200_LAUNCH_PROLOGUE(l)
  if (!l->blocks || !l->coeff || !l->dqcoeff || !l->out || l->kind < 0 || l->kind > B200_ERR_QM ||
      (l->kind == B200_ERR_QM && (!l->qm || !l->scan))) {
    set_error("b200_block_error_launch: missing buffer or bad kind");
    return B200_EINVAL;
  }
  if (l->kind == B200_ERR_HBD && l->bd != 8 && l->bd != 10 && l->bd != 12) {
    set_error("b200_block_error_launch: bd must be 8, 10 or 12 (got %d)", l->bd);
    return B200_EINVAL;
  }
  block_error_kernel<<<grid_for(l->n_blocks), kWarps * 32, 0, s_>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_pix_launch(b200_ctx *ctx, const B200PixLaunch *l, void *stream) {
  B200_LAUNCH_PROLOGUE(l)
  if (!l->blocks || !l->a || !l->b || !l->out || l->op < 0 || l->op > B200_PIX_VAR_SUM ||
      (l->op == B200_PIX_SAD_AVG && !l->aux)) {
    set_error("b200_pix_launch: missing buffer or bad op");
    return B200_EINVAL;
  }
  pix_kernel<<<grid_for(l->n_blocks), kWarps * 32, 0, s_>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_subtract_launch(b200_ctx *ctx, const B200SubtractLaunch *l, void *stream) {
  B200_LAUNCH_PROLOGUE(l)
  if (!l->blocks || !l->a || !l->b || !l->diff) return B200_EINVAL;
  subtract_kernel<<<grid_for(l->n_blocks), kWarps * 32, 0, s_>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_upsampled_pred_launch(b200_ctx *ctx, const B200UpsampleLaunch *l, void *stream) {
  B200_LAUNCH_PROLOGUE(l)
  if (!l->blocks || !l->ref || !l->comp_pred || (l->taps != 2 && l->taps != 4 && l->taps != 8))
    return B200_EINVAL;
  upsample_kernel<<<grid_for(l->n_blocks), kWarps * 32, 0, s_>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}
