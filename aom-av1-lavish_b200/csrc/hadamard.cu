// hadamard.cu -- batched Hadamard / SATD (aom_dsp/avg.c:102-524), lossless WHT
// (av1/encoder/hybrid_fwd_txfm.c:24) and the small block statistics of avg.c (averages,
// min/max, integer projections, vector variance).
//
// Mapping.  8x8 and larger: ONE warp per block (four 8x8 blocks share a warp).  Lane
// (s = lane >> 3, c = lane & 7) owns column c of 8x8 sub-block s (4 tasks per lane at 32x32):
// pass 1 reads its 8 samples straight from global memory (8 independent loads; the 8 lanes of
// a sub-block cover 16 contiguous bytes per row, neighbouring sub-blocks complete the
// sectors), pass 2 reads the transposed intermediate through a pitch-9 shared tile
// (conflict-free, `__syncwarp` only), and everything after that stays in registers: the
// 16x16 stage is two butterfly rounds over lane ^ 8 / lane ^ 16, the 32x32 stage is plain
// arithmetic on the lane's four tasks, the AVX2-order column swap of aom_hadamard_16x16_c is
// a lane ^ 4 exchange, the SATD is a shuffle reduction.  The reference keeps every
// intermediate of hadamard_col8 in an int16_t; wrapping commutes with the add/sub network, so
// the kernel narrows once per pass boundary (see pass 1); the highbd second pass and the
// combination stages stay 32-bit like the reference's.  4x4 (Hadamard and WHT): a lane owns a block.
// Bound: HBM (2 B/sample in, 4 B/sample out); measured 3.0-4.2 TB/s with coefficient output.
#include <string.h>

#include <type_traits>

#include "internal.h"

namespace b200 {
namespace had {

constexpr int kWarps = 4;
constexpr int kThreads = kWarps * 32;

__device__ __forceinline__ int w16(int v) { return (int)(short)v; }

// hadamard_col8 (avg.c:147) with int16 intermediates (I16 = true) or the 32-bit network of
// hadamard_highbd_col8_second_pass (avg.c:383)
template <bool I16>
__device__ __forceinline__ void net8(const int (&x)[8], int (&o)[8]) {
#define B200_W(v) (I16 ? w16(v) : (v))
  const int b0 = B200_W(x[0] + x[1]), b1 = B200_W(x[0] - x[1]), b2 = B200_W(x[2] + x[3]), b3 = B200_W(x[2] - x[3]);
  const int b4 = B200_W(x[4] + x[5]), b5 = B200_W(x[4] - x[5]), b6 = B200_W(x[6] + x[7]), b7 = B200_W(x[6] - x[7]);
  const int c0 = B200_W(b0 + b2), c1 = B200_W(b1 + b3), c2 = B200_W(b0 - b2), c3 = B200_W(b1 - b3);
  const int c4 = B200_W(b4 + b6), c5 = B200_W(b5 + b7), c6 = B200_W(b4 - b6), c7 = B200_W(b5 - b7);
  o[0] = B200_W(c0 + c4); o[7] = B200_W(c1 + c5); o[3] = B200_W(c2 + c6); o[4] = B200_W(c3 + c7);
  o[2] = B200_W(c0 - c4); o[6] = B200_W(c1 - c5); o[1] = B200_W(c2 - c6); o[5] = B200_W(c3 - c7);
#undef B200_W
}

// N in {8, 16, 32}; KIND = B200_HAD_LOWBD / _LP / _HIGHBD.  Blocks per warp: 4 at 8x8 (so
// that all 32 lanes carry a column task), 1 otherwise.
template <int N, int KIND>
__global__ void __launch_bounds__(kThreads) hadamard_kernel(const B200HadamardLaunch l) {
  constexpr int kPerWarp = N == 8 ? 4 : 1;           // blocks per warp
  constexpr int kElems = kPerWarp * N * N;           // coefficients per warp tile
  constexpr int kSub = kElems / 64;                  // 8x8 sub-blocks per warp tile
  constexpr int kTasks = kSub * 8 / 32;              // column tasks per lane (1 or 4)
  constexpr int kSubPerRow = N / 8;
  constexpr int kTmpPitch = 72;                       // 8x8 intermediate at row pitch 9 (+8): conflict-free
  __shared__ int s_tmp[kWarps][kSub * kTmpPitch];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int *tmp = s_tmp[warp];
  const int tiles = (l.n_blocks + kPerWarp - 1) / kPerWarp;
  for (int tile = blockIdx.x * kWarps + warp; tile < tiles; tile += gridDim.x * kWarps) {
    const int blk0 = tile * kPerWarp;
    __syncwarp();
    // first pass: 8-point transform down each column of each 8x8 sub-block (int16 stores),
    // read straight from global memory: the 8 lanes of a sub-block read 16 contiguous bytes per
    // row and horizontally adjacent sub-blocks complete the 32-byte sectors in the same
    // request; all 8 loads of a task are independent (the kernel is latency-bound).
#pragma unroll
    for (int t = 0; t < kTasks; ++t) {
      const int task = t * 32 + lane, sub = task >> 3, c = task & 7;
      const int b = sub / (kSubPerRow * kSubPerRow), sb = sub - b * (kSubPerRow * kSubPerRow);
      // sub-block order of the reference's recursion: quadrant q of a 32x32, then 8x8 i inside it
      const int q = sb >> 2, qi = sb & 3;
      const int sy = ((q >> 1) << 1) + (qi >> 1), sx = ((q & 1) << 1) + (qi & 1);
      int x[8], o[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) x[r] = 0;
      if (blk0 + b < l.n_blocks) {
        const B200HadBlock d = l.blocks[blk0 + b];
        const short *p = l.src + (size_t)d.src_offset + (size_t)(sy * 8) * d.stride + sx * 8 + c;
#pragma unroll
        for (int r = 0; r < 8; ++r) x[r] = p[(size_t)r * d.stride];
      }
      // The reference narrows every intermediate to int16.  Only additions and subtractions
      // are involved up to the end of the second pass, and wrapping is a ring homomorphism
      // (mod 2^16), so narrowing ONCE where the value is next used as a number gives the
      // same bits: low bit depth after pass 2, highbd (32-bit second pass) after pass 1.
      net8<false>(x, o);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        tmp[sub * kTmpPitch + c * 9 + k] = KIND == B200_HAD_HIGHBD ? w16(o[k]) : o[k];
    }
    __syncwarp();
    // second pass down the columns of the intermediate (tmp[c * 8 + k] of the reference is
    // tmp[c * 9 + k] here).  Everything after it stays in registers: lane (s = lane >> 3, c)
    // holds, for task t and k = 0..7, the coefficient of 8x8 sub-block t * 4 + s at storage
    // position k * 8 + c (low bit depth: transposed, avg.c:198-203) or 8 * c + k (highbd).
    int o[kTasks][8];
#pragma unroll
    for (int t = 0; t < kTasks; ++t) {
      const int sub = (t * 32 + lane) >> 3, c = lane & 7;
      int x[8];
#pragma unroll
      for (int r = 0; r < 8; ++r) x[r] = tmp[sub * kTmpPitch + r * 9 + c];
      net8<false>(x, o[t]);
      if (KIND != B200_HAD_HIGHBD) {
#pragma unroll
        for (int k = 0; k < 8; ++k) o[t][k] = w16(o[t][k]);
      }
    }
    const int s4 = lane >> 3, c = lane & 7;
    if constexpr (N >= 16) {
      // 16x16 stage (avg.c:258-276, :298-315, :455-472): the four 8x8 transforms of a 16x16
      // sit in the lanes s4 = 0..3 of the same (t, k, c) -> two butterfly rounds over
      // lane ^ 8 and lane ^ 16
#pragma unroll
      for (int t = 0; t < kTasks; ++t)
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          int v = o[t][k];
          int p1 = __shfl_xor_sync(0xffffffffu, v, 8);
          v = (s4 & 1) ? (p1 - v) >> 1 : (v + p1) >> 1;
          if (KIND == B200_HAD_LP) v = w16(v);
          p1 = __shfl_xor_sync(0xffffffffu, v, 16);
          v = (s4 & 2) ? p1 - v : v + p1;
          if (KIND == B200_HAD_LP) v = w16(v);
          o[t][k] = v;
        }
    }
    if constexpr (N == 32) {
      // 32x32 stage (avg.c:329-347, :485-503): the four 16x16 quadrants are the four tasks of
      // the same lane -> plain register arithmetic (>> 2)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int a0 = o[0][k], a1 = o[1][k], a2 = o[2][k], a3 = o[3][k];
        const int b0 = (a0 + a1) >> 2, b1 = (a0 - a1) >> 2, b2 = (a2 + a3) >> 2, b3 = (a2 - a3) >> 2;
        o[0][k] = b0 + b2;
        o[1][k] = b1 + b3;
        o[2][k] = b0 - b2;
        o[3][k] = b1 - b3;
      }
    }
    if constexpr (KIND == B200_HAD_LOWBD && N >= 16) {
      // the column swap that matches the AVX2 output order (avg.c:279-286): in a row of 16,
      // columns 4..7 <-> 8..11, i.e. (k even, c >= 4) <-> (k odd, c < 4) of the same 8x8.
      // Done as one lane ^ 4 exchange per row pair so that the stores stay 32-byte runs.
#pragma unroll
      for (int t = 0; t < kTasks; ++t)
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const int send = c >= 4 ? o[t][2 * m] : o[t][2 * m + 1];
          const int recv = __shfl_xor_sync(0xffffffffu, send, 4);
          if (c >= 4) o[t][2 * m] = recv;
          else o[t][2 * m + 1] = recv;
        }
    }
    // outputs.  8x8: lane group s4 is block blk0 + s4; larger: one block per warp, 8x8
    // sub-block t * 4 + s4 at offset (t * 4 + s4) * 64.
    const int blk = N == 8 ? blk0 + s4 : blk0;
    const bool live = blk < l.n_blocks;
    int sacc = 0;
#pragma unroll
    for (int t = 0; t < kTasks; ++t)
#pragma unroll
      for (int k = 0; k < 8; ++k) sacc += abs(o[t][k]);
    if (l.coeff && live) {
      const size_t ob = l.blocks[blk].out_offset;
#pragma unroll
      for (int t = 0; t < kTasks; ++t) {
        const int base = N == 8 ? 0 : (t * 4 + s4) * 64;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          int pos;
          if (KIND == B200_HAD_HIGHBD) {
            pos = 8 * c + k;
          } else {
            pos = k * 8 + c;
          }
          if (KIND == B200_HAD_LP) static_cast<short *>(l.coeff)[ob + base + pos] = (short)o[t][k];
          else static_cast<int *>(l.coeff)[ob + base + pos] = o[t][k];
        }
      }
    }
    if (l.satd) {
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 4);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 2);
      sacc += __shfl_xor_sync(0xffffffffu, sacc, 1);
      if (N > 8) {
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 8);
        sacc += __shfl_xor_sync(0xffffffffu, sacc, 16);
      }
      if ((N == 8 ? c == 0 : lane == 0) && live) l.satd[blk] = sacc;
    }
  }
}

// 4x4: a lane owns a block.  WHT = false: aom_hadamard_4x4_c (avg.c:102-145); true:
// av1_fwht4x4_c (hybrid_fwd_txfm.c:24-77).
template <bool WHT>
__global__ void __launch_bounds__(kThreads) had4_kernel(const B200HadamardLaunch l) {
  for (int bi = blockIdx.x * kThreads + threadIdx.x; bi < l.n_blocks; bi += gridDim.x * kThreads) {
    const B200HadBlock d = l.blocks[bi];
    const short *p = l.src + d.src_offset;
    int in[4][4], out[16];
#pragma unroll
    for (int y = 0; y < 4; ++y)
#pragma unroll
      for (int x = 0; x < 4; ++x) in[y][x] = p[(size_t)y * d.stride + x];
    if (!WHT) {
      int t[16], u[16];
#pragma unroll
      for (int c = 0; c < 4; ++c) {  // hadamard_col4 down column c -> t[4c..]
        const int b0 = w16((in[0][c] + in[1][c]) >> 1), b1 = w16((in[0][c] - in[1][c]) >> 1);
        const int b2 = w16((in[2][c] + in[3][c]) >> 1), b3 = w16((in[2][c] - in[3][c]) >> 1);
        t[4 * c] = w16(b0 + b2); t[4 * c + 1] = w16(b1 + b3);
        t[4 * c + 2] = w16(b0 - b2); t[4 * c + 3] = w16(b1 - b3);
      }
#pragma unroll
      for (int c = 0; c < 4; ++c) {  // second pass reads t with stride 4
        const int b0 = w16((t[c] + t[4 + c]) >> 1), b1 = w16((t[c] - t[4 + c]) >> 1);
        const int b2 = w16((t[8 + c] + t[12 + c]) >> 1), b3 = w16((t[8 + c] - t[12 + c]) >> 1);
        u[4 * c] = w16(b0 + b2); u[4 * c + 1] = w16(b1 + b3);
        u[4 * c + 2] = w16(b0 - b2); u[4 * c + 3] = w16(b1 - b3);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) out[i * 4 + j] = u[j * 4 + i];
    } else {
      int m[16];
#pragma unroll
      for (int i = 0; i < 4; ++i) {  // pass 0: down column i, written to row i
        int a1 = in[0][i], b1 = in[1][i], c1 = in[2][i], d1 = in[3][i];
        a1 += b1;
        d1 = d1 - c1;
        const int e1 = (a1 - d1) >> 1;
        b1 = e1 - b1;
        c1 = e1 - c1;
        a1 -= c1;
        d1 += b1;
        m[4 * i] = a1; m[4 * i + 1] = c1; m[4 * i + 2] = d1; m[4 * i + 3] = b1;
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {  // pass 1: down column i of m, scaled by UNIT_QUANT_FACTOR
        int a1 = m[i], b1 = m[4 + i], c1 = m[8 + i], d1 = m[12 + i];
        a1 += b1;
        d1 -= c1;
        const int e1 = (a1 - d1) >> 1;
        b1 = e1 - b1;
        c1 = e1 - c1;
        a1 -= c1;
        d1 += b1;
        out[i] = a1 * 4; out[4 + i] = c1 * 4; out[8 + i] = d1 * 4; out[12 + i] = b1 * 4;
      }
    }
    int acc = 0;
#pragma unroll
    for (int i = 0; i < 16; ++i) acc += abs(out[i]);
    if (l.coeff) {
      int *o = static_cast<int *>(l.coeff) + d.out_offset;
      if ((reinterpret_cast<size_t>(o) & 15) == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          reinterpret_cast<int4 *>(o)[i] = make_int4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i] = out[i];
      }
    }
    if (l.satd) l.satd[bi] = acc;
  }
}

// averages / min-max: a lane owns a block (16 or 64 pixels)
template <bool HBD>
__global__ void __launch_bounds__(kThreads) avg_kernel(const B200AvgLaunch l) {
  typedef typename std::conditional<HBD, unsigned short, unsigned char>::type Px;
  for (int bi = blockIdx.x * kThreads + threadIdx.x; bi < l.n_blocks; bi += gridDim.x * kThreads) {
    const B200AvgBlock d = l.blocks[bi];
    const Px *a = static_cast<const Px *>(l.a) + d.a_offset;
    int o0, o1 = 0;
    if (l.op == B200_MINMAX_8X8) {
      const Px *b = static_cast<const Px *>(l.b) + d.b_offset;
      int mn = 255, mx = 0;
      for (int y = 0; y < 8; ++y)
        for (int x = 0; x < 8; ++x) {
          const int df = abs((int)a[(size_t)y * d.a_stride + x] - (int)b[(size_t)y * d.b_stride + x]);
          mn = min(mn, df);
          mx = max(mx, df);
        }
      o0 = mn;
      o1 = mx;
    } else {
      const int n = l.op == B200_AVG_4X4 ? 4 : 8;
      int sum = 0;
      for (int y = 0; y < n; ++y)
        for (int x = 0; x < n; ++x) sum += a[(size_t)y * d.a_stride + x];
      o0 = (sum + n * n / 2) >> (n == 4 ? 4 : 6);
    }
    l.out[2 * bi] = o0;
    l.out[2 * bi + 1] = o1;
  }
}

// aom_int_pro_row_c (avg.c:528) / aom_int_pro_col_c (:542): one thread per output, int16
// accumulator like the reference.  aom_vector_var_c (:556): one warp.
__global__ void int_pro_kernel(short *out, const unsigned char *ref, int ref_stride, int width,
                               int height, int norm, int col) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (col ? height : width)) return;
  int acc = 0;
  if (col) {
    for (int x = 0; x < width; ++x) acc = w16(acc + ref[(size_t)i * ref_stride + x]);
  } else {
    for (int y = 0; y < height; ++y) acc = w16(acc + ref[(size_t)y * ref_stride + i]);
  }
  out[i] = (short)(acc >> norm);
}
__global__ void vector_var_kernel(const short *ref, const short *src, int bwl, int *out) {
  const int width = 4 << bwl, lane = threadIdx.x;
  int sse = 0, mean = 0;
  for (int i = lane; i < width; i += 32) {
    const int d = ref[i] - src[i];
    mean += d;
    sse += d * d;
  }
  for (int o = 16; o; o >>= 1) {
    sse += __shfl_xor_sync(0xffffffffu, sse, o);
    mean += __shfl_xor_sync(0xffffffffu, mean, o);
  }
  const unsigned m = (unsigned)abs(mean);
  if (lane == 0) *out = sse - (int)((m * m) >> (bwl + 2));
}

// aom_satd_c / aom_satd_lp_c (avg.c:509,518) of one coefficient array: one CTA
__global__ void __launch_bounds__(kThreads) satd_kernel(const void *coeff, int lp, int length, int *out) {
  __shared__ int s_part[kWarps];
  int acc = 0;
  for (int i = threadIdx.x; i < length; i += kThreads)
    acc += abs(lp ? (int)static_cast<const short *>(coeff)[i] : static_cast<const int *>(coeff)[i]);
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < kWarps; ++w) t += s_part[w];
    *out = t;
  }
}

int grid_for(const b200_ctx *ctx, long long units, int per_cta) {
  long long g = (units + per_cta - 1) / per_cta;
  const long long cap = (long long)ctx->sm_count * 16;  // multiple of the SM count; grid-stride beyond
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

}  // namespace had

int int_pro_launch(b200_ctx *ctx, short *out, const unsigned char *ref, int ref_stride, int width,
                   int height, int norm, int col, cudaStream_t s) {
  const int n = col ? height : width;
  had::int_pro_kernel<<<(n + 127) / 128, 128, 0, s>>>(out, ref, ref_stride, width, height, norm, col);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}
int satd_launch(b200_ctx *ctx, const void *coeff, int lp, int length, int *out, cudaStream_t s) {
  had::satd_kernel<<<1, had::kThreads, 0, s>>>(coeff, lp, length, out);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}
int vector_var_launch(b200_ctx *ctx, const short *ref, const short *src, int bwl, int *out, cudaStream_t s) {
  had::vector_var_kernel<<<1, 32, 0, s>>>(ref, src, bwl, out);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_hadamard_launch(b200_ctx *ctx, const B200HadamardLaunch *l, void *stream) {
  using namespace b200::had;
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (l->n_blocks == 0) return B200_OK;
  if (!l->blocks || !l->src || (!l->coeff && !l->satd)) {
    b200::set_error("b200_hadamard_launch: missing buffer (blocks, src, and coeff or satd)");
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  const int n = l->n, k = l->kind;
  bool ok = true;
  if (n == 4 && (k == B200_HAD_LOWBD || k == B200_HAD_WHT)) {
    const int g = grid_for(ctx, l->n_blocks, kThreads);
    if (k == B200_HAD_WHT) had4_kernel<true><<<g, kThreads, 0, s>>>(*l);
    else had4_kernel<false><<<g, kThreads, 0, s>>>(*l);
  } else if (n == 8 && k <= B200_HAD_HIGHBD) {
    const int g = grid_for(ctx, (l->n_blocks + 3) / 4, kWarps);
    if (k == B200_HAD_LOWBD) hadamard_kernel<8, B200_HAD_LOWBD><<<g, kThreads, 0, s>>>(*l);
    else if (k == B200_HAD_LP) hadamard_kernel<8, B200_HAD_LP><<<g, kThreads, 0, s>>>(*l);
    else hadamard_kernel<8, B200_HAD_HIGHBD><<<g, kThreads, 0, s>>>(*l);
  } else if (n == 16 && k <= B200_HAD_HIGHBD) {
    const int g = grid_for(ctx, l->n_blocks, kWarps);
    if (k == B200_HAD_LOWBD) hadamard_kernel<16, B200_HAD_LOWBD><<<g, kThreads, 0, s>>>(*l);
    else if (k == B200_HAD_LP) hadamard_kernel<16, B200_HAD_LP><<<g, kThreads, 0, s>>>(*l);
    else hadamard_kernel<16, B200_HAD_HIGHBD><<<g, kThreads, 0, s>>>(*l);
  } else if (n == 32 && (k == B200_HAD_LOWBD || k == B200_HAD_HIGHBD)) {
    const int g = grid_for(ctx, l->n_blocks, kWarps);
    if (k == B200_HAD_LOWBD) hadamard_kernel<32, B200_HAD_LOWBD><<<g, kThreads, 0, s>>>(*l);
    else hadamard_kernel<32, B200_HAD_HIGHBD><<<g, kThreads, 0, s>>>(*l);
  } else {
    ok = false;
  }
  if (!ok) {
    b200::set_error("b200_hadamard_launch: no %dx%d transform of kind %d in aom_dsp/avg.c", n, n, k);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_avg_launch(b200_ctx *ctx, const B200AvgLaunch *l, void *stream) {
  using namespace b200::had;
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (l->n_blocks == 0) return B200_OK;
  if (!l->blocks || !l->a || !l->out || l->op < B200_AVG_4X4 || l->op > B200_MINMAX_8X8 ||
      (l->op == B200_MINMAX_8X8 && !l->b)) {
    b200::set_error("b200_avg_launch: missing buffer or unknown op %d", l->op);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  const int g = grid_for(ctx, l->n_blocks, kThreads);
  if (l->highbd) avg_kernel<true><<<g, kThreads, 0, s>>>(*l);
  else avg_kernel<false><<<g, kThreads, 0, s>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}
