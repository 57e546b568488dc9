// me.cu -- batched motion search: av1_full_pixel_search (n-step / diamond + exhaustive
// mesh + cost list) and av1_find_best_sub_pixel_tree on the GPU.
//
// Reference semantics restated from av1/encoder/mcomp.c:
//   site tables :369-404,:452-496      mv costs :269-350        diamond_search_sad :1318-1477
//   full_pixel_diamond :1479-1526      exhaustive_mesh_search :1530-1599
//   full_pixel_exhaustive :1603-1679   calc_int_sad_list :791-841
//   av1_full_pixel_search :1755-1895   sub-pel tree :2371-2868, :3128-3196
//
// Mapping: ONE CTA per search (block x reference).  The block's source pixels and the
// whole reachable reference window (mv limits box + block size) are staged in shared
// memory once; every step of the search then evaluates its candidates in parallel,
// one warp per candidate (packed VABSDIFF4 on 4 pixels per lane-iteration), and all
// threads replay the reference's sequential decision rule on the stored totals
// ("first candidate in scan order with the strictly smallest sad + cost", second-best
// = running best just before the winner arrived).  All control flow is CTA-uniform.
#include <limits.h>
#include <string.h>

#include "internal.h"

namespace b200 {
namespace me {

constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kMaxCand = 1024;       // candidate totals kept in shared memory per chunk
constexpr int kMaxSteps = 11;        // MAX_MVSEARCH_STEPS
constexpr int kWindowBudget = 150 * 1024;  // bytes of shared memory for the reference window

struct Mv {
  int row, col;
};

// Search-site tables for the four served methods, built once per launch into global
// memory by the host (see build_sites): site[step][0..16], per_step, radius.
struct Sites {
  short site[2 * kMaxSteps][17][2];
  int n_steps;
  int per_step[2 * kMaxSteps];
  int radius[2 * kMaxSteps];
};

struct Ctx {
  // per-search constants (CTA-uniform)
  const B200MeLaunch *l;
  B200MeBlock d;
  const Sites *sites;
  // pixel views
  const unsigned char *src;  // shared, dense pitch = bw pixels
  const unsigned char *ref0; // reference pixel at mv (0,0): shared window or global plane
  int ref_pitch;             // pixels
  int src_pitch_w;           // source row pitch in 32-bit words (shared copy)
  bool fast;                 // staged window with 16-byte aligned rows: vector-load SAD path
  int px;                    // bytes per pixel
  int sad_shift;             // 0 / 2 / 4: encoder_utils.h:153-208
  unsigned *cand;            // shared: kMaxCand totals
  int *bcast;                // shared scratch for broadcasts
  long long *red;            // shared: 2 * kWarps partial sums (CTA-wide variance)
};

// ------------------------------------------------------------------- mv cost -----
__device__ __forceinline__ int mv_joint(int r, int c) { return (r != 0) * 2 + (c != 0); }

__device__ __forceinline__ int mv_cost_tab(const Ctx &k, int dr, int dc) {
  return k.l->mvjcost[mv_joint(dr, dc)] + k.l->mvcost[0][dr] + k.l->mvcost[1][dc];
}

// mvsad_err_cost (mcomp.c:329)
__device__ __forceinline__ int sad_cost(const Ctx &k, int row, int col) {
  const int rr = k.d.ref_mv_row, rc = k.d.ref_mv_col;
  const int fr = (rr + 3 + (rr >= 0)) >> 3, fc = (rc + 3 + (rc >= 0)) >> 3;  // GET_MV_RAWPEL
  const int dr = (short)((row - fr) * 8), dc = (short)((col - fc) * 8);
  switch (k.d.mv_cost_type) {
    case B200_MV_COST_ENTROPY: {
      const unsigned c = (unsigned)mv_cost_tab(k, dr, dc) * (unsigned)k.d.sad_per_bit;
      return (int)((c + 256u) >> 9);
    }
    case B200_MV_COST_L1_LOWRES: return (32 * (abs(dr) + abs(dc))) >> 3;
    case B200_MV_COST_L1_MIDRES: return (15 * (abs(dr) + abs(dc))) >> 3;
    case B200_MV_COST_L1_HDRES: return (8 * (abs(dr) + abs(dc))) >> 3;
    default: return 0;
  }
}

// mv_err_cost (mcomp.c:290), 1/8-pel mv
__device__ __forceinline__ int err_cost(const Ctx &k, int row, int col) {
  const int dr = (short)(row - k.d.ref_mv_row), dc = (short)(col - k.d.ref_mv_col);
  switch (k.d.mv_cost_type) {
    case B200_MV_COST_ENTROPY:
      if (!k.l->mvcost[0]) return 0;
      return (int)(((long long)mv_cost_tab(k, dr, dc) * k.d.error_per_bit + 8192) >> 14);
    case B200_MV_COST_L1_LOWRES: return (2 * (abs(dr) + abs(dc))) >> 3;
    case B200_MV_COST_L1_HDRES: return (abs(dr) + abs(dc)) >> 3;
    default: return 0;
  }
}

__device__ __forceinline__ bool in_range(const Ctx &k, int row, int col) {
  return col >= k.d.col_min && col <= k.d.col_max && row >= k.d.row_min && row <= k.d.row_max;
}
__device__ __forceinline__ Mv clamp_mv(const Ctx &k, Mv m) {
  m.col = min(max(m.col, (int)k.d.col_min), (int)k.d.col_max);
  m.row = min(max(m.row, (int)k.d.row_min), (int)k.d.row_max);
  return m;
}

// ----------------------------------------------------------------- pixel sums -----
// 4 consecutive bytes at an arbitrary address: two aligned word loads + funnel shift.
__device__ __forceinline__ unsigned load4(const unsigned char *p) {
  const unsigned long long a = reinterpret_cast<unsigned long long>(p);
  const unsigned *w = reinterpret_cast<const unsigned *>(a & ~3ull);
  const unsigned sh = (unsigned)(a & 3) * 8;
  const unsigned lo = w[0];
  if (sh == 0) return lo;
  return __funnelshift_r(lo, w[1], sh);
}

__device__ __forceinline__ unsigned warp_sum_u(unsigned v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ long long warp_sum_ll(long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// One segment of SEGW consecutive 32-bit words of a row: SEGW aligned source words against
// SEGW + 1 aligned reference words funnel-shifted to the candidate's byte alignment.
// PX = 1: four pixels per word (VABSDIFF4 with accumulate); PX = 2: two 16-bit pixels.
template <int SEGW, int PX>
__device__ __forceinline__ unsigned seg_sad(const unsigned *sp, const unsigned char *rbytes) {
  const unsigned long long a = reinterpret_cast<unsigned long long>(rbytes);
  const unsigned *rw = reinterpret_cast<const unsigned *>(a & ~3ull);
  const unsigned sh = (unsigned)(a & 3) * 8;
  unsigned r[SEGW + 1];
#pragma unroll
  for (int i = 0; i <= SEGW; ++i) r[i] = rw[i];
  unsigned acc = 0;
#pragma unroll
  for (int i = 0; i < SEGW; ++i) {
    const unsigned rv = __funnelshift_r(r[i], r[i + 1], sh);
    if (PX == 1) {
      acc = __vsadu4(sp[i], rv) + acc;
    } else {
      const unsigned d = __vabsdiffu2(sp[i], rv);
      acc += (d & 0xffffu) + (d >> 16);
    }
  }
  return acc;
}

template <int SEGW, int PX>
__device__ __forceinline__ unsigned warp_sad_t(const Ctx &k, const unsigned char *rp, int bh,
                                               int rmul, int lane) {
  const int row_words = k.d.bw * PX / 4, spr = row_words / SEGW, nseg = spr * bh;
  const long long pitch_b = (long long)k.ref_pitch * PX;
  unsigned acc = 0;
  for (int s = lane; s < nseg; s += 32) {
    const int y = (s / spr) * rmul, w0 = (s % spr) * SEGW;
    acc += seg_sad<SEGW, PX>(reinterpret_cast<const unsigned *>(k.src) + y * k.src_pitch_w + w0,
                             rp + y * pitch_b + w0 * 4);
  }
  return acc;
}

// ---- staged fast path --------------------------------------------------------------
// Source rows and window rows both have a pitch of 4 (mod 8) words and 16-byte aligned row
// starts, so a quarter-warp reading one uint4 from each of 8 consecutive rows touches all
// 32 banks exactly once.  A lane owns one segment of SEGW words of one row: SEGW/4 vector
// loads of source, SEGW/4 + 1 of reference (16-byte aligned superset of the candidate's
// bytes; WO = warp-uniform word offset of the candidate inside it), then one funnel shift
// + one VABSDIFF4.ACC per 4 pixels.
template <int SEGW, int PX, int WO>
__device__ __forceinline__ unsigned seg_sad_vec(const uint4 *sp, const uint4 *rp, unsigned sh) {
  constexpr int NV = SEGW / 4;
  unsigned sv[SEGW], r[SEGW + 4];
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const uint4 t = sp[v];
    sv[4 * v] = t.x, sv[4 * v + 1] = t.y, sv[4 * v + 2] = t.z, sv[4 * v + 3] = t.w;
  }
#pragma unroll
  for (int v = 0; v <= NV; ++v) {
    const uint4 t = rp[v];
    r[4 * v] = t.x, r[4 * v + 1] = t.y, r[4 * v + 2] = t.z, r[4 * v + 3] = t.w;
  }
  unsigned acc = 0;
#pragma unroll
  for (int i = 0; i < SEGW; ++i) {
    // WO + i + 1 <= SEGW + 3: always inside r[]
    const unsigned rv = __funnelshift_r(r[WO + i], r[(WO + i + 1) < SEGW + 4 ? WO + i + 1 : WO + i], sh);
    if (PX == 1) {
      acc = __vsadu4(sv[i], rv) + acc;
    } else {
      const unsigned d = __vabsdiffu2(sv[i], rv);
      acc += (d & 0xffffu) + (d >> 16);
    }
  }
  return acc;
}

template <int SEGW, int PX, int WO>
__device__ __forceinline__ unsigned warp_sad_vec(const Ctx &k, const unsigned char *rp, int bh,
                                                 int rmul, int lane, unsigned sh) {
  const int row_words = k.d.bw * PX / 4, spr = row_words / SEGW;  // 1, 2, 4 or 8
  const int spp = spr < 4 ? spr : 4;                              // segment columns per pass
  const int rows_per_it = 32 / spp;
  // lane -> (segment, row): 8 consecutive lanes = 8 consecutive rows of one segment column
  const int q = lane >> 3, r8 = lane & 7;
  const int seg0 = q % spp, yl = (q / spp) * 8 + r8;
  const int pitch_w = k.ref_pitch * PX / 4;
  unsigned acc = 0;
  for (int sb = 0; sb < spr; sb += spp) {
    const int seg = sb + seg0;
#pragma unroll 2
    for (int y0 = 0; y0 < bh; y0 += rows_per_it) {
      const int y = y0 + yl;
      if (y < bh) {
        const int yy = y * rmul;
        const uint4 *sp = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned *>(k.src) +
                                                          yy * k.src_pitch_w + seg * SEGW);
        const uint4 *rv = reinterpret_cast<const uint4 *>(reinterpret_cast<const unsigned *>(rp) +
                                                          (long long)yy * pitch_w + seg * SEGW);
        acc += seg_sad_vec<SEGW, PX, WO>(sp, rv, sh);
      }
    }
  }
  return acc;
}

template <int SEGW, int PX>
__device__ __forceinline__ unsigned warp_sad_fast(const Ctx &k, const unsigned char *cand, int bh,
                                                  int rmul, int lane) {
  const unsigned long long a = reinterpret_cast<unsigned long long>(cand);
  const unsigned char *rp = reinterpret_cast<const unsigned char *>(a & ~15ull);
  const unsigned sh = (unsigned)(a & 3) * 8;
  switch ((a >> 2) & 3) {
    case 0: return warp_sad_vec<SEGW, PX, 0>(k, rp, bh, rmul, lane, sh);
    case 1: return warp_sad_vec<SEGW, PX, 1>(k, rp, bh, rmul, lane, sh);
    case 2: return warp_sad_vec<SEGW, PX, 2>(k, rp, bh, rmul, lane, sh);
    default: return warp_sad_vec<SEGW, PX, 3>(k, rp, bh, rmul, lane, sh);
  }
}

// SAD of the source block against the reference block at full-pel mv (row, col), computed
// by ONE warp; every lane returns the total.  skip = rows 0,2,4.. only, doubled.
__device__ __forceinline__ unsigned warp_sad(const Ctx &k, int row, int col, bool skip, int lane) {
  const int bh = skip ? k.d.bh >> 1 : k.d.bh, rmul = skip ? 2 : 1;
  const unsigned char *rp = k.ref0 + ((long long)row * k.ref_pitch + col) * k.px;
  const int row_words = k.d.bw * k.px / 4;
  unsigned acc;
  if (k.fast && row_words >= 4) {
    if (k.px == 1) acc = row_words >= 8 ? warp_sad_fast<8, 1>(k, rp, bh, rmul, lane)
                                        : warp_sad_fast<4, 1>(k, rp, bh, rmul, lane);
    else acc = row_words >= 8 ? warp_sad_fast<8, 2>(k, rp, bh, rmul, lane)
                              : warp_sad_fast<4, 2>(k, rp, bh, rmul, lane);
  } else if (k.px == 1) {
    if (row_words >= 8) acc = warp_sad_t<8, 1>(k, rp, bh, rmul, lane);
    else if (row_words == 4) acc = warp_sad_t<4, 1>(k, rp, bh, rmul, lane);
    else if (row_words == 2) acc = warp_sad_t<2, 1>(k, rp, bh, rmul, lane);
    else acc = warp_sad_t<1, 1>(k, rp, bh, rmul, lane);
  } else {
    if (row_words >= 8) acc = warp_sad_t<8, 2>(k, rp, bh, rmul, lane);
    else if (row_words == 4) acc = warp_sad_t<4, 2>(k, rp, bh, rmul, lane);
    else acc = warp_sad_t<2, 2>(k, rp, bh, rmul, lane);
  }
  acc = warp_sum_u(acc);
  if (skip) acc *= 2;
  return acc >> k.sad_shift;
}

// variance family result of (a - b) sums -> (var, sse) exactly as aom_varianceWxH_c /
// aom_highbd_{8,10,12}_varianceWxH_c (variance.c:123, :361-408)
__device__ __forceinline__ unsigned finish_var(long long sum, unsigned long long sse, int n,
                                               bool highbd, int bd, unsigned *sse_out) {
  unsigned sse32;
  int isum;
  if (highbd && bd > 8) {
    const int s2 = bd - 8, s1 = 2 * s2;
    sse32 = (unsigned)((sse + (1ull << (s1 - 1))) >> s1);
    isum = (int)((sum + (1ll << (s2 - 1))) >> s2);
  } else {
    sse32 = (unsigned)sse;
    isum = (int)sum;
  }
  *sse_out = sse32;
  const long long mean2 = ((long long)isum * isum) / n;
  if (highbd && bd > 8) {
    const long long v = (long long)sse32 - mean2;
    return v >= 0 ? (unsigned)v : 0u;
  }
  return sse32 - (unsigned)mean2;
}

// vf(src, ref @ full-pel mv) by one warp; `ref_first` selects the operand order (the
// rounding of the signed sum at 10/12 bit is not symmetric).
__device__ __forceinline__ unsigned warp_var(const Ctx &k, int row, int col, bool ref_first,
                                             int lane, unsigned *sse_out) {
  const int bw = k.d.bw, bh = k.d.bh, n = bw * bh;
  const unsigned char *rp = k.ref0 + ((long long)row * k.ref_pitch + col) * k.px;
  long long sum = 0;
  unsigned long long sse = 0;
  for (int i = lane; i < n; i += 32) {
    const int y = i / bw, x = i - y * bw;
    int s, r;
    if (k.px == 1) {
      s = k.src[y * k.src_pitch_w * 4 + x];
      r = rp[(long long)y * k.ref_pitch + x];
    } else {
      s = reinterpret_cast<const unsigned short *>(k.src)[y * k.src_pitch_w * 2 + x];
      r = reinterpret_cast<const unsigned short *>(rp)[(long long)y * k.ref_pitch + x];
    }
    const int df = ref_first ? r - s : s - r;
    sum += df;
    sse += (unsigned)(df * df);
  }
  sum = warp_sum_ll(sum);
  sse = (unsigned long long)warp_sum_ll((long long)sse);
  return finish_var(sum, sse, n, k.l->highbd != 0, k.l->bd, sse_out);
}

}  // namespace me
}  // namespace b200

#include "me_fullpel.cuh"
#include "me_subpel.cuh"
#include "me_host.cuh"
