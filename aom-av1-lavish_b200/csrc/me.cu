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
  bool staged;               // reference window is in shared memory (word-aligned rows at least)
  int win0_off;              // staged: byte offset inside me_smem of the reference pixel at mv (0,0)
  int cost_fr, cost_fc;      // full-pel reference mv of the sad-domain mv cost
  int l1_mul;                // 32 / 15 / 8 for the L1 cost types, 0 otherwise
  int px;                    // bytes per pixel
  int sad_shift;             // 0 / 2 / 4: encoder_utils.h:153-208
  unsigned *cand;            // shared: kMaxCand totals
  int *bcast;                // shared scratch for broadcasts
  long long *red;            // shared: 2 * kWarps partial sums (CTA-wide variance)
};

// Dynamic shared memory of both kernels: source block at offset 0, then the staged reference
// window.  The hot SAD loops address it through this array (not through the generic
// pointers of Ctx) so that the compiler emits LDS with 32-bit addresses instead of generic
// loads with 64-bit address arithmetic.
extern __shared__ __align__(16) unsigned char me_smem[];
// The search-site table of the CTA's method (copied from global memory at kernel start).
__shared__ Sites g_sites;
__device__ __forceinline__ const unsigned *smem_words(int byte_off) {
  return reinterpret_cast<const unsigned *>(me_smem + byte_off);
}

// ------------------------------------------------------------------- mv cost -----
__device__ __forceinline__ int mv_joint(int r, int c) { return (r != 0) * 2 + (c != 0); }

__device__ __forceinline__ int mv_cost_tab(const Ctx &k, int dr, int dc) {
  return k.l->mvjcost[mv_joint(dr, dc)] + k.l->mvcost[0][dr] + k.l->mvcost[1][dc];
}

// mvsad_err_cost (mcomp.c:329).  The full-pel reference mv (GET_MV_RAWPEL of ref_mv) and the
// L1 weight of the cost type are per-search constants kept in Ctx (init_costs).
__device__ __forceinline__ void init_costs(Ctx &k) {
  const int rr = k.d.ref_mv_row, rc = k.d.ref_mv_col;
  k.cost_fr = (rr + 3 + (rr >= 0)) >> 3;
  k.cost_fc = (rc + 3 + (rc >= 0)) >> 3;
  const int t = k.d.mv_cost_type;
  k.l1_mul = t == B200_MV_COST_L1_LOWRES ? 32 : (t == B200_MV_COST_L1_MIDRES ? 15 : (t == B200_MV_COST_L1_HDRES ? 8 : 0));
}
__device__ __forceinline__ int sad_cost(const Ctx &k, int row, int col) {
  const int dr = (short)((row - k.cost_fr) * 8), dc = (short)((col - k.cost_fc) * 8);
  if (k.d.mv_cost_type == B200_MV_COST_ENTROPY) {
    const unsigned c = (unsigned)mv_cost_tab(k, dr, dc) * (unsigned)k.d.sad_per_bit;
    return (int)((c + 256u) >> 9);
  }
  return (k.l1_mul * (abs(dr) + abs(dc))) >> 3;
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
// acc + sum of |a.b[i] - b.b[i]| over the four bytes: ONE VABSDIFF4.U8.ACC.  (__vsadu4(a, b)
// + acc compiles to the instruction with a zero accumulator plus an IADD3 tree, i.e. 1.5
// ALU-pipe slots per word.)
__device__ __forceinline__ unsigned sad4_acc(unsigned a, unsigned b, unsigned acc) {
  unsigned d;
  asm("vabsdiff4.u32.u32.u32.add %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(acc));
  return d;
}

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
  return __reduce_add_sync(0xffffffffu, v);  // REDUX.SUM: one instruction instead of 5 shuffles
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

// ---- staged 8-bit path, shared-memory addressing ------------------------------------
// Same lane mapping as warp_sad_vec (8 consecutive lanes = 8 consecutive rows of one
// 32-pixel segment column: conflict-free uint4 reads), but every address is a 32-bit offset
// into me_smem and the per-lane offsets are formed once, so that one 32-pixel row segment
// costs 5 LDS.128 + 8 SHF + 8 VABSDIFF4 and two adds.
template <int SEGW, int WO, bool ALIGNED = false>
__device__ __forceinline__ unsigned sad8_lane(int s_off, int r_off, int n_it, int s_step, int r_step,
                                              unsigned sh) {
  constexpr int NV = SEGW / 4;
  unsigned acc = 0, acc2 = 0;  // two chains: the accumulate form is a dependent sequence
#pragma unroll 2
  for (int it = 0; it < n_it; ++it) {
    unsigned sv[SEGW], r[SEGW + 4];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const uint4 t = reinterpret_cast<const uint4 *>(me_smem + s_off)[v];
      sv[4 * v] = t.x, sv[4 * v + 1] = t.y, sv[4 * v + 2] = t.z, sv[4 * v + 3] = t.w;
    }
#pragma unroll
    for (int v = 0; v <= NV; ++v) {
      const uint4 t = reinterpret_cast<const uint4 *>(me_smem + r_off)[v];
      r[4 * v] = t.x, r[4 * v + 1] = t.y, r[4 * v + 2] = t.z, r[4 * v + 3] = t.w;
    }
#pragma unroll
    for (int i = 0; i < SEGW; i += 2) {
      if (ALIGNED) {  // candidate column is a multiple of 4 pixels: no funnel shifts
        acc = sad4_acc(sv[i], r[WO + i], acc);
        acc2 = sad4_acc(sv[i + 1], r[WO + i + 1], acc2);
      } else {
        acc = sad4_acc(sv[i], __funnelshift_r(r[WO + i], r[WO + i + 1], sh), acc);
        acc2 = sad4_acc(sv[i + 1],
                        __funnelshift_r(r[WO + i + 1], r[(WO + i + 2) < SEGW + 4 ? WO + i + 2 : WO + i + 1], sh), acc2);
      }
    }
    s_off += s_step;
    r_off += r_step;
  }
  return acc + acc2;
}

template <int SEGW>
__device__ __forceinline__ unsigned sad8_wo(int wo, int s_off, int r_off, int n_it, int s_step,
                                            int r_step, unsigned sh) {
  switch (wo) {  // warp-uniform
    case 0: return sad8_lane<SEGW, 0>(s_off, r_off, n_it, s_step, r_step, sh);
    case 1: return sad8_lane<SEGW, 1>(s_off, r_off, n_it, s_step, r_step, sh);
    case 2: return sad8_lane<SEGW, 2>(s_off, r_off, n_it, s_step, r_step, sh);
    default: return sad8_lane<SEGW, 3>(s_off, r_off, n_it, s_step, r_step, sh);
  }
}

// cand_off = byte offset in me_smem of the candidate's top-left reference pixel; the source
// block sits at offset 0.  bw in {16, 32, 64, 128}.  Returns the warp total in every lane.
// `part` of `parts`: the lane's row iterations are cut into `parts` contiguous pieces so that
// several warps can share one candidate (the pieces' totals add up to the SAD).
__device__ __noinline__ unsigned sad8_staged(int cand_off, int src_pitch_b, int win_pitch_b, int bw,
                                             int bh, int rmul, int part, int parts) {
  const int lane = threadIdx.x & 31;
  const unsigned sh = (unsigned)(cand_off & 3) * 8;
  const int wo = (cand_off >> 2) & 3, base = cand_off & ~15;
  unsigned acc;
  if (bw >= 32) {
    const int lspr = 31 - __clz(bw >> 5);         // log2(segments per row): 0, 1, 2
    const int q = lane >> 3;
    const int seg0 = q & ((1 << lspr) - 1), yl = ((q >> lspr) << 3) + (lane & 7);
    const int lrpi = 5 - lspr;                    // log2(rows per iteration)
    const int n_it = yl < bh ? ((bh - 1 - yl) >> lrpi) + 1 : 0;
    const int i0 = (n_it * part) / parts, i1 = (n_it * (part + 1)) / parts;
    const int ys = (yl + (i0 << lrpi)) * rmul, rs = rmul << lrpi;
    acc = sad8_wo<8>(wo, ys * src_pitch_b + seg0 * 32, base + ys * win_pitch_b + seg0 * 32, i1 - i0,
                     rs * src_pitch_b, rs * win_pitch_b, sh);
  } else {
    const int n_it = lane < bh ? ((bh - 1 - lane) >> 5) + 1 : 0;
    const int i0 = (n_it * part) / parts, i1 = (n_it * (part + 1)) / parts;
    const int ys = (lane + (i0 << 5)) * rmul, rs = rmul << 5;
    acc = sad8_wo<4>(wo, ys * src_pitch_b, base + ys * win_pitch_b, i1 - i0, rs * src_pitch_b,
                     rs * win_pitch_b, sh);
  }
  return warp_sum_u(acc);
}

// Variance sums on the same layout: |s - r| per byte (VABSDIFF4), then dot products --
// sse += d.d, sum_s += s.1, sum_r += r.1 (DP4A) -- instead of one scalar lane per pixel.
template <int SEGW, int WO>
__device__ __forceinline__ void var8_lane(int s_off, int r_off, int it0, int n_it, int it_step,
                                          int s_step, int r_step, unsigned sh, unsigned &ss,
                                          unsigned &sr, unsigned &sse) {
  constexpr int NV = SEGW / 4;
  for (int it = it0; it < n_it; it += it_step) {
    unsigned sv[SEGW], r[SEGW + 4];
    const int so = s_off + it * s_step, ro = r_off + it * r_step;
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const uint4 t = reinterpret_cast<const uint4 *>(me_smem + so)[v];
      sv[4 * v] = t.x, sv[4 * v + 1] = t.y, sv[4 * v + 2] = t.z, sv[4 * v + 3] = t.w;
    }
#pragma unroll
    for (int v = 0; v <= NV; ++v) {
      const uint4 t = reinterpret_cast<const uint4 *>(me_smem + ro)[v];
      r[4 * v] = t.x, r[4 * v + 1] = t.y, r[4 * v + 2] = t.z, r[4 * v + 3] = t.w;
    }
#pragma unroll
    for (int i = 0; i < SEGW; ++i) {
      const unsigned rv = __funnelshift_r(r[WO + i], r[(WO + i + 1) < SEGW + 4 ? WO + i + 1 : WO + i], sh);
      const unsigned d = __vabsdiffu4(sv[i], rv);
      sse = __dp4a(d, d, sse);
      ss = __dp4a(sv[i], 0x01010101u, ss);
      sr = __dp4a(rv, 0x01010101u, sr);
    }
  }
}

template <int SEGW>
__device__ __forceinline__ void var8_wo(int wo, int s_off, int r_off, int it0, int n_it, int it_step,
                                        int s_step, int r_step, unsigned sh, unsigned &ss,
                                        unsigned &sr, unsigned &sse) {
  switch (wo) {
    case 0: var8_lane<SEGW, 0>(s_off, r_off, it0, n_it, it_step, s_step, r_step, sh, ss, sr, sse); break;
    case 1: var8_lane<SEGW, 1>(s_off, r_off, it0, n_it, it_step, s_step, r_step, sh, ss, sr, sse); break;
    case 2: var8_lane<SEGW, 2>(s_off, r_off, it0, n_it, it_step, s_step, r_step, sh, ss, sr, sse); break;
    default: var8_lane<SEGW, 3>(s_off, r_off, it0, n_it, it_step, s_step, r_step, sh, ss, sr, sse); break;
  }
}

// Per-lane partial sums of (src - ref) and (src - ref)^2 over the whole block, the rows
// dealt out to all `n_warps` warps of the CTA (iteration `it` goes to warp it % n_warps).
__device__ __noinline__ void var8_staged(int cand_off, int src_pitch_b, int win_pitch_b, int bw, int bh,
                                         int n_warps, int *sum_out, unsigned *sse_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const unsigned sh = (unsigned)(cand_off & 3) * 8;
  const int wo = (cand_off >> 2) & 3, base = cand_off & ~15;
  unsigned ss = 0, sr = 0, sse = 0;
  if (bw >= 32) {
    const int lspr = 31 - __clz(bw >> 5);
    const int q = lane >> 3;
    const int seg0 = q & ((1 << lspr) - 1), yl = ((q >> lspr) << 3) + (lane & 7);
    const int lrpi = 5 - lspr;
    const int n_it = yl < bh ? ((bh - 1 - yl) >> lrpi) + 1 : 0;
    var8_wo<8>(wo, yl * src_pitch_b + seg0 * 32, base + yl * win_pitch_b + seg0 * 32, warp, n_it,
               n_warps, src_pitch_b << lrpi, win_pitch_b << lrpi, sh, ss, sr, sse);
  } else {
    const int n_it = lane < bh ? ((bh - 1 - lane) >> 5) + 1 : 0;
    var8_wo<4>(wo, lane * src_pitch_b, base + lane * win_pitch_b, warp, n_it, n_warps,
               src_pitch_b << 5, win_pitch_b << 5, sh, ss, sr, sse);
  }
  *sum_out = (int)ss - (int)sr;
  *sse_out = sse;
}

// Everything else (window not staged, high bit depth, 4- and 8-pixel wide blocks).
__device__ __noinline__ unsigned warp_sad_slow(const Ctx &k, int row, int col, int bh, int rmul) {
  const int lane = threadIdx.x & 31;
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
  return warp_sum_u(acc);
}

// SAD of the source block against the reference block at full-pel mv (row, col), computed
// by ONE warp; every lane returns the total.  skip = rows 0,2,4.. only, doubled.
// warp_sad_part: piece `part` of `parts` of that sum (before doubling / shifting); only the
// staged 8-bit path splits, elsewhere piece 0 is the whole sum and the others are 0.
__device__ __forceinline__ bool sad_splits(const Ctx &k) { return k.fast && k.px == 1 && k.d.bw >= 16; }
__device__ __forceinline__ unsigned warp_sad_part(const Ctx &k, int row, int col, bool skip, int part,
                                                  int parts) {
  const int bh = skip ? k.d.bh >> 1 : k.d.bh, rmul = skip ? 2 : 1;
  if (sad_splits(k))
    return sad8_staged(k.win0_off + row * k.ref_pitch + col, k.src_pitch_w * 4, k.ref_pitch, k.d.bw, bh,
                       rmul, part, parts);
  return part == 0 ? warp_sad_slow(k, row, col, bh, rmul) : 0u;
}
__device__ __forceinline__ unsigned sad_finish(const Ctx &k, unsigned acc, bool skip) {
  if (skip) acc *= 2;
  return acc >> k.sad_shift;
}
__device__ __forceinline__ unsigned warp_sad(const Ctx &k, int row, int col, bool skip, int lane) {
  return sad_finish(k, warp_sad_part(k, row, col, skip, 0, 1), skip);
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
