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

#include <cuda.h>

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
  const B200MeBlock &d;  // the search descriptor, copied to shared memory at kernel start (a
                         // by-value copy lands in local memory: its byte fields defeat SROA)
  const Sites *sites;
  // pixel views
  const unsigned char *src;  // shared, dense pitch = bw pixels
  const unsigned char *ref0; // reference pixel at mv (0,0): shared window or global plane
  int ref_pitch;             // pixels
  int src_pitch_w;           // source row pitch in 32-bit words (shared copy)
  bool fast;                 // staged window with 16-byte aligned rows: vector-load SAD path
  bool staged;               // reference window is in shared memory (word-aligned rows at least)
  int win0_off;              // staged: byte offset inside me_smem of the reference pixel at mv (0,0)
  bool compound;             // compound search (B200MeLaunch.compound): its operands live in g_cmp
  int win_region_off;        // byte offset inside me_smem where the window region starts (after the source)
  int win_budget;            // bytes of that region (0: none)
  int cost_fr, cost_fc;      // full-pel reference mv of the sad-domain mv cost
  int l1_mul;                // 32 / 15 / 8 for the L1 cost types, 0 otherwise
  int px;                    // bytes per pixel
  int sad_shift;             // 0 / 2 / 4: encoder_utils.h:153-208
  int col_min, col_max, row_min, row_max;  // register copies of d's mv limits (tested per candidate)
  __device__ explicit Ctx(const B200MeBlock &desc) : d(desc) {}
};
__shared__ B200MeBlock g_desc;
// CTA scratch, file scope so that every access is an LDS / STS / ATOMS (through a pointer in
// Ctx the compiler lost the address space: generic loads and ATOM.E on shared memory)
__shared__ unsigned g_cand[kMaxCand + 64];  // candidate totals of a step / a mesh chunk
__shared__ int g_bcast[32];                 // broadcast slots
__shared__ long long g_red[2 * kWarps];     // partial sums of the CTA-wide variance
constexpr int kMaxRuns = 16;                // runs of full_pixel_diamond evaluated side by side
__shared__ int g_cda[kMaxRuns * 16];        // n-step: window byte address of the candidates, or -1
__shared__ int g_cdp[kMaxRuns * 16];        // n-step: their positions, (row & 0xffff) | col << 16
__shared__ int g_runn[kMaxRuns];            // n-step: candidates of run r in this round (0 = finished)
__shared__ int g_runres[kMaxRuns][8];       // n-step: per run {best row, col, second row, col, sad, num00, moved}
// Copies descriptor `bi` into shared memory (11 words) and synchronises the CTA.
// With l.start_from the start mv comes from an earlier launch's result instead (see
// B200MeLaunch.start_from / start_group): full-pel results x 8 for the sub-pel kernel,
// GET_MV_RAWPEL (mv.h:28) of sub-pel results for the full-pel kernel.
template <bool FULLPEL>
__device__ __forceinline__ void load_desc(const B200MeLaunch &l, int bi) {
  static_assert(sizeof(B200MeBlock) == 44, "descriptor is 11 words");
  if (threadIdx.x < 11)
    reinterpret_cast<int *>(&g_desc)[threadIdx.x] = reinterpret_cast<const int *>(l.blocks + bi)[threadIdx.x];
  __syncthreads();
  if (l.start_from) {
    if (threadIdx.x == 0) {
      const B200MeResult &r = l.start_from[l.start_group > 1 ? bi / l.start_group : bi];
      if (FULLPEL) {
        g_desc.start_row = (int16_t)((r.best_row + 3 + (r.best_row >= 0)) >> 3);
        g_desc.start_col = (int16_t)((r.best_col + 3 + (r.best_col >= 0)) >> 3);
      } else {
        g_desc.start_row = (int16_t)(r.best_row * 8);
        g_desc.start_col = (int16_t)(r.best_col * 8);
      }
    }
    __syncthreads();
  }
}

// Dynamic shared memory of both kernels: source block at offset 0, then the staged reference
// window.  The hot SAD loops address it through this array (not through the generic
// pointers of Ctx) so that the compiler emits LDS with 32-bit addresses instead of generic
// loads with 64-bit address arithmetic.
extern __shared__ __align__(128) unsigned char me_smem[];
// The search-site table of the CTA's method (copied from global memory at kernel start).
__shared__ Sites g_sites;
__device__ __forceinline__ const unsigned *smem_words(int byte_off) {
  return reinterpret_cast<const unsigned *>(me_smem + byte_off);
}

// ---- TMA staging -----------------------------------------------------------------------
// With plane extents known (B200MeLaunch.cur_rows / ref_rows), 16-byte aligned bases and
// strides, the source block and the search window are brought into shared memory by the TMA
// engine: one thread arms an mbarrier with the byte count and issues cp.async.bulk.tensor.2d
// boxes (tensor of 32-bit words, so a box row may be up to 1 KB), the CTA waits on the
// barrier's phase.  The boxes have a fixed size per launch (the tensor map holds it): the
// window box is `win_pitch_w` words x kTmaRows rows and a search issues as many as its own
// window height needs; out-of-bounds parts of a box are zero-filled by the hardware and are
// never read.  Destination pitches are 4 (mod 8) words like the cooperative copy's, so the
// vector SAD paths stay conflict-free.
constexpr int kTmaRows = 64;
struct TmaCfg {
  int use;          // 0: cooperative copy
  int win_pitch_w;  // words per window row in shared memory (= box width)
  int src_pitch_w;  // words per source row in shared memory (= box width)
  int src_rows;     // source box height
};
__shared__ __align__(8) unsigned long long g_mbar;

__device__ __forceinline__ unsigned smem_u32(const void *p) {
  return (unsigned)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long *bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, unsigned phase) {
  asm volatile(
      "{\n\t.reg .pred P1;\n\tLAB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n\t"
      "@P1 bra DONE;\n\tbra LAB_WAIT;\n\tDONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(phase)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *dst, const CUtensorMap *map, int x, int y,
                                            unsigned long long *bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
          "r"(smem_u32(dst)),
      "l"(reinterpret_cast<unsigned long long>(map)), "r"(smem_u32(bar)), "r"(x), "r"(y)
      : "memory");
}

// ------------------------------------------------------------------- mv cost -----
__device__ __forceinline__ int mv_joint(int r, int c) { return (r != 0) * 2 + (c != 0); }

__device__ __forceinline__ int mv_cost_tab(const Ctx &k, int dr, int dc) {
  return k.l->mvjcost[mv_joint(dr, dc)] + k.l->mvcost[0][dr] + k.l->mvcost[1][dc];
}

// mvsad_err_cost (mcomp.c:329).  The full-pel reference mv (GET_MV_RAWPEL of ref_mv) and the
// L1 weight of the cost type are per-search constants kept in Ctx (init_costs).
__device__ __forceinline__ void init_costs(Ctx &k) {
  k.col_min = k.d.col_min, k.col_max = k.d.col_max, k.row_min = k.d.row_min, k.row_max = k.d.row_max;
  const int rr = k.d.ref_mv_row, rc = k.d.ref_mv_col;
  k.cost_fr = (rr + 3 + (rr >= 0)) >> 3;
  k.cost_fc = (rc + 3 + (rc >= 0)) >> 3;
  const int t = k.d.mv_cost_type;
  k.l1_mul = t == B200_MV_COST_L1_LOWRES ? 32 : (t == B200_MV_COST_L1_MIDRES ? 15 : (t == B200_MV_COST_L1_HDRES ? 8 : 0));
}
__device__ __forceinline__ int sad_cost(const Ctx &k, int row, int col) {
  const int dr = (short)((row - k.cost_fr) * 8), dc = (short)((col - k.cost_fc) * 8);
  if (k.d.mv_cost_type == B200_MV_COST_ENTROPY) {
    if (!k.l->mvcost[0]) return 0;  // no tables given with the launch (me_check pairs them up)
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
  return col >= k.col_min && col <= k.col_max && row >= k.row_min && row <= k.row_max;
}
__device__ __forceinline__ Mv clamp_mv(const Ctx &k, Mv m) {
  m.col = min(max(m.col, k.col_min), k.col_max);
  m.row = min(max(m.row, k.row_min), k.row_max);
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
                                         int n_warps, int *sum_out, unsigned *sse_out, int warp_in = -1) {
  // warp_in >= 0: the rows are dealt out to a group of n_warps warps, this one being its warp_in-th
  const int lane = threadIdx.x & 31, warp = warp_in >= 0 ? warp_in : (int)(threadIdx.x >> 5);
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
// Operands of a compound search: prediction from the other reference and optional blend mask.  Kept in
// shared memory, not in Ctx: the context is copied and spilled by the hot single-reference paths.
struct CompoundOps {
  const unsigned char *second;
  const unsigned char *cmask;  // nullptr: average
  int cmask_stride;
  int cinv;
};
__shared__ CompoundOps g_cmp;
// compound descriptor of search `bi`; the CTA is synchronised by the staging that follows
__device__ __forceinline__ void init_compound(Ctx &k, const B200MeLaunch &l, int bi) {
  k.compound = l.compound != nullptr && l.second_pred != nullptr;
  if (k.compound && threadIdx.x == 0) {
    const B200MeCompound c = l.compound[bi];
    CompoundOps o;
    o.second = static_cast<const unsigned char *>(l.second_pred) + (size_t)c.second_offset * k.px;
    o.cmask = nullptr;
    o.cmask_stride = 0;
    o.cinv = 0;
    if (c.use_mask && l.mask) {
      o.cmask = l.mask + c.mask_offset;
      o.cmask_stride = c.mask_stride;
      o.cinv = c.invert_mask != 0;
    }
    g_cmp = o;
  }
}
// AOM_BLEND_AVG / AOM_BLEND_A64 (aom_dsp/blend.h:23-40) of predictor sample pv of pixel (x, y), i = y * bw + x
__device__ __forceinline__ int compound_blend(const Ctx &k, int pv, int i, int y, int x) {
  const unsigned char *second = g_cmp.second, *cmask = g_cmp.cmask;
  const int s2 = k.px == 1 ? (int)second[i] : (int)reinterpret_cast<const unsigned short *>(second)[i];
  if (!cmask) return (s2 + pv + 1) >> 1;
  const int m = cmask[y * g_cmp.cmask_stride + x];
  return g_cmp.cinv ? (m * s2 + (64 - m) * pv + 32) >> 6 : (m * pv + (64 - m) * s2 + 32) >> 6;
}
// get_mvpred_compound_sad (mcomp.c:729) with second_pred: vfp->sdaf / vfp->msdf, one warp per candidate
__device__ __noinline__ unsigned warp_sad_compound(const Ctx &k, int row, int col) {
  const int lane = threadIdx.x & 31, bw = k.d.bw, n = bw * k.d.bh;
  const unsigned char *rp = k.ref0 + ((long long)row * k.ref_pitch + col) * k.px;
  const int sp = k.src_pitch_w * 4 / k.px;  // source pitch in pixels (shared copy)
  unsigned acc = 0;
  for (int i = lane; i < n; i += 32) {
    const int y = i / bw, x = i - y * bw;
    int p, s;
    if (k.px == 1) {
      p = rp[(long long)y * k.ref_pitch + x];
      s = k.src[y * sp + x];
    } else {
      p = reinterpret_cast<const unsigned short *>(rp)[(long long)y * k.ref_pitch + x];
      s = reinterpret_cast<const unsigned short *>(k.src)[y * sp + x];
    }
    acc += (unsigned)abs(s - compound_blend(k, p, i, y, x));
  }
  return warp_sum_u(acc) >> k.sad_shift;
}
__device__ __forceinline__ unsigned warp_sad(const Ctx &k, int row, int col, bool skip, int lane) {
  return sad_finish(k, warp_sad_part(k, row, col, skip, 0, 1), skip);
}
// the SAD of av1_refining_search_8p_c: get_mvpred_compound_sad (the only full-pel search served with
// second_pred; the branch stays out of the single-reference paths of the other kernel instances)
__device__ __forceinline__ unsigned warp_sad_refine(const Ctx &k, int row, int col, bool skip, int lane) {
  if (k.compound) return warp_sad_compound(k, row, col);
  return warp_sad(k, row, col, skip, lane);
}

// ---- CTA-cooperative evaluation of a step's candidates (staged 8-bit path) ----------
// One warp per candidate re-reads the source block from shared memory for every candidate
// (4 KB + 6 KB per 64x64 candidate) and pays four dependent row passes per warp: the n-step
// phase was bound by shared-memory bandwidth and latency.  Here a LANE owns one row segment
// of up to 64 pixels (SEGW words) whose source words stay in registers for the whole step,
// task = pass * 32 + lane with a quarter warp = 8 consecutive rows of one segment column
// (pitches of 4 mod 8 words: conflict-free uint4 reads), and a warp walks the step's
// candidates over its rows: SEGW/4 + 1 LDS.128, SEGW SHF, SEGW VABSDIFF4 per candidate.
// A 64x64 block is two passes, so the 8 warps form 4 candidate groups.  Warp totals are added
// into out[i] with shared-memory atomics: out[0 .. n) must be zero on entry.  Candidates
// outside the mv limits are skipped (their slot stays 0; callers test in_range again).
// Lane i of every warp derives candidate i's window address once per step; the loop
// fetches it with a shuffle.  All control flow is warp-uniform.
template <int N>
__device__ __forceinline__ void load_words(const unsigned *p, unsigned (&o)[N]) {
  if constexpr (N >= 4) {
#pragma unroll
    for (int v = 0; v < N / 4; ++v) {
      const uint4 t = reinterpret_cast<const uint4 *>(p)[v];
      o[4 * v] = t.x, o[4 * v + 1] = t.y, o[4 * v + 2] = t.z, o[4 * v + 3] = t.w;
    }
  } else if constexpr (N == 2) {
    const uint2 t = *reinterpret_cast<const uint2 *>(p);
    o[0] = t.x, o[1] = t.y;
  } else {
    o[0] = p[0];
  }
}

#ifndef B200_ME_SEG_ACC
#define B200_ME_SEG_ACC 8
#endif
template <int SEGW, int WO>
__device__ __forceinline__ unsigned seg_sad_regs(const unsigned (&sv)[SEGW], int addr, unsigned sh) {
  constexpr int NV = SEGW / 4 + 1;
  unsigned r[4 * NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const uint4 t = *reinterpret_cast<const uint4 *>(me_smem + addr + 16 * v);
    r[4 * v] = t.x, r[4 * v + 1] = t.y, r[4 * v + 2] = t.z, r[4 * v + 3] = t.w;
  }
  // one accumulator chain per word up to kAcc: a candidate is a single dependent pass over the lane's
  // segment, so the chain depth (SEGW / kAcc VABSDIFF4.ACC) is latency the warp cannot hide
  constexpr int kAcc = SEGW < B200_ME_SEG_ACC ? SEGW : B200_ME_SEG_ACC;
  unsigned a[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; ++i) a[i] = 0;
#pragma unroll
  for (int i = 0; i < SEGW; ++i)
    a[i % kAcc] = sad4_acc(sv[i], __funnelshift_r(r[WO + i], r[WO + i + 1], sh), a[i % kAcc]);
#pragma unroll
  for (int w = kAcc / 2; w > 0; w >>= 1)
#pragma unroll
    for (int i = 0; i < w; ++i) a[i] += a[i + w];
  return a[0];
}

// The same with the loads addressed through a 32-bit shared-window address (no per-call S2UR / ULEA of the
// dynamic shared array) and the PREVIOUS candidate's warp reduction issued while they are in flight: the
// grouped n-step loop is one dependent pass per candidate, so the REDUX latency would otherwise be exposed
// once per candidate.  prev: this lane's sum of the previous candidate; prev_total: its warp total.
// PX == 2 (uint16 pixels): |s - r| per half word = max - min (VIMNMX.U16x2) added into packed 16-bit sums -- a
// lane adds at most 16 terms of at most 12 bits per half (16 * 4095 < 65536) -- folded by one IDP.2A at the end.
template <int SEGW, int WO, int PX = 1>
__device__ __forceinline__ unsigned seg_sad_regs_p(const unsigned (&sv)[SEGW], unsigned saddr, unsigned sh,
                                                   unsigned prev, unsigned &prev_total) {
  static_assert(PX == 1 || SEGW <= 16, "packed 16-bit sums hold 16 terms");
  constexpr int NV = SEGW / 4 + 1;
  unsigned r[4 * NV];
#pragma unroll
  for (int v = 0; v < NV; ++v)
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[4 * v]), "=r"(r[4 * v + 1]), "=r"(r[4 * v + 2]), "=r"(r[4 * v + 3])
                 : "r"(saddr + 16u * v));
  prev_total = __reduce_add_sync(0xffffffffu, prev);
  constexpr int kAcc = SEGW < B200_ME_SEG_ACC ? SEGW : B200_ME_SEG_ACC;
  unsigned a[kAcc];
#pragma unroll
  for (int i = 0; i < kAcc; ++i) a[i] = 0;
#pragma unroll
  for (int i = 0; i < SEGW; ++i) {
    const unsigned rv = __funnelshift_r(r[WO + i], r[WO + i + 1], sh);
    if constexpr (PX == 1) a[i % kAcc] = sad4_acc(sv[i], rv, a[i % kAcc]);
    else a[i % kAcc] = a[i % kAcc] + __vmaxu2(sv[i], rv) - __vminu2(sv[i], rv);
  }
#pragma unroll
  for (int w = kAcc / 2; w > 0; w >>= 1)
#pragma unroll
    for (int i = 0; i < w; ++i) a[i] += a[i + w];
  if constexpr (PX == 2) return __dp2a_lo(a[0], 0x0101u, 0u);
  return a[0];
}

__device__ __forceinline__ bool multi_ok(const Ctx &k) { return k.fast && k.px == 1 && k.d.bw >= 16; }

// A lane's share of the block, fixed for one search and skip mode: hoisted out of the step
// loop (deriving it costs more than evaluating a candidate).  Every block has at most 8
// passes of 32 lane tasks (128 rows x 2 segments), so a warp owns exactly one pass and
// kWarps / passes warps share it, each taking every (kWarps / passes)-th candidate.
template <int SEGW>
struct LaneShare {
  unsigned sv[SEGW];  // the lane's source words
  int l_off;          // byte offset of the lane's segment inside a candidate block
  int c0, cstep;      // candidate subset of the warp
  bool act;
};

template <int SEGW>
__device__ __forceinline__ void lane_share_init(const Ctx &k, bool skip, LaneShare<SEGW> &ls) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int R = skip ? k.d.bh >> 1 : k.d.bh, rmul = skip ? 2 : 1;
  const int lS = k.d.bw * k.px > 4 * SEGW ? 1 : 0;     // log2(segments per row): 128-byte rows have two
  const int np = ((R << lS) + 31) >> 5;                // passes: 1, 2, 4 or 8
  const int lP = np <= 1 ? 0 : 32 - __clz(np - 1);
  const int T = ((warp & ((1 << lP) - 1)) << 5) + lane, q = T >> 3;
  const int seg = q & ((1 << lS) - 1), row = ((q >> lS) << 3) + (T & 7);
  ls.act = row < R;
  const int yy = ls.act ? row * rmul : 0;
  ls.l_off = yy * k.ref_pitch * k.px + seg * (4 * SEGW);
  ls.c0 = warp >> lP;
  ls.cstep = kWarps >> lP;
  load_words<SEGW>(smem_words(yy * k.src_pitch_w * 4 + seg * (4 * SEGW)), ls.sv);
}

// Candidate i of the step lives in g_cda[i] (window byte address, or -1 outside the limits).
template <int SEGW>
__device__ __forceinline__ void cta_sad_multi_t(bool skip, int n, const int *cda, unsigned *out,
                                                const LaneShare<SEGW> &ls) {
  const int lane = threadIdx.x & 31;
  for (int i = ls.c0; i < n; i += ls.cstep) {
    const int a = cda[i];
    if (a < 0) continue;
    const int addr = (a & ~15) + ls.l_off;
    const unsigned sh = (unsigned)(a & 3) * 8;
    unsigned acc;
    switch ((a >> 2) & 3) {
      case 0: acc = seg_sad_regs<SEGW, 0>(ls.sv, addr, sh); break;
      case 1: acc = seg_sad_regs<SEGW, 1>(ls.sv, addr, sh); break;
      case 2: acc = seg_sad_regs<SEGW, 2>(ls.sv, addr, sh); break;
      default: acc = seg_sad_regs<SEGW, 3>(ls.sv, addr, sh); break;
    }
#ifdef EXP_NOREDUX
    if (acc == 0x12345) out[i] = acc;
#else
    acc = warp_sum_u(ls.act ? acc : 0u);
#ifdef EXP_NOATOM
    if (lane == 0 && acc == 0x12345) out[i] = acc;
#else
    if (lane == 0) atomicAdd(&out[i], skip ? 2 * acc : acc);
#endif
#endif
  }
}

// SEGW = 0: no register-resident share (unstaged window, high bit depth, narrow blocks)
template <>
struct LaneShare<0> {};
__device__ __forceinline__ int multi_segw(const Ctx &k) {
  if (!multi_ok(k)) return 0;
  return k.d.bw >= 64 ? 16 : (k.d.bw == 32 ? 8 : 4);
}

// Lane share of the warp-group n-step (diamond_runs_grouped), which also serves uint16 planes: words per lane
// by the row's BYTES; rows of more than 128 bytes (128-pixel wide blocks at high bit depth) do not fit its
// 8 passes and keep the lock-step path.
__device__ __forceinline__ int grouped_segw(const Ctx &k) {
  if (k.px == 1) return multi_segw(k);
  if (!(k.fast && k.d.bw >= 16 && k.d.bw <= 64)) return 0;
  return k.d.bw >= 32 ? 16 : 8;
}

// Small 8-bit blocks on the share-less path (8x8 .. 16x16: the temporal filter's blocks, whose frame-wide
// n-step reads the reference through L1/L2): one warp per candidate leaves 16 or 24 of its lanes idle and pays
// the load latency once per candidate.  Here a lane owns one row and a warp evaluates 32 / bh candidates at a
// time (segmented shuffle sum over the bh lanes of a candidate).  W = words per row.
template <int W>
__device__ __noinline__ void eval_small_rows(const Ctx &k, int n, unsigned *out, int slot0) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bh = k.d.bh, lb = 31 - __clz(bh), cpw = 32 >> lb;  // bh is 8 or 16
  const int sub = lane >> lb, y = lane & (bh - 1);
  unsigned sv[W];
#pragma unroll
  for (int j = 0; j < W; ++j) sv[j] = reinterpret_cast<const unsigned *>(k.src)[y * k.src_pitch_w + j];
  for (int i0 = warp * cpw; i0 < n; i0 += kWarps * cpw) {
    const int i = i0 + sub;
    const bool ok = i < n && g_cda[slot0 + i] >= 0;
    unsigned acc = 0;
    if (ok) {
      const int p = g_cdp[slot0 + i];
      const int row = (short)(p & 0xffff), col = p >> 16;
      acc = seg_sad<W, 1>(sv, k.ref0 + (long long)(row + y) * k.ref_pitch + col);
    }
    for (int o = bh >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (ok && y == 0) out[i] = acc;
  }
}

// The step's candidates (g_cda / g_cdp, n of them) -> out[i] = SAD (finished: doubled when rows
// are skipped, shifted at high bit depth); slots of candidates outside the limits are left
// untouched.  out[0 .. n) must be zero on entry; the caller synchronises the CTA before reading.
template <int SEGW>
__device__ __forceinline__ void eval_candidates(const Ctx &k, bool skip, int n, unsigned *out,
                                                const LaneShare<SEGW> &ls, int slot0 = 0) {
  if constexpr (SEGW > 0) {
    cta_sad_multi_t<SEGW>(skip, n, g_cda + slot0, out, ls);
  } else {
    if (k.px == 1 && !skip && (k.d.bh == 8 || k.d.bh == 16) && (k.d.bw == 8 || k.d.bw == 16)) {  // CTA-uniform
      if (k.d.bw == 8) eval_small_rows<2>(k, n, out, slot0);
      else eval_small_rows<4>(k, n, out, slot0);
      return;
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = warp; i < n; i += kWarps) {
      if (g_cda[slot0 + i] < 0) continue;
      const int p = g_cdp[slot0 + i];
      const unsigned v = warp_sad(k, (short)(p & 0xffff), p >> 16, skip, lane);
      if (lane == 0) out[i] = v;
    }
  }
}

// Warp 0 publishes a step's candidates: positions pos(0 .. n) -> g_cdp, window addresses (or
// 0 on the share-less path) / -1 outside the limits -> g_cda.  n <= 16.
template <int SEGW, class PosF>
__device__ __forceinline__ void publish_candidates(const Ctx &k, int n, PosF pos, int slot0 = 0) {
  const int lane = threadIdx.x & 31;
  if (lane < 16) {
    int a = -1, p = 0;
    if (lane < n) {
      const Mv m = pos(lane);
      p = (m.row & 0xffff) | (m.col << 16);
      if (in_range(k, m.row, m.col)) a = SEGW > 0 ? k.win0_off + m.row * k.ref_pitch + m.col : 0;
    }
    g_cda[slot0 + lane] = a;
    g_cdp[slot0 + lane] = p;
  }
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
