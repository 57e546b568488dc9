// me_subpel.cuh -- av1_find_best_sub_pixel_tree and its _pruned / _pruned_more variants
// (mcomp.c:2907-3126) on one CTA (included by me.cu).
// mcomp.c:2371-2551 (error functions, check_better), :2566-2779 (level checks),
// :2781-2868 (centre error), :3128-3196 (the tree).  Unscaled references,
// last_mv_search_list == NULL.
#pragma once

namespace b200 {
namespace me {

// Interpolation kernels at 1/8-pel positions: even rows of av1_bilinear_filters,
// av1_sub_pel_filters_4, av1_sub_pel_filters_8 (av1/common/filter.h:110-135,210-222)
__constant__ short c_me_taps[3][8][8] = {
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
__constant__ unsigned char c_me_bil[8][2] = { { 128, 0 }, { 112, 16 }, { 96, 32 }, { 80, 48 },
                                              { 64, 64 }, { 48, 80 },  { 32, 96 }, { 16, 112 } };

struct SubCtx {
  Ctx k;
  __device__ explicit SubCtx(const B200MeBlock &desc) : k(desc) {}
  unsigned short *tmp;  // shared: (bh + 7) * bw horizontally filtered samples
  int search_type;      // B200_USE_*
  Mv start;             // 1/8 pel: the descriptor's start mv, or 8 x the chained full-pel result
};

__device__ __forceinline__ int ref_px(const Ctx &k, const unsigned char *p, long long idx) {
  return k.px == 1 ? (int)p[idx] : (int)reinterpret_cast<const unsigned short *>(p)[idx];
}
__device__ __forceinline__ int src_px(const Ctx &k, int i) {
  return k.px == 1 ? (int)k.src[i] : (int)reinterpret_cast<const unsigned short *>(k.src)[i];
}

// CTA-wide: variance of (pred(mv) - src).  mode 0: pred = full-pel reference block
// (setup_center_error, vf(y, src)); 1: 2-tap bilinear svf (estimated_pref_error);
// 2: aom_upsampled_pred with 2/4/8 taps then vf (upsampled_pref_error).
__device__ unsigned cta_pref_error(const SubCtx &s, Mv mv, int mode, unsigned *sse_out) {
  const Ctx &k = s.k;
  const int bw = k.d.bw, bh = k.d.bh, n = bw * bh, lw = 31 - __clz(bw);
  const int fr = mv.row >> 3, fc = mv.col >> 3, sx = mv.col & 7, sy = mv.row & 7;
  const unsigned char *rp = k.ref0 + ((long long)fr * k.ref_pitch + fc) * k.px;
  const int maxv = k.px == 1 ? 255 : (1 << k.l->bd) - 1;
  long long sum = 0;
  unsigned long long sse = 0;
  __syncthreads();
  if (mode == 0) {
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const int y = i >> lw, x = i & (bw - 1);  // block widths are powers of two
      int pv = ref_px(k, rp, (long long)y * k.ref_pitch + x);
      if (k.compound) pv = compound_blend(k, pv, i, y, x);  // setup_center_error (:2781)
      const int df = pv - src_px(k, i);
      sum += df;
      sse += (unsigned)(df * df);
    }
  } else if (mode == 1) {
    const int f0 = c_me_bil[sx][0], f1 = c_me_bil[sx][1], g0 = c_me_bil[sy][0], g1 = c_me_bil[sy][1];
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const int y = i >> lw, x = i & (bw - 1);  // block widths are powers of two
      const long long o = (long long)y * k.ref_pitch + x;
      const int h0 = (ref_px(k, rp, o) * f0 + ref_px(k, rp, o + 1) * f1 + 64) >> 7;
      const int h1 = (ref_px(k, rp, o + k.ref_pitch) * f0 + ref_px(k, rp, o + k.ref_pitch + 1) * f1 + 64) >> 7;
      int pv = (h0 * g0 + h1 * g1 + 64) >> 7;
      if (k.px == 1) pv &= 0xff;  // uint8 store of the second pass (value <= 255 anyway)
      if (k.compound) pv = compound_blend(k, pv, i, y, x);  // svaf / msvf
      const int df = pv - src_px(k, i);
      sum += df;
      sse += (unsigned)(df * df);
    }
  } else {
    const int set = s.search_type == B200_USE_2_TAPS ? 0 : (s.search_type == B200_USE_4_TAPS ? 1 : 2);
    const short *kx = c_me_taps[set][sx], *ky = c_me_taps[set][sy];
    // horizontal pass (clipped to the pixel range like the reference's pixel-typed temp);
    // rows -3..bh+3 only when a vertical filter follows
    const int top = sy ? 3 : 0, ih = sy ? bh + 7 : bh;
    for (int i = threadIdx.x; i < ih * bw; i += kThreads) {
      const int y = i >> lw, x = i & (bw - 1);  // block widths are powers of two
      const long long o = (long long)(y - top) * k.ref_pitch + x;
      int v;
      if (sx) {
        int a = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) a += ref_px(k, rp, o + t - 3) * kx[t];
        v = min(max((a + 64) >> 7, 0), maxv);
      } else {
        v = ref_px(k, rp, o);
      }
      s.tmp[i] = (unsigned short)v;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += kThreads) {
      const int y = i >> lw, x = i & (bw - 1);  // block widths are powers of two
      int pv;
      if (sy) {
        int a = 0;
#pragma unroll
        for (int t = 0; t < 8; ++t) a += (int)s.tmp[(y + t) * bw + x] * ky[t];
        pv = min(max((a + 64) >> 7, 0), maxv);
      } else {
        pv = s.tmp[i];
      }
      if (k.compound) pv = compound_blend(k, pv, i, y, x);  // aom_comp_avg / comp_mask_upsampled_pred
      const int df = pv - src_px(k, i);
      sum += df;
      sse += (unsigned)(df * df);
    }
  }
  sum = warp_sum_ll(sum);
  sse = (unsigned long long)warp_sum_ll((long long)sse);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) {
    g_red[warp] = sum;
    g_red[kWarps + warp] = (long long)sse;
  }
  __syncthreads();
  long long tsum = 0;
  unsigned long long tsse = 0;
  for (int w = 0; w < kWarps; ++w) {
    tsum += g_red[w];
    tsse += (unsigned long long)g_red[kWarps + w];
  }
  return finish_var(tsum, tsse, n, k.l->highbd != 0, k.l->bd, sse_out);
}

// ---- 8-bit fast path of upsampled_pref_error (mode 2) -----------------------------------
// The scalar path above spends ~60 instructions per predicted pixel (byte loads of the
// reference through L1, 16 multiply-adds, one candidate at a time with three CTA barriers
// each).  Here the reachable reference window ((bw + 9) x (bh + 9): the tree never leaves
// +-7/8 pel around its start) and the TRANSPOSED source block are staged in shared memory once
// per search, and up to four candidates are evaluated per pass:
//   stage 1  horizontal 8-tap filter, 4 vertically adjacent outputs per thread: the 8 bytes
//            under the taps come from 3 aligned words + 2 PRMT (per-lane selector), the taps are
//            two packed s8x4 words, 2 DP4A per output; the clipped u8 results are packed into
//            one word of the TRANSPOSED intermediate tmpT[x][y] (odd word pitch: conflict-free)
//   stage 2  vertical filter along tmpT rows: 3 words + 6 PRMT + 8 DP4A for 4 outputs, then
//            VABSDIFF4 against the transposed source word, sse += DP4A(d, d), sum += DP4A(p, 1)
// about 20 instructions per pixel.  sum(src) is formed once per search.  Integer-pel
// components skip their filter stage exactly like aom_upsampled_pred_c (reconinter_enc.c:452).
struct SubFast {
  bool on;
  int win_off, win_pitch;  // byte offset in me_smem of window pixel (r_lo, c_lo); row pitch in bytes
  int r_lo, c_lo;          // full-pel position (relative to the block) of the window's first pixel
  int srcT_off, sp;        // transposed source block: srcT[x * sp + y]
  int tmpT_off, pt, tmpT_size;
  int sum_src;
};
__shared__ unsigned g_sacc[4][2];  // per candidate of a pass: sum of the prediction, sse

__device__ __forceinline__ int dp4a_us(unsigned a, unsigned b, int c) {
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
__device__ __forceinline__ unsigned clip8(int v) { return (unsigned)min(max(v, 0), 255); }
__device__ __forceinline__ unsigned pack_taps(const short *k) {
  return (unsigned)(k[0] & 0xff) | ((unsigned)(k[1] & 0xff) << 8) | ((unsigned)(k[2] & 0xff) << 16) |
         ((unsigned)(k[3] & 0xff) << 24);
}
__host__ __device__ __forceinline__ int odd_word_pitch(int bytes) {
  const int w = (bytes + 3) >> 2;
  return ((w & 1) ? w : w + 1) * 4;
}

// Stages window + transposed source.  `fast_off`: first free byte of me_smem; mbw / mbh: the
// launch's largest block (sizes the regions exactly like the host).
__device__ void subfast_init(const SubCtx &s, SubFast &f, int fast_off, int mbw, int mbh) {
  const Ctx &k = s.k;
  const B200MeLaunch &l = *k.l;
  const int bw = k.d.bw, bh = k.d.bh;
  f.on = false;
  const long long stride_b = l.ref_stride;
  if (k.px != 1 || (stride_b & 3) || bw > mbw || bh > mbh) return;
  f.r_lo = ((s.start.row - 7) >> 3) - 3;
  f.c_lo = ((s.start.col - 7) >> 3) - 3;
  const int wh = bh + 9, ww = bw + 9;
  const unsigned char *g0 = k.ref0 + (long long)f.r_lo * stride_b + f.c_lo;
  const int lead = (int)(reinterpret_cast<unsigned long long>(g0) & 3);
  f.win_pitch = ((mbw + 9 + 3 + 3) & ~3) + 4;
  const int win_bytes = (mbh + 9) * f.win_pitch;
  f.sp = odd_word_pitch(mbh);
  f.pt = odd_word_pitch(mbh + 8);
  f.tmpT_size = mbw * f.pt;
  f.srcT_off = fast_off + win_bytes;
  f.tmpT_off = f.srcT_off + mbw * f.sp;
  unsigned *s_win = reinterpret_cast<unsigned *>(me_smem + fast_off);
  const unsigned *ga = reinterpret_cast<const unsigned *>(g0 - lead);
  const int cw = (lead + ww + 3) >> 2, pw = f.win_pitch >> 2;
  for (int i = threadIdx.x; i < wh * cw; i += kThreads) {
    const int y = i / cw, x = i - y * cw;
    s_win[y * pw + x] = ga[(long long)y * (stride_b >> 2) + x];
  }
  f.win_off = fast_off + lead;
  // transposed source + its sum
  unsigned part = 0;
  for (int i = threadIdx.x; i < bw * bh; i += kThreads) {
    const int y = i / bw, x = i - y * bw;
    const unsigned v = k.src[i];
    me_smem[f.srcT_off + x * f.sp + y] = (unsigned char)v;
    part += v;
  }
  part = warp_sum_u(part);
  if (threadIdx.x == 0) g_bcast[2] = 0;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) atomicAdd(reinterpret_cast<unsigned *>(&g_bcast[2]), part);
  __syncthreads();
  f.sum_src = g_bcast[2];
  f.on = true;
}

// variance of (pred(mv[c]) - src) for the m <= 4 candidates with valid[c]: var_out / sse_out
__device__ void subfast_eval(const SubCtx &s, const SubFast &f, const Mv *mv, const bool *valid, int m,
                             unsigned *var_out, unsigned *sse_out) {
  const Ctx &k = s.k;
  const int bw = k.d.bw, bh = k.d.bh, lbw = 31 - __clz(bw);
  const int set = s.search_type == B200_USE_2_TAPS ? 0 : (s.search_type == B200_USE_4_TAPS ? 1 : 2);
  __syncthreads();  // previous readers of tmpT / g_sacc
  if (threadIdx.x < 8) (&g_sacc[0][0])[threadIdx.x] = 0;
  // ---- stage 1: horizontal
  for (int c = 0; c < m; ++c) {
    if (!valid[c]) continue;
    const int fr = mv[c].row >> 3, fc = mv[c].col >> 3, sx = mv[c].col & 7, sy = mv[c].row & 7;
    const int top = sy ? 3 : 0, ih = sy ? bh + 7 : bh, ngy = (ih + 3) >> 2;
    const int row0 = fr - f.r_lo - top;
    unsigned char *tmpT = me_smem + f.tmpT_off + c * f.tmpT_size;
    if (sx) {
      const unsigned klo = pack_taps(c_me_taps[set][sx]), khi = pack_taps(c_me_taps[set][sx] + 4);
      const int col0 = fc - f.c_lo - 3;
      for (int g = threadIdx.x; g < (ngy << lbw); g += kThreads) {
        const int x = g & (bw - 1), yg = g >> lbw;
        const int b0 = f.win_off + (row0 + 4 * yg) * f.win_pitch + col0 + x;
        const unsigned sel = 0x3210u + 0x1111u * (unsigned)(b0 & 3);
        const unsigned *w = reinterpret_cast<const unsigned *>(me_smem + (b0 & ~3));
        const int pw = f.win_pitch >> 2;
        unsigned out = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const unsigned w0 = w[j * pw], w1 = w[j * pw + 1], w2 = w[j * pw + 2];
          const int a = dp4a_us(__byte_perm(w1, w2, sel), khi, dp4a_us(__byte_perm(w0, w1, sel), klo, 64));
          out |= clip8(a >> 7) << (8 * j);
        }
        *reinterpret_cast<unsigned *>(tmpT + x * f.pt + 4 * yg) = out;
      }
    } else {
      const int col0 = fc - f.c_lo;
      for (int g = threadIdx.x; g < (ngy << lbw); g += kThreads) {
        const int x = g & (bw - 1), yg = g >> lbw;
        const unsigned char *p = me_smem + f.win_off + (row0 + 4 * yg) * f.win_pitch + col0 + x;
        const unsigned out = (unsigned)p[0] | ((unsigned)p[f.win_pitch] << 8) |
                             ((unsigned)p[2 * f.win_pitch] << 16) | ((unsigned)p[3 * f.win_pitch] << 24);
        *reinterpret_cast<unsigned *>(tmpT + x * f.pt + 4 * yg) = out;
      }
    }
  }
  __syncthreads();
  // ---- stage 2: vertical + variance sums
  for (int c = 0; c < m; ++c) {
    if (!valid[c]) continue;
    const int sy = mv[c].row & 7;
    const unsigned char *tmpT = me_smem + f.tmpT_off + c * f.tmpT_size;
    const unsigned klo = pack_taps(c_me_taps[set][sy]), khi = pack_taps(c_me_taps[set][sy] + 4);
    unsigned sum = 0, sse = 0;
    for (int g = threadIdx.x; g < ((bh >> 2) << lbw); g += kThreads) {
      const int x = g & (bw - 1), yg = g >> lbw;
      const unsigned *v = reinterpret_cast<const unsigned *>(tmpT + x * f.pt + 4 * yg);
      unsigned p;
      if (sy) {
        const unsigned v0 = v[0], v1 = v[1], v2 = v[2];
        const int a0 = dp4a_us(v1, khi, dp4a_us(v0, klo, 64));
        const int a1 = dp4a_us(__byte_perm(v1, v2, 0x4321), khi, dp4a_us(__byte_perm(v0, v1, 0x4321), klo, 64));
        const int a2 = dp4a_us(__byte_perm(v1, v2, 0x5432), khi, dp4a_us(__byte_perm(v0, v1, 0x5432), klo, 64));
        const int a3 = dp4a_us(__byte_perm(v1, v2, 0x6543), khi, dp4a_us(__byte_perm(v0, v1, 0x6543), klo, 64));
        p = clip8(a0 >> 7) | (clip8(a1 >> 7) << 8) | (clip8(a2 >> 7) << 16) | (clip8(a3 >> 7) << 24);
      } else {
        p = v[0];
      }
      const unsigned sv = *reinterpret_cast<const unsigned *>(me_smem + f.srcT_off + x * f.sp + 4 * yg);
      const unsigned d = __vabsdiffu4(p, sv);
      sse = __dp4a(d, d, sse);
      sum = __dp4a(p, 0x01010101u, sum);
    }
    sum = warp_sum_u(sum);
    sse = warp_sum_u(sse);
    if ((threadIdx.x & 31) == 0) {
      atomicAdd(&g_sacc[c][0], sum);
      atomicAdd(&g_sacc[c][1], sse);
    }
  }
  __syncthreads();
  for (int c = 0; c < m; ++c) {
    if (!valid[c]) continue;
    unsigned sse32;
    var_out[c] = finish_var((long long)(int)g_sacc[c][0] - f.sum_src, g_sacc[c][1], bw * bh, false, 8, &sse32);
    sse_out[c] = sse32;
  }
}

struct SubState {
  Mv best;
  unsigned besterr, sse;
  int distortion;
};

// check_better / check_better_fast (mcomp.c:2496-2551)
__device__ unsigned check_better(const SubCtx &s, SubState &st, Mv mv, int mode, int *better) {
  const Ctx &k = s.k;
  if (mv.col < k.d.col_min || mv.col > k.d.col_max || mv.row < k.d.row_min || mv.row > k.d.row_max)
    return (unsigned)INT_MAX;
  unsigned sse;
  const int mse = (int)cta_pref_error(s, mv, mode, &sse);
  const unsigned cost = (unsigned)err_cost(k, mv.row, mv.col) + (unsigned)mse;
  if (cost < st.besterr) {
    st.besterr = cost;
    st.best = mv;
    st.distortion = mse;
    st.sse = sse;
    if (better) *better |= 1;
  }
  return cost;
}

// check_better on up to 4 probes at once: the probes inside the limits are evaluated together
// (fast path) and then folded into the state in order, exactly as consecutive check_better calls
// would (the probe positions of a level do not depend on each other's outcome).
__device__ void check_better_n(const SubCtx &s, const SubFast &f, SubState &st, const Mv *mv, int m, int mode,
                               unsigned *cost, int *better) {
  const Ctx &k = s.k;
  if (!(f.on && mode == 2)) {
    for (int c = 0; c < m; ++c) cost[c] = check_better(s, st, mv[c], mode, better);
    return;
  }
  bool valid[4];
  unsigned var[4], sse[4];
  bool any = false;
  for (int c = 0; c < m; ++c) {
    valid[c] = !(mv[c].col < k.d.col_min || mv[c].col > k.d.col_max || mv[c].row < k.d.row_min ||
                 mv[c].row > k.d.row_max);
    any = any || valid[c];
  }
  if (any) subfast_eval(s, f, mv, valid, m, var, sse);
  for (int c = 0; c < m; ++c) {
    if (!valid[c]) {
      cost[c] = (unsigned)INT_MAX;
      continue;
    }
    const int mse = (int)var[c];
    cost[c] = (unsigned)err_cost(k, mv[c].row, mv[c].col) + (unsigned)mse;
    if (cost[c] < st.besterr) {
      st.besterr = cost[c];
      st.best = mv[c];
      st.distortion = mse;
      st.sse = sse[c];
      if (better) *better |= 1;
    }
  }
}

__device__ void subpel_search(const SubCtx &s, const SubFast &f, SubState &st) {
  const Ctx &k = s.k;
  const int forced_stop = k.d.subpel & 3, iters_per_step = (k.d.subpel >> 2) & 3;
  const int allow_hp = (k.d.flags & B200_ME_ALLOW_HP) != 0;
  const int mode = s.search_type != B200_USE_2_TAPS_ORIG ? 2 : 1;
  const int round = min(B200_FULL_PEL - forced_stop, 3 - !allow_hp);
  int hstep = 4;  // INIT_SUBPEL_STEP_SIZE
  st.best = s.start;
  unsigned sse, e;
  if (f.on && mode == 2) {
    const bool one = true;
    subfast_eval(s, f, &st.best, &one, 1, &e, &sse);
  } else {
    e = cta_pref_error(s, st.best, mode == 2 ? 2 : 0, &sse);
  }
  st.distortion = (int)e;
  st.sse = sse;
  st.besterr = e + (unsigned)err_cost(k, st.best.row, st.best.col);
  for (int iter = 0; iter < round; ++iter) {
    const Mv c = st.best;
    // first_level_check(_fast): four cardinals, then the diagonal of the better halves
    const Mv card[4] = { { c.row, c.col - hstep }, { c.row, c.col + hstep }, { c.row - hstep, c.col },
                         { c.row + hstep, c.col } };
    unsigned cost[4];
    check_better_n(s, f, st, card, 4, mode, cost, nullptr);
    const unsigned left = cost[0], right = cost[1], up = cost[2], down = cost[3];
    Mv diag = { up <= down ? -hstep : hstep, left <= right ? -hstep : hstep };
    const Mv dm = { c.row + diag.row, c.col + diag.col };
    check_better_n(s, f, st, &dm, 1, mode, cost, nullptr);
    // second_level_check_v2
    if (!(c.row == st.best.row && c.col == st.best.col) && iters_per_step > 1) {
      if (c.row == st.best.row) diag.row = -diag.row;
      else if (c.col == st.best.col) diag.col = -diag.col;
      const Mv two[2] = { { st.best.row + diag.row, st.best.col }, { st.best.row, st.best.col + diag.col } };
      const Mv db = { st.best.row + diag.row, st.best.col + diag.col };
      int better = 0;
      check_better_n(s, f, st, two, 2, mode, cost, &better);
      if (better) check_better_n(s, f, st, &db, 1, mode, cost, &better);
    }
    hstep >>= 1;
  }
}

// first_level_check(_fast) (mcomp.c:2566,2689): four cardinals, then the diagonal of the
// better halves.  Returns the diagonal step.
__device__ Mv first_level(const SubCtx &s, SubState &st, Mv c, int hstep, int mode) {
  const Mv l = { c.row, c.col - hstep }, r = { c.row, c.col + hstep };
  const Mv t = { c.row - hstep, c.col }, b = { c.row + hstep, c.col };
  const unsigned left = check_better(s, st, l, mode, nullptr);
  const unsigned right = check_better(s, st, r, mode, nullptr);
  const unsigned up = check_better(s, st, t, mode, nullptr);
  const unsigned down = check_better(s, st, b, mode, nullptr);
  const Mv diag = { up <= down ? -hstep : hstep, left <= right ? -hstep : hstep };
  const Mv dm = { c.row + diag.row, c.col + diag.col };
  check_better(s, st, dm, mode, nullptr);
  return diag;
}

// two_level_checks_fast (mcomp.c:2675): first_level_check_fast + second_level_check_fast
// (:2610), every probe through the bilinear estimate (mode 1).
__device__ void two_level_fast(const SubCtx &s, SubState &st, Mv c, int hstep, int iters) {
  const Mv diag = first_level(s, st, c, hstep, 1);
  if (iters <= 1) return;
  const int tr = c.row, tc = c.col, br = st.best.row, bc = st.best.col;
  if (tr != br && tc != bc) {
    const Mv c1 = { br, bc + diag.col }, c2 = { br + diag.row, bc };
    check_better(s, st, c1, 1, nullptr);
    check_better(s, st, c2, 1, nullptr);
  } else if (tr == br && tc != bc) {
    const Mv bl = { br + hstep, bc + diag.col }, tl = { br - hstep, bc + diag.col };
    const Mv rev = { br - diag.row, bc };
    check_better(s, st, bl, 1, nullptr);
    check_better(s, st, tl, 1, nullptr);
    check_better(s, st, rev, 1, nullptr);
  } else if (tr != br && tc == bc) {
    const Mv rl = { br + diag.row, bc + hstep }, ll = { br + diag.row, bc - hstep };
    const Mv rev = { br, bc - diag.col };
    check_better(s, st, rl, 1, nullptr);
    check_better(s, st, ll, 1, nullptr);
    check_better(s, st, rev, 1, nullptr);
  }
}

// divide_and_round (mcomp.c:2853)
__device__ __forceinline__ int div_round(int n, int d) {
  return ((n < 0) ^ (d < 0)) ? ((n - d / 2) / d) : ((n + d / 2) / d);
}

// av1_find_best_sub_pixel_tree_pruned (mcomp.c:2992, tree == 1) and _pruned_more (:2907,
// tree == 2); last_mv_search_list == NULL, unscaled reference.  cl = ms_params->cost_list or
// nullptr.  All control flow is CTA-uniform (every thread holds the same SubState).
__device__ void subpel_search_pruned(const SubCtx &s, SubState &st, int tree, const int *cl) {
  const Ctx &k = s.k;
  const int forced_stop = k.d.subpel & 3, iters_per_step = (k.d.subpel >> 2) & 3;
  const int allow_hp = (k.d.flags & B200_ME_ALLOW_HP) != 0;
  int hstep = 4;
  st.best = s.start;
  const Mv start = st.best;
  unsigned sse;
  const unsigned e = cta_pref_error(s, st.best, 0, &sse);  // setup_center_error :2781
  st.distortion = (int)e;
  st.sse = sse;
  st.besterr = e + (unsigned)err_cost(k, st.best.row, st.best.col);
  if (forced_stop == B200_FULL_PEL) return;
  const bool usable = cl && cl[0] != INT_MAX && cl[1] != INT_MAX && cl[2] != INT_MAX &&
                      cl[3] != INT_MAX && cl[4] != INT_MAX;
  if (tree == B200_SUBPEL_TREE_PRUNED_MORE && usable && cl[0] < cl[1] && cl[0] < cl[2] &&
      cl[0] < cl[3] && cl[0] < cl[4]) {
    // get_cost_surf_min (:2870), bits = 1
    const int ic = div_round(cl[1] - cl[3], cl[1] - 2 * cl[0] + cl[3]);
    const int ir = div_round(cl[4] - cl[2], cl[4] - 2 * cl[0] + cl[2]);
    if (ir != 0 || ic != 0) {
      const Mv m = { start.row + ir * hstep, start.col + ic * hstep };
      check_better(s, st, m, 1, nullptr);
    }
  } else if (tree == B200_SUBPEL_TREE_PRUNED && usable) {
    // the quadrant of the two cheaper full-pel neighbours: horizontal, vertical, diagonal
    const Mv h = { start.row, start.col + (cl[1] < cl[3] ? -hstep : hstep) };
    const Mv v = { start.row + (cl[2] < cl[4] ? hstep : -hstep), start.col };
    const Mv d = { v.row, h.col };
    check_better(s, st, h, 1, nullptr);
    check_better(s, st, v, 1, nullptr);
    check_better(s, st, d, 1, nullptr);
  } else {
    two_level_fast(s, st, start, hstep, iters_per_step);
  }
  if (forced_stop < B200_HALF_PEL) {
    hstep >>= 1;
    two_level_fast(s, st, st.best, hstep, iters_per_step);
  }
  if (allow_hp && forced_stop == B200_EIGHTH_PEL) {
    hstep >>= 1;
    two_level_fast(s, st, st.best, hstep, iters_per_step);
  }
}

__global__ void __launch_bounds__(kThreads, 4) subpel_kernel(const B200MeLaunch l) {
  const int bi = blockIdx.x;
  load_desc<false>(l, bi);
  SubCtx s(g_desc);
  s.k.l = &l;
  init_costs(s.k);
  s.k.sites = nullptr;
  s.k.px = l.highbd ? 2 : 1;
  s.k.sad_shift = 0;
  init_compound(s.k, l, bi);
  s.search_type = (s.k.d.subpel >> 4) & 3;
  s.start = Mv{ s.k.d.start_row, s.k.d.start_col };
  // the limits are sub-pel here: no window staging (window_budget 0), reference via L1/L2
  {
    const int px = s.k.px, bw = s.k.d.bw, bh = s.k.d.bh;
    const unsigned char *g_src = static_cast<const unsigned char *>(l.cur) + (size_t)s.k.d.src_offset * px;
    const int row_bytes = bw * px;
    for (int i = threadIdx.x; i < bh * row_bytes; i += kThreads) {
      const int y = i / row_bytes, x = i - y * row_bytes;
      me_smem[i] = g_src[(size_t)y * l.cur_stride * px + x];
    }
    s.k.src = me_smem;
    s.k.src_pitch_w = row_bytes / 4;  // dense
    s.k.fast = false;
    s.k.ref0 = static_cast<const unsigned char *>(l.ref) + (size_t)s.k.d.ref_offset * px;
    s.k.ref_pitch = l.ref_stride;
    s.tmp = reinterpret_cast<unsigned short *>(me_smem + ((bh * row_bytes + 15) & ~15));
    __syncthreads();
  }
  SubState st;
  const int tree = s.k.d.search_method;
  if (tree == B200_SUBPEL_TREE) {
    SubFast f;
    f.on = false;
    const int mbw = l.max_bw > 0 ? l.max_bw : 128, mbh = l.max_bh > 0 ? l.max_bh : 128;
    if (s.search_type != B200_USE_2_TAPS_ORIG && !l.highbd && s.k.d.bw >= 4 && !s.k.compound)
      subfast_init(s, f, (mbw * mbh + 16 + (mbh + 7) * mbw * 2 + 15) & ~15, mbw, mbh);
    subpel_search(s, f, st);
  } else {
    int cl[5];
    const bool have = (s.k.d.flags & B200_ME_COST_LIST) && l.cost_lists;
    if (have) {
      const int *g = l.cost_lists + (size_t)bi * (l.cost_list_stride > 0 ? l.cost_list_stride : 5);
      for (int i = 0; i < 5; ++i) cl[i] = g[i];
    }
    subpel_search_pruned(s, st, tree, have ? cl : nullptr);
  }
  if (threadIdx.x == 0) {
    B200MeResult r;
    memset(&r, 0, sizeof(r));
    r.best_row = (short)st.best.row;
    r.best_col = (short)st.best.col;
    r.second_row = r.second_col = -32768;
    r.ret = (int)st.besterr;
    r.distortion = st.distortion;
    r.sse = st.sse;
    l.results[bi] = r;
  }
}

}  // namespace me
}  // namespace b200
