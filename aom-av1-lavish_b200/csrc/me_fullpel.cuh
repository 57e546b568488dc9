// me_fullpel.cuh -- av1_full_pixel_search on one CTA (included by me.cu).
#pragma once

namespace b200 {
namespace me {

// One value computed by warp 0 and broadcast to the CTA.
template <class F>
__device__ __forceinline__ unsigned cta_value(const Ctx &k, F f) {
  const int lane = threadIdx.x & 31;
  __syncthreads();
  if (threadIdx.x < 32) {
    const unsigned v = f(lane);
    if (lane == 0) k.bcast[0] = (int)v;
  }
  __syncthreads();
  return (unsigned)k.bcast[0];
}

// get_mvpred_var_cost (mcomp.c:665): vf(src, ref) + sub-pel-domain mv cost.  The whole CTA
// accumulates (s - r) and (s - r)^2, partial sums meet in shared memory.
__device__ int var_cost(const Ctx &k, Mv m) {
  const int bw = k.d.bw, n = bw * k.d.bh;
  const unsigned char *rp = k.ref0 + ((long long)m.row * k.ref_pitch + m.col) * k.px;
  long long sum = 0;
  unsigned long long sse = 0;
  for (int i = threadIdx.x; i < n; i += kThreads) {
    const int y = i / bw, x = i - y * bw;
    int sv, rv;
    if (k.px == 1) {
      sv = k.src[y * k.src_pitch_w * 4 + x];
      rv = rp[(long long)y * k.ref_pitch + x];
    } else {
      sv = reinterpret_cast<const unsigned short *>(k.src)[y * k.src_pitch_w * 2 + x];
      rv = reinterpret_cast<const unsigned short *>(rp)[(long long)y * k.ref_pitch + x];
    }
    const int df = sv - rv;
    sum += df;
    sse += (unsigned)(df * df);
  }
  sum = warp_sum_ll(sum);
  sse = (unsigned long long)warp_sum_ll((long long)sse);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane == 0) {
    k.red[warp] = sum;
    k.red[kWarps + warp] = (long long)sse;
  }
  __syncthreads();
  long long tsum = 0;
  unsigned long long tsse = 0;
  for (int w = 0; w < kWarps; ++w) {
    tsum += k.red[w];
    tsse += (unsigned long long)k.red[kWarps + w];
  }
  unsigned sse32;
  const unsigned v = finish_var(tsum, tsse, n, k.l->highbd != 0, k.l->bd, &sse32);
  return (int)v + err_cost(k, m.row * 8, m.col * 8);
}

// diamond_search_sad (mcomp.c:1318-1477)
__device__ int diamond_search(const Ctx &k, bool skip, Mv start, int search_step, int *num00,
                              Mv *best, Mv *second, bool has_second) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const Sites &cfg = *k.sites;
  int off_center = 0, center_steps = 0;
  start = clamp_mv(k, start);
  const int tot = cfg.n_steps - search_step;
  Mv tmp_second = has_second ? *second : start;
  *best = start;
  unsigned bestsad = cta_value(k, [&](int ln) { return warp_sad(k, start.row, start.col, skip, ln); }) +
                     (unsigned)sad_cost(k, start.row, start.col);
  for (int step = tot - 1; step >= 0; --step) {
    const int n = cfg.per_step[step];
    for (int i = 1 + warp; i <= n; i += kWarps) {
      const int r = best->row + cfg.site[step][i][0], c = best->col + cfg.site[step][i][1];
      unsigned t = 0xffffffffu;
      if (in_range(k, r, c)) t = warp_sad(k, r, c, skip, lane) + (unsigned)sad_cost(k, r, c);
      if (lane == 0) k.cand[i] = t;
    }
    __syncthreads();
    int best_site = 0;
    for (int i = 1; i <= n; ++i) {
      const unsigned t = k.cand[i];
      if (t < bestsad) {
        bestsad = t;
        best_site = i;
      }
    }
    __syncthreads();
    if (best_site) {
      tmp_second = *best;
      best->row += cfg.site[step][best_site][0];
      best->col += cfg.site[step][best_site][1];
      off_center = 1;
    }
    if (!off_center) center_steps++;
    if (best_site == 0 && step > 2) {
      int next = cfg.radius[step - 1];
      while (next == cfg.radius[step] && step > 2) {
        center_steps++;
        --step;
        next = cfg.radius[step - 1];
      }
    }
  }
  *num00 = center_steps;
  if (has_second) *second = tmp_second;
  return (int)bestsad;
}

// calc_int_sad_list (mcomp.c:791) with costlist_has_sad == 0
__device__ void sad_list(const Ctx &k, bool skip, Mv best, int *cl) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (warp < 5) {
    const int dr = warp == 2 ? 1 : (warp == 4 ? -1 : 0);
    const int dc = warp == 1 ? -1 : (warp == 3 ? 1 : 0);
    const int r = best.row + dr, c = best.col + dc;
    int v = INT_MAX;
    if (warp == 0 || in_range(k, r, c)) v = (int)warp_sad(k, r, c, skip, lane) + sad_cost(k, r, c);
    if (lane == 0) k.bcast[1 + warp] = v;
  }
  __syncthreads();
  for (int i = 0; i < 5; ++i) cl[i] = k.bcast[1 + i];
}

// full_pixel_diamond (mcomp.c:1479-1526)
__device__ int full_pixel_diamond(const Ctx &k, bool skip, Mv start, int step_param, int *cl,
                                  Mv *best, Mv *second, bool has_second) {
  int n, num00 = 0;
  diamond_search(k, skip, start, step_param, &n, best, second, has_second);
  int bestsme = var_cost(k, *best);
  const int further = k.sites->n_steps - 1 - step_param;
  while (n < further) {
    ++n;
    Mv tmp;
    diamond_search(k, skip, start, step_param + n, &num00, &tmp, second, has_second);
    const int sme = var_cost(k, tmp);
    if (sme < bestsme) {
      bestsme = sme;
      *best = tmp;
    }
    if (num00) {
      n += num00;
      num00 = 0;
    }
  }
  if (cl) sad_list(k, skip, *best, cl);
  return bestsme;
}

// exhaustive_mesh_search (mcomp.c:1530-1599).  Candidates are visited row-major; at
// step 1 the reference works in groups of four columns and its tail loop stops one short
// of the last column (`i < end_col - c`, :1586) -- reproduced.
__device__ int mesh_search(const Ctx &k, bool skip, Mv start, int range, int step, Mv *best,
                           Mv *second, bool has_second) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  start = clamp_mv(k, start);
  *best = start;
  unsigned best_sad = cta_value(k, [&](int ln) { return warp_sad(k, start.row, start.col, skip, ln); }) +
                      (unsigned)sad_cost(k, start.row, start.col);
  const int r0 = max(-range, k.d.row_min - start.row), c0 = max(-range, k.d.col_min - start.col);
  const int r1 = min(range, k.d.row_max - start.row), c1 = min(range, k.d.col_max - start.col);
  if (r1 < r0 || c1 < c0) return (int)best_sad;
  const int nr = (r1 - r0) / step + 1;
  int nc;
  if (step > 1) {
    nc = (c1 - c0) / step + 1;
  } else {
    const int full = c1 - c0 + 1;
    nc = (full & 3) ? full - 1 : full;
  }
  const int total = nr * nc;
  for (int base = 0; base < total; base += kMaxCand) {
    const int cnt = min(kMaxCand, total - base);
    for (int j = warp; j < cnt; j += kWarps) {
      const int idx = base + j;
      const int ri = idx / nc, ci = idx - ri * nc;
      const int r = start.row + r0 + ri * step, c = start.col + c0 + ci * step;
      const unsigned t = warp_sad(k, r, c, skip, lane) + (unsigned)sad_cost(k, r, c);
      if (lane == 0) k.cand[j] = t;
    }
    __syncthreads();
    // Replay update_mvs_and_sad (:858) over the chunk without walking it serially: the
    // last improvement of the sequential scan is the FIRST occurrence of the chunk minimum
    // (if it beats the running best), and second_best is whatever was best just before it,
    // i.e. the first minimum of the prefix in front of the winner (or the old best).
    if (warp == 0) {
      const int lo = lane * (kMaxCand / 32), hi = min(cnt, lo + kMaxCand / 32);
      unsigned bv = best_sad;
      int bi = INT_MAX;
      for (int j = lo; j < hi; ++j) {
        const unsigned t = k.cand[j];
        if (t < bv) bv = t, bi = j;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov < bv || (ov == bv && oi < bi)) bv = ov, bi = oi;
      }
      const int win = bi;  // INT_MAX: nothing in the chunk beats the running best
      unsigned pv = best_sad;
      int pi = INT_MAX;
      if (win != INT_MAX) {
        for (int j = lo; j < min(hi, win); ++j) {
          const unsigned t = k.cand[j];
          if (t < pv) pv = t, pi = j;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          const unsigned ov = __shfl_xor_sync(0xffffffffu, pv, o);
          const int oi = __shfl_xor_sync(0xffffffffu, pi, o);
          if (ov < pv || (ov == pv && oi < pi)) pv = ov, pi = oi;
        }
      }
      if (lane == 0) {
        k.bcast[8] = win;
        k.bcast[9] = (int)bv;
        k.bcast[10] = pi;
      }
    }
    __syncthreads();
    {
      const int win = k.bcast[8], pi = k.bcast[10];
      if (win != INT_MAX) {
        best_sad = (unsigned)k.bcast[9];
        if (has_second) {
          if (pi != INT_MAX) {
            const int idx = base + pi, ri = idx / nc, ci = idx - ri * nc;
            second->row = start.row + r0 + ri * step;
            second->col = start.col + c0 + ci * step;
          } else {
            *second = *best;
          }
        }
        const int idx = base + win, ri = idx / nc, ci = idx - ri * nc;
        best->row = start.row + r0 + ri * step;
        best->col = start.col + c0 + ci * step;
      }
    }
    __syncthreads();
  }
  return (int)best_sad;
}

// full_pixel_exhaustive (mcomp.c:1603-1679)
__device__ int full_pixel_exhaustive(const Ctx &k, bool skip, Mv start, int *cl, Mv *best,
                                     Mv *second, bool has_second) {
  int interval = k.l->mesh_interval[0], range = k.l->mesh_range[0];
  *best = start;
  if (range < 7 || range > 256 || interval < 1 || interval > range) return INT_MAX;
  const int divisor = range / interval;
  range = max(range, (5 * max(abs(best->row), abs(best->col))) / 4);
  range = min(range, 256);
  interval = max(interval, range / divisor);
  if (k.d.flags & B200_ME_FINE_INTERVAL) interval = min(interval, 4);
  int bestsme = mesh_search(k, skip, *best, range, interval, best, second, has_second);
  if (interval > 1 && range > 7) {
    for (int i = 1; i < 4; ++i) {
      bestsme = mesh_search(k, skip, *best, k.l->mesh_range[i], k.l->mesh_interval[i], best, second,
                            has_second);
      if (k.l->mesh_interval[i] == 1) break;
    }
  }
  if (bestsme < INT_MAX) bestsme = var_cost(k, *best);
  if (cl) sad_list(k, skip, *best, cl);
  return bestsme;
}

__device__ __forceinline__ int ilog2i(int v) { return 31 - __clz(v); }

// av1_full_pixel_search (mcomp.c:1755-1895), n-step / diamond family
__device__ int full_pixel_search(const Ctx &k, bool skip, int *cl, Mv *best, Mv *second) {
  const Mv start = { k.d.start_row, k.d.start_col };
  const int mi_log2 = ilog2i(k.d.bw >> 2) + ilog2i(k.d.bh >> 2);
  int var = 0;
  for (int pass = 0; pass < 2; ++pass) {  // second pass = the reference's redo without skip
    bool run_mesh = (k.d.flags & B200_ME_RUN_MESH) != 0;
    best->row = best->col = -32768;  // MARK_MV_INVALID
    second->row = second->col = -32768;
    for (int i = 0; i < 5; ++i) cl[i] = INT_MAX;
#ifdef B200_ME_TIMING
    const long long tq0 = clock64();
#endif
    var = full_pixel_diamond(k, skip, start, k.d.step_param, cl, best, second, true);
#ifdef B200_ME_TIMING
    k.bcast[12] = (int)((clock64() - tq0) >> 4);
#endif
    if (!run_mesh &&
        (k.d.search_method == B200_ME_NSTEP || k.d.search_method == B200_ME_NSTEP_8PT)) {
      if (var > (k.d.force_mesh_thresh >> (10 - mi_log2))) run_mesh = true;
    }
    if (k.d.flags & B200_ME_PRUNE_MESH) {
      const int diff = max(abs(start.row - best->row), abs(start.col - best->col));
      if (diff <= k.d.mesh_mv_diff_threshold) run_mesh = false;
    }
    if (skip) {
      // verify the row-skipping SAD against the full SAD; redo the whole search without
      // skipping when they disagree by >= 90 % (mcomp.c:1840-1873)
      const int sad = (int)cta_value(k, [&](int ln) { return warp_sad(k, best->row, best->col, false, ln); });
      const int ssad = (int)cta_value(k, [&](int ln) { return warp_sad(k, best->row, best->col, true, ln); });
      if (sad > (1 << mi_log2) && abs(ssad - sad) * 10 >= max(sad, 1) * 9) {
        skip = false;
        continue;
      }
    }
    if (run_mesh) {
      Mv tmp;
      const int var_ex = full_pixel_exhaustive(k, skip, *best, cl, &tmp, second, true);
      if (var_ex < var) {
        var = var_ex;
        *best = tmp;
      }
    }
    break;
  }
  return var;
}

// pitch in words that is 4 (mod 8): conflict-free uint4 access by 8 consecutive rows
__host__ __device__ __forceinline__ int vec_pitch(int words) { return ((words + 3) & ~7) + 4 >= words ? ((words + 3) & ~7) + 4 : ((words + 11) & ~7) + 4; }

// Stage the source block and (when it fits) the reachable reference window; fills in the
// pixel views of `k`.  Layout for the vector SAD path: row pitches of 4 (mod 8) words and
// 16-byte aligned row starts (needs a reference stride that is a multiple of 16 bytes).
__device__ void stage_inputs(Ctx &k, unsigned char *smem, int window_budget, int extra) {
  const B200MeLaunch &l = *k.l;
  const int px = k.px, bw = k.d.bw, bh = k.d.bh;
  unsigned char *s_src = smem;
  const unsigned char *g_src = static_cast<const unsigned char *>(l.cur) + (size_t)k.d.src_offset * px;
  const int row_bytes = bw * px, row_words = row_bytes / 4;
  k.src_pitch_w = row_words >= 4 ? vec_pitch(row_words) : row_words;
  for (int i = threadIdx.x; i < bh * row_bytes; i += kThreads) {
    const int y = i / row_bytes, x = i - y * row_bytes;
    s_src[y * k.src_pitch_w * 4 + x] = g_src[(size_t)y * l.cur_stride * px + x];
  }
  k.src = s_src;
  const int src_bytes = (bh * k.src_pitch_w * 4 + 15) & ~15;
  const unsigned char *g_ref0 = static_cast<const unsigned char *>(l.ref) + (size_t)k.d.ref_offset * px;
  k.ref0 = g_ref0;
  k.ref_pitch = l.ref_stride;
  k.fast = false;
  // window = limits box (+ `extra` pixels on every side) + block
  const int wr0 = k.d.row_min - extra, wc0 = k.d.col_min - extra;
  const int wh = k.d.row_max - k.d.row_min + 2 * extra + bh;
  const int ww = k.d.col_max - k.d.col_min + 2 * extra + bw;
  const long long stride_b = (long long)l.ref_stride * px;
  const unsigned char *g0 = g_ref0 + ((long long)wr0 * l.ref_stride + wc0) * px;
  const bool vec = (stride_b & 15) == 0;
  const int lead = (int)(reinterpret_cast<unsigned long long>(g0) & (vec ? 15 : 3));
  // + 4 words: the vector path reads a 16-byte aligned superset of each candidate segment
  const int words = vec ? vec_pitch((lead + ww * px + 3) / 4 + 4) : (lead + ww * px + 3) / 4 + 1;
  const long long need = (long long)wh * words * 4;
  if ((stride_b & 3) == 0 && need <= window_budget && wh > 0 && ww > 0) {
    unsigned *s_win = reinterpret_cast<unsigned *>(smem + src_bytes);
    const unsigned *ga = reinterpret_cast<const unsigned *>(g0 - lead);
    const int copy_words = (lead + ww * px + 3) / 4 + 1;
    for (int i = threadIdx.x; i < wh * copy_words; i += kThreads) {
      const int y = i / copy_words, x = i - y * copy_words;
      s_win[y * words + x] = ga[(long long)y * (stride_b >> 2) + x];
    }
    k.ref_pitch = words * 4 / px;
    k.ref0 = reinterpret_cast<unsigned char *>(s_win) + lead +
             ((long long)(-wr0) * k.ref_pitch + (-wc0)) * px;
    k.fast = vec;
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kThreads, 3) fullpel_kernel(const B200MeLaunch l, const Sites *sites4,
                                                           int window_budget) {
  extern __shared__ __align__(16) unsigned char me_smem[];
  __shared__ unsigned s_cand[kMaxCand + 16];
  __shared__ int s_bcast[16];
  __shared__ long long s_red[2 * kWarps];
  const int bi = blockIdx.x;
  Ctx k;
  k.red = s_red;
  k.l = &l;
  k.d = l.blocks[bi];
  k.sites = sites4 + (k.d.search_method & 3);
  k.px = l.highbd ? 2 : 1;
  k.sad_shift = l.highbd ? (l.bd == 10 ? 2 : (l.bd == 12 ? 4 : 0)) : 0;
  k.cand = s_cand;
  k.bcast = s_bcast;
#ifdef B200_ME_TIMING
  const long long t_start = clock64();
#endif
  stage_inputs(k, me_smem, window_budget, 0);
#ifdef B200_ME_TIMING
  const long long t_staged = clock64();
#endif
  Mv best, second;
  int cl[5];
  const int var = full_pixel_search(k, (k.d.flags & B200_ME_SKIP_SAD) != 0, cl, &best, &second);
  if (threadIdx.x == 0) {
    B200MeResult r;
    r.best_row = (short)best.row;
    r.best_col = (short)best.col;
    r.second_row = (short)second.row;
    r.second_col = (short)second.col;
    for (int i = 0; i < 5; ++i) r.cost_list[i] = cl[i];
    r.ret = var;
    r.distortion = 0;
    r.sse = 0;
#ifdef B200_ME_TIMING
    r.distortion = k.bcast[12];                          // diamond phase, cycles / 16
    r.sse = (unsigned)((clock64() - t_start) >> 4);      // whole search
    r.cost_list[4] = (int)((t_staged - t_start) >> 4);   // staging
#endif
    l.results[bi] = r;
  }
}

}  // namespace me
}  // namespace b200
