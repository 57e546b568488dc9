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
    if (lane == 0) g_bcast[0] = (int)v;
  }
  __syncthreads();
  return (unsigned)g_bcast[0];
}

// get_mvpred_var_cost (mcomp.c:665): vf(src, ref) + sub-pel-domain mv cost.  The whole CTA
// accumulates (s - r) and (s - r)^2, partial sums meet in shared memory.
__device__ int var_cost(const Ctx &k, Mv m) {
  const int bw = k.d.bw, n = bw * k.d.bh;
  const unsigned char *rp = k.ref0 + ((long long)m.row * k.ref_pitch + m.col) * k.px;
  long long sum = 0;
  unsigned long long sse = 0;
  if (k.fast && k.px == 1 && bw >= 16) {
    int lsum;
    unsigned lsse;
    var8_staged(k.win0_off + m.row * k.ref_pitch + m.col, k.src_pitch_w * 4, k.ref_pitch, bw, k.d.bh,
                kWarps, &lsum, &lsse);
    sum = lsum;
    sse = lsse;
  } else
  if (k.px == 2 && (bw & 1) == 0) {
    // uint16 planes: a thread takes pixel PAIRS (block widths are powers of two: shifts, 32-bit loads where
    // the pair is word aligned -- the source copy always is, the reference when its column is even)
    const int lw = 31 - __clz(bw >> 1), np2 = n >> 1;
    const unsigned short *sp16 = reinterpret_cast<const unsigned short *>(k.src);
    const unsigned short *rp16 = reinterpret_cast<const unsigned short *>(rp);
    const bool ral = ((reinterpret_cast<size_t>(rp16) | (size_t)(k.ref_pitch * 2)) & 3) == 0;
    for (int i = threadIdx.x; i < np2; i += kThreads) {
      const int y = i >> lw, x = (i - (y << lw)) * 2;
      const unsigned s2 = *reinterpret_cast<const unsigned *>(sp16 + y * k.src_pitch_w * 2 + x);
      const unsigned short *rq = rp16 + (long long)y * k.ref_pitch + x;
      const unsigned r2 = ral ? *reinterpret_cast<const unsigned *>(rq) : ((unsigned)rq[0] | ((unsigned)rq[1] << 16));
      const int d0 = (int)(s2 & 0xffff) - (int)(r2 & 0xffff), d1 = (int)(s2 >> 16) - (int)(r2 >> 16);
      sum += d0 + d1;
      sse += (unsigned)(d0 * d0) + (unsigned)(d1 * d1);
    }
  } else
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
  unsigned sse32;
  const unsigned v = finish_var(tsum, tsse, n, k.l->highbd != 0, k.l->bd, &sse32);
  return (int)v + err_cost(k, m.row * 8, m.col * 8);
}

// diamond_search_sad (mcomp.c:1318-1477), ALL runs of full_pixel_diamond side by side.
// full_pixel_diamond (:1479) restarts the n-step search from the same start with a smaller
// first radius up to `further` times; run n + 1 does not depend on run n's result, only on
// whether it is executed at all (num00 skips).  One run per step leaves the CTA waiting on
// two barriers for ~10 SADs (the phase was latency bound: 54 % of the kernel's time for 24 %
// of its SAD work), so here every run advances one step per ROUND: run r's candidates are
// published by warp r % 8 into slots [16 r, 16 r + n), every warp adds its rows of its share
// of ALL runs' candidates into g_cand (eval_candidates), barrier, the owning warp decides its
// runs -- lane i owns candidate i; the reference's in-order strict-< scan = the lexicographic
// minimum of (sad + cost, index) that beats the running best -- moves the centre, picks the
// run's next step and publishes it, barrier.  Runs the sequential code would have skipped are
// computed and then ignored by the replay in full_pixel_diamond_t (about a third more SADs,
// for a quarter of the barriers).  `start_sad`: SAD + cost of the clamped start mv, the same for
// every run.  Results: g_runres[r].  The slots rotate through three groups, cleared two
// rounds ahead.
template <int SEGW>
__device__ void diamond_runs(const Ctx &k, bool skip, Mv start, unsigned start_sad, int step_param,
                             int nruns, const LaneShare<SEGW> &ls) {
  constexpr int kSlots = kMaxRuns / kWarps;  // runs owned by a warp
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const Sites &cfg = g_sites;  // file-scope shared: LDS, not generic loads
  start = clamp_mv(k, start);
  // the owning warp's state of its runs
  Mv cur[kSlots], sec[kSlots];
  unsigned bestsad[kSlots];
  int off_center[kSlots], center_steps[kSlots], step[kSlots];
#pragma unroll
  for (int s = 0; s < kSlots; ++s) {
    const int r = warp + s * kWarps;
    cur[s] = start, sec[s] = start;
    bestsad[s] = start_sad;
    off_center[s] = center_steps[s] = 0;
    step[s] = r < nruns ? cfg.n_steps - (step_param + r) - 1 : -1;
  }
  __syncthreads();  // g_cand / g_cda may still be read by the previous stage
  for (int i = threadIdx.x; i < 3 * kMaxRuns * 16; i += kThreads) g_cand[i] = 0;
  auto publish = [&](int s) {  // the candidates of run (warp, s)'s step around its centre, or its results
    const int r = warp + s * kWarps;
    if (r >= nruns) {
      if (lane == 0 && r < kMaxRuns) g_runn[r] = 0;
      return;
    }
    if (step[s] >= 0) {
      const int st = step[s], n = cfg.per_step[st];
      const Mv c = cur[s];
      publish_candidates<SEGW>(k, n, [&](int i) {
        return Mv{ c.row + cfg.site[st][i + 1][0], c.col + cfg.site[st][i + 1][1] };
      }, r * 16);
      if (lane == 0) g_runn[r] = n, atomicAdd(&g_bcast[24], n);
    } else if (lane == 0) {
      g_runn[r] = 0;
      g_runres[r][0] = cur[s].row, g_runres[r][1] = cur[s].col;
      g_runres[r][2] = sec[s].row, g_runres[r][3] = sec[s].col;
      g_runres[r][4] = (int)bestsad[s], g_runres[r][5] = center_steps[s], g_runres[r][6] = off_center[s];
    }
  };
  // Run r > 0 starts one round per run later than run 0 (it has that many fewer steps, so the
  // rounds do not grow): by then warp 0 knows whether run 0 is still on its centre after r
  // steps -- if so num00 >= r and the reference never executes run r -- and either publishes
  // run r's first candidates on its owner's behalf or leaves it out.
  if (threadIdx.x < kMaxRuns) g_runn[threadIdx.x] = 0;
  __syncthreads();
  if (warp == 0) publish(0);
  __syncthreads();
  int ring = 0;
#ifdef B200_ME_TIMING3
  long long t3_eval = 0, t3_ctrl = 0, t3_a, t3_b;
  int t3_rounds = 0;
#endif
  for (int round = 0;; ++round) {
    if (round >= nruns && !__any_sync(0xffffffffu, lane < nruns && g_runn[lane] > 0)) break;
#ifdef B200_ME_TIMING3
    t3_a = clock64();
#endif
    unsigned *cand = g_cand + ring * (kMaxRuns * 16);
    for (int r = 0; r < nruns; ++r) {
      const int n = g_runn[r];
#ifndef EXP_NOEVAL
      if (n) eval_candidates<SEGW>(k, skip, n, cand + r * 16, ls, r * 16);
#endif
    }
    __syncthreads();
#ifdef B200_ME_TIMING3
    t3_b = clock64();
    t3_eval += t3_b - t3_a;
#endif
#pragma unroll
    for (int s = 0; s < kSlots; ++s) {
      const int r = warp + s * kWarps;
      if (r >= nruns || step[s] < 0 || r > round) continue;  // warp-uniform
      if (r == round && r > 0 && g_runn[r] == 0) {  // left out by warp 0
        step[s] = -1;
        continue;
      }
      const int st = step[s], n = cfg.per_step[st];
#ifdef B200_ME_TIMING
      if (lane == 0) atomicAdd(&g_bcast[11], 1 + (n << 8));  // steps | candidates << 8
#endif
      unsigned t = 0xffffffffu;
      if (lane < n && g_cda[r * 16 + lane] >= 0) {
        const int p = g_cdp[r * 16 + lane];
        t = cand[r * 16 + lane] + (unsigned)sad_cost(k, (short)(p & 0xffff), p >> 16);
      }
      const unsigned tmin = __reduce_min_sync(0xffffffffu, t);
      const bool moved = tmin < bestsad[s];
      if (moved) {
        const int site = 1 + (int)__reduce_min_sync(0xffffffffu, t == tmin ? (unsigned)lane : 99u);
        bestsad[s] = tmin;
        sec[s] = cur[s];
        cur[s].row += cfg.site[st][site][0];
        cur[s].col += cfg.site[st][site][1];
        off_center[s] = 1;
      }
      if (!off_center[s]) center_steps[s]++;
      int ns = st;
      if (!moved && ns > 2) {
        int next = cfg.radius[ns - 1];
        while (next == cfg.radius[ns] && ns > 2) {
          center_steps[s]++;
          --ns;
          next = cfg.radius[ns - 1];
        }
      }
      step[s] = ns - 1;
      if (lane < 16) g_cand[((ring + 2) % 3) * (kMaxRuns * 16) + r * 16 + lane] = 0;
      publish(s);
    }
    if (warp == 0 && round + 1 < nruns && off_center[0]) {
      // run 0 has left its centre within round + 1 steps: run round + 1 may be needed
      const int r1 = round + 1, st1 = cfg.n_steps - (step_param + r1) - 1, n1 = cfg.per_step[st1];
      publish_candidates<SEGW>(k, n1, [&](int i) {
        return Mv{ start.row + cfg.site[st1][i + 1][0], start.col + cfg.site[st1][i + 1][1] };
      }, r1 * 16);
      if (lane == 0) g_runn[r1] = n1, atomicAdd(&g_bcast[24], n1);
    }
    ring = (ring + 1) % 3;
    __syncthreads();
#ifdef B200_ME_TIMING3
    t3_ctrl += clock64() - t3_b;
    ++t3_rounds;
#endif
  }
#ifdef B200_ME_TIMING3
  if (threadIdx.x == 0) g_bcast[13] = (int)(t3_eval >> 4), g_bcast[15] = (int)(t3_ctrl >> 4), g_bcast[14] = t3_rounds;
#endif
}

// ---- the same runs, one WARP GROUP per run (staged 8-bit windows, SEGW > 0) -------------
// diamond_runs keeps the whole CTA in lock step: two CTA-wide barriers per round, one warp deciding
// while seven wait -- measured (tools/me_phase_timing.py) the n-step phase took 58 % of a 64x64
// search for a quarter of its SADs, 2.8x its shared-memory / ALU floor.  Here the np warps that
// together cover one candidate block (LaneShare: np = passes) form a group that owns a run from
// its first step to its last: per step the lanes derive the step's candidates in registers
// (lane i = candidate i, fetched by shuffle: nothing is published), every warp adds its rows of
// each candidate, the group meets at ITS named barrier (a __syncwarp when np == 1), every warp
// sums the np partial totals and replays the decision redundantly -- no CTA-wide barrier, no
// atomics, no idle warps; the kWarps / np groups run different runs side by side (run r on group
// r mod G, then r + G, ...).  Run r > 0 is wanted only if run 0 left its centre within its first r
// steps (else num00 >= r and the reference never executes it): group 0 publishes run 0's
// progress in g_bcast[25] = steps done << 16 | step at which it left the centre (0xffff: not
// yet), the other groups poll it before they start a run.  Results: g_runres[r], as diamond_runs.
__device__ __forceinline__ void group_sync(int id, int nthreads) {
  if (nthreads <= 32) __syncwarp();
  else asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

template <int SEGW, int PX>
__device__ void diamond_runs_grouped(const Ctx &k, bool skip, Mv start, unsigned start_sad, int step_param,
                                     int nruns, const LaneShare<SEGW> &ls) {
  static_assert(SEGW > 0, "register-resident lane share only");
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const Sites &cfg = g_sites;
  start = clamp_mv(k, start);
  const int G = ls.cstep, np = kWarps / G;  // groups, warps per group (= passes)
  const int grp = ls.c0, pass = warp & (np - 1);
  const bool leader = pass == 0 && lane == 0;
  unsigned *part = g_cand;  // [2][kWarps][16] partial totals of a step
  volatile int *progress = &g_bcast[25];
  const unsigned smem_base = (unsigned)__cvta_generic_to_shared(me_smem);
  __syncthreads();  // g_cand may still be read by the previous stage
  if (threadIdx.x == 0) *progress = 0xffff;
  __syncthreads();
  // runs -> groups: longest run first onto the least loaded group (run r has n_steps - step_param - r steps),
  // lane g < G keeps group g's load and run set; every warp derives the same table
  unsigned my_runs;
  {
    unsigned load = lane < G ? 0u : 0xfffffu, mask = 0;
    for (int r = 0; r < nruns; ++r) {
      const int g = (int)(__reduce_min_sync(0xffffffffu, (load << 5) | (unsigned)lane) & 31u);
      if (lane == g) load += (unsigned)max(cfg.n_steps - (step_param + r), 1), mask |= 1u << r;
    }
    my_runs = __shfl_sync(0xffffffffu, mask, grp);
  }
  int buf = 0;
  while (my_runs) {
    const int r = __ffs(my_runs) - 1;
    my_runs &= my_runs - 1;
    if (r > 0) {
      // group 0 finished run 0 before it gets here; the others wait until run 0 has done r steps
      int v = *progress;
      while ((v & 0xffff) > r && (v >> 16) < r) {
        __nanosleep(64);
        v = *progress;
      }
      if ((v & 0xffff) > r) {  // still on its centre after r steps: never executed
        if (leader) g_runres[r][5] = 0, g_runres[r][6] = 0;
        continue;
      }
    }
    Mv cur = start, sec = start;
    unsigned bestsad = start_sad;
    int off_center = 0, center_steps = 0, steps_done = 0, left_at = 0xffff, n_cand = 0;
    int step = cfg.n_steps - (step_param + r) - 1;
    while (step >= 0) {
      const int st = step, n = cfg.per_step[st];
      // lane i < n: candidate i of the step around the run's centre
      int a = -1, prow = 0, pcol = 0;
      if (lane < n) {
        prow = cur.row + cfg.site[st][lane + 1][0];
        pcol = cur.col + cfg.site[st][lane + 1][1];
        if (in_range(k, prow, pcol)) a = k.win0_off + (prow * k.ref_pitch + pcol) * PX;
      }
      // candidate j's address comes from lane j one iteration ahead; its lane sums are reduced while
      // candidate j + 1's window rows are being loaded
      unsigned mine = 0, pend = 0;
      int pend_j = -1;
      int aj = __shfl_sync(0xffffffffu, a, 0);
      for (int j = 0; j < n; ++j) {
        const int an = __shfl_sync(0xffffffffu, a, j + 1);
        if (aj >= 0) {  // warp-uniform
          const unsigned saddr = smem_base + (unsigned)((aj & ~15) + ls.l_off);
          const unsigned sh = (unsigned)(aj & 3) * 8;
          unsigned acc, tot;
          switch ((aj >> 2) & 3) {
            case 0: acc = seg_sad_regs_p<SEGW, 0, PX>(ls.sv, saddr, sh, pend, tot); break;
            case 1: acc = seg_sad_regs_p<SEGW, 1, PX>(ls.sv, saddr, sh, pend, tot); break;
            case 2: acc = seg_sad_regs_p<SEGW, 2, PX>(ls.sv, saddr, sh, pend, tot); break;
            default: acc = seg_sad_regs_p<SEGW, 3, PX>(ls.sv, saddr, sh, pend, tot); break;
          }
          if (lane == pend_j) mine = tot;
          pend = ls.act ? acc : 0u;
          pend_j = j;
        }
        aj = an;
      }
      {
        const unsigned tot = warp_sum_u(pend);
        if (lane == pend_j) mine = tot;
      }
      unsigned total = mine;
      if (np > 1) {
        if (lane < 16) part[(buf * kWarps + warp) * 16 + lane] = mine;
        group_sync(1 + grp, np * 32);
        total = 0;
        if (lane < 16)
          for (int p = 0; p < np; ++p) total += part[(buf * kWarps + grp * np + p) * 16 + lane];
        buf ^= 1;  // the next step writes the other half: a slower warp of the group may still be reading
      }
      if (skip) total *= 2;
      if constexpr (PX == 2) total >>= k.sad_shift;  // aom_highbd_sadMxN_bits10 / _bits12 (encoder_utils.h:153-208)
      n_cand += n;
#ifdef B200_ME_TIMING
      if (leader) atomicAdd(&g_bcast[11], 1 + (n << 8));  // steps | candidates << 8
#endif
      const unsigned t = (lane < n && a >= 0) ? total + (unsigned)sad_cost(k, prow, pcol) : 0xffffffffu;
      const unsigned tmin = __reduce_min_sync(0xffffffffu, t);
      const bool moved = tmin < bestsad;
      if (moved) {
        const int site = 1 + (int)__reduce_min_sync(0xffffffffu, t == tmin ? (unsigned)lane : 99u);
        bestsad = tmin;
        sec = cur;
        cur.row += cfg.site[st][site][0];
        cur.col += cfg.site[st][site][1];
        off_center = 1;
      }
      if (!off_center) center_steps++;
      int ns = st;
      if (!moved && ns > 2) {
        int next = cfg.radius[ns - 1];
        while (next == cfg.radius[ns] && ns > 2) {
          center_steps++;
          --ns;
          next = cfg.radius[ns - 1];
        }
      }
      step = ns - 1;
      ++steps_done;
      if (r == 0) {
        if (off_center && left_at == 0xffff) left_at = steps_done;
        if (leader) {
          __threadfence_block();
          *progress = ((step >= 0 ? steps_done : 0x7fff) << 16) | left_at;
        }
      }
    }
    if (r == 0 && steps_done == 0 && leader) *progress = (0x7fff << 16) | 0xffff;  // run 0 had no step at all
    // get_mvpred_var_cost (mcomp.c:665) of the run's result, which the replay of full_pixel_diamond compares
    // between runs: measured here by the group (its rows dealt out to the np warps) instead of by the whole
    // CTA once the runs are over
    // (uint16 planes: the replay measures the few distinct results with the CTA-wide var_cost instead)
    int lsum = 0;
    unsigned lsse = 0;
    if constexpr (PX == 1)
      var8_staged(k.win0_off + cur.row * k.ref_pitch + cur.col, k.src_pitch_w * 4, k.ref_pitch, k.d.bw, k.d.bh, np,
                  &lsum, &lsse, pass);
    long long vsum = warp_sum_ll((long long)lsum);
    unsigned long long vsse = (unsigned long long)warp_sum_ll((long long)lsse);
    if (PX == 1 && np > 1) {
      if (lane == 0) g_red[warp] = vsum, g_red[kWarps + warp] = (long long)vsse;
      group_sync(1 + grp, np * 32);
      vsum = 0, vsse = 0;
      for (int p = 0; p < np; ++p)
        vsum += g_red[grp * np + p], vsse += (unsigned long long)g_red[kWarps + grp * np + p];
    }
    if (leader) {
      unsigned sse32;
      const unsigned v = finish_var(vsum, vsse, k.d.bw * k.d.bh, false, 8, &sse32);
      g_runres[r][0] = cur.row, g_runres[r][1] = cur.col;
      g_runres[r][2] = sec.row, g_runres[r][3] = sec.col;
      g_runres[r][4] = (int)bestsad, g_runres[r][5] = center_steps, g_runres[r][6] = off_center;
      g_runres[r][7] = (int)v + err_cost(k, cur.row * 8, cur.col * 8);
      atomicAdd(&g_bcast[24], n_cand);
    }
  }
  __syncthreads();
}

// calc_int_sad_list (mcomp.c:791) with costlist_has_sad == 0
__device__ void sad_list(const Ctx &k, bool skip, Mv best, int *cl) {
  // centre first, then left, bottom, right, top
  auto pos = [&](int i) {
    const int dr = i == 2 ? 1 : (i == 4 ? -1 : 0), dc = i == 1 ? -1 : (i == 3 ? 1 : 0);
    return Mv{ best.row + dr, best.col + dc };
  };
  const int segw = multi_segw(k);  // CTA-uniform
  __syncthreads();
  if (threadIdx.x < 8) g_cand[48 + threadIdx.x] = 0;
  if (threadIdx.x < 32) {
    if (segw) publish_candidates<16>(k, 5, pos);  // any SEGW > 0 publishes window addresses
    else publish_candidates<0>(k, 5, pos);
  }
  __syncthreads();
  switch (segw) {
    case 16: { LaneShare<16> ls; lane_share_init(k, skip, ls); eval_candidates<16>(k, skip, 5, g_cand + 48, ls); } break;
    case 8: { LaneShare<8> ls; lane_share_init(k, skip, ls); eval_candidates<8>(k, skip, 5, g_cand + 48, ls); } break;
    case 4: { LaneShare<4> ls; lane_share_init(k, skip, ls); eval_candidates<4>(k, skip, 5, g_cand + 48, ls); } break;
    default: eval_candidates<0>(k, skip, 5, g_cand + 48, LaneShare<0>()); break;
  }
  __syncthreads();
  for (int i = 0; i < 5; ++i) {
    const Mv m = pos(i);
    cl[i] = in_range(k, m.row, m.col) ? (int)g_cand[48 + i] + sad_cost(k, m.row, m.col) : INT_MAX;
  }
}

// full_pixel_diamond (mcomp.c:1479-1526): all runs at once (diamond_runs), then the reference's
// sequential bookkeeping replayed on their results -- which runs count (num00 skips), the
// variance comparison between them, and second_best_mv handed from run to run (a run that
// never moved leaves it as it found it).
template <int SEGW, int PX = 1>
__device__ int full_pixel_diamond_t(const Ctx &k, bool skip, Mv start, int step_param, int *cl,
                                    Mv *best, Mv *second, bool has_second) {
  LaneShare<SEGW> ls;
  if constexpr (SEGW > 0) lane_share_init(k, skip, ls);
  const Mv cs = clamp_mv(k, start);
  const unsigned start_sad = cta_value(k, [&](int ln) { return warp_sad(k, cs.row, cs.col, skip, ln); }) +
                             (unsigned)sad_cost(k, cs.row, cs.col);
  const int further = k.sites->n_steps - 1 - step_param;
  const int nruns = min(kMaxRuns, 1 + max(further, 0));
#ifdef B200_ME_TIMING2
  const long long tr0 = clock64();
#endif
  if constexpr (SEGW > 0) diamond_runs_grouped<SEGW, PX>(k, skip, start, start_sad, step_param, nruns, ls);
  else diamond_runs<SEGW>(k, skip, start, start_sad, step_param, nruns, ls);
#ifdef B200_ME_TIMING2
  const long long tr1 = clock64();
  if (threadIdx.x == 0) g_bcast[14] = (int)((tr1 - tr0) >> 4);
#endif
  auto take_second = [&](int r) {
    if (has_second && g_runres[r][6]) second->row = g_runres[r][2], second->col = g_runres[r][3];
  };
  int n = g_runres[0][5];
  best->row = g_runres[0][0], best->col = g_runres[0][1];
  take_second(0);
  // the grouped runs left their variance in g_runres[r][7]; the lock-step path measures it here
  constexpr bool kGroupVar = SEGW > 0 && PX == 1;
  int bestsme = kGroupVar ? g_runres[0][7] : var_cost(k, *best);
  // runs that end on an mv already measured (the usual case: they converge) reuse its variance
  Mv last = *best;
  int last_sme = bestsme;
  while (n < further) {
    ++n;
    const Mv tmp = { g_runres[n][0], g_runres[n][1] };
    take_second(n);
    int sme;
    if constexpr (kGroupVar) {
      sme = g_runres[n][7];
    } else if (tmp.row == best->row && tmp.col == best->col) {
      sme = bestsme;
    } else if (tmp.row == last.row && tmp.col == last.col) {
      sme = last_sme;
    } else {
      sme = var_cost(k, tmp);
      last = tmp, last_sme = sme;
    }
    if (sme < bestsme) {
      bestsme = sme;
      *best = tmp;
    }
    n += g_runres[n][5];
  }
#ifdef B200_ME_TIMING2
  const long long tr2 = clock64();
#endif
  if (cl) sad_list(k, skip, *best, cl);
#ifdef B200_ME_TIMING2
  if (threadIdx.x == 0) g_bcast[13] = (int)((tr2 - tr1) >> 4), g_bcast[15] = (int)((clock64() - tr2) >> 4);
#endif
  return bestsme;
}

// HBD: the kernel instance of uint16 planes (fullpel_kernel<3>); the 8-bit instance does not carry its code
template <bool HBD>
__device__ int full_pixel_diamond(const Ctx &k, bool skip, Mv start, int step_param, int *cl,
                                  Mv *best, Mv *second, bool has_second) {
  if constexpr (HBD) {
    switch (grouped_segw(k)) {  // CTA-uniform
      case 16: return full_pixel_diamond_t<16, 2>(k, skip, start, step_param, cl, best, second, has_second);
      case 8: return full_pixel_diamond_t<8, 2>(k, skip, start, step_param, cl, best, second, has_second);
      default: return full_pixel_diamond_t<0>(k, skip, start, step_param, cl, best, second, has_second);
    }
  } else {
    switch (multi_segw(k)) {
      case 16: return full_pixel_diamond_t<16>(k, skip, start, step_param, cl, best, second, has_second);
      case 8: return full_pixel_diamond_t<8>(k, skip, start, step_param, cl, best, second, has_second);
      case 4: return full_pixel_diamond_t<4>(k, skip, start, step_param, cl, best, second, has_second);
      default: return full_pixel_diamond_t<0>(k, skip, start, step_param, cl, best, second, has_second);
    }
  }
}

// ---- pattern_search family (mcomp.c:1017-1316) --------------------------------------
// Candidate j of scale s, generated by rule from the reference's tables (:498-660): scale s
// reaches 2^s pixels.  kind 0: hex (8-point ring at scale 0, then 6-point hexagons), 1:
// bigdia (4 nearest points at scale 0, then 8-point diamonds), 2: square.
__device__ __forceinline__ int pat_count(int kind, int s) {
  return kind == 2 ? 8 : (kind == 0 ? (s == 0 ? 8 : 6) : (s == 0 ? 4 : 8));
}
__constant__ signed char c_pat_ring[2][8] = { { -1, 0, 1, 1, 1, 0, -1, -1 }, { -1, -1, -1, 0, 1, 1, 1, 0 } };
__constant__ signed char c_pat_hex[2][6] = { { -1, 1, 2, 1, -1, -2 }, { -2, -2, 0, 2, 2, 0 } };
__constant__ signed char c_pat_dia[2][8] = { { -1, 0, 1, 2, 1, 0, -1, -2 }, { -1, -2, -1, 0, 1, 2, 1, 0 } };
__constant__ signed char c_pat_near[2][4] = { { 0, 1, 0, -1 }, { -1, 0, 1, 0 } };
__device__ __forceinline__ Mv pat_site(int kind, int s, int j) {
  const int d = 1 << s, h = d >> 1;
  if (kind == 2 || (kind == 0 && s == 0)) return Mv{ c_pat_ring[0][j] * d, c_pat_ring[1][j] * d };
  if (kind == 0) return Mv{ c_pat_hex[0][j] * h, c_pat_hex[1][j] * h };
  if (s == 0) return Mv{ c_pat_near[0][j], c_pat_near[1][j] };
  return Mv{ c_pat_dia[0][j] * h, c_pat_dia[1][j] * h };
}

struct PatState {
  int kind, br, bc, pp;
  unsigned bestsad, raw_bestsad;
};

// One round: candidates idx0, idx0 + 1, ... (n of them, or the three check points when
// chk != nullptr) of scale s around (br, bc), one warp each; then every thread replays
// the reference's in-order update (update_mvs_and_sad :858).  Returns the position of the
// last improvement or -1.  cl (optional): raw SADs at cl[index + 1]; `marks`: INT_MAX for
// points outside the limits (calc_sad_update_bestmv_with_indices :983).
__device__ int pat_round(const Ctx &k, PatState &st, bool skip, int s, const int *chk, int n, int *cl,
                         bool marks) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned *cand = g_cand + st.pp;
  st.pp ^= 32;
  if (warp < n) {
    const Mv o = pat_site(st.kind, s, chk ? chk[warp] : warp);
    const int r = st.br + o.row, c = st.bc + o.col;
    unsigned v = 0xffffffffu;
    if (in_range(k, r, c)) v = warp_sad(k, r, c, skip, lane);
    if (lane == 0) cand[warp] = v;
  }
  __syncthreads();
  int best = -1;
  for (int i = 0; i < n; ++i) {
    const int index = chk ? chk[i] : i;
    const unsigned v = cand[i];
    if (v == 0xffffffffu) {
      if (cl && marks) cl[index + 1] = INT_MAX;
      continue;
    }
    if (cl) cl[index + 1] = (int)v;
    if (v >= st.bestsad) continue;
    const Mv o = pat_site(st.kind, s, index);
    const unsigned t = v + (unsigned)sad_cost(k, st.br + o.row, st.bc + o.col);
    if (t < st.bestsad) {
      st.raw_bestsad = v;
      st.bestsad = t;
      best = i;
    }
  }
  return best;
}

// All candidates of a scale; inside the limits the reference only reaches the first
// n & ~3 of them (its scalar tail loop never runs, :1065-1078) -- kept.
__device__ __forceinline__ int pat_round_all(const Ctx &k, PatState &st, bool skip, int s) {
  const int range = 1 << s, n = pat_count(st.kind, s);
  const bool inside = st.br - range >= k.d.row_min && st.br + range <= k.d.row_max &&
                      st.bc - range >= k.d.col_min && st.bc + range <= k.d.col_max;
  return pat_round(k, st, skip, s, nullptr, inside ? n & ~3 : n, nullptr, false);
}

__device__ __forceinline__ void pat_move(PatState &st, int s, int j) {
  const Mv o = pat_site(st.kind, s, j);
  st.br += o.row;
  st.bc += o.col;
}

__device__ __noinline__ int pattern_search(const Ctx &k, bool skip, Mv start, int search_step, bool do_init, int *cl,
                              Mv *best_mv) {
  PatState st;
  const int m = k.d.search_method;
  st.kind = (m == B200_ME_HEX || m == B200_ME_FAST_HEX) ? 0 : (m == B200_ME_SQUARE ? 2 : 1);
  st.pp = 0;
  const bool last_is_4 = st.kind == 1;
  search_step = min(search_step, kMaxSteps - 1);
  int best_init_s = kMaxSteps - 1 - search_step, s, kk = -1;
  start = clamp_mv(k, start);
  st.br = start.row;
  st.bc = start.col;
  if (cl)
    for (int i = 0; i < 5; ++i) cl[i] = INT_MAX;
  bool has_sad = false;
  st.raw_bestsad = cta_value(k, [&](int ln) { return warp_sad(k, start.row, start.col, skip, ln); });
  st.bestsad = st.raw_bestsad + (unsigned)sad_cost(k, start.row, start.col);
  if (do_init) {
    s = best_init_s;
    best_init_s = -1;
    for (int t = 0; t <= s; ++t) {
      const int bs = pat_round_all(k, st, skip, t);
      if (bs != -1) {
        best_init_s = t;
        kk = bs;
      }
    }
    if (best_init_s != -1) pat_move(st, best_init_s, kk);
  }
  if (best_init_s != -1) {
    const int last_s = last_is_4 && cl != nullptr;
    int best_site = -1;
    for (s = best_init_s; s >= last_s; s--) {
      const int n = pat_count(st.kind, s);
      if (!do_init || s != best_init_s) {
        best_site = pat_round_all(k, st, skip, s);
        if (best_site == -1) continue;
        pat_move(st, s, best_site);
        kk = best_site;
      }
      do {
        const int chk[3] = { kk == 0 ? n - 1 : kk - 1, kk, kk == n - 1 ? 0 : kk + 1 };
        best_site = pat_round(k, st, skip, s, chk, 3, nullptr, false);
        if (best_site != -1) {
          kk = chk[best_site];
          pat_move(st, s, kk);
        }
      } while (best_site != -1);
    }
    if (s == 0) {  // only with a cost list: the 4-point scale fills it on the way
      cl[0] = (int)st.raw_bestsad;
      has_sad = true;
      if (!do_init || s != best_init_s) {
        best_site = pat_round(k, st, skip, 0, nullptr, 4, cl, false);
        if (best_site != -1) {
          pat_move(st, 0, best_site);
          kk = best_site;
        }
      }
      while (best_site != -1) {
        const int chk[3] = { kk == 0 ? 3 : kk - 1, kk, kk == 3 ? 0 : kk + 1 };
        cl[1] = cl[2] = cl[3] = cl[4] = INT_MAX;
        cl[((kk + 2) & 3) + 1] = cl[0];
        cl[0] = (int)st.raw_bestsad;
        best_site = pat_round(k, st, skip, 0, chk, 3, cl, true);
        if (best_site != -1) {
          kk = chk[best_site];
          pat_move(st, 0, kk);
        }
      }
    }
  }
  best_mv->row = st.br;
  best_mv->col = st.bc;
  if (cl) {
    if (!has_sad) {
      sad_list(k, skip, *best_mv, cl);
    } else {  // calc_int_sad_list with costlist_has_sad: only the mv costs are added
      cl[0] += sad_cost(k, st.br, st.bc);
      if (cl[1] != INT_MAX) cl[1] += sad_cost(k, st.br, st.bc - 1);
      if (cl[2] != INT_MAX) cl[2] += sad_cost(k, st.br + 1, st.bc);
      if (cl[3] != INT_MAX) cl[3] += sad_cost(k, st.br, st.bc + 1);
      if (cl[4] != INT_MAX) cl[4] += sad_cost(k, st.br - 1, st.bc);
    }
  }
  return var_cost(k, *best_mv);
}

// ---- register-blocked SAD maps for the mesh stages -----------------------------------
// The generic path above spends one warp per candidate and re-reads source and reference
// from shared memory for every candidate (5 LDS.128 per 8 VABSDIFF4).  A mesh stage is a
// regular grid of candidates, so neighbouring candidates can share loaded words:
//   * step 1 (`mesh1_band`): a lane owns RB vertically adjacent candidates of one column
//     offset; it walks the reference rows once, funnel-shifts each loaded row once, and
//     compares it against a rotating window of RB source rows held in registers
//     (SEGW shifts + RB * SEGW VABSDIFF4.ACC per SEGW + 1 reference words loaded);
//   * step 4 / 8 / 16 (`mesh_strided_band`): candidates of one grid row are whole words
//     apart, so a lane loads one long aligned strip of a reference row, shifts it once and
//     slides the source row over it for JB horizontally adjacent candidates.
// Both write `SAD + mv cost` of every candidate of a band of grid rows into g_cand[] in
// scan order (shared-memory atomics join the 32-pixel column segments of a block).
// Preconditions (checked by the caller): 8-bit, window staged in shared memory, no row
// skipping, bw >= 4 * SEGW, bh a multiple of 4.
constexpr int kRB = 4;  // vertically adjacent candidates per lane (step 1)
constexpr int kJB = 4;  // horizontally adjacent candidates per lane (step 4 / 8 / 16), 8-bit

// One word of source against one word of reference.  PX = 1: four 8-bit pixels, one
// VABSDIFF4.U8.ACC into a 32-bit total.  PX = 2 (uint16 planes): two pixels; sm_100 has no packed
// 16-bit absolute difference (vabsdiff2 and the scalar half-word vabsdiff are emulated), but
// |a - b| = max(a, b) - min(a, b) per half word, both VIMNMX.U16x2, and `acc + max - min` is ONE
// IADD3 on the packed pair: no borrow crosses the halves because every half-word term is >= 0.
// The packed halves hold at most 16 terms of 12 bits (16 * 4095 < 65536) before `sad_flush`
// (IDP.2A) moves them into the 32-bit total.
template <int PX>
__device__ __forceinline__ unsigned sadw_acc(unsigned s, unsigned r, unsigned acc) {
  if constexpr (PX == 1) return sad4_acc(s, r, acc);
  else return acc + __vmaxu2(s, r) - __vminu2(s, r);
}
template <int PX>
__device__ __forceinline__ void sad_flush(unsigned &acc, unsigned &tot) {
  if constexpr (PX == 2) {
    tot = __dp2a_lo(acc, 0x0101u, tot);
    acc = 0;
  }
}

// PHASE 0: first RB reference rows (source window filling), 1: steady state, 2: last RB - 1
// rows (window draining).
template <int SEGW, int PHASE, int PX, int RB>
__device__ __forceinline__ void mesh1_chunk(int s_off, int src_pitch_b, int r_off, int pitch_b,
                                            unsigned sh, int t0, int tmax,
                                            unsigned (&sw)[RB][SEGW], unsigned (&acc)[RB]) {
#pragma unroll
  for (int u = 0; u < RB; ++u) {
    if (PHASE == 2 && u == RB - 1) break;
    const int t = t0 + u;
    if (PHASE == 2 && t >= tmax) break;  // warp-uniform
    if (PHASE != 2) load_words<SEGW>(smem_words(s_off + t * src_pitch_b), sw[u]);
    unsigned r[SEGW + 1], rv[SEGW];
    const unsigned *rp = smem_words(r_off + t * pitch_b);
#pragma unroll
    for (int i = 0; i <= SEGW; ++i) r[i] = rp[i];
#pragma unroll
    for (int i = 0; i < SEGW; ++i) rv[i] = __funnelshift_r(r[i], r[i + 1], sh);
#pragma unroll
    for (int c = 0; c < RB; ++c) {
      if (PHASE == 0 && c > u) continue;
      if (PHASE == 2 && c <= u) continue;
#pragma unroll
      for (int i = 0; i < SEGW; ++i) acc[c] = sadw_acc<PX>(sw[(u - c) & (RB - 1)][i], rv[i], acc[c]);
    }
  }
}

template <int SEGW, int PX, int RB>
__device__ void mesh1_band(const Ctx &k, int row0, int nrows, int col0, int nc) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int bh = k.d.bh, nseg = k.d.bw * PX / (4 * SEGW);
  const int nrg = (nrows + RB - 1) / RB, ncc = (nc + 31) >> 5;
  const int ntask = nrg * nseg * ncc;
  const int pitch_b = k.ref_pitch * PX, src_pitch_b = k.src_pitch_w * 4;
  for (int task = warp; task < ntask; task += kWarps) {
    const int cc = task % ncc, seg = (task / ncc) % nseg, rg = task / (ncc * nseg);
    const int nk = min(RB, nrows - rg * RB);
    const int ci = cc * 32 + lane, cia = min(ci, nc - 1);
    const int row = row0 + rg * RB, col = col0 + cia;
    const int a = k.win0_off + row * pitch_b + col * PX + seg * SEGW * 4;
    const int r_off = a & ~3, s_off = seg * SEGW * 4;
    const unsigned sh = (unsigned)(a & 3) * 8;
    unsigned sw[RB][SEGW], acc[RB], tot[RB];
#pragma unroll
    for (int c = 0; c < RB; ++c) acc[c] = 0, tot[c] = 0;
    const int tmax = bh + nk - 1;
    mesh1_chunk<SEGW, 0, PX, RB>(s_off, src_pitch_b, r_off, pitch_b, sh, 0, tmax, sw, acc);
    for (int t0 = RB; t0 < bh; t0 += RB) {
#pragma unroll
      for (int c = 0; c < RB; ++c) sad_flush<PX>(acc[c], tot[c]);
      mesh1_chunk<SEGW, 1, PX, RB>(s_off, src_pitch_b, r_off, pitch_b, sh, t0, tmax, sw, acc);
    }
#pragma unroll
    for (int c = 0; c < RB; ++c) sad_flush<PX>(acc[c], tot[c]);
    mesh1_chunk<SEGW, 2, PX, RB>(s_off, src_pitch_b, r_off, pitch_b, sh, bh, tmax, sw, acc);
#pragma unroll
    for (int c = 0; c < RB; ++c) sad_flush<PX>(acc[c], tot[c]);
    if (ci < nc) {
#pragma unroll
      for (int c = 0; c < RB; ++c)
        if (c < nk) {
          // PX = 2: raw partial sums; the caller shifts the block total (>> 2 / >> 4, the encoder's
          // high-bit-depth SAD wrappers) and adds the mv cost once all segments have arrived
          unsigned v = PX == 1 ? acc[c] : tot[c];
          if (PX == 1 && seg == 0) v += (unsigned)sad_cost(k, row + c, col);
          atomicAdd(&g_cand[(rg * RB + c) * nc + ci], v);
        }
    }
  }
}

template <int SEGW, int STEPW, int WO, int PX, int JB>
__device__ void mesh_strided_band_t(const Ctx &k, int row0, int nrows, int col0, int nc, int step,
                                    unsigned sh) {
  constexpr int NSH = SEGW + (JB - 1) * STEPW;  // shifted words needed
  constexpr int NV = (WO + NSH + 1 + 3) / 4;    // aligned uint4 loads covering them (+1 for the funnel)
  const int lane = threadIdx.x & 31, ylane = threadIdx.x & 7;
  const unsigned gmask = 0xffu << (lane & 24);
  const int bh = k.d.bh, nseg = k.d.bw * PX / (4 * SEGW);
  const int ncg = (nc + JB - 1) / JB;
  const int ntask = nrows * ncg * nseg;
  const int pitch_b = k.ref_pitch * PX, src_pitch_b = k.src_pitch_w * 4;
  for (int task = threadIdx.x >> 3; task < ntask; task += kThreads / 8) {
    const int seg = task % nseg, cg = (task / nseg) % ncg, ri = task / (nseg * ncg);
    const int row = row0 + ri * step, col = col0 + cg * JB * step;
    const int r_off = (k.win0_off + row * pitch_b + col * PX + seg * SEGW * 4) & ~15;
    const int s_off = seg * SEGW * 4;
    unsigned acc[JB], tot[JB];
#pragma unroll
    for (int j = 0; j < JB; ++j) acc[j] = 0, tot[j] = 0;
    for (int y = ylane; y < bh; y += 8) {
      unsigned s[SEGW], r[4 * NV], v[NSH];
      load_words<SEGW>(smem_words(s_off + y * src_pitch_b), s);
      load_words<4 * NV>(smem_words(r_off + y * pitch_b), r);
#pragma unroll
      for (int i = 0; i < NSH; ++i) v[i] = __funnelshift_r(r[WO + i], r[WO + i + 1], sh);
#pragma unroll
      for (int j = 0; j < JB; ++j) {
#pragma unroll
        for (int i = 0; i < SEGW; ++i) acc[j] = sadw_acc<PX>(s[i], v[j * STEPW + i], acc[j]);
        sad_flush<PX>(acc[j], tot[j]);  // SEGW <= 8 terms per half word and row
      }
    }
#pragma unroll
    for (int j = 0; j < JB; ++j) {
      if (PX == 2) acc[j] = tot[j];
      acc[j] += __shfl_xor_sync(gmask, acc[j], 1);
      acc[j] += __shfl_xor_sync(gmask, acc[j], 2);
      acc[j] += __shfl_xor_sync(gmask, acc[j], 4);
    }
    if (ylane == 0) {
#pragma unroll
      for (int j = 0; j < JB; ++j) {
        const int ci = cg * JB + j;
        if (ci < nc) {
          unsigned t = acc[j];
          if (PX == 1 && seg == 0) t += (unsigned)sad_cost(k, row, col + j * step);
          atomicAdd(&g_cand[ri * nc + ci], t);
        }
      }
    }
  }
}

template <int SEGW, int STEPW, int PX, int JB>
__device__ __forceinline__ void mesh_strided_band(const Ctx &k, int row0, int nrows, int col0, int nc,
                                                  int step) {
  // all candidates of a stage share the byte alignment (their columns differ by whole
  // multiples of 16 bytes within a lane group and of 4 bytes overall)
  const int a = k.win0_off + col0 * PX;
  const unsigned sh = (unsigned)(a & 3) * 8;
  // word offset inside the 16-byte chunk: candidates of one lane group start JB * step
  // pixels apart (a multiple of 16 bytes), grid rows start step * pitch apart (pitch % 16 == 0)
  switch ((a >> 2) & 3) {
    case 0: mesh_strided_band_t<SEGW, STEPW, 0, PX, JB>(k, row0, nrows, col0, nc, step, sh); break;
    case 1: mesh_strided_band_t<SEGW, STEPW, 1, PX, JB>(k, row0, nrows, col0, nc, step, sh); break;
    case 2: mesh_strided_band_t<SEGW, STEPW, 2, PX, JB>(k, row0, nrows, col0, nc, step, sh); break;
    default: mesh_strided_band_t<SEGW, STEPW, 3, PX, JB>(k, row0, nrows, col0, nc, step, sh); break;
  }
}

// 0: no fast path for this stage; otherwise fills g_cand[0..nrows*nc) (must be zeroed)
// uint16 planes: 2 pixels per word, candidates step * 2 bytes apart; two candidates per lane in the
// strided stages.  Not inlined: its register allocation stays out of the 8-bit paths of the kernel.
__device__ __noinline__ void mesh_band_fast_hbd(const Ctx &k, int row0, int nrows, int col0, int nc, int step) {
  if (step == 1) mesh1_band<4, 2, 4>(k, row0, nrows, col0, nc);
  else if (k.d.bw >= 16) {
    if (step == 4) mesh_strided_band<8, 2, 2, 2>(k, row0, nrows, col0, nc, step);
    else if (step == 8) mesh_strided_band<8, 4, 2, 2>(k, row0, nrows, col0, nc, step);
    else mesh_strided_band<8, 8, 2, 2>(k, row0, nrows, col0, nc, step);
  } else {  // 8 pixels = 4 words
    if (step == 4) mesh_strided_band<4, 2, 2, 2>(k, row0, nrows, col0, nc, step);
    else if (step == 8) mesh_strided_band<4, 4, 2, 2>(k, row0, nrows, col0, nc, step);
    else mesh_strided_band<4, 8, 2, 2>(k, row0, nrows, col0, nc, step);
  }
  // raw block sums -> encoder SAD (>> 2 / >> 4, encoder_utils.h:153-208) + mv cost
  __syncthreads();
  for (int j = threadIdx.x; j < nrows * nc; j += kThreads) {
    const int ri = j / nc, ci = j - ri * nc;
    g_cand[j] = (g_cand[j] >> k.sad_shift) + (unsigned)sad_cost(k, row0 + ri * step, col0 + ci * step);
  }
}

__device__ __forceinline__ bool mesh_band_fast(const Ctx &k, int row0, int nrows, int col0, int nc,
                                               int step) {
  const int bw = k.d.bw;
  if (k.px == 2) {
    mesh_band_fast_hbd(k, row0, nrows, col0, nc, step);
    return true;
  }
  if (step == 1) {
    // (8 vertically adjacent candidates per lane -- mesh1_band<4, 1, 8> -- halve the loads and funnel shifts
    // per accumulation but were slower on B200: 0.430 vs 0.409 ms per 1080p frame, the kernel's 80-register
    // budget spills the 32-word source window)
    if (bw >= 16) mesh1_band<4, 1, 4>(k, row0, nrows, col0, nc);
    else mesh1_band<2, 1, 4>(k, row0, nrows, col0, nc);
    return true;
  }
  if (bw >= 32) {
    if (step == 4) mesh_strided_band<8, 1, 1, kJB>(k, row0, nrows, col0, nc, step);
    else if (step == 8) mesh_strided_band<8, 2, 1, kJB>(k, row0, nrows, col0, nc, step);
    else mesh_strided_band<8, 4, 1, kJB>(k, row0, nrows, col0, nc, step);
  } else if (bw >= 16) {
    if (step == 4) mesh_strided_band<4, 1, 1, kJB>(k, row0, nrows, col0, nc, step);
    else if (step == 8) mesh_strided_band<4, 2, 1, kJB>(k, row0, nrows, col0, nc, step);
    else mesh_strided_band<4, 4, 1, kJB>(k, row0, nrows, col0, nc, step);
  } else {  // 8-pixel blocks (the temporal filter's sub-blocks): 2-word segments
    if (step == 4) mesh_strided_band<2, 1, 1, kJB>(k, row0, nrows, col0, nc, step);
    else if (step == 8) mesh_strided_band<2, 2, 1, kJB>(k, row0, nrows, col0, nc, step);
    else mesh_strided_band<2, 4, 1, kJB>(k, row0, nrows, col0, nc, step);
  }
  return true;
}

__device__ __forceinline__ bool mesh_fast_ok(const Ctx &k, bool skip, int step, int nc) {
  if (skip || !k.staged || (k.d.bh & 3) || nc > kMaxCand) return false;
  if (k.d.bw < 8) return false;
  if (step == 1) return true;  // 8 bytes = one 2-word segment (uint16: 16 bytes = one 4-word segment)
  // strided: 16-byte aligned window rows (vector loads), whole-word candidate spacing
  return k.fast && (step == 4 || step == 8 || step == 16);
}

// exhaustive_mesh_search (mcomp.c:1530-1599).  Candidates are visited row-major; at
// step 1 the reference works in groups of four columns and its tail loop stops one short
// of the last column (`i < end_col - c`, :1586) -- reproduced.
__device__ bool stage_window(Ctx &k, int wr0, int wc0, int wh, int ww);

__device__ int mesh_search(const Ctx &k_in, bool skip, Mv start, int range, int step, Mv *best,
                           Mv *second, bool has_second) {
  // A search whose mv limits are frame wide (temporal filter, TPL, first pass) cannot keep its whole
  // reachable window on chip, but one mesh stage only visits start +- range: stage that box here and run
  // the stage on the register-blocked maps instead of one warp per candidate through L1/L2 (ncu on the
  // temporal-filter queue: 3.6-9.1 GB of L2->L1 traffic per launch, 0.5-0.9 us per 8x8 / 16x16 search).
  Ctx km(k_in);
  if (!k_in.staged && k_in.win_budget > 0) {
    const Mv s0 = clamp_mv(k_in, start);
    const int br0 = max(-range, k_in.d.row_min - s0.row), bc0 = max(-range, k_in.d.col_min - s0.col);
    const int br1 = min(range, k_in.d.row_max - s0.row), bc1 = min(range, k_in.d.col_max - s0.col);
    if (br1 >= br0 && bc1 >= bc0)
      stage_window(km, s0.row + br0, s0.col + bc0, br1 - br0 + k_in.d.bh, bc1 - bc0 + k_in.d.bw);
    __syncthreads();
  }
  const Ctx &k = km;
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
  if (threadIdx.x == 0) g_bcast[24] += total + 1;  // measurement: SAD candidates of this search
  // chunks of the scan: whole grid rows ("bands") for the register-blocked SAD maps,
  // kMaxCand candidates otherwise.  Chunk boundaries do not influence the decisions.
  const bool fast = mesh_fast_ok(k, skip, step, nc);
  int band_rows = 0;
  if (fast) {
    band_rows = min(nr, kMaxCand / nc);
    if (band_rows > kRB) band_rows &= ~(kRB - 1);
  }
  const int chunk = fast ? band_rows * nc : kMaxCand;
  for (int base = 0; base < total; base += chunk) {
    const int cnt = min(chunk, total - base);
    if (fast) {
      for (int j = threadIdx.x; j < cnt; j += kThreads) g_cand[j] = 0;
      __syncthreads();
      const int ri0 = base / nc;
      mesh_band_fast(k, start.row + r0 + ri0 * step, cnt / nc, start.col + c0, nc, step);
    } else {
      for (int j = warp; j < cnt; j += kWarps) {
        const int idx = base + j;
        const int ri = idx / nc, ci = idx - ri * nc;
        const int r = start.row + r0 + ri * step, c = start.col + c0 + ci * step;
        const unsigned t = warp_sad(k, r, c, skip, lane) + (unsigned)sad_cost(k, r, c);
        if (lane == 0) g_cand[j] = t;
      }
    }
    __syncthreads();
    // Replay update_mvs_and_sad (:858) over the chunk without walking it serially: the
    // last improvement of the sequential scan is the FIRST occurrence of the chunk minimum
    // (if it beats the running best), and second_best is whatever was best just before it,
    // i.e. the first minimum of the prefix in front of the winner (or the old best).
    // Every thread owns kMaxCand / kThreads consecutive candidates; (value, index) pairs
    // are reduced lexicographically per warp (REDUX) and then over the warps' slots.
    {
      constexpr int PER = kMaxCand / kThreads;
      const int lo = threadIdx.x * PER;
      unsigned cv[PER];
#pragma unroll
      for (int j = 0; j < PER; ++j) cv[j] = lo + j < cnt ? g_cand[lo + j] : 0xffffffffu;
      int win = INT_MAX, pi = INT_MAX;
      unsigned wv = best_sad;
      for (int pass = 0; pass < 2; ++pass) {
        // pass 0: whole chunk; pass 1: the prefix in front of the winner
        const int lim = pass == 0 ? cnt : win;
        unsigned bv = best_sad;
        int bi = INT_MAX;
#pragma unroll
        for (int j = 0; j < PER; ++j)
          if (lo + j < lim && cv[j] < bv) bv = cv[j], bi = lo + j;
        const unsigned mv = __reduce_min_sync(0xffffffffu, bv);
        const int mi = (int)__reduce_min_sync(0xffffffffu, bv == mv ? (unsigned)bi : 0x7fffffffu);
        if (lane == 0) {
          g_red[warp] = (long long)(((unsigned long long)mv << 32) | (unsigned)mi);
        }
        __syncthreads();
        unsigned long long m = ~0ull;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
          const unsigned long long t = (unsigned long long)g_red[w];
          m = t < m ? t : m;
        }
        const int idx = (int)(unsigned)m;
        if (pass == 0) {
          win = idx;
          wv = (unsigned)(m >> 32);
          if (win == INT_MAX || !has_second) break;  // CTA-uniform
          __syncthreads();  // g_red is reused by the second pass
        } else {
          pi = idx;
        }
      }
      if (threadIdx.x == 0) {
        g_bcast[8] = win;
        g_bcast[9] = (int)wv;
        g_bcast[10] = pi;
      }
    }
    __syncthreads();
    {
      const int win = g_bcast[8], pi = g_bcast[10];
      if (win != INT_MAX) {
        best_sad = (unsigned)g_bcast[9];
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
#ifdef B200_ME_TIMING
  long long tm = clock64();
#endif
  int bestsme = mesh_search(k, skip, *best, range, interval, best, second, has_second);
#if defined(B200_ME_TIMING) && !defined(B200_ME_TIMING2) && !defined(B200_ME_TIMING3)
  g_bcast[13] = (int)((clock64() - tm) >> 4);
#endif
  if (interval > 1 && range > 7) {
    for (int i = 1; i < 4; ++i) {
#ifdef B200_ME_TIMING
      tm = clock64();
#endif
      bestsme = mesh_search(k, skip, *best, k.l->mesh_range[i], k.l->mesh_interval[i], best, second,
                            has_second);
#if defined(B200_ME_TIMING) && !defined(B200_ME_TIMING2) && !defined(B200_ME_TIMING3)
      if (i < 3) g_bcast[13 + i] = (int)((clock64() - tm) >> 4);
#endif
      if (k.l->mesh_interval[i] == 1) break;
    }
  }
  if (bestsme < INT_MAX) bestsme = var_cost(k, *best);
  if (cl) sad_list(k, skip, *best, cl);
  return bestsme;
}

__device__ __forceinline__ int ilog2i(int v) { return 31 - __clz(v); }

// av1_full_pixel_search (mcomp.c:1755-1895).  FAM 0: n-step / diamond family, FAM 1: the
// pattern_search family -- separate kernel instances, so that neither carries the other's
// registers (one kernel for everything spilled 450 bytes).
template <int FAM>
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
    {
      const int m = k.d.search_method, sp = k.d.step_param;
      int *pcl = (k.d.flags & B200_ME_COST_LIST) ? cl : nullptr;  // changes the BIGDIA family's last scale
      if constexpr (FAM == 1) {
        if (m == B200_ME_HEX || m == B200_ME_BIGDIA || m == B200_ME_SQUARE)
          var = pattern_search(k, skip, start, sp, true, pcl, best);
        else if (m == B200_ME_FAST_HEX || m == B200_ME_FAST_DIAMOND)
          var = pattern_search(k, skip, start, max(kMaxSteps - 2, sp), false, pcl, best);
        else if (m == B200_ME_FAST_BIGDIA)
          var = pattern_search(k, skip, start, max(kMaxSteps - 3, sp), false, pcl, best);
        else
          var = pattern_search(k, skip, start, max(kMaxSteps - 1, sp), false, pcl, best);
      } else {
        (void)m, (void)pcl;
        var = full_pixel_diamond<FAM == 3>(k, skip, start, sp, cl, best, second, true);
      }
    }
#ifdef B200_ME_TIMING
    g_bcast[12] = (int)((clock64() - tq0) >> 4);
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

// av1_refining_search_8p_c (mcomp.c:1683-1753), single reference: up to SEARCH_RANGE_8P = 3
// rounds over the 8 neighbours of the running best, each grid position evaluated once.
// One warp per neighbour; every thread replays the reference's in-order double test
// (raw SAD first, then SAD + cost).
__device__ int refine_8p(const Ctx &k, bool skip, Mv *bestp) {
  constexpr int kRange = 3, kStride = 2 * kRange + 1;
  const int nb[8][2] = { { -1, 0 }, { 0, -1 }, { 0, 1 }, { 1, 0 }, { -1, -1 }, { 1, -1 }, { -1, 1 }, { 1, 1 } };
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const Mv start = { k.d.start_row, k.d.start_col };
  Mv best = clamp_mv(k, start);
  unsigned best_sad = cta_value(k, [&](int ln) { return warp_sad_refine(k, best.row, best.col, skip, ln); });
  best_sad += (unsigned)sad_cost(k, best.row, best.col);
  int centre = kRange * kStride + kRange;
  unsigned long long visited = 1ull << centre;
  for (int i = 0; i < kRange; ++i) {
    __syncthreads();
    if (warp < 8) {
      const int g = centre + nb[warp][0] * kStride + nb[warp][1];
      const int r = best.row + nb[warp][0], c = best.col + nb[warp][1];
      unsigned v = 0xffffffffu;
      if (!((visited >> g) & 1) && in_range(k, r, c)) v = warp_sad_refine(k, r, c, skip, lane);
      if (lane == 0) g_cand[warp] = v;
    }
    __syncthreads();
    int best_site = -1;
    for (int j = 0; j < 8; ++j) {
      const int g = centre + nb[j][0] * kStride + nb[j][1];
      if ((visited >> g) & 1) continue;
      visited |= 1ull << g;
      const unsigned v = g_cand[j];
      if (v == 0xffffffffu || v >= best_sad) continue;
      const unsigned t = v + (unsigned)sad_cost(k, best.row + nb[j][0], best.col + nb[j][1]);
      if (t < best_sad) {
        best_sad = t;
        best_site = j;
      }
    }
    if (best_site < 0) break;
    best.row += nb[best_site][0];
    best.col += nb[best_site][1];
    centre += nb[best_site][0] * kStride + nb[best_site][1];
  }
  *bestp = best;
  return (int)best_sad;
}

// pitch in words that is 4 (mod 8): conflict-free uint4 access by 8 consecutive rows
__host__ __device__ __forceinline__ int vec_pitch(int words) { return ((words + 3) & ~7) + 4 >= words ? ((words + 3) & ~7) + 4 : ((words + 11) & ~7) + 4; }

// Cooperative copy of the reference window whose top-left pixel is at full-pel mv (wr0, wc0), wh x ww
// pixels, into the window region of shared memory (when it fits k.win_budget); on success the pixel views
// of `k` address the copy.  The caller synchronises the CTA afterwards.
__device__ bool stage_window(Ctx &k, int wr0, int wc0, int wh, int ww) {
  const B200MeLaunch &l = *k.l;
  const int px = k.px;
  const long long stride_b = (long long)l.ref_stride * px;
  const unsigned char *g_ref0 = static_cast<const unsigned char *>(l.ref) + (size_t)k.d.ref_offset * px;
  const unsigned char *g0 = g_ref0 + ((long long)wr0 * l.ref_stride + wc0) * px;
  const bool vec = (stride_b & 15) == 0;
  const int lead = (int)(reinterpret_cast<unsigned long long>(g0) & (vec ? 15 : 3));
  // + 4 words: the vector path reads a 16-byte aligned superset of each candidate segment
  const int words = vec ? vec_pitch((lead + ww * px + 3) / 4 + 4) : (lead + ww * px + 3) / 4 + 1;
  const long long need = (long long)wh * words * 4;
  if (!((stride_b & 3) == 0 && need <= k.win_budget && wh > 0 && ww > 0)) return false;
  unsigned *s_win = reinterpret_cast<unsigned *>(me_smem + k.win_region_off);
  const unsigned *ga = reinterpret_cast<const unsigned *>(g0 - lead);
  // exactly the words that hold window bytes: nothing is read past the window's right edge
  // (the funnel shifts also touch the following word; its value is shifted out)
  const int copy_words = (lead + ww * px + 3) / 4;
  for (int i = threadIdx.x; i < wh * (copy_words + 1); i += kThreads) {
    const int y = i / (copy_words + 1), x = i - y * (copy_words + 1);
    s_win[y * words + x] = x < copy_words ? ga[(long long)y * (stride_b >> 2) + x] : 0u;
  }
  k.ref_pitch = words * 4 / px;
  k.ref0 = reinterpret_cast<unsigned char *>(s_win) + lead + ((long long)(-wr0) * k.ref_pitch + (-wc0)) * px;
  k.fast = vec;
  k.staged = true;
  k.win0_off = k.win_region_off + lead + ((-wr0) * k.ref_pitch + (-wc0)) * px;
  return true;
}

// Stage the source block and (when it fits) the reachable reference window; fills in the
// pixel views of `k`.  Layout for the vector SAD path: row pitches of 4 (mod 8) words and
// 16-byte aligned row starts (needs a reference stride that is a multiple of 16 bytes).
__device__ void stage_inputs(Ctx &k, unsigned char *smem, int window_budget, int extra,
                             const TmaCfg &tc, const CUtensorMap *tm_cur, const CUtensorMap *tm_ref) {
  const B200MeLaunch &l = *k.l;
  const int px = k.px, bw = k.d.bw, bh = k.d.bh;
  unsigned char *s_src = smem;
  const unsigned char *g_src = static_cast<const unsigned char *>(l.cur) + (size_t)k.d.src_offset * px;
  const int row_bytes = bw * px, row_words = row_bytes / 4;
  const long long stride_b = (long long)l.ref_stride * px;
  // window = limits box (+ `extra` pixels on every side) + block
  const int wr0 = k.d.row_min - extra, wc0 = k.d.col_min - extra;
  const int wh = k.d.row_max - k.d.row_min + 2 * extra + bh;
  const int ww = k.d.col_max - k.d.col_min + 2 * extra + bw;
  const unsigned char *g_ref0 = static_cast<const unsigned char *>(l.ref) + (size_t)k.d.ref_offset * px;
  k.src = s_src;
  k.ref0 = g_ref0;
  k.ref_pitch = l.ref_stride;
  k.fast = false;
  k.staged = false;
  const long long src_b = (long long)k.d.src_offset * px, cur_stride_b = (long long)l.cur_stride * px;
  // byte position of the window's first pixel inside the reference buffer
  const long long win_b = ((long long)k.d.ref_offset + (long long)wr0 * l.ref_stride + wc0) * px;
  if (tc.use && row_words >= 4 && (src_b % cur_stride_b) % 16 == 0 && bh <= tc.src_rows &&
      vec_pitch(row_words) <= tc.src_pitch_w && win_b >= 0 && wh > 0 && ww > 0) {
    const int wy = (int)(win_b / stride_b), wxb = (int)(win_b - (long long)wy * stride_b);
    const int lead = wxb & 15;
    const int nbox = (wh + kTmaRows - 1) / kTmaRows;
    const int src_bytes = (tc.src_rows * tc.src_pitch_w * 4 + 127) & ~127;
    const int box_bytes = kTmaRows * tc.win_pitch_w * 4;
    if ((lead + ww * px + 3) / 4 + 4 <= tc.win_pitch_w && (long long)nbox * box_bytes <= window_budget) {
      if (threadIdx.x == 0) mbar_init(&g_mbar, 1);
      __syncthreads();
      if (threadIdx.x == 0) {
        mbar_expect_tx(&g_mbar, (unsigned)(tc.src_rows * tc.src_pitch_w * 4 + nbox * box_bytes));
        const int sy = (int)(src_b / cur_stride_b), sxb = (int)(src_b - (long long)sy * cur_stride_b);
        tma_load_2d(smem, tm_cur, sxb >> 2, sy, &g_mbar);
        for (int b = 0; b < nbox; ++b)
          tma_load_2d(smem + src_bytes + b * box_bytes, tm_ref, (wxb - lead) >> 2, wy + b * kTmaRows, &g_mbar);
      }
      k.src_pitch_w = tc.src_pitch_w;
      k.win_region_off = src_bytes;
      k.win_budget = window_budget;
      k.ref_pitch = tc.win_pitch_w * 4 / px;
      k.ref0 = smem + src_bytes + lead + ((long long)(-wr0) * k.ref_pitch + (-wc0)) * px;
      k.fast = true;
      k.staged = true;
      k.win0_off = src_bytes + lead + ((-wr0) * k.ref_pitch + (-wc0)) * px;
      mbar_wait(&g_mbar, 0);
      return;
    }
  }
  // cooperative copy
  k.src_pitch_w = row_words >= 4 ? vec_pitch(row_words) : row_words;
  for (int i = threadIdx.x; i < bh * row_bytes; i += kThreads) {
    const int y = i / row_bytes, x = i - y * row_bytes;
    s_src[y * k.src_pitch_w * 4 + x] = g_src[(size_t)y * l.cur_stride * px + x];
  }
  const int src_bytes = (bh * k.src_pitch_w * 4 + 15) & ~15;
  k.win_region_off = src_bytes;
  k.win_budget = window_budget;
  stage_window(k, wr0, wc0, wh, ww);
  __syncthreads();
}

__host__ __device__ __forceinline__ int me_family(int method) {
  return method <= B200_ME_CLAMPED_DIAMOND ? 0 : (method == B200_ME_REFINE_8P ? 2 : 1);
}

template <int FAM>
__global__ void __launch_bounds__(kThreads, FAM == 3 ? 2 : 3) fullpel_kernel(const B200MeLaunch l, const Sites *sites4,
                                                           int window_budget, const TmaCfg tc,
                                                           const __grid_constant__ CUtensorMap tm_cur,
                                                           const __grid_constant__ CUtensorMap tm_ref) {
  const int bi = blockIdx.x;
  load_desc<true>(l, bi);
  Ctx k(g_desc);
  k.l = &l;
  if (me_family(k.d.search_method) != (FAM == 3 ? 0 : FAM)) return;  // instance 3 = family 0 on uint16 planes
  if (threadIdx.x == 0) g_bcast[24] = 0;  // served by another instance of the launch
  init_costs(k);
  {  // the method's site table: read at every step of the search, keep it on chip
    const int *g = reinterpret_cast<const int *>(sites4 + (k.d.search_method & 3));
    int *d = reinterpret_cast<int *>(&g_sites);
    for (int i = threadIdx.x; i < (int)(sizeof(Sites) / 4); i += kThreads) d[i] = g[i];
  }
  k.sites = &g_sites;
  k.px = l.highbd ? 2 : 1;
  k.sad_shift = l.highbd ? (l.bd == 10 ? 2 : (l.bd == 12 ? 4 : 0)) : 0;
  if constexpr (FAM == 2) init_compound(k, l, bi);  // compound refinement (av1_refining_search_8p_c)
  else k.compound = false;
#ifdef B200_ME_TIMING
  const long long t_start = clock64();
#endif
  // a block larger than the launch declared would overrun the shared memory sized from
  // max_bw / max_bh: its search is refused (ret = INT_MAX, mv = INVALID)
  if (k.d.bw > (l.max_bw > 0 ? l.max_bw : 128) || k.d.bh > (l.max_bh > 0 ? l.max_bh : 128)) {
    if (threadIdx.x == 0) {
      B200MeResult r;
      memset(&r, 0, sizeof(r));
      r.best_row = r.best_col = r.second_row = r.second_col = -32768;
      r.ret = INT_MAX;
      l.results[bi] = r;
    }
    return;
  }
  stage_inputs(k, me_smem, window_budget, 0, tc, &tm_cur, &tm_ref);
#ifdef B200_ME_TIMING
  if (threadIdx.x == 0) g_bcast[11] = 0, g_bcast[13] = 0, g_bcast[14] = 0, g_bcast[15] = 0;
  const long long t_staged = clock64();
#endif
  Mv best, second;
  int cl[5];
  int var;
  if constexpr (FAM == 2) {
    second.row = second.col = -32768;
    for (int i = 0; i < 5; ++i) cl[i] = 0;
    var = refine_8p(k, (k.d.flags & B200_ME_SKIP_SAD) != 0, &best);
  } else {
    var = full_pixel_search<FAM>(k, (k.d.flags & B200_ME_SKIP_SAD) != 0, cl, &best, &second);
  }
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
    r.distortion = g_bcast[12];                          // diamond phase, cycles / 16
    r.sse = (unsigned)((clock64() - t_start) >> 4);      // whole search
    r.cost_list[4] = (int)((t_staged - t_start) >> 4);   // staging
    r.cost_list[0] = g_bcast[11];
    r.cost_list[1] = g_bcast[13], r.cost_list[2] = g_bcast[14], r.cost_list[3] = g_bcast[15];  // mesh stages
#endif
    l.results[bi] = r;
    if (l.stats) {
      atomicAdd(l.stats, (unsigned long long)g_bcast[24]);
      atomicAdd(l.stats + 1, (unsigned long long)g_bcast[24] * (unsigned)(k.d.bw * k.d.bh));
    }
  }
}

// Back-to-back VABSDIFF4.ACC on 8 independent accumulators per thread: the integer pipe's
// measured SAD rate (b200_measure_sad_peak).
__global__ void __launch_bounds__(256) sad_peak_kernel(unsigned *out, int iters, unsigned seed) {
  unsigned a[8], x = seed + threadIdx.x, y = seed * 3 + blockIdx.x;
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
#pragma unroll
      for (int i = 0; i < 8; ++i) a[i] = sad4_acc(x + i, y, a[i]);
    }
    x += a[0] & 1;
  }
  unsigned t = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) t += a[i];
  if (t == 0x12345678u) out[0] = t;
}

}  // namespace me
}  // namespace b200
