// me_host.cuh -- host side of the motion search kernels (included by me.cu).
#pragma once

namespace b200 {
namespace me {

static void put_sites(Sites *s, int stage, int radius, int tan_radius, int npts) {
  const short m[13][2] = { { 0, 0 }, { (short)-radius, 0 }, { (short)radius, 0 },
                           { 0, (short)-radius }, { 0, (short)radius },
                           { (short)-radius, (short)-tan_radius }, { (short)radius, (short)tan_radius },
                           { (short)-tan_radius, (short)radius }, { (short)tan_radius, (short)-radius },
                           { (short)-radius, (short)tan_radius }, { (short)radius, (short)-tan_radius },
                           { (short)tan_radius, (short)radius }, { (short)-tan_radius, (short)-radius } };
  for (int i = 0; i <= npts; ++i) {
    s->site[stage][i][0] = m[i][0];
    s->site[stage][i][1] = m[i][1];
  }
  s->per_step[stage] = npts;
  s->radius[stage] = radius;
}

// av1_init_motion_compensation_nstep (mcomp.c:452) / av1_init_dsmotion_compensation (:369)
static void build_sites(Sites *s, int method) {
  memset(s, 0, sizeof(*s));
  if (method == B200_ME_NSTEP || method == B200_ME_NSTEP_8PT) {
    const int level = method == B200_ME_NSTEP_8PT;
    const int stages = level ? 16 : 15;
    int radius = 1;
    for (int st = 0; st < stages; ++st) {
      int tanr = (int)(0.41 * radius);
      if (tanr < 1) tanr = 1;
      int npts = 12;
      if (radius <= 5 || level) {
        tanr = radius;
        npts = 8;
      }
      put_sites(s, st, radius, tanr, npts);
      if (st < 12) {
        const int grown = (int)(radius * 1.5 + 0.5);
        radius = grown > radius + 1 ? grown : radius + 1;
      }
    }
    s->n_steps = stages;
  } else {
    const int level = method == B200_ME_CLAMPED_DIAMOND;
    int stage = kMaxSteps - 1, n = 0;
    for (int radius = level ? 256 : 1024; radius > 0;) {
      put_sites(s, stage, radius, radius, 8);
      if (!level || stage < 9) radius /= 2;
      --stage;
      ++n;
    }
    s->n_steps = n;
  }
}

}  // namespace me

int me_init_device(b200_ctx *ctx) {
  me::Sites *h = new me::Sites[4];
  for (int m = 0; m < 4; ++m) me::build_sites(&h[m], m);
  B200_CUDA_TRY(cudaMalloc(&ctx->d_me_sites, sizeof(me::Sites) * 4));
  B200_CUDA_TRY(cudaMemcpy(ctx->d_me_sites, h, sizeof(me::Sites) * 4, cudaMemcpyHostToDevice));
  delete[] h;
  B200_CUDA_TRY(cudaFuncSetAttribute(me::fullpel_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  B200_CUDA_TRY(cudaFuncSetAttribute(me::fullpel_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  B200_CUDA_TRY(cudaFuncSetAttribute(me::fullpel_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  B200_CUDA_TRY(cudaFuncSetAttribute(me::fullpel_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  B200_CUDA_TRY(cudaFuncSetAttribute(me::subpel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     200 * 1024));
  return B200_OK;
}

}  // namespace b200

static int me_check(const B200MeLaunch *l) {
  if (!l->blocks || !l->cur || !l->ref || !l->results || l->cur_stride <= 0 || l->ref_stride <= 0) {
    b200::set_error("b200_me_*_launch: missing buffer or stride");
    return B200_EINVAL;
  }
  if (l->compound && !l->second_pred) {
    b200::set_error("b200_me_*_launch: compound descriptors without second_pred");
    return B200_EINVAL;
  }
  if ((l->mvcost[0] == nullptr) != (l->mvcost[1] == nullptr) || (l->mvcost[0] && !l->mvjcost)) {
    b200::set_error("b200_me_*_launch: mvjcost / mvcost[0] / mvcost[1] must be given together");
    return B200_EINVAL;
  }
  if (l->highbd && l->bd != 8 && l->bd != 10 && l->bd != 12) {
    b200::set_error("b200_me_*_launch: bd must be 8, 10 or 12");
    return B200_EINVAL;
  }
  return B200_OK;
}

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link against libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
        q != cudaDriverEntryPointSuccess)
      p = nullptr;
    cudaGetLastError();
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// A plane of `rows` rows of `stride_b` bytes as a 2-D tensor of 32-bit words with a box of
// box_w words x box_h rows.  False when the plane cannot be described (alignment, sizes).
static bool make_plane_map(CUtensorMap *m, const void *base, long long stride_b, int rows, int box_w,
                           int box_h) {
  EncodeTiledFn enc = encode_tiled_fn();
  if (!enc || !base || rows <= 0 || (stride_b & 15) || ((uintptr_t)base & 15) || box_w > 256 ||
      box_h > 256 || box_w <= 0 || box_h <= 0)
    return false;
  const cuuint64_t dims[2] = { (cuuint64_t)(stride_b / 4), (cuuint64_t)rows };
  const cuuint64_t strides[1] = { (cuuint64_t)stride_b };
  const cuuint32_t box[2] = { (cuuint32_t)box_w, (cuuint32_t)box_h };
  const cuuint32_t estr[2] = { 1, 1 };
  return enc(m, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, const_cast<void *>(base), dims, strides, box, estr,
             CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
             CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

extern "C" int b200_me_fullpel_launch(b200_ctx *ctx, const B200MeLaunch *l, void *stream) {
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (l->n_blocks == 0) return B200_OK;
  int rc = me_check(l);
  if (rc != B200_OK) return rc;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  int budget = l->window_bytes;
  if (budget < 0) budget = 0;
  if (budget > b200::me::kWindowBudget) budget = b200::me::kWindowBudget;
  const int px = l->highbd ? 2 : 1;
  const int mbw = l->max_bw > 0 ? l->max_bw : 128, mbh = l->max_bh > 0 ? l->max_bh : 128;
  // source block with padded row pitch (<= row words + 11), rounded for the TMA destination,
  // + window + 128: the strip loads of the strided mesh path may run a few words past it
  const size_t smem = (((size_t)(mbw * px + 44) * mbh + 127) & ~(size_t)127) + (size_t)budget + 256;
  // TMA staging when the caller told us the plane extents and the largest window
  b200::me::TmaCfg tc = { 0, 0, 0, 0 };
  CUtensorMap tm_cur, tm_ref;
  memset(&tm_cur, 0, sizeof(tm_cur));
  memset(&tm_ref, 0, sizeof(tm_ref));
  if (budget > 0 && l->cur_rows > 0 && l->ref_rows > 0 && l->max_col_span > 0 && l->max_row_span > 0 &&
      mbw * px >= 16) {
    const int win_w = b200::me::vec_pitch((15 + (l->max_col_span + mbw) * px + 3) / 4 + 4);
    const int src_w = b200::me::vec_pitch(mbw * px / 4);
    if (make_plane_map(&tm_ref, l->ref, (long long)l->ref_stride * px, l->ref_rows, win_w, b200::me::kTmaRows) &&
        make_plane_map(&tm_cur, l->cur, (long long)l->cur_stride * px, l->cur_rows, src_w, mbh)) {
      tc.use = 1;
      tc.win_pitch_w = win_w;
      tc.src_pitch_w = src_w;
      tc.src_rows = mbh;
    }
  }
  // one kernel instance per method family present in the batch (the descriptors live in
  // device memory: without the caller's hint all three are launched and the CTAs of the
  // other families exit at once)
  const unsigned fam = (l->families & 7u) ? (l->families & 7u) : 7u;
  const b200::me::Sites *sites = static_cast<const b200::me::Sites *>(ctx->d_me_sites);
  if ((fam & B200_ME_FAMILY_NSTEP) && !l->highbd)
    b200::me::fullpel_kernel<0><<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l, sites, budget, tc, tm_cur, tm_ref);
  if ((fam & B200_ME_FAMILY_NSTEP) && l->highbd)  // uint16 planes: the n-step family's own instance
    b200::me::fullpel_kernel<3><<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l, sites, budget, tc, tm_cur, tm_ref);
  if (fam & B200_ME_FAMILY_PATTERN)
    b200::me::fullpel_kernel<1><<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l, sites, budget, tc, tm_cur, tm_ref);
  if (fam & B200_ME_FAMILY_REFINE)
    b200::me::fullpel_kernel<2><<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l, sites, budget, tc, tm_cur, tm_ref);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_me_subpel_launch(b200_ctx *ctx, const B200MeLaunch *l, void *stream) {
  if (!ctx || !l || l->n_blocks < 0) return B200_EINVAL;
  if (l->n_blocks == 0) return B200_OK;
  int rc = me_check(l);
  if (rc != B200_OK) return rc;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
  // source block + horizontally filtered rows (bh + 7) * bw * 2 bytes
  const int mbw = l->max_bw > 0 ? l->max_bw : 128, mbh = l->max_bh > 0 ? l->max_bh : 128;
  size_t smem = (size_t)mbw * mbh * (l->highbd ? 2 : 1) + 16 + (size_t)(mbh + 7) * mbw * 2;
  if (!l->highbd) {
    // 8-bit fast path (me_subpel.cuh): reference window + transposed source + 4 transposed
    // intermediates, laid out after the generic regions
    smem = (smem + 15) & ~(size_t)15;
    smem += (size_t)(mbh + 9) * ((((mbw + 9 + 3 + 3) & ~3) + 4)) + (size_t)mbw * b200::me::odd_word_pitch(mbh) +
            4 * (size_t)mbw * b200::me::odd_word_pitch(mbh + 8) + 64;
  }
  b200::me::subpel_kernel<<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

extern "C" int b200_me_init_search_range(int size) {
  int sr = 0;
  if (size < 16) size = 16;
  while ((size << sr) < 1023) sr++;  // MAX_FULL_PEL_VAL
  return sr < b200::me::kMaxSteps - 2 ? sr : b200::me::kMaxSteps - 2;
}

extern "C" int b200_measure_sad_peak(b200_ctx *ctx, double *gabsdiff_per_s) {
  if (!ctx || !gabsdiff_per_s) return B200_EINVAL;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  unsigned *d = nullptr;
  B200_CUDA_TRY(cudaMalloc(&d, 16));
  cudaEvent_t e0, e1;
  B200_CUDA_TRY(cudaEventCreate(&e0));
  B200_CUDA_TRY(cudaEventCreate(&e1));
  const int grid = ctx->sm_count * 8, iters = 4096;
  float best = 1e30f;
  for (int rep = 0; rep < 5; ++rep) {
    B200_CUDA_TRY(cudaEventRecord(e0, ctx->stream));
    b200::me::sad_peak_kernel<<<grid, 256, 0, ctx->stream>>>(d, iters, 17u + rep);
    B200_CUDA_TRY(cudaEventRecord(e1, ctx->stream));
    B200_CUDA_TRY(cudaEventSynchronize(e1));
    float ms = 0;
    B200_CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best) best = ms;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  const double ops = (double)grid * 256 * iters * 32 * 4;  // 32 VABSDIFF4 per iteration, 4 pixels each
  *gabsdiff_per_s = ops / (best * 1e-3) / 1e9;
  return B200_OK;
}

// b200_me_batch_host pipelines the call: the descriptors are cut into up to kMeHostChunks chunks of at least
// two waves of CTAs; the copy stream brings each chunk's part of the planes up (the byte range its searches
// can touch: limits box + block + the sub-pel filter margin; ranges already on the device are not copied
// again) while the two compute streams, alternating, run the previous chunks' full-pel + sub-pel kernels and
// read their results back -- so the step costs max(copy, compute) instead of their sum, and the tail of one
// chunk's kernels overlaps the head of the next chunk's.
namespace {
constexpr int kMeHostChunks = 8;
constexpr long long kMeHostMargin = 9;  // pixels around the limits box: 8-tap filter (4) + the last sub-pel step, rounded up
struct ByteSpan {
  long long lo, hi;
};
// copies what [want.lo, want.hi) adds to the contiguous span `have` already on the device (256-byte granules)
int copy_new_part(char *dst, const char *src, long long total, ByteSpan want, ByteSpan &have, cudaStream_t s,
                  size_t *copied) {
  want.lo = want.lo < 0 ? 0 : (want.lo & ~255ll);
  want.hi = want.hi > total ? total : ((want.hi + 255) & ~255ll);
  if (want.hi > total) want.hi = total;
  if (want.lo >= want.hi) return B200_OK;
  auto put = [&](long long lo, long long hi) -> int {
    if (lo >= hi) return B200_OK;
    B200_CUDA_TRY(cudaMemcpyAsync(dst + lo, src + lo, (size_t)(hi - lo), cudaMemcpyHostToDevice, s));
    *copied += (size_t)(hi - lo);
    return B200_OK;
  };
  int rc = B200_OK;
  if (have.lo >= have.hi) {
    rc = put(want.lo, want.hi);
    have = want;
    return rc;
  }
  if (want.lo < have.lo) {
    rc = put(want.lo, have.lo);
    have.lo = want.lo;
  }
  if (rc == B200_OK && want.hi > have.hi) {
    rc = put(have.hi, want.hi);
    have.hi = want.hi;
  }
  return rc;
}
}  // namespace

extern "C" int b200_me_batch_host(b200_ctx *ctx, const B200MeLaunch *hl, size_t cur_bytes, size_t ref_bytes,
                                  const B200MeBlock *subpel_blocks, B200MeResult *fullpel_out,
                                  B200MeResult *subpel_out) {
  if (!ctx || !hl || hl->n_blocks < 0 || !hl->blocks || !hl->cur || !hl->ref) return B200_EINVAL;
  if (hl->compound || hl->second_pred || hl->mask) {
    b200::set_error("b200_me_batch_host: compound searches go through b200_me_fullpel_launch / b200_me_subpel_launch");
    return B200_EINVAL;
  }
  if (hl->n_blocks == 0) return B200_OK;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const size_t n = (size_t)hl->n_blocks;
  const size_t tab = (size_t)(2 * 16383 + 1) * sizeof(int);
  const bool costs = hl->mvcost[0] && hl->mvcost[1] && hl->mvjcost;
  size_t off = 0;
  auto carve = [&](size_t bytes) {
    const size_t o = off;
    off += (bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t o_cur = carve(cur_bytes), o_ref = carve(ref_bytes), o_blk = carve(n * sizeof(B200MeBlock));
  const size_t o_sub = carve(subpel_blocks ? n * sizeof(B200MeBlock) : 0);
  const size_t o_rf = carve(n * sizeof(B200MeResult)), o_rs = carve(subpel_blocks ? n * sizeof(B200MeResult) : 0);
  const size_t o_j = carve(costs ? 4 * sizeof(int) : 0), o_c0 = carve(costs ? tab : 0), o_c1 = carve(costs ? tab : 0);
  int rc = b200::scratch_reserve(ctx, off);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  cudaStream_t cs = ctx->copy_stream;
  cudaStream_t st[2] = { ctx->stream, ctx->aux_stream };
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_blk, hl->blocks, n * sizeof(B200MeBlock), cudaMemcpyHostToDevice, cs));
  if (subpel_blocks)
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_sub, subpel_blocks, n * sizeof(B200MeBlock), cudaMemcpyHostToDevice, cs));
  B200MeLaunch dl = *hl;
  dl.cur = base + o_cur;
  dl.ref = base + o_ref;
  dl.cost_lists = nullptr;
  dl.start_from = nullptr;
  dl.stats = nullptr;
  dl.mvjcost = nullptr;
  dl.mvcost[0] = dl.mvcost[1] = nullptr;
  if (costs) {
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_j, hl->mvjcost, 4 * sizeof(int), cudaMemcpyHostToDevice, cs));
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_c0, hl->mvcost[0] - 16383, tab, cudaMemcpyHostToDevice, cs));
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_c1, hl->mvcost[1] - 16383, tab, cudaMemcpyHostToDevice, cs));
    dl.mvjcost = (const int *)(base + o_j);
    dl.mvcost[0] = (const int *)(base + o_c0) + 16383;
    dl.mvcost[1] = (const int *)(base + o_c1) + 16383;
  }
  // chunks of >= two waves of resident CTAs; a sub-pel queue that does not mirror the full-pel one
  // (other block positions) is served unpipelined: every plane byte is up before the first kernel
  const long long px = hl->highbd ? 2 : 1;
  bool mirror = true;
  if (subpel_blocks)
    for (size_t i = 0; i < n && mirror; ++i)
      mirror = subpel_blocks[i].src_offset == hl->blocks[i].src_offset &&
               subpel_blocks[i].ref_offset == hl->blocks[i].ref_offset;
  const size_t waves2 = (size_t)ctx->sm_count * 6;
  int n_chunks = !mirror ? 1 : (int)std::min<size_t>(kMeHostChunks, std::max<size_t>(1, n / waves2));
  const size_t chunk = (n + n_chunks - 1) / n_chunks;
  ByteSpan have_cur = { 0, 0 }, have_ref = { 0, 0 };
  size_t copied = 0;
  for (int c = 0; c < n_chunks; ++c) {
    const size_t b0 = chunk * c, b1 = std::min(n, b0 + chunk);
    if (b0 >= b1) break;
    ByteSpan wc = { 0, (long long)cur_bytes }, wr = { 0, (long long)ref_bytes };
    if (n_chunks > 1) {
      wc.lo = wr.lo = LLONG_MAX, wc.hi = wr.hi = 0;
      const long long mbh = hl->max_bh > 0 ? hl->max_bh : 128, mbw = hl->max_bw > 0 ? hl->max_bw : 128;
      for (size_t i = b0; i < b1; ++i) {
        const B200MeBlock &d = hl->blocks[i];
        // the source box the kernels stage is max_bh x max_bw from the block's origin
        const long long s0 = (long long)d.src_offset, s1 = s0 + (mbh - 1) * hl->cur_stride + mbw + 16;
        const long long r0 = (long long)d.ref_offset + (d.row_min - kMeHostMargin) * hl->ref_stride + d.col_min - kMeHostMargin;
        const long long r1 = (long long)d.ref_offset + (d.row_max + d.bh + kMeHostMargin) * hl->ref_stride + d.col_max +
                             d.bw + kMeHostMargin;
        wc.lo = std::min(wc.lo, s0), wc.hi = std::max(wc.hi, s1);
        wr.lo = std::min(wr.lo, r0), wr.hi = std::max(wr.hi, r1);
      }
      wc.lo *= px, wc.hi *= px, wr.lo *= px, wr.hi *= px;
    }
    rc = copy_new_part(base + o_cur, (const char *)hl->cur, (long long)cur_bytes, wc, have_cur, cs, &copied);
    if (rc == B200_OK)
      rc = copy_new_part(base + o_ref, (const char *)hl->ref, (long long)ref_bytes, wr, have_ref, cs, &copied);
    if (rc != B200_OK) return rc;
    B200_CUDA_TRY(cudaEventRecord(ctx->group_done[c], cs));
    cudaStream_t s = st[c & 1];
    B200_CUDA_TRY(cudaStreamWaitEvent(s, ctx->group_done[c], 0));
    dl.n_blocks = (int)(b1 - b0);
    dl.blocks = (const B200MeBlock *)(base + o_blk) + b0;
    dl.results = (B200MeResult *)(base + o_rf) + b0;
    rc = b200_me_fullpel_launch(ctx, &dl, s);
    if (rc != B200_OK) return rc;
    if (fullpel_out)
      B200_CUDA_TRY(cudaMemcpyAsync(fullpel_out + b0, base + o_rf + b0 * sizeof(B200MeResult),
                                    (b1 - b0) * sizeof(B200MeResult), cudaMemcpyDeviceToHost, s));
    if (subpel_blocks) {
      B200MeLaunch sl = dl;
      sl.blocks = (const B200MeBlock *)(base + o_sub) + b0;
      sl.start_from = (const B200MeResult *)(base + o_rf) + b0;
      sl.results = (B200MeResult *)(base + o_rs) + b0;
      // the pruned trees read the full-pel cost list in place
      sl.cost_lists = &((const B200MeResult *)(base + o_rf) + b0)->cost_list[0];
      sl.cost_list_stride = sizeof(B200MeResult) / sizeof(int32_t);
      rc = b200_me_subpel_launch(ctx, &sl, s);
      if (rc != B200_OK) return rc;
      if (subpel_out)
        B200_CUDA_TRY(cudaMemcpyAsync(subpel_out + b0, base + o_rs + b0 * sizeof(B200MeResult),
                                      (b1 - b0) * sizeof(B200MeResult), cudaMemcpyDeviceToHost, s));
    }
  }
  B200_CUDA_TRY(cudaStreamSynchronize(cs));
  B200_CUDA_TRY(cudaStreamSynchronize(st[0]));
  B200_CUDA_TRY(cudaStreamSynchronize(st[1]));
  return B200_OK;
}
