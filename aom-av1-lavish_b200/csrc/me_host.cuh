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
  B200_CUDA_TRY(cudaFuncSetAttribute(me::subpel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     100 * 1024));
  return B200_OK;
}

}  // namespace b200

static int me_check(const B200MeLaunch *l) {
  if (!l->blocks || !l->cur || !l->ref || !l->results || l->cur_stride <= 0 || l->ref_stride <= 0) {
    b200::set_error("b200_me_*_launch: missing buffer or stride");
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
  if (fam & B200_ME_FAMILY_NSTEP)
    b200::me::fullpel_kernel<0><<<l->n_blocks, b200::me::kThreads, smem, s>>>(*l, sites, budget, tc, tm_cur, tm_ref);
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
  const size_t smem = (size_t)mbw * mbh * (l->highbd ? 2 : 1) + 16 + (size_t)(mbh + 7) * mbw * 2;
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
