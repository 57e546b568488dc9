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
  B200_CUDA_TRY(cudaFuncSetAttribute(me::fullpel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
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
  if (l->highbd && l->bd != 8 && l->bd != 10 && l->bd != 12) {
    b200::set_error("b200_me_*_launch: bd must be 8, 10 or 12");
    return B200_EINVAL;
  }
  return B200_OK;
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
  const int mbw = l->max_bw > 0 ? l->max_bw : 128, mbh = l->max_bh > 0 ? l->max_bh : 128;
  // source block with padded row pitch (<= row words + 11) + window
  // + 128: the strip loads of the strided mesh path may run a few words past the window
  const size_t smem = (size_t)(mbw * (l->highbd ? 2 : 1) + 44) * mbh + 16 + (size_t)budget + 128;
  b200::me::fullpel_kernel<<<l->n_blocks, b200::me::kThreads, smem, s>>>(
      *l, static_cast<const b200::me::Sites *>(ctx->d_me_sites), budget);
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
