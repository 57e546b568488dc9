// queue.cu -- frame tier and the encoder-side block queue (SURVEY.md section 8b tier 3, rows
// a25 / a26 / f1): the per-block loops of the encoder's look-ahead stages turned into one
// descriptor list per frame and a chained launch sequence.
//
//   b200_frame_*                  a luma plane on the device with its replicated border
//   b200_queue_tf_motion_search   tf_motion_search for every block of a frame pair
//                                 (av1/encoder/temporal_filter.c:88-260, loop :804-872)
//   b200_queue_tpl_*              TPL motion_estimation + txfm_quant_rdcost
//                                 (av1/encoder/tpl_model.c:228-304)
#include <limits.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "internal.h"

struct b200_frame {
  void *d_plane;   // (height + 2*border) rows of `pitch` pixels
  int pitch;       // pixels, a multiple of 16
  int rows;        // height + 2 * border
  int width, height, border, bd;
  int origin;      // pixel offset of picture pixel (0,0) from d_plane
};

namespace b200 {
namespace {

// aom_extend_frame_borders (aom_scale/generic/yv12extend.c:223, extend_plane :24-64): every
// pixel outside the picture takes the value of the nearest picture pixel.  The picture itself
// was copied in by cudaMemcpy2D; one thread per 4 border pixels of a row.
template <class T>
__global__ void extend_border_kernel(T *plane, int pitch, int width, int height, int border) {
  const int y = blockIdx.y - border;  // row relative to the picture
  const int cy = min(max(y, 0), height - 1);
  T *row = plane + (size_t)(y + border) * pitch + border;
  const T *src = plane + (size_t)(cy + border) * pitch + border;
  for (int x = blockIdx.x * blockDim.x + threadIdx.x - border; x < width + border; x += gridDim.x * blockDim.x) {
    if (y >= 0 && y < height && x >= 0 && x < width) continue;
    row[x] = src[min(max(x, 0), width - 1)];
  }
}

inline int imin(int a, int b) { return a < b ? a : b; }
inline int imax(int a, int b) { return a > b ? a : b; }

struct Limits {
  int col_min, col_max, row_min, row_max;
};

// av1_set_mv_row_limits / av1_set_mv_col_limits (mcomp.h:227-251); AOM_INTERP_EXTEND = 4
// (av1/common/filter.h), MI_SIZE = 4.
Limits block_limits(int mi_rows, int mi_cols, int mi_row, int mi_col, int mi_h, int mi_w, int border) {
  const int ext = 2 * 4;
  Limits l;
  l.row_min = imax(-(mi_row * 4 + border - ext), -((mi_row + mi_h) * 4 + ext));
  l.row_max = imin((mi_rows - mi_row - mi_h) * 4 + border - ext, (mi_rows - mi_row) * 4 + ext);
  l.col_min = imax(-(mi_col * 4 + border - ext), -((mi_col + mi_w) * 4 + ext));
  l.col_max = imin((mi_cols - mi_col - mi_w) * 4 + border - ext, (mi_cols - mi_col) * 4 + ext);
  return l;
}

constexpr int kMaxFullPel = 1023;            // MAX_FULL_PEL_VAL (mcomp.h:30)
constexpr int kMvLow = -(1 << 14), kMvUpp = 1 << 14;  // MV_LOW / MV_UPP (mv.h)

// av1_set_mv_search_range (mcomp.c:204-237)
Limits fullpel_range(Limits l, int mv_row, int mv_col) {
  int col_min = ((mv_col + 7) >> 3) - kMaxFullPel, row_min = ((mv_row + 7) >> 3) - kMaxFullPel;
  int col_max = (mv_col >> 3) + kMaxFullPel, row_max = (mv_row >> 3) + kMaxFullPel;
  col_min = imax(col_min, (kMvLow >> 3) + 1);
  row_min = imax(row_min, (kMvLow >> 3) + 1);
  col_max = imin(col_max, (kMvUpp >> 3) - 1);
  row_max = imin(row_max, (kMvUpp >> 3) - 1);
  if (l.col_min < col_min) l.col_min = col_min;
  if (l.col_max > col_max) l.col_max = col_max;
  if (l.row_min < row_min) l.row_min = row_min;
  if (l.row_max > row_max) l.row_max = row_max;
  return l;
}

// av1_set_subpel_mv_search_range (mcomp.h:357-373)
Limits subpel_range(const Limits &l, int mv_row, int mv_col) {
  const int max_mv = kMaxFullPel * 8;
  int minc = imax(l.col_min * 8, mv_col - max_mv), maxc = imin(l.col_max * 8, mv_col + max_mv);
  int minr = imax(l.row_min * 8, mv_row - max_mv), maxr = imin(l.row_max * 8, mv_row + max_mv);
  maxc = imax(minc, maxc);
  maxr = imax(minr, maxr);
  Limits s;
  s.col_min = imax(kMvLow + 1, minc);
  s.col_max = imin(kMvUpp - 1, maxc);
  s.row_min = imax(kMvLow + 1, minr);
  s.row_max = imin(kMvUpp - 1, maxr);
  return s;
}

inline int rawpel(int v) { return (v + 3 + (v >= 0)) >> 3; }  // GET_MV_RAWPEL (mv.h:28)

void put_limits(B200MeBlock *d, const Limits &l) {
  d->col_min = (int16_t)l.col_min;
  d->col_max = (int16_t)l.col_max;
  d->row_min = (int16_t)l.row_min;
  d->row_max = (int16_t)l.row_max;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" int b200_frame_upload(b200_ctx *ctx, const void *y, int stride, int width, int height, int border,
                                 int bd, b200_frame **out) {
  if (!ctx || !y || !out || width <= 0 || height <= 0 || border < 0 || stride < width) return B200_EINVAL;
  if (bd != 8 && bd != 10 && bd != 12) {
    set_error("b200_frame_upload: bd must be 8, 10 or 12");
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const int px = bd > 8 ? 2 : 1;
  b200_frame *f = (b200_frame *)calloc(1, sizeof(*f));
  if (!f) return B200_ECUDA;
  f->width = width, f->height = height, f->border = border, f->bd = bd;
  f->pitch = (width + 2 * border + 15) & ~15;
  f->rows = height + 2 * border;
  f->origin = border * f->pitch + border;
  if (cudaMalloc(&f->d_plane, (size_t)f->pitch * f->rows * px) != cudaSuccess) {
    free(f);
    set_error("b200_frame_upload: out of device memory");
    return B200_ECUDA;
  }
  cudaStream_t s = ctx->stream;
  char *dst = (char *)f->d_plane + (size_t)f->origin * px;
  cudaError_t e = cudaMemcpy2DAsync(dst, (size_t)f->pitch * px, y, (size_t)stride * px, (size_t)width * px, height,
                                    cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess && border > 0) {
    const dim3 grid(4, f->rows);
    if (px == 1)
      extend_border_kernel<unsigned char><<<grid, 256, 0, s>>>((unsigned char *)f->d_plane, f->pitch, width, height, border);
    else
      extend_border_kernel<unsigned short><<<grid, 256, 0, s>>>((unsigned short *)f->d_plane, f->pitch, width, height, border);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) {
    set_error("b200_frame_upload: %s", cudaGetErrorString(e));
    cudaFree(f->d_plane);
    free(f);
    return B200_ECUDA;
  }
  *out = f;
  return B200_OK;
}

extern "C" int b200_frame_free(b200_ctx *ctx, b200_frame *f) {
  if (!ctx || !f) return B200_EINVAL;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  cudaFree(f->d_plane);
  free(f);
  return B200_OK;
}

extern "C" int b200_frame_download(b200_ctx *ctx, const b200_frame *f, void *out) {
  if (!ctx || !f || !out) return B200_EINVAL;
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const int px = f->bd > 8 ? 2 : 1, w = f->width + 2 * f->border;
  B200_CUDA_TRY(cudaMemcpy2DAsync(out, (size_t)w * px, f->d_plane, (size_t)f->pitch * px, (size_t)w * px, f->rows,
                                  cudaMemcpyDeviceToHost, ctx->stream));
  B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
  return B200_OK;
}

// ------------------------------------------------------------------ temporal filter ----
extern "C" void b200_queue_tf_default_params(int width, int height, B200TfParams *p) {
  memset(p, 0, sizeof(*p));
  const int min_dim = width < height ? width : height, max_dim = width > height ? width : height;
  p->bw = p->bh = 16;                                  // TF_BLOCK_SIZE, temporal_filter.h:31
  p->border_in_pixels = 160;                           // AOM_ENC_NO_SCALE_BORDER, aom_scale/yv12config.h:28-32
  p->search_method = B200_ME_NSTEP;                    // temporal_filter.c:134
  p->step_param = b200_me_init_search_range(max_dim);  // :112
  p->mv_cost_type = min_dim >= 720 ? B200_MV_COST_L1_HDRES
                                   : (min_dim >= 480 ? B200_MV_COST_L1_MIDRES : B200_MV_COST_L1_LOWRES);  // :116-119
  const int range[4] = { 64, 28, 15, 7 }, interval[4] = { 8, 4, 1, 1 };  // good_quality_mesh_patterns[0], speed_features.c:26
  for (int i = 0; i < 4; ++i) p->mesh_range[i] = range[i], p->mesh_interval[i] = interval[i];
  p->force_mesh_thresh = 1 << 25;                      // speed_features.c:356 (>= 720p is 1 << 25, below 1 << 20)
  if (min_dim < 720) p->force_mesh_thresh = 1 << 20;
  p->prune_mesh_search = 0;
  p->mesh_search_mv_diff_threshold = 4;                // mcomp.c:147
  p->use_downsampled_sad = 0;
  p->sad_per_bit = 16;                                 // av1_set_sad_per_bit at a mid qindex
  p->error_per_bit = 10;
  p->allow_hp = 0;                                     // high precision mv is off above qindex 128 / for TF frames
  p->iters_per_step = 2;                               // speed_features.c:1929
  p->subpel_tree = B200_SUBPEL_TREE;
  p->want_cost_list = 0;
}

static int tf_grid(int width, int height, const B200TfParams *p, int *rows, int *cols) {
  if (!p || p->bw < 8 || p->bh < 8 || (p->bw & (p->bw - 1)) || (p->bh & (p->bh - 1)) || p->bw > 128 || p->bh > 128)
    return B200_EINVAL;
  *rows = (height + p->bh - 1) / p->bh;  // get_num_blocks, temporal_filter.h
  *cols = (width + p->bw - 1) / p->bw;
  return B200_OK;
}

extern "C" int b200_queue_tf_describe(int width, int height, int pitch, int origin, const B200TfParams *p,
                                      const int16_t *ref_mvs, int stage, B200MeBlock *full, B200MeBlock *sub) {
  int mb_rows, mb_cols;
  if (tf_grid(width, height, p, &mb_rows, &mb_cols) != B200_OK || !full || !sub || stage < 0 || stage > 1)
    return -1;
  // cm->mi_params: the picture rounded up to 8 pixels, in 4x4 units (av1_set_mb_mi / enc_set_mb_mi)
  const int mi_rows = ((height + 7) & ~7) >> 2, mi_cols = ((width + 7) & ~7) >> 2;
  const int sw = p->bw >> 1, sh = p->bh >> 1;
  int n = 0;
  for (int r = 0; r < mb_rows; ++r)
    for (int c = 0; c < mb_cols; ++c) {
      // mb->mv_limits of the WHOLE block; the sub-block searches keep them (temporal_filter.c:854-859)
      const Limits lim = block_limits(mi_rows, mi_cols, r * (p->bh >> 2), c * (p->bw >> 2), p->bh >> 2, p->bw >> 2,
                                      p->border_in_pixels);
      const Limits fl = fullpel_range(lim, 0, 0);  // baseline_mv = kZeroMv (:124)
      const Limits sl = subpel_range(lim, 0, 0);
      const int parts = stage == 0 ? 1 : 4;
      for (int k = 0; k < parts; ++k) {
        const int bw = stage == 0 ? p->bw : sw, bh = stage == 0 ? p->bh : sh;
        const int x = c * p->bw + (stage ? (k & 1) * sw : 0), y = r * p->bh + (stage ? (k >> 1) * sh : 0);
        B200MeBlock d;
        memset(&d, 0, sizeof(d));
        d.src_offset = d.ref_offset = (uint32_t)(origin + y * pitch + x);
        d.bw = (uint8_t)bw, d.bh = (uint8_t)bh;
        d.sad_per_bit = p->sad_per_bit, d.error_per_bit = p->error_per_bit;
        d.force_mesh_thresh = p->force_mesh_thresh;
        d.mesh_mv_diff_threshold = (uint8_t)p->mesh_search_mv_diff_threshold;
        B200MeBlock f = d;
        put_limits(&f, fl);
        if (stage == 0 && ref_mvs) {  // start_mv = get_fullmv_from_mv(ref_mv) (:122); stage 1 is chained on the device
          f.start_row = (int16_t)rawpel(ref_mvs[2 * (r * mb_cols + c)]);
          f.start_col = (int16_t)rawpel(ref_mvs[2 * (r * mb_cols + c) + 1]);
        }
        f.search_method = (uint8_t)p->search_method;
        f.step_param = (uint8_t)p->step_param;
        f.mv_cost_type = (uint8_t)p->mv_cost_type;
        f.flags = B200_ME_RUN_MESH | (p->prune_mesh_search ? B200_ME_PRUNE_MESH : 0) |
                  (p->use_downsampled_sad && bh >= 16 ? B200_ME_SKIP_SAD : 0) |
                  (p->want_cost_list ? B200_ME_COST_LIST : 0);
        full[n] = f;
        B200MeBlock s = d;
        put_limits(&s, sl);
        s.search_method = (uint8_t)p->subpel_tree;
        s.mv_cost_type = B200_MV_COST_NONE;  // :166, :217
        s.flags = (p->allow_hp ? B200_ME_ALLOW_HP : 0) | (p->want_cost_list ? B200_ME_COST_LIST : 0);
        s.subpel = (uint8_t)(B200_EIGHTH_PEL | (p->iters_per_step << 2) | (B200_USE_8_TAPS << 4));  // :161-163
        sub[n] = s;
        ++n;
      }
    }
  return n;
}

// Shared memory for the mesh-local window of a frame-wide search (me_fullpel.cuh mesh_search): the
// largest mesh stage box is (2 * range + block) pixels square, rows padded to the vector pitch.
static int mesh_window_bytes(const int *mesh_range, int bw, int bh, int px) {
  int r = 0;
  for (int i = 0; i < 4; ++i) r = mesh_range[i] > r ? mesh_range[i] : r;
  if (r <= 0) return 0;
  if (r > 64) r = 64;  // larger first stages (range scaled by the start mv, mcomp.c:1620) fall back to L1/L2
  const int rows = 2 * r + bh, words = ((15 + (2 * r + bw) * px + 3) / 4 + 4 + 11) & ~3;
  return rows * (words + 8) * 4;
}

extern "C" int b200_queue_tf_motion_search(b200_ctx *ctx, const b200_frame *cur, const b200_frame *ref,
                                           const B200TfParams *p, const int16_t *ref_mvs, B200TfResult *out) {
  if (!ctx || !cur || !ref || !p || !out) return B200_EINVAL;
  if (cur->width != ref->width || cur->height != ref->height || cur->pitch != ref->pitch ||
      cur->border != ref->border || cur->bd != ref->bd) {
    set_error("b200_queue_tf_motion_search: the two frames differ in geometry");
    return B200_EINVAL;
  }
  int mb_rows, mb_cols;
  if (tf_grid(cur->width, cur->height, p, &mb_rows, &mb_cols) != B200_OK) return B200_EINVAL;
  if (cur->border < p->bw || cur->border < p->border_in_pixels) {
    set_error("b200_queue_tf_motion_search: frame border %d smaller than border_in_pixels %d", cur->border,
              p->border_in_pixels);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const int n = mb_rows * mb_cols;
  std::vector<B200MeBlock> desc((size_t)n * 10);  // full0[n] sub0[n] full1[4n] sub1[4n]
  B200MeBlock *full0 = desc.data(), *sub0 = full0 + n, *full1 = sub0 + n, *sub1 = full1 + 4 * n;
  if (b200_queue_tf_describe(cur->width, cur->height, cur->pitch, cur->origin, p, ref_mvs, 0, full0, sub0) != n ||
      b200_queue_tf_describe(cur->width, cur->height, cur->pitch, cur->origin, p, ref_mvs, 1, full1, sub1) != 4 * n)
    return B200_EINVAL;
  const size_t desc_bytes = desc.size() * sizeof(B200MeBlock), res_bytes = (size_t)n * 10 * sizeof(B200MeResult);
  int rc = scratch_reserve(ctx, desc_bytes + 256 + res_bytes);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  B200MeBlock *d_desc = (B200MeBlock *)base;
  B200MeResult *d_res = (B200MeResult *)(base + ((desc_bytes + 255) & ~(size_t)255));
  B200MeResult *r_full0 = d_res, *r_sub0 = r_full0 + n, *r_full1 = r_sub0 + n, *r_sub1 = r_full1 + 4 * n;
  cudaStream_t s = ctx->stream;
  B200_CUDA_TRY(cudaMemcpyAsync(d_desc, desc.data(), desc_bytes, cudaMemcpyHostToDevice, s));

  B200MeLaunch l;
  memset(&l, 0, sizeof(l));
  l.highbd = cur->bd > 8, l.bd = cur->bd;
  l.cur = cur->d_plane, l.ref = ref->d_plane;
  l.cur_stride = l.ref_stride = cur->pitch;
  l.cur_rows = l.ref_rows = cur->rows;
  for (int i = 0; i < 4; ++i) l.mesh_range[i] = p->mesh_range[i], l.mesh_interval[i] = p->mesh_interval[i];
  l.families = p->search_method <= B200_ME_CLAMPED_DIAMOND ? B200_ME_FAMILY_NSTEP : B200_ME_FAMILY_PATTERN;
  l.cost_list_stride = sizeof(B200MeResult) / sizeof(int32_t);
  for (int stage = 0; stage < 2; ++stage) {
    const int cnt = stage ? 4 * n : n;
    B200MeLaunch fl = l;
    fl.n_blocks = cnt;
    fl.max_bw = stage ? p->bw >> 1 : p->bw, fl.max_bh = stage ? p->bh >> 1 : p->bh;
    fl.window_bytes = mesh_window_bytes(p->mesh_range, fl.max_bw, fl.max_bh, cur->bd > 8 ? 2 : 1);
    fl.blocks = d_desc + (stage ? 2 * n : 0);
    fl.results = stage ? r_full1 : r_full0;
    if (stage) fl.start_from = r_sub0, fl.start_group = 4;  // start_mv = get_fullmv_from_mv(ref_mv) (:198)
    rc = b200_me_fullpel_launch(ctx, &fl, s);
    if (rc != B200_OK) return rc;
    B200MeLaunch sl = fl;
    sl.blocks = d_desc + (stage ? 6 * n : n);
    sl.results = stage ? r_sub1 : r_sub0;
    sl.start_from = fl.results, sl.start_group = 1;
    sl.cost_lists = p->want_cost_list ? &fl.results->cost_list[0] : nullptr;
    rc = b200_me_subpel_launch(ctx, &sl, s);
    if (rc != B200_OK) return rc;
  }
  std::vector<B200MeResult> res((size_t)n * 10);
  B200_CUDA_TRY(cudaMemcpyAsync(res.data(), d_res, res_bytes, cudaMemcpyDeviceToHost, s));
  B200_CUDA_TRY(cudaStreamSynchronize(s));

  const B200MeResult *h_sub0 = res.data() + n, *h_sub1 = res.data() + 6 * n;
  const int min_dim = cur->width < cur->height ? cur->width : cur->height;
  const int pels = p->bw * p->bh, sub_pels = pels >> 2;
  for (int i = 0; i < n; ++i) {
    B200TfResult &o = out[i];
    // block_mse = DIVIDE_AND_ROUND(error, mb_pels) (:172)
    const int block_mse = (int)(((unsigned)h_sub0[i].ret + (unsigned)(pels >> 1)) / (unsigned)pels);
    o.block_mv[0] = h_sub0[i].best_row, o.block_mv[1] = h_sub0[i].best_col;
    o.block_mse = block_mse;
    long long sum = 0;
    int mn = INT32_MAX, mx = INT32_MIN;
    for (int k = 0; k < 4; ++k) {
      const B200MeResult &r = h_sub1[4 * i + k];
      o.sub_mv[k][0] = r.best_row, o.sub_mv[k][1] = r.best_col;
      o.sub_mse[k] = (int)(((unsigned)r.ret + (unsigned)(sub_pels >> 1)) / (unsigned)sub_pels);  // :234
      sum += o.sub_mse[k];
      mn = o.sub_mse[k] < mn ? o.sub_mse[k] : mn;
      mx = o.sub_mse[k] > mx ? o.sub_mse[k] : mx;
    }
    // tf_determine_block_partition (:275-297)
    if (((long long)block_mse * 15 < sum * 4 && mx - mn < 48) || ((long long)block_mse * 14 < sum * 4 && mx - mn < 24))
      for (int k = 0; k < 4; ++k) {
        o.sub_mv[k][0] = o.block_mv[0], o.sub_mv[k][1] = o.block_mv[1];
        o.sub_mse[k] = block_mse;
      }
    // do not pass the reference mv down when the error is too large (:255-259)
    const int thresh = (min_dim >= 720 ? 12 : 3) << (cur->bd - 8);
    o.ref_mv[0] = block_mse > thresh ? 0 : o.block_mv[0];
    o.ref_mv[1] = block_mse > thresh ? 0 : o.block_mv[1];
  }
  return B200_OK;
}

// ---------------------------------------------------------------------------- TPL ----
extern "C" void b200_queue_tpl_default_params(int width, int height, B200TplParams *p) {
  memset(p, 0, sizeof(*p));
  const int min_dim = width < height ? width : height;
  p->bw = p->bh = 16;                   // tpl_data->tpl_bsize_1d = 16 (tpl_model.c:158)
  p->border_in_pixels = 160;
  p->search_method = B200_ME_NSTEP;     // tpl_sf.search_method (speed_features.c init_tpl_sf)
  p->step_param = 0;                    // tpl_sf.reduce_first_step_size = 0
  p->forced_stop = B200_EIGHTH_PEL;     // tpl_sf.subpel_force_stop
  p->mv_cost_type = B200_MV_COST_ENTROPY;
  const int range[4] = { 64, 28, 15, 7 }, interval[4] = { 8, 4, 1, 1 };
  for (int i = 0; i < 4; ++i) p->mesh_range[i] = range[i], p->mesh_interval[i] = interval[i];
  p->force_mesh_thresh = min_dim >= 720 ? 1 << 25 : 1 << 20;
  p->sad_per_bit = 16, p->error_per_bit = 10;
  p->iters_per_step = 2;
  p->subpel_tree = B200_SUBPEL_TREE;
  p->want_cost_list = 0;                // cond_cost_list(cpi, cost_list): sf.mv_sf.use_fullpel_costlist (:283-289)
}

static int tpl_grid(int width, int height, const B200TplParams *p, int *rows, int *cols) {
  if (!p || p->bw < 4 || p->bh < 4 || (p->bw & (p->bw - 1)) || (p->bh & (p->bh - 1)) || p->bw > 128 || p->bh > 128)
    return B200_EINVAL;
  const int mi_rows = ((height + 7) & ~7) >> 2, mi_cols = ((width + 7) & ~7) >> 2;
  *rows = (mi_rows * 4 + p->bh - 1) / p->bh;  // for (mi_row = 0; mi_row < mi_rows; mi_row += mi_height) (:1376)
  *cols = (mi_cols * 4 + p->bw - 1) / p->bw;
  return B200_OK;
}

extern "C" int b200_queue_tpl_describe(int width, int height, int pitch, int origin, const B200TplParams *p,
                                       const int16_t *center_mvs, B200MeBlock *full, B200MeBlock *sub) {
  int rows, cols;
  if (tpl_grid(width, height, p, &rows, &cols) != B200_OK || !full || !sub) return -1;
  const int mi_rows = ((height + 7) & ~7) >> 2, mi_cols = ((width + 7) & ~7) >> 2;
  int n = 0;
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c, ++n) {
      const int mv_row = center_mvs ? center_mvs[2 * n] : 0, mv_col = center_mvs ? center_mvs[2 * n + 1] : 0;
      const Limits lim = block_limits(mi_rows, mi_cols, r * (p->bh >> 2), c * (p->bw >> 2), p->bh >> 2, p->bw >> 2,
                                      p->border_in_pixels);
      B200MeBlock d;
      memset(&d, 0, sizeof(d));
      d.src_offset = d.ref_offset = (uint32_t)(origin + r * p->bh * pitch + c * p->bw);
      d.bw = (uint8_t)p->bw, d.bh = (uint8_t)p->bh;
      d.sad_per_bit = p->sad_per_bit, d.error_per_bit = p->error_per_bit;
      d.force_mesh_thresh = p->force_mesh_thresh;
      d.mesh_mv_diff_threshold = 4;
      d.ref_mv_row = (int16_t)mv_row, d.ref_mv_col = (int16_t)mv_col;  // ref_mv = center_mv (:278)
      B200MeBlock f = d;
      put_limits(&f, fullpel_range(lim, mv_row, mv_col));
      f.start_row = (int16_t)rawpel(mv_row), f.start_col = (int16_t)rawpel(mv_col);  // :259
      f.search_method = (uint8_t)p->search_method;
      f.step_param = (uint8_t)p->step_param;
      f.mv_cost_type = (uint8_t)p->mv_cost_type;
      f.flags = (p->prune_mesh_search ? B200_ME_PRUNE_MESH : 0) |
                (p->use_downsampled_sad && p->bh >= 16 ? B200_ME_SKIP_SAD : 0) |
                (p->want_cost_list ? B200_ME_COST_LIST : 0);
      full[n] = f;
      B200MeBlock s = d;
      put_limits(&s, subpel_range(lim, mv_row, mv_col));
      s.search_method = (uint8_t)p->subpel_tree;
      s.mv_cost_type = B200_MV_COST_NONE;  // :294
      s.flags = (p->allow_hp ? B200_ME_ALLOW_HP : 0) | (p->want_cost_list ? B200_ME_COST_LIST : 0);
      s.subpel = (uint8_t)(p->forced_stop | (p->iters_per_step << 2) | (B200_USE_2_TAPS << 4));  // :292-293
      sub[n] = s;
    }
  return n;
}

extern "C" int b200_queue_tpl_motion_estimation(b200_ctx *ctx, const b200_frame *cur, const b200_frame *ref,
                                                const B200TplParams *p, const int16_t *center_mvs,
                                                B200TplMeResult *out) {
  if (!ctx || !cur || !ref || !p || !out) return B200_EINVAL;
  if (cur->width != ref->width || cur->height != ref->height || cur->pitch != ref->pitch ||
      cur->border != ref->border || cur->bd != ref->bd) {
    set_error("b200_queue_tpl_motion_estimation: the two frames differ in geometry");
    return B200_EINVAL;
  }
  int rows, cols;
  if (tpl_grid(cur->width, cur->height, p, &rows, &cols) != B200_OK) return B200_EINVAL;
  if (cur->border < p->border_in_pixels || cur->border < p->bw) {
    set_error("b200_queue_tpl_motion_estimation: frame border %d smaller than border_in_pixels %d", cur->border,
              p->border_in_pixels);
    return B200_EINVAL;
  }
  B200_CUDA_TRY(cudaSetDevice(ctx->device));
  const int n = rows * cols;
  std::vector<B200MeBlock> desc((size_t)2 * n);
  if (b200_queue_tpl_describe(cur->width, cur->height, cur->pitch, cur->origin, p, center_mvs, desc.data(),
                              desc.data() + n) != n)
    return B200_EINVAL;
  const bool costs = p->mvjcost && p->mvcost[0] && p->mvcost[1];
  const size_t tab = (size_t)(2 * 16383 + 1) * sizeof(int);
  size_t off = 0;
  auto carve = [&](size_t bytes) {
    const size_t o = off;
    off += (bytes + 255) & ~(size_t)255;
    return o;
  };
  const size_t o_desc = carve(desc.size() * sizeof(B200MeBlock)), o_res = carve((size_t)2 * n * sizeof(B200MeResult));
  const size_t o_j = carve(costs ? 16 : 0), o_c0 = carve(costs ? tab : 0), o_c1 = carve(costs ? tab : 0);
  int rc = scratch_reserve(ctx, off);
  if (rc != B200_OK) return rc;
  char *base = (char *)ctx->d_scratch;
  cudaStream_t s = ctx->stream;
  B200_CUDA_TRY(cudaMemcpyAsync(base + o_desc, desc.data(), desc.size() * sizeof(B200MeBlock), cudaMemcpyHostToDevice, s));
  B200MeLaunch l;
  memset(&l, 0, sizeof(l));
  if (costs) {
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_j, p->mvjcost, 16, cudaMemcpyHostToDevice, s));
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_c0, p->mvcost[0] - 16383, tab, cudaMemcpyHostToDevice, s));
    B200_CUDA_TRY(cudaMemcpyAsync(base + o_c1, p->mvcost[1] - 16383, tab, cudaMemcpyHostToDevice, s));
    l.mvjcost = (const int *)(base + o_j);
    l.mvcost[0] = (const int *)(base + o_c0) + 16383;
    l.mvcost[1] = (const int *)(base + o_c1) + 16383;
  }
  l.highbd = cur->bd > 8, l.bd = cur->bd;
  l.cur = cur->d_plane, l.ref = ref->d_plane;
  l.cur_stride = l.ref_stride = cur->pitch;
  l.cur_rows = l.ref_rows = cur->rows;
  for (int i = 0; i < 4; ++i) l.mesh_range[i] = p->mesh_range[i], l.mesh_interval[i] = p->mesh_interval[i];
  l.families = p->search_method <= B200_ME_CLAMPED_DIAMOND ? B200_ME_FAMILY_NSTEP : B200_ME_FAMILY_PATTERN;
  l.n_blocks = n;
  l.max_bw = p->bw, l.max_bh = p->bh;
  l.window_bytes = mesh_window_bytes(p->mesh_range, p->bw, p->bh, cur->bd > 8 ? 2 : 1);
  l.blocks = (const B200MeBlock *)(base + o_desc);
  l.results = (B200MeResult *)(base + o_res);
  rc = b200_me_fullpel_launch(ctx, &l, s);
  if (rc != B200_OK) return rc;
  B200MeLaunch sl = l;
  sl.blocks = l.blocks + n;
  sl.results = l.results + n;
  sl.start_from = l.results;  // subpel_start_mv = get_mv_from_fullmv(&best_mv->as_fullmv) (:295)
  sl.cost_lists = p->want_cost_list ? &l.results->cost_list[0] : nullptr;
  sl.cost_list_stride = sizeof(B200MeResult) / sizeof(int32_t);
  rc = b200_me_subpel_launch(ctx, &sl, s);
  if (rc != B200_OK) return rc;
  std::vector<B200MeResult> res((size_t)n);
  B200_CUDA_TRY(cudaMemcpyAsync(res.data(), sl.results, (size_t)n * sizeof(B200MeResult), cudaMemcpyDeviceToHost, s));
  B200_CUDA_TRY(cudaStreamSynchronize(s));
  for (int i = 0; i < n; ++i) {
    out[i].mv[0] = res[i].best_row, out[i].mv[1] = res[i].best_col;
    out[i].sme = (uint32_t)res[i].ret;
  }
  return B200_OK;
}

extern "C" int b200_queue_tpl_txfm_quant_rdcost(b200_ctx *ctx, const void *src, int src_stride, const void *pred,
                                                int pred_stride, int rows, int cols, int bd, int tx_size,
                                                const B200QRow *qrow, B200TplRdResult *out) {
  if (!ctx || !src || !pred || !qrow || !out || rows <= 0 || cols <= 0) return B200_EINVAL;
  if (tx_size < B200_TX_4X4 || tx_size > B200_TX_32X32 || (bd != 8 && bd != 10 && bd != 12)) {
    set_error("b200_queue_tpl_txfm_quant_rdcost: square tx sizes 4x4..32x32 and bd 8/10/12 only");
    return B200_EINVAL;
  }
  const int w = tx_info(tx_size).w, n = rows * cols, per = w * w;
  if (src_stride < cols * w || pred_stride < cols * w || src_stride > 65535 || pred_stride > 65535) return B200_EINVAL;
  std::vector<B200TxBlock> blocks((size_t)n);
  std::vector<B200TxPred> preds((size_t)n);
  for (int r = 0; r < rows; ++r)
    for (int c = 0; c < cols; ++c) {
      const int i = r * cols + c;
      memset(&blocks[i], 0, sizeof(B200TxBlock));
      memset(&preds[i], 0, sizeof(B200TxPred));
      blocks[i].src_offset = (uint32_t)(r * w * src_stride + c * w);
      blocks[i].out_offset = (uint32_t)i * per;
      blocks[i].src_stride = (uint16_t)src_stride;
      blocks[i].tx_size = (uint8_t)tx_size;
      blocks[i].tx_type = 0;  // DCT_DCT: av1_quick_txfm (hybrid_fwd_txfm.c), scan_order of DCT_DCT (:103)
      preds[i].pred_offset = (uint32_t)(r * w * pred_stride + c * w);
      preds[i].pred_stride = (uint16_t)pred_stride;
    }
  const bool hbd = bd > 8;
  const size_t out_elems = (size_t)n * per;
  std::vector<unsigned char> levels(out_elems * (hbd ? 4 : 2));
  std::vector<uint32_t> offs((size_t)n + 1);
  std::vector<uint16_t> eob((size_t)n);
  std::vector<int64_t> dist((size_t)2 * n);
  const int rc = b200_txq_batch_host_compact(
      ctx, hbd ? B200_IN_PIXELS16 : B200_IN_PIXELS8, hbd ? B200_Q_FP_HBD : B200_Q_FP, bd, B200_TXQ_BLOCK_ERROR,
      blocks.data(), preds.data(), n, src, (size_t)src_stride * rows * w, pred, (size_t)pred_stride * rows * w, qrow, 1,
      out_elems, levels.data(), out_elems, offs.data(), eob.data(), dist.data(), nullptr, nullptr);
  if (rc != B200_OK) return rc;
  const int shift = tx_size == B200_TX_32X32 ? 0 : 2;  // get_quantize_error (:107)
  for (int i = 0; i < n; ++i) {
    // rate_estimator (:207-226): the levels are already in scan order
    int rate = 1;
    for (int k = 0; k < eob[i]; ++k) {
      const int q = hbd ? ((const int32_t *)levels.data())[offs[i] + k] : ((const int16_t *)levels.data())[offs[i] + k];
      const unsigned a = (unsigned)(q < 0 ? -q : q);
      int msb = 0;
      for (unsigned v = a + 1; v >>= 1;) ++msb;  // get_msb(abs_level + 1)
      rate += msb + 1 + (a > 0);
    }
    out[i].rate_cost = rate << 9;  // AV1_PROB_COST_SHIFT
    out[i].eob = eob[i];
    const int64_t err = dist[2 * i] >> shift, sse = dist[2 * i + 1] >> shift;
    out[i].recon_error = err > 1 ? err : 1;
    out[i].sse = sse > 1 ? sse : 1;
  }
  return B200_OK;
}
