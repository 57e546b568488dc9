// rtcd.cu -- drop-in tier: the reference's RTCD prototypes (`_b200` flavour), each
// running the batched kernels on a batch of ONE block with host buffers.
// See include/b200av1_rtcd.h.  Latency-bound by construction (H2D + launch + D2H per
// call); the batched tier is the throughput path.
#include <stdlib.h>
#include <string.h>

#include <mutex>

#include "../../include/b200av1_rtcd.h"
#include "internal.h"

using namespace b200;

namespace {

std::mutex g_mu;  // drop-in calls share the default context's scratch

[[noreturn]] void fatal(const char *where) {
  fprintf(stderr, "libb200av1: fatal in %s: %s\n", where, b200_last_error());
  abort();  // RTCD signatures have no error channel
}
#define CK(expr)                                   \
  do {                                             \
    if ((expr) != B200_OK) fatal(__func__);        \
  } while (0)
#define CKCUDA(expr)                                                              \
  do {                                                                            \
    cudaError_t e_ = (expr);                                                      \
    if (e_ != cudaSuccess) {                                                      \
      set_error("%s: %s", #expr, cudaGetErrorString(e_));                         \
      fatal(__func__);                                                            \
    }                                                                             \
  } while (0)

// bump allocator over the context's device scratch
struct Arena {
  b200_ctx *ctx;
  size_t off = 0;
  explicit Arena(b200_ctx *c, size_t need) : ctx(c) {
    if (scratch_reserve(c, need + 4096) != B200_OK) fatal("scratch");
  }
  template <class T>
  T *take(size_t count) {
    T *p = reinterpret_cast<T *>(static_cast<char *>(ctx->d_scratch) + off);
    off += (count * sizeof(T) + 255) & ~(size_t)255;
    return p;
  }
};

b200_ctx *ctx0() {
  b200_ctx *c = b200_default_ctx();
  if (cudaSetDevice(c->device) != cudaSuccess) fatal("cudaSetDevice");
  return c;
}

int tx_size_of(int w, int h) {
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    if (b200_tx_wide(ts) == w && b200_tx_high(ts) == h) return ts;
  return -1;
}

// ------------------------------------------------------------------ transforms ----
void fwd_txfm_one(int ts, const int16_t *input, int32_t *output, int stride, int tx_type) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const int w = b200_tx_wide(ts), h = b200_tx_high(ts), n = b200_tx_max_eob(ts);
  if (!b200_tx_valid(ts, tx_type)) {
    set_error("av1_fwd_txfm2d_%dx%d_b200: tx_type %d has no %dx%d transform", w, h, tx_type, w, h);
    fatal(__func__);
  }
  Arena a(c, (size_t)w * h * 2 + (size_t)n * 4 + 1024);
  int16_t *d_src = a.take<int16_t>((size_t)w * h);
  int32_t *d_out = a.take<int32_t>(n);
  B200TxBlock *d_blk = a.take<B200TxBlock>(1);
  B200TxBlock blk;
  memset(&blk, 0, sizeof(blk));
  blk.src_stride = (uint16_t)w;
  blk.tx_size = (uint8_t)ts;
  blk.tx_type = (uint8_t)tx_type;
  CKCUDA(cudaMemcpy2DAsync(d_src, (size_t)w * 2, input, (size_t)stride * 2, (size_t)w * 2, h,
                           cudaMemcpyHostToDevice, c->stream));
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, c->stream));
  B200TxqLaunch l;
  memset(&l, 0, sizeof(l));
  l.tx_size = ts;
  l.in_kind = B200_IN_RESIDUAL;
  l.q_kind = B200_Q_NONE;
  l.bd = 8;
  l.flags = B200_TXQ_STORE_COEFF;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.src = d_src;
  l.coeff = d_out;
  CK(txq_launch(c, l, c->stream));
  CKCUDA(cudaMemcpyAsync(output, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
  CKCUDA(cudaStreamSynchronize(c->stream));
}

// ------------------------------------------------------------------ quantisers ----
void quant_one(int kind, const void *coeff, intptr_t n, const int16_t *zbin, const int16_t *round,
               const int16_t *quant, const int16_t *quant_shift, void *qcoeff, void *dqcoeff,
               const int16_t *dequant, uint16_t *eob, const int16_t *iscan, const uint8_t *qm,
               const uint8_t *iqm, int log_scale) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t esz = kind == B200_QUANT_LP ? 2 : 4;
  Arena a(c, (size_t)n * (esz * 3 + 2 + 2) + 4096);
  char *d_coeff = a.take<char>((size_t)n * esz);
  char *d_q = a.take<char>((size_t)n * esz);
  char *d_dq = a.take<char>((size_t)n * esz);
  int16_t *d_iscan = a.take<int16_t>(n);
  uint8_t *d_qm = qm ? a.take<uint8_t>(n) : nullptr;
  uint8_t *d_iqm = iqm ? a.take<uint8_t>(n) : nullptr;
  B200QRow *d_row = a.take<B200QRow>(1);
  B200QuantBlock *d_blk = a.take<B200QuantBlock>(1);
  uint16_t *d_eob = a.take<uint16_t>(1);
  B200QRow row;
  memset(&row, 0, sizeof(row));
  const bool fp = kind == B200_QUANT_FP || kind == B200_QUANT_FP_HBD || kind == B200_QUANT_LP;
  for (int k = 0; k < 2; ++k) {
    if (fp) {  // the FP entry points receive round_fp / quant_fp in the round / quant slots
      row.round_fp[k] = round[k];
      row.quant_fp[k] = quant[k];
    } else {
      row.round[k] = round[k];
      row.quant[k] = quant[k];
    }
    if (zbin) row.zbin[k] = zbin[k];
    if (quant_shift) row.quant_shift[k] = quant_shift[k];
    row.dequant[k] = dequant[k];
  }
  B200QuantBlock blk;
  memset(&blk, 0, sizeof(blk));
  blk.n_coeffs = (uint16_t)n;
  blk.log_scale = (uint8_t)log_scale;
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpyAsync(d_coeff, coeff, (size_t)n * esz, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_iscan, iscan, (size_t)n * 2, cudaMemcpyHostToDevice, s));
  if (qm) CKCUDA(cudaMemcpyAsync(d_qm, qm, n, cudaMemcpyHostToDevice, s));
  if (iqm) CKCUDA(cudaMemcpyAsync(d_iqm, iqm, n, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_row, &row, sizeof(row), cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, s));
  B200QuantLaunch l;
  memset(&l, 0, sizeof(l));
  l.kind = kind;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.coeff = d_coeff;
  l.qrows = d_row;
  l.iscan = d_iscan;
  l.qm = d_qm;
  l.iqm = d_iqm;
  l.qcoeff = d_q;
  l.dqcoeff = d_dq;
  l.eob = d_eob;
  CK(b200_quant_launch(c, &l, s));
  CKCUDA(cudaMemcpyAsync(qcoeff, d_q, (size_t)n * esz, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaMemcpyAsync(dqcoeff, d_dq, (size_t)n * esz, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaMemcpyAsync(eob, d_eob, 2, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

int64_t block_error_one(int kind, const void *coeff, const void *dqcoeff, intptr_t n, int64_t *ssz,
                        int bd) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t esz = kind == B200_ERR_LP ? 2 : 4;
  Arena a(c, (size_t)n * esz * 2 + 4096);
  char *d_c = a.take<char>((size_t)n * esz);
  char *d_q = a.take<char>((size_t)n * esz);
  B200ErrBlock *d_blk = a.take<B200ErrBlock>(1);
  int64_t *d_out = a.take<int64_t>(2);
  B200ErrBlock blk = { 0, 0, (uint32_t)n, 0 };
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpyAsync(d_c, coeff, (size_t)n * esz, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_q, dqcoeff, (size_t)n * esz, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, s));
  B200ErrLaunch l;
  memset(&l, 0, sizeof(l));
  l.kind = kind;
  l.bd = bd;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.coeff = d_c;
  l.dqcoeff = d_q;
  l.out = d_out;
  CK(b200_block_error_launch(c, &l, s));
  int64_t out[2];
  CKCUDA(cudaMemcpyAsync(out, d_out, 16, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
  if (ssz) *ssz = out[1];
  return out[0];
}

// ---------------------------------------------------------------- pixel blocks ----
// Copies the w x h (plus margins) windows of `a` and up to four `b` operands to the
// device as dense tiles and runs one pix launch over them.
struct PixResult {
  uint32_t v[4][2];
};

const void *unbyte(const uint8_t *p, int highbd) {
  // CONVERT_TO_SHORTPTR (aom_ports/mem.h:79): highbd pointers travel as (addr >> 1)
  return highbd ? reinterpret_cast<const void *>(reinterpret_cast<uintptr_t>(p) << 1)
                : static_cast<const void *>(p);
}

PixResult pix_run(int op, int highbd, int bd, int w, int h, const uint8_t *a8, int a_stride,
                  const uint8_t *const *b8, int n_b, int b_stride, int xoff, int yoff,
                  const uint8_t *aux8) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t px = highbd ? 2 : 1;
  // sub-pel variance interpolates `a`: it needs one extra row and column
  const int aw = w + (op == B200_PIX_SUBPEL_VAR ? 1 : 0), ah = h + (op == B200_PIX_SUBPEL_VAR ? 1 : 0);
  Arena ar(c, ((size_t)aw * ah + (size_t)w * h * (n_b + 1)) * px + 8192);
  char *d_a = ar.take<char>((size_t)aw * ah * px);
  char *d_b = ar.take<char>((size_t)w * h * px * n_b);
  char *d_aux = aux8 ? ar.take<char>((size_t)w * h * px) : nullptr;
  B200PixBlock *d_blk = ar.take<B200PixBlock>(n_b);
  uint32_t *d_out = ar.take<uint32_t>(2 * n_b);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpy2DAsync(d_a, aw * px, unbyte(a8, highbd), a_stride * px, aw * px, ah,
                           cudaMemcpyHostToDevice, s));
  B200PixBlock blk[4];
  memset(blk, 0, sizeof(blk));
  for (int i = 0; i < n_b; ++i) {
    CKCUDA(cudaMemcpy2DAsync(d_b + (size_t)i * w * h * px, w * px, unbyte(b8[i], highbd),
                             b_stride * px, w * px, h, cudaMemcpyHostToDevice, s));
    blk[i].a_offset = 0;
    blk[i].b_offset = (uint32_t)((size_t)i * w * h);
    blk[i].a_stride = (uint16_t)aw;
    blk[i].b_stride = (uint16_t)w;
    blk[i].w = (uint8_t)w;
    blk[i].h = (uint8_t)h;
    blk[i].xoff = (uint8_t)xoff;
    blk[i].yoff = (uint8_t)yoff;
  }
  if (aux8)
    CKCUDA(cudaMemcpyAsync(d_aux, unbyte(aux8, highbd), (size_t)w * h * px, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_blk, blk, sizeof(B200PixBlock) * n_b, cudaMemcpyHostToDevice, s));
  B200PixLaunch l;
  memset(&l, 0, sizeof(l));
  l.op = op;
  l.highbd = highbd;
  l.bd = bd;
  l.n_blocks = n_b;
  l.blocks = d_blk;
  l.a = d_a;
  l.b = d_b;
  l.aux = d_aux;
  l.out = d_out;
  CK(b200_pix_launch(c, &l, s));
  PixResult r;
  memset(&r, 0, sizeof(r));
  CKCUDA(cudaMemcpyAsync(r.v, d_out, sizeof(uint32_t) * 2 * n_b, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
  return r;
}

// SAD family: a = src, b = ref
unsigned sad1(int op, int hb, int w, int h, const uint8_t *src, int ss, const uint8_t *ref, int rs,
              const uint8_t *second = nullptr) {
  // 128-wide / tall blocks exceed the uint8 w/h fields only at 256; 128 fits
  return pix_run(op, hb, 8, w, h, src, ss, &ref, 1, rs, 0, 0, second).v[0][0];
}
void sad4(int op, int hb, int w, int h, const uint8_t *src, int ss, const uint8_t *const ref[4],
          int rs, uint32_t out[4]) {
  const PixResult r = pix_run(op, hb, 8, w, h, src, ss, ref, 4, rs, 0, 0, nullptr);
  for (int i = 0; i < 4; ++i) out[i] = r.v[i][0];
}
// variance family: reference argument order (src_ptr, src_stride, ref_ptr, ref_stride)
unsigned var1(int op, int hb, int bd, int w, int h, const uint8_t *a, int as, int xo, int yo,
              const uint8_t *b, int bs, unsigned *sse) {
  const PixResult r = pix_run(op, hb, bd, w, h, a, as, &b, 1, bs, xo, yo, nullptr);
  if (sse) *sse = r.v[0][1];
  return r.v[0][0];
}

void upsample_one(int hb, int bd, uint8_t *comp_pred8, int w, int h, int sx, int sy,
                  const uint8_t *ref8, int ref_stride, int subpel_search) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  // SUBPEL_SEARCH_TYPE (av1/encoder/mcomp.h): USE_2_TAPS = 1, USE_4_TAPS = 2, USE_8_TAPS = 3
  const int taps = subpel_search <= 1 ? 2 : (subpel_search == 2 ? 4 : 8);
  const size_t px = hb ? 2 : 1;
  // the reference touches taps -3..+4 only in the directions it filters
  const int left = sx ? 3 : 0, right = sx ? 4 : 0, top = sy ? 3 : 0, bottom = sy ? 4 : 0;
  const int ww = w + left + right, hh = h + top + bottom;
  Arena ar(c, ((size_t)ww * hh + (size_t)w * h) * px + 4096);
  char *d_ref = ar.take<char>((size_t)ww * hh * px);
  char *d_out = ar.take<char>((size_t)w * h * px);
  B200PixBlock *d_blk = ar.take<B200PixBlock>(1);
  cudaStream_t s = c->stream;
  const char *href = static_cast<const char *>(unbyte(ref8, hb)) -
                     ((ptrdiff_t)top * ref_stride + left) * (ptrdiff_t)px;
  CKCUDA(cudaMemcpy2DAsync(d_ref, ww * px, href, ref_stride * px, ww * px, hh,
                           cudaMemcpyHostToDevice, s));
  B200PixBlock blk;
  memset(&blk, 0, sizeof(blk));
  blk.a_offset = (uint32_t)(top * ww + left);
  blk.a_stride = (uint16_t)ww;
  blk.w = (uint8_t)w;
  blk.h = (uint8_t)h;
  blk.xoff = (uint8_t)sx;
  blk.yoff = (uint8_t)sy;
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, s));
  B200UpsampleLaunch l;
  memset(&l, 0, sizeof(l));
  l.highbd = hb;
  l.bd = bd;
  l.taps = taps;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.ref = d_ref;
  l.comp_pred = d_out;
  CK(b200_upsampled_pred_launch(c, &l, s));
  void *hout = const_cast<void *>(unbyte(comp_pred8, hb));
  CKCUDA(cudaMemcpyAsync(hout, d_out, (size_t)w * h * px, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

void subtract_one(int hb, int rows, int cols, int16_t *diff, ptrdiff_t diff_stride,
                  const uint8_t *src8, ptrdiff_t ss, const uint8_t *pred8, ptrdiff_t ps) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t px = hb ? 2 : 1;
  Arena ar(c, (size_t)rows * cols * (2 * px + 2) + 4096);
  char *d_a = ar.take<char>((size_t)rows * cols * px);
  char *d_b = ar.take<char>((size_t)rows * cols * px);
  int16_t *d_diff = ar.take<int16_t>((size_t)rows * cols);
  B200PixBlock *d_blk = ar.take<B200PixBlock>(1);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpy2DAsync(d_a, cols * px, unbyte(src8, hb), ss * px, cols * px, rows,
                           cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpy2DAsync(d_b, cols * px, unbyte(pred8, hb), ps * px, cols * px, rows,
                           cudaMemcpyHostToDevice, s));
  B200PixBlock blk;
  memset(&blk, 0, sizeof(blk));
  blk.a_stride = blk.b_stride = (uint16_t)cols;
  blk.w = (uint8_t)cols;
  blk.h = (uint8_t)rows;
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, s));
  B200SubtractLaunch l;
  memset(&l, 0, sizeof(l));
  l.highbd = hb;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.a = d_a;
  l.b = d_b;
  l.diff = d_diff;
  CK(b200_subtract_launch(c, &l, s));
  CKCUDA(cudaMemcpy2DAsync(diff, diff_stride * 2, d_diff, cols * 2, cols * 2, rows,
                           cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

// ------------------------------------------------------------ inverse transforms ----
// one block: dst (host, uint8 or uint16) += inverse(dqcoeff), through b200_itx_launch in place
void inv_txfm_one(int ts, int tx_type, const int32_t *dqcoeff, void *dst, int stride, int bd, int highbd,
                  int lossless, int eob) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const int w = b200_tx_wide(ts), h = b200_tx_high(ts), n = lossless ? 16 : b200_tx_max_eob(ts);
  const size_t px = highbd ? 2 : 1;
  if (!lossless && !b200_tx_valid(ts, tx_type)) {
    set_error("inverse transform: tx_type %d has no %dx%d transform", tx_type, w, h);
    fatal(__func__);
  }
  if (eob == 0) return;  // av1_inverse_transform_block (idct.c:307)
  Arena a(c, (size_t)w * h * px + (size_t)n * 4 + 1024);
  char *d_dst = a.take<char>((size_t)w * h * px);
  int32_t *d_dq = a.take<int32_t>(n);
  B200TxBlock *d_blk = a.take<B200TxBlock>(1);
  B200TxPred *d_pred = a.take<B200TxPred>(1);
  uint16_t *d_eob = a.take<uint16_t>(1);
  B200TxBlock blk;
  memset(&blk, 0, sizeof(blk));
  blk.tx_size = (uint8_t)ts;
  blk.tx_type = (uint8_t)tx_type;
  B200TxPred pd;
  memset(&pd, 0, sizeof(pd));
  pd.pred_stride = (uint16_t)w;
  const uint16_t e16 = (uint16_t)(eob < 0 ? n : eob);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpy2DAsync(d_dst, w * px, dst, (size_t)stride * px, w * px, h, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_dq, dqcoeff, (size_t)n * 4, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_blk, &blk, sizeof(blk), cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_pred, &pd, sizeof(pd), cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_eob, &e16, sizeof(e16), cudaMemcpyHostToDevice, s));
  B200ItxLaunch l;
  memset(&l, 0, sizeof(l));
  l.tx_size = ts;
  l.highbd = highbd;
  l.bd = bd;
  l.n_blocks = 1;
  l.blocks = d_blk;
  l.preds = d_pred;
  l.dqcoeff = d_dq;
  l.eob = d_eob;
  l.pred = d_dst;
  l.recon = d_dst;
  l.lossless = lossless;
  CK(itx_launch(c, l, s));
  CKCUDA(cudaMemcpy2DAsync(dst, (size_t)stride * px, d_dst, w * px, w * px, h, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

}  // namespace

// ================================================================== exported =====
extern "C" {

#define B200_DEF_FWD_TXFM(W, H)                                                              \
  void av1_fwd_txfm2d_##W##x##H##_b200(const int16_t *input, int32_t *output, int stride,    \
                                       uint8_t tx_type, int bd) {                            \
    (void)bd; /* only feeds debug range checks in the reference (av1_fwd_txfm2d.c:42-54) */ \
    fwd_txfm_one(tx_size_of(W, H), input, output, stride, tx_type);                          \
  }
B200_DEF_FWD_TXFM(4, 4)
B200_DEF_FWD_TXFM(8, 8)
B200_DEF_FWD_TXFM(16, 16)
B200_DEF_FWD_TXFM(32, 32)
B200_DEF_FWD_TXFM(64, 64)
B200_DEF_FWD_TXFM(4, 8)
B200_DEF_FWD_TXFM(8, 4)
B200_DEF_FWD_TXFM(8, 16)
B200_DEF_FWD_TXFM(16, 8)
B200_DEF_FWD_TXFM(16, 32)
B200_DEF_FWD_TXFM(32, 16)
B200_DEF_FWD_TXFM(32, 64)
B200_DEF_FWD_TXFM(64, 32)
B200_DEF_FWD_TXFM(4, 16)
B200_DEF_FWD_TXFM(16, 4)
B200_DEF_FWD_TXFM(8, 32)
B200_DEF_FWD_TXFM(32, 8)
B200_DEF_FWD_TXFM(16, 64)
B200_DEF_FWD_TXFM(64, 16)

void av1_fwht4x4_b200(const int16_t *input, b200_tran_low_t *output, int stride);
void av1_lowbd_fwd_txfm_b200(const int16_t *src_diff, b200_tran_low_t *coeff, int diff_stride,
                             B200TxfmParam *p) {
  if (p->lossless) {  // av1_lowbd_fwd_txfm_c -> av1_highbd_fwd_txfm -> highbd_fwd_txfm_4x4 (hybrid_fwd_txfm.c:85-93)
    av1_fwht4x4_b200(src_diff, coeff, diff_stride);
    return;
  }
  fwd_txfm_one(p->tx_size, src_diff, coeff, diff_stride, p->tx_type);
}

#define B200_DEF_INV_TXFM(W, H)                                                                  \
  void av1_inv_txfm2d_add_##W##x##H##_b200(const int32_t *input, uint16_t *output, int stride,  \
                                           uint8_t tx_type, int bd) {                            \
    inv_txfm_one(tx_size_of(W, H), tx_type, input, output, stride, bd, 1, 0, -1);                \
  }
B200_DEF_INV_TXFM(4, 4)
B200_DEF_INV_TXFM(8, 8)
B200_DEF_INV_TXFM(16, 16)
B200_DEF_INV_TXFM(32, 32)
B200_DEF_INV_TXFM(64, 64)
B200_DEF_INV_TXFM(4, 8)
B200_DEF_INV_TXFM(8, 4)
B200_DEF_INV_TXFM(8, 16)
B200_DEF_INV_TXFM(16, 8)
B200_DEF_INV_TXFM(16, 32)
B200_DEF_INV_TXFM(32, 16)
B200_DEF_INV_TXFM(32, 64)
B200_DEF_INV_TXFM(64, 32)
B200_DEF_INV_TXFM(4, 16)
B200_DEF_INV_TXFM(16, 4)
B200_DEF_INV_TXFM(8, 32)
B200_DEF_INV_TXFM(32, 8)
B200_DEF_INV_TXFM(16, 64)
B200_DEF_INV_TXFM(64, 16)
// av1_inv_txfm_add (idct.c:283) / av1_highbd_inv_txfm_add (:216): dst holds the prediction
void av1_inv_txfm_add_b200(const b200_tran_low_t *dqcoeff, uint8_t *dst, int stride, const B200TxfmParam *p) {
  inv_txfm_one(p->tx_size, p->tx_type, dqcoeff, dst, stride, p->bd, 0, p->lossless, p->lossless ? p->eob : -1);
}
void av1_highbd_inv_txfm_add_b200(const b200_tran_low_t *input, uint8_t *dest8, int stride,
                                  const B200TxfmParam *p) {
  // CONVERT_TO_SHORTPTR (aom_ports/mem.h:79)
  inv_txfm_one(p->tx_size, p->tx_type, input, (void *)(((uintptr_t)dest8) << 1), stride, p->bd, 1, p->lossless,
               p->lossless ? p->eob : -1);
}

#define QA coeff_ptr, n_coeffs, zbin_ptr, round_ptr, quant_ptr, quant_shift_ptr, qcoeff_ptr, \
           dqcoeff_ptr, dequant_ptr, eob_ptr, iscan
#define B200_DEF_QUANT(NAME, KIND, LS)                         \
  void NAME(B200_QUANT_ARGS) {                                 \
    (void)scan;                                                \
    quant_one(KIND, QA, nullptr, nullptr, LS);                 \
  }
B200_DEF_QUANT(av1_quantize_fp_b200, B200_QUANT_FP, 0)
B200_DEF_QUANT(av1_quantize_fp_32x32_b200, B200_QUANT_FP, 1)
B200_DEF_QUANT(av1_quantize_fp_64x64_b200, B200_QUANT_FP, 2)
B200_DEF_QUANT(aom_quantize_b_b200, B200_QUANT_B, 0)
B200_DEF_QUANT(aom_quantize_b_32x32_b200, B200_QUANT_B, 1)
B200_DEF_QUANT(aom_quantize_b_64x64_b200, B200_QUANT_B, 2)
B200_DEF_QUANT(aom_quantize_b_adaptive_b200, B200_QUANT_B_ADAPT, 0)
B200_DEF_QUANT(aom_quantize_b_32x32_adaptive_b200, B200_QUANT_B_ADAPT, 1)
B200_DEF_QUANT(aom_quantize_b_64x64_adaptive_b200, B200_QUANT_B_ADAPT, 2)
B200_DEF_QUANT(aom_highbd_quantize_b_b200, B200_QUANT_B_HBD, 0)
B200_DEF_QUANT(aom_highbd_quantize_b_32x32_b200, B200_QUANT_B_HBD, 1)
B200_DEF_QUANT(aom_highbd_quantize_b_64x64_b200, B200_QUANT_B_HBD, 2)
B200_DEF_QUANT(aom_highbd_quantize_b_adaptive_b200, B200_QUANT_B_ADAPT_HBD, 0)
B200_DEF_QUANT(aom_highbd_quantize_b_32x32_adaptive_b200, B200_QUANT_B_ADAPT_HBD, 1)
B200_DEF_QUANT(aom_highbd_quantize_b_64x64_adaptive_b200, B200_QUANT_B_ADAPT_HBD, 2)

void av1_highbd_quantize_fp_b200(B200_QUANT_ARGS, int log_scale) {
  (void)scan;
  quant_one(B200_QUANT_FP_HBD, QA, nullptr, nullptr, log_scale);
}
void av1_quantize_lp_b200(const int16_t *coeff_ptr, intptr_t n_coeffs, const int16_t *round_ptr,
                          const int16_t *quant_ptr, int16_t *qcoeff_ptr, int16_t *dqcoeff_ptr,
                          const int16_t *dequant_ptr, uint16_t *eob_ptr, const int16_t *scan,
                          const int16_t *iscan) {
  (void)scan;
  quant_one(B200_QUANT_LP, coeff_ptr, n_coeffs, nullptr, round_ptr, quant_ptr, nullptr, qcoeff_ptr,
            dqcoeff_ptr, dequant_ptr, eob_ptr, iscan, nullptr, nullptr, 0);
}
void aom_quantize_b_helper_b200(B200_QUANT_ARGS, const b200_qm_val_t *qm_ptr,
                                const b200_qm_val_t *iqm_ptr, const int log_scale) {
  (void)scan;
  quant_one(B200_QUANT_B, QA, qm_ptr, iqm_ptr, log_scale);
}
void aom_highbd_quantize_b_helper_b200(B200_QUANT_ARGS, const b200_qm_val_t *qm_ptr,
                                       const b200_qm_val_t *iqm_ptr, const int log_scale) {
  (void)scan;
  quant_one(B200_QUANT_B_HBD, QA, qm_ptr, iqm_ptr, log_scale);
}

int64_t av1_block_error_b200(const b200_tran_low_t *coeff, const b200_tran_low_t *dqcoeff,
                             intptr_t block_size, int64_t *ssz) {
  return block_error_one(B200_ERR_LOWBD, coeff, dqcoeff, block_size, ssz, 8);
}
int64_t av1_block_error_lp_b200(const int16_t *coeff, const int16_t *dqcoeff, intptr_t block_size) {
  return block_error_one(B200_ERR_LP, coeff, dqcoeff, block_size, nullptr, 8);
}
int64_t av1_highbd_block_error_b200(const b200_tran_low_t *coeff, const b200_tran_low_t *dqcoeff,
                                    intptr_t block_size, int64_t *ssz, int bd) {
  return block_error_one(B200_ERR_HBD, coeff, dqcoeff, block_size, ssz, bd);
}

void aom_subtract_block_b200(int rows, int cols, int16_t *diff_ptr, ptrdiff_t diff_stride,
                             const uint8_t *src_ptr, ptrdiff_t src_stride, const uint8_t *pred_ptr,
                             ptrdiff_t pred_stride) {
  subtract_one(0, rows, cols, diff_ptr, diff_stride, src_ptr, src_stride, pred_ptr, pred_stride);
}
void aom_highbd_subtract_block_b200(int rows, int cols, int16_t *diff_ptr, ptrdiff_t diff_stride,
                                    const uint8_t *src_ptr, ptrdiff_t src_stride,
                                    const uint8_t *pred_ptr, ptrdiff_t pred_stride) {
  subtract_one(1, rows, cols, diff_ptr, diff_stride, src_ptr, src_stride, pred_ptr, pred_stride);
}
int64_t aom_sse_b200(const uint8_t *a, int a_stride, const uint8_t *b, int b_stride, int width,
                     int height) {
  const PixResult r = pix_run(B200_PIX_SSE, 0, 8, width, height, a, a_stride, &b, 1, b_stride, 0, 0, nullptr);
  return (int64_t)(((uint64_t)r.v[0][1] << 32) | r.v[0][0]);
}
int64_t aom_highbd_sse_b200(const uint8_t *a8, int a_stride, const uint8_t *b8, int b_stride,
                            int width, int height) {
  const PixResult r = pix_run(B200_PIX_SSE, 1, 8, width, height, a8, a_stride, &b8, 1, b_stride, 0, 0, nullptr);
  return (int64_t)(((uint64_t)r.v[0][1] << 32) | r.v[0][0]);
}

#define B200_DEF_BLOCK_FNS(W, H)                                                                   \
  unsigned int aom_sad##W##x##H##_b200(const uint8_t *s, int ss, const uint8_t *r, int rs) {        \
    return sad1(B200_PIX_SAD, 0, W, H, s, ss, r, rs);                                               \
  }                                                                                                 \
  unsigned int aom_sad_skip_##W##x##H##_b200(const uint8_t *s, int ss, const uint8_t *r, int rs) {  \
    return sad1(B200_PIX_SAD_SKIP, 0, W, H, s, ss, r, rs);                                          \
  }                                                                                                 \
  unsigned int aom_sad##W##x##H##_avg_b200(const uint8_t *s, int ss, const uint8_t *r, int rs,      \
                                           const uint8_t *second_pred) {                            \
    return sad1(B200_PIX_SAD_AVG, 0, W, H, s, ss, r, rs, second_pred);                              \
  }                                                                                                 \
  void aom_sad##W##x##H##x4d_b200(const uint8_t *s, int ss, const uint8_t *const r[4], int rs,      \
                                  uint32_t out[4]) {                                                \
    sad4(B200_PIX_SAD, 0, W, H, s, ss, r, rs, out);                                                 \
  }                                                                                                 \
  void aom_sad##W##x##H##x3d_b200(const uint8_t *s, int ss, const uint8_t *const r[4], int rs,      \
                                  uint32_t out[4]) {                                                \
    sad4(B200_PIX_SAD, 0, W, H, s, ss, r, rs, out); /* the C x3d fills all four (sad.c:124) */      \
  }                                                                                                 \
  void aom_sad_skip_##W##x##H##x4d_b200(const uint8_t *s, int ss, const uint8_t *const r[4],        \
                                        int rs, uint32_t out[4]) {                                  \
    sad4(B200_PIX_SAD_SKIP, 0, W, H, s, ss, r, rs, out);                                            \
  }                                                                                                 \
  unsigned int aom_highbd_sad##W##x##H##_b200(const uint8_t *s, int ss, const uint8_t *r, int rs) { \
    return sad1(B200_PIX_SAD, 1, W, H, s, ss, r, rs);                                               \
  }                                                                                                 \
  unsigned int aom_highbd_sad_skip_##W##x##H##_b200(const uint8_t *s, int ss, const uint8_t *r,     \
                                                    int rs) {                                       \
    return sad1(B200_PIX_SAD_SKIP, 1, W, H, s, ss, r, rs);                                          \
  }                                                                                                 \
  void aom_highbd_sad##W##x##H##x4d_b200(const uint8_t *s, int ss, const uint8_t *const r[4],       \
                                         int rs, uint32_t out[4]) {                                 \
    sad4(B200_PIX_SAD, 1, W, H, s, ss, r, rs, out);                                                 \
  }                                                                                                 \
  unsigned int aom_variance##W##x##H##_b200(const uint8_t *a, int as, const uint8_t *b, int bs,     \
                                            unsigned int *sse) {                                    \
    return var1(B200_PIX_VAR, 0, 8, W, H, a, as, 0, 0, b, bs, sse);                                 \
  }                                                                                                 \
  uint32_t aom_sub_pixel_variance##W##x##H##_b200(const uint8_t *a, int as, int xo, int yo,         \
                                                  const uint8_t *b, int bs, uint32_t *sse) {        \
    return var1(B200_PIX_SUBPEL_VAR, 0, 8, W, H, a, as, xo, yo, b, bs, sse);                        \
  }                                                                                                 \
  unsigned int aom_highbd_8_variance##W##x##H##_b200(const uint8_t *a, int as, const uint8_t *b,    \
                                                     int bs, unsigned int *sse) {                   \
    return var1(B200_PIX_VAR, 1, 8, W, H, a, as, 0, 0, b, bs, sse);                                 \
  }                                                                                                 \
  unsigned int aom_highbd_10_variance##W##x##H##_b200(const uint8_t *a, int as, const uint8_t *b,   \
                                                      int bs, unsigned int *sse) {                  \
    return var1(B200_PIX_VAR, 1, 10, W, H, a, as, 0, 0, b, bs, sse);                                \
  }                                                                                                 \
  unsigned int aom_highbd_12_variance##W##x##H##_b200(const uint8_t *a, int as, const uint8_t *b,   \
                                                      int bs, unsigned int *sse) {                  \
    return var1(B200_PIX_VAR, 1, 12, W, H, a, as, 0, 0, b, bs, sse);                                \
  }                                                                                                 \
  uint32_t aom_highbd_8_sub_pixel_variance##W##x##H##_b200(const uint8_t *a, int as, int xo,        \
                                                           int yo, const uint8_t *b, int bs,        \
                                                           uint32_t *sse) {                         \
    return var1(B200_PIX_SUBPEL_VAR, 1, 8, W, H, a, as, xo, yo, b, bs, sse);                        \
  }                                                                                                 \
  uint32_t aom_highbd_10_sub_pixel_variance##W##x##H##_b200(const uint8_t *a, int as, int xo,       \
                                                            int yo, const uint8_t *b, int bs,       \
                                                            uint32_t *sse) {                        \
    return var1(B200_PIX_SUBPEL_VAR, 1, 10, W, H, a, as, xo, yo, b, bs, sse);                       \
  }                                                                                                 \
  uint32_t aom_highbd_12_sub_pixel_variance##W##x##H##_b200(const uint8_t *a, int as, int xo,       \
                                                            int yo, const uint8_t *b, int bs,       \
                                                            uint32_t *sse) {                        \
    return var1(B200_PIX_SUBPEL_VAR, 1, 12, W, H, a, as, xo, yo, b, bs, sse);                       \
  }
B200_DEF_BLOCK_FNS(128, 128)
B200_DEF_BLOCK_FNS(128, 64)
B200_DEF_BLOCK_FNS(64, 128)
B200_DEF_BLOCK_FNS(64, 64)
B200_DEF_BLOCK_FNS(64, 32)
B200_DEF_BLOCK_FNS(32, 64)
B200_DEF_BLOCK_FNS(32, 32)
B200_DEF_BLOCK_FNS(32, 16)
B200_DEF_BLOCK_FNS(16, 32)
B200_DEF_BLOCK_FNS(16, 16)
B200_DEF_BLOCK_FNS(16, 8)
B200_DEF_BLOCK_FNS(8, 16)
B200_DEF_BLOCK_FNS(8, 8)
B200_DEF_BLOCK_FNS(8, 4)
B200_DEF_BLOCK_FNS(4, 8)
B200_DEF_BLOCK_FNS(4, 4)
B200_DEF_BLOCK_FNS(4, 16)
B200_DEF_BLOCK_FNS(16, 4)
B200_DEF_BLOCK_FNS(8, 32)
B200_DEF_BLOCK_FNS(32, 8)
B200_DEF_BLOCK_FNS(16, 64)
B200_DEF_BLOCK_FNS(64, 16)

#define B200_DEF_MSE(W, H)                                                                     \
  unsigned int aom_mse##W##x##H##_b200(const uint8_t *a, int as, const uint8_t *b, int bs,      \
                                       unsigned int *sse) {                                     \
    return var1(B200_PIX_MSE, 0, 8, W, H, a, as, 0, 0, b, bs, sse);                             \
  }
B200_DEF_MSE(16, 16)
B200_DEF_MSE(16, 8)
B200_DEF_MSE(8, 16)
B200_DEF_MSE(8, 8)

void aom_upsampled_pred_b200(void *xd, const void *cm, int mi_row, int mi_col, const void *mv,
                             uint8_t *comp_pred, int width, int height, int subpel_x_q3,
                             int subpel_y_q3, const uint8_t *ref, int ref_stride,
                             int subpel_search) {
  (void)xd, (void)cm, (void)mi_row, (void)mi_col, (void)mv;
  upsample_one(0, 8, comp_pred, width, height, subpel_x_q3, subpel_y_q3, ref, ref_stride,
               subpel_search);
}
void aom_highbd_upsampled_pred_b200(void *xd, const void *cm, int mi_row, int mi_col,
                                    const void *mv, uint8_t *comp_pred8, int width, int height,
                                    int subpel_x_q3, int subpel_y_q3, const uint8_t *ref8,
                                    int ref_stride, int bd, int subpel_search) {
  (void)xd, (void)cm, (void)mi_row, (void)mi_col, (void)mv;
  upsample_one(1, bd, comp_pred8, width, height, subpel_x_q3, subpel_y_q3, ref8, ref_stride,
               subpel_search);
}

}  // extern "C"


// ------------------------------------------- Hadamard / SATD / averages (avg.c) ----
namespace {

// one n x n Hadamard (or `count` side-by-side 8x8 LP blocks) of a host residual
void hadamard_one(int n, int kind, const int16_t *src_diff, ptrdiff_t stride, void *coeff, int count = 1) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t esz = kind == B200_HAD_LP ? 2 : 4, nn = (size_t)n * n;
  Arena a(c, nn * count * (2 + esz) + 4096);
  int16_t *d_src = a.take<int16_t>(nn * count);
  char *d_out = a.take<char>(nn * count * esz);
  B200HadBlock *d_blk = a.take<B200HadBlock>(count);
  B200HadBlock blk[2];
  cudaStream_t s = c->stream;
  const int w = n * count;  // the blocks sit side by side in the caller's buffer
  CKCUDA(cudaMemcpy2DAsync(d_src, (size_t)w * 2, src_diff, (size_t)stride * 2, (size_t)w * 2, n,
                           cudaMemcpyHostToDevice, s));
  for (int i = 0; i < count; ++i) {
    blk[i].src_offset = (uint32_t)(i * n);
    blk[i].out_offset = (uint32_t)(i * nn);
    blk[i].stride = w;
  }
  CKCUDA(cudaMemcpyAsync(d_blk, blk, sizeof(B200HadBlock) * count, cudaMemcpyHostToDevice, s));
  B200HadamardLaunch l;
  memset(&l, 0, sizeof(l));
  l.n = n;
  l.kind = kind;
  l.n_blocks = count;
  l.blocks = d_blk;
  l.src = d_src;
  l.coeff = d_out;
  CK(b200_hadamard_launch(c, &l, s));
  CKCUDA(cudaMemcpyAsync(coeff, d_out, nn * count * esz, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

int satd_one(const void *coeff, int lp, int length) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t esz = lp ? 2 : 4;
  Arena a(c, (size_t)length * esz + 1024);
  char *d_c = a.take<char>((size_t)length * esz);
  int *d_o = a.take<int>(1);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpyAsync(d_c, coeff, (size_t)length * esz, cudaMemcpyHostToDevice, s));
  CK(satd_launch(c, d_c, lp, length, d_o, s));
  int r = 0;
  CKCUDA(cudaMemcpyAsync(&r, d_o, sizeof(r), cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
  return r;
}

// `count` n x n blocks at (x, y) pixel offsets inside one w x h host window
void avg_run(int op, int highbd, const uint8_t *a8, int a_stride, const uint8_t *b8, int b_stride,
             int w, int h, int count, const int (*xy)[2], int (*out)[2]) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const size_t px = highbd ? 2 : 1;
  Arena ar(c, (size_t)w * h * px * 2 + 4096);
  char *d_a = ar.take<char>((size_t)w * h * px);
  char *d_b = b8 ? ar.take<char>((size_t)w * h * px) : nullptr;
  B200AvgBlock *d_blk = ar.take<B200AvgBlock>(count);
  int32_t *d_out = ar.take<int32_t>(2 * count);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpy2DAsync(d_a, w * px, unbyte(a8, highbd), a_stride * px, w * px, h,
                           cudaMemcpyHostToDevice, s));
  if (b8)
    CKCUDA(cudaMemcpy2DAsync(d_b, w * px, unbyte(b8, highbd), b_stride * px, w * px, h,
                             cudaMemcpyHostToDevice, s));
  B200AvgBlock blk[4];
  for (int i = 0; i < count; ++i) {
    blk[i].a_offset = blk[i].b_offset = (uint32_t)(xy[i][1] * w + xy[i][0]);
    blk[i].a_stride = blk[i].b_stride = w;
  }
  CKCUDA(cudaMemcpyAsync(d_blk, blk, sizeof(B200AvgBlock) * count, cudaMemcpyHostToDevice, s));
  B200AvgLaunch l;
  memset(&l, 0, sizeof(l));
  l.op = op;
  l.highbd = highbd;
  l.n_blocks = count;
  l.blocks = d_blk;
  l.a = d_a;
  l.b = d_b;
  l.out = d_out;
  CK(b200_avg_launch(c, &l, s));
  CKCUDA(cudaMemcpyAsync(out, d_out, sizeof(int32_t) * 2 * count, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

unsigned avg_single(int op, int highbd, const uint8_t *s8, int p) {
  const int n = op == B200_AVG_4X4 ? 4 : 8;
  const int xy[1][2] = { { 0, 0 } };
  int out[1][2];
  avg_run(op, highbd, s8, p, nullptr, 0, n, n, 1, xy, out);
  return (unsigned)out[0][0];
}

void int_pro_one(int16_t *buf, const uint8_t *ref, int ref_stride, int width, int height, int norm,
                 int col) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const int n = col ? height : width;
  Arena a(c, (size_t)width * height + (size_t)n * 2 + 1024);
  unsigned char *d_ref = a.take<unsigned char>((size_t)width * height);
  short *d_out = a.take<short>(n);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpy2DAsync(d_ref, width, ref, ref_stride, width, height, cudaMemcpyHostToDevice, s));
  CK(int_pro_launch(c, d_out, d_ref, width, width, height, norm, col, s));
  CKCUDA(cudaMemcpyAsync(buf, d_out, (size_t)n * 2, cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
}

}  // namespace

extern "C" {
void aom_hadamard_4x4_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(4, B200_HAD_LOWBD, s, st, c); }
void aom_hadamard_8x8_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(8, B200_HAD_LOWBD, s, st, c); }
void aom_hadamard_16x16_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(16, B200_HAD_LOWBD, s, st, c); }
void aom_hadamard_32x32_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(32, B200_HAD_LOWBD, s, st, c); }
void aom_hadamard_lp_8x8_b200(const int16_t *s, ptrdiff_t st, int16_t *c) { hadamard_one(8, B200_HAD_LP, s, st, c); }
void aom_hadamard_lp_16x16_b200(const int16_t *s, ptrdiff_t st, int16_t *c) { hadamard_one(16, B200_HAD_LP, s, st, c); }
void aom_hadamard_lp_8x8_dual_b200(const int16_t *s, ptrdiff_t st, int16_t *c) { hadamard_one(8, B200_HAD_LP, s, st, c, 2); }
void aom_highbd_hadamard_8x8_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(8, B200_HAD_HIGHBD, s, st, c); }
void aom_highbd_hadamard_16x16_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(16, B200_HAD_HIGHBD, s, st, c); }
void aom_highbd_hadamard_32x32_b200(const int16_t *s, ptrdiff_t st, b200_tran_low_t *c) { hadamard_one(32, B200_HAD_HIGHBD, s, st, c); }
int aom_satd_b200(const b200_tran_low_t *coeff, int length) { return satd_one(coeff, 0, length); }
int aom_satd_lp_b200(const int16_t *coeff, int length) { return satd_one(coeff, 1, length); }
void av1_fwht4x4_b200(const int16_t *input, b200_tran_low_t *output, int stride) {
  hadamard_one(4, B200_HAD_WHT, input, stride, output);
}
unsigned int aom_avg_4x4_b200(const uint8_t *s, int p) { return avg_single(B200_AVG_4X4, 0, s, p); }
unsigned int aom_avg_8x8_b200(const uint8_t *s, int p) { return avg_single(B200_AVG_8X8, 0, s, p); }
unsigned int aom_highbd_avg_4x4_b200(const uint8_t *s8, int p) { return avg_single(B200_AVG_4X4, 1, s8, p); }
unsigned int aom_highbd_avg_8x8_b200(const uint8_t *s8, int p) { return avg_single(B200_AVG_8X8, 1, s8, p); }
void aom_avg_8x8_quad_b200(const uint8_t *s, int p, int x16_idx, int y16_idx, int *avg) {
  const int xy[4][2] = { { 0, 0 }, { 8, 0 }, { 0, 8 }, { 8, 8 } };
  int out[4][2];
  avg_run(B200_AVG_8X8, 0, s + (size_t)y16_idx * p + x16_idx, p, nullptr, 0, 16, 16, 4, xy, out);
  for (int k = 0; k < 4; ++k) avg[k] = out[k][0];
}
void aom_minmax_8x8_b200(const uint8_t *s, int p, const uint8_t *d, int dp, int *min, int *max) {
  const int xy[1][2] = { { 0, 0 } };
  int out[1][2];
  avg_run(B200_MINMAX_8X8, 0, s, p, d, dp, 8, 8, 1, xy, out);
  *min = out[0][0];
  *max = out[0][1];
}
void aom_highbd_minmax_8x8_b200(const uint8_t *s8, int p, const uint8_t *d8, int dp, int *min, int *max) {
  const int xy[1][2] = { { 0, 0 } };
  int out[1][2];
  avg_run(B200_MINMAX_8X8, 1, s8, p, d8, dp, 8, 8, 1, xy, out);
  *min = out[0][0];
  *max = out[0][1];
}
void aom_int_pro_row_b200(int16_t *hbuf, const uint8_t *ref, const int ref_stride, const int width,
                          const int height, int norm_factor) {
  int_pro_one(hbuf, ref, ref_stride, width, height, norm_factor, 0);
}
void aom_int_pro_col_b200(int16_t *vbuf, const uint8_t *ref, const int ref_stride, const int width,
                          const int height, int norm_factor) {
  int_pro_one(vbuf, ref, ref_stride, width, height, norm_factor, 1);
}
int aom_vector_var_b200(const int16_t *ref, const int16_t *src, int bwl) {
  std::lock_guard<std::mutex> lk(g_mu);
  b200_ctx *c = ctx0();
  const int width = 4 << bwl;
  Arena a(c, (size_t)width * 4 + 1024);
  short *d_r = a.take<short>(width), *d_s = a.take<short>(width);
  int *d_o = a.take<int>(1);
  cudaStream_t s = c->stream;
  CKCUDA(cudaMemcpyAsync(d_r, ref, (size_t)width * 2, cudaMemcpyHostToDevice, s));
  CKCUDA(cudaMemcpyAsync(d_s, src, (size_t)width * 2, cudaMemcpyHostToDevice, s));
  CK(vector_var_launch(c, d_r, d_s, bwl, d_o, s));
  int r = 0;
  CKCUDA(cudaMemcpyAsync(&r, d_o, sizeof(r), cudaMemcpyDeviceToHost, s));
  CKCUDA(cudaStreamSynchronize(s));
  return r;
}
}  // extern "C"

#include "rtcd_compound.inc"
