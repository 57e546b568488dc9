// txq.cu -- host side of the fused transform+quantise path: static configuration
// tables, scan orders, kernel registry and the launcher.  The kernel itself lives in
// txq_kernel.cuh and is instantiated per transform size in txq_inst_*.cu.
#include <math.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "txq_kernel.cuh"

namespace b200 {

static const TxInfo kTxInfo[B200_TX_SIZES_ALL] = {
#define TXI(W, H)                                                                          \
  { W, H, TxCfg<W, H>::Wk, TxCfg<W, H>::Hk, { TxCfg<W, H>::kUp, -TxCfg<W, H>::kDown1,      \
    -TxCfg<W, H>::kDown2 }, TxCfg<W, H>::kCosCol, TxCfg<W, H>::kCosRow, TxCfg<W, H>::kRect, \
    TxCfg<W, H>::kLogScale }
  TXI(4, 4),   TXI(8, 8),   TXI(16, 16), TXI(32, 32), TXI(64, 64), TXI(4, 8),  TXI(8, 4),
  TXI(8, 16),  TXI(16, 8),  TXI(16, 32), TXI(32, 16), TXI(32, 64), TXI(64, 32), TXI(4, 16),
  TXI(16, 4),  TXI(8, 32),  TXI(32, 8),  TXI(16, 64), TXI(64, 16),
#undef TXI
};
const TxInfo &tx_info(int tx_size) { return kTxInfo[tx_size]; }

// kernel registry, filled by the instance translation units
int txq_group_a(TxqKernelInfo (*table)[3], bool upload);
int txq_group_b(TxqKernelInfo (*table)[3], bool upload);
int txq_group_c(TxqKernelInfo (*table)[3], bool upload);
int txq_group_d(TxqKernelInfo (*table)[3], bool upload);
int txq_group_e(TxqKernelInfo (*table)[3], bool upload);
int txq_group_f(TxqKernelInfo (*table)[3], bool upload);
int txq_group_g(TxqKernelInfo (*table)[3], bool upload);
int txq_group_h(TxqKernelInfo (*table)[3], bool upload);
int txq_group_i(TxqKernelInfo (*table)[3], bool upload);
typedef int (*GroupFn)(TxqKernelInfo (*)[3], bool);
static const GroupFn kGroups[] = { txq_group_a, txq_group_b, txq_group_c,
                                   txq_group_d, txq_group_e, txq_group_f,
                                   txq_group_g, txq_group_h, txq_group_i };

static TxqKernelInfo g_kernels[B200_TX_SIZES_ALL][3];
static int g_ctas_per_sm[B200_TX_SIZES_ALL][3];
static std::once_flag g_entries_once;
static void init_entries() {
  for (GroupFn g : kGroups) g(g_kernels, false);
}

// Scan orders.  The codec stores coefficients column-major (position c*Hk + r).  Three
// classes exist (av1_scan_orders, scan.c:1666): "default" walks the anti-diagonals
// (tall blocks with r increasing, wide blocks with c increasing, square blocks
// alternating, odd diagonals r-increasing), "mrow" (V_* types) walks row by row,
// "mcol" (H_* types) is storage order.  64-point sizes use their kept 32-point region.
static std::vector<int16_t> g_scan[B200_TX_SIZES_ALL][3], g_iscan[B200_TX_SIZES_ALL][3];
static size_t g_iscan_dev_offset[B200_TX_SIZES_ALL];
static size_t g_iscan_total;
// prefix masks for txq_small.cuh: entry [cls][k] has bit p set for each storage position p
// among the first k scan entries (k = 0..N), sizes with N <= 64 only
static std::vector<unsigned long long> g_pmask;
static size_t g_pmask_offset[B200_TX_SIZES_ALL];
static std::once_flag g_scan_once;

static void build_scans() {
  size_t off = 0;
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts) {
    const int wk = kTxInfo[ts].wk, hk = kTxInfo[ts].hk, n = wk * hk;
    for (int cls = 0; cls < 3; ++cls) {
      std::vector<int16_t> s;
      s.reserve(n);
      if (cls == 2) {
        for (int i = 0; i < n; ++i) s.push_back((int16_t)i);
      } else if (cls == 1) {
        for (int r = 0; r < hk; ++r)
          for (int c = 0; c < wk; ++c) s.push_back((int16_t)(c * hk + r));
      } else {
        for (int dgn = 0; dgn < wk + hk - 1; ++dgn) {
          const bool r_up = hk > wk ? true : (wk > hk ? false : (dgn & 1) != 0);
          for (int k = 0; k <= dgn; ++k) {
            const int r = r_up ? k : dgn - k, c = dgn - r;
            if (r < hk && c < wk) s.push_back((int16_t)(c * hk + r));
          }
        }
      }
      std::vector<int16_t> inv(n);
      for (int i = 0; i < n; ++i) inv[s[i]] = (int16_t)i;
      g_scan[ts][cls] = s;
      g_iscan[ts][cls] = inv;
    }
    g_iscan_dev_offset[ts] = off;
    off += 3 * (size_t)n;
    g_pmask_offset[ts] = g_pmask.size();
    if (n <= 64) {
      for (int cls = 0; cls < 3; ++cls) {
        unsigned long long m = 0;
        g_pmask.push_back(0);
        for (int k = 0; k < n; ++k) {
          m |= 1ull << g_scan[ts][cls][k];
          g_pmask.push_back(m);
        }
      }
    }
  }
  g_iscan_total = off;
}

size_t scan_dev_offset(int tx_size) {
  std::call_once(g_scan_once, build_scans);
  return g_iscan_dev_offset[tx_size];
}

static int scan_class(int tx_type) { return tx_type < 10 ? 0 : ((tx_type & 1) ? 2 : 1); }

int txq_init_device(b200_ctx *ctx) {
  std::call_once(g_entries_once, init_entries);
  std::call_once(g_scan_once, build_scans);
  for (GroupFn g : kGroups) {
    const int rc = g(g_kernels, true);  // per-TU constant tables on the current device
    if (rc != B200_OK) return rc;
  }
  std::vector<int16_t> all(g_iscan_total);
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int cls = 0; cls < 3; ++cls)
      memcpy(all.data() + g_iscan_dev_offset[ts] + cls * g_iscan[ts][cls].size(),
             g_iscan[ts][cls].data(), g_iscan[ts][cls].size() * sizeof(int16_t));
  B200_CUDA_TRY(cudaMalloc(&ctx->d_iscan, all.size() * sizeof(int16_t)));
  B200_CUDA_TRY(cudaMemcpy(ctx->d_iscan, all.data(), all.size() * sizeof(int16_t),
                           cudaMemcpyHostToDevice));
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int cls = 0; cls < 3; ++cls)
      memcpy(all.data() + g_iscan_dev_offset[ts] + cls * g_scan[ts][cls].size(),
             g_scan[ts][cls].data(), g_scan[ts][cls].size() * sizeof(int16_t));
  B200_CUDA_TRY(cudaMalloc(&ctx->d_scan, all.size() * sizeof(int16_t)));
  B200_CUDA_TRY(cudaMemcpy(ctx->d_scan, all.data(), all.size() * sizeof(int16_t),
                           cudaMemcpyHostToDevice));
  B200_CUDA_TRY(cudaMalloc(&ctx->d_pmask, g_pmask.size() * sizeof(unsigned long long)));
  B200_CUDA_TRY(cudaMemcpy(ctx->d_pmask, g_pmask.data(), g_pmask.size() * sizeof(unsigned long long),
                           cudaMemcpyHostToDevice));
  for (int ts = 0; ts < B200_TX_SIZES_ALL; ++ts)
    for (int k = 0; k < 3; ++k) {
      const TxqKernelInfo &e = g_kernels[ts][k];
      B200_CUDA_TRY(cudaFuncSetAttribute(e.fn, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         e.smem_bytes));
      int nb = 0;
      B200_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, e.fn, kWarpsPerCta * 32,
                                                                  e.smem_bytes));
      g_ctas_per_sm[ts][k] = nb > 0 ? nb : 1;
    }
  return B200_OK;
}

int txq_launch(b200_ctx *ctx, const B200TxqLaunch &l, cudaStream_t s) {
  if (l.tx_size < 0 || l.tx_size >= B200_TX_SIZES_ALL || l.n_blocks < 0 || !l.blocks || !l.src) {
    set_error("b200_txq_launch: bad tx_size / n_blocks / null blocks or src");
    return B200_EINVAL;
  }
  if (l.in_kind != B200_IN_RESIDUAL && (!l.preds || !l.pred)) {
    set_error("b200_txq_launch: pixel input needs preds and pred");
    return B200_EINVAL;
  }
  if (l.q_kind != B200_Q_NONE && (!l.qrows || !l.qcoeff || !l.dqcoeff || !l.eob)) {
    set_error("b200_txq_launch: quantiser selected but qrows/qcoeff/dqcoeff/eob missing");
    return B200_EINVAL;
  }
  if ((l.flags & B200_TXQ_STORE_COEFF) && !l.coeff) {
    set_error("b200_txq_launch: STORE_COEFF without coeff buffer");
    return B200_EINVAL;
  }
  if (l.bd != 8 && l.bd != 10 && l.bd != 12) {
    set_error("b200_txq_launch: bd must be 8, 10 or 12 (got %d)", l.bd);
    return B200_EINVAL;
  }
  if (l.q_kind < B200_Q_NONE || l.q_kind > B200_Q_B_HBD) {
    set_error("b200_txq_launch: bad q_kind %d", l.q_kind);
    return B200_EINVAL;
  }
  if (l.n_blocks == 0) return B200_OK;
  if (l.in_kind < 0 || l.in_kind > 2) {
    set_error("b200_txq_launch: bad in_kind %d", l.in_kind);
    return B200_EINVAL;
  }
  const TxqKernelInfo &e = g_kernels[l.tx_size][l.in_kind];
  TxqArgs a;
  a.blocks = l.blocks;
  a.preds = l.preds;
  a.src = l.src;
  a.pred = l.pred;
  a.qrows = l.qrows;
  a.iscan3 = ctx->d_iscan + g_iscan_dev_offset[l.tx_size];
  a.pmask3 = ctx->d_pmask + g_pmask_offset[l.tx_size];
  a.coeff = l.coeff;
  a.qcoeff = l.qcoeff;
  a.dqcoeff = l.dqcoeff;
  a.eob = l.eob;
  a.dist = l.dist;
  a.n_blocks = l.n_blocks;
  a.q_kind = l.q_kind;
  a.bd = l.bd;
  a.flags = l.flags;
  const int n_tiles = (l.n_blocks + e.bpw - 1) / e.bpw;
  int grid = (n_tiles + kWarpsPerCta - 1) / kWarpsPerCta;
  const int resident = ctx->sm_count * g_ctas_per_sm[l.tx_size][l.in_kind];
  if (grid > resident) grid = resident;  // persistent: warps stride over the tiles
  e.fn<<<grid, kWarpsPerCta * 32, e.smem_bytes, s>>>(a);
  B200_CUDA_TRY(cudaGetLastError());
  return B200_OK;
}

}  // namespace b200

// host-visible table accessors (tests, host mirrors)
extern "C" int b200_tx_wide(int tx_size) { return b200::kTxInfo[tx_size].w; }
extern "C" int b200_tx_high(int tx_size) { return b200::kTxInfo[tx_size].h; }
extern "C" int b200_tx_max_eob(int tx_size) {
  return b200::kTxInfo[tx_size].wk * b200::kTxInfo[tx_size].hk;
}
extern "C" int b200_tx_valid(int tx_size, int tx_type) {
  if (tx_size < 0 || tx_size >= B200_TX_SIZES_ALL || tx_type < 0 || tx_type >= 16) return 0;
  const int dims[2] = { b200::kTxInfo[tx_size].h, b200::kTxInfo[tx_size].w };
  const int kinds[2] = { b200::vtx_of(tx_type), b200::htx_of(tx_type) };
  for (int i = 0; i < 2; ++i) {
    if (dims[i] == 64 && kinds[i] != 0) return 0;
    if (dims[i] == 32 && (kinds[i] == 1 || kinds[i] == 2)) return 0;
  }
  return 1;
}
extern "C" const int32_t *b200_cospi(int bit) {
  static int32_t t[4][64];
  static const bool ok = b200::fill_cospi_tables(t);
  return ok && bit >= 10 && bit <= 13 ? t[bit - 10] : nullptr;
}
extern "C" const int16_t *b200_scan(int tx_size, int tx_type) {
  std::call_once(b200::g_scan_once, b200::build_scans);
  return b200::g_scan[tx_size][b200::scan_class(tx_type)].data();
}
extern "C" const int16_t *b200_iscan(int tx_size, int tx_type) {
  std::call_once(b200::g_scan_once, b200::build_scans);
  return b200::g_iscan[tx_size][b200::scan_class(tx_type)].data();
}
