// ubench.cu -- integer-pipe throughput probe for sm_100a (development tool, not product).
// Prints ops / clk / SM for the instruction mixes the transform / SAD kernels rely on.
#include <cstdio>
#include <cuda_runtime.h>

#define ITERS 512
#define DEF_KERNEL(NAME, BODY)                                                   \
  __global__ void NAME(int *out, int seed) {                                     \
    int a0 = threadIdx.x + seed, a1 = a0 * 3 + 1, a2 = a0 ^ 0x55, a3 = a0 + 7;   \
    int a4 = a0 * 5, a5 = a0 - 9, a6 = a0 ^ 0x1234, a7 = a0 + 99;                \
    long long w0 = a0, w1 = a1, w2 = a2, w3 = a3, w4 = a4, w5 = a5, w6 = a6, w7 = a7; \
    int b = seed * 7 + 3, c = seed + 11;                                         \
    (void)w0; (void)w1; (void)w2; (void)w3; (void)w4; (void)w5; (void)w6; (void)w7; \
    for (int i = 0; i < ITERS; ++i) { BODY BODY }                                 \
    out[blockIdx.x * blockDim.x + threadIdx.x] =                                 \
        a0 + a1 + a2 + a3 + a4 + a5 + a6 + a7 + (int)(w0 + w1 + w2 + w3 + w4 + w5 + w6 + w7); \
  }

#define OP8(ASM, C)                                           \
  asm volatile(ASM : "+" C(a0) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a1) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a2) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a3) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a4) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a5) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a6) : "r"(b), "r"(c));             \
  asm volatile(ASM : "+" C(a7) : "r"(b), "r"(c));
#define R "r"
#define OPW8(ASM)                                             \
  asm volatile(ASM : "+l"(w0) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w1) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w2) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w3) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w4) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w5) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w6) : "r"(b), "r"(c));              \
  asm volatile(ASM : "+l"(w7) : "r"(b), "r"(c));

DEF_KERNEL(k_imad, OP8("mad.lo.s32 %0, %0, %1, %2;", R))
DEF_KERNEL(k_imadwide, OPW8("{ .reg .b32 lo, hi; mov.b64 {lo, hi}, %0; mad.wide.s32 %0, lo, %1, %0; }"))
DEF_KERNEL(k_iadd, OP8("add.s32 %0, %0, %1;", R))
DEF_KERNEL(k_shf, OP8("shf.r.clamp.b32 %0, %0, %1, %2;", R))
DEF_KERNEL(k_shr, OP8("shr.s32 %0, %0, %1;", R))
DEF_KERNEL(k_shl_imm, OP8("{ .reg .b32 t; shl.b32 t, %0, 3; add.s32 %0, t, %1; }", R))
DEF_KERNEL(k_lop3, OP8("lop3.b32 %0, %0, %1, %2, 0x96;", R))
DEF_KERNEL(k_vabsdiff4, OP8("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;", R))
DEF_KERNEL(k_vabsdiff2, OP8("vabsdiff2.u32.u32.u32.add %0, %1, %2, %0;", R))
DEF_KERNEL(k_dp4a, OP8("dp4a.u32.u32 %0, %1, %2, %0;", R))
DEF_KERNEL(k_dp2a, OP8("dp2a.lo.u32.u32 %0, %1, %2, %0;", R))
DEF_KERNEL(k_min, OP8("min.s32 %0, %0, %1;", R))
DEF_KERNEL(k_abs, OP8("{ .reg .b32 t; abs.s32 t, %0; sub.s32 %0, t, %1; }", R))
DEF_KERNEL(k_setp_sel, OP8("{ .reg .pred p; setp.gt.s32 p, %0, %1; selp.b32 %0, %2, %0, p; }", R))
DEF_KERNEL(k_iadd3, OP8("{ .reg .b32 t; add.s32 t, %0, %1; add.s32 %0, t, %2; }", R))
DEF_KERNEL(k_funnel64, OPW8("{ .reg .b32 lo, hi, r; mov.b64 {lo, hi}, %0; shf.r.clamp.b32 r, lo, hi, %1; mov.b64 %0, {r, lo}; }"))
DEF_KERNEL(k_prmt, OP8("prmt.b32 %0, %0, %1, %2;", R))
DEF_KERNEL(k_add64, OPW8("{ .reg .s64 t; cvt.s64.s32 t, %1; add.s64 %0, %0, t; }"))
DEF_KERNEL(k_sad, OP8("sad.u32 %0, %1, %2, %0;", R))
// mixed: IMAD + IADD interleaved (fma pipe + alu pipe dual issue?)
#define WRING(A, B) asm volatile("{ .reg .b32 lo, hi; mov.b64 {lo, hi}, %1; mad.wide.s32 %0, hi, %2, %0; }" : "+l"(A) : "l"(B), "r"(b));
DEF_KERNEL(k_imadwide_ring, WRING(w0, w7) WRING(w1, w0) WRING(w2, w1) WRING(w3, w2) WRING(w4, w3) WRING(w5, w4) WRING(w6, w5) WRING(w7, w6))
DEF_KERNEL(k_mix_imad_iadd,
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a0) : "r"(b), "r"(c));
           asm volatile("add.s32 %0, %0, %1;" : "+r"(a1) : "r"(b));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a2) : "r"(b), "r"(c));
           asm volatile("add.s32 %0, %0, %1;" : "+r"(a3) : "r"(b));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a4) : "r"(b), "r"(c));
           asm volatile("add.s32 %0, %0, %1;" : "+r"(a5) : "r"(b));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a6) : "r"(b), "r"(c));
           asm volatile("add.s32 %0, %0, %1;" : "+r"(a7) : "r"(b));)
DEF_KERNEL(k_mix_vabs_iadd,
           asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a0) : "r"(b), "r"(c));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a1) : "r"(b), "r"(c));
           asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a2) : "r"(b), "r"(c));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a3) : "r"(b), "r"(c));
           asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a4) : "r"(b), "r"(c));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a5) : "r"(b), "r"(c));
           asm volatile("vabsdiff4.u32.u32.u32.add %0, %1, %2, %0;" : "+r"(a6) : "r"(b), "r"(c));
           asm volatile("mad.lo.s32 %0, %0, %1, %2;" : "+r"(a7) : "r"(b), "r"(c));)

template <class K>
void run(const char *name, K kern, int ops_per_body, int *d_out, int sms, double clk_ghz) {
  const int threads = 1024, blocks = sms * 2;
  kern<<<blocks, threads>>>(d_out, 1);
  cudaDeviceSynchronize();
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  cudaEventRecord(e0);
  for (int i = 0; i < 5; ++i) kern<<<blocks, threads>>>(d_out, i);
  cudaEventRecord(e1);
  cudaEventSynchronize(e1);
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  double ops = 5.0 * blocks * threads * (double)ITERS * 2 * ops_per_body;
  double per_s = ops / (ms * 1e-3);
  printf("%-18s %8.3f ms  %7.2f Tops/s  %6.1f ops/clk/SM @%.3f GHz\n", name, ms, per_s / 1e12,
         per_s / sms / (clk_ghz * 1e9), clk_ghz);
}

int main() {
  cudaDeviceProp p;
  cudaGetDeviceProperties(&p, 0);
  int *d_out;
  cudaMalloc(&d_out, 1024 * p.multiProcessorCount * 2 * sizeof(int));
  int khz = 0;
  cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  double ghz = khz / 1e6;
  printf("%s, %d SMs, clock attr %.3f GHz (ops/clk uses this nominal clock)\n", p.name,
         p.multiProcessorCount, ghz);
#define RUN(K) run(#K, K, 8, d_out, p.multiProcessorCount, ghz)
  RUN(k_imad); RUN(k_imadwide); RUN(k_imadwide_ring); RUN(k_iadd); RUN(k_shf); RUN(k_shr); RUN(k_shl_imm); RUN(k_setp_sel); RUN(k_iadd3); RUN(k_funnel64); RUN(k_lop3);
  RUN(k_vabsdiff4); RUN(k_vabsdiff2); RUN(k_dp4a); RUN(k_dp2a); RUN(k_min); RUN(k_abs);
  RUN(k_prmt); RUN(k_add64); RUN(k_sad); RUN(k_mix_imad_iadd); RUN(k_mix_vabs_iadd);
  return 0;
}
