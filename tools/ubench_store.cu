// Store-pattern microbenchmark: how fast can a B200 SM array write when every lane owns a
// private contiguous run (lane-owns-block layout) versus warp-contiguous stores?
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ void stg256(int *p, int v) {
  asm volatile("st.global.v8.b32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(p), "r"(v) : "memory");
}
// MODE 0: lane owns RUN ints, writes them with 256-bit stores (lane stride = RUN*4 bytes)
// MODE 1: same ownership, 128-bit stores; MODE 2: warp-contiguous 128-bit; MODE 3: warp-contiguous 32-bit
template <int MODE, int RUN>
__global__ void k(int *out, size_t n_ints) {
  const size_t warps = (size_t)gridDim.x * blockDim.x / 32;
  const size_t w = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) / 32;
  const int lane = threadIdx.x & 31;
  const size_t per_iter = 32 * RUN;
  for (size_t base = w * per_iter; base + per_iter <= n_ints; base += warps * per_iter) {
    if (MODE == 0) {
#pragma unroll
      for (int g = 0; g < RUN / 8; ++g) stg256(out + base + lane * RUN + 8 * g, lane + g);
    } else if (MODE == 1) {
#pragma unroll
      for (int g = 0; g < RUN / 4; ++g)
        *reinterpret_cast<int4 *>(out + base + lane * RUN + 4 * g) = make_int4(lane, g, 0, 1);
    } else if (MODE == 2) {
#pragma unroll
      for (int g = 0; g < RUN / 4; ++g)
        *reinterpret_cast<int4 *>(out + base + g * 128 + lane * 4) = make_int4(lane, g, 0, 1);
    } else {
#pragma unroll
      for (int g = 0; g < RUN; ++g) out[base + g * 32 + lane] = lane + g;
    }
  }
}
template <int MODE, int RUN>
void run(int *d, size_t n, int ctas) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  for (int i = 0; i < 2; ++i) k<MODE, RUN><<<ctas, 128>>>(d, n);
  cudaEventRecord(a);
  for (int i = 0; i < 5; ++i) k<MODE, RUN><<<ctas, 128>>>(d, n);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  printf("mode %d run %3d ctas %5d: %.1f GB/s\n", MODE, RUN, ctas, 5.0 * n * 4 / ms / 1e6);
}
int main() {
  size_t n = (size_t)256 << 20;  // 1 GiB of ints
  int *d; cudaMalloc(&d, n * 4);
  for (int ctas : {148 * 2, 148 * 4, 148 * 16}) {
    run<0, 16>(d, n, ctas); run<0, 32>(d, n, ctas); run<0, 64>(d, n, ctas);
    run<1, 64>(d, n, ctas); run<2, 64>(d, n, ctas); run<3, 64>(d, n, ctas);
  }
  return 0;
}
