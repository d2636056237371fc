// mma.sync m8n8k16 u8*s8 latency / throughput, against m16n8k32 (tools/cuda/immabench.cu).
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ void mma8(int (&c)[2], uint32_t a, uint32_t b) {
  asm volatile("mma.sync.aligned.m8n8k16.row.col.s32.u8.s8.s32 {%0,%1}, {%2}, {%3}, {%0,%1};\n" : "+r"(c[0]), "+r"(c[1]) : "r"(a), "r"(b));
}
template <int ILP>
__global__ void k(long long *out, int iters, uint32_t seed) {
  int c[ILP][2];
  for (int i = 0; i < ILP; i++) c[i][0] = c[i][1] = 0;
  uint32_t a = seed + threadIdx.x, b = seed * 3 + threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) mma8(c[i], a + i, b);
  }
  long long t1 = clock64();
  int s = 0;
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1];
  if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = s; }
}
int main() {
  long long *out; cudaMallocManaged(&out, 64);
  const int iters = 4096;
#define RUN(ILP, warps) do { k<ILP><<<1, 32 * warps>>>(out, iters, 5); cudaDeviceSynchronize(); k<ILP><<<1, 32 * warps>>>(out, iters, 5); cudaDeviceSynchronize(); \
    printf("m8n8k16 ILP %d warps %2d: %.1f cycles per mma per warp, %.2f mma/cycle/SM\n", ILP, warps, (double)out[0] / iters / ILP, (double)iters * ILP * warps / out[0]); } while (0)
  RUN(1, 1); RUN(2, 1); RUN(4, 1); RUN(1, 4); RUN(4, 4); RUN(1, 16); RUN(4, 16);
  return 0;
}
