// Latency / throughput of mma.sync m16n8k32 u8*s8 and of dependent LDS on this GPU.
#include <cstdio>
#include <cstdint>
__device__ __forceinline__ void mma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <int ILP>
__global__ void k(long long *out, int iters, uint32_t seed) {
  int c[ILP][4];
  for (int i = 0; i < ILP; i++) c[i][0] = c[i][1] = c[i][2] = c[i][3] = 0;
  uint32_t a = seed + threadIdx.x, b = seed * 3 + threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < ILP; i++) mma(c[i], a, a + 1, a + 2, a + 3, b, b + 1);
  }
  long long t1 = clock64();
  int s = 0;
  for (int i = 0; i < ILP; i++) s += c[i][0] + c[i][1] + c[i][2] + c[i][3];
  if (threadIdx.x == 0 && blockIdx.x == 0) { out[0] = t1 - t0; out[1] = s; }
}
__global__ void lds_chain(long long *out, int iters) {
  __shared__ uint32_t tab[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) tab[i] = (i * 37 + 11) & 1023;
  __syncthreads();
  uint32_t p = threadIdx.x;
  long long t0 = clock64();
  for (int it = 0; it < iters; it++) p = tab[p];
  long long t1 = clock64();
  if (threadIdx.x == 0) { out[0] = t1 - t0; out[1] = p; }
}
int main() {
  long long *out; cudaMallocManaged(&out, 64);
  const int iters = 4096;
#define RUN(ILP, warps) do { k<ILP><<<1, 32 * warps>>>(out, iters, 5); cudaDeviceSynchronize(); k<ILP><<<1, 32 * warps>>>(out, iters, 5); cudaDeviceSynchronize(); \
    printf("imma ILP %d warps %2d: %.1f cycles per mma per warp, %.2f mma/cycle/SM\n", ILP, warps, (double)out[0] / iters / ILP, (double)iters * ILP * warps / out[0]); } while (0)
  RUN(1, 1); RUN(2, 1); RUN(4, 1); RUN(1, 4); RUN(4, 4); RUN(1, 16); RUN(4, 16);
  lds_chain<<<1, 32>>>(out, iters); cudaDeviceSynchronize();
  printf("dependent LDS: %.1f cycles\n", (double)out[0] / iters);
  return 0;
}
