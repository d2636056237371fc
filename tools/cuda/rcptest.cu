// Which arithmetic forms reproduce the 2048-entry rcpps table on this GPU?  nvcc -arch=sm_100a -fmad=false
#include <cstdio>
#include <cstdint>
#include "../../lpcnet_b200/csrc/bitexact.cuh"
__device__ __forceinline__ float mufu_rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ uint32_t round12(float r) { return (__float_as_uint(r) + 0x400u) & 0xfffff800u; }
__global__ void k(int *cnt, int *ex) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 2048) return;
  const uint32_t kk = 4097u + 2u * i;
  const uint32_t q = (uint32_t)(((1ull << 26) / kk + 1) >> 1);
  const uint32_t tab = 0x3f000000u + ((q - 4096u) << 11);
  for (uint32_t e = 8; e <= 246; e++) {
    const float mid = __uint_as_float((e << 23) | (i << 12) | 0x800u);
    const uint32_t want = tab - ((e - 127u) << 23);
    const float r0 = mufu_rcp(mid);
    const float r1 = fmaf(r0, fmaf(-mid, r0, 1.0f), r0);
    const float r2 = __frcp_rn(mid);
    const uint32_t a = round12(r0), b = round12(r1), c = round12(r2);
    if (a != want) { int n = atomicAdd(&cnt[0], 1); if (n < 8) { ex[3 * n] = i; ex[3 * n + 1] = e; ex[3 * n + 2] = (int)(__float_as_uint(r0) - __float_as_uint(r2)); } }
    if (b != want) atomicAdd(&cnt[1], 1);
    if (c != want) atomicAdd(&cnt[2], 1);
    if (__float_as_uint(lpcb::rcp_estimate(mid, lpcb::RcpArith{})) != want) atomicAdd(&cnt[5], 1);
    if (((__float_as_uint(r0) + 0x3ffu) & 0xfffff800u) != want) { int n = atomicAdd(&cnt[6], 1); if (n < 4) { ex[24 + 2 * n] = i; ex[25 + 2 * n] = e; } }
    if (e == 127 && a != want) atomicAdd(&cnt[3], 1);
    if (e == 127 && b != want) atomicAdd(&cnt[4], 1);
  }
}
int main() {
  int *cnt, *ex; cudaMallocManaged(&cnt, 64); cudaMallocManaged(&ex, 4 * 40);
  for (int i = 0; i < 16; i++) cnt[i] = 0;
  k<<<8, 256>>>(cnt, ex); cudaDeviceSynchronize();
  printf("mismatches of 2048*239: mufu %d  newton %d  frcp_rn %d ; at e=127: mufu %d newton %d ; library form %d\n", cnt[0], cnt[1], cnt[2], cnt[3], cnt[4], cnt[5]);
  printf("bias 0x3ff without special case: %d mismatches", cnt[6]);
  for (int n = 0; n < 4 && n < cnt[6]; n++) printf("  (bucket %d e %d)", ex[24 + 2 * n], ex[25 + 2 * n]);
  printf("\n");
  for (int n = 0; n < 8 && n < cnt[0]; n++) printf("  bucket %d e %d  mufu-rn ulps %d\n", ex[3 * n], ex[3 * n + 1], ex[3 * n + 2]);
  return 0;
}
