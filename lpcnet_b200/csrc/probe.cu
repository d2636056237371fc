// On-chip roofline probe: how fast can every SM gather embedding-row-sized (4608 B) rows from an
// L2-resident table?  This is the dominant data movement of the sample kernel (three such rows per
// stream per sample, src/nnet.c:484-491), so its ceiling is the kernel's bandwidth roofline.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/lpcnet_b200.h"

namespace {

constexpr int kRowFloats = 1152;
constexpr int kRows = 768;  // 3 tables x 256 rows = 3.5 MB

__global__ void __launch_bounds__(512) gather_probe(const float4 *__restrict__ table, float *sink, int iters, uint32_t seed) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  uint32_t x = seed ^ (warp * 2654435761u);
  float4 acc = make_float4(0, 0, 0, 0);
  for (int it = 0; it < iters; it++) {
    x = x * 1664525u + 1013904223u;
    const float4 *row = table + (size_t)((x >> 10) % kRows) * (kRowFloats / 4);
#pragma unroll
    for (int k = 0; k < kRowFloats / 4 / 32; k++) {   // 9 x LDG.128 per lane
      const float4 v = __ldg(row + k * 32 + lane);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  if (acc.x + acc.y + acc.z + acc.w == 12345.678f) sink[0] = acc.x;
}

}  // namespace

extern "C" int lpcnet_b200_measure_l2_gather(int device, double *gbs) {
  if (cudaSetDevice(device) != cudaSuccess) return -1;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1;
  float4 *table = nullptr;
  float *sink = nullptr;
  const size_t bytes = (size_t)kRows * kRowFloats * 4;
  if (cudaMalloc(&table, bytes) != cudaSuccess || cudaMalloc(&sink, 4) != cudaSuccess) return -1;
  cudaMemset(table, 0, bytes);
  const int grid = prop.multiProcessorCount * 2, block = 512, iters = 400;
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  gather_probe<<<grid, block>>>(table, sink, 50, 1u);
  double best = 0;
  for (int rep = 0; rep < 5; rep++) {
    cudaEventRecord(e0);
    gather_probe<<<grid, block>>>(table, sink, iters, 7u + rep);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(table); cudaFree(sink); return -1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    const double moved = (double)grid * (block / 32) * iters * kRowFloats * 4;
    const double v = moved / (ms * 1e-3) / 1e9;
    if (v > best) best = v;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(table); cudaFree(sink);
  if (gbs) *gbs = best;
  return 0;
}
