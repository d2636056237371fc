// An entry point that allocates nothing: cepstrum -> LPC (the work of lpc_kernel in frame.cu: lpc_from_cepstrum,
// src/freq.c:310-320 and everything below it) for callers that cannot call cudaMalloc -- code under CUDA graph
// capture, tight serving loops, harnesses that forbid it.  No counterpart in the reference.
// Nothing else in the library uses this file, and this file changes no other code path.
//
// lpc_kernel's 320-point transform runs on global scratch (one column of 640 floats per row, interleaved over the
// threads).  Here that scratch is a file-scope device array of a fixed maximum size, and so are the DSP tables:
//   g_scratch  kNoAllocMaxRows x 640 floats = 10.0 MiB, held by every GPU that loads the library until exit
//   g_tw / g_bitrev / g_dct   4.4 KB of tables, uploaded once per device with cudaMemcpyToSymbol
// In host code `g_scratch` is only a stand-in for the device array, not a device address: the address a kernel can
// use comes from cudaGetSymbolAddress, resolved once per device with that device current (set_up below).
// All calls share the array.  A row's kernel thread zero-fills its own 640-float column before it uses it
// (lpc_from_cepstrum_dev clears all 320 bins), so what an earlier call left behind is never read.  Two calls whose
// kernels overlap on the GPU (different streams, no ordering between them) RACE on the array: keep calls on one
// stream, or order them with events.
#include <mutex>
#include <string>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "dsp.cuh"
#include "engine.h"

namespace lpcb {
namespace {      // (inside lpcb like dsp.cuh's: one unnamed namespace per translation unit)

constexpr int kNoAllocMaxRows = LPCNET_B200_NOALLOC_MAX_ROWS;
static_assert(kNoAllocMaxRows % 64 == 0, "whole blocks of 64 rows");

static __device__ __align__(256) float g_scratch[(size_t)kNoAllocMaxRows * 640];
static __device__ __align__(256) float g_tw[640];
static __device__ __align__(256) int16_t g_bitrev[320];
static __device__ __align__(256) float g_dct[18 * 18];

// one thread per row; scratch: the device address of g_scratch, [320][2][stride] with stride = threads of the grid
__global__ void lpc_noalloc_kernel(const float *__restrict__ cepstrum, int stride_in, float *__restrict__ lpc_out,
                                   int n_rows, float *scratch) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  const size_t nthreads = (size_t)gridDim.x * blockDim.x;
  if (idx >= n_rows) return;
  Scratch S{scratch + idx, nthreads};
  float lpc[16];
  // device code may name the table arrays directly
  lpc_from_cepstrum_dev(DspTables{g_tw, g_bitrev, g_dct}, cepstrum + (size_t)idx * stride_in, lpc, S);
  for (int i = 0; i < 16; i++) lpc_out[(size_t)idx * 16 + i] = lpc[i];
}

constexpr int kMaxDevices = 64;
std::mutex g_mu;
float *g_scratch_addr[kMaxDevices] = {nullptr};   // device address of g_scratch on every device that was set up

// once per device, with that device current: the scratch's device address and the tables.  No allocation.
int set_up(int device, float **addr) {
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_scratch_addr[device]) { *addr = g_scratch_addr[device]; return 0; }
  float *p = nullptr;
  cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_scratch);
  if (e != cudaSuccess || !p)
    return set_last_error(std::string("lpcnet_b200_lpc_from_cepstrum_noalloc: cudaGetSymbolAddress failed: ") + cudaGetErrorString(e));
  std::vector<float> tw, dct;
  std::vector<int16_t> br;
  host_dsp_tables(tw, br, dct);     // host vectors only
  if ((e = cudaMemcpyToSymbol(g_tw, tw.data(), sizeof(float) * 640)) != cudaSuccess ||
      (e = cudaMemcpyToSymbol(g_bitrev, br.data(), sizeof(int16_t) * 320)) != cudaSuccess ||
      (e = cudaMemcpyToSymbol(g_dct, dct.data(), sizeof(float) * 18 * 18)) != cudaSuccess)
    return set_last_error(std::string("lpcnet_b200_lpc_from_cepstrum_noalloc: table upload failed: ") + cudaGetErrorString(e));
  g_scratch_addr[device] = p;
  *addr = p;
  return 0;
}

}  // namespace
}  // namespace lpcb

using namespace lpcb;

extern "C" int lpcnet_b200_lpc_from_cepstrum_noalloc(int device, const float *d_cepstrum, int stride, float *d_lpc,
                                                     int n_rows, void *cuda_stream) {
  if (n_rows < 0) return set_last_error("lpcnet_b200_lpc_from_cepstrum_noalloc: n_rows is negative");
  if (n_rows > kNoAllocMaxRows)
    return set_last_error("lpcnet_b200_lpc_from_cepstrum_noalloc: n_rows exceeds LPCNET_B200_NOALLOC_MAX_ROWS (" +
                          std::to_string(kNoAllocMaxRows) + " rows: the static scratch array holds no more); split the call");
  if (n_rows > 0 && (!d_cepstrum || !d_lpc)) return set_last_error("lpcnet_b200_lpc_from_cepstrum_noalloc: null buffer");
  if (n_rows > 0 && stride < 18) return set_last_error("lpcnet_b200_lpc_from_cepstrum_noalloc: stride must be >= 18");
  cudaError_t e;
  if (device < 0) {
    if ((e = cudaGetDevice(&device)) != cudaSuccess)
      return set_last_error(std::string("lpcnet_b200_lpc_from_cepstrum_noalloc: no CUDA device: ") + cudaGetErrorString(e));
  } else if ((e = cudaSetDevice(device)) != cudaSuccess) {
    return set_last_error(std::string("lpcnet_b200_lpc_from_cepstrum_noalloc: cudaSetDevice: ") + cudaGetErrorString(e));
  }
  if (device >= kMaxDevices) return set_last_error("lpcnet_b200_lpc_from_cepstrum_noalloc: device index out of range");
  float *scratch = nullptr;
  if (set_up(device, &scratch)) return -1;
  if (n_rows == 0) return 0;        // a set-up call (e.g. before a graph capture starts)
  const int blocks = (n_rows + 63) / 64;
  lpc_noalloc_kernel<<<blocks, 64, 0, static_cast<cudaStream_t>(cuda_stream)>>>(d_cepstrum, stride, d_lpc, n_rows, scratch);
  if ((e = cudaGetLastError()) != cudaSuccess)
    return set_last_error(std::string("lpcnet_b200_lpc_from_cepstrum_noalloc: launch failed: ") + cudaGetErrorString(e));
  return 0;
}
