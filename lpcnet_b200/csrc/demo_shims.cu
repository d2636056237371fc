// What the reference's own command-line program needs from the library beyond include/lpcnet.h, so that the UNMODIFIED
// src/lpcnet_demo.c links against liblpcnet_b200.so (oracle/Makefile target lpcnet_demo_b200, tests/test_gpu_demo.py):
//   float lpc_from_cepstrum(float *lpc, const float *cepstrum)     src/freq.h:57, src/freq.c:310-320 (its -addlpc mode)
// One row through the frame network's cepstrum -> LPC code on the GPU (no CPU fallback here either).
#include <mutex>
#include <string>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "dsp.cuh"
#include "engine.h"

namespace lpcb {
namespace {

__global__ void lpc_single_kernel(DspTables T, const float *cepstrum, float *out17, float *scratch) {
  if (threadIdx.x != 0) return;
  Scratch S{scratch, 1};
  lpc_from_cepstrum_dev(T, cepstrum, out17, S, out17 + 16);
}

struct Shim {
  float *d_in = nullptr, *d_out = nullptr, *d_scratch = nullptr;
  void *d_tab = nullptr;
  DspTables T{};
  bool ok = false;
};
std::mutex g_mu;
Shim g_shim[64];

bool set_up(Shim &s) {
  std::vector<float> tw, dct;
  std::vector<int16_t> br;
  host_dsp_tables(tw, br, dct);
  const size_t o_br = tw.size() * 4, o_dct = (o_br + br.size() * 2 + 15) & ~size_t(15), total = o_dct + dct.size() * 4;
  if (cudaMalloc(&s.d_tab, total) != cudaSuccess || cudaMalloc(&s.d_in, 18 * 4) != cudaSuccess ||
      cudaMalloc(&s.d_out, 17 * 4) != cudaSuccess || cudaMalloc(&s.d_scratch, 640 * 4) != cudaSuccess)
    return false;
  char *p = static_cast<char *>(s.d_tab);
  if (cudaMemcpy(p, tw.data(), tw.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(p + o_br, br.data(), br.size() * 2, cudaMemcpyHostToDevice) != cudaSuccess ||
      cudaMemcpy(p + o_dct, dct.data(), dct.size() * 4, cudaMemcpyHostToDevice) != cudaSuccess)
    return false;
  s.T.fft_tw = reinterpret_cast<const float *>(p);
  s.T.fft_bitrev = reinterpret_cast<const int16_t *>(p + o_br);
  s.T.dct = reinterpret_cast<const float *>(p + o_dct);
  return s.ok = true;
}

}  // namespace
}  // namespace lpcb

using namespace lpcb;

extern "C" __attribute__((visibility("default"))) float lpc_from_cepstrum(float *lpc, const float *cepstrum) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev >= 64) { set_last_error("lpc_from_cepstrum: no CUDA device"); return 0.f; }
  std::lock_guard<std::mutex> lk(g_mu);
  Shim &s = g_shim[dev];
  if (!s.ok && !set_up(s)) { set_last_error(std::string("lpc_from_cepstrum: ") + cudaGetErrorString(cudaGetLastError())); return 0.f; }
  float out[17] = {0.f};
  if (cudaMemcpy(s.d_in, cepstrum, 18 * 4, cudaMemcpyHostToDevice) != cudaSuccess) { set_last_error("lpc_from_cepstrum: H2D failed"); return 0.f; }
  lpc_single_kernel<<<1, 32>>>(s.T, s.d_in, s.d_out, s.d_scratch);
  if (cudaMemcpy(out, s.d_out, 17 * 4, cudaMemcpyDeviceToHost) != cudaSuccess) { set_last_error("lpc_from_cepstrum: kernel or D2H failed"); return 0.f; }
  for (int i = 0; i < 16; i++) lpc[i] = out[i];
  return out[16];
}
