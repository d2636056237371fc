// Instantiation of the sample kernel for two 8-stream tiles per CTA (two groups of 8 warps; own translation unit so the
// variants compile in parallel).
#include "sample_kernel.cuh"
namespace lpcb {
cudaError_t launch_sample_pp(const SampleArgs &a, int n_sms, cudaStream_t s) { return launch_tile_pairs(a, n_sms, s); }
}  // namespace lpcb
