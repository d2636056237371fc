// Instantiation of the sample kernel for 32 streams per CTA (own translation unit so the variants compile in parallel).
#include "sample_kernel.cuh"
namespace lpcb {
cudaError_t launch_sample_t32(const SampleArgs &a, int n_sms, cudaStream_t s) { return launch_tile<2, 16>(a, n_sms, s); }
}  // namespace lpcb
