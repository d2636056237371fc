// Instantiation of the sample kernel for 8 streams per CTA (own translation unit so the variants compile in parallel).
#include "sample_kernel.cuh"
namespace lpcb {
cudaError_t launch_sample_t8(const SampleArgs &a, int n_sms, cudaStream_t s) { return launch_tile<1, 8>(a, n_sms, s); }
}  // namespace lpcb
