// Host-side dispatch of the persistent sample kernel (sample_kernel.cuh) by streams per CTA.
#include "sample_kernel.cuh"

namespace lpcb {

cudaError_t launch_sample_t8(const SampleArgs &a, int n_sms, cudaStream_t s);
cudaError_t launch_sample_t16(const SampleArgs &a, int n_sms, cudaStream_t s);
cudaError_t launch_sample_t32(const SampleArgs &a, int n_sms, cudaStream_t s);

size_t sample_kernel_smem(int tile_streams, const DevModel &m) {
  const int VR = tile_streams == 8 ? 8 : 16;
  const int MT = tile_streams / VR;
  return make_layout(tile_streams, MT, m.a_rec.n_chunks, m.b_in.n_chunks, m.b_rec.n_chunks).total;
}

cudaError_t launch_sample_kernel(const SampleArgs &a, int tile_streams, int n_sms, cudaStream_t s, int *n_launches) {
  cudaError_t e;
  switch (tile_streams) {
    case 8: e = launch_sample_t8(a, n_sms, s); break;
    case 16: e = launch_sample_t16(a, n_sms, s); break;
    case 32: e = launch_sample_t32(a, n_sms, s); break;
    default: return cudaErrorInvalidValue;
  }
  if (e == cudaSuccess && n_launches) ++*n_launches;
  return e;
}

}  // namespace lpcb
