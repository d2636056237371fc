// Host-side dispatch of the persistent sample kernel (sample_kernel.cuh) by streams per CTA.
#include "sample_kernel.cuh"

namespace lpcb {

cudaError_t launch_sample_t8(const SampleArgs &a, int n_sms, cudaStream_t s);
cudaError_t launch_sample_t16(const SampleArgs &a, int n_sms, cudaStream_t s);
cudaError_t launch_sample_t32(const SampleArgs &a, int n_sms, cudaStream_t s);
cudaError_t launch_sample_pp(const SampleArgs &a, int n_sms, cudaStream_t s);

// tile code 28: two 8-stream tiles per CTA (sample_kernel G = 2)
constexpr int kTilePairs = 28;

size_t sample_kernel_smem(int tile_streams, const DevModel &m) {
  if (tile_streams == kTilePairs) return make_layout(8, 1, m.a_rec.n_chunks, m.a_rec.slots[2] != 0, m.b_rec.n_chunks, 2).total;
  const int VR = tile_streams == 8 ? 8 : 16;
  const int MT = tile_streams / VR;
  return make_layout(tile_streams, MT, m.a_rec.n_chunks, m.a_rec.slots[2] != 0 && !m.has_sat, m.b_rec.n_chunks).total;
}

cudaError_t launch_sample_kernel(const SampleArgs &a, int tile_streams, int n_sms, cudaStream_t s, int *n_launches) {
  cudaError_t e;
  if (tile_streams == 0) {
    const int B = a.n_streams, full = B / (16 * n_sms) * (16 * n_sms), rest = B - full;
    if (full == 0) return launch_sample_kernel(a, rest <= 8 * n_sms ? 8 : 16, n_sms, s, n_launches);
    if (rest == 0 || rest > 8 * n_sms || a.trace || a.logits) return launch_sample_kernel(a, 16, n_sms, s, n_launches);
    SampleArgs head = a, tail = a;
    head.stream_end = full;
    tail.stream_begin = full;
    if ((e = launch_sample_kernel(head, 16, n_sms, s, n_launches)) != cudaSuccess) return e;
    return launch_sample_kernel(tail, 8, n_sms, s, n_launches);
  }
  switch (tile_streams) {
    case 8: e = launch_sample_t8(a, n_sms, s); break;
    case 16: e = launch_sample_t16(a, n_sms, s); break;
    case 32: e = launch_sample_t32(a, n_sms, s); break;
    case kTilePairs:
      // the trace / logit kernels and models on the exact s16 pass have no two-group variant: plain 16-stream tiles
      e = (a.trace || a.logits || a.m.has_sat) ? launch_sample_t16(a, n_sms, s) : launch_sample_pp(a, n_sms, s);
      break;
    default: return cudaErrorInvalidValue;
  }
  if (e == cudaSuccess && n_launches) ++*n_launches;
  return e;
}

}  // namespace lpcb
