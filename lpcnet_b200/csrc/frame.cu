// The 100 Hz frame network (run_frame_network, src/lpcnet.c:82-120) for a chunk of frames of
// every stream at once, plus the cepstrum -> LPC chain (src/freq.c:202-320, src/kiss_fft.c) and
// the 1.6 kb/s packet decoder (src/lpcnet_dec.c:59-155).
//
// The k=3 convolutions of the reference are dense layers over [two previous inputs | input]
// (src/nnet.c:452-470), so nothing is recurrent across frames except those two-row histories:
// every (stream, frame) row of the chunk is evaluated in parallel, with the histories carried
// in the stream state between chunks.  All matvecs accumulate with fmaf in ascending input
// order starting from the bias, which is exactly sgemv_accum16 (src/vec_avx.h:618-643).
#include "bitexact.cuh"
#include "engine.h"
#include "dsp.cuh"
#include "dsp_warp.cuh"

namespace lpcb {

namespace {

constexpr int kRowsPerBlock = 8;
constexpr int kFT = 128;  // threads per block of the dense kernels
constexpr int kPrefetch = 16;   // weight loads in flight per thread in the dense loops

__device__ __forceinline__ int frame_count_at(int fc0, int f) { return min(fc0 + f, 1000); }

// streams this launch covers: the whole batch, or the list FrameArgs::rows (k-th entry -> stream index)
__device__ __forceinline__ int n_active(const FrameArgs &A) { return A.rows ? A.n_rows : A.n_streams; }
__device__ __forceinline__ int stream_of(const FrameArgs &A, int k) { return A.rows ? __ldg(A.rows + k) : k; }

// K0: in_vec rows (history + features|pitch embedding) and the per-frame active flags.
__global__ void frame_inputs_kernel(FrameArgs A) {
  const int B = A.n_streams, F = A.n_frames;
  const int row = blockIdx.x;  // (stream, r) with r in [0, F+2)
  const int s = stream_of(A, row / (F + 2)), r = row % (F + 2);
  float *dst = A.fb.in_vec + ((size_t)s * (F + 2) + r) * kConvIn;
  const int tid = threadIdx.x;
  if (r < 2) {
    // feature_conv1_state = [older input | newer input] (src/nnet.c:460-469)
    if (tid < kConvIn) dst[tid] = A.st.conv1_state[(size_t)s * 2 * kConvIn + r * kConvIn + tid];
    // conv2 history rows of c1
    for (int i = tid; i < kC; i += blockDim.x)
      A.fb.c1[((size_t)s * (F + 2) + r) * kC + i] = A.st.conv2_state[(size_t)s * 2 * kC + r * kC + i];
    return;
  }
  const int f = r - 2;
  const float *feat = A.features + ((size_t)s * A.total_frames + A.frame0 + f) * A.feature_stride;
  // pitch = (int)floor(.1 + 50*features[18] + 100), clamped to [33,255] (src/lpcnet.c:93-94)
  // (the conversion of an out-of-range or NaN double is x86's cvttsd2si, INT_MIN, so that a huge or infinite
  // features[18] clamps to 33 as in the reference build -- CUDA's own conversion saturates to INT_MAX)
  const double pd = floor(.1 + (double)(50.f * feat[18]) + 100);
  int pitch = (pd != pd || pd >= 2147483648.0 || pd <= -2147483649.0) ? (int)0x80000000 : (int)pd;
  pitch = pitch > 255 ? 255 : (pitch < 33 ? 33 : pitch);
  if (tid < kFeat) dst[tid] = feat[tid];
  else if (tid < kConvIn) dst[tid] = A.m.embed_pitch[pitch * kPitchEmb + tid - kFeat];
  if (tid == 0) {
    const int fc0 = A.st.frame_count[s];
    // lpcnet_synthesize_tail_impl runs after frame_count++ (src/lpcnet.c:119, 239)
    const bool on = !A.stream_mask || A.stream_mask[s];
    A.fb.active[(size_t)f * B + s] = (on && frame_count_at(fc0, f + 1) > A.m.features_delay) ? 1 : 0;
  }
}

// out[r][i] = act(bias[i] + sum_j W[j*N+i] * x[r][j]), j ascending, fmaf: sgemv_accum16 order.
// The R input rows are interleaved in shared memory, xt[j][r], so one weight load meets the R activations
// of its input with two 16-byte broadcast loads (R = 8).  OUT_T: write the result interleaved too (it feeds
// the next layer), else as rows with stride out_stride.
template <int R, bool OUT_T>
__device__ __forceinline__ void dense_rows(const float *__restrict__ W, const float *__restrict__ bias, int K, int N,
                                           const float *xt /*smem [K][R]*/, float *out, int out_stride, bool do_tanh,
                                           const uint32_t *rcp, int tid, int nthreads, const bool *zero_row) {
  static_assert(R == 8, "two float4 per input");
  for (int i = tid; i < N; i += nthreads) {
    float2 acc2[R / 2];                       // rows (2k, 2k+1): packed fmaf, each half the same IEEE operation
    const float b = __ldg(bias + i);
#pragma unroll
    for (int r = 0; r < R / 2; r++) acc2[r] = make_float2(b, b);
    // The weights come from L2 and a block is alone on its rows, so the loop is bound by the load round trip:
    // kPrefetch weights are in flight per thread while the fmaf chain keeps the reference's order j = 0..K-1.
    auto step = [&](float w, int j) {
      const float4 x0 = *reinterpret_cast<const float4 *>(xt + j * R), x1 = *reinterpret_cast<const float4 *>(xt + j * R + 4);
      const float2 w2 = make_float2(w, w);
      acc2[0] = f2_fma(w2, make_float2(x0.x, x0.y), acc2[0]); acc2[1] = f2_fma(w2, make_float2(x0.z, x0.w), acc2[1]);
      acc2[2] = f2_fma(w2, make_float2(x1.x, x1.y), acc2[2]); acc2[3] = f2_fma(w2, make_float2(x1.z, x1.w), acc2[3]);
    };
    int j = 0;
    for (; j + kPrefetch <= K; j += kPrefetch) {
      float w[kPrefetch];
#pragma unroll
      for (int u = 0; u < kPrefetch; u++) w[u] = __ldg(W + (size_t)(j + u) * N + i);
#pragma unroll
      for (int u = 0; u < kPrefetch; u++) step(w[u], j + u);
    }
    for (; j < K; j++) step(__ldg(W + (size_t)j * N + i), j);
    const float acc[R] = {acc2[0].x, acc2[0].y, acc2[1].x, acc2[1].y, acc2[2].x, acc2[2].y, acc2[3].x, acc2[3].y};
#pragma unroll
    for (int r = 0; r < R; r++) {
      float v = do_tanh ? tanh_ref(acc[r], rcp) : acc[r];
      if (zero_row && zero_row[r]) v = 0.f;
      if (OUT_T) out[i * R + r] = v;
      else out[(size_t)r * out_stride + i] = v;
    }
  }
}

// K1: conv1 for R rows per block.

// true when every row of this block belongs to a stream the mask leaves out (nothing to compute or store)
__device__ __forceinline__ bool block_masked_out(const FrameArgs &A, int R) {
  if (!A.stream_mask) return false;
  const int F = A.n_frames, total = n_active(A) * F;
  for (int r = 0; r < R; r++) {
    const int row = blockIdx.x * R + r;
    if (row < total && A.stream_mask[stream_of(A, row / F)]) return false;
  }
  return true;
}

__global__ void __launch_bounds__(kFT) conv1_kernel(FrameArgs A) {
  constexpr int R = kRowsPerBlock;
  __shared__ __align__(16) float xt[3 * kConvIn][R];
  __shared__ float y[R][kC];
  __shared__ bool zero[R];
  __shared__ uint32_t rcp[2048];
  const int F = A.n_frames, total = n_active(A) * F;
  const int tid = threadIdx.x;
  if (block_masked_out(A, R)) return;
  for (int i = tid; i < 2048; i += kFT) rcp[i] = A.m.rcp_tab[i];
  for (int r = 0; r < R; r++) {
    const int row = min(blockIdx.x * R + r, total - 1);
    const int s = stream_of(A, row / F), f = row % F;
    const float *src = A.fb.in_vec + ((size_t)s * (F + 2) + f) * kConvIn;  // rows f, f+1, f+2 are contiguous
    for (int i = tid; i < 3 * kConvIn; i += kFT) xt[i][r] = src[i];
    if (tid == 0) zero[r] = frame_count_at(A.st.frame_count[s], f) < 1;  // FEATURE_CONV1_DELAY (src/lpcnet.c:99)
  }
  __syncthreads();
  dense_rows<R, false>(A.m.conv1_w, A.m.conv1_b, 3 * kConvIn, kC, &xt[0][0], &y[0][0], kC, true, rcp, tid, kFT, zero);
  __syncthreads();
  for (int r = 0; r < R; r++) {
    const int row = blockIdx.x * R + r;
    if (row >= total) break;
    const int s = stream_of(A, row / F), f = row % F;
    for (int i = tid; i < kC; i += kFT) A.fb.c1[((size_t)s * (F + 2) + f + 2) * kC + i] = y[r][i];
  }
}

// K2: conv2 -> dense1 -> dense2 -> gru_a/gru_b conditioning for R rows per block.
__global__ void __launch_bounds__(kFT) cond_kernel(FrameArgs A) {
  constexpr int R = kRowsPerBlock;
  __shared__ __align__(16) float xt[3 * kC][R];
  __shared__ __align__(16) float yt[kC][R];
  __shared__ __align__(16) float zt[kC][R];
  __shared__ bool zero[R];
  __shared__ int row_s[R], row_f[R];
  __shared__ uint32_t rcp[2048];
  const int F = A.n_frames, B = A.n_streams, total = n_active(A) * F;
  const int tid = threadIdx.x;
  if (block_masked_out(A, R)) return;
  for (int i = tid; i < 2048; i += kFT) rcp[i] = A.m.rcp_tab[i];
  for (int r = 0; r < R; r++) {
    const int row = min(blockIdx.x * R + r, total - 1);
    const int s = stream_of(A, row / F), f = row % F;
    const float *src = A.fb.c1 + ((size_t)s * (F + 2) + f) * kC;
    for (int i = tid; i < 3 * kC; i += kFT) xt[i][r] = src[i];
    if (tid == 0) {
      zero[r] = frame_count_at(A.st.frame_count[s], f) < A.m.features_delay;  // src/lpcnet.c:101
      row_s[r] = s; row_f[r] = f;
    }
  }
  __syncthreads();
  dense_rows<R, true>(A.m.conv2_w, A.m.conv2_b, 3 * kC, kC, &xt[0][0], &yt[0][0], 0, true, rcp, tid, kFT, zero);
  __syncthreads();
  dense_rows<R, true>(A.m.dense1_w, A.m.dense1_b, kC, kC, &yt[0][0], &zt[0][0], 0, true, rcp, tid, kFT, nullptr);
  __syncthreads();
  dense_rows<R, true>(A.m.dense2_w, A.m.dense2_b, kC, kC, &zt[0][0], &yt[0][0], 0, true, rcp, tid, kFT, nullptr);
  __syncthreads();
  // END2END models (src/lpcnet.c:57-78, 108): the frame's LPC are the step-up recursion over the first 16 outputs of
  // dense2, taken as reflection coefficients -- no cepstrum, no delay line.  One thread per row; separate multiply and
  // add as in the reference build.
  if (A.m.end2end && tid < R && blockIdx.x * R + tid < total && !(A.stream_mask && !A.stream_mask[row_s[tid]])) {
    float cur[16], nxt[16];
#pragma unroll
    for (int i = 0; i < 16; i++) cur[i] = yt[i][tid];
#pragma unroll
    for (int i = 1; i < 16; i++) {
#pragma unroll
      for (int j = 0; j < i; j++) nxt[j] = cur[j] + cur[i] * cur[i - j - 1];
#pragma unroll
      for (int j = 0; j < i; j++) cur[j] = nxt[j];
    }
    float *dst = A.fb.lpc + ((size_t)row_f[tid] * B + row_s[tid]) * 16;
#pragma unroll
    for (int i = 0; i < 16; i++) dst[i] = cur[i];
  }
  // linear heads, written straight to [F][B][*]; two output columns per thread share the activation loads
  constexpr int NH = kRowsA + kRowsB;
  for (int i0 = tid; i0 < NH; i0 += 2 * kFT) {
    const int i1 = i0 + kFT < NH ? i0 + kFT : i0;   // the last pass may have one column only (computed twice, stored once)
    const float *Wc[2]; int colc[2], Nc[2]; float2 acc2[2][R / 2];
#pragma unroll
    for (int c = 0; c < 2; c++) {
      const int i = c ? i1 : i0;
      const bool isA = i < kRowsA;
      colc[c] = isA ? i : i - kRowsA; Nc[c] = isA ? kRowsA : kRowsB;
      Wc[c] = (isA ? A.m.a_feat_w : A.m.b_feat_w) + colc[c];
      const float b = __ldg((isA ? A.m.a_feat_b : A.m.b_feat_b) + colc[c]);
#pragma unroll
      for (int r = 0; r < R / 2; r++) acc2[c][r] = make_float2(b, b);
    }
    static_assert(kC % kPrefetch == 0, "prefetch depth must divide the layer width");
    for (int j0 = 0; j0 < kC; j0 += kPrefetch) {
      float w0[kPrefetch], w1[kPrefetch];
#pragma unroll
      for (int u = 0; u < kPrefetch; u++) {
        w0[u] = __ldg(Wc[0] + (size_t)(j0 + u) * Nc[0]);
        w1[u] = __ldg(Wc[1] + (size_t)(j0 + u) * Nc[1]);
      }
#pragma unroll
      for (int u = 0; u < kPrefetch; u++) {
        const int j = j0 + u;
        const float4 x0 = *reinterpret_cast<const float4 *>(&yt[j][0]), x1 = *reinterpret_cast<const float4 *>(&yt[j][4]);
        const float2 xp[R / 2] = {make_float2(x0.x, x0.y), make_float2(x0.z, x0.w), make_float2(x1.x, x1.y), make_float2(x1.z, x1.w)};
        const float2 wa = make_float2(w0[u], w0[u]), wb = make_float2(w1[u], w1[u]);
#pragma unroll
        for (int r = 0; r < R / 2; r++) { acc2[0][r] = f2_fma(wa, xp[r], acc2[0][r]); acc2[1][r] = f2_fma(wb, xp[r], acc2[1][r]); }
      }
    }
#pragma unroll
    for (int c = 0; c < 2; c++) {
      if (c == 1 && i1 == i0) break;
      const bool isA = Nc[c] == kRowsA;
#pragma unroll
      for (int r = 0; r < R; r++) {
        if (blockIdx.x * R + r >= total) break;
        // a stream the mask leaves out keeps the conditioning of its own last frame network (the sample segments
        // of the PLC run on it later: lpcnet_synthesize_tail_impl reads lpcnet->gru_a_condition)
        if (A.stream_mask && !A.stream_mask[row_s[r]]) continue;
        float *dst = isA ? A.fb.cond_a + ((size_t)row_f[r] * B + row_s[r]) * kRowsAP
                         : A.fb.cond_b + ((size_t)row_f[r] * B + row_s[r]) * kRowsB;
        dst[isA ? gru_a_col_index(colc[c]) : colc[c]] = (r & 1) ? acc2[c][r >> 1].y : acc2[c][r >> 1].x;   // GRU-A conditioning in the sample kernel's gather order
      }
    }
  }
}

// K3a: raw LPC per (stream, frame) into fb.lpc[f][s] (delay line applied by K3b).  One warp per row (dsp_warp.cuh).
constexpr int kDspWarps = 8;
__device__ __forceinline__ float lane_pick16(const float (&v)[16], int lane) {   // v[lane] without a local-memory array
  float r = v[0];
#pragma unroll
  for (int i = 1; i < 16; i++) r = lane == i ? v[i] : r;
  return r;
}
__global__ void __launch_bounds__(kDspWarps * 32) lpc_kernel(FrameArgs A) {
  __shared__ __align__(16) float sm[kDspWarps * kWarpDspFloats];
  const int B = A.n_streams, F = A.n_frames;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int idx = blockIdx.x * kDspWarps + w;
  if (idx >= n_active(A) * F) return;
  const int s = stream_of(A, idx / F), f = idx % F;
  if (A.stream_mask && !A.stream_mask[s]) return;      // keeps its LPC (see cond_kernel)
  const float *feat = A.features + ((size_t)s * A.total_frames + A.frame0 + f) * A.feature_stride;
  float lpc[16];
  w_lpc_from_cepstrum(DspTables{A.m.fft_tw, A.m.fft_bitrev, A.m.dct}, warp_dsp(sm, w, lane), feat, lpc);
  if (lane < 16) A.fb.lpc[((size_t)f * B + s) * 16 + lane] = lane_pick16(lpc, lane);
}

// K3b: the FEATURES_DELAY delay line (src/lpcnet.c:109-115), lpc_weighting (freq.c:299-308), and the
// per-stream frame-state update (conv histories, old_lpc, frame_count).  One WARP per stream: lane i < 16 owns LPC
// coefficient i through the whole chunk (every operation here is per coefficient; frames walked newest to oldest so
// the in-place shift never reads an overwritten row), all 32 lanes copy the history rows.  (One thread per stream
// took 94 us per launch whatever the batch: a serial walk over 440 floats of L2 round trips.)
__global__ void frame_finalize_kernel(FrameArgs A) {
  const int B = A.n_streams, F = A.n_frames, D = A.m.end2end ? 0 : A.m.features_delay;   // (END2END has no delay line)
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= n_active(A)) return;
  const int s = stream_of(A, k);
  if (A.stream_mask && !A.stream_mask[s]) return;   // this stream did not run a frame network
  if (lane < 16) {
    const int i = lane;
    if (D > 0) {
      float newest[kMaxDelay];   // new_lpc[F-1-k][i]
      for (int q = 0; q < D; q++)
        newest[q] = (F - 1 - q >= 0) ? A.fb.lpc[((size_t)(F - 1 - q) * B + s) * 16 + i]
                                     : A.st.old_lpc[((size_t)s * kMaxDelay + (q - F)) * 16 + i];
      for (int f = F - 1; f >= 0; f--)
        A.fb.lpc[((size_t)f * B + s) * 16 + i] = (f - D >= 0) ? A.fb.lpc[((size_t)(f - D) * B + s) * 16 + i]
                                                              : A.st.old_lpc[((size_t)s * kMaxDelay + (D - 1 - f)) * 16 + i];
      for (int q = 0; q < D; q++) A.st.old_lpc[((size_t)s * kMaxDelay + q) * 16 + i] = newest[q];
    }
    float gamma_i = A.m.lpc_gamma;                   // gamma^(i+1) by the reference's running product
    for (int q = 0; q < i; q++) gamma_i = gamma_i * A.m.lpc_gamma;
    for (int f = 0; f < F; f++) {
      float *dst = A.fb.lpc + ((size_t)f * B + s) * 16 + i;
      *dst = *dst * gamma_i;
    }
    // current-frame conditioning kept in the stream state like the reference (lpcnet_private.h:43-47)
    if (!A.state_only) A.st.lpc[(size_t)s * 16 + i] = A.fb.lpc[((size_t)(F - 1) * B + s) * 16 + i];
  }
  // histories: the last two in_vec / c1 rows (rows F and F+1 of the [F+2] buffers), contiguous in both places
  const float *iv = A.fb.in_vec + ((size_t)s * (F + 2) + F) * kConvIn;
  for (int i = lane; i < 2 * kConvIn; i += 32) A.st.conv1_state[(size_t)s * 2 * kConvIn + i] = iv[i];
  const float *cv = A.fb.c1 + ((size_t)s * (F + 2) + F) * kC;
  for (int i = lane; i < 2 * kC; i += 32) A.st.conv2_state[(size_t)s * 2 * kC + i] = cv[i];
  if (lane == 0) A.st.frame_count[s] = frame_count_at(A.st.frame_count[s], F);
}

// ---- packet decoder (src/lpcnet_dec.c:59-155, src/common.c:37-65), one thread per stream ----
__device__ __forceinline__ unsigned bits_take(const unsigned char *p, int &bitpos, int n) {
  unsigned d = 0;
  for (int k = 0; k < n; k++, bitpos++) d = (d << 1) | ((p[bitpos >> 3] >> (7 - (bitpos & 7))) & 1u);
  return d;
}

__device__ __forceinline__ void single_interp_dev(float *x, const float *left, const float *right, int id) {
  for (int i = 0; i < 18; i++) {
    if (id == 0) x[i] = .5f * (left[i] + right[i]);
    else if (id == 1) x[i] = left[i];
    else x[i] = right[i];
  }
}

__global__ void decode_packets_kernel(DecodeArgs A) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= A.n_streams) return;
  const float *cb1 = A.m.codebooks, *cb2 = cb1 + 1024 * 17, *cb3 = cb2 + 1024 * 17, *diff4 = cb3 + 1024 * 17;
  float vq_mem[18];
  for (int i = 0; i < 18; i++) vq_mem[i] = A.st.vq_mem[(size_t)s * 18 + i];
  for (int pk = 0; pk < A.n_packets; pk++) {
    const unsigned char *buf = A.packets + ((size_t)s * A.n_packets + pk) * 8;
    float *F = A.features + ((size_t)s * A.n_packets + pk) * 4 * 36;
    int bp = 0;
    const int c0_id = bits_take(buf, bp, 7);
    const int main_pitch = bits_take(buf, bp, 6);
    int modulation = bits_take(buf, bp, 3);
    const int corr_id = bits_take(buf, bp, 2);
    int vq_end[3];
    vq_end[0] = bits_take(buf, bp, 10); vq_end[1] = bits_take(buf, bp, 10); vq_end[2] = bits_take(buf, bp, 10);
    int vq_mid = bits_take(buf, bp, 13);
    int interp_id = bits_take(buf, bp, 3);
    for (int i = 0; i < 4 * 36; i++) F[i] = 0.f;
    int voiced = 1;
    modulation -= 4;
    if (modulation == -4) { voiced = 0; modulation = 0; }
    const float frame_corr = voiced ? 0.3875f + .175f * (float)corr_id : 0.0375f + .075f * (float)corr_id;
    for (int sub = 0; sub < 4; sub++) {
      float p = (float)(pow((double)2.f, main_pitch / 21.) * 32);
      p *= 1.f + (float)modulation / 16.f / 7.f * (float)(2 * sub - 3);
      p = 33 > p ? 33 : p;
      p = 255 < p ? 255 : p;
      F[sub * 36 + 18] = .02f * (p - 100.f);
      F[sub * 36 + 19] = frame_corr - .5f;
    }
    float *F0 = F, *F1 = F + 36, *F2 = F + 72, *F3 = F + 108;
    F3[0] = (float)(c0_id - 64) / 4.f;
    for (int i = 0; i < 17; i++) F3[i + 1] = cb1[vq_end[0] * 17 + i] + cb2[vq_end[1] * 17 + i] + cb3[vq_end[2] * 17 + i];
    float sign = 1;
    if (vq_mid >= 4096) { vq_mid -= 4096; sign = -1; }
    for (int i = 0; i < 18; i++) F1[i] = sign * diff4[vq_mid * 18 + i];
    if ((vq_mid & 3) < 2) {
      for (int i = 0; i < 18; i++) F1[i] += .5f * (vq_mem[i] + F3[i]);
    } else if ((vq_mid & 3) == 2) {
      for (int i = 0; i < 18; i++) F1[i] += vq_mem[i];
    } else {
      for (int i = 0; i < 18; i++) F1[i] += F3[i];
    }
    interp_id += (interp_id >= 7);
    single_interp_dev(F0, vq_mem, F1, interp_id / 3);
    single_interp_dev(F2, F1, F3, interp_id % 3);
    for (int i = 0; i < 18; i++) vq_mem[i] = F3[i];
  }
  for (int i = 0; i < 18; i++) A.st.vq_mem[(size_t)s * 18 + i] = vq_mem[i];
}

}  // namespace

cudaError_t launch_frame_network(const FrameArgs &a, cudaStream_t s, int *n_launches, const FrameFork *fork) {
  const int B = a.rows ? a.n_rows : a.n_streams, F = a.n_frames;     // streams this launch covers
  if (B <= 0) return cudaSuccess;
  const int rows = B * F;
  cudaError_t e;
  const bool lpc = !a.m.end2end;          // (END2END: cond_kernel writes the LPC itself)
  const bool forked = lpc && fork && fork->aux;
  if (lpc) {
    if (forked) {
      if ((e = cudaEventRecord(fork->begin, s)) != cudaSuccess) return e;
      if ((e = cudaStreamWaitEvent(fork->aux, fork->begin, 0)) != cudaSuccess) return e;
      lpc_kernel<<<(rows + kDspWarps - 1) / kDspWarps, kDspWarps * 32, 0, fork->aux>>>(a);
      if ((e = cudaEventRecord(fork->done, fork->aux)) != cudaSuccess) return e;
    }
  }
  frame_inputs_kernel<<<B * (F + 2), 128, 0, s>>>(a);
  conv1_kernel<<<(rows + kRowsPerBlock - 1) / kRowsPerBlock, kFT, 0, s>>>(a);
  cond_kernel<<<(rows + kRowsPerBlock - 1) / kRowsPerBlock, kFT, 0, s>>>(a);
  if (lpc && !forked) lpc_kernel<<<(rows + kDspWarps - 1) / kDspWarps, kDspWarps * 32, 0, s>>>(a);
  if (forked && (e = cudaStreamWaitEvent(s, fork->done, 0)) != cudaSuccess) return e;
  frame_finalize_kernel<<<(B * 32 + 127) / 128, 128, 0, s>>>(a);
  if (n_launches) *n_launches += lpc ? 5 : 4;
  return cudaGetLastError();
}

cudaError_t launch_decode_packets(const DecodeArgs &a, cudaStream_t s, int *n_launches) {
  decode_packets_kernel<<<(a.n_streams + 63) / 64, 64, 0, s>>>(a);
  if (n_launches) *n_launches += 1;
  return cudaGetLastError();
}

}  // namespace lpcb
