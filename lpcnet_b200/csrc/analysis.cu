// Per-frame feature analysis for many streams at once: the reference's
// lpcnet_compute_single_frame_features = preemphasis + compute_frame_features + process_single_frame
// (src/lpcnet_enc.c:872-880, 498-577, 814-870, 913-933) -- the step before the synthesis path ("-features"
// per frame) and an inner step of the PLC (src/lpcnet_plc.c:252-254, 424-432).
//
// What is sequential in the reference and what is not:
//  * pre-emphasis, the analysis window, the LPC residual (an FIR over the 80-sample-delayed signal) and its
//    one-tap smoothing are feed-forward: every sample / frame of a call is computed in parallel from the
//    signal and a 160-sample history (P), with the frame's own LPC;
//  * the pitch search is a Viterbi pass over half-frames: one block per stream walks the frames in order,
//    256 threads = 256 lags (cross-correlation, 3x interpolation, octave penalty, path update, argmax);
//  * the running energy under the correlation window is a double-precision recurrence over the lags
//    (lpcnet_enc.c:539-546): one thread walks it so that every addition rounds as in the reference.
// Every float operation is a separate IEEE operation in the reference's order (-fmad=false), so the features
// are bit-identical to the reference's parity build; log10() is evaluated in double as there.
#include <math.h>

#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "dsp.cuh"
#include "dsp_warp.cuh"
#include "engine.h"

namespace lpcb {

namespace {

constexpr int kWin = 320, kOverlap = 160, kTOff = 80, kPMax = 256, kPMin = 32, kBands = 18, kNF = 36;

struct EncState {        // per stream, device
  float *p_hist;         // [B][160]  analysis_mem: the last 160 pre-emphasised samples
  float *mem_preemph;    // [B]
  float *s_last;         // [B]       pitch_filt: the LPC residual of the last sample
  float *e_hist;         // [B][256]  exc_buf[160..416): the last 256 smoothed residual samples
  float *path;           // [B][256]  pitch_max_path[0]
  float *path_all;       // [B]
  int *best_i;           // [B]
  float *vq_mem;         // [B][18]   the encoder's vq_mem (superframe quantisation only)
};

struct AnalysisArgs {
  EncState st;
  DspTables tab;
  const float *half_window;   // [160]
  int n_streams, n_frames;    // frames in this chunk
  int frame0, total_frames;   // position of the chunk in the call (pcm / feature row strides)
  const short *pcm16;         // device [B][total_frames*160], or NULL
  const float *pcmf;          // device [B][total_frames*160], or NULL
  float *features;            // device [B][total_frames][36]
  float *P;                   // [B][160 + 160F]  history | pre-emphasised signal
  float *S;                   // [B][1 + 160F]    previous residual | LPC residual
  float *fft_scratch;         // [320][2][B*F]
  const uint8_t *stream_mask; // nullable [B]: streams with 0 keep their state and their feature rows
  const int *rows;            // nullable: device list of the n_rows streams to process -- the grids then cover the list
  int n_rows;                 //   (the PLC's small classes: a masked grid over the whole batch costs a wave per kernel)
  // superframe modes (lpcnet_encode / lpcnet_compute_features, lpcnet_enc.c:882-911): frames come in packets of four
  int quantize;               // 1: process_superframe(encode = quantize = 1), 0: (0, 0)
  int *ids;                   // [B][total_frames/4][kIds] bit-field values of the packet (quantize mode)
  const float *pitch_pow;     // [64] (float)(pow(2.f, main_pitch/21.)*PITCH_MIN_PERIOD), by the host's libm
};

constexpr int kIds = 10;      // c0_id+64, main_pitch, modulation field, corr_id, vq_end[3], vq_mid, interp_id, (spare)

// (int)x of the reference build for a double that may be NaN or out of range: x86 cvttsd2si gives INT_MIN
__device__ __forceinline__ int cvt_x86(double v) {
  return (v != v || v >= 2147483648.0 || v <= -2147483649.0) ? (int)0x80000000 : (int)v;
}

__device__ __forceinline__ float pcm_at(const AnalysisArgs &A, int s, int n) {
  const size_t i = (size_t)s * A.total_frames * kFrame + (size_t)A.frame0 * kFrame + n;
  return A.pcm16 ? (float)A.pcm16[i] : A.pcmf[i];
}

__device__ __forceinline__ int an_count(const AnalysisArgs &A) { return A.rows ? A.n_rows : A.n_streams; }
__device__ __forceinline__ int an_stream(const AnalysisArgs &A, int k) { return A.rows ? __ldg(A.rows + k) : k; }

// preemphasis (lpcnet_enc.c:872-880): y[n] = x[n] + mem, mem = -0.85f * x[n]
__global__ void preemph_kernel(AnalysisArgs A) {
  const int L = kOverlap + A.n_frames * kFrame;
  const size_t total = (size_t)an_count(A) * L;
  for (size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i0 < total; i0 += (size_t)gridDim.x * blockDim.x) {
    const int s = an_stream(A, (int)(i0 / L)), k = (int)(i0 % L);
    if (A.stream_mask && !A.stream_mask[s]) continue;   // (its rows may belong to a concurrent call on another CUDA stream)
    const size_t i = (size_t)s * L + k;
    float v;
    if (k < kOverlap) {
      v = A.st.p_hist[(size_t)s * kOverlap + k];
    } else {
      const int n = k - kOverlap;
      const float mem = n == 0 ? A.st.mem_preemph[s] : -0.85f * pcm_at(A, s, n - 1);
      v = pcm_at(A, s, n) + mem;
    }
    A.P[i] = v;
  }
}

// frame_analysis + the spectral half of compute_frame_features (lpcnet_enc.c:488-527): window, forward FFT,
// band energies, log smoothing, DCT, LPC from the cepstrum.  One WARP per (stream, frame) (dsp_warp.cuh): lanes share
// the transform's butterflies, one band per lane for the band sums / logs / DCT.
constexpr int kSpecWarps = 8;
__global__ void __launch_bounds__(kSpecWarps * 32) spectral_kernel(AnalysisArgs A) {
  __shared__ __align__(16) float sm[kSpecWarps * kWarpDspFloats];
  const int F = A.n_frames;
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const int idx = blockIdx.x * kSpecWarps + wi;
  if (idx >= an_count(A) * F) return;
  const int s = an_stream(A, idx / F), f = idx % F;
  if (A.stream_mask && !A.stream_mask[s]) return;
  const WarpDsp W = warp_dsp(sm, wi, lane);
  const float *x = A.P + (size_t)s * (kOverlap + F * kFrame) + (size_t)f * kFrame;   // [analysis_mem | frame]
  // apply_window (freq.c:322-328) and opus_fft_c's scaled bit-reversed copy (kiss_fft.c:566-585)
  const float scale = 1.f / 320;
  for (int k = lane; k < kWin; k += 32) {
    const float w = __ldg(A.half_window + (k < kOverlap ? k : kWin - 1 - k));
    W.x[__ldg(A.tab.fft_bitrev + k)] = make_float2(scale * (x[k] * w), scale * 0.f);
  }
  __syncwarp();
  wfft320(W, A.tab.fft_tw);
  // lpcn_compute_band_energy (freq.c:128-150), band `lane`; its log (lpcnet_enc.c:516)
  const float sum = w_band_energy<false>(W);
  float *Ly = W.f + 18, *ceps = Ly;                  // (W.f[0..18) and [36..53) belong to w_lpc_from_cepstrum)
  if (lane < kBands) Ly[lane] = (float)log10(1e-2 + (double)sum);
  __syncwarp();
  // lpcnet_enc.c:514-522: the smoothing walks the bands in order (every lane runs it; lane i keeps band i's value)
  float mine = 0.f, logMax = -2, follow = -2;
  for (int i = 0; i < kBands; i++) {
    float l = Ly[i];
    l = fmaxf(logMax - 8, fmaxf(follow - 2.5f, l));
    if (i == lane) mine = l;
    logMax = fmaxf(logMax, l);
    follow = fmaxf(follow - 2.5f, l);
  }
  __syncwarp();
  if (lane < kBands) Ly[lane] = mine;
  __syncwarp();
  // dct (freq.c:218-228)
  float c = w_dct(A.tab.dct, Ly, lane);
  if (lane == 0) c -= 4;
  __syncwarp();
  if (lane < kBands) ceps[lane] = c;
  __syncwarp();
  float lpc[16];
  w_lpc_from_cepstrum(A.tab, W, ceps, lpc);
  float *out = A.features + ((size_t)s * A.total_frames + A.frame0 + f) * kNF;
  if (lane < kBands) out[lane] = c;
  if (lane < 16) {
    float v = lpc[0];
#pragma unroll
    for (int i = 1; i < 16; i++) v = lane == i ? lpc[i] : v;
    out[kBands + 2 + lane] = v;
  }
}

// The LPC residual of the 80-sample-delayed signal (lpcnet_enc.c:528-539): an FIR, so every sample is independent.
// S[s][1 + n] = a[n] + sum_j lpc[j] * a[n-1-j], a[n] = P[80 + n]; S[s][0] = the previous call's last value.
__global__ void residual_kernel(AnalysisArgs A) {
  const int F = A.n_frames, L = F * kFrame;
  const size_t total = (size_t)an_count(A) * L;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int s = an_stream(A, (int)(i / L)), n = (int)(i % L), f = n / kFrame;
    if (A.stream_mask && !A.stream_mask[s]) continue;
    const float *a = A.P + (size_t)s * (kOverlap + L) + kTOff + n;
    const float *lpc = A.features + ((size_t)s * A.total_frames + A.frame0 + f) * kNF + kBands + 2;
    float sum = a[0];
    for (int j = 0; j < 16; j++) sum += lpc[j] * a[-1 - j];
    A.S[(size_t)s * (L + 1) + 1 + n] = sum;
    if (n == 0) A.S[(size_t)s * (L + 1)] = A.st.s_last[s];
  }
}

// The pitch half of compute_frame_features (lpcnet_enc.c:540-576) and the Viterbi pass over half-frames; one block
// per stream.  SUPER = false: process_single_frame (814-870), two half-frames per frame.  SUPER = true:
// process_superframe (579-743), eight half-frames per packet of four frames, then the pitch contour regression and
// -- in quantize mode -- the pitch / correlation fields of the packet.
//
// Only the Viterbi pass is sequential over the frames.  The cross-correlations and the energy recurrences of the
// chunk's half-frames ("rows") depend on the residual alone, so a group of up to 16 frames is prepared at once:
//  * warps 0-7, thread t = lag t: the 80-term correlation of every row of the group as a j-ascending float sum
//    (celt_pitch_xcorr's order), the frame's samples read as broadcast 16-byte loads;
//  * warp 8, lane r = row r: the row's energies -- the double-precision recurrence over the 256 lags
//    (lpcnet_enc.c:542-551) is a serial chain, so the 32 rows walk it side by side instead of one after the other
//    (one thread per block did that before: 72 % of the kernel's stall samples sat at the barrier behind it).
//    Rows are 80 floats apart, i.e. 16 banks: lanes run skewed by lane/2 steps so that a warp load touches 32 banks.
// Then the frames of the group follow in order: normalisation, 3x interpolation, octave penalty, path update.
constexpr int kPT = 256 + 32, kGroup = 16;
template <bool SUPER>
__global__ void __launch_bounds__(kPT) pitch_kernel(AnalysisArgs A) {
  constexpr int NR = SUPER ? 8 : 2;         // correlation rows (half-frames) per Viterbi pass
  extern __shared__ float sm[];
  const int s = an_stream(A, blockIdx.x), t = threadIdx.x, F = A.n_frames, L = F * kFrame;
  if (A.stream_mask && !A.stream_mask[s]) return;
  const int GR = F < kGroup ? F : kGroup;   // frames per group
  const bool lagthr = t < kPMax;
  float *ebuf = sm;                         // [256 + L]
  float *xcg = ebuf + kPMax + L;            // [2*GR][256] correlations, then the normalised / interpolated rows
  float *enr = xcg + 2 * GR * kPMax;        // [2*GR][256] 1 + ener0 + ener1 of every lag, as the reference's float
  float *path0 = enr + 2 * GR * kPMax;      // [256]
  float *path1 = path0 + kPMax;             // [256]
  int *prev = reinterpret_cast<int *>(path1 + kPMax);          // [NR][256]
  float *redv = reinterpret_cast<float *>(prev + NR * kPMax);  // [8]
  int *redi = reinterpret_cast<int *>(redv + 8);               // [8]
  float *e0s = reinterpret_cast<float *>(redi + 8);            // [2*GR] frame_weight before normalisation
  __shared__ float path_all_s;
  __shared__ int best_i_s;

  const float *Ss = A.S + (size_t)s * (L + 1);
  if (lagthr) {
    ebuf[t] = A.st.e_hist[(size_t)s * kPMax + t];
    path0[t] = A.st.path[(size_t)s * kPMax + t];
    path1[t] = 0.f;
  }
  for (int n = t; n < L; n += kPT) ebuf[kPMax + n] = Ss[n + 1] + .7f * Ss[n];
  if (t == 0) { path_all_s = A.st.path_all[s]; best_i_s = A.st.best_i[s]; }
  __syncthreads();

  for (int g0 = 0; g0 < F; g0 += GR) {
    const int gf = F - g0 < GR ? F - g0 : GR, rows = 2 * gf;
    if (lagthr) {
      // celt_pitch_xcorr (pitch.c:43-83, pitch.h:42-105): every lag is a j-ascending sum from zero
      for (int r = 0; r < rows; r++) {
        const float *ys = ebuf + g0 * kFrame + r * (kFrame / 2), *xs = ys + kPMax;
        float acc = 0;
#pragma unroll 5
        for (int j = 0; j < kFrame / 2; j += 4) {
          const float4 x4 = *reinterpret_cast<const float4 *>(xs + j);
          acc = acc + x4.x * ys[t + j];
          acc = acc + x4.y * ys[t + j + 1];
          acc = acc + x4.z * ys[t + j + 2];
          acc = acc + x4.w * ys[t + j + 3];
        }
        xcg[r * kPMax + t] = acc;
      }
    } else {
      const int r = t - kPMax, skew = r >> 1;
      const bool on = r < rows;
      const float *ys = ebuf + g0 * kFrame + (on ? r : 0) * (kFrame / 2), *xs = ys + kPMax;
      // ener0 = celt_inner_prod(x, x, 80), ener1 = celt_inner_prod(y, y, 79) (lpcnet_enc.c:540-541): float sums
      float e0 = 0, e1f = 0;
      for (int st = 0; st < kFrame / 2 + 16; st++) {
        const int j = st - skew;
        if (on && j >= 0 && j < kFrame / 2) {
          e0 = e0 + xs[j] * xs[j];
          if (j < kFrame / 2 - 1) e1f = e1f + ys[j] * ys[j];
        }
      }
      // ener1 (double) walks the lags: += y[i+79]^2 before lag i, -= y[i]^2 after it (lpcnet_enc.c:542-551)
      double e1 = e1f;
      float *out = enr + (on ? r : 0) * kPMax;
      for (int st = 0; st < kPMax + 16; st++) {
        const int i = st - skew;
        if (on && i >= 0 && i < kPMax) {
          const float a = ys[i + kFrame / 2 - 1], b = ys[i];
          e1 += (double)(a * a);
          out[i] = (float)((double)(1 + e0) + e1);
          e1 -= (double)(b * b);
        }
      }
      if (on) e0s[r] = e0;
    }
    __syncthreads();
    for (int fl = 0; fl < gf; fl++) {
      const int f = g0 + fl;
      // rows 2+2*pcount+sub of the reference's xc / frame_weight: the packet's (SUPER) or the frame's rows
      const int pr0 = SUPER ? 2 * (fl & ~3) : 2 * fl;
      float *xc = xcg + pr0 * kPMax;
      const float *ener0_s = e0s + pr0;
      for (int sub = 0; sub < 2; sub++) {
        const int row = 2 * fl + sub;
        float *c = xcg + row * kPMax;
        if (lagthr) c[t] = 2 * c[t] / enr[row * kPMax + t];
        __syncthreads();
        // 3x upsampled correlation, keep the max (lpcnet_enc.c:552-569)
        float up = 0.f;
        if (t >= 4 && t < kPMax - 4) {
          const float interp[7] = {0.026184f, -0.098339f, 0.369938f, 0.837891f, -0.184969f, 0.070242f, -0.020947f};
          float v1 = 0, v2 = 0;
#pragma unroll
          for (int j = 0; j < 7; j++) {
            v1 += c[t - 3 + j] * interp[j];
            v2 += c[t + 3 - j] * interp[j];
          }
          up = fmaxf(c[t], fmaxf(v1, v2));
        }
        __syncthreads();
        if (t >= 4 && t < kPMax - 4) c[t] = up;
      }
      __syncthreads();
      if (SUPER && (fl & 3) != 3) continue;     // the superframe pass runs after the packet's fourth frame
      // process_single_frame (lpcnet_enc.c:814-870) / process_superframe (579-630): the same pass over 2 / 8 rows
      float fws = 1e-15f;
#pragma unroll
      for (int r = 0; r < NR; r++) fws += ener0_s[r];
      float fw[NR];
#pragma unroll
      for (int r = 0; r < NR; r++) fw[r] = ener0_s[r] * ((float)NR / fws);
#pragma unroll
      for (int sub = 0; sub < NR; sub++) {
        float *c = xc + sub * kPMax;
        // octave penalty: the values it compares against sit at higher lags than the one it scales
        float cv = lagthr ? c[t] : 0.f;
        if (t < kPMax - 2 * kPMin) {
          const float half = fmaxf(fmaxf(c[(kPMax + t) / 2], c[(kPMax + t + 2) / 2]), c[(kPMax + t - 1) / 2]);
          if (cv < half * 1.1f) cv *= .8f;
        }
        __syncthreads();
        if (lagthr) c[t] = cv;
        float v = -3.0e38f;
        if (t < kPMax - kPMin) {
          float max_prev = path_all_s - 6.f;
          int pp = best_i_s;
          for (int j = (t < 4 ? -t : -4); j <= 4 && t + j < kPMax - kPMin; j++) {
            const int aj = j < 0 ? -j : j;
            const float cand = path0[t + j] - .02f * (float)aj * (float)aj;
            if (cand > max_prev) { max_prev = cand; pp = t + j; }
          }
          prev[sub * kPMax + t] = pp;
          v = max_prev + fw[sub] * cv;
          path1[t] = v;
        }
        // first maximum over the lags (strict '>' in lag order, starting from -1e15 / lag 0)
        int bi = t;
#pragma unroll
        for (int o = 16; o; o >>= 1) {
          const float ov = __shfl_down_sync(0xffffffffu, v, o);
          const int oi = __shfl_down_sync(0xffffffffu, bi, o);
          if (ov > v || (ov == v && oi < bi)) { v = ov; bi = oi; }
        }
        if (lagthr && (t & 31) == 0) { redv[t >> 5] = v; redi[t >> 5] = bi; }
        __syncthreads();
        if (t == 0) {
          float mv = -1e15f;
          int mi = 0;
          for (int w = 0; w < kPMax / 32; w++)
            if (redv[w] > mv) { mv = redv[w]; mi = redi[w]; }
          path_all_s = mv; best_i_s = mi;
        }
        __syncthreads();
        if (t < kPMax - kPMin) path1[t] -= path_all_s;
        if (lagthr) path0[t] = path1[t];
        __syncthreads();
      }
      if (t == 0) {
        int bi = best_i_s, best[NR];
        float frame_corr = 0;
#pragma unroll
        for (int sub = NR - 1; sub >= 0; sub--) {
          best[sub] = kPMax - bi;
          frame_corr += fw[sub] * xc[sub * kPMax + bi];
          bi = prev[sub * kPMax + bi];
        }
        frame_corr /= (float)NR;
        if (!SUPER) {
          float *out = A.features + ((size_t)s * A.total_frames + A.frame0 + f) * kNF;
          out[kBands] = .01f * (float)(max(66, min(510, best[0] + best[1])) - 200);
          out[kBands + 1] = frame_corr - .5f;
        } else {
          // lpcnet_enc.c:640-699: weighted linear regression of the pitch contour, then its quantisation
          const int quantize = A.quantize;
          if (quantize && frame_corr < 0) frame_corr = 0;
          float sx = 0, sxx = 0, sxy = 0, sy = 0, sw = 0;
#pragma unroll
          for (int sub = 2; sub < 10; sub++) {
            const float w = fw[sub - 2];
            sw += w;
            sx += w * (float)sub;
            sxx += w * (float)sub * (float)sub;
            sxy += w * (float)sub * (float)best[sub - 2];
            sy += w * (float)best[sub - 2];
          }
          const int voiced = (double)frame_corr >= .3;
          float best_a = (sw * sxy - sx * sy) / (sw * sxx - sx * sx);
          int corr_id;
          if (voiced) {
            const float mean_pitch = sy / sw;
            const float max_a = mean_pitch / 32.f;
            const float lo = -max_a > best_a ? -max_a : best_a;
            best_a = max_a < lo ? max_a : lo;
            corr_id = cvt_x86(floor((double)((frame_corr - .3f) / .175f)));
            if (quantize) frame_corr = 0.3875f + .175f * (float)corr_id;
          } else {
            best_a = 0;
            corr_id = cvt_x86(floor((double)(frame_corr / .075f)));
            if (quantize) frame_corr = 0.0375f + .075f * (float)corr_id;
          }
          const float best_b = (sy - best_a * sx) / sw;
          const float center_pitch = best_b + 5.5f * best_a;
          int main_pitch = cvt_x86(floor(.5 + 21. * 1.442695041 * log((double)(center_pitch / 32.f))));
          main_pitch = max(0, min(63, main_pitch));
          int modulation = cvt_x86(floor(.5 + (double)(112.f * best_a / center_pitch)));
          modulation = max(-3, min(3, modulation));
          const int fr0 = A.frame0 + f - 3;
          for (int sub = 0; sub < 4; sub++) {
            float *out = A.features + ((size_t)s * A.total_frames + fr0 + sub) * kNF;
            if (quantize) {
              float p = __ldg(A.pitch_pow + main_pitch);
              p *= 1.f + (float)modulation / 16.f / 7.f * (float)(2 * sub - 3);
              const float lo = 33.f > p ? 33.f : p;
              p = 255.f < lo ? 255.f : lo;
              out[kBands] = .02f * (p - 100.f);
            } else {
              out[kBands] = .01f * (float)(max(66, min(510, best[2 * sub] + best[2 * sub + 1])) - 200);
            }
            out[kBands + 1] = frame_corr - .5f;
          }
          if (quantize) {
            int *id = A.ids + ((size_t)s * (A.total_frames / 4) + fr0 / 4) * kIds;
            id[1] = main_pitch;
            id[2] = voiced ? modulation + 4 : 0;
            id[3] = corr_id;
          }
        }
      }
      __syncthreads();
    }
  }
  // carry the stream state to the next call
  if (lagthr) {
    A.st.e_hist[(size_t)s * kPMax + t] = ebuf[L + t];
    A.st.path[(size_t)s * kPMax + t] = path0[t];
  }
  if (t < kOverlap) A.st.p_hist[(size_t)s * kOverlap + t] = A.P[(size_t)s * (kOverlap + L) + L + t];
  if (t == 0) {
    A.st.path_all[s] = path_all_s;
    A.st.best_i[s] = best_i_s;
    A.st.s_last[s] = Ss[L];
    A.st.mem_preemph[s] = -0.85f * pcm_at(A, s, L - 1);
  }
}

size_t pitch_smem(int F, int rows) {
  const int GR = F < kGroup ? F : kGroup;
  return (size_t)(kPMax + F * kFrame + 2 * (2 * GR * kPMax) + 2 * kPMax + rows * kPMax + 16 + 2 * GR) * 4 + 64;
}

#include "vq.cuh"

}  // namespace

}  // namespace lpcb

using namespace lpcb;

// ---- host side: the batched handle and the reference's encoder-state entry points -------------------------
struct LPCNetFeatureBatch {
  int device = 0, n_streams = 0;
  EncState st{};
  void *st_alloc = nullptr;
  size_t st_bytes = 0;
  void *tab_alloc = nullptr;
  DspTables tab{};
  const float *half_window = nullptr;
  cudaStream_t stream = nullptr;
  void *work = nullptr; size_t work_bytes = 0;
  void *d_pcm = nullptr; size_t d_pcm_bytes = 0;
  void *d_feat = nullptr; size_t d_feat_bytes = 0;
  // 1.6 kb/s encoder (lpcnet_encode): codebooks, per-packet field values, packets
  float *d_cb = nullptr;
  const float *pitch_pow = nullptr;
  void *d_ids = nullptr; size_t d_ids_bytes = 0;
  void *d_left0 = nullptr; size_t d_left0_bytes = 0;
  void *d_packets = nullptr; size_t d_packets_bytes = 0;
  long long launches = 0;
};

namespace {

#define CKA(call)                                                                                      \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess) return set_last_error(std::string(#call) + ": " + cudaGetErrorString(e_));  \
  } while (0)

int grow_buf(void **p, size_t *have, size_t want) {
  if (*have >= want) return 0;
  if (*p) cudaFree(*p);
  *p = nullptr; *have = 0;
  CKA(cudaMalloc(p, want));
  *have = want;
  return 0;
}

int feature_chunk_frames(int B, int n_frames, bool super) {
  int F = 32768 / (B > 0 ? B : 1);
  F = F < 1 ? 1 : (F > 32 ? 32 : F);
  if (super) F = F < 4 ? 4 : (F & ~3);     // whole packets of four frames
  return F < n_frames ? F : n_frames;
}

// mode 0: single-frame analysis; 1: superframes (n_frames a multiple of 4), unquantised (lpcnet_compute_features);
// 2: superframes with the pitch / correlation fields quantised into b->d_ids (lpcnet_encode)
int analysis_device(LPCNetFeatureBatch *b, const short *d16, const float *df, int n_frames, float *d_features,
                    const uint8_t *d_mask = nullptr, int mode = 0, const int *d_rows = nullptr, int n_rows = 0) {
  const int B = b->n_streams;
  const int NA = d_rows ? n_rows : B;            // streams the grids cover
  if (NA <= 0) return 0;
  CKA(cudaSetDevice(b->device));
  const int Fc = feature_chunk_frames(B, n_frames, mode != 0);
  const size_t nP = (size_t)B * (kOverlap + Fc * kFrame), nS = (size_t)B * (1 + Fc * kFrame);
  const size_t want = (nP + nS) * 4 + 1024;      // (the transforms' scratch lives in shared memory: dsp.cuh block_scratch)
  if (grow_buf(&b->work, &b->work_bytes, want)) return -1;
  for (int f0 = 0; f0 < n_frames; f0 += Fc) {
    const int F = n_frames - f0 < Fc ? n_frames - f0 : Fc;
    AnalysisArgs A{};
    A.st = b->st; A.tab = b->tab; A.half_window = b->half_window;
    A.n_streams = B; A.n_frames = F; A.frame0 = f0; A.total_frames = n_frames;
    A.pcm16 = d16; A.pcmf = df; A.features = d_features; A.stream_mask = d_rows ? nullptr : d_mask;
    A.rows = d_rows; A.n_rows = n_rows;
    A.quantize = mode == 2; A.ids = static_cast<int *>(b->d_ids); A.pitch_pow = b->pitch_pow;
    A.P = static_cast<float *>(b->work);
    A.S = A.P + (((size_t)B * (kOverlap + F * kFrame) + 63) & ~size_t(63));
    A.fft_scratch = A.S + (((size_t)B * (1 + F * kFrame) + 63) & ~size_t(63));
    const size_t np = (size_t)NA * (kOverlap + F * kFrame);
    preemph_kernel<<<(unsigned)((np + 255) / 256 > 65535 * 16 ? 65535 * 16 : (np + 255) / 256), 256, 0, b->stream>>>(A);
    spectral_kernel<<<(NA * F + kSpecWarps - 1) / kSpecWarps, kSpecWarps * 32, 0, b->stream>>>(A);
    const size_t nr = (size_t)NA * F * kFrame;
    residual_kernel<<<(unsigned)((nr + 255) / 256 > 65535 * 16 ? 65535 * 16 : (nr + 255) / 256), 256, 0, b->stream>>>(A);
    if (mode) pitch_kernel<true><<<NA, kPT, pitch_smem(F, 8), b->stream>>>(A);
    else pitch_kernel<false><<<NA, kPT, pitch_smem(F, 2), b->stream>>>(A);    // (> 48 KB: opted in at create)
    b->launches += 4;
    CKA(cudaGetLastError());
  }
  return 0;
}

int analysis_host(LPCNetFeatureBatch *b, const short *p16, const float *pf, int n_frames, float *features) {
  if (!b || (!p16 && !pf) || !features || n_frames < 0) return set_last_error("lpcnet_feature_batch_compute: bad arguments");
  if (n_frames == 0) return 0;
  const size_t ns = (size_t)b->n_streams * n_frames * kFrame, nf = (size_t)b->n_streams * n_frames * kNF;
  CKA(cudaSetDevice(b->device));
  if (grow_buf(&b->d_pcm, &b->d_pcm_bytes, ns * (p16 ? 2 : 4))) return -1;
  if (grow_buf(&b->d_feat, &b->d_feat_bytes, nf * 4)) return -1;
  CKA(cudaMemcpyAsync(b->d_pcm, p16 ? (const void *)p16 : (const void *)pf, ns * (p16 ? 2 : 4), cudaMemcpyHostToDevice, b->stream));
  CKA(cudaMemsetAsync(b->d_feat, 0, nf * 4, b->stream));
  if (analysis_device(b, p16 ? static_cast<const short *>(b->d_pcm) : nullptr,
                      p16 ? nullptr : static_cast<const float *>(b->d_pcm), n_frames, static_cast<float *>(b->d_feat)))
    return -1;
  CKA(cudaMemcpyAsync(features, b->d_feat, nf * 4, cudaMemcpyDeviceToHost, b->stream));
  CKA(cudaStreamSynchronize(b->stream));
  return 0;
}

// lpcnet_encode / lpcnet_compute_features for every stream: n_packets packets of 640 samples (lpcnet_enc.c:882-911)
int superframe_host(LPCNetFeatureBatch *b, const short *pcm, int n_packets, unsigned char *packets, float *features) {
  const bool quantize = packets != nullptr;
  if (!b || !pcm || n_packets < 0 || (!packets && !features)) return set_last_error("lpcnet_feature_batch_encode: bad arguments");
  if (quantize && !b->d_cb)
    return set_last_error("lpcnet_feature_batch_encode: no codebooks (lpcnet_feature_batch_set_codebooks / lpcnet_b200_set_codebooks)");
  if (n_packets == 0) return 0;
  const int B = b->n_streams, n_frames = 4 * n_packets;
  const size_t ns = (size_t)B * n_frames * kFrame, nf = (size_t)B * n_frames * kNF;
  CKA(cudaSetDevice(b->device));
  if (grow_buf(&b->d_pcm, &b->d_pcm_bytes, ns * 2)) return -1;
  if (grow_buf(&b->d_feat, &b->d_feat_bytes, nf * 4)) return -1;
  if (quantize) {
    if (grow_buf(&b->d_ids, &b->d_ids_bytes, (size_t)B * n_packets * kIds * 4)) return -1;
    if (grow_buf(&b->d_left0, &b->d_left0_bytes, (size_t)B * kBands * 4)) return -1;
    if (grow_buf(&b->d_packets, &b->d_packets_bytes, (size_t)B * n_packets * 8)) return -1;
  }
  CKA(cudaMemcpyAsync(b->d_pcm, pcm, ns * 2, cudaMemcpyHostToDevice, b->stream));
  CKA(cudaMemsetAsync(b->d_feat, 0, nf * 4, b->stream));
  float *d_feat = static_cast<float *>(b->d_feat);
  if (analysis_device(b, static_cast<const short *>(b->d_pcm), nullptr, n_frames, d_feat, nullptr, quantize ? 2 : 1)) return -1;
  if (quantize) {
    VqArgs V{};
    V.cb = b->d_cb; V.features = d_feat; V.ids = static_cast<int *>(b->d_ids); V.vq_mem = b->st.vq_mem;
    V.left0 = static_cast<float *>(b->d_left0); V.packets = static_cast<unsigned char *>(b->d_packets);
    V.n_streams = B; V.n_packets = n_packets;
    const unsigned grid = (unsigned)(((size_t)B * n_packets + kVqWarps - 1) / kVqWarps);
    vq_end_kernel<<<grid, kVqThreads, kVqEndSmem, b->stream>>>(V);
    vq_mid_kernel<<<grid, kVqThreads, kVqMidSmem, b->stream>>>(V);
    b->launches += 2;
    CKA(cudaGetLastError());
    CKA(cudaMemcpyAsync(packets, b->d_packets, (size_t)B * n_packets * 8, cudaMemcpyDeviceToHost, b->stream));
  }
  if (features) CKA(cudaMemcpyAsync(features, b->d_feat, nf * 4, cudaMemcpyDeviceToHost, b->stream));
  CKA(cudaStreamSynchronize(b->stream));
  return 0;
}

int upload_codebooks(LPCNetFeatureBatch *b, const float *codebooks, int n_floats) {
  const int want = 3 * kCbN * kNb1 + 4096 * kBands;
  if (!b || !codebooks || n_floats != want)
    return set_last_error("codebooks: expected ceps_codebook1|2|3 [1024][17] + ceps_codebook_diff4 [4096][18] = 125952 floats");
  CKA(cudaSetDevice(b->device));
  if (!b->d_cb) CKA(cudaMalloc(&b->d_cb, (size_t)want * 4));
  CKA(cudaMemcpy(b->d_cb, codebooks, (size_t)want * 4, cudaMemcpyHostToDevice));
  return 0;
}

std::mutex g_cb_mu;
std::vector<float> g_codebooks;     // process-wide default (the reference links ceps_codebooks.c into the library)

}  // namespace

int lpcb::feature_batch_frame_masked(LPCNetFeatureBatch *b, const short *d_pcm16, const uint8_t *d_mask, float *d_features,
                                     cudaStream_t stream, const int *d_rows, int n_rows) {
  cudaStream_t keep = b->stream;
  b->stream = stream;
  const int r = analysis_device(b, d_pcm16, nullptr, 1, d_features, d_mask, 0, d_rows, n_rows);
  b->stream = keep;
  return r;
}

extern "C" {

LPCNetFeatureBatch *lpcnet_feature_batch_create(int n_streams, int device) {
  if (n_streams <= 0) { set_last_error("lpcnet_feature_batch_create: n_streams must be positive"); return nullptr; }
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
    set_last_error("lpcnet_feature_batch_create: no such CUDA device (there is no CPU fallback)");
    return nullptr;
  }
  LPCNetFeatureBatch *b = new (std::nothrow) LPCNetFeatureBatch();
  if (!b) return nullptr;
  b->device = device; b->n_streams = n_streams;
  auto bail = [&](const char *what, cudaError_t e) -> LPCNetFeatureBatch * {
    set_last_error(std::string(what) + ": " + cudaGetErrorString(e));
    lpcnet_feature_batch_destroy(b);
    return nullptr;
  };
  cudaError_t e;
  if ((e = cudaSetDevice(device)) != cudaSuccess) return bail("cudaSetDevice", e);
  if ((e = cudaStreamCreateWithFlags(&b->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail("cudaStreamCreate", e);
  // tables: kiss_fft twiddles / bit reversal, DCT, and the analysis half-window (dump_lpcnet_tables.c:83-84)
  std::vector<float> tw, dct, hw(kOverlap);
  std::vector<int16_t> br;
  host_dsp_tables(tw, br, dct);
  for (int i = 0; i < kOverlap; i++)
    hw[i] = (float)sin(.5 * M_PI * sin(.5 * M_PI * (i + .5) / kOverlap) * sin(.5 * M_PI * (i + .5) / kOverlap));
  // (float)(pow(2.f, main_pitch/21.)*PITCH_MIN_PERIOD) (lpcnet_enc.c:687), by the host's libm as in the reference
  std::vector<float> pp(64);
  for (int i = 0; i < 64; i++) pp[i] = (float)(pow(2.f, i / 21.) * kPMin);
  const size_t o_tw = 0, o_dct = o_tw + tw.size() * 4, o_hw = o_dct + dct.size() * 4, o_pp = o_hw + hw.size() * 4,
               o_br = o_pp + pp.size() * 4;
  if ((e = cudaMalloc(&b->tab_alloc, o_br + br.size() * 2)) != cudaSuccess) return bail("cudaMalloc(tables)", e);
  char *tp = static_cast<char *>(b->tab_alloc);
  cudaMemcpy(tp + o_pp, pp.data(), pp.size() * 4, cudaMemcpyHostToDevice);
  b->pitch_pow = (const float *)(tp + o_pp);
  if ((e = cudaFuncSetAttribute(pitch_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pitch_smem(32, 8))) != cudaSuccess ||
      (e = cudaFuncSetAttribute(pitch_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pitch_smem(32, 2))) != cudaSuccess)
    return bail("cudaFuncSetAttribute(pitch kernels)", e);
  if ((e = cudaFuncSetAttribute(vq_end_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVqEndSmem)) != cudaSuccess ||
      (e = cudaFuncSetAttribute(vq_mid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kVqMidSmem)) != cudaSuccess)
    return bail("cudaFuncSetAttribute(vq kernels)", e);
  cudaMemcpy(tp + o_tw, tw.data(), tw.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(tp + o_dct, dct.data(), dct.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(tp + o_hw, hw.data(), hw.size() * 4, cudaMemcpyHostToDevice);
  if ((e = cudaMemcpy(tp + o_br, br.data(), br.size() * 2, cudaMemcpyHostToDevice)) != cudaSuccess) return bail("cudaMemcpy(tables)", e);
  b->tab = DspTables{(const float *)(tp + o_tw), (const int16_t *)(tp + o_br), (const float *)(tp + o_dct)};
  b->half_window = (const float *)(tp + o_hw);
  // stream state, zero = lpcnet_encoder_init (lpcnet_enc.c:471-475; exc_mem is encoder-only)
  const size_t B = n_streams;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
  const size_t o_p = take(B * kOverlap * 4), o_m = take(B * 4), o_s = take(B * 4), o_e = take(B * kPMax * 4),
               o_pa = take(B * kPMax * 4), o_al = take(B * 4), o_bi = take(B * 4), o_vq = take(B * kBands * 4);
  if ((e = cudaMalloc(&b->st_alloc, off)) != cudaSuccess) return bail("cudaMalloc(state)", e);
  b->st_bytes = off;
  if ((e = cudaMemset(b->st_alloc, 0, off)) != cudaSuccess) return bail("cudaMemset(state)", e);
  char *p = static_cast<char *>(b->st_alloc);
  b->st.p_hist = (float *)(p + o_p); b->st.mem_preemph = (float *)(p + o_m); b->st.s_last = (float *)(p + o_s);
  b->st.e_hist = (float *)(p + o_e); b->st.path = (float *)(p + o_pa); b->st.path_all = (float *)(p + o_al);
  b->st.best_i = (int *)(p + o_bi); b->st.vq_mem = (float *)(p + o_vq);
  {
    // the VQ tables: lpcnet_b200_set_codebooks, else the ceps_codebook* arrays of the default blob file (the stand-in
    // for what the reference compiles in, ceps_codebooks.c; api.cu file_codebooks), else none (lpcnet_encode fails)
    std::lock_guard<std::mutex> lk(g_cb_mu);
    const std::vector<float> from_file = g_codebooks.empty() ? file_codebooks() : std::vector<float>();
    const std::vector<float> &cb = g_codebooks.empty() ? from_file : g_codebooks;
    if (!cb.empty() && upload_codebooks(b, cb.data(), (int)cb.size())) {
      lpcnet_feature_batch_destroy(b);
      return nullptr;
    }
  }
  return b;
}

void lpcnet_feature_batch_destroy(LPCNetFeatureBatch *b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) { cudaStreamSynchronize(b->stream); cudaStreamDestroy(b->stream); }
  cudaFree(b->st_alloc); cudaFree(b->tab_alloc); cudaFree(b->work); cudaFree(b->d_pcm); cudaFree(b->d_feat);
  cudaFree(b->d_cb); cudaFree(b->d_ids); cudaFree(b->d_left0); cudaFree(b->d_packets);
  delete b;
}

int lpcnet_feature_batch_n_streams(const LPCNetFeatureBatch *b) { return b ? b->n_streams : 0; }
long long lpcnet_feature_batch_launch_count(const LPCNetFeatureBatch *b) { return b ? b->launches : 0; }

int lpcnet_feature_batch_reset(LPCNetFeatureBatch *b, int stream) {
  if (!b || stream < -1 || stream >= b->n_streams) return set_last_error("lpcnet_feature_batch_reset: bad arguments");
  CKA(cudaSetDevice(b->device));
  if (stream < 0) {
    CKA(cudaMemsetAsync(b->st_alloc, 0, b->st_bytes, b->stream));
  } else {
    const size_t s = stream;
    CKA(cudaMemsetAsync(b->st.p_hist + s * kOverlap, 0, kOverlap * 4, b->stream));
    CKA(cudaMemsetAsync(b->st.mem_preemph + s, 0, 4, b->stream));
    CKA(cudaMemsetAsync(b->st.s_last + s, 0, 4, b->stream));
    CKA(cudaMemsetAsync(b->st.e_hist + s * kPMax, 0, kPMax * 4, b->stream));
    CKA(cudaMemsetAsync(b->st.path + s * kPMax, 0, kPMax * 4, b->stream));
    CKA(cudaMemsetAsync(b->st.path_all + s, 0, 4, b->stream));
    CKA(cudaMemsetAsync(b->st.best_i + s, 0, 4, b->stream));
    CKA(cudaMemsetAsync(b->st.vq_mem + s * kBands, 0, kBands * 4, b->stream));
  }
  CKA(cudaStreamSynchronize(b->stream));
  return 0;
}

int lpcnet_feature_batch_compute(LPCNetFeatureBatch *b, const short *pcm, int n_frames, float *features) {
  return analysis_host(b, pcm, nullptr, n_frames, features);
}
int lpcnet_feature_batch_compute_float(LPCNetFeatureBatch *b, const float *pcm, int n_frames, float *features) {
  return analysis_host(b, nullptr, pcm, n_frames, features);
}
int lpcnet_feature_batch_compute_device(LPCNetFeatureBatch *b, const short *d_pcm16, const float *d_pcmf, int n_frames,
                                        float *d_features, void *cuda_stream) {
  if (!b || (!d_pcm16 && !d_pcmf) || !d_features || n_frames < 0)
    return set_last_error("lpcnet_feature_batch_compute_device: bad arguments");
  cudaStream_t keep = b->stream;
  if (cuda_stream) b->stream = static_cast<cudaStream_t>(cuda_stream);
  const int r = n_frames ? analysis_device(b, d_pcm16, d_pcmf, n_frames, d_features) : 0;
  b->stream = keep;
  return r;
}

int lpcnet_feature_batch_set_codebooks(LPCNetFeatureBatch *b, const float *codebooks, int n_floats) {
  return upload_codebooks(b, codebooks, n_floats);
}

int lpcnet_feature_batch_encode(LPCNetFeatureBatch *b, const short *pcm, int n_packets, unsigned char *packets,
                                float *quantised_features) {
  if (!packets) return set_last_error("lpcnet_feature_batch_encode: packets must not be NULL");
  return superframe_host(b, pcm, n_packets, packets, quantised_features);
}

int lpcnet_feature_batch_compute_features(LPCNetFeatureBatch *b, const short *pcm, int n_packets, float *features) {
  if (!features) return set_last_error("lpcnet_feature_batch_compute_features: features must not be NULL");
  return superframe_host(b, pcm, n_packets, nullptr, features);
}

}  // extern "C"
std::vector<float> lpcb::default_codebooks() {
  std::lock_guard<std::mutex> lk(g_cb_mu);
  return g_codebooks;
}
extern "C" {

int lpcnet_b200_set_codebooks(const float *codebooks, int n_floats) {
  std::lock_guard<std::mutex> lk(g_cb_mu);
  if (!codebooks) { g_codebooks.clear(); return 0; }
  if (n_floats != 3 * kCbN * kNb1 + 4096 * kBands) return set_last_error("lpcnet_b200_set_codebooks: expected 125952 floats");
  g_codebooks.assign(codebooks, codebooks + n_floats);
  return 0;
}

// ---- the reference's encoder-state API for this path (include/lpcnet.h:98-158), one stream -----------------
struct LPCNetEncState {
  uint32_t magic;
  LPCNetFeatureBatch *batch;
};
static const uint32_t kEncMagic = 0x4C50454Eu;

int lpcnet_encoder_get_size(void) { return (int)sizeof(LPCNetEncState); }

int lpcnet_encoder_init(LPCNetEncState *st) {
  if (!st) return -1;
  st->magic = kEncMagic;
  st->batch = lpcnet_feature_batch_create(1, current_device());   // like lpcnet_load_model: the caller's current GPU
  return st->batch ? 0 : -1;
}

LPCNetEncState *lpcnet_encoder_create(void) {
  LPCNetEncState *st = static_cast<LPCNetEncState *>(calloc(1, sizeof(LPCNetEncState)));
  if (st) lpcnet_encoder_init(st);
  return st;
}

void lpcnet_encoder_destroy(LPCNetEncState *st) {
  if (!st) return;
  if (st->magic == kEncMagic) lpcnet_feature_batch_destroy(st->batch);
  free(st);
}

int lpcnet_compute_single_frame_features(LPCNetEncState *st, const short *pcm, float *features) {
  if (!st || st->magic != kEncMagic || !st->batch) return set_last_error("lpcnet_compute_single_frame_features: state not initialised");
  return lpcnet_feature_batch_compute(st->batch, pcm, 1, features);
}

int lpcnet_compute_single_frame_features_float(LPCNetEncState *st, const float *pcm, float *features) {
  if (!st || st->magic != kEncMagic || !st->batch) return set_last_error("lpcnet_compute_single_frame_features_float: state not initialised");
  return lpcnet_feature_batch_compute_float(st->batch, pcm, 1, features);
}

int lpcnet_encode(LPCNetEncState *st, const short *pcm, unsigned char *buf) {
  if (!st || st->magic != kEncMagic || !st->batch) return set_last_error("lpcnet_encode: state not initialised");
  return lpcnet_feature_batch_encode(st->batch, pcm, 1, buf, nullptr);
}

int lpcnet_compute_features(LPCNetEncState *st, const short *pcm, float *features) {
  if (!st || st->magic != kEncMagic || !st->batch) return set_last_error("lpcnet_compute_features: state not initialised");
  return lpcnet_feature_batch_compute_features(st->batch, pcm, 1, features);
}

}  // extern "C"
