// Packet-loss concealment for many streams at once: the reference's non-causal mode
// (lpcnet_plc_update_non_causal / lpcnet_plc_conceal_non_causal / process_queued_update / compute_plc_pred /
// clear_state, src/lpcnet_plc.c:349-491, 342-347, 135-145, 175-181), LPCNET_PLC_NONCAUSAL with or without
// LPCNET_PLC_DC_FILTER, for a model built with FEATURES_DELAY 0 (lpcnet_plc.c:357-361).
//
// Every stream follows its own loss pattern, so in one 10 ms frame the streams of a batch are in different
// branches of the reference's state machine.  The branch a stream takes depends only on the loss flags (lost now,
// lost before, an update queued), which the host knows: it sorts the streams of every frame into classes and
// uploads one byte mask per stage; all streams then walk the same sequence of stages, and a stage moves only
// the streams of its class:
//   FG  first good frame after a loss: prediction, 80 free samples, the backward pass on the reversed frame
//       (clear_state, frame network, 160 forced + 80 free samples), cross-fade, state restore, analysis
//   Q   the update FG queues for the stream's next call: frame network + 160 teacher-forced samples, run at once
//   M   good frame (2 x 80 forced samples) or lost frame (80 forced-or-free + 80 free samples)
// FG and Q move few streams (those leaving a loss burst) through a long chain of sample steps (320 + 160), M moves
// all the others through 160: the FG + Q chain runs on a side CUDA stream of higher priority, beside M on the batch's
// stream, with segment buffers of its own; events order it against the stages that feed it and the frame's end.
// The sample steps are the synthesis engine's persistent kernel (sample_kernel.cuh) with per-stream flags:
// inactive / free-running / teacher-forced.  Frame-rate arithmetic (DC filter in double, Burg cepstrum, the
// int8 predictor GRUs, cross-fade) follows the reference operation by operation; -fmad=false.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <new>
#include <string>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "bitexact.cuh"
#include "dsp.cuh"
#include "dsp_warp.cuh"
#include "engine.h"

namespace lpcb {
namespace {

constexpr int kTOff = 80, kBands = 18;

// stream classes of one frame (host side) and the per-stage byte masks derived from them
enum Cls : uint8_t { kG0 = 0, kFG = 1, kL0 = 2, kL1 = 3 };   // good, first good after loss, first lost, lost again
enum Mask { mUpdM = 0 /* update && !FG */, mUpd, mFG, mMain, mL1, fFgFree, fFgForced, fA, fB, mCls, kNMasks };

struct PlcState {        // per stream, device
  short *pcm;            // [B][160]  st->pcm[0..160)
  float *features;       // [B][20]   st->features
  float *gru1, *gru2;    // [B][256]  plc_net
  double *dc_mem, *syn_dc;   // [B]
  short *dc_buf;         // [B][80]
  short *queued;         // [B][160]  queued_samples
};

struct PlcWork {         // per frame scratch, device
  short *in;             // [B][160]  the frame after DC removal
  short *save;           // [B][160]  pcm_save
  short *lp;             // [B][160]
  double *mem_bak;       // [B]
  int *delta;            // [B]
  float *burg;           // [B][36]
  float *encf;           // [B][36]   analysis of the input frame
  float *encf2;          // [B][36]   analysis of the concealment buffer (state update only)
  float *fnfeat;         // [B][20]   frame-network input
  short *segf;           // [B][160]  teacher-forcing input of a sample segment
  short *sego, *sego2;   // [B][160]  sample segment outputs
  short *out;            // [B][160]  the frame's output
};

struct PlcArgs {
  PlcState st;
  PlcWork w;
  DevState lp;           // the synthesis state (LPCNetState) and its snapshot
  DevState lp_copy;
  DevPlc net;
  const uint32_t *rcp_tab;
  DspTables tab;
  float *fft_scratch;    // [320][2][B]
  const float *fade;     // [80] cross-fade weights
  const uint8_t *masks;  // [kNMasks][B] of this frame
  const uint16_t *lossc; // [B] loss_count before this frame
  short *pcm_io;         // [B][total_frames*160]
  int n_streams, frame, total_frames, remove_dc;
  // side chain: the device list of the frame's first-good-frame streams -- its per-stream kernels run over the list
  // (a grid over the whole batch that mostly returns at the mask costs a wave of blocks per kernel, and the side
  // chain has only the SMs the main class's persistent kernels leave)
  const int *list;
  int n_list;
};
__device__ __forceinline__ int plc_count(const PlcArgs &A) { return A.list ? A.n_list : A.n_streams; }
__device__ __forceinline__ int plc_stream(const PlcArgs &A, int k) { return A.list ? __ldg(A.list + k) : k; }

#define DC_CONST 0.003

// ---- update streams: DC removal (lpcnet_plc.c:361-371), pcm_save ----
// One warp per stream (4 streams per block of 128): the frame moves in coalesced 64-byte rows, the DC filter's
// recurrence (sequential in double precision) runs on lane 0 out of shared memory.
constexpr int kWpsThreads = 128;
__device__ __forceinline__ int wps_stream(int &lane, int &w) {
  lane = threadIdx.x & 31; w = threadIdx.x >> 5;
  return blockIdx.x * (kWpsThreads / 32) + w;
}
inline int wps_grid(int n_streams) { return (n_streams + kWpsThreads / 32 - 1) / (kWpsThreads / 32); }

__global__ void __launch_bounds__(kWpsThreads) plc_pre_kernel(PlcArgs A) {
  __shared__ short xs[kWpsThreads / 32][kFrame], ls[kWpsThreads / 32][kFrame];
  int lane, w;
  const int s = wps_stream(lane, w);
  if (s >= A.n_streams || !A.masks[mUpd * A.n_streams + s]) return;
  const short *x = A.pcm_io + ((size_t)s * A.total_frames + A.frame) * kFrame;
  short *in = A.w.in + (size_t)s * kFrame, *save = A.w.save + (size_t)s * kFrame, *lp = A.w.lp + (size_t)s * kFrame;
  for (int i = lane; i < kFrame; i += 32) xs[w][i] = x[i];
  __syncwarp();
  if (lane == 0) {
    A.w.delta[s] = (int)A.st.syn_dc[s];
    if (A.remove_dc) {
      double dc = A.st.dc_mem[s] + A.st.syn_dc[s];
      A.st.syn_dc[s] = 0;
      A.w.mem_bak[s] = dc;
      for (int i = 0; i < kFrame; i++) {
        ls[w][i] = (short)(int)floor(.5 + dc);
        dc += DC_CONST * ((double)xs[w][i] - dc);
      }
      A.st.dc_mem[s] = dc;
    }
  }
  __syncwarp();
  for (int i = lane; i < kFrame; i += 32) {
    const short l = A.remove_dc ? ls[w][i] : (short)0;
    const short v = (short)(xs[w][i] - l);
    lp[i] = l; in[i] = v; save[i] = v;
  }
}

// ---- burg_cepstral_analysis (freq.c:152-199, burg.c:97-245), one thread per update stream ----
__device__ double inner_flp(const float *a, const float *b, int n) {
  double r = 0.0;
  int i;
  for (i = 0; i < n - 3; i += 4)
    r += (double)a[i + 0] * (double)b[i + 0] + (double)a[i + 1] * (double)b[i + 1] + (double)a[i + 2] * (double)b[i + 2] +
         (double)a[i + 3] * (double)b[i + 3];
  for (; i < n; i++) r += (double)a[i] * (double)b[i];
  return r;
}

// corr (optional): the 1 + D autocorrelation sums inner_flp(x, x + n, len - n), n = 0..D, formed elsewhere (one lag
// per lane in plc_burg_kernel: each is a serial sum in the reference's order, and they do not depend on each other)
__device__ float burg_analysis(float *Aout, const float *x, float minInvGain, int len, int D, const double *corr = nullptr) {
  double C0, invGain, num, nrg_f, nrg_b, rc, Atmp, tmp1, tmp2;
  double Cf[16], Cl[16], CAf[17], CAb[17], Af[16];
  bool reached = false;
  C0 = corr ? corr[0] : inner_flp(x, x, len);
  for (int n = 0; n < 16; n++) Cf[n] = 0;
  for (int n = 1; n < D + 1; n++) Cf[n - 1] += corr ? corr[n] : inner_flp(x, x + n, len - n);
  for (int n = 0; n < 16; n++) Cl[n] = Cf[n];
  CAb[0] = CAf[0] = C0 + (double)1e-5f * C0 + (double)1e-9f;
  invGain = 1.0;
  for (int n = 0; n < D; n++) {
    tmp1 = x[n];
    tmp2 = x[len - n - 1];
    for (int k = 0; k < n; k++) {
      Cf[k] -= (double)(x[n] * x[n - k - 1]);
      Cl[k] -= (double)(x[len - n - 1] * x[len - n + k]);
      Atmp = Af[k];
      tmp1 += (double)x[n - k - 1] * Atmp;
      tmp2 += (double)x[len - n + k] * Atmp;
    }
    for (int k = 0; k <= n; k++) {
      CAf[k] -= tmp1 * (double)x[n - k];
      CAb[k] -= tmp2 * (double)x[len - n + k - 1];
    }
    tmp1 = Cf[n];
    tmp2 = Cl[n];
    for (int k = 0; k < n; k++) {
      Atmp = Af[k];
      tmp1 += Cl[n - k - 1] * Atmp;
      tmp2 += Cf[n - k - 1] * Atmp;
    }
    CAf[n + 1] = tmp1;
    CAb[n + 1] = tmp2;
    num = CAb[n + 1];
    nrg_b = CAb[0];
    nrg_f = CAf[0];
    for (int k = 0; k < n; k++) {
      Atmp = Af[k];
      num += CAb[n - k] * Atmp;
      nrg_b += CAb[k + 1] * Atmp;
      nrg_f += CAf[k + 1] * Atmp;
    }
    rc = -2.0 * num / (nrg_f + nrg_b);
    tmp1 = invGain * (1.0 - rc * rc);
    if (tmp1 <= (double)minInvGain) {
      rc = sqrt(1.0 - (double)minInvGain / invGain);
      if (num > 0) rc = -rc;
      invGain = minInvGain;
      reached = true;
    } else {
      invGain = tmp1;
    }
    for (int k = 0; k < (n + 1) >> 1; k++) {
      tmp1 = Af[k];
      tmp2 = Af[n - k - 1];
      Af[k] = tmp1 + rc * tmp2;
      Af[n - k - 1] = tmp2 + rc * tmp1;
    }
    Af[n] = rc;
    if (reached) {
      for (int k = n + 1; k < D; k++) Af[k] = 0.0;
      break;
    }
    for (int k = 0; k <= n + 1; k++) {
      tmp1 = CAf[k];
      CAf[k] += rc * CAb[n - k + 1];
      CAb[n - k + 1] += rc * tmp1;
    }
  }
  if (reached) {
    for (int k = 0; k < D; k++) Aout[k] = (float)(-Af[k]);
    C0 -= inner_flp(x, x, D);
    nrg_f = C0 * invGain;
  } else {
    nrg_f = CAf[0];
    tmp1 = 1.0;
    for (int k = 0; k < D; k++) {
      Atmp = Af[k];
      nrg_f += CAf[k + 1] * Atmp;
      tmp1 += Atmp * Atmp;
      Aout[k] = (float)(-Atmp);
    }
    nrg_f -= (double)1e-5f * C0 * tmp1;
  }
  return (float)nrg_f;
}

__constant__ double c_pow995[16];   // pow(.995, i+1), filled by the host's libm at start-up

// burg_cepstral_analysis (freq.c:152-186), kBurgWarps half frames per block of as many warps, in three steps:
//  1. warp w, half frame w: pre-emphasis, then the 17 autocorrelation sums of Burg's method one lag per lane (each a
//     serial double-precision sum in the reference's order; they do not depend on each other);
//  2. warp 0, lane l: Burg's recursion for half frame l -- serial and in double precision, so the 32 of the block run
//     side by side on the lanes of one warp (one per warp leaves 31 lanes idle: measured 2.5x slower per frame);
//  3. warp w again: the transform, band sums, logs and DCT as in dsp_warp.cuh.
constexpr int kBurgWarps = 16;   // (32 warps per block cap the recursion at 64 registers: 157 us per launch against 1xx)
constexpr size_t kBurgSmem = (size_t)kBurgWarps * kWarpDspFloats * sizeof(float);
__global__ void __launch_bounds__(kBurgWarps * 32) plc_burg_kernel(PlcArgs A) {
  extern __shared__ __align__(16) float burg_sm[];
  __shared__ int valid_s[kBurgWarps];
  const int lane = threadIdx.x & 31, wi = threadIdx.x >> 5;
  const int idx = blockIdx.x * kBurgWarps + wi;
  const int s = idx >> 1, half = idx & 1;
  const bool valid = s < A.n_streams && A.masks[mUpd * A.n_streams + s];
  const WarpDsp W = warp_dsp(burg_sm, wi, lane);
  // per-warp layout until step 3 overwrites it: bin [79] floats, corr [17] doubles behind it, in the transform's buffer
  auto bin_of = [&](int w) { return burg_sm + (size_t)w * kWarpDspFloats; };
  auto corr_of = [&](int w) { return reinterpret_cast<double *>(burg_sm + (size_t)w * kWarpDspFloats) + 48; };
  auto blpc_of = [&](int w) { return burg_sm + (size_t)w * kWarpDspFloats + 640 + 18; };      // W.f + 18: [16], then g
  if (lane == 0) valid_s[wi] = valid;
  if (valid) {
    const short *pcm = A.w.in + (size_t)s * kFrame + half * kTOff;
    float *bin = bin_of(wi);
    for (int i = lane; i < kTOff - 1; i += 32) bin[i] = (float)pcm[i + 1] - 0.85f * (float)pcm[i];
    __syncwarp();
    if (lane <= 16) corr_of(wi)[lane] = inner_flp(bin, bin + lane, kTOff - 1 - lane);
  }
  __syncthreads();
  if (wi == 0 && lane < kBurgWarps && valid_s[lane]) {
    float a[16];
    float g = burg_analysis(a, bin_of(lane), (float)1e-3, kTOff - 1, 16, corr_of(lane));
    g /= (float)(kTOff - 2 * (16 - 1));
    float *o = blpc_of(lane);
    for (int i = 0; i < 16; i++) o[i] = a[i];
    o[16] = g;
  }
  __syncthreads();
  if (!valid) return;
  const float *blpc = blpc_of(wi);
  const float g = blpc[16];
  float *out = A.w.encf2 + (size_t)s * 36 + half * kBands;
  // forward_transform of [1, -lpc[i]*.995^(i+1), 0...]: opus_fft_c's scaled bit-reversed copy
  const float scale = 1.f / 320;
  float vin[10];
#pragma unroll
  for (int q = 0; q < 10; q++) {        // (read before the buffer is overwritten: bin / corr live in it)
    const int k = lane + 32 * q;
    float v = 0.f;
    if (k == 0) v = 1.f;
    else if (k <= 16) v = (float)((double)(-blpc[k - 1]) * c_pow995[k - 1]);
    vin[q] = v;
  }
  __syncwarp();
#pragma unroll
  for (int q = 0; q < 10; q++) W.x[__ldg(A.tab.fft_bitrev + lane + 32 * q)] = make_float2(scale * vin[q], scale * 0.f);
  __syncwarp();
  wfft320(W, A.tab.fft_tw);
  // compute_band_energy_inverse (freq.c:59-84), band `lane`
  const float sum = w_band_energy<true>(W);
  const double gain = (.45 * (double)g) * (double)(1.f / ((float)320 * 320 * 320));
  float *Ly = W.f + 36;
  if (lane < kBands) {
    const float E = (float)((double)sum * gain);
    Ly[lane] = (float)log10(1e-2 + (double)E);
  }
  __syncwarp();
  float mine = 0.f, logMax = -2, follow = -2;
  for (int i = 0; i < kBands; i++) {
    float l = Ly[i];
    // freq.c:180-183: the constant 2.5 is a double there, so these comparisons run in double
    const double f25 = (double)follow - 2.5;
    const double inner = f25 > (double)l ? f25 : (double)l;
    const float lm8 = logMax - 8;
    l = (float)((double)lm8 > inner ? (double)lm8 : inner);
    if (i == lane) mine = l;
    logMax = fmaxf(logMax, l);
    follow = (float)(f25 > (double)l ? f25 : (double)l);
  }
  __syncwarp();
  if (lane < kBands) Ly[lane] = mine;
  __syncwarp();
  float c = w_dct(A.tab.dct, Ly, lane);
  if (lane == 0) c += -4;
  if (lane < kBands) out[lane] = c;
}

__global__ void plc_burg_combine_kernel(PlcArgs A) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_streams * kBands) return;
  const int s = i / kBands, k = i % kBands;
  if (!A.masks[mUpd * A.n_streams + s]) return;
  const float c0 = A.w.encf2[(size_t)s * 36 + k], c1 = A.w.encf2[(size_t)s * 36 + kBands + k];
  A.w.burg[(size_t)s * 36 + k] = (float)(.5 * (double)(c0 + c1));
  A.w.burg[(size_t)s * 36 + kBands + k] = c0 - c1;
}

// ---- compute_plc_pred (lpcnet_plc.c:135-145): one block of 256 threads per stream ----
// mode per stream from its class: FG pass -> [burg | 0 | 1]; main pass, good -> [burg | features | 1]; main
// pass, lost -> zeros, then the loss attenuation of features[0] (lpcnet_plc.c:457-458).
// u8 x s8 dot product of four byte pairs on an int32 accumulator (vpmaddubsw + vpmaddwd without the s16 saturation,
// which the loader has excluded)
__device__ __forceinline__ int32_t dp4a_u8s8(uint32_t a, uint32_t b, int32_t c) {
  int32_t d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}

// the pairs of `row` that can saturate vpmaddubsw's s16 (engine.h DevSat; none for clipped models): their weights are
// zero in the dp4a words and the saturated pair sums are added here
__device__ __forceinline__ int32_t plc_sat_row(const DevSat &sat, int row, const uint32_t *xq) {
  int32_t add = 0;
  if (sat.n) {
    const int e1 = (int)__ldg(sat.start + row + 1);
    for (int e = (int)__ldg(sat.start + row); e < e1; e++) add += sat16_pair(__ldg(sat.e + e), reinterpret_cast<const unsigned char *>(xq));
  }
  return add;
}

__device__ __forceinline__ float plc_gru_row(const uint32_t *__restrict__ w_in, const uint32_t *xq_in, int nin4,
                                              const uint32_t *__restrict__ w_rec, const uint32_t *xq_st, float bias_in,
                                              float bias_rec, int row, float &recur_out, const DevSat &sat_in,
                                              const DevSat &sat_rec) {
  // compute_gruB (nnet.c:326-372): zrh = subias[row] + 0 -> int8 dot products on int32 accumulators seeded
  // with cvtps2dq(16256*t), back to float by * 1/16256 (vec_avx.h:690-755, 790-860)
  int32_t acc = cvt_rne_ref((bias_in + 0.f) * LPCB_SCALE);
  for (int k = 0; k < nin4; k++) acc = dp4a_u8s8(xq_in[k], __ldg(w_in + (size_t)k * kPlcRows + row), acc);
  acc = (int32_t)((uint32_t)acc + (uint32_t)plc_sat_row(sat_in, row, xq_in));
  int32_t acr = cvt_rne_ref(bias_rec * LPCB_SCALE);
  for (int k = 0; k < kPlcUnits / 4; k++) acr = dp4a_u8s8(xq_st[k], __ldg(w_rec + (size_t)k * kPlcRows + row), acr);
  acr = (int32_t)((uint32_t)acr + (uint32_t)plc_sat_row(sat_rec, row, xq_st));
  recur_out = (float)acr * LPCB_SCALE_1;
  return (float)acc * LPCB_SCALE_1;
}

__device__ void plc_gru(const uint32_t *w_in, const uint32_t *w_rec, const float *subias, const float *x_in, int nin,
                        float *state /*global [256]*/, uint32_t *xq_in, uint32_t *xq_st, float *s_state,
                        const uint32_t *rcp, int t, const DevSat &sat_in, const DevSat &sat_rec) {
  // vector_ps_to_epi8 of the input and of the state
  unsigned char *qi = reinterpret_cast<unsigned char *>(xq_in), *qs = reinterpret_cast<unsigned char *>(xq_st);
  if (t < nin) qi[t] = (unsigned char)quant_u8_ref(x_in[t]);
  const float st = state[t];
  qs[t] = (unsigned char)quant_u8_ref(st);
  __syncthreads();
  float rz, rr, rh;
  float z = plc_gru_row(w_in, xq_in, nin / 4, w_rec, xq_st, subias[t], subias[kPlcRows + t], t, rz, sat_in, sat_rec);
  float r = plc_gru_row(w_in, xq_in, nin / 4, w_rec, xq_st, subias[kPlcUnits + t], subias[kPlcRows + kPlcUnits + t],
                        kPlcUnits + t, rr, sat_in, sat_rec);
  float h = plc_gru_row(w_in, xq_in, nin / 4, w_rec, xq_st, subias[2 * kPlcUnits + t],
                        subias[kPlcRows + 2 * kPlcUnits + t], 2 * kPlcUnits + t, rh, sat_in, sat_rec);
  z = sigmoid_ref(z + rz, rcp);
  r = sigmoid_ref(r + rr, rcp);
  h = tanh_ref(h + rh * r, rcp);
  const float nv = z * st + (1.f - z) * h;
  __syncthreads();
  state[t] = nv;
  s_state[t] = nv;
  __syncthreads();
}

// compute_plc_pred for kPredR streams per block: thread t owns unit t of the GRUs for all of them, so every weight word
// is loaded once per kPredR streams (one stream per block was bound by the L1/TEX pipe: 0.7 MB of weights per stream,
// `profiles/r2_v3_plc_pred_full.txt`: l1tex 82 %, 22 long-scoreboard stalls per issue).  Per stream the arithmetic and
// its order are those of plc_gru / plc_gru_row above.
constexpr int kPredR = 4;
struct PredSmem {
  float in[kPredR][kPlcIn + 3], d1[kPredR][kPlcCond], g1[kPredR][kPlcUnits], g2[kPredR][kPlcUnits];
  uint32_t xq_in[kPredR][kPlcUnits / 4], xq_st[kPredR][kPlcUnits / 4];
};

__device__ void plc_gru_r(const uint32_t *__restrict__ w_in, const uint32_t *__restrict__ w_rec, const float *subias,
                          const float *x_in, int x_stride, int nin, float *const (&state)[kPredR], const bool (&valid)[kPredR],
                          PredSmem &S, float *s_out, const uint32_t *rcp, int t, const DevSat &sat_in, const DevSat &sat_rec) {
  float st[kPredR];
#pragma unroll
  for (int r = 0; r < kPredR; r++) {
    if (t < nin) reinterpret_cast<unsigned char *>(S.xq_in[r])[t] = (unsigned char)quant_u8_ref(x_in[r * x_stride + t]);
    st[r] = valid[r] ? state[r][t] : 0.f;
    reinterpret_cast<unsigned char *>(S.xq_st[r])[t] = (unsigned char)quant_u8_ref(st[r]);
  }
  __syncthreads();
  float gi[3][kPredR], gr[3][kPredR];
#pragma unroll
  for (int kind = 0; kind < 3; kind++) {
    const int row = kind * kPlcUnits + t;
    int32_t acc[kPredR], acr[kPredR];
    const int32_t seed_in = cvt_rne_ref((subias[row] + 0.f) * LPCB_SCALE), seed_rec = cvt_rne_ref(subias[kPlcRows + row] * LPCB_SCALE);
#pragma unroll
    for (int r = 0; r < kPredR; r++) { acc[r] = seed_in; acr[r] = seed_rec; }
    // (the loads do not depend on each other: 16 in flight per thread, or the loops run at L2 latency per word)
#pragma unroll 16
    for (int k = 0; k < nin / 4; k++) {
      const uint32_t w = __ldg(w_in + (size_t)k * kPlcRows + row);
#pragma unroll
      for (int r = 0; r < kPredR; r++) acc[r] = dp4a_u8s8(S.xq_in[r][k], w, acc[r]);
    }
#pragma unroll 16
    for (int k = 0; k < kPlcUnits / 4; k++) {
      const uint32_t w = __ldg(w_rec + (size_t)k * kPlcRows + row);
#pragma unroll
      for (int r = 0; r < kPredR; r++) acr[r] = dp4a_u8s8(S.xq_st[r][k], w, acr[r]);
    }
#pragma unroll
    for (int r = 0; r < kPredR; r++) {
      acc[r] = (int32_t)((uint32_t)acc[r] + (uint32_t)plc_sat_row(sat_in, row, S.xq_in[r]));
      acr[r] = (int32_t)((uint32_t)acr[r] + (uint32_t)plc_sat_row(sat_rec, row, S.xq_st[r]));
      gi[kind][r] = (float)acc[r] * LPCB_SCALE_1;
      gr[kind][r] = (float)acr[r] * LPCB_SCALE_1;
    }
  }
  float nv[kPredR];
#pragma unroll
  for (int r = 0; r < kPredR; r++) {
    const float z = sigmoid_ref(gi[0][r] + gr[0][r], rcp);
    const float rg = sigmoid_ref(gi[1][r] + gr[1][r], rcp);
    const float h = tanh_ref(gi[2][r] + gr[2][r] * rg, rcp);
    nv[r] = z * st[r] + (1.f - z) * h;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kPredR; r++) {
    if (valid[r]) state[r][t] = nv[r];
    s_out[r * kPlcUnits + t] = nv[r];
  }
  __syncthreads();
}

__global__ void __launch_bounds__(256) plc_pred_kernel(PlcArgs A, int fg_pass) {
  __shared__ PredSmem S;
  const int t = threadIdx.x, B = A.n_streams, n = plc_count(A);
  int sid[kPredR], mode[kPredR];   // mode: 0 not this pass, 1 zeros, 2 burg only, 3 burg + features
  bool valid[kPredR];
  float *g1[kPredR], *g2[kPredR];
  bool any = false;
#pragma unroll
  for (int r = 0; r < kPredR; r++) {
    const int e = blockIdx.x * kPredR + r;
    sid[r] = e < n ? plc_stream(A, e) : 0;
    const uint8_t cls = A.masks[mCls * B + sid[r]];
    mode[r] = e >= n ? 0 : fg_pass ? (cls == kFG ? 2 : 0) : (cls == kFG ? 0 : (cls == kG0 ? 3 : 1));
    valid[r] = mode[r] != 0;
    any |= valid[r];
    g1[r] = A.st.gru1 + (size_t)sid[r] * kPlcUnits;
    g2[r] = A.st.gru2 + (size_t)sid[r] * kPlcUnits;
  }
  if (!any) return;
  if (t < kPlcIn) {
#pragma unroll
    for (int r = 0; r < kPredR; r++) {
      float v = 0.f;
      if (mode[r] >= 2 && t < 2 * kBands) v = A.w.burg[(size_t)sid[r] * 2 * kBands + t];
      else if (mode[r] == 3 && t < 2 * kBands + 20) v = A.w.encf[(size_t)sid[r] * 36 + t - 2 * kBands];
      else if (mode[r] >= 2 && t == kPlcIn - 1) v = 1.f;
      S.in[r][t] = v;
    }
  }
  __syncthreads();
  // _lpcnet_compute_dense, 57 -> 128, tanh: sgemv_accum16 = fmaf in input order from the bias
  if (t < kPlcCond) {
    float acc[kPredR];
    const float b = __ldg(A.net.dense1_b + t);
#pragma unroll
    for (int r = 0; r < kPredR; r++) acc[r] = b;
#pragma unroll 19
    for (int j = 0; j < kPlcIn; j++) {
      const float w = __ldg(A.net.dense1_w + j * kPlcCond + t);
#pragma unroll
      for (int r = 0; r < kPredR; r++) acc[r] = fmaf(w, S.in[r][j], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < kPredR; r++) S.d1[r][t] = tanh_ref(acc[r], A.rcp_tab);
  }
  __syncthreads();
  plc_gru_r(A.net.g1_in, A.net.g1_rec, A.net.g1_subias, &S.d1[0][0], kPlcCond, kPlcCond, g1, valid, S, &S.g1[0][0], A.rcp_tab, t,
            A.net.sat_g1_in, A.net.sat_g1_rec);
  plc_gru_r(A.net.g2_in, A.net.g2_rec, A.net.g2_subias, &S.g1[0][0], kPlcUnits, kPlcUnits, g2, valid, S, &S.g2[0][0], A.rcp_tab, t,
            A.net.sat_g2_in, A.net.sat_g2_rec);
  // plc_out, 256 -> 20, linear: 20 rows is not a multiple of 16, so the reference takes sgemv_accum's scalar
  // loop (vec_avx.h:73-86 of nnet.c) -- separate multiply and add
  if (t < 20 * kPredR) {
    const int r = t / 20, o = t % 20;
    if (valid[r]) {
      float acc = __ldg(A.net.out_b + o);
#pragma unroll 16
      for (int j = 0; j < kPlcUnits; j++) acc += __ldg(A.net.out_w + j * 20 + o) * S.g2[r][j];
      if (o == 19) acc = fminf(.5f, acc + .1f);
      if (o == 0 && mode[r] == 1) {
        const int lc = A.lossc[sid[r]];
        const float att[10] = {0, 0, -.2f, -.2f, -.4f, -.4f, -.8f, -.8f, -1.6f, -1.6f};
        float v;
        if (lc >= 10) v = acc + att[9] - (float)(2 * (lc - 9));
        else v = acc + att[lc];
        acc = fmaxf(-10.f, v);
      }
      A.st.features[(size_t)sid[r] * 20 + o] = acc;
    }
  }
}

// ---- small per-stream moves ----
// frame-network input rows: `from_enc` streams take the analysis features, the others st->features
__global__ void plc_gather_feat_kernel(PlcArgs A, int mask_row, int main_pass) {
  const int i0 = blockIdx.x * blockDim.x + threadIdx.x;
  if (i0 >= plc_count(A) * 20) return;
  const int s = plc_stream(A, i0 / 20), k = i0 % 20, i = s * 20 + k;
  if (!A.masks[mask_row * A.n_streams + s]) return;
  const bool enc = main_pass && A.masks[mCls * A.n_streams + s] == kG0;
  A.w.fnfeat[i] = enc ? A.w.encf[(size_t)s * 36 + k] : A.st.features[i];
}

// snapshot / restore of the synthesis state (`copy = st->lpcnet`, `st->lpcnet = copy`), and clear_state
__global__ void plc_state_copy_kernel(DevState dst, DevState src, const uint8_t *mask, int B, const int *list = nullptr) {
  const int s = list ? __ldg(list + blockIdx.x) : blockIdx.x, t = threadIdx.x;
  if (!mask[s]) return;
  for (int i = t; i < 2 * kConvIn; i += blockDim.x) dst.conv1_state[(size_t)s * 2 * kConvIn + i] = src.conv1_state[(size_t)s * 2 * kConvIn + i];
  for (int i = t; i < 2 * kC; i += blockDim.x) dst.conv2_state[(size_t)s * 2 * kC + i] = src.conv2_state[(size_t)s * 2 * kC + i];
  for (int i = t; i < kN; i += blockDim.x) dst.gru_a[(size_t)s * kN + i] = src.gru_a[(size_t)s * kN + i];
  if (t < kNB) dst.gru_b[(size_t)s * kNB + t] = src.gru_b[(size_t)s * kNB + t];
  if (t < 16) {
    dst.last_sig[(size_t)s * 16 + t] = src.last_sig[(size_t)s * 16 + t];
    dst.lpc[(size_t)s * 16 + t] = src.lpc[(size_t)s * 16 + t];
  }
  if (t < 4) dst.rng[(size_t)s * 4 + t] = src.rng[(size_t)s * 4 + t];
  if (t < kMaxDelay * 16 && dst.old_lpc && src.old_lpc)       // the FEATURES_DELAY line (causal modes)
    dst.old_lpc[(size_t)s * kMaxDelay * 16 + t] = src.old_lpc[(size_t)s * kMaxDelay * 16 + t];
  if (t == 0) {
    dst.deemph[s] = src.deemph[s];
    dst.last_exc[s] = src.last_exc[s];
    dst.frame_count[s] = src.frame_count[s];
  }
}

__global__ void plc_clear_state_kernel(DevState st, const uint8_t *mask, int B, const int *list = nullptr) {
  const int s = list ? __ldg(list + blockIdx.x) : blockIdx.x, t = threadIdx.x;
  if (!mask[s]) return;
  for (int i = t; i < kN; i += blockDim.x) st.gru_a[(size_t)s * kN + i] = 0.f;
  if (t < kNB) st.gru_b[(size_t)s * kNB + t] = 0.f;
  if (t < 16) st.last_sig[(size_t)s * 16 + t] = 0.f;
  if (t == 0) { st.deemph[s] = 0.f; st.last_exc[s] = 128; }   // lin2ulaw(0.f)
}

// FG: after the 80 free samples (lpcnet_plc.c:385-402): keep them, redo the DC removal so that it accounts for
// the synthesised 5 ms, and lay out the reversed frame for the backward pass
__global__ void __launch_bounds__(kWpsThreads) plc_fg_mid_kernel(PlcArgs A) {
  __shared__ short xs[kWpsThreads / 32][kFrame], ls[kWpsThreads / 32][kFrame], ys[kWpsThreads / 32][kTOff];
  int lane, w;
  const int k = wps_stream(lane, w);
  if (k >= plc_count(A)) return;
  const int s = plc_stream(A, k);
  if (!A.masks[mFG * A.n_streams + s]) return;
  short *pcm = A.st.pcm + (size_t)s * kFrame, *in = A.w.in + (size_t)s * kFrame, *lp = A.w.lp + (size_t)s * kFrame;
  const short *syn = A.w.sego + (size_t)s * kTOff;   // an 80-sample segment: rows of 80
  for (int i = lane; i < kTOff; i += 32) { const short v = syn[i]; ys[w][i] = v; pcm[kFrame - kTOff + i] = v; }
  for (int i = lane; i < kFrame; i += 32) xs[w][i] = (short)(in[i] + lp[i]);
  __syncwarp();
  if (A.remove_dc) {
    if (lane == 0) {
      double dc = A.w.mem_bak[s], sd = A.st.syn_dc[s];
      for (int i = 0; i < kTOff; i++) sd += DC_CONST * ((double)ys[w][i] - sd);
      dc += sd;
      A.w.delta[s] = (int)((double)A.w.delta[s] + sd);
      A.st.syn_dc[s] = 0;
      for (int i = 0; i < kFrame; i++) {
        ls[w][i] = (short)(int)floor(.5 + dc);
        dc += DC_CONST * ((double)xs[w][i] - dc);
      }
      A.st.dc_mem[s] = dc;
    }
    __syncwarp();
    for (int i = lane; i < kFrame; i += 32) {
      const short l = ls[w][i], v = (short)(xs[w][i] - l);
      lp[i] = l; in[i] = v; A.w.save[(size_t)s * kFrame + i] = v;
      xs[w][i] = v;
    }
  } else {
    for (int i = lane; i < kFrame; i += 32) xs[w][i] = in[i];
  }
  __syncwarp();
  for (int i = lane; i < kFrame; i += 32) A.w.segf[(size_t)s * kFrame + i] = xs[w][kFrame - 1 - i];
}

// FG: cross-fade of the forward and the backward synthesis, and the queued update (lpcnet_plc.c:403-419)
__global__ void plc_fg_end_kernel(PlcArgs A) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= plc_count(A)) return;
  const int s = plc_stream(A, k);
  if (!A.masks[mFG * A.n_streams + s]) return;
  short *pcm = A.st.pcm + (size_t)s * kFrame;
  const short *rev = A.w.sego2 + (size_t)s * kTOff, *in = A.w.in + (size_t)s * kFrame;
  const int delta = A.w.delta[s];
  for (int i = 0; i < kTOff; i++) {
    const float w = A.fade[i];
    const float a = w * (float)pcm[kFrame - 1 - i];
    const float b = (1.f - w) * (float)(rev[i] + delta);
    pcm[kFrame - 1 - i] = (short)(int)floor((.5 + (double)a) + (double)b);
  }
  short *q = A.st.queued + (size_t)s * kFrame;
  for (int i = 0; i < kTOff; i++) q[i] = pcm[kFrame - kTOff + i];
  for (int i = 0; i < kFrame - kTOff; i++) q[kTOff + i] = in[i];
}

// main pass, the teacher-forcing input of the two 80-sample segments
__global__ void plc_seg_prep_kernel(PlcArgs A, int second) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_streams * kTOff) return;
  const int s = i / kTOff, k = i % kTOff;
  const uint8_t cls = A.masks[mCls * A.n_streams + s];
  short v = 0;
  if (!second) { if (cls == kG0 || cls == kL0) v = A.st.pcm[(size_t)s * kFrame + kTOff + k]; }
  else if (cls == kG0) v = A.w.in[(size_t)s * kFrame + k];
  A.w.segf[(size_t)s * kTOff + k] = v;
}

// lost streams after the two segments, before the analysis (lpcnet_plc.c:460-471)
__global__ void plc_lost_mid_kernel(PlcArgs A) {
  const int s = blockIdx.x * blockDim.x + threadIdx.x;
  if (s >= A.n_streams) return;
  const uint8_t cls = A.masks[mCls * A.n_streams + s];
  if (cls != kL0 && cls != kL1) return;
  short *pcm = A.st.pcm + (size_t)s * kFrame, *out = A.w.out + (size_t)s * kFrame;
  const short *a = A.w.sego + (size_t)s * kTOff, *b = A.w.sego2 + (size_t)s * kTOff;
  for (int i = 0; i < kTOff; i++) out[i] = cls == kL0 ? pcm[kTOff + i] : a[i];
  for (int i = 0; i < kTOff; i++) out[kTOff + i] = b[i];
  if (cls == kL1) for (int i = 0; i < kTOff; i++) pcm[kTOff + i] = out[i];
}

// end of the frame for every stream: outputs, buffers, DC (lpcnet_plc.c:436-446, 473-489)
__global__ void __launch_bounds__(kWpsThreads) plc_finish_kernel(PlcArgs A) {
  __shared__ short os[kWpsThreads / 32][kFrame];
  int lane, w;
  const int s = wps_stream(lane, w);
  if (s >= A.n_streams) return;
  const uint8_t cls = A.masks[mCls * A.n_streams + s];
  short *pcm = A.st.pcm + (size_t)s * kFrame, *out = A.w.out + (size_t)s * kFrame;
  short *dcb = A.st.dc_buf + (size_t)s * kTOff;
  short *io = A.pcm_io + ((size_t)s * A.total_frames + A.frame) * kFrame;
  constexpr int NI = kFrame / 32;                     // 5 samples per lane: i = lane + 32 k
  short o[NI];
  if (cls == kG0 || cls == kFG) {
    const short *in = A.w.in + (size_t)s * kFrame, *save = A.w.save + (size_t)s * kFrame, *lp = A.w.lp + (size_t)s * kFrame;
    short nd[NI];
#pragma unroll
    for (int k = 0; k < NI; k++) {
      const int i = lane + 32 * k;
      o[k] = i < kTOff ? pcm[kTOff + i] : in[i - kTOff];
      nd[k] = 0;
      if (A.remove_dc) {
        o[k] = (short)(o[k] + (i < kTOff ? dcb[i] : lp[i - kTOff]));
        if (i < kTOff) nd[k] = lp[kFrame - kTOff + i];
      }
    }
    __syncwarp();                                     // every read of pcm / dc_buf is done
#pragma unroll
    for (int k = 0; k < NI; k++) {
      const int i = lane + 32 * k;
      pcm[i] = save[i];
      if (A.remove_dc && i < kTOff) dcb[i] = nd[k];
    }
  } else {
#pragma unroll
    for (int k = 0; k < NI; k++) { const int i = lane + 32 * k; o[k] = out[i]; os[w][i] = o[k]; }
    __syncwarp();
#pragma unroll
    for (int k = 0; k < NI; k++) { const int i = lane + 32 * k; if (i < kTOff) pcm[i] = os[w][kTOff + i]; }
    if (A.remove_dc) {
      const int dc = (int)floor(.5 + A.st.dc_mem[s]);
      if (lane == 0) {
        double sd = A.st.syn_dc[s];
        for (int i = (cls == kL0 ? kTOff : 0); i < kFrame; i++) sd += DC_CONST * ((double)os[w][i] - sd);
        A.st.syn_dc[s] = sd;
      }
#pragma unroll
      for (int k = 0; k < NI; k++) {
        const int i = lane + 32 * k;
        o[k] = (short)(o[k] + (i < kTOff ? dcb[i] : (short)dc));
      }
      __syncwarp();
#pragma unroll
      for (int k = 0; k < NI; k++) { const int i = lane + 32 * k; if (i < kTOff) dcb[i] = (short)dc; }
    }
  }
#pragma unroll
  for (int k = 0; k < NI; k++) { const int i = lane + 32 * k; out[i] = o[k]; io[i] = o[k]; }
}

#include "plc_causal.cuh"
#include "dred.cuh"

}  // namespace
}  // namespace lpcb

using namespace lpcb;

// ---- host side ------------------------------------------------------------------------------------------------
struct LPCNetPLCBatch {
  int device = 0, n_streams = 0, options = 0, remove_dc = 0;
  LPCNetBatch *syn = nullptr;           // synthesis engine: model, LPCNetState of every stream, sample kernel
  LPCNetFeatureBatch *ana = nullptr;    // LPCNetEncState of every stream
  void *alloc = nullptr;                // PLC state + per-frame scratch + state snapshot
  size_t state_bytes = 0;               // leading part of `alloc` that lpcnet_plc_reset clears
  PlcState st{};
  PlcWork w{};
  DevState snap{};
  float *fft_scratch = nullptr;
  float *fade = nullptr;
  void *tab_alloc = nullptr;
  DspTables tab{};
  std::vector<int> loss_count;          // host mirrors of the control state (functions of the loss pattern only)
  std::vector<uint8_t> queued;
  void *d_io = nullptr; size_t d_io_bytes = 0;
  void *d_ctl = nullptr; size_t d_ctl_bytes = 0;
  long long frames_done = 0;
  // side chain (Q and FG streams): CUDA stream, segment buffers [B][160] x 3, events
  cudaStream_t side = nullptr;
  PlcWork w_side{};
  cudaEvent_t ev_pre = nullptr, ev_side = nullptr, ev_fin = nullptr, ev_q[2] = {nullptr, nullptr};
  bool q_pending = false;
  int q_slot = 0;
  // causal / codec modes (lpcnet_plc.c:185-340): device buffers and the host's copy of the control state
  bool causal = false;
  int enable_blending = 1, delay = 0;
  CausalState cst{};
  void *causal_alloc = nullptr; size_t causal_bytes = 0;
  struct Ctl {
    int blend = 0, skip = 0, pcm_fill = 0, fbfill = 0;
    int fcount = 0;                     // the synthesiser's frame_count (lpcnet.c:119): frames it has conditioned on
    int fec_keep = 0, fec_read = 0, fec_fill = 0, fec_skip = 0;
    std::vector<float> fec;             // [fec_fill][20]
  };
  std::vector<Ctl> ctl;
};

namespace {

#define CKP(call)                                                                                      \
  do {                                                                                                 \
    cudaError_t e_ = (call);                                                                           \
    if (e_ != cudaSuccess) return set_last_error(std::string(#call) + ": " + cudaGetErrorString(e_));  \
  } while (0)

int grow_dev(void **p, size_t *have, size_t want) {
  if (*have >= want) return 0;
  if (*p) cudaFree(*p);
  *p = nullptr; *have = 0;
  CKP(cudaMalloc(p, want));
  *have = want;
  return 0;
}

int plc_alloc(LPCNetPLCBatch *p) {
  const size_t B = p->n_streams;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
  const size_t o_pcm = take(B * kFrame * 2), o_feat = take(B * 20 * 4), o_g1 = take(B * kPlcUnits * 4),
               o_g2 = take(B * kPlcUnits * 4), o_dc = take(B * 8), o_sd = take(B * 8), o_dcb = take(B * kTOff * 2),
               o_q = take(B * kFrame * 2);
  const size_t o_in = take(B * kFrame * 2), o_save = take(B * kFrame * 2), o_lp = take(B * kFrame * 2),
               o_mb = take(B * 8), o_de = take(B * 4), o_burg = take(B * 36 * 4), o_e1 = take(B * 36 * 4),
               o_e2 = take(B * 36 * 4), o_fn = take(B * 20 * 4), o_sf = take(B * kFrame * 2), o_so = take(B * kFrame * 2),
               o_so2 = take(B * kFrame * 2), o_out = take(B * kFrame * 2), o_sf1 = take(B * kFrame * 2),
               o_so1 = take(B * kFrame * 2), o_so21 = take(B * kFrame * 2);
  const size_t o_c1 = take(B * 2 * kConvIn * 4), o_c2 = take(B * 2 * kC * 4), o_ga = take(B * kN * 4), o_gb = take(B * kNB * 4),
               o_ls = take(B * 16 * 4), o_lpc = take(B * 16 * 4), o_rng = take(B * 16), o_dm = take(B * 4),
               o_le = take(B * 4), o_fc = take(B * 4), o_ol = take(B * kMaxDelay * 16 * 4);
  const size_t n_fft = ((2 * B + 63) / 64) * 64;   // one transform per (stream, half frame)
  const size_t o_fft = take(n_fft * 640 * 4), o_fade = take(kTOff * 4);
  CKP(cudaMalloc(&p->alloc, off));
  CKP(cudaMemset(p->alloc, 0, off));
  p->state_bytes = o_fft;
  char *b = static_cast<char *>(p->alloc);
  p->st.pcm = (short *)(b + o_pcm); p->st.features = (float *)(b + o_feat); p->st.gru1 = (float *)(b + o_g1);
  p->st.gru2 = (float *)(b + o_g2); p->st.dc_mem = (double *)(b + o_dc); p->st.syn_dc = (double *)(b + o_sd);
  p->st.dc_buf = (short *)(b + o_dcb); p->st.queued = (short *)(b + o_q);
  p->w.in = (short *)(b + o_in); p->w.save = (short *)(b + o_save); p->w.lp = (short *)(b + o_lp);
  p->w.mem_bak = (double *)(b + o_mb); p->w.delta = (int *)(b + o_de); p->w.burg = (float *)(b + o_burg);
  p->w.encf = (float *)(b + o_e1); p->w.encf2 = (float *)(b + o_e2); p->w.fnfeat = (float *)(b + o_fn);
  p->w.segf = (short *)(b + o_sf); p->w.sego = (short *)(b + o_so); p->w.sego2 = (short *)(b + o_so2);
  p->w.out = (short *)(b + o_out);
  p->w_side = p->w;
  p->w_side.segf = (short *)(b + o_sf1); p->w_side.sego = (short *)(b + o_so1); p->w_side.sego2 = (short *)(b + o_so21);
  int lo = 0, hi = 0;
  CKP(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CKP(cudaStreamCreateWithPriority(&p->side, cudaStreamNonBlocking, hi));
  for (cudaEvent_t *e : {&p->ev_pre, &p->ev_side, &p->ev_fin, &p->ev_q[0], &p->ev_q[1]}) CKP(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
  p->snap.conv1_state = (float *)(b + o_c1); p->snap.conv2_state = (float *)(b + o_c2); p->snap.gru_a = (float *)(b + o_ga);
  p->snap.gru_b = (float *)(b + o_gb); p->snap.last_sig = (float *)(b + o_ls); p->snap.lpc = (float *)(b + o_lpc);
  p->snap.rng = (uint32_t *)(b + o_rng); p->snap.deemph = (float *)(b + o_dm); p->snap.last_exc = (int *)(b + o_le);
  p->snap.frame_count = (int *)(b + o_fc); p->snap.old_lpc = (float *)(b + o_ol);
  p->fft_scratch = (float *)(b + o_fft);
  p->fade = (float *)(b + o_fade);
  // cross-fade weights, w = .5 - .5*cos(M_PI*i/TRAINING_OFFSET) as a float (lpcnet_plc.c:406), by the host's libm
  float fade[kTOff];
  for (int i = 0; i < kTOff; i++) fade[i] = (float)(.5 - .5 * cos(M_PI * i / (kTOff)));
  CKP(cudaMemcpy(p->fade, fade, sizeof(fade), cudaMemcpyHostToDevice));
  double pw[16];
  for (int i = 0; i < 16; i++) pw[i] = pow(.995, i + 1);
  CKP(cudaMemcpyToSymbol(c_pow995, pw, sizeof(pw)));
  // DSP tables (the model's copy lives inside the engine; a private copy keeps this file independent)
  std::vector<float> tw, dct;
  std::vector<int16_t> br;
  host_dsp_tables(tw, br, dct);
  const size_t t_tw = 0, t_dct = tw.size() * 4, t_br = t_dct + dct.size() * 4;
  CKP(cudaMalloc(&p->tab_alloc, t_br + br.size() * 2));
  char *tp = static_cast<char *>(p->tab_alloc);
  CKP(cudaMemcpy(tp + t_tw, tw.data(), tw.size() * 4, cudaMemcpyHostToDevice));
  CKP(cudaMemcpy(tp + t_dct, dct.data(), dct.size() * 4, cudaMemcpyHostToDevice));
  CKP(cudaMemcpy(tp + t_br, br.data(), br.size() * 2, cudaMemcpyHostToDevice));
  p->tab = DspTables{(const float *)(tp + t_tw), (const int16_t *)(tp + t_br), (const float *)(tp + t_dct)};
  return 0;
}

// One call: n_frames frames of every stream, in lock step.  lost: [B][n_frames]; d_io: device [B][n_frames*160].
// frames [f0, f0 + n_frames) of a call of total_frames frames (the control rows of a long call are built in pieces)
int plc_run(LPCNetPLCBatch *p, const unsigned char *lost, int f0, int n_frames, int total_frames) {
  const int B = p->n_streams;
  BatchView V = batch_view(p->syn);
  cudaStream_t cs = V.stream;
  // classes, masks and loss counters of every frame (they depend on the loss pattern only)
  const size_t lc_off = ((size_t)kNMasks * B + 15) & ~size_t(15);           // the uint16 counters sit aligned
  // stream lists of the two sample-segment groups (first good frame, main), for the compact path
  const size_t perm_off = (lc_off + (size_t)B * 2 + 15) & ~size_t(15);
  const size_t per_frame = (perm_off + (size_t)2 * B * 4 + 15) & ~size_t(15);
  std::vector<uint8_t> ctl(per_frame * n_frames);
  struct Any { bool upd, fg, main_, l1, updm; int nfg, nmain; };
  std::vector<Any> any(n_frames);
  for (int f = 0; f < n_frames; f++) {
    uint8_t *m = ctl.data() + per_frame * f;
    uint16_t *lc = reinterpret_cast<uint16_t *>(m + lc_off);
    int *perm = reinterpret_cast<int *>(m + perm_off);
    Any a{false, false, false, false, false, 0, 0};
    for (int s = 0; s < B; s++) {
      const bool is_lost = lost[(size_t)s * total_frames + f0 + f] != 0;
      const int c = p->loss_count[s];
      const uint8_t cls = is_lost ? (c == 0 ? kL0 : kL1) : (c == 0 ? kG0 : kFG);
      m[mCls * B + s] = cls;
      m[mUpd * B + s] = !is_lost;
      m[mFG * B + s] = cls == kFG;
      m[fFgFree * B + s] = cls == kFG ? 1 : 0;
      m[fFgForced * B + s] = cls == kFG ? 3 : 0;
      m[mMain * B + s] = cls != kFG;
      m[mL1 * B + s] = cls == kL1;
      m[mUpdM * B + s] = !is_lost && cls != kFG;
      m[fA * B + s] = cls == kG0 || cls == kL0 ? 3 : (cls == kL1 ? 1 : 0);
      m[fB * B + s] = cls == kG0 ? 3 : (cls == kFG ? 0 : 1);
      lc[s] = (uint16_t)(c > 60000 ? 60000 : c);
      a.upd |= !is_lost; a.fg |= cls == kFG; a.main_ |= cls != kFG; a.l1 |= cls == kL1;
      a.updm |= !is_lost && cls != kFG;
      if (cls == kFG) perm[a.nfg++] = s; else perm[B + a.nmain++] = s;
      p->loss_count[s] = is_lost ? c + 1 : 0;
    }
    any[f] = a;
  }
  if (grow_dev(&p->d_ctl, &p->d_ctl_bytes, ctl.size())) return -1;
  CKP(cudaMemcpyAsync(p->d_ctl, ctl.data(), ctl.size(), cudaMemcpyHostToDevice, cs));

  PlcArgs A{};
  A.st = p->st; A.w = p->w; A.lp = V.st; A.lp_copy = p->snap; A.net = V.model->d.plc; A.rcp_tab = V.model->d.rcp_tab;
  A.tab = p->tab; A.fft_scratch = p->fft_scratch; A.fade = p->fade;
  A.pcm_io = static_cast<short *>(p->d_io); A.n_streams = B; A.total_frames = total_frames; A.remove_dc = p->remove_dc;
  const int tb = 128, gs = (B + tb - 1) / tb;
  // The side chain: Q and FG streams on p->side (or on the batch's stream: LPCNET_B200_PLC_ONE_STREAM, the sequential
  // order of the reference's branches, kept for A/B runs).  A1 = the stage arguments with the side chain's buffers.
  const char *one = getenv("LPCNET_B200_PLC_ONE_STREAM");
  const bool two = !(one && atoi(one));
  cudaStream_t ss = two ? p->side : cs;
  int main_tile = 0;
  if (const char *e = getenv("LPCNET_B200_PLC_MAIN_TILE")) main_tile = atoi(e);
  PlcArgs A1 = A;
  if (two) A1.w = p->w_side;
  // LPCNET_B200_PLC_TIMING=1 (development): CUDA events between the stages of the batch's stream, printed per call
  const bool timing = getenv("LPCNET_B200_PLC_TIMING") != nullptr;
  static const char *kMark[] = {"start", "pre+burg", "upd analysis", "pred+gather+fn", "segment A", "segment B",
                                "lost_mid+L1 analysis", "wait side", "finish"};
  constexpr int kMarks = sizeof(kMark) / sizeof(kMark[0]);
  std::vector<cudaEvent_t> tev;
  auto mark = [&](int) { if (timing) { cudaEvent_t e; cudaEventCreate(&e); cudaEventRecord(e, cs); tev.push_back(e); } };
  if (two) CKP(cudaEventRecord(p->ev_fin, cs));            // the control rows are uploaded, the last call's frames done
  for (int f = 0; f < n_frames; f++) {
    const uint8_t *M = static_cast<const uint8_t *>(p->d_ctl) + per_frame * f;
    A.masks = A1.masks = M;
    A.lossc = A1.lossc = reinterpret_cast<const uint16_t *>(M + lc_off);
    A.frame = A1.frame = f0 + f;
    const Any a = any[f];
    const int *permFG = reinterpret_cast<const int *>(M + perm_off), *permMain = permFG + B;
    const bool side_work = two && a.fg;
    bool side_recorded = false;
    // an update of an earlier frame may still be running (two events, used in turn: this frame's side chain records
    // the other one before the batch's stream gets to wait for this one)
    const bool q_before = p->q_pending;
    cudaEvent_t q_event = p->ev_q[p->q_slot];
    p->q_pending = false;
    mark(0);
    if (side_work) CKP(cudaStreamWaitEvent(ss, p->ev_fin, 0));
    if (a.upd) {
      plc_pre_kernel<<<wps_grid(B), kWpsThreads, 0, cs>>>(A);
      CKP(cudaFuncSetAttribute(plc_burg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBurgSmem));
      plc_burg_kernel<<<(2 * B + kBurgWarps - 1) / kBurgWarps, kBurgWarps * 32, kBurgSmem, cs>>>(A);
      plc_burg_combine_kernel<<<(B * kBands + 255) / 256, 256, 0, cs>>>(A);
    }
    mark(1);
    if (a.fg) {
      if (two) { CKP(cudaEventRecord(p->ev_pre, cs)); CKP(cudaStreamWaitEvent(ss, p->ev_pre, 0)); }
      // every kernel of the chain runs over the list of the frame's first-good-frame streams
      A1.list = permFG; A1.n_list = a.nfg;
      const int nl = a.nfg;
      const uint8_t *mfg = M + (size_t)mFG * B;
      plc_pred_kernel<<<(nl + kPredR - 1) / kPredR, 256, 0, ss>>>(A1, 1);
      plc_state_copy_kernel<<<nl, 128, 0, ss>>>(p->snap, V.st, mfg, B, permFG);
      plc_gather_feat_kernel<<<(nl * 20 + 255) / 256, 256, 0, ss>>>(A1, mFG, 0);
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, mfg, false, permFG, a.nfg, two ? ss : nullptr)) return -1;
      if (batch_sample_segment(p->syn, kTOff, M + (size_t)fFgFree * B, nullptr, A1.w.sego, permFG, a.nfg, two ? ss : nullptr)) return -1;
      plc_fg_mid_kernel<<<wps_grid(nl), kWpsThreads, 0, ss>>>(A1);
      plc_clear_state_kernel<<<nl, 128, 0, ss>>>(V.st, mfg, B, permFG);
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, mfg, false, permFG, a.nfg, two ? ss : nullptr)) return -1;
      if (batch_sample_segment(p->syn, kFrame, M + (size_t)fFgForced * B, A1.w.segf, A1.w.sego, permFG, a.nfg, two ? ss : nullptr)) return -1;
      if (batch_sample_segment(p->syn, kTOff, M + (size_t)fFgFree * B, nullptr, A1.w.sego2, permFG, a.nfg, two ? ss : nullptr)) return -1;
      plc_fg_end_kernel<<<(nl + tb - 1) / tb, tb, 0, ss>>>(A1);
      plc_state_copy_kernel<<<nl, 128, 0, ss>>>(V.st, p->snap, mfg, B, permFG);
      if (feature_batch_frame_masked(p->ana, p->st.pcm, mfg, p->w.encf2, ss, permFG, a.nfg)) return -1;
      // (the analysis of an FG stream's input frame follows that of its concealment buffer: same chain)
      if (two && feature_batch_frame_masked(p->ana, p->w.in, mfg, p->w.encf, ss, permFG, a.nfg)) return -1;
      // Q: process_queued_update (lpcnet_plc.c:342-347) -- the update the FG branch queues for the stream's next call
      // (frame network on st->features + 160 samples forced on queued_samples) touches the synthesis state only, and
      // nothing between here and that call does: it runs now, at the end of the side chain.  The frame's end does not
      // wait for it (ev_side is recorded here): it runs on beside plc_finish and the next frame's first stages, and
      // the batch's stream joins it (ev_q) before the next main-class predictor, the first kernel that touches what
      // the update reads (st->features) or writes (the synthesis state) for these streams.
      if (two) { CKP(cudaEventRecord(p->ev_side, ss)); side_recorded = true; }
      plc_gather_feat_kernel<<<(nl * 20 + 255) / 256, 256, 0, ss>>>(A1, mFG, 0);
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, mfg, false, permFG, a.nfg, two ? ss : nullptr)) return -1;
      if (batch_sample_segment(p->syn, kFrame, M + (size_t)fFgForced * B, p->st.queued, A1.w.sego, permFG, a.nfg, two ? ss : nullptr)) return -1;
      if (two) { p->q_slot ^= 1; CKP(cudaEventRecord(p->ev_q[p->q_slot], ss)); p->q_pending = true; }
    }
    if (side_work && !side_recorded) CKP(cudaEventRecord(p->ev_side, ss));
    if (two) {
      if (a.updm && feature_batch_frame_masked(p->ana, p->w.in, M + (size_t)mUpdM * B, p->w.encf, cs)) return -1;
    } else if (a.upd && feature_batch_frame_masked(p->ana, p->w.in, M + (size_t)mUpd * B, p->w.encf, cs)) return -1;
    mark(2);
    if (a.main_) {
      if (q_before) CKP(cudaStreamWaitEvent(cs, q_event, 0));
      plc_pred_kernel<<<(B + kPredR - 1) / kPredR, 256, 0, cs>>>(A, 0);
      plc_gather_feat_kernel<<<(B * 20 + 255) / 256, 256, 0, cs>>>(A, mMain, 1);
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, M + (size_t)mMain * B, false, permMain, a.nmain)) return -1;
      plc_seg_prep_kernel<<<(B * kTOff + 255) / 256, 256, 0, cs>>>(A, 0);
      mark(3);
      if (batch_sample_segment(p->syn, kTOff, M + (size_t)fA * B, p->w.segf, p->w.sego, permMain, a.nmain, nullptr, main_tile)) return -1;
      plc_seg_prep_kernel<<<(B * kTOff + 255) / 256, 256, 0, cs>>>(A, 1);
      mark(4);
      if (batch_sample_segment(p->syn, kTOff, M + (size_t)fB * B, p->w.segf, p->w.sego2, permMain, a.nmain, nullptr, main_tile)) return -1;
      mark(5);
      plc_lost_mid_kernel<<<gs, tb, 0, cs>>>(A);
      if (a.l1 && feature_batch_frame_masked(p->ana, p->st.pcm, M + (size_t)mL1 * B, p->w.encf2, cs)) return -1;
    } else { mark(3); mark(4); mark(5); }
    mark(6);
    if (q_before && !a.main_) CKP(cudaStreamWaitEvent(cs, q_event, 0));
    if (side_work) CKP(cudaStreamWaitEvent(cs, p->ev_side, 0));
    mark(7);
    plc_finish_kernel<<<wps_grid(B), kWpsThreads, 0, cs>>>(A);
    mark(8);
    if (two) CKP(cudaEventRecord(p->ev_fin, cs));
    CKP(cudaGetLastError());
    p->frames_done++;
  }
  if (p->q_pending) {       // the call returns with every stream's state final
    CKP(cudaStreamWaitEvent(cs, p->ev_q[p->q_slot], 0));
    p->q_pending = false;
  }
  if (timing && !tev.empty()) {
    CKP(cudaStreamSynchronize(cs));
    double sum[kMarks] = {0};
    for (int f = 0; f < n_frames; f++)
      for (int k = 1; k < kMarks; k++) {
        float ms = 0;
        cudaEventElapsedTime(&ms, tev[(size_t)f * kMarks + k - 1], tev[(size_t)f * kMarks + k]);
        sum[k] += ms;
      }
    double tot = 0;
    for (int k = 1; k < kMarks; k++) tot += sum[k];
    fprintf(stderr, "plc_run %d frames x %d streams, ms per frame on the batch's stream (%s): total %.3f |", n_frames, B,
            two ? "side stream on" : "one stream", tot / n_frames);
    for (int k = 1; k < kMarks; k++) fprintf(stderr, " %s %.3f |", kMark[k], sum[k] / n_frames);
    fprintf(stderr, "\n");
    for (cudaEvent_t e : tev) cudaEventDestroy(e);
  }
  return 0;
}


// ---- causal / codec modes -----------------------------------------------------------------------------------------
// control rows of one frame, [row][B] bytes (cU sits at mUpd's index so that the Burg kernels serve both drivers)
enum CRow { cC = 0, cU = 1, cSB, cSnB, cNS, cFl0, cFl1, cFl2, cFl3, cA0, cA1, cA2, pmA0, pmA1, pmA2, pmC, pmSB, pmNB,
            fA0, fA1, fA2, fSBfree, fSBforced, fC, fC2, pushSB0, pushSB1, pushEnc, kCRows };
static_assert(cU == mUpd, "plc_burg_kernel reads the update mask at mUpd");

int plc_alloc_causal(LPCNetPLCBatch *p) {
  const size_t B = p->n_streams;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
  const size_t o_pcm = take(B * kCpcm * 2), o_cp = take(3 * B * 512 * 4), o_fb = take(B * 80 * 4), o_fe = take(4 * B * 20 * 4);
  CKP(cudaMalloc(&p->causal_alloc, off));
  CKP(cudaMemset(p->causal_alloc, 0, off));
  p->causal_bytes = off;
  char *b = static_cast<char *>(p->causal_alloc);
  p->cst.cpcm = (short *)(b + o_pcm); p->cst.pcopy = (float *)(b + o_cp); p->cst.fbuf = (float *)(b + o_fb);
  p->cst.fecst = (float *)(b + o_fe);
  return 0;
}

void plc_reset_causal_ctl(LPCNetPLCBatch *p) {
  p->ctl.assign(p->n_streams, LPCNetPLCBatch::Ctl());
  for (auto &c : p->ctl) c.pcm_fill = p->delay * kFrame + kTOff;      // lpcnet_plc_reset: PLC_BUF_SIZE
}

// One call of the causal / codec modes: the host walks the reference's control flow for every stream and frame
// (no signal data is involved), the device follows the resulting rows.
int plc_run_causal(LPCNetPLCBatch *p, const unsigned char *lost, int f0, int n_frames, int total_frames) {
  const int B = p->n_streams, D = p->delay, BUF = D * kFrame + kTOff;
  BatchView V = batch_view(p->syn);
  cudaStream_t cs = V.stream;
  const size_t u16_off = ((size_t)kCRows * B + 15) & ~size_t(15);          // appendOff[B], lossc[B]
  const size_t perm_off = (u16_off + (size_t)2 * B * 2 + 15) & ~size_t(15); // 5 stream lists: A0 A1 A2 SB C
  const size_t per_frame = (perm_off + (size_t)5 * B * 4 + 15) & ~size_t(15);
  std::vector<uint8_t> ctl(per_frame * n_frames, 0);
  struct Any { int fl[4], a[3], nA[3], nSB, nC, u, sb, snb, ns, nb, pushsb[2], pushenc, fec[4]; };
  std::vector<Any> any(n_frames);
  struct FecUse { int frame, slot, stream; float v[20]; };
  std::vector<FecUse> fec_use;
  auto push_code = [](LPCNetPLCBatch::Ctl &c) -> uint8_t {     // run_frame_network_deferred (lpcnet.c:122-133)
    if (c.fbfill == 4) return (uint8_t)(4 | 8);
    c.fbfill++;
    return (uint8_t)c.fbfill;
  };
  for (int f = 0; f < n_frames; f++) {
    uint8_t *m = ctl.data() + per_frame * f;
    uint16_t *append_off = reinterpret_cast<uint16_t *>(m + u16_off), *lossc = append_off + B;
    int *perm = reinterpret_cast<int *>(m + perm_off);
    Any a{};
    for (int s = 0; s < B; s++) {
      LPCNetPLCBatch::Ctl &c = p->ctl[s];
      int &lc = p->loss_count[s];
      // get_fec_or_pred (lpcnet_plc.c:147-166): mode 4 with the vector staged for `slot`, or mode 1
      auto fec_or_pred = [&](int slot) -> bool {
        if (c.fec_read != c.fec_fill && c.fec_skip == 0) {
          FecUse u; u.frame = f; u.slot = slot; u.stream = s;
          memcpy(u.v, &c.fec[(size_t)c.fec_read * 20], sizeof(u.v));
          fec_use.push_back(u);
          a.fec[slot] = 1;
          c.fec_read++;
          c.fec_keep = std::max(0, std::max(c.fec_keep, c.fec_read - D - 1));
          return true;
        }
        if (c.fec_skip > 0) c.fec_skip--;
        return false;
      };
      // lpcnet_synthesize_tail_impl writes zeros and leaves the state alone until the synthesiser has run more than
      // FEATURES_DELAY frame networks (lpcnet.c:239-243): a loss in the first frames of a stream
      auto ran_network = [&]() { if (c.fcount < 1000) c.fcount++; };
      auto warm = [&]() -> uint8_t { return c.fcount > D ? 1 : 0; };
      if (lost[(size_t)s * total_frames + f0 + f]) {
        m[cC * B + s] = 1;
        for (int k = 0; k < c.fbfill; k++) { m[(cFl0 + k) * B + s] = 1; a.fl[k] = 1; ran_network(); }
        c.fbfill = 0;
        const int n_iter = (c.pcm_fill + kFrame - 1) / kFrame;
        for (int it = 0; it < n_iter; it++) {
          const int j = 3 - n_iter + it;
          m[(cA0 + j) * B + s] = 1;
          m[(pmA0 + j) * B + s] = fec_or_pred(j) ? 4 : 1;
          ran_network();
          m[(fA0 + j) * B + s] = (uint8_t)(3 * warm());
          perm[j * B + a.nA[j]++] = s;
          a.a[j] = 1;
          c.skip++;
        }
        c.pcm_fill = 0;
        m[fC * B + s] = warm();
        const bool used = fec_or_pred(3);
        ran_network();
        m[fC2 * B + s] = warm();
        lc = used ? 0 : lc + 1;
        m[pmC * B + s] = (uint8_t)((used ? 4 : 1) | 8);
        lossc[s] = (uint16_t)(lc > 60000 ? 60000 : lc);
        perm[4 * B + a.nC++] = s;
        c.blend = 1;
      } else {
        m[cU * B + s] = 1; a.u = 1;
        if (c.skip) {
          if (c.blend) {
            m[cSB * B + s] = 1; a.sb = 1;
            if (p->enable_blending) {
              ran_network();               // twice in the reference, the first on a copy that is thrown away
              m[pmSB * B + s] = 2; m[fSBfree * B + s] = warm(); m[fSBforced * B + s] = (uint8_t)(3 * warm());
              for (int i = 0; i < D; i++) { m[(pushSB0 + i) * B + s] = push_code(c); a.pushsb[i] = 1; }
              perm[3 * B + a.nSB++] = s;
            } else {
              c.fec_read -= D;                                    // fec_rewind (lpcnet_plc.c:168-173)
              if (c.fec_read < c.fec_keep) c.fec_read = c.fec_keep;
            }
            c.pcm_fill = kTOff;
          } else {
            m[cSnB * B + s] = 1; a.snb = 1;
            append_off[s] = (uint16_t)c.pcm_fill;
            c.pcm_fill += kFrame;
          }
        }
        if (!c.blend) {
          m[pmNB * B + s] = 3; a.nb = 1;
          if (c.fec_skip) c.fec_skip--;
          else if (c.fec_read < c.fec_fill) c.fec_read++;
          c.fec_keep = std::max(0, std::max(c.fec_keep, c.fec_read - D - 1));
        }
        if (c.skip) {
          if (p->enable_blending) { m[pushEnc * B + s] = push_code(c); a.pushenc = 1; }
          c.skip--;
        } else {
          m[cNS * B + s] = 1; a.ns = 1;
          m[pushEnc * B + s] = push_code(c); a.pushenc = 1;
        }
        lc = 0;
        c.blend = 0;
      }
    }
    any[f] = a;
  }
  if (grow_dev(&p->d_ctl, &p->d_ctl_bytes, ctl.size())) return -1;
  CKP(cudaMemcpyAsync(p->d_ctl, ctl.data(), ctl.size(), cudaMemcpyHostToDevice, cs));

  PlcArgs A{};
  A.st = p->st; A.w = p->w; A.lp = V.st; A.lp_copy = p->snap; A.net = V.model->d.plc; A.rcp_tab = V.model->d.rcp_tab;
  A.tab = p->tab; A.fft_scratch = p->fft_scratch; A.fade = p->fade;
  A.pcm_io = static_cast<short *>(p->d_io); A.n_streams = B; A.total_frames = total_frames; A.remove_dc = p->remove_dc;
  const int g20 = (B * 20 + 255) / 256;
  const CausalState C = p->cst;
  std::vector<float> stage((size_t)4 * B * 20);
  size_t fu = 0;
  for (int f = 0; f < n_frames; f++) {
    const uint8_t *M = static_cast<const uint8_t *>(p->d_ctl) + per_frame * f;
    auto row = [&](int r) { return M + (size_t)r * B; };
    const uint16_t *append_off = reinterpret_cast<const uint16_t *>(M + u16_off);
    const int *perm = reinterpret_cast<const int *>(M + perm_off);
    A.masks = M; A.lossc = append_off + B; A.frame = f0 + f;
    const Any a = any[f];
    if (fu < fec_use.size() && fec_use[fu].frame == f) {        // the FEC vectors this frame consumes
      for (; fu < fec_use.size() && fec_use[fu].frame == f; fu++)
        memcpy(&stage[((size_t)fec_use[fu].slot * B + fec_use[fu].stream) * 20], fec_use[fu].v, 80);
      CKP(cudaMemcpyAsync(C.fecst, stage.data(), stage.size() * 4, cudaMemcpyHostToDevice, cs));
      CKP(cudaStreamSynchronize(cs));                           // `stage` is reused by the next frame
    }
    // ---- concealed frames (lpcnet_plc_conceal_causal) ----
    for (int k = 0; k < 4; k++)                                 // run_frame_network_flush
      if (a.fl[k] && batch_frame_network_masked(p->syn, C.fbuf + k * 20, 80, row(cFl0 + k), true)) return -1;
    for (int j = 0; j < 3; j++) {                               // catch up on the buffered samples
      if (!a.a[j]) continue;
      const int n = j < 2 ? kFrame : kTOff;
      c_push_copy_kernel<<<B, 128, 0, cs>>>(A.st, C, row(cA0 + j), B, D);
      c_pred_kernel<<<B, 256, 0, cs>>>(A, row(pmA0 + j), C.fecst + (size_t)j * B * 20);
      c_feat_kernel<<<g20, 256, 0, cs>>>(A, row(cA0 + j));
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, row(cA0 + j))) return -1;
      c_pcm_load_kernel<<<B, 64, 0, cs>>>(C, row(cA0 + j), p->w.segf, n);
      if (batch_sample_segment(p->syn, n, row(fA0 + j), p->w.segf, p->w.sego, perm + (size_t)j * B, a.nA[j])) return -1;
      c_pcm_shift_kernel<<<B, 128, 0, cs>>>(C, row(cA0 + j), BUF);
    }
    if (a.nC) {
      c_push_copy_kernel<<<B, 128, 0, cs>>>(A.st, C, row(cC), B, D);
      if (batch_sample_segment(p->syn, kTOff, row(fC), nullptr, p->w.sego, perm + (size_t)4 * B, a.nC)) return -1;
      c_pred_kernel<<<B, 256, 0, cs>>>(A, row(pmC), C.fecst + (size_t)3 * B * 20);
      c_feat_kernel<<<g20, 256, 0, cs>>>(A, row(cC));
      if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, row(cC))) return -1;
      if (batch_sample_segment(p->syn, kTOff, row(fC2), nullptr, p->w.sego2, perm + (size_t)4 * B, a.nC)) return -1;
      c_finish_kernel<<<wps_grid(B), kWpsThreads, 0, cs>>>(A, row(cU), row(cC), 0);
    }
    // ---- received frames (lpcnet_plc_update_causal) ----
    if (a.u) {
      c_pre_kernel<<<wps_grid(B), kWpsThreads, 0, cs>>>(A, row(cU));
      CKP(cudaFuncSetAttribute(plc_burg_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBurgSmem));
      plc_burg_kernel<<<(2 * B + kBurgWarps - 1) / kBurgWarps, kBurgWarps * 32, kBurgSmem, cs>>>(A);
      plc_burg_combine_kernel<<<(B * kBands + 255) / 256, 256, 0, cs>>>(A);
    }
    if (a.sb) {
      if (p->enable_blending) {
        c_set_net_kernel<<<B, 128, 0, cs>>>(A.st, C, row(cSB), B, D);
        c_pred_kernel<<<B, 256, 0, cs>>>(A, row(pmSB), nullptr);
        for (int i = 0; i < D; i++)
          if (a.pushsb[i]) c_fb_push_kernel<<<g20, 256, 0, cs>>>(C, row(pushSB0 + i), A.st.features, 20, B);
        plc_state_copy_kernel<<<B, 128, 0, cs>>>(p->snap, V.st, row(cSB), B);
        c_feat_kernel<<<g20, 256, 0, cs>>>(A, row(cSB));
        if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, row(cSB))) return -1;
        if (batch_sample_segment(p->syn, kTOff, row(fSBfree), nullptr, p->w.sego, perm + (size_t)3 * B, a.nSB)) return -1;
        c_crossfade_kernel<<<(B * kTOff + 255) / 256, 256, 0, cs>>>(A, row(cSB));
        plc_state_copy_kernel<<<B, 128, 0, cs>>>(V.st, p->snap, row(cSB), B);
        if (batch_frame_network_masked(p->syn, p->w.fnfeat, 20, row(cSB))) return -1;
        if (batch_sample_segment(p->syn, kTOff, row(fSBforced), p->w.segf, p->w.sego, perm + (size_t)3 * B, a.nSB)) return -1;
      } else {
        if (D > 0) c_set_net_kernel<<<B, 128, 0, cs>>>(A.st, C, row(cSB), B, D - 1);
        plc_clear_state_kernel<<<B, 128, 0, cs>>>(V.st, row(cSB), B);      // lpcnet_reset_signal (lpcnet.c:226-233)
      }
      c_pcm_store_kernel<<<B, 64, 0, cs>>>(C, row(cSB), nullptr, 0, p->w.in, kFrame, kFrame - kTOff, kTOff);
    }
    if (a.snb) c_pcm_store_kernel<<<B, 64, 0, cs>>>(C, row(cSnB), append_off, 0, p->w.in, kFrame, 0, kFrame);
    // the analysis of every stream's frame (w.in: the received frame, or the concealment just synthesised)
    if (feature_batch_frame_masked(p->ana, p->w.in, nullptr, p->w.encf, cs)) return -1;
    if (a.nb) c_pred_kernel<<<B, 256, 0, cs>>>(A, row(pmNB), nullptr);
    if (a.pushenc) c_fb_push_kernel<<<g20, 256, 0, cs>>>(C, row(pushEnc), p->w.encf, 36, B);
    if (a.ns) {
      c_pcm_store_kernel<<<B, 64, 0, cs>>>(C, row(cNS), nullptr, BUF, p->w.in, kFrame, 0, kFrame);
      c_pcm_shift_kernel<<<B, 128, 0, cs>>>(C, row(cNS), BUF);
    }
    if (a.u) c_finish_kernel<<<wps_grid(B), kWpsThreads, 0, cs>>>(A, row(cU), row(cC), 1);
    CKP(cudaGetLastError());
    p->frames_done++;
  }
  return 0;
}

}  // namespace

extern "C" {

LPCNetPLCBatch *lpcnet_plc_batch_create(const unsigned char *blob, int len, int n_streams, int device, int options,
                                        float lpc_gamma) {
  if ((options & 3) == 3 || (options & ~7)) {
    set_last_error("lpcnet_plc_batch_create: options must be LPCNET_PLC_CAUSAL (0), _NONCAUSAL (1) or _CODEC (2), optionally | LPCNET_PLC_DC_FILTER (4)");
    return nullptr;
  }
  LPCNetPLCBatch *p = new (std::nothrow) LPCNetPLCBatch();
  if (!p) return nullptr;
  p->n_streams = n_streams; p->options = options; p->remove_dc = !!(options & 4);
  p->causal = (options & 3) != 1; p->enable_blending = (options & 3) != 2;
  // the non-causal mode needs a no-lookahead model: FEATURES_DELAY 0 (lpcnet_plc.c:357-361); the causal and codec
  // modes run on the model's own delay (lpcnet_b200_set_model_constants, default 2)
  p->delay = p->causal ? default_features_delay() : 0;
  p->syn = lpcnet_batch_create(blob, len, n_streams, device, p->delay, lpc_gamma);
  if (!p->syn) { delete p; return nullptr; }
  BatchView V = batch_view(p->syn);
  p->device = V.device;
  if (!V.model->has_plc) {
    set_last_error("lpcnet_plc_batch_create: the blob has no plc_* arrays (lpcnet_plc_load_model would fail too)");
    lpcnet_plc_batch_destroy(p);
    return nullptr;
  }
  p->ana = lpcnet_feature_batch_create(n_streams, p->device);
  if (!p->ana || plc_alloc(p) || (p->causal && plc_alloc_causal(p))) { lpcnet_plc_batch_destroy(p); return nullptr; }
  if (p->causal) plc_reset_causal_ctl(p);
  p->loss_count.assign(n_streams, 0);
  p->queued.assign(n_streams, 0);
  return p;
}

void lpcnet_plc_batch_destroy(LPCNetPLCBatch *p) {
  if (!p) return;
  if (p->syn) lpcnet_batch_sync(p->syn);
  if (p->side) { cudaStreamSynchronize(p->side); cudaStreamDestroy(p->side); }
  for (cudaEvent_t e : {p->ev_pre, p->ev_side, p->ev_fin, p->ev_q[0], p->ev_q[1]}) if (e) cudaEventDestroy(e);
  cudaFree(p->alloc); cudaFree(p->tab_alloc); cudaFree(p->d_io); cudaFree(p->d_ctl); cudaFree(p->causal_alloc);
  if (p->ana) lpcnet_feature_batch_destroy(p->ana);
  if (p->syn) lpcnet_batch_destroy(p->syn);
  delete p;
}

int lpcnet_plc_batch_n_streams(const LPCNetPLCBatch *p) { return p ? p->n_streams : 0; }

int lpcnet_plc_batch_set_seed(LPCNetPLCBatch *p, int stream, const unsigned char *data, int n) {
  if (!p) return set_last_error("null PLC batch");
  return lpcnet_batch_set_seed(p->syn, stream, data, n);
}

long long lpcnet_plc_batch_launch_count(const LPCNetPLCBatch *p) {
  return p ? lpcnet_batch_launch_count(p->syn) + lpcnet_feature_batch_launch_count(p->ana) + 12 * p->frames_done : 0;
}

// lpcnet_plc_reset (lpcnet_plc.c:46-60) for every stream: lpcnet_reset, lpcnet_encoder_init, PLC state cleared
int lpcnet_plc_batch_reset(LPCNetPLCBatch *p) {
  if (!p) return set_last_error("lpcnet_plc_batch_reset: null batch");
  if (lpcnet_batch_reset(p->syn, -1) || lpcnet_feature_batch_reset(p->ana, -1)) return -1;
  CKP(cudaSetDevice(p->device));
  CKP(cudaMemset(p->alloc, 0, p->state_bytes));
  p->loss_count.assign(p->n_streams, 0);
  p->queued.assign(p->n_streams, 0);
  if (p->causal) {
    CKP(cudaMemset(p->causal_alloc, 0, p->causal_bytes));
    plc_reset_causal_ctl(p);
  }
  return 0;
}

// lpcnet_plc_fec_add / lpcnet_plc_fec_clear (lpcnet_plc.c:111-132) for one stream of the batch; features NULL means
// "skip one".  The queue lives on the host (its positions are control state); the non-causal mode never reads it.
int lpcnet_plc_batch_fec_add(LPCNetPLCBatch *p, int stream, const float *features) {
  if (!p || stream < 0 || stream >= p->n_streams) return set_last_error("lpcnet_plc_batch_fec_add: bad arguments");
  if (!p->causal) return 0;
  LPCNetPLCBatch::Ctl &c = p->ctl[stream];
  if (!features) { c.fec_skip++; return 0; }
  const int kMaxFec = 100;                      // PLC_MAX_FEC
  if (c.fec_fill == kMaxFec) {
    if (c.fec_keep == 0) return 0;              // "FEC buffer full": the reference drops the vector too
    c.fec.erase(c.fec.begin(), c.fec.begin() + (size_t)c.fec_keep * 20);
    c.fec_fill -= c.fec_keep;
    c.fec_read -= c.fec_keep;
    c.fec_keep = 0;
  }
  c.fec.resize((size_t)(c.fec_fill + 1) * 20);
  memcpy(&c.fec[(size_t)c.fec_fill * 20], features, 80);
  c.fec_fill++;
  return 0;
}

int lpcnet_plc_batch_fec_clear(LPCNetPLCBatch *p, int stream) {
  if (!p || stream < 0 || stream >= p->n_streams) return set_last_error("lpcnet_plc_batch_fec_clear: bad arguments");
  if (!p->causal) return 0;
  LPCNetPLCBatch::Ctl &c = p->ctl[stream];
  c.fec_keep = c.fec_read = c.fec_fill = c.fec_skip = 0;
  c.fec.clear();
  return 0;
}

int lpcnet_plc_batch_process(LPCNetPLCBatch *p, const unsigned char *lost, short *pcm, int n_frames) {
  if (!p || !lost || !pcm || n_frames < 0) return set_last_error("lpcnet_plc_batch_process: bad arguments");
  if (n_frames == 0) return 0;
  CKP(cudaSetDevice(p->device));
  const size_t bytes = (size_t)p->n_streams * n_frames * kFrame * 2;
  if (grow_dev(&p->d_io, &p->d_io_bytes, bytes)) return -1;
  cudaStream_t cs = batch_view(p->syn).stream;
  CKP(cudaMemcpyAsync(p->d_io, pcm, bytes, cudaMemcpyHostToDevice, cs));
  // the per-frame control rows cost ~50 bytes per stream: build them for at most 32 MiB worth of frames at a time
  int step = std::max(1, (int)std::min<long long>(n_frames, (32ll << 20) / (64ll * p->n_streams)));
  if (const char *e = getenv("LPCNET_B200_PLC_CTL_FRAMES")) step = std::max(1, atoi(e));      // tests: force small pieces
  for (int f0 = 0; f0 < n_frames; f0 += step) {
    const int nf = std::min(step, n_frames - f0);
    if (p->causal ? plc_run_causal(p, lost, f0, nf, n_frames) : plc_run(p, lost, f0, nf, n_frames)) return -1;
  }
  CKP(cudaMemcpyAsync(pcm, p->d_io, bytes, cudaMemcpyDeviceToHost, cs));
  CKP(cudaStreamSynchronize(cs));
  return 0;
}

// observable frame-rate state of one stream (tests): st->features[0..20), the predictor's GRU states, dc_mem, syn_dc
int lpcnet_plc_batch_peek(LPCNetPLCBatch *p, int stream, float *features20, float *gru1, float *gru2, double *dc2) {
  if (!p || stream < 0 || stream >= p->n_streams) return set_last_error("lpcnet_plc_batch_peek: bad arguments");
  CKP(cudaSetDevice(p->device));
  CKP(cudaStreamSynchronize(batch_view(p->syn).stream));
  const size_t s = stream;
  if (features20) CKP(cudaMemcpy(features20, p->st.features + s * 20, 20 * 4, cudaMemcpyDeviceToHost));
  if (gru1) CKP(cudaMemcpy(gru1, p->st.gru1 + s * kPlcUnits, kPlcUnits * 4, cudaMemcpyDeviceToHost));
  if (gru2) CKP(cudaMemcpy(gru2, p->st.gru2 + s * kPlcUnits, kPlcUnits * 4, cudaMemcpyDeviceToHost));
  if (dc2) {
    CKP(cudaMemcpy(dc2, p->st.dc_mem + s, 8, cudaMemcpyDeviceToHost));
    CKP(cudaMemcpy(dc2 + 1, p->st.syn_dc + s, 8, cudaMemcpyDeviceToHost));
  }
  return 0;
}

// the last frame's Burg cepstrum and analysis features of one stream (tests)
int lpcnet_plc_batch_peek_frame(LPCNetPLCBatch *p, int stream, float *burg36, float *enc36) {
  if (!p || stream < 0 || stream >= p->n_streams) return set_last_error("lpcnet_plc_batch_peek_frame: bad arguments");
  CKP(cudaSetDevice(p->device));
  CKP(cudaStreamSynchronize(batch_view(p->syn).stream));
  const size_t s = stream;
  if (burg36) CKP(cudaMemcpy(burg36, p->w.burg + s * 36, 36 * 4, cudaMemcpyDeviceToHost));
  if (enc36) CKP(cudaMemcpy(enc36, p->w.encf + s * 36, 36 * 4, cudaMemcpyDeviceToHost));
  return 0;
}

// ---- the reference's single-stream PLC API (include/lpcnet.h:190-215), non-causal mode ----
struct LPCNetPLCState {
  uint32_t magic;
  int options;
  LPCNetPLCBatch *batch;
};
static const uint32_t kPlcMagic = 0x4C50504Cu;

int lpcnet_plc_get_size(void) { return (int)sizeof(LPCNetPLCState); }

int lpcnet_plc_init(LPCNetPLCState *st, int options) {
  if (!st) return -1;
  st->magic = kPlcMagic; st->options = options; st->batch = nullptr;
  if ((options & 3) == 3) {                    // lpcnet_plc.c:64-76
    set_last_error("lpcnet_plc_init: options must be LPCNET_PLC_CAUSAL, _NONCAUSAL or _CODEC, optionally | LPCNET_PLC_DC_FILTER");
    return -1;
  }
  return 0;
}

LPCNetPLCState *lpcnet_plc_create(int options) {
  LPCNetPLCState *st = static_cast<LPCNetPLCState *>(calloc(1, sizeof(LPCNetPLCState)));
  if (st) lpcnet_plc_init(st, options);
  return st;
}

void lpcnet_plc_destroy(LPCNetPLCState *st) {
  if (!st) return;
  if (st->magic == kPlcMagic && st->batch) lpcnet_plc_batch_destroy(st->batch);
  free(st);
}

int lpcnet_plc_load_model(LPCNetPLCState *st, const unsigned char *data, int len) {
  if (!st || st->magic != kPlcMagic) return set_last_error("lpcnet_plc_load_model: state not initialised");
  if (st->batch) { lpcnet_plc_batch_destroy(st->batch); st->batch = nullptr; }
  st->batch = lpcnet_plc_batch_create(data, len, 1, current_device(), st->options, default_lpc_gamma());   // as lpcnet_load_model
  return st->batch ? 0 : -1;
}

void lpcnet_plc_reset(LPCNetPLCState *st) {
  if (st && st->magic == kPlcMagic && st->batch) lpcnet_plc_batch_reset(st->batch);
}

// The FEC queue (lpcnet_plc.c:111-132) feeds get_fec_or_pred, which only the causal entry points call
// (lpcnet_plc.c:316, 325); the non-causal mode never reads it, so -- as in the reference -- features added in that
// mode have no effect on the output.
void lpcnet_plc_fec_add(LPCNetPLCState *st, const float *features) {
  if (st && st->magic == kPlcMagic && st->batch) lpcnet_plc_batch_fec_add(st->batch, 0, features);
}
void lpcnet_plc_fec_clear(LPCNetPLCState *st) {
  if (st && st->magic == kPlcMagic && st->batch) lpcnet_plc_batch_fec_clear(st->batch, 0);
}

int lpcnet_plc_update(LPCNetPLCState *st, short *pcm) {
  if (!st || st->magic != kPlcMagic || !st->batch) return set_last_error("lpcnet_plc_update: no model loaded");
  const unsigned char lost = 0;
  return lpcnet_plc_batch_process(st->batch, &lost, pcm, 1);
}

int lpcnet_plc_conceal(LPCNetPLCState *st, short *pcm) {
  if (!st || st->magic != kPlcMagic || !st->batch) return set_last_error("lpcnet_plc_conceal: no model loaded");
  const unsigned char lost = 1;
  return lpcnet_plc_batch_process(st->batch, &lost, pcm, 1);
}

// ---- DRED: the RDOVAE decoder (src/dred_rdovae.c:37-51, src/dred_rdovae_dec.c) -------------------------------------
struct LPCNetDredDecoder {
  LPCNetBatch *host = nullptr;     // owns the device model (the blob carries every model, as write_lpcnet_weights.c writes it)
  void *d_buf = nullptr; size_t d_bytes = 0;
  long long launches = 0;
};

LPCNetDredDecoder *lpcnet_dred_decoder_create(const unsigned char *blob, int len, int device) {
  LPCNetDredDecoder *d = new (std::nothrow) LPCNetDredDecoder();
  if (!d) return nullptr;
  d->host = lpcnet_batch_create(blob, len, 1, device, default_features_delay(), 1.0f);
  if (!d->host) { delete d; return nullptr; }
  if (!batch_view(d->host).model->has_dred) {
    set_last_error("lpcnet_dred_decoder_create: the blob has no RDOVAE decoder arrays (dec_dense*, dec_final, state1..3)");
    lpcnet_dred_decoder_destroy(d);
    return nullptr;
  }
  return d;
}

void lpcnet_dred_decoder_destroy(LPCNetDredDecoder *d) {
  if (!d) return;
  if (d->host) {
    lpcnet_batch_sync(d->host);
    cudaSetDevice(batch_view(d->host).device);
    cudaFree(d->d_buf);
    lpcnet_batch_destroy(d->host);
  }
  delete d;
}

long long lpcnet_dred_decoder_launch_count(const LPCNetDredDecoder *d) { return d ? d->launches : 0; }

int lpcnet_dred_decode_batch(LPCNetDredDecoder *d, int n_streams, float *features, const float *states, const float *latents,
                             int nb_latents) {
  if (!d || n_streams <= 0 || nb_latents < 0 || !features || !states || !latents)
    return set_last_error("lpcnet_dred_decode_batch: bad arguments");
  if (nb_latents == 0) return 0;
  BatchView V = batch_view(d->host);
  CKP(cudaSetDevice(V.device));
  const size_t n_st = (size_t)n_streams * kDredState, n_z = (size_t)n_streams * nb_latents * kDredLatent,
               n_f = (size_t)n_streams * nb_latents * kDredOut;
  if (grow_dev(&d->d_buf, &d->d_bytes, (n_st + n_z + n_f) * 4)) return -1;
  float *d_st = static_cast<float *>(d->d_buf), *d_z = d_st + n_st, *d_f = d_z + n_z;
  CKP(cudaMemcpyAsync(d_st, states, n_st * 4, cudaMemcpyHostToDevice, V.stream));
  CKP(cudaMemcpyAsync(d_z, latents, n_z * 4, cudaMemcpyHostToDevice, V.stream));
  DredArgs A{};
  A.net = V.model->d.dred; A.rcp_tab = V.model->d.rcp_tab; A.states = d_st; A.latents = d_z; A.features = d_f;
  A.n_latents = nb_latents;
  dred_decode_kernel<<<n_streams, 256, 0, V.stream>>>(A);
  CKP(cudaGetLastError());
  d->launches++;
  CKP(cudaMemcpyAsync(features, d_f, n_f * 4, cudaMemcpyDeviceToHost, V.stream));
  CKP(cudaStreamSynchronize(V.stream));
  return 0;
}

// DRED_rdovae_decode_all (include/dred_rdovae.h:40): one stream
int lpcnet_dred_decode_all(LPCNetDredDecoder *d, float *features, const float *state, const float *latents, int nb_latents) {
  return lpcnet_dred_decode_batch(d, 1, features, state, latents, nb_latents);
}

}  // extern "C"
