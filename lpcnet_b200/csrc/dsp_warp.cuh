// Frame-rate DSP, one WARP per transform (round 2).  dsp.cuh runs every 320-point transform and the band arithmetic
// around it on one thread, the way the reference's scalar code does; a single-frame launch then costs the serial
// latency of ~10 k dependent instructions (76-130 us) whatever the batch.  Here the 32 lanes of a warp share one
// transform: the butterflies of a stage are independent, so lanes take different butterflies and run the SAME
// arithmetic on them (same operations, same order per output: bit-identical to the per-thread form and to kiss_fft),
// the per-band sums go one band per lane (each sum still walks its bins in the reference's order), and only the
// Levinson recursion stays serial (every lane runs it redundantly on the same 17 values).
//   W.x  [320] complex, this warp's, in shared memory          W.f  [64] floats of per-warp scratch
#pragma once
#include "dsp.cuh"

namespace lpcb {
namespace {

constexpr int kWarpDspFloats = 640 + 64;
struct WarpDsp {
  float2 *x;
  float *f;
  int lane;
};
__device__ __forceinline__ WarpDsp warp_dsp(float *block_smem, int warp_in_block, int lane) {
  float *p = block_smem + (size_t)warp_in_block * kWarpDspFloats;
  return WarpDsp{reinterpret_cast<float2 *>(p), p + 640, lane};
}
__device__ __forceinline__ Cpx wld(const float2 *x, int k) { const float2 v = x[k]; Cpx c; c.r = v.x; c.i = v.y; return c; }
__device__ __forceinline__ void wst(float2 *x, int k, Cpx c) { x[k] = make_float2(c.r, c.i); }

// band of the bins 4q .. 4q+3 (c_eband5ms = {0,1,2,3,4,5,6,7,8,10,12,14,16,20,24,28,34,40})
__constant__ unsigned char c_band_of_q[40] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 12, 12,
                                              13, 13, 13, 13, 14, 14, 14, 14, 15, 15, 15, 15, 15, 15, 16, 16, 16, 16, 16, 16};

// opus_fft_impl (kiss_fft.c:518-563) on bit-reversed, scaled input: stages {4,1} {4,4} {4,16} {5,64}; within a stage
// every butterfly owns its elements, so the lanes' order does not matter
__device__ void wfft320(const WarpDsp &W, const float *tw) {
  float2 *x = W.x;
  for (int b = W.lane; b < 80; b += 32) {          // kf_bfly4, m == 1 (kiss_fft.c:103-122)
    const int o = 4 * b;
    Cpx f0 = wld(x, o), f1 = wld(x, o + 1), f2 = wld(x, o + 2), f3 = wld(x, o + 3), s0, s1;
    C_SUB(s0, f0, f2);
    C_ADDTO(f0, f2);
    C_ADD(s1, f1, f3);
    C_SUB(f2, f0, s1);
    C_ADDTO(f0, s1);
    C_SUB(s1, f1, f3);
    f1.r = s0.r + s1.i; f1.i = s0.i - s1.r;
    f3.r = s0.r - s1.i; f3.i = s0.i + s1.r;
    wst(x, o, f0); wst(x, o + 1, f1); wst(x, o + 2, f2); wst(x, o + 3, f3);
  }
  __syncwarp();
#pragma unroll
  for (int stage = 0; stage < 2; stage++) {        // kf_bfly4, m = 4 (fstride 20, N 20, mm 16) and m = 16 (5, 5, 64)
    const int m = stage ? 16 : 4, fstride = stage ? 5 : 20, mm = 4 * m;
    for (int b = W.lane; b < 80; b += 32) {
      const int i = b / m, j = b % m, o = i * mm + j;
      Cpx s[6];
      Cpx f0 = wld(x, o), fm = wld(x, o + m), fm2 = wld(x, o + 2 * m), fm3 = wld(x, o + 3 * m);
      const Cpx t1 = twd(tw, j * fstride), t2 = twd(tw, j * fstride * 2), t3 = twd(tw, j * fstride * 3);
      C_MUL(s[0], fm, t1);
      C_MUL(s[1], fm2, t2);
      C_MUL(s[2], fm3, t3);
      C_SUB(s[5], f0, s[1]);
      C_ADDTO(f0, s[1]);
      C_ADD(s[3], s[0], s[2]);
      C_SUB(s[4], s[0], s[2]);
      C_SUB(fm2, f0, s[3]);
      C_ADDTO(f0, s[3]);
      fm.r = s[5].r + s[4].i; fm.i = s[5].i - s[4].r;
      fm3.r = s[5].r - s[4].i; fm3.i = s[5].i + s[4].r;
      wst(x, o, f0); wst(x, o + m, fm); wst(x, o + 2 * m, fm2); wst(x, o + 3 * m, fm3);
    }
    __syncwarp();
  }
  {                                                // kf_bfly5, m = 64, fstride 1 (kiss_fft.c:229-295)
    constexpr int m = 64;
    const Cpx ya = twd(tw, m), yb = twd(tw, 2 * m);
    for (int u = W.lane; u < m; u += 32) {
      Cpx s[13];
      Cpx F0 = wld(x, u), F1 = wld(x, u + m), F2 = wld(x, u + 2 * m), F3 = wld(x, u + 3 * m), F4 = wld(x, u + 4 * m);
      s[0] = F0;
      C_MUL(s[1], F1, twd(tw, u));
      C_MUL(s[2], F2, twd(tw, 2 * u));
      C_MUL(s[3], F3, twd(tw, 3 * u));
      C_MUL(s[4], F4, twd(tw, 4 * u));
      C_ADD(s[7], s[1], s[4]);
      C_SUB(s[10], s[1], s[4]);
      C_ADD(s[8], s[2], s[3]);
      C_SUB(s[9], s[2], s[3]);
      F0.r = F0.r + (s[7].r + s[8].r);
      F0.i = F0.i + (s[7].i + s[8].i);
      s[5].r = s[0].r + (s[7].r * ya.r + s[8].r * yb.r);
      s[5].i = s[0].i + (s[7].i * ya.r + s[8].i * yb.r);
      s[6].r = s[10].i * ya.i + s[9].i * yb.i;
      s[6].i = -(s[10].r * ya.i + s[9].r * yb.i);
      C_SUB(F1, s[5], s[6]);
      C_ADD(F4, s[5], s[6]);
      s[11].r = s[0].r + (s[7].r * yb.r + s[8].r * ya.r);
      s[11].i = s[0].i + (s[7].i * yb.r + s[8].i * ya.r);
      s[12].r = s[9].i * ya.i - s[10].i * yb.i;
      s[12].i = s[10].r * yb.i - s[9].r * ya.i;
      C_ADD(F2, s[11], s[12]);
      C_SUB(F3, s[11], s[12]);
      wst(x, u, F0); wst(x, u + m, F1); wst(x, u + 2 * m, F2); wst(x, u + 3 * m, F3); wst(x, u + 4 * m, F4);
    }
    __syncwarp();
  }
}

// lpcn_compute_band_energy (freq.c:128-150) / compute_band_energy_inverse (freq.c:59-84) of the spectrum in W.x:
// lane b < 18 returns band b's sum.  The reference adds band i's bins into sum[i] and sum[i+1] as it walks them; each
// sum therefore receives band b-1's `frac` terms (bins ascending) and then band b's `1 - frac` terms (bins ascending).
template <bool INVERSE>
__device__ float w_band_energy(const WarpDsp &W) {
  const int b = W.lane;
  float sum = 0.f;
  if (b < 18) {
#pragma unroll 1
    for (int part = 0; part < 2; part++) {
      const int i = part ? b : b - 1;               // the band whose bins are walked
      if (i < 0 || i > 16) continue;
      const int band_size = (c_eband5ms[i + 1] - c_eband5ms[i]) * 4;
      for (int j = 0; j < band_size; j++) {
        const float frac = (float)j / (float)band_size;
        const Cpx X = wld(W.x, c_eband5ms[i] * 4 + j);
        float tmp = X.r * X.r;
        tmp += X.i * X.i;
        if (INVERSE) tmp = (float)((double)1.f / ((double)tmp + 1e-9));
        sum += part ? (1.f - frac) * tmp : frac * tmp;
      }
    }
    if (b == 0 || b == 17) sum *= 2;
  }
  return sum;
}

// dct (freq.c:218-228) of the 18 values in `in` (shared memory): lane i < 18 returns output i
__device__ float w_dct(const float *dct, const float *in, int lane) {
  float acc = 0;
  if (lane < 18)
    for (int j = 0; j < 18; j++) acc += in[j] * __ldg(dct + j * 18 + lane);
  return (float)((double)acc * sqrt(2. / 18));
}

// lpc_from_cepstrum (src/freq.c:310-320) and everything below it, by a warp.  cepstrum: 18 values every lane can read
// (not in W.f[0..53)).  Every lane ends up with the 16 coefficients; err_out as in lpc_from_cepstrum_dev.
__device__ void w_lpc_from_cepstrum(const DspTables &M, const WarpDsp &W, const float *cepstrum, float (&lpc)[16],
                                    float *err_out = nullptr) {
  const int lane = W.lane;
  float *Ex = W.f, *acs = W.f + 36;
  // idct (freq.c:230-240), one band per lane
  if (lane < 18) {
    float sum = 0;
    for (int j = 0; j < 18; j++) {
      float t = cepstrum[j];
      if (j == 0) t += 4;
      sum += t * __ldg(M.dct + lane * 18 + j);
    }
    const float e = (float)((double)sum * sqrt(2. / 18));
    Ex[lane] = (float)(pow((double)10.f, (double)e) * (double)c_compensation[lane]);
  }
  for (int k = lane; k < 320; k += 32) W.x[k] = make_float2(0.f, 0.f);
  __syncwarp();
  // interp_band_gain (freq.c:202-215) into the bit-reversed FFT input (kiss_fft.c:566-585), one bin per lane and pass;
  // inverse_transform (freq.c:256-273): x[320-k] = conj(x[k])
  const float scale = 1.f / 320;
  for (int k = lane; k < 160; k += 32) {
    const int i = c_band_of_q[k >> 2];
    const int band_size = (c_eband5ms[i + 1] - c_eband5ms[i]) * 4, j = k - c_eband5ms[i] * 4;
    const float frac = (float)j / (float)band_size;
    const float v = (1.f - frac) * Ex[i] + frac * Ex[i + 1];
    W.x[__ldg(M.fft_bitrev + k)] = make_float2(scale * v, scale * 0.f);
    if (k > 0) W.x[__ldg(M.fft_bitrev + 320 - k)] = make_float2(scale * v, scale * (-0.f));
  }
  __syncwarp();
  wfft320(W, M.fft_tw);
  // lpc_from_bands (freq.c:290-296): noise floor and lag window in double
  if (lane < 17) {
    float a = 320.f * W.x[lane ? 320 - lane : 0].x;
    if (lane == 0) a = (float)((double)a + ((double)a * 1e-4 + 320 / 12 / 38.));
    else a = (float)((double)a * (1 - 6e-5 * lane * lane));
    acs[lane] = a;
  }
  __syncwarp();
  // lpcn_lpc (freq.c:86-127), every lane on the same values
  float ac[17];
#pragma unroll
  for (int i = 0; i < 17; i++) ac[i] = acs[i];
  float error = ac[0];
#pragma unroll
  for (int i = 0; i < 16; i++) lpc[i] = 0;
  if (ac[0] != 0) {
#pragma unroll
    for (int i = 0; i < 16; i++) {
      float rr = 0;
#pragma unroll
      for (int j = 0; j < i; j++) rr += lpc[j] * ac[i - j];
      rr += ac[i + 1];
      const float r = -rr / error;
      lpc[i] = r;
#pragma unroll
      for (int j = 0; j < (i + 1) >> 1; j++) {
        const float t1 = lpc[j], t2 = lpc[i - 1 - j];
        lpc[j] = t1 + r * t2;
        lpc[i - 1 - j] = t2 + r * t1;
      }
      error = error - (r * r) * error;
      if (error < .001f * ac[0]) break;
    }
  }
  if (err_out && lane == 0) *err_out = error;
  __syncwarp();
}

}  // namespace
}  // namespace lpcb
