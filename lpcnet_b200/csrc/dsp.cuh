// Frame-rate DSP shared by the frame network and the feature analysis.
// kiss_fft for nfft = 320 = 5*4*4*4 (float build of src/kiss_fft.c), one transform per thread with the scratch
// interleaved over threads, shared by the frame network (cepstrum -> LPC) and the feature analysis.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace lpcb {
namespace {

// ---- kiss_fft, nfft = 320 = 5*4*4*4 (float build), per-thread, scratch interleaved over threads ----
struct Cpx { float r, i; };
#define C_MUL(m, a, b) do { (m).r = (a).r * (b).r - (a).i * (b).i; (m).i = (a).r * (b).i + (a).i * (b).r; } while (0)
#define C_ADD(res, a, b) do { (res).r = (a).r + (b).r; (res).i = (a).i + (b).i; } while (0)
#define C_SUB(res, a, b) do { (res).r = (a).r - (b).r; (res).i = (a).i - (b).i; } while (0)
#define C_ADDTO(res, a) do { (res).r += (a).r; (res).i += (a).i; } while (0)

struct Scratch {
  float *base;   // [320][2][stride]
  size_t stride;
  __device__ __forceinline__ Cpx ld(int k) const { Cpx c; c.r = base[(size_t)(2 * k) * stride]; c.i = base[(size_t)(2 * k + 1) * stride]; return c; }
  __device__ __forceinline__ void st(int k, Cpx c) const { base[(size_t)(2 * k) * stride] = c.r; base[(size_t)(2 * k + 1) * stride] = c.i; }
};

// The transforms run one per thread with the scratch of a block in SHARED memory, interleaved over its threads
// ([320][2][kFftBlock]: conflict-free, 80 KB per block): every butterfly access is a 30-cycle shared-memory access
// instead of an L2 round trip (the scratch used to be a global array: 133 us per single-frame lpc_kernel launch,
// latency-bound whatever the row count).  Launch with kFftBlock threads and kFftSmem bytes; fft_opt_in() first.
constexpr int kFftBlock = 32;
constexpr size_t kFftSmem = (size_t)kFftBlock * 640 * sizeof(float);
__device__ __forceinline__ Scratch block_scratch() {
  extern __shared__ __align__(16) float fft_smem[];
  return Scratch{fft_smem + threadIdx.x, (size_t)kFftBlock};
}
template <class K>
inline cudaError_t fft_opt_in(K kernel) {   // > 48 KB of dynamic shared memory needs the opt-in (per device; cheap to repeat)
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFftSmem);
}

__device__ __forceinline__ Cpx twd(const float *tw, int k) { Cpx c; c.r = __ldg(tw + 2 * k); c.i = __ldg(tw + 2 * k + 1); return c; }

// kf_bfly4 (src/kiss_fft.c:96-162)
__device__ void bfly4(const Scratch &S, const float *tw, int fstride, int m, int N, int mm) {
  if (m == 1) {
    for (int i = 0; i < N; i++) {
      const int o = 4 * i;
      Cpx f0 = S.ld(o), f1 = S.ld(o + 1), f2 = S.ld(o + 2), f3 = S.ld(o + 3), s0, s1;
      C_SUB(s0, f0, f2);
      C_ADDTO(f0, f2);
      C_ADD(s1, f1, f3);
      C_SUB(f2, f0, s1);
      C_ADDTO(f0, s1);
      C_SUB(s1, f1, f3);
      f1.r = s0.r + s1.i; f1.i = s0.i - s1.r;
      f3.r = s0.r - s1.i; f3.i = s0.i + s1.r;
      S.st(o, f0); S.st(o + 1, f1); S.st(o + 2, f2); S.st(o + 3, f3);
    }
  } else {
    const int m2 = 2 * m, m3 = 3 * m;
    for (int i = 0; i < N; i++) {
      const int base = i * mm;
      for (int j = 0; j < m; j++) {
        const int o = base + j;
        Cpx s[6];
        Cpx f0 = S.ld(o), fm = S.ld(o + m), fm2 = S.ld(o + m2), fm3 = S.ld(o + m3);
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
        S.st(o, f0); S.st(o + m, fm); S.st(o + m2, fm2); S.st(o + m3, fm3);
      }
    }
  }
}

// kf_bfly5 (src/kiss_fft.c:229-295)
__device__ void bfly5(const Scratch &S, const float *tw, int fstride, int m, int N, int mm) {
  const Cpx ya = twd(tw, fstride * m), yb = twd(tw, fstride * 2 * m);
  for (int i = 0; i < N; i++) {
    const int base = i * mm;
    for (int u = 0; u < m; ++u) {
      const int o = base + u;
      Cpx s[13];
      Cpx F0 = S.ld(o), F1 = S.ld(o + m), F2 = S.ld(o + 2 * m), F3 = S.ld(o + 3 * m), F4 = S.ld(o + 4 * m);
      s[0] = F0;
      C_MUL(s[1], F1, twd(tw, u * fstride));
      C_MUL(s[2], F2, twd(tw, 2 * u * fstride));
      C_MUL(s[3], F3, twd(tw, 3 * u * fstride));
      C_MUL(s[4], F4, twd(tw, 4 * u * fstride));
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
      S.st(o, F0); S.st(o + m, F1); S.st(o + 2 * m, F2); S.st(o + 3 * m, F3); S.st(o + 4 * m, F4);
    }
  }
}

__constant__ short c_eband5ms[18] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 34, 40};
__constant__ float c_compensation[18] = {0.8f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 0.666667f, 0.5f, 0.5f,
                                         0.5f, 0.333333f, 0.25f, 0.25f, 0.2f, 0.166667f, 0.173913f};

// opus_fft_impl (kiss_fft.c:518-563) on bit-reversed input already scaled by 1/320; factors {5,64, 4,16, 4,4, 4,1}
__device__ __forceinline__ void fft320_butterflies(const Scratch &S, const float *tw) {
  bfly4(S, tw, 80, 1, 80, 1);    // i=3: p=4, m=1,  fstride=5*4*4, N=fstride, mm=m2=4
  bfly4(S, tw, 20, 4, 20, 16);   // i=2: p=4, m=4
  bfly4(S, tw, 5, 16, 5, 64);    // i=1: p=4, m=16
  bfly5(S, tw, 1, 64, 1, 1);     // i=0: p=5, m=64
}

struct DspTables {
  const float *fft_tw;          // [320][2]
  const int16_t *fft_bitrev;    // [320]
  const float *dct;             // [18*18]
};

// lpc_from_cepstrum (src/freq.c:310-320) and everything below it; one thread per (stream, frame).
// err_out (optional): the prediction error lpcn_lpc returns (the return value of the reference's lpc_from_cepstrum)
__device__ void lpc_from_cepstrum_dev(const DspTables &M, const float *cepstrum, float *lpc, const Scratch &S,
                                      float *err_out = nullptr) {
  float tmp[18], Ex[18];
  for (int i = 0; i < 18; i++) tmp[i] = cepstrum[i];
  tmp[0] += 4;
  // idct (freq.c:230-240)
  for (int i = 0; i < 18; i++) {
    float sum = 0;
    for (int j = 0; j < 18; j++) sum += tmp[j] * __ldg(M.dct + i * 18 + j);
    Ex[i] = (float)((double)sum * sqrt(2. / 18));
  }
  for (int i = 0; i < 18; i++) Ex[i] = (float)(pow((double)10.f, (double)Ex[i]) * (double)c_compensation[i]);
  // interp_band_gain (freq.c:202-215) written straight into the bit-reversed FFT input
  // (opus_fft_c, kiss_fft.c:566-585: fout[bitrev[i]] = scale * fin[i]); bins 160 and above band 17 are 0.
  const float scale = 1.f / 320;
  for (int k = 0; k < 320; k++) { Cpx zc; zc.r = 0.f; zc.i = 0.f; S.st(k, zc); }
  for (int i = 0; i < 17; i++) {
    const int band_size = (c_eband5ms[i + 1] - c_eband5ms[i]) * 4;
    for (int j = 0; j < band_size; j++) {
      const float frac = (float)j / (float)band_size;
      const float v = (1.f - frac) * Ex[i] + frac * Ex[i + 1];
      const int k = c_eband5ms[i] * 4 + j;   // 0..159
      Cpx c;
      c.r = scale * v; c.i = scale * 0.f;
      S.st(__ldg(M.fft_bitrev + k), c);
      if (k > 0) {
        // inverse_transform (freq.c:256-273): x[320-k] = conj(x[k])
        Cpx cc;
        cc.r = scale * v; cc.i = scale * (-0.f);
        S.st(__ldg(M.fft_bitrev + 320 - k), cc);
      }
    }
  }
  // opus_fft_impl (kiss_fft.c:518-563), factors {5,64, 4,16, 4,4, 4,1}
  fft320_butterflies(S, M.fft_tw);
  float ac[17];
  ac[0] = 320.f * S.ld(0).r;
  for (int i = 1; i < 17; i++) ac[i] = 320.f * S.ld(320 - i).r;
  // lpc_from_bands (freq.c:290-296): noise floor and lag window in double
  ac[0] = (float)((double)ac[0] + ((double)ac[0] * 1e-4 + 320 / 12 / 38.));
  for (int i = 1; i < 17; i++) ac[i] = (float)((double)ac[i] * (1 - 6e-5 * i * i));
  // lpcn_lpc (freq.c:86-127)
  float error = ac[0];
  for (int i = 0; i < 16; i++) lpc[i] = 0;
  if (ac[0] != 0) {
    for (int i = 0; i < 16; i++) {
      float rr = 0;
      for (int j = 0; j < i; j++) rr += lpc[j] * ac[i - j];
      rr += ac[i + 1];
      const float r = -rr / error;
      lpc[i] = r;
      for (int j = 0; j < (i + 1) >> 1; j++) {
        const float t1 = lpc[j], t2 = lpc[i - 1 - j];
        lpc[j] = t1 + r * t2;
        lpc[i - 1 - j] = t2 + r * t1;
      }
      error = error - (r * r) * error;
      if (error < .001f * ac[0]) break;
    }
  }
  if (err_out) *err_out = error;
}

}  // namespace
}  // namespace lpcb
