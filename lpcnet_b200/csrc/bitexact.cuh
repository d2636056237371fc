// Device arithmetic that must reproduce the reference's AVX2 DOT_PROD build bit for bit.
// This translation unit is compiled with -fmad=false: every `*` and `+` below is one IEEE
// binary32 operation, and a fused multiply-add appears only where the reference itself uses
// _mm256_fmadd_ps (written fmaf()).
#pragma once
#include <stdint.h>

namespace lpcb {

// `rcpps` (src/vec_avx.h:408,437): exactly a table over the top 11 mantissa bits; the
// exponent just negates.  tab[i] = bits of rcpps(1 + i/2048).
__device__ __forceinline__ float rcp_estimate(float x, const uint32_t *__restrict__ tab) {
  uint32_t b = __float_as_uint(x);
  uint32_t e = (b >> 23) & 0xffu;
  return __uint_as_float(tab[(b >> 12) & 0x7ffu] - ((e - 127u) << 23));
}

// The default table without the lookup: entry i is 1/(mid-point of bucket i) rounded half-up to 12
// mantissa bits, so MUFU.RCP of the mid-point (exponent kept) plus an integer rounding gives the same
// bits whenever the estimate lands on the right side of the rounding boundary.  Three buckets put the
// binary32 reciprocal exactly ON the boundary (the exact quotient 2^25/k is 0.13 ulp below it for bucket
// 1984, 0.13 and 0.59 ulp above it for buckets 2047 and 550), and they need different tie-breaks: down for
// 1984, up for the other two.  Bit 1 of the bucket number (mantissa bit 13) is 0 for 1984 and 1 for 2047
// and 550, so the rounding bias 0x3ff + that bit resolves all three; for every other bucket the two biases
// round alike.  The next-closest buckets are 1.05 ulp from their boundary, inside MUFU.RCP's error on
// paper -- so model.cu compares this function with the table for every bucket and exponent on the GPU at
// load, and the kernel variants built on it are used only when all of them agree.
struct RcpArith {};
__device__ __forceinline__ float rcp_estimate(float x, RcpArith) {
  const uint32_t b = __float_as_uint(x);
  const float mid = __uint_as_float((b & 0xfffff000u) | 0x800u);
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(mid));
  return __uint_as_float((__float_as_uint(r) + 0x3ffu + ((b >> 13) & 1u)) & 0xfffff800u);
}

// tanh8_approx (src/vec_avx.h:393-411)
template <class R>
__device__ __forceinline__ float tanh_ref(float x, R tab) {
  const float N0 = 952.52801514f, N1 = 96.39235687f, N2 = 0.60863042f;
  const float D0 = 952.72399902f, D1 = 413.36801147f, D2 = 11.88600922f;
  float X2 = x * x;
  float num = fmaf(fmaf(N2, X2, N1), X2, N0);
  float den = fmaf(fmaf(D2, X2, D1), X2, D0);
  num = num * x;
  den = rcp_estimate(den, tab);
  num = num * den;
  // _mm256_max_ps(-1, _mm256_min_ps(1, num)); num is never NaN here, so FMNMX is equivalent
  return fmaxf(-1.f, fminf(1.f, num));
}

// sigmoid8_approx (src/vec_avx.h:421-440)
template <class R>
__device__ __forceinline__ float sigmoid_ref(float x, R tab) {
  const float N0 = 238.13200378f, N1 = 6.02452230f, N2 = 0.00950985f;
  const float D0 = 952.72399902f, D1 = 103.34200287f, D2 = 0.74287558f;
  float X2 = x * x;
  float num = fmaf(fmaf(N2, X2, N1), X2, N0);
  float den = fmaf(fmaf(D2, X2, D1), X2, D0);
  num = num * x;
  den = rcp_estimate(den, tab);
  num = fmaf(num, den, 0.5f);
  // fma(.,.,0.5) is never -0, so fmaxf(0, .) equals _mm256_max_ps(0, .)
  return fmaxf(0.f, fminf(1.f, num));
}

// log2_approx + lin2ulaw (src/common.h:18-33, 47-58)
__device__ __forceinline__ float log2_approx_ref(float x) {
  int32_t i = (int32_t)__float_as_uint(x);
  int integer = (i >> 23) - 127;
  i -= integer << 23;
  float frac = __uint_as_float((uint32_t)i) - 1.5f;
  frac = -0.41445418f + frac * (0.95909232f + frac * (-0.33951290f + frac * 0.16541097f));
  return (float)(1 + integer) + frac;
}

// (int)floor(.5 + (double)u) for |u| <= 32768 without double precision: u - floorf(u) is exact, so
// the comparison implements round-half-up exactly (verified exhaustively over every binary32 in
// [-32768, 32768] against the double formula; u + 0.5f alone fails for u = 0.49999997).
__device__ __forceinline__ int round_half_up_ref(float u) {
  const float r = floorf(u);
  return (int)(r + ((u - r >= 0.5f) ? 1.f : 0.f));
}

__device__ __forceinline__ int lin2ulaw_ref(float x) {
  const float scale = 255.f / 32768.f;
  float s = x >= 0 ? 1.f : -1.f;
  x = fabsf(x);
  float u = s * (128.f * (0.69315f * log2_approx_ref(1.f + scale * x)) / 5.5451774445f);
  u = 128.f + u;
  if (u < 0) u = 0;
  if (u > 255) u = 255;
  return round_half_up_ref(u);
}

// kiss99_rand (src/kiss99.c:59-81); st = {z, w, jsr, jcong}
__device__ __forceinline__ uint32_t kiss99_next(uint32_t &z, uint32_t &w, uint32_t &jsr, uint32_t &jcong) {
  z = 36969u * (z & 0xFFFFu) + (z >> 16);
  w = 18000u * (w & 0xFFFFu) + (w >> 16);
  uint32_t mwc = (z << 16) + w;
  jsr ^= jsr << 13;
  jsr ^= jsr >> 17;
  jsr ^= jsr << 5;
  jcong = 69069u * jcong + 1234567u;
  return (mwc ^ jcong) + jsr;
}

// cvtps2dq (round to nearest even; out of range -> 0x80000000), as used to seed the int32
// accumulators (src/vec_avx.h:703-705, 804-806).
__device__ __forceinline__ int32_t cvt_rne_ref(float x) {
  if (!(x >= -2147483648.f && x < 2147483648.f)) return (int32_t)0x80000000;
  return __float2int_rn(x);
}

// vector_ps_to_epi8 (src/vec_avx.h:321-336): u8 = packus(packus(cvtps2dq(fma(x,127,127))))
__device__ __forceinline__ uint32_t quant_u8_ref(float x) {
  int32_t v = cvt_rne_ref(fmaf(x, 127.f, 127.f));
  int32_t t = v < 0 ? 0 : (v > 65535 ? 65535 : v);   // packus_epi32
  int32_t s = (int32_t)(int16_t)(uint16_t)t;        // packus_epi16 reads it as signed 16
  return (uint32_t)(s < 0 ? 0 : (s > 255 ? 255 : s));
}

// Hot-loop variants.  cvt.rni saturates where cvtps2dq returns 0x80000000; 0x7fffffff is not a
// binary32 value, so it can only be the saturated result and is mapped back.  (A NaN input would
// give 0 instead of 0x80000000; the gate arithmetic cannot produce NaN from finite weights.)
__device__ __forceinline__ int32_t cvt_rne_fast(float x) {
  const int32_t r = __float2int_rn(x);
  return r == 0x7fffffff ? (int32_t)0x80000000 : r;
}
// For |x| < 257 the packus chain of vector_ps_to_epi8 is a plain clamp to [0,255]; GRU states are
// convex combinations of values in [-1,1].
__device__ __forceinline__ uint32_t quant_u8_fast(float x) {
  return (uint32_t)min(max(__float2int_rn(fmaf(x, 127.f, 127.f)), 0), 255);
}

// The same for a GRU-A state produced by the blend z*a + (1-z)*h with |a|, |h| <= 1 (so |x| <= 1 + a few
// ulp and fma(x,127,127) lies in (-1e-4, 254.0001)): adding 1.5*2^23 rounds to the nearest-even integer
// exactly like the conversion does, and the integer is the low byte.  No conversion instruction, no clamp.
// (lpcnet_batch_set_state refuses GRU-A states outside [-1, 1], which zero-initialised streams never reach.)
__device__ __forceinline__ uint32_t quant_u8_unit(float x) {
  return __float_as_uint(fmaf(x, 127.f, 127.f) + 12582912.f) & 0xffu;
}

#define LPCB_SCALE (128.f * 127.f)
#define LPCB_SCALE_1 (1.f / 128.f / 127.f)

// The loader proves |16256*t| < 2^31 for every reachable pre-activation (model.cu), so the seed of the
// int32 accumulators (src/vec_avx.h:804-806) is a plain round-to-nearest-even conversion.
__device__ __forceinline__ int32_t acc_seed(float v) { return __float2int_rn(v * LPCB_SCALE); }
__device__ __forceinline__ float acc_out(int32_t acc) { return (float)acc * LPCB_SCALE_1; }

// The same without conversion instructions (F2I / I2F share the 16-lane pipe of MUFU; these run on the FP32 / integer
// pipes), for models whose int32 sums stay inside (-2^22, 2^22) -- model.cu proves it per model (DevModel::small_acc).
// v + 1.5*2^23 has ulp 1, so the add itself rounds v to the nearest-even integer k and the bits are 0x4B400000 + k; the
// integer sums ride on those bits, and subtracting 1.5*2^23 from the result read as a float returns float(sum), exact.
#define LPCB_MAGIC 12582912.f
__device__ __forceinline__ int32_t acc_seed_small(float v) {   // rne(v*16256) + 0x4B400000
  return (int32_t)__float_as_uint(__fadd_rn(v * LPCB_SCALE, LPCB_MAGIC));
}
__device__ __forceinline__ float acc_out_small(int32_t biased) {   // float(sum) / 16256 for sum = biased - 0x4B400000
  return __fadd_rn(__uint_as_float((uint32_t)biased), -LPCB_MAGIC) * LPCB_SCALE_1;
}

// D += A(u8, 16x32 row) * B(s8, 32x8 col), int32 accumulate.  SASS: IMMA.16832.U8.S8
__device__ __forceinline__ void mma_u8s8(int32_t (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
  asm(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// ---- packed binary32 pairs (FFMA2 / FADD2 / FMUL2 on sm_100): each half is the same IEEE operation ----
// (operands travel as binary32 registers and are paired inside the asm: no type punning in C++)
__device__ __forceinline__ float2 f2_fma(float2 a, float2 b, float2 c) {
  float2 d;
  asm("{ .reg .b64 ra, rb, rc, rd;\n\t"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5}; mov.b64 rc, {%6, %7};\n\t"
      "fma.rn.f32x2 rd, ra, rb, rc;\n\t"
      "mov.b64 {%0, %1}, rd; }"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y), "f"(c.x), "f"(c.y));
  return d;
}
__device__ __forceinline__ float2 f2_add(float2 a, float2 b) {
  float2 d;
  asm("{ .reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5};\n\t"
      "add.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd; }"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 f2_mul(float2 a, float2 b) {
  float2 d;
  asm("{ .reg .b64 ra, rb, rd;\n\t"
      "mov.b64 ra, {%2, %3}; mov.b64 rb, {%4, %5};\n\t"
      "mul.rn.f32x2 rd, ra, rb;\n\t"
      "mov.b64 {%0, %1}, rd; }"
      : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
  return d;
}
__device__ __forceinline__ float2 f2_set(float v) { return make_float2(v, v); }

// tanh8_approx / sigmoid8_approx on a pair
template <class R>
__device__ __forceinline__ float2 tanh2_ref(float2 x, R tab) {
  const float2 X2 = f2_mul(x, x);
  float2 num = f2_fma(f2_fma(f2_set(0.60863042f), X2, f2_set(96.39235687f)), X2, f2_set(952.52801514f));
  float2 den = f2_fma(f2_fma(f2_set(11.88600922f), X2, f2_set(413.36801147f)), X2, f2_set(952.72399902f));
  num = f2_mul(num, x);
  den.x = rcp_estimate(den.x, tab); den.y = rcp_estimate(den.y, tab);
  num = f2_mul(num, den);
  num.x = fmaxf(-1.f, fminf(1.f, num.x)); num.y = fmaxf(-1.f, fminf(1.f, num.y));
  return num;
}
template <class R>
__device__ __forceinline__ float2 sigmoid2_ref(float2 x, R tab) {
  const float2 X2 = f2_mul(x, x);
  float2 num = f2_fma(f2_fma(f2_set(0.00950985f), X2, f2_set(6.02452230f)), X2, f2_set(238.13200378f));
  float2 den = f2_fma(f2_fma(f2_set(0.74287558f), X2, f2_set(103.34200287f)), X2, f2_set(952.72399902f));
  num = f2_mul(num, x);
  den.x = rcp_estimate(den.x, tab); den.y = rcp_estimate(den.y, tab);
  num = f2_fma(num, den, f2_set(0.5f));
  num.x = fmaxf(0.f, fminf(1.f, num.x)); num.y = fmaxf(0.f, fminf(1.f, num.y));
  return num;
}

}  // namespace lpcb
