/* Oracle primitives: the reference's vector kernels restated as scalar C.
 * TEST INFRASTRUCTURE ONLY (see lpcnet_oracle.h).  Compile with -ffp-contract=off. */
#include <math.h>
#include <string.h>
#include <stdlib.h>
#include "lpcnet_oracle.h"

static uint32_t f2u(float f) { uint32_t u; memcpy(&u, &f, 4); return u; }
static float u2f(uint32_t u) { float f; memcpy(&f, &u, 4); return f; }

/* ---- rcpps model ----------------------------------------------------------------
 * _mm256_rcp_ps (vec_avx.h:408,437) is a hardware estimate.  Measured on the fixture
 * CPU: the result depends only on the top 11 mantissa bits of the input and the
 * exponent simply negates, i.e. it is a 2048-entry table; on Intel parts the entry for
 * bucket i equals round(2^25 / (4097 + 2 i)) / 8192  (= 1/midpoint rounded to 12 bits). */
static uint32_t g_rcp_tab[2048];
static int g_rcp_init = 0;

void orc_default_rcp_table(uint32_t *tab) {
  int i;
  for (i = 0; i < 2048; i++) {
    uint32_t k = 4097u + 2u * (uint32_t)i;
    uint32_t q = (uint32_t)(((1ull << 26) / k + 1) >> 1); /* round(2^25/k), no ties (k odd) */
    tab[i] = 0x3f000000u + ((q - 4096u) << 11);            /* q/8192 in [0.5,1) */
  }
}

void orc_set_rcp_table(const uint32_t *tab) {
  if (tab) memcpy(g_rcp_tab, tab, sizeof(g_rcp_tab));
  else orc_default_rcp_table(g_rcp_tab);
  g_rcp_init = 1;
}

float orc_rcp(float x) {
  uint32_t b, e, r;
  if (!g_rcp_init) orc_set_rcp_table(NULL);
  b = f2u(x);
  e = (b >> 23) & 0xff;             /* callers only pass positive normal numbers */
  r = g_rcp_tab[(b >> 12) & 0x7ff] - ((e - 127u) << 23);
  return u2f(r);
}

/* ---- activations: tanh8_approx / sigmoid8_approx (vec_avx.h:393-440) ------------- */
float orc_tanh_approx(float x) {
  const float N0 = 952.52801514f, N1 = 96.39235687f, N2 = 0.60863042f;
  const float D0 = 952.72399902f, D1 = 413.36801147f, D2 = 11.88600922f;
  float X2, num, den;
  X2 = x * x;
  num = fmaf(fmaf(N2, X2, N1), X2, N0);
  den = fmaf(fmaf(D2, X2, D1), X2, D0);
  num = num * x;
  den = orc_rcp(den);
  num = num * den;
  /* _mm256_max_ps(min_out, _mm256_min_ps(max_out, num)) */
  num = (1.f < num) ? 1.f : num;
  num = (-1.f > num) ? -1.f : num;
  return num;
}

float orc_sigmoid_approx(float x) {
  const float N0 = 238.13200378f, N1 = 6.02452230f, N2 = 0.00950985f;
  const float D0 = 952.72399902f, D1 = 103.34200287f, D2 = 0.74287558f;
  float X2, num, den;
  X2 = x * x;
  num = fmaf(fmaf(N2, X2, N1), X2, N0);
  den = fmaf(fmaf(D2, X2, D1), X2, D0);
  num = num * x;
  den = orc_rcp(den);
  num = fmaf(num, den, 0.5f);
  num = (1.f < num) ? 1.f : num;
  num = (0.f > num) ? 0.f : num;
  return num;
}

/* ---- mu-law (common.h:18-58) ---------------------------------------------------- */
static float log2_approx(float x) {
  int integer;
  float frac;
  uint32_t bi = f2u(x);
  int32_t i = (int32_t)bi;
  integer = (i >> 23) - 127;
  i -= integer << 23;
  frac = u2f((uint32_t)i) - 1.5f;
  frac = -0.41445418f + frac * (0.95909232f + frac * (-0.33951290f + frac * 0.16541097f));
  return (float)(1 + integer) + frac;
}

int orc_lin2ulaw(float x) {
  float u;
  float scale = 255.f / 32768.f;
  int s = x >= 0 ? 1 : -1;
  x = fabsf(x);
  u = ((float)s * ((float)128 * (0.69315f * log2_approx((float)1 + scale * x)) / 5.5451774445f));
  u = (float)128 + u;
  if (u < 0) u = 0;
  if (u > 255) u = 255;
  return (int)floor(.5 + (double)u);
}

float orc_ulaw2lin(float u) {
  float s;
  float scale_1 = 32768.f / 255.f;
  u = u - 128.f;
  s = u >= 0.f ? 1.f : -1.f;
  u = fabsf(u);
  return (float)((double)(s * scale_1) * (exp((double)u / 128. * (double)5.5451774445f) - 1));
}

/* ---- kiss99 (kiss99.c:32-81) ---------------------------------------------------- */
uint32_t orc_kiss99_rand(uint32_t st[4]) {
  uint32_t z = st[0], w = st[1], jsr = st[2], jcong = st[3];
  uint32_t znew = 36969u * (z & 0xFFFF) + (z >> 16);
  uint32_t wnew = 18000u * (w & 0xFFFF) + (w >> 16);
  uint32_t mwc = (znew << 16) + wnew;
  uint32_t shr3 = jsr ^ (jsr << 13);
  uint32_t cong;
  shr3 ^= shr3 >> 17;
  shr3 ^= shr3 << 5;
  cong = 69069u * jcong + 1234567u;
  st[0] = znew; st[1] = wnew; st[2] = shr3; st[3] = cong;
  return (mwc ^ cong) + shr3;
}

void orc_kiss99_srand(uint32_t st[4], const unsigned char *d, int n) {
  int i;
  st[0] = 362436069u; st[1] = 521288629u; st[2] = 123456789u; st[3] = 380116160u;
  for (i = 3; i < n; i += 4) {
    st[0] ^= d[i - 3]; st[1] ^= d[i - 2]; st[2] ^= d[i - 1]; st[3] ^= d[i];
    orc_kiss99_rand(st);
  }
  if (i - 3 < n) st[0] ^= d[i - 3];
  if (i - 2 < n) st[1] ^= d[i - 2];
  if (i - 1 < n) st[2] ^= d[i - 1];
  if (st[0] == 0 || st[0] == 0x9068FFFFu) st[0]++;
  if (st[1] == 0 || st[1] == 0x464FFFFFu) st[1]++;
  if (st[2] == 0) st[2]++;
}

/* ---- int8 path (vec_avx.h:321-336, 690-858) ------------------------------------- */
/* cvtps2dq: round-to-nearest-even; out of range gives the "integer indefinite". */
static int32_t cvt_rne(float x) {
  if (!(x >= -2147483648.f && x < 2147483648.f)) return INT32_MIN;
  return (int32_t)nearbyintf(x);
}

static unsigned char quant_u8(float x) {
  int32_t v = cvt_rne(fmaf(x, 127.f, 127.f));
  int32_t t = v < 0 ? 0 : (v > 65535 ? 65535 : v);  /* packus_epi32 */
  int16_t s = (int16_t)(uint16_t)t;                 /* packus_epi16 reads it as signed */
  return (unsigned char)(s < 0 ? 0 : (s > 255 ? 255 : s));
}

void orc_ps_to_epi8(unsigned char *x, const float *in, int len) {
  int i;
  for (i = 0; i < len; i++) x[i] = quant_u8(in[i]);
}

static int32_t sat16(int32_t v) { return v > 32767 ? 32767 : (v < -32768 ? -32768 : v); }

/* one 8x4 block: acc[r] += madd(maddubs(x4, w[r][0..3]), 1) */
static void block8x4(int32_t *acc, const signed char *w, const unsigned char *x4) {
  int r;
  for (r = 0; r < 8; r++) {
    int32_t p0 = sat16((int32_t)x4[0] * w[4 * r + 0] + (int32_t)x4[1] * w[4 * r + 1]);
    int32_t p1 = sat16((int32_t)x4[2] * w[4 * r + 2] + (int32_t)x4[3] * w[4 * r + 3]);
    acc[r] = (int32_t)((uint32_t)acc[r] + (uint32_t)(p0 + p1));
  }
}

#define ORC_SCALE (128.f * 127.f)
#define ORC_SCALE_1 (1.f / 128.f / 127.f)

void orc_sparse_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const int *idx, const float *xin) {
  unsigned char *x = malloc(cols);
  int i, j, r;
  orc_ps_to_epi8(x, xin, cols);
  for (i = 0; i < rows; i += 8) {
    int32_t acc[8];
    int colblocks = *idx++;
    for (r = 0; r < 8; r++) acc[r] = cvt_rne(out[i + r] * ORC_SCALE);
    for (j = 0; j < colblocks; j++) {
      int pos = *idx++;
      block8x4(acc, w, &x[pos]);
      w += 32;
    }
    for (r = 0; r < 8; r++) out[i + r] = (float)acc[r] * ORC_SCALE_1;
  }
  free(x);
}

void orc_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const float *xin) {
  unsigned char *x = malloc(cols);
  int i, j, r;
  orc_ps_to_epi8(x, xin, cols);
  for (i = 0; i < rows; i += 8) {
    int32_t acc[8];
    for (r = 0; r < 8; r++) acc[r] = cvt_rne(out[i + r] * ORC_SCALE);
    for (j = 0; j < cols; j += 4) {
      block8x4(acc, w, &x[j]);
      w += 32;
    }
    for (r = 0; r < 8; r++) out[i + r] = (float)acc[r] * ORC_SCALE_1;
  }
  free(x);
}

/* ---- fp32 matvec: sgemv_accum (nnet.c:73-86) over sgemv_accum16 (vec_avx.h:618-643) */
void orc_sgemv_accum(float *out, const float *w, int rows, int cols, int col_stride, const float *x) {
  int i, j;
  if (rows % 16 == 0) {
    for (i = 0; i < rows; i++) {
      float y = out[i];
      for (j = 0; j < cols; j++) y = fmaf(w[j * col_stride + i], x[j], y);
      out[i] = y;
    }
  } else {
    for (i = 0; i < rows; i++)
      for (j = 0; j < cols; j++) out[i] += w[j * col_stride + i] * x[j];
  }
}
