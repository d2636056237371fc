/* Oracle: per-frame feature analysis (the step before the synthesis path, and an inner step of the PLC):
 * lpcnet_compute_single_frame_features = preemphasis + compute_frame_features + process_single_frame
 * (src/lpcnet_enc.c:872-880, 498-577, 814-870, 913-933), and burg_cepstral_analysis (src/freq.c:152-199,
 * src/burg.c).  TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off: every C operator below is one
 * IEEE operation, in the reference's order. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "orc_internal.h"

#define FRAME 160
#define OVERLAP 160
#define WINDOW 320
#define NBANDS 18
#define PMAX 256
#define PMIN 32
#define TOFF 80

static const int16_t eband5ms[NBANDS] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 34, 40};


static float g_half_window[OVERLAP];
static int g_hw = 0;
/* dump_lpcnet_tables.c:83-84 (the table in lpcnet_tables.c:207 is this formula printed with 9 digits) */
const float *orc_half_window(void) {
  int i;
  if (!g_hw) {
    for (i = 0; i < OVERLAP; i++)
      g_half_window[i] = (float)sin(.5 * M_PI * sin(.5 * M_PI * (i + .5) / OVERLAP) * sin(.5 * M_PI * (i + .5) / OVERLAP));
    g_hw = 1;
  }
  return g_half_window;
}

OrcEncState *orc_enc_new(void) { return (OrcEncState *)calloc(1, sizeof(OrcEncState)); }   /* lpcnet_encoder_init */
void orc_enc_free(OrcEncState *s) { free(s); }

/* forward_transform (freq.c:242-254): real input, bins 0..160 of opus_fft (scaled by 1/320) */
static void forward_transform(OrcCpx *out, const float *in) {
  OrcCpx x[WINDOW], y[WINDOW];
  int i;
  for (i = 0; i < WINDOW; i++) { x[i].r = in[i]; x[i].i = 0; }
  orc_fft320(x, y);
  for (i = 0; i < WINDOW / 2 + 1; i++) out[i] = y[i];
}

/* lpcn_compute_band_energy (freq.c:128-150) */
static void band_energy(float *bandE, const OrcCpx *X) {
  float sum[NBANDS] = {0};
  int i, j;
  for (i = 0; i < NBANDS - 1; i++) {
    int band_size = (eband5ms[i + 1] - eband5ms[i]) * 4;
    for (j = 0; j < band_size; j++) {
      float frac = (float)j / band_size;
      float tmp = X[eband5ms[i] * 4 + j].r * X[eband5ms[i] * 4 + j].r;
      tmp += X[eband5ms[i] * 4 + j].i * X[eband5ms[i] * 4 + j].i;
      sum[i] += (1 - frac) * tmp;
      sum[i + 1] += frac * tmp;
    }
  }
  sum[0] *= 2;
  sum[NBANDS - 1] *= 2;
  for (i = 0; i < NBANDS; i++) bandE[i] = sum[i];
}

/* dct (freq.c:218-228) */
static void dct18(float *out, const float *in) {
  const float *T = orc_dct_table();
  int i, j;
  for (i = 0; i < NBANDS; i++) {
    float sum = 0;
    for (j = 0; j < NBANDS; j++) sum += in[j] * T[j * NBANDS + i];
    out[i] = (float)(sum * sqrt(2. / NBANDS));
  }
}

static float max16(float a, float b) { return a > b ? a : b; }

/* the log-energy smoothing shared by compute_frame_features (lpcnet_enc.c:516-522) and compute_burg_cepstrum */
static void log_bands(float *Ly, const float *Ex, int float_consts) {
  float logMax = -2, follow = -2;
  int i;
  for (i = 0; i < NBANDS; i++) {
    Ly[i] = (float)log10(1e-2 + Ex[i]);
    if (float_consts) {
      Ly[i] = max16(logMax - 8, max16(follow - 2.5f, Ly[i]));
      logMax = max16(logMax, Ly[i]);
      follow = max16(follow - 2.5f, Ly[i]);
    } else {   /* freq.c:180-183 writes 2.5 (double): those comparisons and selections run in double */
      const double f25 = (double)follow - 2.5;
      const double inner = f25 > (double)Ly[i] ? f25 : (double)Ly[i];
      const float lm8 = logMax - 8;
      Ly[i] = (float)((double)lm8 > inner ? (double)lm8 : inner);
      logMax = max16(logMax, Ly[i]);
      follow = (float)(f25 > (double)Ly[i] ? f25 : (double)Ly[i]);
    }
  }
}

/* compute_frame_features (lpcnet_enc.c:498-577); writes rows 0, 1 of st->xc / st->frame_weight and st->features
 * (the reference's rows 2+2*pcount+sub and features[pcount]: the superframe encoder copies them out, orc_encode.c) */
void orc_enc_frame_features(OrcEncState *st, const float *in) {
  const float *hw = orc_half_window();
  float aligned_in[FRAME], x[WINDOW], Ex[NBANDS], Ly[NBANDS], xcorr[PMAX];
  OrcCpx X[WINDOW / 2 + 1];
  int i, j, sub;
  memcpy(aligned_in, &st->analysis_mem[OVERLAP - TOFF], TOFF * sizeof(float));
  /* frame_analysis (lpcnet_enc.c:488-496), apply_window (freq.c:322-328) */
  memcpy(x, st->analysis_mem, OVERLAP * sizeof(float));
  memcpy(&x[OVERLAP], in, FRAME * sizeof(float));
  memcpy(st->analysis_mem, &in[FRAME - OVERLAP], OVERLAP * sizeof(float));
  for (i = 0; i < OVERLAP; i++) { x[i] *= hw[i]; x[WINDOW - 1 - i] *= hw[i]; }
  forward_transform(X, x);
  band_energy(Ex, X);
  log_bands(Ly, Ex, 1);
  dct18(st->features, Ly);
  st->features[0] -= 4;
  orc_lpc_from_cepstrum(st->lpc, st->features);
  for (i = 0; i < ORC_LPC_ORDER; i++) st->features[NBANDS + 2 + i] = st->lpc[i];
  memmove(st->exc_buf, &st->exc_buf[FRAME], PMAX * sizeof(float));
  memcpy(&aligned_in[TOFF], in, (FRAME - TOFF) * sizeof(float));
  for (i = 0; i < FRAME; i++) {
    float sum = aligned_in[i];
    for (j = 0; j < ORC_LPC_ORDER; j++) sum += st->lpc[j] * st->pitch_mem[j];
    memmove(st->pitch_mem + 1, st->pitch_mem, (ORC_LPC_ORDER - 1) * sizeof(float));
    st->pitch_mem[0] = aligned_in[i];
    st->exc_buf[PMAX + i] = sum + .7f * st->pitch_filt;
    st->pitch_filt = sum;
  }
  for (sub = 0; sub < 2; sub++) {
    static const float interp[7] = {0.026184f, -0.098339f, 0.369938f, 0.837891f, -0.184969f, 0.070242f, -0.020947f};
    const int off = sub * FRAME / 2;
    const float *xs = &st->exc_buf[PMAX + off], *ys = st->exc_buf + off;
    float interpolated[PMAX] = {0};
    float ener0 = 0, ener, e1f = 0;
    double ener1;
    /* celt_pitch_xcorr / xcorr_kernel (pitch.c:43-83, pitch.h:42-105): each lag is a j-ascending sum from 0 */
    for (i = 0; i < PMAX; i++) {
      float s = 0;
      for (j = 0; j < FRAME / 2; j++) s = s + xs[j] * ys[i + j];
      xcorr[i] = s;
    }
    for (j = 0; j < FRAME / 2; j++) ener0 = ener0 + xs[j] * xs[j];
    for (j = 0; j < FRAME / 2 - 1; j++) e1f = e1f + ys[j] * ys[j];
    ener1 = e1f;
    st->frame_weight[sub] = ener0;
    for (i = 0; i < PMAX; i++) {
      ener1 += st->exc_buf[i + off + FRAME / 2 - 1] * st->exc_buf[i + off + FRAME / 2 - 1];
      ener = (float)((1 + ener0) + ener1);
      st->xc[sub][i] = 2 * xcorr[i] / ener;
      ener1 -= st->exc_buf[i + off] * st->exc_buf[i + off];
    }
    for (i = 4; i < PMAX - 4; i++) {
      float val1 = 0, val2 = 0;
      for (j = 0; j < 7; j++) {
        val1 += st->xc[sub][i - 3 + j] * interp[j];
        val2 += st->xc[sub][i + 3 - j] * interp[j];
      }
      interpolated[i] = max16(st->xc[sub][i], max16(val1, val2));
    }
    for (i = 4; i < PMAX - 4; i++) st->xc[sub][i] = interpolated[i];
  }
}

/* process_single_frame (lpcnet_enc.c:814-870) with pcount = 0 */
static void single_frame_pitch(OrcEncState *st) {
  int pitch_prev[2][PMAX];
  int best[2];
  int i, j, sub, best_i;
  float frame_corr, frame_weight_sum = 1e-15f;
  for (sub = 0; sub < 2; sub++) frame_weight_sum += st->frame_weight[sub];
  for (sub = 0; sub < 2; sub++) st->frame_weight[sub] *= (2.f / frame_weight_sum);
  for (sub = 0; sub < 2; sub++) {
    float max_path_all = -1e15f;
    float *xc = st->xc[sub];
    best_i = 0;
    for (i = 0; i < PMAX - 2 * PMIN; i++) {
      float xc_half = max16(max16(xc[(PMAX + i) / 2], xc[(PMAX + i + 2) / 2]), xc[(PMAX + i - 1) / 2]);
      if (xc[i] < xc_half * 1.1f) xc[i] *= .8f;
    }
    for (i = 0; i < PMAX - PMIN; i++) {
      float max_prev = st->pitch_max_path_all - 6.f;
      pitch_prev[sub][i] = st->best_i;
      for (j = (-4 > -i ? -4 : -i); j <= 4 && i + j < PMAX - PMIN; j++) {
        float c = st->pitch_max_path[0][i + j] - .02f * abs(j) * abs(j);
        if (c > max_prev) { max_prev = c; pitch_prev[sub][i] = i + j; }
      }
      st->pitch_max_path[1][i] = max_prev + st->frame_weight[sub] * xc[i];
      if (st->pitch_max_path[1][i] > max_path_all) { max_path_all = st->pitch_max_path[1][i]; best_i = i; }
    }
    for (i = 0; i < PMAX - PMIN; i++) st->pitch_max_path[1][i] -= max_path_all;
    memcpy(st->pitch_max_path[0], st->pitch_max_path[1], PMAX * sizeof(float));
    st->pitch_max_path_all = max_path_all;
    st->best_i = best_i;
  }
  best_i = st->best_i;
  frame_corr = 0;
  for (sub = 1; sub >= 0; sub--) {
    best[sub] = PMAX - best_i;
    frame_corr += st->frame_weight[sub] * st->xc[sub][best_i];
    best_i = pitch_prev[sub][best_i];
  }
  frame_corr /= 2;
  {
    int p = best[0] + best[1];
    p = p < 510 ? p : 510;
    p = p > 66 ? p : 66;
    st->features[NBANDS] = .01f * (p - 200);
  }
  st->features[NBANDS + 1] = frame_corr - .5f;
}

/* lpcnet_compute_single_frame_features_float (lpcnet_enc.c:913-933); pcm as floats (shorts convert exactly) */
void orc_enc_single_frame(OrcEncState *st, const float *pcm, float *features) {
  float x[FRAME];
  int i;
  for (i = 0; i < FRAME; i++) {            /* preemphasis (lpcnet_enc.c:872-880), coef 0.85f */
    float yi = pcm[i] + st->mem_preemph;
    st->mem_preemph = -0.85f * pcm[i];
    x[i] = yi;
  }
  orc_enc_frame_features(st, x);
  single_frame_pitch(st);
  memcpy(features, st->features, sizeof(st->features));
}

/* ---- burg_cepstral_analysis (freq.c:152-199) over silk_burg_analysis (burg.c:97-245) ------------------------ */

/* silk_energy_FLP / silk_inner_product_FLP (burg.c:41-92): 4-way grouped double sums */
static double inner_flp(const float *a, const float *b, int n) {
  double r = 0.0;
  int i;
  for (i = 0; i < n - 3; i += 4)
    r += a[i + 0] * (double)b[i + 0] + a[i + 1] * (double)b[i + 1] + a[i + 2] * (double)b[i + 2] + a[i + 3] * (double)b[i + 3];
  for (; i < n; i++) r += a[i] * (double)b[i];
  return r;
}

/* silk_burg_analysis with nb_subfr = 1 (the only use, freq.c:168) */
static float burg_analysis(float *A, const float *x, float minInvGain, int len, int D) {
  double C0, invGain, num, nrg_f, nrg_b, rc, Atmp, tmp1, tmp2;
  double Cf[16], Cl[16], CAf[17], CAb[17], Af[16];
  int k, n, reached = 0;
  C0 = inner_flp(x, x, len);
  memset(Cf, 0, sizeof(Cf));
  for (n = 1; n < D + 1; n++) Cf[n - 1] += inner_flp(x, x + n, len - n);
  memcpy(Cl, Cf, sizeof(Cf));
  CAb[0] = CAf[0] = C0 + 1e-5f * C0 + 1e-9f;
  invGain = 1.0f;
  for (n = 0; n < D; n++) {
    tmp1 = x[n];
    tmp2 = x[len - n - 1];
    for (k = 0; k < n; k++) {
      Cf[k] -= x[n] * x[n - k - 1];                     /* float product */
      Cl[k] -= x[len - n - 1] * x[len - n + k];
      Atmp = Af[k];
      tmp1 += x[n - k - 1] * Atmp;
      tmp2 += x[len - n + k] * Atmp;
    }
    for (k = 0; k <= n; k++) {
      CAf[k] -= tmp1 * x[n - k];
      CAb[k] -= tmp2 * x[len - n + k - 1];
    }
    tmp1 = Cf[n];
    tmp2 = Cl[n];
    for (k = 0; k < n; k++) {
      Atmp = Af[k];
      tmp1 += Cl[n - k - 1] * Atmp;
      tmp2 += Cf[n - k - 1] * Atmp;
    }
    CAf[n + 1] = tmp1;
    CAb[n + 1] = tmp2;
    num = CAb[n + 1];
    nrg_b = CAb[0];
    nrg_f = CAf[0];
    for (k = 0; k < n; k++) {
      Atmp = Af[k];
      num += CAb[n - k] * Atmp;
      nrg_b += CAb[k + 1] * Atmp;
      nrg_f += CAf[k + 1] * Atmp;
    }
    rc = -2.0 * num / (nrg_f + nrg_b);
    tmp1 = invGain * (1.0 - rc * rc);
    if (tmp1 <= minInvGain) {
      rc = sqrt(1.0 - minInvGain / invGain);
      if (num > 0) rc = -rc;
      invGain = minInvGain;
      reached = 1;
    } else {
      invGain = tmp1;
    }
    for (k = 0; k < (n + 1) >> 1; k++) {
      tmp1 = Af[k];
      tmp2 = Af[n - k - 1];
      Af[k] = tmp1 + rc * tmp2;
      Af[n - k - 1] = tmp2 + rc * tmp1;
    }
    Af[n] = rc;
    if (reached) {
      for (k = n + 1; k < D; k++) Af[k] = 0.0;
      break;
    }
    for (k = 0; k <= n + 1; k++) {
      tmp1 = CAf[k];
      CAf[k] += rc * CAb[n - k + 1];
      CAb[n - k + 1] += rc * tmp1;
    }
  }
  if (reached) {
    for (k = 0; k < D; k++) A[k] = (float)(-Af[k]);
    C0 -= inner_flp(x, x, D);
    nrg_f = C0 * invGain;
  } else {
    nrg_f = CAf[0];
    tmp1 = 1.0;
    for (k = 0; k < D; k++) {
      Atmp = Af[k];
      nrg_f += CAf[k + 1] * Atmp;
      tmp1 += Atmp * Atmp;
      A[k] = (float)(-Atmp);
    }
    nrg_f -= 1e-5f * C0 * tmp1;
  }
  return (float)nrg_f;
}

/* compute_burg_cepstrum (freq.c:152-186) for len = 80, order = 16 */
static void burg_cepstrum(const float *pcm, float *out) {
  float burg_in[FRAME], burg_lpc[ORC_LPC_ORDER], x[WINDOW], Eburg[NBANDS], Ly[NBANDS], sum[NBANDS] = {0}, g;
  OrcCpx X[WINDOW / 2 + 1];
  const int len = FRAME / 2, order = ORC_LPC_ORDER;
  int i, j;
  for (i = 0; i < len - 1; i++) burg_in[i] = pcm[i + 1] - 0.85f * pcm[i];
  g = burg_analysis(burg_lpc, burg_in, 1e-3, len - 1, order);
  g /= len - 2 * (order - 1);
  memset(x, 0, sizeof(x));
  x[0] = 1;
  for (i = 0; i < order; i++) x[i + 1] = (float)(-burg_lpc[i] * pow(.995, i + 1));
  forward_transform(X, x);
  /* compute_band_energy_inverse (freq.c:59-84) */
  for (i = 0; i < NBANDS - 1; i++) {
    int band_size = (eband5ms[i + 1] - eband5ms[i]) * 4;
    for (j = 0; j < band_size; j++) {
      float frac = (float)j / band_size;
      float tmp = X[eband5ms[i] * 4 + j].r * X[eband5ms[i] * 4 + j].r;
      tmp += X[eband5ms[i] * 4 + j].i * X[eband5ms[i] * 4 + j].i;
      tmp = (float)(1.f / (tmp + 1e-9));
      sum[i] += (1 - frac) * tmp;
      sum[i + 1] += frac * tmp;
    }
  }
  sum[0] *= 2;
  sum[NBANDS - 1] *= 2;
  for (i = 0; i < NBANDS; i++) Eburg[i] = (float)(sum[i] * (.45 * g * (1.f / ((float)WINDOW * WINDOW * WINDOW))));
  log_bands(Ly, Eburg, 0);
  dct18(out, Ly);
  out[0] += -4;
}

/* burg_cepstral_analysis (freq.c:188-199): mean and difference of the two half-frame cepstra */
void orc_burg_cepstral_analysis(float *ceps /*[36]*/, const float *x /*[160]*/) {
  int i;
  burg_cepstrum(x, &ceps[0]);
  burg_cepstrum(&x[FRAME / 2], &ceps[NBANDS]);
  for (i = 0; i < NBANDS; i++) {
    float c0 = ceps[i], c1 = ceps[NBANDS + i];
    ceps[i] = (float)(.5 * (c0 + c1));
    ceps[NBANDS + i] = (c0 - c1);
  }
}
