/* Oracle: the 100 Hz frame network and the cepstrum -> LPC chain.
 * TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off. */
#include <math.h>
#include <string.h>
#include "orc_internal.h"

/* vec_tanh / vec_sigmoid / linear (nnet.c:88-120, vec_avx.h:552-582) */
void orc_vec_act(float *y, const float *x, int n, int act) {
  int i;
  if (act == ORC_TANH) for (i = 0; i < n; i++) y[i] = orc_tanh_approx(x[i]);
  else if (act == ORC_SIGMOID) for (i = 0; i < n; i++) y[i] = orc_sigmoid_approx(x[i]);
  else for (i = 0; i < n; i++) y[i] = x[i];
}

/* _lpcnet_compute_dense (nnet.c:122-135) */
void orc_dense(const OrcDense *l, float *out, const float *in) {
  int i;
  for (i = 0; i < l->nout; i++) out[i] = l->bias[i];
  orc_sgemv_accum(out, l->w, l->nout, l->nin, l->nout, in);
  orc_vec_act(out, out, l->nout, l->act);
}

/* compute_conv1d (nnet.c:452-470) */
void orc_conv1d(const OrcConv *l, float *out, float *mem, const float *in) {
  float tmp[384];
  int i, hist = l->nin * (l->k - 1);
  memcpy(tmp, mem, hist * sizeof(float));
  memcpy(tmp + hist, in, l->nin * sizeof(float));
  for (i = 0; i < l->nout; i++) out[i] = l->bias[i];
  orc_sgemv_accum(out, l->w, l->nout, l->nin * l->k, l->nout, tmp);
  orc_vec_act(out, out, l->nout, l->act);
  memcpy(mem, tmp + l->nin, hist * sizeof(float));
}

/* ---- kiss_fft, nfft = 320, float build ----------------------------------------------
 * Tables as dump_lpcnet_tables.c:58-99 generates them (lpcnet_tables.c stores them with
 * %#0.9g, which round-trips binary32): twiddles from kiss_fft.c:409-424, bit-reversal
 * from kiss_fft.c:312-345, factors {5,64, 4,16, 4,4, 4,1} from kf_factor (351-405). */
typedef OrcCpx cpx;
static cpx g_tw[320];
static int16_t g_bitrev[320];
static float g_dct[18 * 18];
static int g_tables = 0;
static const int16_t g_factors[8] = {5, 64, 4, 16, 4, 4, 4, 1};

static void bitrev_rec(int Fout, int16_t *f, int fstride, const int16_t *factors) {
  int p = *factors++, m = *factors++, j;
  if (m == 1) {
    for (j = 0; j < p; j++) { *f = (int16_t)(Fout + j); f += fstride; }
  } else {
    for (j = 0; j < p; j++) { bitrev_rec(Fout, f, fstride * p, factors); f += fstride; Fout += m; }
  }
}

static void init_tables(void) {
  int i, j;
  if (g_tables) return;
  for (i = 0; i < 320; i++) {
    const double pi = 3.14159265358979323846264338327;
    double phase = (-2 * pi / 320) * i;
    g_tw[i].r = (float)cos(phase);
    g_tw[i].i = (float)sin(phase);
  }
  bitrev_rec(0, g_bitrev, 1, g_factors);
  for (i = 0; i < 18; i++)
    for (j = 0; j < 18; j++) {
      float v = (float)cos((i + .5) * j * M_PI / 18);
      if (j == 0) v = (float)((double)v * sqrt(.5));
      g_dct[i * 18 + j] = v;
    }
  g_tables = 1;
}

const float *orc_dct_table(void) { init_tables(); return g_dct; }

void orc_fft_tables(float *tw, int16_t *br, float *dct) {
  init_tables();
  memcpy(tw, g_tw, sizeof(g_tw)); memcpy(br, g_bitrev, sizeof(g_bitrev)); memcpy(dct, g_dct, sizeof(g_dct));
}

#define CMUL(m, a, b) do { (m).r = (a).r * (b).r - (a).i * (b).i; (m).i = (a).r * (b).i + (a).i * (b).r; } while (0)
#define CADD(res, a, b) do { (res).r = (a).r + (b).r; (res).i = (a).i + (b).i; } while (0)
#define CSUB(res, a, b) do { (res).r = (a).r - (b).r; (res).i = (a).i - (b).i; } while (0)
#define CADDTO(res, a) do { (res).r += (a).r; (res).i += (a).i; } while (0)

/* kf_bfly4 (kiss_fft.c:96-162) */
static void bfly4(cpx *Fout, int fstride, int m, int N, int mm) {
  int i, j;
  if (m == 1) {
    for (i = 0; i < N; i++) {
      cpx s0, s1;
      CSUB(s0, Fout[0], Fout[2]);
      CADDTO(Fout[0], Fout[2]);
      CADD(s1, Fout[1], Fout[3]);
      CSUB(Fout[2], Fout[0], s1);
      CADDTO(Fout[0], s1);
      CSUB(s1, Fout[1], Fout[3]);
      Fout[1].r = s0.r + s1.i;
      Fout[1].i = s0.i - s1.r;
      Fout[3].r = s0.r - s1.i;
      Fout[3].i = s0.i + s1.r;
      Fout += 4;
    }
  } else {
    cpx s[6];
    const int m2 = 2 * m, m3 = 3 * m;
    cpx *beg = Fout;
    for (i = 0; i < N; i++) {
      const cpx *tw1, *tw2, *tw3;
      Fout = beg + i * mm;
      tw3 = tw2 = tw1 = g_tw;
      for (j = 0; j < m; j++) {
        CMUL(s[0], Fout[m], *tw1);
        CMUL(s[1], Fout[m2], *tw2);
        CMUL(s[2], Fout[m3], *tw3);
        CSUB(s[5], *Fout, s[1]);
        CADDTO(*Fout, s[1]);
        CADD(s[3], s[0], s[2]);
        CSUB(s[4], s[0], s[2]);
        CSUB(Fout[m2], *Fout, s[3]);
        tw1 += fstride; tw2 += fstride * 2; tw3 += fstride * 3;
        CADDTO(*Fout, s[3]);
        Fout[m].r = s[5].r + s[4].i;
        Fout[m].i = s[5].i - s[4].r;
        Fout[m3].r = s[5].r - s[4].i;
        Fout[m3].i = s[5].i + s[4].r;
        ++Fout;
      }
    }
  }
}

/* kf_bfly5 (kiss_fft.c:229-295) */
static void bfly5(cpx *Fout, int fstride, int m, int N, int mm) {
  cpx *F0, *F1, *F2, *F3, *F4, *beg = Fout;
  cpx s[13], ya, yb;
  int i, u;
  ya = g_tw[fstride * m];
  yb = g_tw[fstride * 2 * m];
  for (i = 0; i < N; i++) {
    Fout = beg + i * mm;
    F0 = Fout; F1 = F0 + m; F2 = F0 + 2 * m; F3 = F0 + 3 * m; F4 = F0 + 4 * m;
    for (u = 0; u < m; ++u) {
      s[0] = *F0;
      CMUL(s[1], *F1, g_tw[u * fstride]);
      CMUL(s[2], *F2, g_tw[2 * u * fstride]);
      CMUL(s[3], *F3, g_tw[3 * u * fstride]);
      CMUL(s[4], *F4, g_tw[4 * u * fstride]);
      CADD(s[7], s[1], s[4]);
      CSUB(s[10], s[1], s[4]);
      CADD(s[8], s[2], s[3]);
      CSUB(s[9], s[2], s[3]);
      F0->r = F0->r + (s[7].r + s[8].r);
      F0->i = F0->i + (s[7].i + s[8].i);
      s[5].r = s[0].r + (s[7].r * ya.r + s[8].r * yb.r);
      s[5].i = s[0].i + (s[7].i * ya.r + s[8].i * yb.r);
      s[6].r = s[10].i * ya.i + s[9].i * yb.i;
      s[6].i = -(s[10].r * ya.i + s[9].r * yb.i);
      CSUB(*F1, s[5], s[6]);
      CADD(*F4, s[5], s[6]);
      s[11].r = s[0].r + (s[7].r * yb.r + s[8].r * ya.r);
      s[11].i = s[0].i + (s[7].i * yb.r + s[8].i * ya.r);
      s[12].r = s[9].i * ya.i - s[10].i * yb.i;
      s[12].i = s[10].r * yb.i - s[9].r * ya.i;
      CADD(*F2, s[11], s[12]);
      CSUB(*F3, s[11], s[12]);
      ++F0; ++F1; ++F2; ++F3; ++F4;
    }
  }
}

/* opus_fft_c + opus_fft_impl (kiss_fft.c:518-585) for the fixed factorisation */
void orc_fft320(const cpx *fin, cpx *fout) {
  int fstride[5], i, L = 4, m, m2;
  const float scale = 1.f / 320;
  init_tables();
  for (i = 0; i < 320; i++) {
    fout[g_bitrev[i]].r = scale * fin[i].r;
    fout[g_bitrev[i]].i = scale * fin[i].i;
  }
  fstride[0] = 1;
  for (i = 0; i < L; i++) fstride[i + 1] = fstride[i] * g_factors[2 * i];
  m = g_factors[2 * L - 1];
  for (i = L - 1; i >= 0; i--) {
    m2 = i != 0 ? g_factors[2 * i - 1] : 1;
    if (g_factors[2 * i] == 4) bfly4(fout, fstride[i], m, fstride[i], m2);
    else bfly5(fout, fstride[i], m, fstride[i], m2);
    m = m2;
  }
}

/* lpcn_lpc (freq.c:86-127), float build: MULT32_32_Q31 = *, SHR32/SHL32 = identity */
void orc_levinson(float *lpc, const float *ac, int p) {
  int i, j;
  float r, error = ac[0];
  memset(lpc, 0, p * sizeof(float));
  if (ac[0] != 0) {
    for (i = 0; i < p; i++) {
      float rr = 0;
      for (j = 0; j < i; j++) rr += lpc[j] * ac[i - j];
      rr += ac[i + 1];
      r = -rr / error;
      lpc[i] = r;
      for (j = 0; j < (i + 1) >> 1; j++) {
        float t1 = lpc[j], t2 = lpc[i - 1 - j];
        lpc[j] = t1 + r * t2;
        lpc[i - 1 - j] = t2 + r * t1;
      }
      error = error - (r * r) * error;
      if (error < .001f * ac[0]) break;
    }
  }
}

static const int16_t eband5ms[18] = {0, 1, 2, 3, 4, 5, 6, 7, 8, 10, 12, 14, 16, 20, 24, 28, 34, 40};
static const float compensation[18] = {0.8f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 1.f, 0.666667f, 0.5f, 0.5f,
                                       0.5f, 0.333333f, 0.25f, 0.25f, 0.2f, 0.166667f, 0.173913f};

/* lpc_from_cepstrum -> idct -> lpc_from_bands -> interp_band_gain -> inverse_transform
 * (freq.c:310-320, 230-240, 275-297, 202-215, 256-273) */
void orc_lpc_from_cepstrum(float *lpc, const float *cepstrum) {
  float tmp[18], Ex[18], Xr[161], ac[17];
  cpx x[320], y[320];
  int i, j;
  init_tables();
  memcpy(tmp, cepstrum, sizeof(tmp));
  tmp[0] += 4;
  for (i = 0; i < 18; i++) {
    float sum = 0;
    for (j = 0; j < 18; j++) sum += tmp[j] * g_dct[i * 18 + j];
    Ex[i] = (float)((double)sum * sqrt(2. / 18));
  }
  for (i = 0; i < 18; i++) Ex[i] = (float)(pow((double)10.f, (double)Ex[i]) * (double)compensation[i]);
  memset(Xr, 0, sizeof(Xr));
  for (i = 0; i < 17; i++) {
    int band_size = (eband5ms[i + 1] - eband5ms[i]) * 4;
    for (j = 0; j < band_size; j++) {
      float frac = (float)j / (float)band_size;
      Xr[eband5ms[i] * 4 + j] = ((float)1 - frac) * Ex[i] + frac * Ex[i + 1];
    }
  }
  Xr[160] = 0;
  for (i = 0; i < 161; i++) { x[i].r = Xr[i]; x[i].i = 0; }
  for (; i < 320; i++) { x[i].r = x[320 - i].r; x[i].i = -x[320 - i].i; }
  orc_fft320(x, y);
  ac[0] = (float)320 * y[0].r;
  for (i = 1; i < 17; i++) ac[i] = (float)320 * y[320 - i].r;
  ac[0] = (float)((double)ac[0] + ((double)ac[0] * 1e-4 + 320 / 12 / 38.));
  for (i = 1; i < 17; i++) ac[i] = (float)((double)ac[i] * (1 - 6e-5 * i * i));
  orc_levinson(lpc, ac, 16);
}

/* rc2lpc (lpcnet.c:57-78, the END2END build): step-up recursion from reflection coefficients; the pass for i = 0 of
 * the reference's loop moves nothing */
static void orc_rc2lpc(float *lpc, const float *rc) {
  float cur[16], nxt[16];
  int i, j;
  memcpy(cur, rc, sizeof(cur));
  for (i = 1; i < 16; i++) {
    for (j = 0; j < i; j++) nxt[j] = cur[j] + cur[i] * cur[i - j - 1];
    memcpy(cur, nxt, i * sizeof(float));
  }
  memcpy(lpc, cur, sizeof(cur));
}

void orc_model_set_end2end(OrcModel *m, int on) { m->end2end = on ? 1 : 0; }

/* run_frame_network (lpcnet.c:82-120) */
void orc_frame_network(const OrcModel *m, OrcState *s, const float *features, float *ga, float *gb, float *lpc) {
  float in[84], conv1_out[128], conv2_out[128], dense1_out[128], condition[128];
  int pitch, i;
  pitch = (int)floor(.1 + (double)(50 * features[18]) + 100);
  pitch = pitch > 255 ? 255 : (pitch < 33 ? 33 : pitch);
  memcpy(in, features, 20 * sizeof(float));
  memcpy(in + 20, &m->e_pitch.w[pitch * 64], 64 * sizeof(float));
  orc_conv1d(&m->conv1, conv1_out, s->conv1_state, in);
  if (s->frame_count < 1) memset(conv1_out, 0, sizeof(conv1_out));
  orc_conv1d(&m->conv2, conv2_out, s->conv2_state, conv1_out);
  if (s->frame_count < m->features_delay) memset(conv2_out, 0, sizeof(conv2_out));
  orc_dense(&m->dense1, dense1_out, conv2_out);
  orc_dense(&m->dense2, condition, dense1_out);
  orc_dense(&m->a_feat, ga, condition);
  orc_dense(&m->b_feat, gb, condition);
  if (m->end2end) {                       /* lpcnet.c:107-108 */
    orc_rc2lpc(lpc, condition);
  } else if (m->features_delay > 0) {
    int D = m->features_delay;
    memcpy(lpc, s->old_lpc[D - 1], 16 * sizeof(float));
    memmove(s->old_lpc[1], s->old_lpc[0], (D - 1) * 16 * sizeof(float));
    orc_lpc_from_cepstrum(s->old_lpc[0], features);
  } else {
    orc_lpc_from_cepstrum(lpc, features);
  }
  { /* lpc_weighting (freq.c:299-308) */
    float gamma_i = m->lpc_gamma;
    for (i = 0; i < 16; i++) { lpc[i] *= gamma_i; gamma_i *= m->lpc_gamma; }
  }
  if (s->frame_count < 1000) s->frame_count++;
}
