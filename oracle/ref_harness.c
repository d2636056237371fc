/* Test harness around the UNMODIFIED reference sources (compiled in place from
 * /root/reference/src by oracle/Makefile into oracle/_ref/).  TEST INFRASTRUCTURE ONLY:
 * it exists to pin oracle/lpcnet_oracle.c and to generate tests/golden/ fixtures; the
 * product (lpcnet_b200/) never links or calls it.
 *
 * Everything computed here is computed by reference code: the sample loop is driven
 * one sample at a time through the reference's own lpcnet_synthesize_tail_impl(N=1)
 * (src/lpcnet.c:235-271), so per-step state can be observed between steps without
 * restating the loop.  The only arithmetic restated in this file is (a) the LPC
 * prediction / mu-law index pair that the loop computes as locals (lpcnet.c:252-254),
 * recomputed only to *record* them, and (b) the all-255-node dual-FC logit sweep used
 * by the pdf check, which follows sample_mdense's inner loop (src/nnet.c:192-202) with
 * the reference's own tanh_approx.
 */
#include <stdlib.h>
#include <string.h>
#include <stdint.h>
#include <math.h>
#include <xmmintrin.h>
#include "lpcnet_private.h"
#include "nnet.h"
#include "common.h"

typedef struct {
  LPCNetState *st;
  unsigned char *blob;
} RefHandle;

int ref_features_delay(void) { return FEATURES_DELAY; }
int ref_state_size(void) { return lpcnet_get_size(); }

void *ref_new(const unsigned char *blob, int len) {
  RefHandle *h = calloc(1, sizeof(*h));
  h->blob = malloc(len);
  memcpy(h->blob, blob, len);
  h->st = lpcnet_create();
  if (lpcnet_load_model(h->st, h->blob, len) != 0) {
    lpcnet_destroy(h->st); free(h->blob); free(h);
    return NULL;
  }
  return h;
}

void ref_free(void *hp) {
  RefHandle *h = hp;
  if (!h) return;
  lpcnet_destroy(h->st); free(h->blob); free(h);
}

void ref_seed(void *hp, const unsigned char *data, int n) {
  RefHandle *h = hp;
  kiss99_srand(&h->st->rng, data, n);
}

void ref_reset(void *hp) { lpcnet_reset(((RefHandle *)hp)->st); }

void ref_synthesize(void *hp, const float *features, short *out, int N, int preload) {
  lpcnet_synthesize_impl(((RefHandle *)hp)->st, features, out, N, preload);
}

void ref_public_synthesize(void *hp, const float *features, short *out, int N) {
  lpcnet_synthesize(((RefHandle *)hp)->st, features, out, N);
}

/* ---- state snapshot (for cross-loading into the CUDA engine and the C oracle) ---- */
typedef struct {
  float conv1_state[FEATURE_CONV1_STATE_SIZE];
  float conv2_state[FEATURE_CONV2_STATE_SIZE];
  float gru_a_state[GRU_A_STATE_SIZE];
  float gru_b_state[GRU_B_STATE_SIZE];
  float last_sig[LPC_ORDER];
  float lpc[LPC_ORDER];
  float old_lpc[2][LPC_ORDER];
  float gru_a_condition[3 * GRU_A_STATE_SIZE];
  float gru_b_condition[3 * GRU_B_STATE_SIZE];
  float deemph_mem;
  int32_t last_exc;
  int32_t frame_count;
  uint32_t rng[4];
} RefSnapshot;

int ref_snapshot_size(void) { return sizeof(RefSnapshot); }

void ref_get(void *hp, RefSnapshot *s) {
  LPCNetState *st = ((RefHandle *)hp)->st;
  memset(s, 0, sizeof(*s));
  memcpy(s->conv1_state, st->nnet.feature_conv1_state, sizeof(s->conv1_state));
  memcpy(s->conv2_state, st->nnet.feature_conv2_state, sizeof(s->conv2_state));
  memcpy(s->gru_a_state, st->nnet.gru_a_state, sizeof(s->gru_a_state));
  memcpy(s->gru_b_state, st->nnet.gru_b_state, sizeof(s->gru_b_state));
  memcpy(s->last_sig, st->last_sig, sizeof(s->last_sig));
  memcpy(s->lpc, st->lpc, sizeof(s->lpc));
#if FEATURES_DELAY > 0
  memcpy(s->old_lpc, st->old_lpc, sizeof(float) * LPC_ORDER * (FEATURES_DELAY < 2 ? FEATURES_DELAY : 2));
#endif
  memcpy(s->gru_a_condition, st->gru_a_condition, sizeof(s->gru_a_condition));
  memcpy(s->gru_b_condition, st->gru_b_condition, sizeof(s->gru_b_condition));
  s->deemph_mem = st->deemph_mem;
  s->last_exc = st->last_exc;
  s->frame_count = st->frame_count;
  s->rng[0] = st->rng.z; s->rng[1] = st->rng.w; s->rng[2] = st->rng.jsr; s->rng[3] = st->rng.jcong;
}

void ref_set(void *hp, const RefSnapshot *s) {
  LPCNetState *st = ((RefHandle *)hp)->st;
  memcpy(st->nnet.feature_conv1_state, s->conv1_state, sizeof(s->conv1_state));
  memcpy(st->nnet.feature_conv2_state, s->conv2_state, sizeof(s->conv2_state));
  memcpy(st->nnet.gru_a_state, s->gru_a_state, sizeof(s->gru_a_state));
  memcpy(st->nnet.gru_b_state, s->gru_b_state, sizeof(s->gru_b_state));
  memcpy(st->last_sig, s->last_sig, sizeof(s->last_sig));
  memcpy(st->lpc, s->lpc, sizeof(s->lpc));
#if FEATURES_DELAY > 0
  memcpy(st->old_lpc, s->old_lpc, sizeof(float) * LPC_ORDER * (FEATURES_DELAY < 2 ? FEATURES_DELAY : 2));
#endif
  memcpy(st->gru_a_condition, s->gru_a_condition, sizeof(s->gru_a_condition));
  memcpy(st->gru_b_condition, s->gru_b_condition, sizeof(s->gru_b_condition));
  st->deemph_mem = s->deemph_mem;
  st->last_exc = s->last_exc;
  st->frame_count = s->frame_count;
  st->rng.z = s->rng[0]; st->rng.w = s->rng[1]; st->rng.jsr = s->rng[2]; st->rng.jcong = s->rng[3];
}

/* ---- per-step trace ---- */
typedef struct {
  int32_t exc;        /* st->last_exc after the step */
  int32_t u_sig;      /* lin2ulaw(last_sig[0]) fed to the step (lpcnet.c:253) */
  int32_t u_pred;     /* lin2ulaw(pred) fed to the step (lpcnet.c:254) */
  uint32_t rng_z;     /* kiss99 state after the step */
  uint32_t rng_w;
  uint32_t xa_hash;   /* FNV-1a over vector_ps_to_epi8(gru_a_state) after the step */
  uint32_t xb_hash;   /* same for gru_b_state */
  int32_t pcm;        /* output sample (or the forced one) */
} RefTrace;

static uint32_t fnv1a(const unsigned char *p, int n) {
  uint32_t h = 2166136261u;
  int i;
  for (i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; }
  return h;
}

void ref_all_node_logits(void *hp, float *logits /*[256], [0] unused = 0*/) {
  LPCNetState *st = ((RefHandle *)hp)->st;
  const MDenseLayer *layer = &st->model.dual_fc;
  const float *input = st->nnet.gru_b_state;
  int M = layer->nb_inputs, N = layer->nb_neurons, C = layer->nb_channels;
  int stride = M * C;
  int i, j;
  logits[0] = 0;
  for (i = 1; i < N; i++) {
    float sum1 = layer->bias[i];
    float sum2 = layer->bias[i + N];
    for (j = 0; j < M; j++) {
      sum1 += layer->input_weights[i * stride + j] * input[j];
      sum2 += layer->input_weights[i * stride + j + M] * input[j];
    }
    sum1 = layer->factor[i] * tanh_approx(sum1);
    sum2 = layer->factor[N + i] * tanh_approx(sum2);
    sum1 += sum2;
    logits[i] = sum1;
  }
}

/* Run n_frames frames of one stream.  feats: [n_frames][feat_stride] floats.
 * forced: NULL (free-running) or [n_frames*160] teacher-forcing PCM, applied through the
 * reference's own preload mechanism (lpcnet.c:256-258).  pcm: [n_frames*160] out.
 * trace: NULL or [n_frames*160]; logits: NULL or [n_frames*160][256].
 * Steps are issued as tail_impl(N=1) so state is observable between samples. */
void ref_run(void *hp, const float *feats, int feat_stride, int n_frames, short *pcm,
             const short *forced, RefTrace *trace, float *logits) {
  LPCNetState *st = ((RefHandle *)hp)->st;
  int f, i, j;
  for (f = 0; f < n_frames; f++) {
    short *out = &pcm[f * FRAME_SIZE];
    run_frame_network(st, st->gru_a_condition, st->gru_b_condition, st->lpc, &feats[f * feat_stride]);
    if (forced) memcpy(out, &forced[f * FRAME_SIZE], FRAME_SIZE * sizeof(short));
    for (i = 0; i < FRAME_SIZE; i++) {
      int t = f * FRAME_SIZE + i;
      int active = st->frame_count > FEATURES_DELAY;
      RefTrace tr;
      memset(&tr, 0, sizeof(tr));
      if (active && trace) {
        float pred = 0;
        for (j = 0; j < LPC_ORDER; j++) pred -= st->last_sig[j] * st->lpc[j];
        tr.u_sig = lin2ulaw(st->last_sig[0]);
        tr.u_pred = lin2ulaw(pred);
      }
      lpcnet_synthesize_tail_impl(st, &out[i], 1, forced ? 1 : 0);
      if (!active && forced) out[i] = 0; /* tail_impl zeroes output during warm-up */
      if (trace) {
        if (active) {
          unsigned char xa[GRU_A_STATE_SIZE + 32], xb[GRU_B_STATE_SIZE + 32]; /* the AVX2 routine stores 32 bytes per 8 inputs */
          vector_ps_to_epi8(xa, st->nnet.gru_a_state, GRU_A_STATE_SIZE);
          vector_ps_to_epi8(xb, st->nnet.gru_b_state, GRU_B_STATE_SIZE);
          tr.exc = st->last_exc;
          tr.rng_z = st->rng.z; tr.rng_w = st->rng.w;
          tr.xa_hash = fnv1a(xa, GRU_A_STATE_SIZE);
          tr.xb_hash = fnv1a(xb, GRU_B_STATE_SIZE);
        }
        tr.pcm = out[i];
        trace[t] = tr;
      }
      if (logits) {
        if (active) ref_all_node_logits(hp, &logits[(size_t)t * 256]);
        else memset(&logits[(size_t)t * 256], 0, 256 * sizeof(float));
      }
    }
  }
}

/* ---- kernel-level known-answer helpers (the reference's inline functions) ---- */
float ref_tanh_approx(float x) { return tanh_approx(x); }
float ref_sigmoid_approx(float x) { return sigmoid_approx(x); }
int ref_lin2ulaw(float x) { return lin2ulaw(x); }
float ref_ulaw2lin(float u) { return ulaw2lin(u); }
float ref_rcp_ss(float x) { return _mm_cvtss_f32(_mm_rcp_ss(_mm_set_ss(x))); }
void ref_vec_tanh(float *y, const float *x, int n) { vec_tanh(y, x, n); }
void ref_vec_sigmoid(float *y, const float *x, int n) { vec_sigmoid(y, x, n); }
void ref_ps_to_epi8(unsigned char *x, const float *in, int len) { vector_ps_to_epi8(x, in, len); }
void ref_sparse_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const int *idx, const float *x) {
  sparse_sgemv_accum8x4(out, w, rows, cols, idx, x);
}
void ref_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const float *x) {
  sgemv_accum8x4(out, w, rows, cols, rows, x);
}
void ref_sgemv_accum16(float *out, const float *w, int rows, int cols, int col_stride, const float *x) {
  sgemv_accum16(out, w, rows, cols, col_stride, x);
}
void ref_kiss99_srand(uint32_t st[4], const unsigned char *d, int n) {
  kiss99_ctx c; kiss99_srand(&c, d, n);
  st[0] = c.z; st[1] = c.w; st[2] = c.jsr; st[3] = c.jcong;
}
uint32_t ref_kiss99_rand(uint32_t st[4]) {
  kiss99_ctx c; uint32_t r;
  c.z = st[0]; c.w = st[1]; c.jsr = st[2]; c.jcong = st[3];
  r = kiss99_rand(&c);
  st[0] = c.z; st[1] = c.w; st[2] = c.jsr; st[3] = c.jcong;
  return r;
}
void ref_lpc_from_cepstrum(float *lpc, const float *ceps) { lpc_from_cepstrum(lpc, ceps); }
void ref_sampling_logit_table(void *hp, float *t) {
  memcpy(t, ((RefHandle *)hp)->st->sampling_logit_table, 256 * sizeof(float));
}
void ref_frame_network(void *hp, const float *features, float *ga, float *gb, float *lpc) {
  LPCNetState *st = ((RefHandle *)hp)->st;
  run_frame_network(st, ga, gb, lpc, features);
}
void ref_decode_packet(float *features /*[4][36]*/, float *vq_mem, const unsigned char *buf) {
  decode_packet((float(*)[NB_TOTAL_FEATURES])features, vq_mem, buf);
}

/* ---- feature analysis / PLC (lpcnet_enc.c, lpcnet_plc.c): thin handles over the reference's own API ---------- */
#include "lpcnet.h"
#include "plc_data.h"
#include "freq.h"

void ref_burg_cepstral_analysis(float *ceps, const float *x) { burg_cepstral_analysis(ceps, x); }

typedef struct {
  LPCNetPLCState *st;
  unsigned char *blob;
} RefPLC;

/* lpcnet_plc_create + lpcnet_plc_load_model (lpcnet_plc.c:100-109, 88-98); the blob is borrowed, so keep a copy */
void *ref_plc_new(const unsigned char *blob, int len, int options) {
  RefPLC *h = calloc(1, sizeof(*h));
  h->blob = malloc(len);
  memcpy(h->blob, blob, len);
  h->st = lpcnet_plc_create(options);
  if (!h->st || lpcnet_plc_load_model(h->st, h->blob, len) != 0) {
    if (h->st) lpcnet_plc_destroy(h->st);
    free(h->blob); free(h);
    return NULL;
  }
  return h;
}
void ref_plc_free(void *hp) {
  RefPLC *h = hp;
  if (!h) return;
  lpcnet_plc_destroy(h->st); free(h->blob); free(h);
}
void ref_plc_seed(void *hp, const unsigned char *data, int n) { kiss99_srand(&((RefPLC *)hp)->st->lpcnet.rng, data, n); }
int ref_plc_update(void *hp, short *pcm) { return lpcnet_plc_update(((RefPLC *)hp)->st, pcm); }
int ref_plc_conceal(void *hp, short *pcm) { return lpcnet_plc_conceal(((RefPLC *)hp)->st, pcm); }
void ref_plc_fec_add(void *hp, const float *features) { lpcnet_plc_fec_add(((RefPLC *)hp)->st, features); }
void ref_plc_fec_clear(void *hp) { lpcnet_plc_fec_clear(((RefPLC *)hp)->st); }
/* observable PLC state for step-by-step comparison */
void ref_plc_peek(void *hp, float *features20, float *gru1, float *gru2, int *loss_count, double *dc /*[2]*/) {
  LPCNetPLCState *st = ((RefPLC *)hp)->st;
  memcpy(features20, st->features, 20 * sizeof(float));
  memcpy(gru1, st->plc_net.plc_gru1_state, sizeof(st->plc_net.plc_gru1_state));
  memcpy(gru2, st->plc_net.plc_gru2_state, sizeof(st->plc_net.plc_gru2_state));
  *loss_count = st->loss_count;
  dc[0] = st->dc_mem; dc[1] = st->syn_dc;
}
/* compute_plc_pred (lpcnet_plc.c:135-145) is static there: the same five calls on the reference's layers */
void ref_plc_pred(void *hp, float *gru1, float *gru2, float *out, const float *in) {
  LPCNetPLCState *st = ((RefPLC *)hp)->st;
  float zeros[3 * PLC_MAX_RNN_NEURONS] = {0};
  float dense_out[PLC_DENSE1_OUT_SIZE];
  _lpcnet_compute_dense(&st->model.plc_dense1, dense_out, in);
  compute_gruB(&st->model.plc_gru1, zeros, gru1, dense_out);
  compute_gruB(&st->model.plc_gru2, zeros, gru2, gru1);
  _lpcnet_compute_dense(&st->model.plc_out, out, gru2);
  out[19] = MIN16(.5f, out[19] + .1f);
}

/* ---- DRED: the RDOVAE decoder (dred_rdovae_dec.c) on a model parsed from the blob ------------------------------- */
#include "dred_rdovae_dec.h"

/* DRED_rdovae_decode_all (dred_rdovae.c:37-51) over the reference's dred_rdovae_dec_init_states /
 * dred_rdovae_decode_qframe: features [nb_latents][4*20]; returns 0, or -1 if the blob lacks the decoder arrays */
int ref_dred_decode_all(const unsigned char *blob, int len, float *features, const float *state, const float *latents,
                        int nb_latents) {
  WeightArray *list;
  RDOVAEDec model;
  RDOVAEDecState dec;
  int i;
  parse_weights(&list, blob, len);
  if (!list || init_rdovaedec(&model, list) != 0) { free(list); return -1; }
  memset(&dec, 0, sizeof(dec));
  dred_rdovae_dec_init_states(&dec, &model, state);
  for (i = 0; i < 2 * nb_latents; i += 2)
    dred_rdovae_decode_qframe(&dec, &model, &features[2 * i * DRED_NUM_FEATURES], &latents[(i / 2) * DRED_LATENT_DIM]);
  free(list);
  return 0;
}
