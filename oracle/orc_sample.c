/* Oracle: the 16 kHz sample network and the synthesis loop.
 * TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off. */
#include <math.h>
#include <string.h>
#include "orc_internal.h"

/* compute_sparse_gru (nnet.c:410-448), USE_SU_BIAS build; `input` = gru_a_input[3N] */
static void sparse_gru_a(const OrcModel *m, float *state, const float *input) {
  const OrcSparseGRU *g = &m->gru_a;
  enum { N = 384 };
  float recur[3 * N];
  float *z = recur, *r = recur + N, *h = recur + 2 * N;
  const float *bias = &g->subias[3 * N];
  int i, k;
  for (k = 0; k < 2; k++)
    for (i = 0; i < N; i++)
      recur[k * N + i] = bias[k * N + i] + g->diag[k * N + i] * state[i] + input[k * N + i];
  for (; k < 3; k++)
    for (i = 0; i < N; i++)
      recur[k * N + i] = bias[k * N + i] + g->diag[k * N + i] * state[i];
  orc_sparse_sgemv_accum8x4(recur, g->w, 3 * N, N, g->idx, state);
  orc_vec_act(recur, recur, 2 * N, ORC_SIGMOID);
  for (i = 0; i < N; i++) h[i] = h[i] * r[i] + input[2 * N + i];
  orc_vec_act(h, h, N, ORC_TANH);
  for (i = 0; i < N; i++) state[i] = z[i] * state[i] + ((float)1 - z[i]) * h[i];
}

/* compute_gruB (nnet.c:326-372), USE_SU_BIAS build; any size up to 256 units (GRU-B: 16, PLC GRUs: 256) */
void orc_gru_b(const OrcGRU *g, const float *cond, float *state, const float *input) {
  const int N = g->n;
  float zrh[3 * 256], recur[3 * 256];
  float *z = zrh, *r = zrh + N, *h = zrh + 2 * N;
  int i;
  for (i = 0; i < 3 * N; i++) zrh[i] = g->subias[i] + cond[i];
  orc_sparse_sgemv_accum8x4(zrh, g->w_in, 3 * N, g->nin, g->idx_in, input);
  for (i = 0; i < 3 * N; i++) recur[i] = g->subias[3 * N + i];
  orc_sgemv_accum8x4(recur, g->w_rec, 3 * N, N, state);
  for (i = 0; i < 2 * N; i++) zrh[i] += recur[i];
  orc_vec_act(zrh, zrh, 2 * N, ORC_SIGMOID);
  for (i = 0; i < N; i++) h[i] += recur[2 * N + i] * r[i];
  orc_vec_act(h, h, N, ORC_TANH);
  for (i = 0; i < N; i++) h[i] = z[i] * state[i] + ((float)1 - z[i]) * h[i];
  for (i = 0; i < N; i++) state[i] = h[i];
}

/* one dual-FC node logit (nnet.c:192-202) */
static float node_logit(const OrcMDense *l, const float *in, int i) {
  int M = l->nin, N = l->n, stride = l->nin * l->c, j;
  float sum1 = l->bias[i], sum2 = l->bias[i + N];
  for (j = 0; j < M; j++) {
    sum1 += l->w[i * stride + j] * in[j];
    sum2 += l->w[i * stride + j + M] * in[j];
  }
  sum1 = l->factor[i] * orc_tanh_approx(sum1);
  sum2 = l->factor[N + i] * orc_tanh_approx(sum2);
  sum1 += sum2;
  return sum1;
}

/* sample_mdense (nnet.c:163-214) */
static int sample_tree(const OrcModel *m, OrcState *s) {
  float thr[8];
  int b, val = 0;
  for (b = 0; b < 8; b += 4) {
    uint32_t r = orc_kiss99_rand(s->rng);
    thr[b] = m->logit_table[r & 0xFF];
    thr[b + 1] = m->logit_table[(r >> 8) & 0xFF];
    thr[b + 2] = m->logit_table[(r >> 16) & 0xFF];
    thr[b + 3] = m->logit_table[(r >> 24) & 0xFF];
  }
  for (b = 0; b < 8; b++) {
    int i = (1 << b) | val;
    int bit = thr[b] < node_logit(&m->dual_fc, s->gru_b_state, i);
    val = (val << 1) | bit;
  }
  return val;
}

void orc_all_node_logits(const OrcModel *m, const OrcState *s, float *logits) {
  int i;
  logits[0] = 0;
  for (i = 1; i < 256; i++) logits[i] = node_logit(&m->dual_fc, s->gru_b_state, i);
}

/* run_sample_network (lpcnet.c:146-167) + compute_gru_a_input (nnet.c:484-491) */
int orc_sample_network(const OrcModel *m, OrcState *s, int last_exc, int u_sig, int u_pred) {
  enum { N = 384 };
  float gru_a_input[3 * N];
  int i;
  const float *e1 = &m->e_sig.w[u_sig * 3 * N];
  const float *e2 = &m->e_pred.w[u_pred * 3 * N];
  const float *e3 = &m->e_exc.w[last_exc * 3 * N];
  for (i = 0; i < 3 * N; i++) gru_a_input[i] = s->gru_a_condition[i] + e1[i] + e2[i] + e3[i];
  sparse_gru_a(m, s->gru_a_state, gru_a_input);
  orc_gru_b(&m->gru_b, s->gru_b_condition, s->gru_b_state, s->gru_a_state);
  return sample_tree(m, s);
}

/* lpcnet_synthesize_tail_impl (lpcnet.c:235-271) */
void orc_sample_tail(const OrcModel *m, OrcState *s, int16_t *output, int N, int preload) {
  int i, j;
  if (s->frame_count <= m->features_delay) {
    memset(output, 0, N * sizeof(int16_t));
    return;
  }
  for (i = 0; i < N; i++) {
    float pcm, pred = 0;
    int exc, u_sig, u_pred;
    for (j = 0; j < 16; j++) pred -= s->last_sig[j] * s->lpc[j];
    u_sig = orc_lin2ulaw(s->last_sig[0]);
    u_pred = orc_lin2ulaw(pred);
    exc = orc_sample_network(m, s, s->last_exc, u_sig, u_pred);
    if (i < preload) {
      exc = orc_lin2ulaw((float)output[i] - 0.85f * s->deemph_mem - pred);
      pcm = (float)output[i] - 0.85f * s->deemph_mem;
    } else {
      pcm = pred + orc_ulaw2lin((float)exc);
    }
    memmove(&s->last_sig[1], &s->last_sig[0], 15 * sizeof(float));
    s->last_sig[0] = pcm;
    s->last_exc = exc;
    pcm += 0.85f * s->deemph_mem;
    s->deemph_mem = pcm;
    if (pcm < -32767) pcm = -32767;
    if (pcm > 32767) pcm = 32767;
    if (i >= preload) output[i] = (int16_t)(int)floor(.5 + (double)pcm);
  }
}

/* lpcnet_synthesize_impl (lpcnet.c:273-277) */
void orc_synthesize(const OrcModel *m, OrcState *s, const float *features, int16_t *out, int N, int preload) {
  orc_frame_network(m, s, features, s->gru_a_condition, s->gru_b_condition, s->lpc);
  orc_sample_tail(m, s, out, N, preload);
}

static uint32_t fnv1a(const unsigned char *p, int n) {
  uint32_t h = 2166136261u;
  int i;
  for (i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; }
  return h;
}

/* Same contract as ref_run in ref_harness.c. */
void orc_run(const OrcModel *m, OrcState *s, const float *feats, int feat_stride, int n_frames,
             int16_t *pcm, const int16_t *forced, OrcTrace *trace, float *logits) {
  int f, i, j;
  for (f = 0; f < n_frames; f++) {
    int16_t *out = &pcm[f * 160];
    orc_frame_network(m, s, &feats[f * feat_stride], s->gru_a_condition, s->gru_b_condition, s->lpc);
    if (forced) memcpy(out, &forced[f * 160], 160 * sizeof(int16_t));
    for (i = 0; i < 160; i++) {
      int t = f * 160 + i;
      int active = s->frame_count > m->features_delay;
      OrcTrace tr;
      memset(&tr, 0, sizeof(tr));
      if (active && trace) {
        float pred = 0;
        for (j = 0; j < 16; j++) pred -= s->last_sig[j] * s->lpc[j];
        tr.u_sig = orc_lin2ulaw(s->last_sig[0]);
        tr.u_pred = orc_lin2ulaw(pred);
      }
      orc_sample_tail(m, s, &out[i], 1, forced ? 1 : 0);
      if (trace) {
        if (active) {
          unsigned char xa[384], xb[16];
          orc_ps_to_epi8(xa, s->gru_a_state, 384);
          orc_ps_to_epi8(xb, s->gru_b_state, 16);
          tr.exc = s->last_exc;
          tr.rng_z = s->rng[0]; tr.rng_w = s->rng[1];
          tr.xa_hash = fnv1a(xa, 384);
          tr.xb_hash = fnv1a(xb, 16);
        }
        tr.pcm = out[i];
        trace[t] = tr;
      }
      if (logits) {
        if (active) orc_all_node_logits(m, s, &logits[(size_t)t * 256]);
        else memset(&logits[(size_t)t * 256], 0, 256 * sizeof(float));
      }
    }
  }
}
