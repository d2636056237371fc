/* Oracle: weight-blob parsing and model/state set-up.  TEST INFRASTRUCTURE ONLY.
 * Follows src/parse_lpcnet_weights.c:36-221 (record validation, lookup by name and
 * exact size, idx stream validation) and the init call list of the generated
 * nnet_data.c (training_tf2/dump_lpcnet.py). */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include "orc_internal.h"

typedef struct { const char *name; int size; const void *data; } Arr;

static int parse(const unsigned char *d, int len, Arr *arr, int cap) {
  int n = 0;
  while (len > 0) {
    int32_t size, block;
    const char *name;
    if (len < 64) return -1;
    memcpy(&size, d + 12, 4);
    memcpy(&block, d + 16, 4);
    name = (const char *)d + 20;
    if (block < size || block > len - 64 || name[43] != 0 || size <= 0) return -1;
    if (n >= cap) return -1;
    arr[n].name = name; arr[n].size = size; arr[n].data = d + 64;
    n++;
    d += block + 64; len -= block + 64;
  }
  return n;
}

static const void *find(const Arr *a, int n, const char *name, int size) {
  int i;
  for (i = 0; i < n; i++)
    if (strcmp(a[i].name, name) == 0) return a[i].size == size ? a[i].data : NULL;
  return NULL;
}

/* find_idx_check (parse_lpcnet_weights.c:90-113) */
static const int *find_idx(const Arr *a, int n, const char *name, int nb_in, int nb_out, int *total) {
  int i, remain;
  const int *idx, *base;
  *total = 0;
  for (i = 0; i < n; i++) if (strcmp(a[i].name, name) == 0) break;
  if (i == n) return NULL;
  base = idx = a[i].data;
  remain = a[i].size / 4;
  while (remain > 0) {
    int nb = *idx++, k;
    if (remain < nb + 1) return NULL;
    for (k = 0; k < nb; k++) { int pos = *idx++; if (pos + 3 >= nb_in || (pos & 3)) return NULL; }
    nb_out -= 8; remain -= nb + 1; *total += nb;
  }
  return nb_out == 0 ? base : NULL;
}

static int dense(OrcDense *l, const Arr *a, int n, const char *b, const char *w, int nin, int nout, int act) {
  l->bias = find(a, n, b, nout * 4); l->w = find(a, n, w, nin * nout * 4);
  l->nin = nin; l->nout = nout; l->act = act;
  return !l->bias || !l->w;
}
static int conv(OrcConv *l, const Arr *a, int n, const char *b, const char *w, int nin, int k, int nout) {
  l->bias = find(a, n, b, nout * 4); l->w = find(a, n, w, k * nin * nout * 4);
  l->nin = nin; l->k = k; l->nout = nout; l->act = ORC_TANH;
  return !l->bias || !l->w;
}
static int embed(OrcEmbed *l, const Arr *a, int n, const char *w, int rows, int dim) {
  l->w = find(a, n, w, rows * dim * 4); l->n = rows; l->dim = dim;
  return !l->w;
}

OrcModel *orc_model_load(const unsigned char *blob, int len, int features_delay, float lpc_gamma,
                         const uint32_t *rcp_table) {
  enum { N = 384, NB = 16, C = 128 };
  Arr arr[192];
  int n, i, bad = 0;
  OrcModel *m;
  if (features_delay < 0 || features_delay > ORC_MAX_DELAY) return NULL;
  m = calloc(1, sizeof(*m));
  m->blob = malloc(len); memcpy(m->blob, blob, len); m->len = len;
  m->features_delay = features_delay; m->lpc_gamma = lpc_gamma;
  n = parse(m->blob, len, arr, 192);
  if (n < 0) { orc_model_free(m); return NULL; }
  bad |= embed(&m->e_sig, arr, n, "gru_a_embed_sig_weights", 256, 3 * N);
  bad |= embed(&m->e_pred, arr, n, "gru_a_embed_pred_weights", 256, 3 * N);
  bad |= embed(&m->e_exc, arr, n, "gru_a_embed_exc_weights", 256, 3 * N);
  bad |= dense(&m->a_feat, arr, n, "gru_a_dense_feature_bias", "gru_a_dense_feature_weights", C, 3 * N, ORC_LINEAR);
  bad |= dense(&m->b_feat, arr, n, "gru_b_dense_feature_bias", "gru_b_dense_feature_weights", C, 3 * NB, ORC_LINEAR);
  /* gru_init (parse_lpcnet_weights.c:146-170) */
  m->gru_b.bias = find(arr, n, "gru_b_bias", 6 * NB * 4);
  m->gru_b.subias = find(arr, n, "gru_b_subias", 6 * NB * 4);
  m->gru_b.idx_in = find_idx(arr, n, "gru_b_weights_idx", N, 3 * NB, &m->gru_b.nblocks);
  m->gru_b.w_in = find(arr, n, "gru_b_weights", 32 * m->gru_b.nblocks);
  m->gru_b.w_rec = find(arr, n, "gru_b_recurrent_weights", 3 * NB * NB);
  m->gru_b.nin = N; m->gru_b.n = NB;
  bad |= !m->gru_b.bias || !m->gru_b.subias || !m->gru_b.idx_in || !m->gru_b.w_in || !m->gru_b.w_rec;
  bad |= conv(&m->conv1, arr, n, "feature_conv1_bias", "feature_conv1_weights", 84, 3, C);
  bad |= conv(&m->conv2, arr, n, "feature_conv2_bias", "feature_conv2_weights", C, 3, C);
  bad |= embed(&m->e_pitch, arr, n, "embed_pitch_weights", 256, 64);
  bad |= find(arr, n, "embed_sig_weights", 256 * 128 * 4) == NULL;
  bad |= dense(&m->dense1, arr, n, "feature_dense1_bias", "feature_dense1_weights", C, C, ORC_TANH);
  bad |= dense(&m->dense2, arr, n, "feature_dense2_bias", "feature_dense2_weights", C, C, ORC_TANH);
  m->dual_fc.bias = find(arr, n, "dual_fc_bias", 2 * 256 * 4);
  m->dual_fc.w = find(arr, n, "dual_fc_weights", NB * 2 * 256 * 4);
  m->dual_fc.factor = find(arr, n, "dual_fc_factor", 2 * 256 * 4);
  m->dual_fc.nin = NB; m->dual_fc.n = 256; m->dual_fc.c = 2;
  bad |= !m->dual_fc.bias || !m->dual_fc.w || !m->dual_fc.factor;
  /* sparse_gru_init (parse_lpcnet_weights.c:172-192) */
  m->gru_a.bias = find(arr, n, "sparse_gru_a_bias", 6 * N * 4);
  m->gru_a.subias = find(arr, n, "sparse_gru_a_subias", 6 * N * 4);
  m->gru_a.diag = find(arr, n, "sparse_gru_a_recurrent_weights_diag", 3 * N * 4);
  m->gru_a.idx = find_idx(arr, n, "sparse_gru_a_recurrent_weights_idx", N, 3 * N, &m->gru_a.nblocks);
  m->gru_a.w = find(arr, n, "sparse_gru_a_recurrent_weights", 32 * m->gru_a.nblocks);
  m->gru_a.n = N;
  bad |= !m->gru_a.bias || !m->gru_a.subias || !m->gru_a.diag || !m->gru_a.idx || !m->gru_a.w;
  if (bad) { orc_model_free(m); return NULL; }
  /* init_plc_model (plc_data.c stand-in, oracle/gen/plc_data.c): optional */
  if (find(arr, n, "plc_dense1_bias", 128 * 4)) {
    OrcGRU *gs[2] = {&m->plc_gru1, &m->plc_gru2};
    const char *nm[2][5] = {{"plc_gru1_bias", "plc_gru1_subias", "plc_gru1_weights_idx", "plc_gru1_weights", "plc_gru1_recurrent_weights"},
                            {"plc_gru2_bias", "plc_gru2_subias", "plc_gru2_weights_idx", "plc_gru2_weights", "plc_gru2_recurrent_weights"}};
    const int nin[2] = {128, 256};
    int k, pbad = 0;
    pbad |= dense(&m->plc_dense1, arr, n, "plc_dense1_bias", "plc_dense1_weights", 57, 128, ORC_TANH);
    pbad |= dense(&m->plc_out, arr, n, "plc_out_bias", "plc_out_weights", 256, 20, ORC_LINEAR);
    for (k = 0; k < 2; k++) {
      OrcGRU *g = gs[k];
      g->bias = find(arr, n, nm[k][0], 6 * 256 * 4);
      g->subias = find(arr, n, nm[k][1], 6 * 256 * 4);
      g->idx_in = find_idx(arr, n, nm[k][2], nin[k], 3 * 256, &g->nblocks);
      g->w_in = find(arr, n, nm[k][3], 32 * g->nblocks);
      g->w_rec = find(arr, n, nm[k][4], 3 * 256 * 256);
      g->nin = nin[k]; g->n = 256;
      pbad |= !g->bias || !g->subias || !g->idx_in || !g->w_in || !g->w_rec;
    }
    if (pbad) { orc_model_free(m); return NULL; }
    m->has_plc = 1;
  }
  /* init_rdovaedec (dred_rdovae_dec_data.c stand-in, oracle/gen/dred_rdovae_dec_data.c): optional */
  if (find(arr, n, "dec_dense1_bias", 256 * 4)) {
    static const char *dn[5] = {"dec_dense1", "dec_dense3", "dec_dense5", "dec_dense7", "dec_dense8"};
    static const char *gn[3] = {"dec_dense2", "dec_dense4", "dec_dense6"};
    char a[64], b[64];
    int k, dbad = 0;
    for (k = 0; k < 3; k++) {
      snprintf(a, sizeof(a), "state%d_bias", k + 1); snprintf(b, sizeof(b), "state%d_weights", k + 1);
      dbad |= dense(&m->dred_state[k], arr, n, a, b, 24, 256, ORC_TANH);
    }
    for (k = 0; k < 5; k++) {
      snprintf(a, sizeof(a), "%s_bias", dn[k]); snprintf(b, sizeof(b), "%s_weights", dn[k]);
      dbad |= dense(&m->dred_dense[k], arr, n, a, b, k == 0 ? 80 : 256, 256, ORC_TANH);
    }
    dbad |= dense(&m->dred_final, arr, n, "dec_final_bias", "dec_final_weights", 2048, 80, ORC_LINEAR);
    for (k = 0; k < 3; k++) {
      OrcGRU *g = &m->dred_gru[k];
      snprintf(a, sizeof(a), "%s_bias", gn[k]); g->bias = find(arr, n, a, 6 * 256 * 4);
      snprintf(a, sizeof(a), "%s_subias", gn[k]); g->subias = find(arr, n, a, 6 * 256 * 4);
      snprintf(a, sizeof(a), "%s_weights_idx", gn[k]); g->idx_in = find_idx(arr, n, a, 256, 3 * 256, &g->nblocks);
      snprintf(a, sizeof(a), "%s_weights", gn[k]); g->w_in = find(arr, n, a, 32 * g->nblocks);
      snprintf(a, sizeof(a), "%s_recurrent_weights", gn[k]); g->w_rec = find(arr, n, a, 3 * 256 * 256);
      g->nin = 256; g->n = 256;
      dbad |= !g->bias || !g->subias || !g->idx_in || !g->w_in || !g->w_rec;
    }
    if (dbad) { orc_model_free(m); return NULL; }
    m->has_dred = 1;
  }
  /* sampling_logit_table (lpcnet.c:188-191): float prob, double log */
  for (i = 0; i < 256; i++) {
    float prob = .025f + .95f * (float)i / 255.f;
    m->logit_table[i] = (float)(-log((double)(((float)1 - prob) / prob)));
  }
  orc_set_rcp_table(rcp_table);
  return m;
}

void orc_model_free(OrcModel *m) { if (m) { free(m->blob); free(m); } }
int orc_model_nblocks_a(const OrcModel *m) { return m->gru_a.nblocks; }
int orc_model_nblocks_b(const OrcModel *m) { return m->gru_b.nblocks; }

/* lpcnet_reset (lpcnet.c:174-182) */
void orc_state_reset(OrcState *s) {
  const OrcModel *m = s->m;
  memset(s, 0, sizeof(*s));
  s->m = m;
  s->last_exc = orc_lin2ulaw(0.f);
  orc_kiss99_srand(s->rng, (const unsigned char *)"LPCNet", 6);
}

OrcState *orc_state_new(const OrcModel *m) {
  OrcState *s = calloc(1, sizeof(*s));
  s->m = m;
  orc_state_reset(s);
  return s;
}
void orc_state_free(OrcState *s) { free(s); }
void orc_seed(OrcState *s, const unsigned char *d, int n) { orc_kiss99_srand(s->rng, d, n); }

void orc_get(const OrcState *s, OrcSnapshot *p) {
  memset(p, 0, sizeof(*p));
  memcpy(p->conv1_state, s->conv1_state, sizeof(p->conv1_state));
  memcpy(p->conv2_state, s->conv2_state, sizeof(p->conv2_state));
  memcpy(p->gru_a_state, s->gru_a_state, sizeof(p->gru_a_state));
  memcpy(p->gru_b_state, s->gru_b_state, sizeof(p->gru_b_state));
  memcpy(p->last_sig, s->last_sig, sizeof(p->last_sig));
  memcpy(p->lpc, s->lpc, sizeof(p->lpc));
  memcpy(p->old_lpc, s->old_lpc, sizeof(float) * 16 * s->m->features_delay);
  memcpy(p->gru_a_condition, s->gru_a_condition, sizeof(p->gru_a_condition));
  memcpy(p->gru_b_condition, s->gru_b_condition, sizeof(p->gru_b_condition));
  p->deemph_mem = s->deemph_mem; p->last_exc = s->last_exc; p->frame_count = s->frame_count;
  memcpy(p->rng, s->rng, sizeof(p->rng));
}

void orc_set(OrcState *s, const OrcSnapshot *p) {
  memcpy(s->conv1_state, p->conv1_state, sizeof(p->conv1_state));
  memcpy(s->conv2_state, p->conv2_state, sizeof(p->conv2_state));
  memcpy(s->gru_a_state, p->gru_a_state, sizeof(p->gru_a_state));
  memcpy(s->gru_b_state, p->gru_b_state, sizeof(p->gru_b_state));
  memcpy(s->last_sig, p->last_sig, sizeof(p->last_sig));
  memcpy(s->lpc, p->lpc, sizeof(p->lpc));
  memcpy(s->old_lpc, p->old_lpc, sizeof(float) * 16 * s->m->features_delay);
  memcpy(s->gru_a_condition, p->gru_a_condition, sizeof(p->gru_a_condition));
  memcpy(s->gru_b_condition, p->gru_b_condition, sizeof(p->gru_b_condition));
  s->deemph_mem = p->deemph_mem; s->last_exc = p->last_exc; s->frame_count = p->frame_count;
  memcpy(s->rng, p->rng, sizeof(p->rng));
}
