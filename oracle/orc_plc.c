/* Oracle: packet-loss concealment.
 *  - the non-causal mode (LPCNET_PLC_NONCAUSAL, with or without the DC filter): lpcnet_plc_update_non_causal /
 *    lpcnet_plc_conceal_non_causal / process_queued_update / compute_plc_pred / clear_state
 *    (src/lpcnet_plc.c:349-491, 342-347, 135-145, 175-181), for a model built with FEATURES_DELAY 0
 *    (lpcnet_plc.c:357-361); it never reads the FEC queue;
 *  - the causal and codec modes (LPCNET_PLC_CAUSAL, LPCNET_PLC_CODEC): lpcnet_plc_update_causal /
 *    lpcnet_plc_conceal_causal with the FEC queue, get_fec_or_pred, fec_rewind and the deferred frame network
 *    (src/lpcnet_plc.c:111-173, 185-340; src/lpcnet.c:122-144, 226-233), any FEATURES_DELAY <= 2, PLC_SKIP_UPDATES on.
 * TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off. */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "orc_internal.h"

#define FRAME 160
#define TOFF 80
#define NBANDS 18
#define PLC_IN (2 * NBANDS + ORC_NB_FEATURES + 1)   /* 57 */
#define DC_CONST 0.003

#define MAX_FEC 100
#define MAX_DELAY 2
typedef struct { float gru1[256], gru2[256]; } PlcNet;

struct OrcPLC {
  const OrcModel *m;
  OrcState lp;
  OrcEncState enc;
  int remove_dc, non_causal, enable_blending;
  /* causal / codec modes */
  float fec[MAX_FEC][ORC_NB_FEATURES];
  int fec_keep_pos, fec_read_pos, fec_fill_pos, fec_skip;
  int pcm_fill, skip_analysis, blend;
  PlcNet plc_copy[MAX_DELAY + 1];
  float feature_buffer[4 * ORC_NB_FEATURES];      /* LPCNetState::feature_buffer (lpcnet_private.h:36-38) */
  int feature_buffer_fill;
  int16_t pcm[MAX_DELAY * FRAME + TOFF + FRAME];   /* PLC_BUF_SIZE + FRAME_SIZE */
  float features[36];
  int loss_count;
  float gru1[256], gru2[256];
  double dc_mem, syn_dc;
  int16_t dc_buf[TOFF];
  int queued_update;
  int16_t queued_samples[FRAME];
};


/* lpcnet_plc_create(options) + lpcnet_plc_load_model: options = 1 (non-causal) | 4 (DC filter) */
OrcPLC *orc_plc_new(const OrcModel *m, int options) {
  OrcPLC *p;
  if (!m->has_plc || (options & 3) == 3) return NULL;
  if ((options & 3) == 1 && m->features_delay != 0) return NULL;    /* lpcnet_plc.c:357-361 */
  if (m->features_delay > MAX_DELAY) return NULL;
  p = calloc(1, sizeof(*p));
  p->m = m;
  p->lp.m = m;
  orc_state_reset(&p->lp);
  p->remove_dc = !!(options & 4);
  p->non_causal = (options & 3) == 1;
  p->enable_blending = (options & 3) != 2;
  p->pcm_fill = m->features_delay * FRAME + TOFF;                    /* lpcnet_plc_reset: PLC_BUF_SIZE */
  return p;
}
void orc_plc_free(OrcPLC *p) { free(p); }
void orc_plc_seed(OrcPLC *p, const unsigned char *d, int n) { orc_kiss99_srand(p->lp.rng, d, n); }

/* compute_plc_pred (lpcnet_plc.c:135-145) */
void orc_plc_pred(const OrcModel *m, float *gru1, float *gru2, float *out, const float *in) {
  float zeros[3 * 256] = {0};
  float dense_out[128];
  orc_dense(&m->plc_dense1, dense_out, in);
  orc_gru_b(&m->plc_gru1, zeros, gru1, dense_out);
  orc_gru_b(&m->plc_gru2, zeros, gru2, gru1);
  orc_dense(&m->plc_out, out, gru2);
  out[19] = (.5f < out[19] + .1f) ? .5f : out[19] + .1f;
}

static void analysis(OrcPLC *p, const int16_t *pcm) {
  float x[FRAME], f[36];
  int i;
  for (i = 0; i < FRAME; i++) x[i] = pcm[i];
  orc_enc_single_frame(&p->enc, x, f);
}

static void process_queued_update(OrcPLC *p) {
  if (p->queued_update) {
    orc_synthesize(p->m, &p->lp, p->features, p->queued_samples, FRAME, FRAME);
    p->queued_update = 0;
  }
}

static const float att_table[10] = {0, 0, -.2, -.2, -.4, -.4, -.8, -.8, -1.6, -1.6};

static void attenuate(OrcPLC *p) {
  float v;
  if (p->loss_count >= 10) v = p->features[0] + att_table[9] - 2 * (p->loss_count - 9);
  else v = p->features[0] + att_table[p->loss_count];
  p->features[0] = -10 > v ? -10 : v;
}

static int orc_plc_update_causal(OrcPLC *p, int16_t *pcm);
static int orc_plc_conceal_causal(OrcPLC *p, int16_t *pcm);

/* lpcnet_plc_update_non_causal (lpcnet_plc.c:349-447) */
int orc_plc_update(OrcPLC *p, int16_t *pcm) {
  float x[FRAME], plc_features[PLC_IN];
  int16_t pcm_save[FRAME], lp[FRAME] = {0};
  double mem_bak = 0;
  int delta = (int)p->syn_dc;
  int i;
  if (!p->non_causal) return orc_plc_update_causal(p, pcm);
  process_queued_update(p);
  if (p->remove_dc) {
    p->dc_mem += p->syn_dc;
    p->syn_dc = 0;
    mem_bak = p->dc_mem;
    for (i = 0; i < FRAME; i++) {
      lp[i] = (int16_t)(int)floor(.5 + p->dc_mem);
      p->dc_mem += DC_CONST * (pcm[i] - p->dc_mem);
      pcm[i] -= lp[i];
    }
  }
  memcpy(pcm_save, pcm, sizeof(pcm_save));
  for (i = 0; i < FRAME; i++) x[i] = pcm[i];
  orc_burg_cepstral_analysis(plc_features, x);
  if (p->loss_count > 0) {
    OrcState copy;
    float zeros[PLC_IN] = {0};
    int16_t rev[FRAME];
    memcpy(zeros, plc_features, 2 * NBANDS * sizeof(float));
    zeros[2 * NBANDS + ORC_NB_FEATURES] = 1;
    orc_plc_pred(p->m, p->gru1, p->gru2, p->features, zeros);
    copy = p->lp;
    orc_synthesize(p->m, &p->lp, p->features, &p->pcm[FRAME - TOFF], TOFF, 0);
    if (p->remove_dc) {
      for (i = 0; i < FRAME; i++) pcm[i] += lp[i];
      p->dc_mem = mem_bak;
      for (i = 0; i < TOFF; i++) p->syn_dc += DC_CONST * (p->pcm[FRAME - TOFF + i] - p->syn_dc);
      p->dc_mem += p->syn_dc;
      delta += p->syn_dc;                 /* int += double: converts the sum back to int */
      p->syn_dc = 0;
      for (i = 0; i < FRAME; i++) {
        lp[i] = (int16_t)(int)floor(.5 + p->dc_mem);
        p->dc_mem += DC_CONST * (pcm[i] - p->dc_mem);
        pcm[i] -= lp[i];
      }
      memcpy(pcm_save, pcm, sizeof(pcm_save));
    }
    for (i = 0; i < FRAME; i++) rev[i] = pcm[FRAME - i - 1];
    /* clear_state (lpcnet_plc.c:175-181) */
    memset(p->lp.last_sig, 0, sizeof(p->lp.last_sig));
    p->lp.last_exc = orc_lin2ulaw(0.f);
    p->lp.deemph_mem = 0;
    memset(p->lp.gru_a_state, 0, sizeof(p->lp.gru_a_state));
    memset(p->lp.gru_b_state, 0, sizeof(p->lp.gru_b_state));
    orc_synthesize(p->m, &p->lp, p->features, rev, FRAME, FRAME);
    orc_sample_tail(p->m, &p->lp, rev, TOFF, 0);
    for (i = 0; i < TOFF; i++) {
      float w = (float)(.5 - .5 * cos(M_PI * i / (TOFF)));
      p->pcm[FRAME - 1 - i] = (int16_t)(int)floor(.5 + w * p->pcm[FRAME - 1 - i] + (1 - w) * (rev[i] + delta));
    }
    p->lp = copy;
    p->queued_update = 1;
    memcpy(&p->queued_samples[0], &p->pcm[FRAME - TOFF], TOFF * sizeof(int16_t));
    memcpy(&p->queued_samples[TOFF], pcm, (FRAME - TOFF) * sizeof(int16_t));
    analysis(p, p->pcm);
  }
  analysis(p, pcm);
  if (p->loss_count == 0) {
    memcpy(&plc_features[2 * NBANDS], p->enc.features, ORC_NB_FEATURES * sizeof(float));
    plc_features[2 * NBANDS + ORC_NB_FEATURES] = 1;
    orc_plc_pred(p->m, p->gru1, p->gru2, p->features, plc_features);
    orc_synthesize(p->m, &p->lp, p->enc.features, &p->pcm[FRAME - TOFF], TOFF, TOFF);
    orc_sample_tail(p->m, &p->lp, pcm, FRAME - TOFF, FRAME - TOFF);
  }
  memcpy(&pcm[FRAME - TOFF], pcm, TOFF * sizeof(int16_t));
  memcpy(pcm, &p->pcm[TOFF], (FRAME - TOFF) * sizeof(int16_t));
  memcpy(p->pcm, pcm_save, FRAME * sizeof(int16_t));
  p->loss_count = 0;
  if (p->remove_dc) {
    for (i = 0; i < TOFF; i++) pcm[i] += p->dc_buf[i];
    for (; i < FRAME; i++) pcm[i] += lp[i - TOFF];
    for (i = 0; i < TOFF; i++) p->dc_buf[i] = lp[FRAME - TOFF + i];
  }
  return 0;
}

/* lpcnet_plc_conceal_non_causal (lpcnet_plc.c:449-491) */
int orc_plc_conceal(OrcPLC *p, int16_t *pcm) {
  float zeros[PLC_IN] = {0};
  int i;
  if (!p->non_causal) return orc_plc_conceal_causal(p, pcm);
  process_queued_update(p);
  orc_plc_pred(p->m, p->gru1, p->gru2, p->features, zeros);
  attenuate(p);
  if (p->loss_count == 0) {
    memcpy(pcm, &p->pcm[FRAME - TOFF], TOFF * sizeof(int16_t));
    orc_synthesize(p->m, &p->lp, p->features, &p->pcm[FRAME - TOFF], TOFF, TOFF);
    orc_sample_tail(p->m, &p->lp, &pcm[TOFF], FRAME - TOFF, 0);
  } else {
    orc_synthesize(p->m, &p->lp, p->features, pcm, TOFF, 0);
    orc_sample_tail(p->m, &p->lp, &pcm[TOFF], FRAME - TOFF, 0);
    memcpy(&p->pcm[FRAME - TOFF], pcm, TOFF * sizeof(int16_t));
    analysis(p, p->pcm);
  }
  memcpy(p->pcm, &pcm[TOFF], (FRAME - TOFF) * sizeof(int16_t));
  if (p->remove_dc) {
    int dc = (int)floor(.5 + p->dc_mem);
    if (p->loss_count == 0) {
      for (i = TOFF; i < FRAME; i++) p->syn_dc += DC_CONST * (pcm[i] - p->syn_dc);
    } else {
      for (i = 0; i < FRAME; i++) p->syn_dc += DC_CONST * (pcm[i] - p->syn_dc);
    }
    for (i = 0; i < TOFF; i++) pcm[i] += p->dc_buf[i];
    for (; i < FRAME; i++) pcm[i] += dc;
    for (i = 0; i < TOFF; i++) p->dc_buf[i] = (int16_t)dc;
  }
  p->loss_count++;
  return 0;
}

void orc_plc_peek(const OrcPLC *p, float *features20, float *gru1, float *gru2, int *loss_count, double *dc) {
  memcpy(features20, p->features, 20 * sizeof(float));
  memcpy(gru1, p->gru1, sizeof(p->gru1));
  memcpy(gru2, p->gru2, sizeof(p->gru2));
  *loss_count = p->loss_count;
  dc[0] = p->dc_mem; dc[1] = p->syn_dc;
}

/* ---- causal / codec modes ------------------------------------------------------------------------------------ */

/* lpcnet_plc_fec_add / lpcnet_plc_fec_clear (lpcnet_plc.c:111-132); features NULL = "skip one" */
void orc_plc_fec_add(OrcPLC *p, const float *features) {
  if (features == NULL) { p->fec_skip++; return; }
  if (p->fec_fill_pos == MAX_FEC) {
    if (p->fec_keep_pos == 0) return;                       /* "FEC buffer full" */
    memmove(&p->fec[0][0], &p->fec[p->fec_keep_pos][0], (p->fec_fill_pos - p->fec_keep_pos) * ORC_NB_FEATURES * sizeof(float));
    p->fec_fill_pos = p->fec_fill_pos - p->fec_keep_pos;
    p->fec_read_pos -= p->fec_keep_pos;
    p->fec_keep_pos = 0;
  }
  memcpy(&p->fec[p->fec_fill_pos][0], features, ORC_NB_FEATURES * sizeof(float));
  p->fec_fill_pos++;
}
void orc_plc_fec_clear(OrcPLC *p) { p->fec_keep_pos = p->fec_read_pos = p->fec_fill_pos = p->fec_skip = 0; }

static int imax(int a, int b) { return a > b ? a : b; }

/* get_fec_or_pred (lpcnet_plc.c:147-166) */
static int get_fec_or_pred(OrcPLC *p, float *out) {
  const int D = p->m->features_delay;
  if (p->fec_read_pos != p->fec_fill_pos && p->fec_skip == 0) {
    float plc_features[PLC_IN] = {0}, discard[ORC_NB_FEATURES];
    memcpy(out, &p->fec[p->fec_read_pos][0], ORC_NB_FEATURES * sizeof(float));
    p->fec_read_pos++;
    p->fec_keep_pos = imax(0, imax(p->fec_keep_pos, p->fec_read_pos - D - 1));
    memcpy(&plc_features[2 * NBANDS], out, ORC_NB_FEATURES * sizeof(float));
    plc_features[2 * NBANDS + ORC_NB_FEATURES] = -1;
    orc_plc_pred(p->m, p->gru1, p->gru2, discard, plc_features);
    return 1;
  } else {
    float zeros[PLC_IN] = {0};
    orc_plc_pred(p->m, p->gru1, p->gru2, out, zeros);
    if (p->fec_skip > 0) p->fec_skip--;
    return 0;
  }
}

/* fec_rewind (lpcnet_plc.c:168-173) */
static void fec_rewind(OrcPLC *p, int offset) {
  p->fec_read_pos -= offset;
  if (p->fec_read_pos < p->fec_keep_pos) p->fec_read_pos = p->fec_keep_pos;
}

/* run_frame_network_deferred / run_frame_network_flush (lpcnet.c:122-144); kernel sizes 3 + 3 - 2 = 4 rows */
static void frame_network_deferred(OrcPLC *p, const float *features) {
  const int max_buffer_size = 4;
  if (p->feature_buffer_fill == max_buffer_size)
    memmove(p->feature_buffer, &p->feature_buffer[ORC_NB_FEATURES], (max_buffer_size - 1) * ORC_NB_FEATURES * sizeof(float));
  else
    p->feature_buffer_fill++;
  memcpy(&p->feature_buffer[(p->feature_buffer_fill - 1) * ORC_NB_FEATURES], features, ORC_NB_FEATURES * sizeof(float));
}
static void frame_network_flush(OrcPLC *p) {
  int i;
  for (i = 0; i < p->feature_buffer_fill; i++) {
    float lpc[16], ga[1152], gb[48];
    orc_frame_network(p->m, &p->lp, &p->feature_buffer[i * ORC_NB_FEATURES], ga, gb, lpc);
  }
  p->feature_buffer_fill = 0;
}

static void plc_net_get(const OrcPLC *p, PlcNet *n) { memcpy(n->gru1, p->gru1, sizeof(n->gru1)); memcpy(n->gru2, p->gru2, sizeof(n->gru2)); }
static void plc_net_set(OrcPLC *p, const PlcNet *n) { memcpy(p->gru1, n->gru1, sizeof(n->gru1)); memcpy(p->gru2, n->gru2, sizeof(n->gru2)); }
static void push_plc_copy(OrcPLC *p) {     /* RNN_MOVE(&plc_copy[1], &plc_copy[0], FEATURES_DELAY); plc_copy[0] = plc_net */
  const int D = p->m->features_delay;
  if (D > 0) memmove(&p->plc_copy[1], &p->plc_copy[0], D * sizeof(PlcNet));
  plc_net_get(p, &p->plc_copy[0]);
}

/* lpcnet_plc_update_causal (lpcnet_plc.c:185-286), PLC_SKIP_UPDATES defined */
static int orc_plc_update_causal(OrcPLC *p, int16_t *pcm) {
  const int D = p->m->features_delay, BUF = D * FRAME + TOFF;
  float x[FRAME], plc_features[PLC_IN];
  int16_t lp[FRAME] = {0};
  int i, delta = 0;
  if (p->remove_dc) {
    p->dc_mem += p->syn_dc;
    delta = (int)p->syn_dc;
    p->syn_dc = 0;
    for (i = 0; i < FRAME; i++) {
      lp[i] = (int16_t)(int)floor(.5 + p->dc_mem);
      p->dc_mem += DC_CONST * (pcm[i] - p->dc_mem);
      pcm[i] -= lp[i];
    }
  }
  for (i = 0; i < FRAME; i++) x[i] = pcm[i];
  orc_burg_cepstral_analysis(plc_features, x);
  if (p->skip_analysis) {
    if (p->blend) {
      int16_t tmp[FRAME - TOFF];
      float zeros[PLC_IN] = {0};
      memcpy(zeros, plc_features, 2 * NBANDS * sizeof(float));
      zeros[2 * NBANDS + ORC_NB_FEATURES] = 1;
      if (p->enable_blending) {
        OrcState copy;
        plc_net_set(p, &p->plc_copy[D]);
        orc_plc_pred(p->m, p->gru1, p->gru2, p->features, zeros);
        for (i = 0; i < D; i++) frame_network_deferred(p, p->features);
        copy = p->lp;
        orc_synthesize(p->m, &p->lp, p->features, tmp, FRAME - TOFF, 0);
        for (i = 0; i < FRAME - TOFF; i++) {
          float w = (float)(.5 - .5 * cos(M_PI * i / (FRAME - TOFF)));
          pcm[i] = (int16_t)(int)floor(.5 + w * pcm[i] + (1 - w) * (tmp[i] - delta));
        }
        p->lp = copy;
        orc_synthesize(p->m, &p->lp, p->features, pcm, FRAME - TOFF, FRAME - TOFF);
      } else {
        if (D > 0) plc_net_set(p, &p->plc_copy[D - 1]);
        fec_rewind(p, D);
        /* lpcnet_reset_signal (lpcnet.c:226-233) */
        p->lp.deemph_mem = 0;
        p->lp.last_exc = orc_lin2ulaw(0.f);
        memset(p->lp.last_sig, 0, sizeof(p->lp.last_sig));
        memset(p->lp.gru_a_state, 0, sizeof(p->lp.gru_a_state));
        memset(p->lp.gru_b_state, 0, sizeof(p->lp.gru_b_state));
      }
      memcpy(p->pcm, &pcm[FRAME - TOFF], TOFF * sizeof(int16_t));
      p->pcm_fill = TOFF;
    } else {
      memcpy(&p->pcm[p->pcm_fill], pcm, FRAME * sizeof(int16_t));
      p->pcm_fill += FRAME;
    }
  }
  analysis(p, pcm);
  if (!p->blend) {
    memcpy(&plc_features[2 * NBANDS], p->enc.features, ORC_NB_FEATURES * sizeof(float));
    plc_features[2 * NBANDS + ORC_NB_FEATURES] = 1;
    orc_plc_pred(p->m, p->gru1, p->gru2, p->features, plc_features);
    if (p->fec_skip) p->fec_skip--;
    else if (p->fec_read_pos < p->fec_fill_pos) p->fec_read_pos++;
    p->fec_keep_pos = imax(0, imax(p->fec_keep_pos, p->fec_read_pos - D - 1));
  }
  if (p->skip_analysis) {
    if (p->enable_blending) frame_network_deferred(p, p->enc.features);
    p->skip_analysis--;
  } else {
    for (i = 0; i < FRAME; i++) p->pcm[BUF + i] = pcm[i];
    frame_network_deferred(p, p->enc.features);
    memmove(p->pcm, &p->pcm[FRAME], BUF * sizeof(int16_t));
  }
  p->loss_count = 0;
  if (p->remove_dc) for (i = 0; i < FRAME; i++) pcm[i] += lp[i];
  p->blend = 0;
  return 0;
}

/* lpcnet_plc_conceal_causal (lpcnet_plc.c:289-340) */
static int orc_plc_conceal_causal(OrcPLC *p, int16_t *pcm) {
  const int D = p->m->features_delay, BUF = D * FRAME + TOFF;
  int16_t output[FRAME];
  int i;
  frame_network_flush(p);
  while (p->pcm_fill > 0) {
    const int update_count = p->pcm_fill < FRAME ? p->pcm_fill : FRAME;
    memcpy(output, p->pcm, update_count * sizeof(int16_t));
    push_plc_copy(p);
    get_fec_or_pred(p, p->features);
    orc_synthesize(p->m, &p->lp, p->features, output, update_count, update_count);
    memmove(p->pcm, &p->pcm[FRAME], BUF * sizeof(int16_t));
    p->pcm_fill -= update_count;
    p->skip_analysis++;
  }
  push_plc_copy(p);
  orc_sample_tail(p->m, &p->lp, pcm, FRAME - TOFF, 0);
  if (get_fec_or_pred(p, p->features)) p->loss_count = 0;
  else p->loss_count++;
  attenuate(p);
  orc_synthesize(p->m, &p->lp, p->features, &pcm[FRAME - TOFF], TOFF, 0);
  analysis(p, pcm);
  p->blend = 1;
  if (p->remove_dc) {
    for (i = 0; i < FRAME; i++) {
      p->syn_dc += DC_CONST * (pcm[i] - p->syn_dc);
      pcm[i] += (int16_t)(int)floor(.5 + p->dc_mem);
    }
  }
  return 0;
}
