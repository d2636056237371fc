/* Oracle: the 1.6 kb/s encoder (the wire format's other half): lpcnet_encode and
 * lpcnet_compute_features = 4 x (preemphasis + compute_frame_features) + process_superframe
 * (src/lpcnet_enc.c:882-911, 579-743) with the 3-stage M-best VQ, the interpolation search and the bit packer
 * (lpcnet_enc.c:53-463).  TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off: every C operator below is
 * one IEEE operation, in the reference's order.  Float-to-int conversions of NaN / out-of-range values are left
 * to the host CPU on purpose (x86 cvttsd2si gives INT_MIN, as in the reference build this is pinned against). */
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "orc_internal.h"

#define FRAME 160
#define NBANDS 18
#define NB1 17
#define PMAX 256
#define PMIN 32
#define NTOT 36
#define SURVIVORS 5
#define FORBIDDEN_INTERP 7

static float max16(float a, float b) { return a > b ? a : b; }
static float min16(float a, float b) { return a < b ? a : b; }
static int imax(int a, int b) { return a > b ? a : b; }
static int imin(int a, int b) { return a < b ? a : b; }

/* vq_quantize_mbest (lpcnet_enc.c:53-78) */
static void vq_mbest(const float *cb, int n, const float *x, int ndim, int mbest, float *dist, int *index) {
  int i, j;
  for (i = 0; i < mbest; i++) { dist[i] = 1e15f; index[i] = 0; }
  for (i = 0; i < n; i++) {
    float d = 0;
    for (j = 0; j < ndim; j++) d += (x[j] - cb[i * ndim + j]) * (x[j] - cb[i * ndim + j]);
    if (d < dist[mbest - 1]) {
      int pos;
      for (j = 0; j < mbest - 1; j++) if (d < dist[j]) break;
      pos = j;
      for (j = mbest - 1; j >= pos + 1; j--) { dist[j] = dist[j - 1]; index[j] = index[j - 1]; }
      dist[pos] = d;
      index[pos] = i;
    }
  }
}

/* the survivor merge of quantize_3stage_mbest (lpcnet_enc.c:160-178, 200-220): nidx leading indices per survivor */
static void merge_survivors(float *glob_dist, int (*glob)[3], const float *curr_dist, const int *curr_index,
                            const int *lead, int nlead) {
  int pos, m = 0, j, c;
  for (pos = 0; pos < SURVIVORS; pos++) {
    if (curr_dist[m] < glob_dist[pos]) {
      for (j = SURVIVORS - 1; j >= pos + 1; j--) {
        glob_dist[j] = glob_dist[j - 1];
        for (c = 0; c <= nlead; c++) glob[j][c] = glob[j - 1][c];
      }
      glob_dist[pos] = curr_dist[m];
      for (c = 0; c < nlead; c++) glob[pos][c] = lead[c];
      glob[pos][nlead] = curr_index[m];
      m++;
    }
  }
}

/* quantize_3stage_mbest (lpcnet_enc.c:134-248) */
static void quantize_3stage_mbest(float *x, int entry[3], const float *cb1, const float *cb2, const float *cb3) {
  int i, k, m;
  int curr_index[SURVIVORS], index1[SURVIVORS][3], index2[SURVIVORS][3], index3[SURVIVORS][3];
  float curr_dist[SURVIVORS], glob_dist[SURVIVORS];
  vq_mbest(cb1, 1024, x, NB1, SURVIVORS, curr_dist, curr_index);
  for (k = 0; k < SURVIVORS; k++) index1[k][0] = curr_index[k];
  for (k = 0; k < SURVIVORS; k++) {
    float diff[NB1];
    for (i = 0; i < NB1; i++) diff[i] = x[i] - cb1[index1[k][0] * NB1 + i];
    vq_mbest(cb2, 1024, diff, NB1, SURVIVORS, curr_dist, curr_index);
    if (k == 0) {
      for (m = 0; m < SURVIVORS; m++) { index2[m][0] = index1[k][0]; index2[m][1] = curr_index[m]; glob_dist[m] = curr_dist[m]; }
    } else if (curr_dist[0] < glob_dist[SURVIVORS - 1]) {
      merge_survivors(glob_dist, index2, curr_dist, curr_index, index1[k], 1);
    }
  }
  for (k = 0; k < SURVIVORS; k++) {
    float diff[NB1];
    for (i = 0; i < NB1; i++) diff[i] = x[i] - cb1[index2[k][0] * NB1 + i] - cb2[index2[k][1] * NB1 + i];
    vq_mbest(cb3, 1024, diff, NB1, SURVIVORS, curr_dist, curr_index);
    if (k == 0) {
      for (m = 0; m < SURVIVORS; m++) {
        index3[m][0] = index2[k][0]; index3[m][1] = index2[k][1]; index3[m][2] = curr_index[m];
        glob_dist[m] = curr_dist[m];
      }
    } else if (curr_dist[0] < glob_dist[SURVIVORS - 1]) {
      merge_survivors(glob_dist, index3, curr_dist, curr_index, index2[k], 2);
    }
  }
  entry[0] = index3[0][0]; entry[1] = index3[0][1]; entry[2] = index3[0][2];
  for (i = 0; i < NB1; i++)
    x[i] = cb1[entry[0] * NB1 + i] + cb2[entry[1] * NB1 + i] + cb3[entry[2] * NB1 + i];
}

/* quantize_diff over find_nearest_multi (lpcnet_enc.c:250-321), bits = 12, sign = 1 */
static void quantize_diff(float *x, const float *left, const float *right, const float *cb, int *entry) {
  float pred[4 * NBANDS], target[4 * NBANDS], min_dist = 1e15f, s = 1;
  int i, j, nearest = 0, id;
  const int nb_entries = 1 << 12;
  for (i = 0; i < NBANDS; i++) pred[i] = pred[NBANDS + i] = .5f * (left[i] + right[i]);
  for (i = 0; i < NBANDS; i++) pred[2 * NBANDS + i] = left[i];
  for (i = 0; i < NBANDS; i++) pred[3 * NBANDS + i] = right[i];
  for (i = 0; i < 4 * NBANDS; i++) target[i] = x[i % NBANDS] - pred[i];
  for (i = 0; i < nb_entries; i++) {
    const int offset = (i & 3) * NBANDS;
    float dist = 0;
    for (j = 0; j < NBANDS; j++) dist += (target[offset + j] - cb[i * NBANDS + j]) * (target[offset + j] - cb[i * NBANDS + j]);
    if (dist < min_dist) { min_dist = dist; nearest = i; }
  }
  for (i = 0; i < nb_entries; i++) {
    const int offset = (i & 3) * NBANDS;
    float dist = 0;
    for (j = 0; j < NBANDS; j++) dist += (target[offset + j] + cb[i * NBANDS + j]) * (target[offset + j] + cb[i * NBANDS + j]);
    if (dist < min_dist) { min_dist = dist; nearest = i + nb_entries; }
  }
  id = nearest;
  *entry = id;
  if (id >= nb_entries) { s = -1; id -= nb_entries; }
  for (i = 0; i < NBANDS; i++) x[i] = pred[(id & 3) * NBANDS + i] + s * cb[id * NBANDS + i];
}

/* interp_search (lpcnet_enc.c:323-341) */
static void interp_search(const float *x, const float *left, const float *right, float *dist_out) {
  float pred[4 * NBANDS];
  int i, k;
  for (i = 0; i < NBANDS; i++) pred[i] = pred[NBANDS + i] = .5f * (left[i] + right[i]);
  for (i = 0; i < NBANDS; i++) pred[2 * NBANDS + i] = left[i];
  for (i = 0; i < NBANDS; i++) pred[3 * NBANDS + i] = right[i];
  for (k = 1; k < 4; k++) {
    float dist = 0;
    for (i = 0; i < NBANDS; i++) dist += (x[i] - pred[k * NBANDS + i]) * (x[i] - pred[k * NBANDS + i]);
    dist_out[k - 1] = dist;
  }
}

/* double_interp_search (lpcnet_enc.c:378-400) */
static int double_interp_search(float (*F)[NTOT], const float *mem) {
  float dist[2][3], min_dist = 1e15f;
  int i, j, best_id = 0;
  interp_search(F[0], mem, F[1], dist[0]);
  interp_search(F[2], F[1], F[3], dist[1]);
  for (i = 0; i < 3; i++)
    for (j = 0; j < 3; j++) {
      const int id = 3 * i + j;
      const float d = dist[0][i] + dist[1][j];
      if (d < min_dist && id != FORBIDDEN_INTERP) { min_dist = d; best_id = id; }
    }
  return best_id - (best_id >= FORBIDDEN_INTERP);
}

/* single_interp / perform_double_interp (common.c:37-65) */
static void single_interp(float *x, const float *left, const float *right, int id) {
  int i;
  for (i = 0; i < NBANDS; i++) x[i] = id == 0 ? .5f * (left[i] + right[i]) : (id == 1 ? left[i] : right[i]);
}

/* bits_pack (lpcnet_enc.c:441-463): MSB first */
typedef struct { int byte_pos, bit_pos; unsigned char *chars; } Packer;
static void bits_pack(Packer *b, unsigned int data, int nb_bits) {
  while (nb_bits) {
    const int bit = (data >> (nb_bits - 1)) & 1;
    if (b->byte_pos == 8) return;
    b->chars[b->byte_pos] |= bit << (8 - 1 - b->bit_pos);
    if (++b->bit_pos == 8) { b->bit_pos = 0; b->byte_pos++; }
    nb_bits--;
  }
}

/* process_superframe (lpcnet_enc.c:579-743).  xc/fw: rows 2..9 of the reference's xc / frame_weight. */
static void process_superframe(OrcEncState *st, float (*xc)[PMAX + 1], float *fw, float (*F)[NTOT], unsigned char *buf,
                               int encode, int quantize, const float *codebooks) {
  int prev_local[8][PMAX];
  int best[10], i, j, sub, best_i, voiced, main_pitch, modulation, corr_id = 0;
  int c0_id = 0, vq_end[3] = {0, 0, 0}, vq_mid = 0, interp_id = 0;
  float best_a = 0, best_b = 0, w, sx = 0, sxx = 0, sxy = 0, sy = 0, sw = 0, frame_corr, center_pitch;
  float frame_weight_sum = 1e-15f;
  for (sub = 0; sub < 8; sub++) frame_weight_sum += fw[sub];
  for (sub = 0; sub < 8; sub++) fw[sub] *= (8.f / frame_weight_sum);
  for (sub = 0; sub < 8; sub++) {
    float max_path_all = -1e15f;
    float *c = xc[sub];
    best_i = 0;
    for (i = 0; i < PMAX - 2 * PMIN; i++) {
      float xc_half = max16(max16(c[(PMAX + i) / 2], c[(PMAX + i + 2) / 2]), c[(PMAX + i - 1) / 2]);
      if (c[i] < xc_half * 1.1f) c[i] *= .8f;
    }
    for (i = 0; i < PMAX - PMIN; i++) {
      float max_prev = st->pitch_max_path_all - 6.f;
      prev_local[sub][i] = st->best_i;
      for (j = (-4 > -i ? -4 : -i); j <= 4 && i + j < PMAX - PMIN; j++) {
        float cand = st->pitch_max_path[0][i + j] - .02f * abs(j) * abs(j);
        if (cand > max_prev) { max_prev = cand; prev_local[sub][i] = i + j; }
      }
      st->pitch_max_path[1][i] = max_prev + fw[sub] * c[i];
      if (st->pitch_max_path[1][i] > max_path_all) { max_path_all = st->pitch_max_path[1][i]; best_i = i; }
    }
    for (i = 0; i < PMAX - PMIN; i++) st->pitch_max_path[1][i] -= max_path_all;
    memcpy(st->pitch_max_path[0], st->pitch_max_path[1], PMAX * sizeof(float));
    st->pitch_max_path_all = max_path_all;
    st->best_i = best_i;
  }
  best_i = st->best_i;
  frame_corr = 0;
  for (sub = 7; sub >= 0; sub--) {
    best[2 + sub] = PMAX - best_i;
    frame_corr += fw[sub] * xc[sub][best_i];
    best_i = prev_local[sub][best_i];
  }
  frame_corr /= 8;
  if (quantize && frame_corr < 0) frame_corr = 0;
  for (sub = 2; sub < 10; sub++) {
    w = fw[sub - 2];
    sw += w;
    sx += w * sub;
    sxx += w * sub * sub;
    sxy += w * sub * best[sub];
    sy += w * best[sub];
  }
  voiced = frame_corr >= .3;
  best_a = (sw * sxy - sx * sy) / (sw * sxx - sx * sx);
  if (voiced) {
    float mean_pitch = sy / sw;
    float max_a = mean_pitch / 32;
    best_a = min16(max_a, max16(-max_a, best_a));
    corr_id = (int)floor((frame_corr - .3f) / .175f);
    if (quantize) frame_corr = 0.3875f + .175f * corr_id;
  } else {
    best_a = 0;
    corr_id = (int)floor(frame_corr / .075f);
    if (quantize) frame_corr = 0.0375f + .075f * corr_id;
  }
  best_b = (sy - best_a * sx) / sw;
  center_pitch = best_b + 5.5f * best_a;
  main_pitch = (int)floor(.5 + 21. * 1.442695041 * log(center_pitch / PMIN));
  main_pitch = imax(0, imin(63, main_pitch));
  modulation = (int)floor(.5 + 16 * 7 * best_a / center_pitch);
  modulation = imax(-3, imin(3, modulation));
  for (sub = 0; sub < 4; sub++) {
    if (quantize) {
      float p = (float)(pow(2.f, main_pitch / 21.) * PMIN);
      p *= 1.f + modulation / 16.f / 7.f * (2 * sub - 3);
      p = min16(255, max16(33, p));
      F[sub][NBANDS] = .02f * (p - 100);
      F[sub][NBANDS + 1] = frame_corr - .5f;
    } else {
      F[sub][NBANDS] = .01f * (imax(66, imin(510, best[2 + 2 * sub] + best[2 + 2 * sub + 1])) - 200);
      F[sub][NBANDS + 1] = frame_corr - .5f;
    }
  }
  if (quantize) {
    const float *cb1 = codebooks, *cb2 = cb1 + 1024 * NB1, *cb3 = cb2 + 1024 * NB1, *diff4 = cb3 + 1024 * NB1;
    c0_id = (int)floor(.5 + F[3][0] * 4);
    c0_id = imax(-64, imin(63, c0_id));
    F[3][0] = c0_id / 4.f;
    quantize_3stage_mbest(&F[3][1], vq_end, cb1, cb2, cb3);
    quantize_diff(&F[1][0], st->vq_mem, &F[3][0], diff4, &vq_mid);
    interp_id = double_interp_search(F, st->vq_mem);
    {
      int id = interp_id + (interp_id >= FORBIDDEN_INTERP);
      single_interp(F[0], st->vq_mem, F[1], id / 3);
      single_interp(F[2], F[1], F[3], id % 3);
    }
  }
  for (sub = 0; sub < 4; sub++) {
    orc_lpc_from_cepstrum(st->lpc, F[sub]);
    for (i = 0; i < ORC_LPC_ORDER; i++) F[sub][NBANDS + 2 + i] = st->lpc[i];
  }
  memcpy(st->vq_mem, F[3], NBANDS * sizeof(float));
  if (encode) {
    Packer bits = {0, 0, buf};
    memset(buf, 0, 8);
    bits_pack(&bits, c0_id + 64, 7);
    bits_pack(&bits, main_pitch, 6);
    bits_pack(&bits, voiced ? modulation + 4 : 0, 3);
    bits_pack(&bits, corr_id, 2);
    bits_pack(&bits, vq_end[0], 10);
    bits_pack(&bits, vq_end[1], 10);
    bits_pack(&bits, vq_end[2], 10);
    bits_pack(&bits, vq_mid, 13);
    bits_pack(&bits, interp_id, 3);
  }
}

/* lpcnet_encode (quantize = encode = 1, buf[8]) and lpcnet_compute_features (both 0) (lpcnet_enc.c:882-911); pcm as
 * floats (shorts convert exactly).  features: [4][36] out (the encoder's quantised rows, or the analysis rows). */
void orc_enc_superframe(OrcEncState *st, const float *pcm /*[640]*/, unsigned char *buf, float *features, int quantize,
                        const float *codebooks) {
  float xc[8][PMAX + 1], fw[8], F[4][NTOT], x[FRAME];
  int i, k;
  for (k = 0; k < 4; k++) {
    for (i = 0; i < FRAME; i++) {            /* preemphasis (lpcnet_enc.c:872-880), coef 0.85f */
      float yi = pcm[k * FRAME + i] + st->mem_preemph;
      st->mem_preemph = -0.85f * pcm[k * FRAME + i];
      x[i] = yi;
    }
    orc_enc_frame_features(st, x);
    memcpy(xc[2 * k], st->xc[0], sizeof(xc[0]));
    memcpy(xc[2 * k + 1], st->xc[1], sizeof(xc[0]));
    fw[2 * k] = st->frame_weight[0];
    fw[2 * k + 1] = st->frame_weight[1];
    memcpy(F[k], st->features, sizeof(F[k]));
  }
  process_superframe(st, xc, fw, F, buf, quantize, quantize, codebooks);
  if (features) memcpy(features, F, sizeof(F));
}
