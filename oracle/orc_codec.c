/* Oracle: 1.6 kb/s packet decoder (frame-rate pre-step of lpcnet_decode).
 * TEST INFRASTRUCTURE ONLY.  Follows src/lpcnet_dec.c:59-155 and src/common.c:37-65. */
#include <math.h>
#include <string.h>
#include "orc_internal.h"

typedef struct { const unsigned char *chars; int byte_pos, bit_pos; } Unpacker;

/* bits_unpack (lpcnet_dec.c:59-78): MSB-first */
static unsigned int bits_unpack(Unpacker *u, int nb_bits) {
  unsigned int d = 0;
  while (nb_bits) {
    d <<= 1;
    d |= (u->chars[u->byte_pos] >> (8 - 1 - u->bit_pos)) & 1;
    u->bit_pos++;
    if (u->bit_pos == 8) { u->bit_pos = 0; u->byte_pos++; }
    nb_bits--;
  }
  return d;
}

/* single_interp / perform_double_interp (common.c:37-65) */
static void single_interp(float *x, const float *left, const float *right, int id) {
  int i;
  for (i = 0; i < 18; i++) {
    if (id == 0) x[i] = .5f * (left[i] + right[i]);
    else if (id == 1) x[i] = left[i];
    else x[i] = right[i];
  }
}

/* decode_packet (lpcnet_dec.c:81-155); codebooks = cb1[1024*17] | cb2 | cb3 | diff4[4096*18] */
void orc_decode_packet(float *features, float *vq_mem, const unsigned char *buf, const float *codebooks) {
  const float *cb1 = codebooks, *cb2 = cb1 + 1024 * 17, *cb3 = cb2 + 1024 * 17, *diff4 = cb3 + 1024 * 17;
  float (*F)[36] = (float (*)[36])features;
  Unpacker bits = {buf, 0, 0};
  int c0_id, main_pitch, modulation, corr_id, vq_end[3], vq_mid, interp_id, i, sub, voiced = 1, id0, id1;
  float frame_corr, sign;
  c0_id = bits_unpack(&bits, 7);
  main_pitch = bits_unpack(&bits, 6);
  modulation = bits_unpack(&bits, 3);
  corr_id = bits_unpack(&bits, 2);
  vq_end[0] = bits_unpack(&bits, 10);
  vq_end[1] = bits_unpack(&bits, 10);
  vq_end[2] = bits_unpack(&bits, 10);
  vq_mid = bits_unpack(&bits, 13);
  interp_id = bits_unpack(&bits, 3);
  memset(features, 0, 4 * 36 * sizeof(float));
  modulation -= 4;
  if (modulation == -4) { voiced = 0; modulation = 0; }
  if (voiced) frame_corr = 0.3875f + .175f * (float)corr_id;
  else frame_corr = 0.0375f + .075f * (float)corr_id;
  for (sub = 0; sub < 4; sub++) {
    float p = (float)(pow((double)2.f, main_pitch / 21.) * 32);
    p *= 1.f + (float)modulation / 16.f / 7.f * (float)(2 * sub - 3);
    p = 33 > p ? 33 : p;
    p = 255 < p ? 255 : p;
    F[sub][18] = .02f * (p - 100.f);
    F[sub][19] = frame_corr - .5f;
  }
  F[3][0] = (float)(c0_id - 64) / 4.f;
  for (i = 0; i < 17; i++)
    F[3][i + 1] = cb1[vq_end[0] * 17 + i] + cb2[vq_end[1] * 17 + i] + cb3[vq_end[2] * 17 + i];
  sign = 1;
  if (vq_mid >= 4096) { vq_mid -= 4096; sign = -1; }
  for (i = 0; i < 18; i++) F[1][i] = sign * diff4[vq_mid * 18 + i];
  if ((vq_mid & 3) < 2) {
    for (i = 0; i < 18; i++) F[1][i] += .5f * (vq_mem[i] + F[3][i]);
  } else if ((vq_mid & 3) == 2) {
    for (i = 0; i < 18; i++) F[1][i] += vq_mem[i];
  } else {
    for (i = 0; i < 18; i++) F[1][i] += F[3][i];
  }
  interp_id += (interp_id >= 7);
  id0 = interp_id / 3; id1 = interp_id % 3;
  single_interp(F[0], vq_mem, F[1], id0);
  single_interp(F[2], F[1], F[3], id1);
  memcpy(vq_mem, F[3], 18 * sizeof(float));
}
