/* Oracle: the RDOVAE decoder of DRED: DRED_rdovae_decode_all = dred_rdovae_dec_init_states +
 * dred_rdovae_decode_qframe per latent vector (src/dred_rdovae.c:37-51, src/dred_rdovae_dec.c:37-98): dense and
 * int8 compute_gruB layers only.  Its output rows are what a caller hands to lpcnet_plc_fec_add.
 * TEST INFRASTRUCTURE ONLY.  Compile with -ffp-contract=off. */
#include <string.h>
#include "orc_internal.h"

#define C 256

/* features: [nb_latents][4*20] (each quadruple frame = four feature vectors in reverse order, as the reference
 * writes them); state: [24]; latents: [nb_latents][80].  Returns -1 if the model has no decoder arrays. */
int orc_dred_decode_all(const OrcModel *m, float *features, const float *state, const float *latents, int nb_latents) {
  float g[3][C], zero[3 * C] = {0}, buffer[8 * C];
  int i, k;
  if (!m->has_dred) return -1;
  /* dred_rdovae_dec_init_states (dred_rdovae_dec.c:37-48) */
  for (k = 0; k < 3; k++) orc_dense(&m->dred_state[k], g[k], state);
  for (i = 0; i < nb_latents; i++) {
    /* dred_rdovae_decode_qframe (dred_rdovae_dec.c:51-98): the eight layer outputs are concatenated in `buffer` */
    orc_dense(&m->dred_dense[0], &buffer[0], &latents[i * 80]);
    orc_gru_b(&m->dred_gru[0], zero, g[0], &buffer[0]);
    memcpy(&buffer[C], g[0], sizeof(g[0]));
    orc_dense(&m->dred_dense[1], &buffer[2 * C], &buffer[C]);
    orc_gru_b(&m->dred_gru[1], zero, g[1], &buffer[2 * C]);
    memcpy(&buffer[3 * C], g[1], sizeof(g[1]));
    orc_dense(&m->dred_dense[2], &buffer[4 * C], &buffer[3 * C]);
    orc_gru_b(&m->dred_gru[2], zero, g[2], &buffer[4 * C]);
    memcpy(&buffer[5 * C], g[2], sizeof(g[2]));
    orc_dense(&m->dred_dense[3], &buffer[6 * C], &buffer[5 * C]);
    orc_dense(&m->dred_dense[4], &buffer[7 * C], &buffer[6 * C]);
    orc_dense(&m->dred_final, &features[i * 80], buffer);
  }
  return 0;
}
