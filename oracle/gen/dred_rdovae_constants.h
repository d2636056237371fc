/* Stand-in for the generated dred_rdovae_constants.h (torch/rdovae/export_rdovae_weights.py:226-245,
 * training_tf2/dump_rdovae.py:307-325): the sizes of the random-init RDOVAE decoder the oracle builds
 * (cond_size = cond_size2 = 256, latent_dim 80, state_dim 24), and the two maxima nnet.c reads (nnet.c:324, 450). */
#ifndef DRED_RDOVAE_CONSTANTS_H
#define DRED_RDOVAE_CONSTANTS_H
#define DRED_NUM_FEATURES 20
#define DRED_LATENT_DIM 80
#define DRED_STATE_DIM 24
#define DRED_NUM_QUANTIZATION_LEVELS 16
#define DRED_MAX_RNN_NEURONS 256
#define DRED_MAX_CONV_INPUTS 128
#endif
