/* Stand-in for the generated dred_rdovae_dec_data.c: init_rdovaedec over the weight-array list, as init_plc_model. */
#include "dred_rdovae.h"
#include "dred_rdovae_dec_data.h"

int init_rdovaedec(RDOVAEDec *model, const WeightArray *arrays) {
  if (dense_init(&model->state1, arrays, "state1_bias", "state1_weights", 24, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->state2, arrays, "state2_bias", "state2_weights", 24, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->state3, arrays, "state3_bias", "state3_weights", 24, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_dense1, arrays, "dec_dense1_bias", "dec_dense1_weights", 80, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_dense3, arrays, "dec_dense3_bias", "dec_dense3_weights", 256, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_dense5, arrays, "dec_dense5_bias", "dec_dense5_weights", 256, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_dense7, arrays, "dec_dense7_bias", "dec_dense7_weights", 256, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_dense8, arrays, "dec_dense8_bias", "dec_dense8_weights", 256, 256, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->dec_final, arrays, "dec_final_bias", "dec_final_weights", 2048, 80, ACTIVATION_LINEAR)) return 1;
  if (gru_init(&model->dec_dense2, arrays, "dec_dense2_bias", "dec_dense2_subias", "dec_dense2_weights", "dec_dense2_weights_idx", "dec_dense2_recurrent_weights", 256, 256, ACTIVATION_TANH, 1)) return 1;
  if (gru_init(&model->dec_dense4, arrays, "dec_dense4_bias", "dec_dense4_subias", "dec_dense4_weights", "dec_dense4_weights_idx", "dec_dense4_recurrent_weights", 256, 256, ACTIVATION_TANH, 1)) return 1;
  if (gru_init(&model->dec_dense6, arrays, "dec_dense6_bias", "dec_dense6_subias", "dec_dense6_weights", "dec_dense6_weights_idx", "dec_dense6_recurrent_weights", 256, 256, ACTIVATION_TANH, 1)) return 1;
  return 0;
}
