/* Stand-in for the generated plc_data.c (training_tf2/dump_plc.py:160-208, 331-368). */
#include "nnet.h"
#include "plc_data.h"

int init_plc_model(PLCModel *model, const WeightArray *arrays) {
  if (dense_init(&model->plc_dense1, arrays, "plc_dense1_bias", "plc_dense1_weights", 57, 128, ACTIVATION_TANH)) return 1;
  if (gru_init(&model->plc_gru1, arrays, "plc_gru1_bias", "plc_gru1_subias", "plc_gru1_weights", "plc_gru1_weights_idx", "plc_gru1_recurrent_weights", 128, 256, ACTIVATION_TANH, 1)) return 1;
  if (gru_init(&model->plc_gru2, arrays, "plc_gru2_bias", "plc_gru2_subias", "plc_gru2_weights", "plc_gru2_weights_idx", "plc_gru2_recurrent_weights", 256, 256, ACTIVATION_TANH, 1)) return 1;
  if (dense_init(&model->plc_out, arrays, "plc_out_bias", "plc_out_weights", 256, 20, ACTIVATION_LINEAR)) return 1;
  return 0;
}
