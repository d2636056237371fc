/* Stand-in for the generated nnet_data.c: the init call list that dump_lpcnet.py
 * writes through model_init (dump_lpcnet.py:157-169, 205-218, 240-244, 271-282,
 * 308-318, 329-333).  Built with -DUSE_WEIGHTS_FILE so no array literals exist. */
#include "nnet.h"
#include "nnet_data.h"

int init_lpcnet_model(LPCNetModel *model, const WeightArray *arrays) {
  if (embedding_init(&model->gru_a_embed_sig, arrays, "gru_a_embed_sig_weights", 256, 1152)) return 1;
  if (embedding_init(&model->gru_a_embed_pred, arrays, "gru_a_embed_pred_weights", 256, 1152)) return 1;
  if (embedding_init(&model->gru_a_embed_exc, arrays, "gru_a_embed_exc_weights", 256, 1152)) return 1;
  if (dense_init(&model->gru_a_dense_feature, arrays, "gru_a_dense_feature_bias", "gru_a_dense_feature_weights", 128, 1152, ACTIVATION_LINEAR)) return 1;
  if (dense_init(&model->gru_b_dense_feature, arrays, "gru_b_dense_feature_bias", "gru_b_dense_feature_weights", 128, 48, ACTIVATION_LINEAR)) return 1;
  if (gru_init(&model->gru_b, arrays, "gru_b_bias", "gru_b_subias", "gru_b_weights", "gru_b_weights_idx", "gru_b_recurrent_weights", 384, 16, ACTIVATION_TANH, 1)) return 1;
  if (conv1d_init(&model->feature_conv1, arrays, "feature_conv1_bias", "feature_conv1_weights", 84, 3, 128, ACTIVATION_TANH)) return 1;
  if (conv1d_init(&model->feature_conv2, arrays, "feature_conv2_bias", "feature_conv2_weights", 128, 3, 128, ACTIVATION_TANH)) return 1;
  if (embedding_init(&model->embed_pitch, arrays, "embed_pitch_weights", 256, 64)) return 1;
  if (embedding_init(&model->embed_sig, arrays, "embed_sig_weights", 256, 128)) return 1;
  if (dense_init(&model->feature_dense1, arrays, "feature_dense1_bias", "feature_dense1_weights", 128, 128, ACTIVATION_TANH)) return 1;
  if (dense_init(&model->feature_dense2, arrays, "feature_dense2_bias", "feature_dense2_weights", 128, 128, ACTIVATION_TANH)) return 1;
  if (mdense_init(&model->dual_fc, arrays, "dual_fc_bias", "dual_fc_weights", "dual_fc_factor", 16, 256, 2, ACTIVATION_SIGMOID)) return 1;
  if (sparse_gru_init(&model->sparse_gru_a, arrays, "sparse_gru_a_bias", "sparse_gru_a_subias", "sparse_gru_a_recurrent_weights_diag", "sparse_gru_a_recurrent_weights", "sparse_gru_a_recurrent_weights_idx", 384, ACTIVATION_TANH, 1)) return 1;
  return 0;
}
