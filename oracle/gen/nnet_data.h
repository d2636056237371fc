/* Stand-in for the generated nnet_data.h that training_tf2/dump_lpcnet.py would emit
 * (dump_lpcnet.py:420-509) for the default model (rnn_units1=384, rnn_units2=16,
 * cond_size=128).  The real file ships in a model tarball that cannot be downloaded
 * offline.  Test infrastructure only: used to compile the unmodified reference
 * sources into oracle/_ref/.  FEATURES_DELAY is selectable (-DORACLE_FEATURES_DELAY=0
 * for the non-causal PLC build, lpcnet_plc.c:357-361). */
#ifndef RNN_DATA_H
#define RNN_DATA_H

#include "nnet.h"

/* This is *not* an end-to-end model */
/* #define END2END */

#define LPC_GAMMA 1.0f

#ifndef ORACLE_FEATURES_DELAY
#define ORACLE_FEATURES_DELAY 2
#endif
#define FEATURES_DELAY ORACLE_FEATURES_DELAY

#define GRU_A_EMBED_SIG_OUT_SIZE 1152
#define GRU_A_EMBED_PRED_OUT_SIZE 1152
#define GRU_A_EMBED_EXC_OUT_SIZE 1152
#define GRU_A_DENSE_FEATURE_OUT_SIZE 1152
#define GRU_B_DENSE_FEATURE_OUT_SIZE 48
#define FEATURE_CONV1_OUT_SIZE 128
#define FEATURE_CONV1_STATE_SIZE (84*2)
#define FEATURE_CONV1_DELAY 1
#define FEATURE_CONV2_OUT_SIZE 128
#define FEATURE_CONV2_STATE_SIZE (128*2)
#define FEATURE_CONV2_DELAY 1
#define EMBED_PITCH_OUT_SIZE 64
#define EMBED_SIG_OUT_SIZE 128
#define FEATURE_DENSE1_OUT_SIZE 128
#define FEATURE_DENSE2_OUT_SIZE 128
#define GRU_A_OUT_SIZE 384
#define GRU_A_STATE_SIZE 384
#define GRU_B_OUT_SIZE 16
#define GRU_B_STATE_SIZE 16
#define DUAL_FC_OUT_SIZE 256
#define SPARSE_GRU_A_OUT_SIZE 384
#define SPARSE_GRU_A_STATE_SIZE 384

#define MAX_RNN_NEURONS 384
#define MAX_CONV_INPUTS 384
#define MAX_MDENSE_TMP 512

typedef struct {
  float feature_conv1_state[FEATURE_CONV1_STATE_SIZE];
  float feature_conv2_state[FEATURE_CONV2_STATE_SIZE];
  float gru_a_state[GRU_A_STATE_SIZE];
  float gru_b_state[GRU_B_STATE_SIZE];
} NNetState;

typedef struct {
  EmbeddingLayer gru_a_embed_sig;
  EmbeddingLayer gru_a_embed_pred;
  EmbeddingLayer gru_a_embed_exc;
  DenseLayer gru_a_dense_feature;
  DenseLayer gru_b_dense_feature;
  GRULayer gru_b;
  Conv1DLayer feature_conv1;
  Conv1DLayer feature_conv2;
  EmbeddingLayer embed_pitch;
  EmbeddingLayer embed_sig;
  DenseLayer feature_dense1;
  DenseLayer feature_dense2;
  MDenseLayer dual_fc;
  SparseGRULayer sparse_gru_a;
} LPCNetModel;

int init_lpcnet_model(LPCNetModel *model, const WeightArray *arrays);

#endif
