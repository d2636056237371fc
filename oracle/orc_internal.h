/* Internal structures of the CPU oracle.  TEST INFRASTRUCTURE ONLY. */
#ifndef ORC_INTERNAL_H
#define ORC_INTERNAL_H
#include "lpcnet_oracle.h"

typedef struct { const float *bias, *w; int nin, nout, act; } OrcDense;      /* nnet.h:64-70 */
typedef struct { const float *bias, *w; int nin, k, nout, act; } OrcConv;    /* nnet.h:105-112 */
typedef struct { const float *w; int n, dim; } OrcEmbed;                      /* nnet.h:114-118 */
typedef struct {                                                              /* nnet.h:82-92 */
  const float *bias, *subias; const signed char *w_in; const int *idx_in; const signed char *w_rec;
  int nin, n, nblocks;
} OrcGRU;
typedef struct {                                                              /* nnet.h:94-103 */
  const float *bias, *subias, *diag; const signed char *w; const int *idx; int n, nblocks;
} OrcSparseGRU;
typedef struct { const float *bias, *w, *factor; int nin, n, c; } OrcMDense; /* nnet.h:72-80 */

enum { ORC_LINEAR = 0, ORC_SIGMOID = 1, ORC_TANH = 2 };

struct OrcModel {
  unsigned char *blob; int len;
  int features_delay; float lpc_gamma;
  int end2end;             /* the reference's -DEND2END build (lpcnet.c:57-78, 108) */
  OrcEmbed e_sig, e_pred, e_exc, e_pitch;
  OrcDense a_feat, b_feat, dense1, dense2;
  OrcConv conv1, conv2;
  OrcGRU gru_b;
  OrcSparseGRU gru_a;
  OrcMDense dual_fc;
  float logit_table[256];
  /* PLC predictor (plc_data.c stand-in, training_tf2/dump_plc.py), present when the blob carries plc_* arrays */
  int has_plc;
  OrcDense plc_dense1, plc_out;
  OrcGRU plc_gru1, plc_gru2;
  /* RDOVAE decoder of DRED (dred_rdovae_dec_data.c stand-in), present when the blob carries the dec_* arrays */
  int has_dred;
  OrcDense dred_state[3], dred_dense[5], dred_final;   /* state1..3; dec_dense1,3,5,7,8; dec_final */
  OrcGRU dred_gru[3];                                   /* dec_dense2,4,6 */
};

/* LPCNetEncState (lpcnet_private.h:55-75), the fields single-frame analysis uses */
#define PMAX 256
#define OVERLAP 160
struct OrcEncState {
  float analysis_mem[OVERLAP];
  float mem_preemph;
  float pitch_mem[ORC_LPC_ORDER];
  float pitch_filt;
  float xc[2][PMAX + 1];            /* rows 2, 3 of the reference's xc[10]: pcount stays 0 in single-frame mode */
  float frame_weight[2];
  float exc_buf[PMAX + 320];
  float pitch_max_path[2][PMAX];
  float pitch_max_path_all;
  int best_i;
  float lpc[ORC_LPC_ORDER];
  float features[36];
  float vq_mem[18];                 /* superframe encoder only (orc_encode.c) */
};
void orc_enc_frame_features(struct OrcEncState *st, const float *in /*[160], pre-emphasised*/);
#undef PMAX
#undef OVERLAP

struct OrcState {
  const OrcModel *m;
  uint32_t rng[4];
  /* everything below is zeroed by reset (lpcnet_private.h:33, lpcnet.c:174-182) */
  float conv1_state[168], conv2_state[256], gru_a_state[384], gru_b_state[16];
  int last_exc;
  float last_sig[ORC_LPC_ORDER];
  float old_lpc[ORC_MAX_DELAY][ORC_LPC_ORDER];
  float gru_a_condition[1152], gru_b_condition[48];
  int frame_count;
  float deemph_mem;
  float lpc[ORC_LPC_ORDER];
};

typedef struct { float r, i; } OrcCpx;
void orc_fft320(const OrcCpx *fin, OrcCpx *fout);            /* opus_fft_c, nfft = 320, scale 1/320 */
void orc_levinson(float *lpc, const float *ac, int p);
const float *orc_dct_table(void);
void orc_vec_act(float *y, const float *x, int n, int act);
void orc_dense(const OrcDense *l, float *out, const float *in);
void orc_conv1d(const OrcConv *l, float *out, float *mem, const float *in);
void orc_gru_b(const OrcGRU *g, const float *cond, float *state, const float *input);   /* compute_gruB */
void orc_sample_tail(const OrcModel *m, OrcState *s, int16_t *out, int N, int preload);
int orc_sample_network(const OrcModel *m, OrcState *s, int last_exc, int u_sig, int u_pred);
void orc_all_node_logits(const OrcModel *m, const OrcState *s, float *logits);
#endif
