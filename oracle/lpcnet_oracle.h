/* CPU oracle for the LPCNet synthesis hot path: a plain-C restatement of the reference
 * algorithm (AVX2 DOT_PROD + USE_SU_BIAS semantics, src/vec_avx.h:39-41), written from
 * the reference's behaviour, each function citing the file:line it follows.
 *
 * TEST INFRASTRUCTURE ONLY.  Only tests/, __graft_entry__.smoke() and bench.py's
 * cpu_baseline leg may link or call this.  The product (lpcnet_b200/) never does and
 * has no CPU fallback.
 *
 * PINNED: checked bit-for-bit against the unmodified reference sources compiled in this
 * container (oracle/_ref/, flags -O2 -mavx2 -mfma -ffp-contract=off) and against the
 * committed outputs of that build in tests/golden/ (tests/test_oracle_*.py).  The
 * reference's own tests hold no golden vectors for this path.
 *
 * Pinned floating-point semantics: every C operator is one IEEE binary32/binary64
 * operation (no contraction); the reference's explicit _mm256_fmadd_ps are fmaf();
 * `rcpps` is a 2048-entry table indexed by the top 11 mantissa bits (measured from the
 * CPU the fixtures were generated on and passed in at model load).
 */
#ifndef LPCNET_ORACLE_H
#define LPCNET_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORC_LPC_ORDER 16
#define ORC_NB_FEATURES 20
#define ORC_NB_BANDS 18
#define ORC_FRAME_SIZE 160
#define ORC_MAX_DELAY 2

typedef struct OrcModel OrcModel;
typedef struct OrcState OrcState;

/* same layout as RefTrace in ref_harness.c */
typedef struct {
  int32_t exc, u_sig, u_pred;
  uint32_t rng_z, rng_w, xa_hash, xb_hash;
  int32_t pcm;
} OrcTrace;

/* same layout as RefSnapshot in ref_harness.c (default model sizes) */
typedef struct {
  float conv1_state[168];
  float conv2_state[256];
  float gru_a_state[384];
  float gru_b_state[16];
  float last_sig[16];
  float lpc[16];
  float old_lpc[2][16];
  float gru_a_condition[1152];
  float gru_b_condition[48];
  float deemph_mem;
  int32_t last_exc;
  int32_t frame_count;
  uint32_t rng[4];
} OrcSnapshot;

/* rcp_table: 2048 binary32 bit patterns, rcpps(1 + i/2048); NULL = built-in model of the
 * Intel estimate table (verified against the container's CPU in tests). */
OrcModel *orc_model_load(const unsigned char *blob, int len, int features_delay, float lpc_gamma,
                         const uint32_t *rcp_table);
void orc_model_free(OrcModel *m);
/* the reference's -DEND2END build: LPC = rc2lpc(condition[0..16)) (lpcnet.c:57-78, 108); default off */
void orc_model_set_end2end(OrcModel *m, int on);
int orc_model_nblocks_a(const OrcModel *m);
int orc_model_nblocks_b(const OrcModel *m);

OrcState *orc_state_new(const OrcModel *m);
void orc_state_free(OrcState *s);
void orc_state_reset(OrcState *s);
void orc_seed(OrcState *s, const unsigned char *data, int n);
void orc_get(const OrcState *s, OrcSnapshot *snap);
void orc_set(OrcState *s, const OrcSnapshot *snap);

/* One lpcnet_synthesize_impl (src/lpcnet.c:273-277): frame network, then N samples. */
void orc_synthesize(const OrcModel *m, OrcState *s, const float *features, int16_t *out, int N, int preload);
/* n_frames frames of one stream; forced/trace/logits nullable (see ref_run). */
void orc_run(const OrcModel *m, OrcState *s, const float *feats, int feat_stride, int n_frames,
             int16_t *pcm, const int16_t *forced, OrcTrace *trace, float *logits);
void orc_frame_network(const OrcModel *m, OrcState *s, const float *features, float *ga, float *gb, float *lpc);

/* primitives (known-answer tests against the reference's inline functions) */
void orc_set_rcp_table(const uint32_t *tab);
void orc_default_rcp_table(uint32_t *tab);
float orc_rcp(float x);
float orc_tanh_approx(float x);
float orc_sigmoid_approx(float x);
int orc_lin2ulaw(float x);
float orc_ulaw2lin(float u);
void orc_kiss99_srand(uint32_t st[4], const unsigned char *data, int n);
uint32_t orc_kiss99_rand(uint32_t st[4]);
void orc_ps_to_epi8(unsigned char *x, const float *in, int len);
void orc_sparse_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const int *idx, const float *x);
void orc_sgemv_accum8x4(float *out, const signed char *w, int rows, int cols, const float *x);
void orc_sgemv_accum(float *out, const float *w, int rows, int cols, int col_stride, const float *x);
void orc_lpc_from_cepstrum(float *lpc, const float *cepstrum);
void orc_fft_tables(float *twiddles /*[320][2]*/, int16_t *bitrev /*[320]*/, float *dct /*[18*18]*/);
void orc_decode_packet(float *features /*[4][36]*/, float *vq_mem /*[18]*/, const unsigned char *buf,
                       const float *codebooks /*cb1|cb2|cb3|diff4*/);

/* Feature analysis (lpcnet_compute_single_frame_features, src/lpcnet_enc.c:913-933): 160 samples -> 36 features */
typedef struct OrcEncState OrcEncState;
OrcEncState *orc_enc_new(void);
void orc_enc_free(OrcEncState *s);
void orc_enc_single_frame(OrcEncState *s, const float *pcm /*[160]*/, float *features /*[36]*/);
/* lpcnet_encode (quantize 1: 640 samples -> 8 bytes, features = the quantised rows) and lpcnet_compute_features
 * (quantize 0, buf unused) (src/lpcnet_enc.c:882-911); codebooks = cb1|cb2|cb3|diff4 as for orc_decode_packet */
void orc_enc_superframe(OrcEncState *s, const float *pcm /*[640]*/, unsigned char *buf /*[8]*/, float *features /*[4][36]*/,
                        int quantize, const float *codebooks);
const float *orc_half_window(void);   /* [160] */
/* burg_cepstral_analysis (src/freq.c:188-199): 160 samples -> 2 x 18 Burg cepstra (mean, difference) */
void orc_burg_cepstral_analysis(float *ceps /*[36]*/, const float *x /*[160]*/);

/* RDOVAE decoder of DRED (src/dred_rdovae.c:37-51): features [nb_latents][80], state [24], latents [nb_latents][80] */
int orc_dred_decode_all(const OrcModel *m, float *features, const float *state, const float *latents, int nb_latents);

/* Packet-loss concealment, non-causal mode (src/lpcnet_plc.c:349-491); the model must carry the plc_* arrays and
 * be loaded with features_delay 0.  options: 1 (LPCNET_PLC_NONCAUSAL) | 4 (LPCNET_PLC_DC_FILTER). */
typedef struct OrcPLC OrcPLC;
OrcPLC *orc_plc_new(const OrcModel *m, int options);
void orc_plc_free(OrcPLC *p);
void orc_plc_seed(OrcPLC *p, const unsigned char *data, int n);
int orc_plc_update(OrcPLC *p, int16_t *pcm /*[160] in/out*/);
int orc_plc_conceal(OrcPLC *p, int16_t *pcm /*[160] out*/);
/* lpcnet_plc_fec_add (features NULL = skip one) / lpcnet_plc_fec_clear: read by the causal and codec modes only */
void orc_plc_fec_add(OrcPLC *p, const float *features /*[20] or NULL*/);
void orc_plc_fec_clear(OrcPLC *p);
void orc_plc_pred(const OrcModel *m, float *gru1 /*[256]*/, float *gru2 /*[256]*/, float *out /*[20]*/, const float *in /*[57]*/);
void orc_plc_peek(const OrcPLC *p, float *features20, float *gru1, float *gru2, int *loss_count, double *dc /*[2]*/);

#ifdef __cplusplus
}
#endif
#endif
