/* lpcnet_b200.h -- C ABI of the B200-native LPCNet synthesis engine (liblpcnet_b200.so).
 *
 * Two groups of entry points, all `extern "C"`, plain pointers and sizes:
 *
 *  1. Drop-in symbols: same names, argument meaning and return conventions as the
 *     reference's include/lpcnet.h for the synthesis path, so a program compiled against
 *     the reference header links against this library unchanged.  Each declaration cites
 *     the reference line it replaces.  The opaque state is a small host handle; the real
 *     per-stream state lives in HBM.
 *
 *  2. The batched multi-stream API (lpcnet_batch_*): the entry point BASELINE.json's
 *     north_star adds next to the reference API.  One batch = n_streams independent
 *     LPCNetStates advanced in lock step by one persistent CUDA kernel.
 *
 * There is no CPU fallback: every call that needs the GPU fails (non-zero return /
 * latched error) when no CUDA device is present.
 */
#ifndef LPCNET_B200_H
#define LPCNET_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef LPCNET_B200_EXPORT
#define LPCNET_B200_EXPORT __attribute__((visibility("default")))
#endif

/* Same values as include/lpcnet.h:45-53 of the reference. */
#define LPCNET_B200_NB_FEATURES 20
#define LPCNET_B200_NB_TOTAL_FEATURES 36
#define LPCNET_B200_FRAME_SIZE 160
#define LPCNET_B200_PACKET_BYTES 8
#define LPCNET_B200_PACKET_SAMPLES 640

/* ------------------------------------------------------------------------------------
 * 1. Drop-in symbols (reference: include/lpcnet.h; implementation src/lpcnet.c)
 * ------------------------------------------------------------------------------------ */
typedef struct LPCNetState LPCNetState;       /* include/lpcnet.h:55 */
typedef struct LPCNetDecState LPCNetDecState; /* include/lpcnet.h:57 */

/* include/lpcnet.h:160, src/lpcnet.c:169-172.  Size of the host handle. */
LPCNET_B200_EXPORT int lpcnet_get_size(void);
/* include/lpcnet.h:169, src/lpcnet.c:184-200.  Returns 0.  Like the reference's
 * USE_WEIGHTS_FILE build, the state has no model until lpcnet_load_model is called. */
LPCNET_B200_EXPORT int lpcnet_init(LPCNetState *st);
/* include/lpcnet.h:174, src/lpcnet.c:213-219. */
LPCNET_B200_EXPORT LPCNetState *lpcnet_create(void);
/* include/lpcnet.h:179, src/lpcnet.c:221-224. */
LPCNET_B200_EXPORT void lpcnet_destroy(LPCNetState *st);
/* include/lpcnet.h:78, src/lpcnet.c:174-182: zero the stream state, last_exc=128,
 * reseed kiss99 with "LPCNet". */
LPCNET_B200_EXPORT void lpcnet_reset(LPCNetState *st);
/* include/lpcnet.h:214, src/lpcnet.c:202-210.  0 on success, -1 on a malformed blob
 * (same validation as src/parse_lpcnet_weights.c:36-113) or when no GPU is available.
 * The blob is copied to the device; it need not outlive the call (a superset of the
 * reference, which borrows it). */
LPCNET_B200_EXPORT int lpcnet_load_model(LPCNetState *st, const unsigned char *data, int len);
/* include/lpcnet.h:188, src/lpcnet.c:279-281: one frame-network evaluation on
 * features[0..20), then N (<=160) samples into output.  void like the reference: a CUDA
 * failure is latched and reported by lpcnet_b200_last_error(). */
LPCNET_B200_EXPORT void lpcnet_synthesize(LPCNetState *st, const float *features, short *output, int N);

/* include/lpcnet.h:67-96, src/lpcnet.c:285-319. */
LPCNET_B200_EXPORT int lpcnet_decoder_get_size(void);
LPCNET_B200_EXPORT int lpcnet_decoder_init(LPCNetDecState *st);
LPCNET_B200_EXPORT LPCNetDecState *lpcnet_decoder_create(void);
LPCNET_B200_EXPORT void lpcnet_decoder_destroy(LPCNetDecState *st);
/* buf: 8 bytes; pcm: 640 samples.  Returns 0.  The VQ codebooks are compiled into the
 * reference (ceps_codebooks.c, absent from the tree); here they are looked up in the
 * blob under their C names (ceps_codebook1/2/3, ceps_codebook_diff4) or installed with
 * lpcnet_b200_set_codebooks().  Extension, not in the reference header: */
LPCNET_B200_EXPORT int lpcnet_decoder_load_model(LPCNetDecState *st, const unsigned char *data, int len);
LPCNET_B200_EXPORT int lpcnet_decode(LPCNetDecState *st, const unsigned char *buf, short *pcm);

/* ------------------------------------------------------------------------------------
 * 2. Batched multi-stream API (new; SURVEY.md 8b last row)
 * ------------------------------------------------------------------------------------ */
typedef struct LPCNetBatch LPCNetBatch;

/* Host-visible per-stream state: the mutable part of the reference's LPCNetState
 * (src/lpcnet_private.h:28-48) for the default model sizes, used for snapshot/restore
 * and for cross-loading states to and from the reference. */
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
  uint32_t rng[4]; /* kiss99: z, w, jsr, jcong */
} LPCNetB200StreamState;

/* Per-step debug record (teacher-forced parity runs). */
typedef struct {
  int32_t exc;      /* last_exc after the step */
  int32_t u_sig;    /* lin2ulaw(last_sig[0]) fed to the step (src/lpcnet.c:253) */
  int32_t u_pred;   /* lin2ulaw(pred) fed to the step (src/lpcnet.c:254) */
  uint32_t rng_z;   /* kiss99 state after the step */
  uint32_t rng_w;
  uint32_t xa_hash; /* FNV-1a over the u8-quantised GRU-A state after the step */
  uint32_t xb_hash; /* same for GRU-B */
  int32_t pcm;      /* output sample (or the forced one) */
} LPCNetB200Trace;

/* features_delay and lpc_gamma are FEATURES_DELAY / LPC_GAMMA of the generated
 * nnet_data.h (training_tf2/dump_lpcnet.py:437-439); they are not stored in the blob.
 * device < 0 selects the current CUDA device.  Returns NULL on error (see
 * lpcnet_b200_last_error). */
LPCNET_B200_EXPORT LPCNetBatch *lpcnet_batch_create(const unsigned char *blob, int len, int n_streams,
                                                   int device, int features_delay, float lpc_gamma);
LPCNET_B200_EXPORT void lpcnet_batch_destroy(LPCNetBatch *b);
LPCNET_B200_EXPORT int lpcnet_batch_n_streams(const LPCNetBatch *b);
/* lpcnet_reset for every stream (stream < 0) or one stream. */
LPCNET_B200_EXPORT int lpcnet_batch_reset(LPCNetBatch *b, int stream);
/* kiss99_srand(&st->rng, data, n) (src/kiss99.c:32-57) for one stream. */
LPCNET_B200_EXPORT int lpcnet_batch_set_seed(LPCNetBatch *b, int stream, const unsigned char *data, int n);
/* Run on this cudaStream_t (as void*) instead of the batch's own stream. */
LPCNET_B200_EXPORT int lpcnet_batch_set_cuda_stream(LPCNetBatch *b, void *cuda_stream);
/* Streams per CTA tile: 0 = automatic (by n_streams), else one of 8, 16, 32. */
LPCNET_B200_EXPORT int lpcnet_batch_set_tile(LPCNetBatch *b, int streams_per_cta);

/* HOST buffers.  features: [n_streams][n_frames][feature_stride] floats (only [0,20) of
 * each frame is read, src/lpcnet_demo.c:213-215); pcm: [n_streams][n_frames*160].
 * forced_pcm (nullable): teacher forcing through the reference's preload rule
 * (src/lpcnet.c:256-258) on every step.  trace / node_logits (nullable, debug):
 * [n_streams][n_frames*160] records / [n_streams][n_frames*160][256] floats
 * (entry 0 unused).  Copies host<->device inside the call; synchronous. */
LPCNET_B200_EXPORT int lpcnet_batch_synthesize(LPCNetBatch *b, const float *features, int feature_stride,
                                               int n_frames, short *pcm, const short *forced_pcm,
                                               LPCNetB200Trace *trace, float *node_logits);
/* Same with DEVICE buffers, asynchronous on the batch's CUDA stream; samples_per_frame
 * (<=160) is the N of lpcnet_synthesize.  pcm: [n_streams][n_frames*samples_per_frame]. */
LPCNET_B200_EXPORT int lpcnet_batch_synthesize_device(LPCNetBatch *b, const float *d_features,
                                                      int feature_stride, int n_frames,
                                                      int samples_per_frame, short *d_pcm,
                                                      const short *d_forced_pcm, LPCNetB200Trace *d_trace,
                                                      float *d_node_logits);
/* Frame network only (debug/parity): evaluates n_frames frames, advancing the conv
 * states, old_lpc and frame_count, and returns the conditioning per frame (host buffers,
 * [n_streams][n_frames][1152|48|16]; any may be NULL). */
LPCNET_B200_EXPORT int lpcnet_batch_frame_network(LPCNetBatch *b, const float *features, int feature_stride,
                                                  int n_frames, float *gru_a_condition,
                                                  float *gru_b_condition, float *lpc);
/* packets: [n_streams][n_packets][8]; pcm: [n_streams][n_packets*640] (host buffers).
 * lpcnet_decode for every stream (src/lpcnet.c:310-319).  features_out (nullable):
 * [n_streams][n_packets*4][36] decoded features. */
LPCNET_B200_EXPORT int lpcnet_batch_decode(LPCNetBatch *b, const unsigned char *packets, int n_packets,
                                           short *pcm, float *features_out);
/* codebooks: ceps_codebook1|2|3 [1024*17 each] then ceps_codebook_diff4 [4096*18]. */
LPCNET_B200_EXPORT int lpcnet_batch_set_codebooks(LPCNetBatch *b, const float *codebooks, int n_floats);
LPCNET_B200_EXPORT int lpcnet_batch_get_state(LPCNetBatch *b, int stream, LPCNetB200StreamState *out);
LPCNET_B200_EXPORT int lpcnet_batch_set_state(LPCNetBatch *b, int stream, const LPCNetB200StreamState *in);
LPCNET_B200_EXPORT int lpcnet_batch_sync(LPCNetBatch *b);
/* Model facts.  which = 0 / 1: number of block-sparse 8x4 blocks in GRU-A / GRU-B input (for the op
 * counts); 2: 1 when the rcpps estimate is computed arithmetically (the table was verified equal on this
 * GPU), 0 when looked up; 3: 1 when GRU-A runs the unrolled fixed-slot path; 4: GRU-A chunks of 8 blocks; 5: streams per CTA in effect;
 * 8 / 16 / 32: bytes of shared memory the sample kernel needs per CTA at that many streams per CTA. */
LPCNET_B200_EXPORT int lpcnet_batch_nblocks(const LPCNetBatch *b, int which);
/* Kernels launched by this batch so far (for bench.py's gpu_launches). */
LPCNET_B200_EXPORT long long lpcnet_batch_launch_count(const LPCNetBatch *b);

/* Profiling: when on, CUDA events bracket every sample-kernel launch and every frame-network
 * launch group on the batch's stream; read_profile synchronises, returns the summed
 * durations and counts since the last read, and clears them. */
LPCNET_B200_EXPORT int lpcnet_batch_set_profiling(LPCNetBatch *b, int on);
/* After read_profile: duration (ms) and number of sample steps of each sample-kernel launch it summed.
 * Returns the number of launches; fills at most `max` entries (either array may be NULL). */
LPCNET_B200_EXPORT int lpcnet_batch_profile_launches(LPCNetBatch *b, float *ms, int *steps, int max);
LPCNET_B200_EXPORT int lpcnet_batch_read_profile(LPCNetBatch *b, double *sample_ms, int *sample_launches,
                                                 double *frame_ms, int *frame_groups);
/* Micro-benchmark used for the on-chip roofline: GB/s achieved by all SMs gathering random
 * 4608-byte rows (one embedding row) from an L2-resident 3.5 MB table with 128-bit loads. */
LPCNET_B200_EXPORT int lpcnet_b200_measure_l2_gather(int device, double *gbytes_per_s);

/* One handle over several GPUs of one box (no counterpart in the reference, which is single-stream
 * and CPU-only): n_streams are split into contiguous shards (sizes differ by at most 1) over the devices whose bit is
 * set in device_mask (0 = every visible device), one lpcnet_batch per GPU, one feeder thread per GPU.  Streams never
 * interact, so nothing crosses NVLink; a call hands every feeder its rows of the caller's HOST buffers (layouts of
 * lpcnet_batch_synthesize / _decode over all n_streams) and returns when every GPU is done.  Pinned buffers let each
 * GPU stream its chunks out while it computes.  Errors: -1 / NULL with the failing GPU named in lpcnet_b200_last_error(). */
typedef struct LPCNetMultiBatch LPCNetMultiBatch;
LPCNET_B200_EXPORT LPCNetMultiBatch *lpcnet_batch_create_multi(const unsigned char *blob, int len, int n_streams,
                                                               unsigned device_mask, int features_delay, float lpc_gamma);
LPCNET_B200_EXPORT void lpcnet_multi_destroy(LPCNetMultiBatch *mb);
LPCNET_B200_EXPORT int lpcnet_multi_n_streams(const LPCNetMultiBatch *mb);
LPCNET_B200_EXPORT int lpcnet_multi_n_devices(const LPCNetMultiBatch *mb);
/* shard k: the GPU it runs on and its stream range */
LPCNET_B200_EXPORT int lpcnet_multi_shard(const LPCNetMultiBatch *mb, int k, int *device, int *first_stream, int *n_streams);
LPCNET_B200_EXPORT int lpcnet_multi_set_seed(LPCNetMultiBatch *mb, int stream, const unsigned char *data, int n);
LPCNET_B200_EXPORT int lpcnet_multi_reset(LPCNetMultiBatch *mb, int stream /* -1: all */);
LPCNET_B200_EXPORT int lpcnet_multi_set_codebooks(LPCNetMultiBatch *mb, const float *codebooks, int n_floats);
LPCNET_B200_EXPORT int lpcnet_multi_synthesize(LPCNetMultiBatch *mb, const float *features, int feature_stride,
                                               int n_frames, short *pcm);
LPCNET_B200_EXPORT int lpcnet_multi_decode(LPCNetMultiBatch *mb, const unsigned char *packets, int n_packets, short *pcm);

/* Cepstrum -> LPC without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Computes what the frame network computes per frame before its delay line (lpc_from_cepstrum and
 * everything below it): d_cepstrum [n_rows][stride >= 18] floats and d_lpc [n_rows][16] floats are DEVICE buffers of
 * the caller; the launch is asynchronous on cuda_stream (NULL = the default stream).  The kernel's FFT scratch is a
 * static device array inside the library (10 MiB, held by every GPU that loads the library until exit), so
 * n_rows <= LPCNET_B200_NOALLOC_MAX_ROWS; more is rejected (-1, lpcnet_b200_last_error()): split the call.
 * device -1 = the current device.  The first call on a device looks up the array's device address and uploads 4 KB
 * of tables (no allocation either); n_rows = 0 does only that, e.g. before a capture starts.  Every call shares the
 * array; each row clears its own part before use, so nothing is carried from call to call -- but calls whose kernels
 * OVERLAP on the GPU (different streams, no ordering) race on it. */
#define LPCNET_B200_NOALLOC_MAX_ROWS 4096
LPCNET_B200_EXPORT int lpcnet_b200_lpc_from_cepstrum_noalloc(int device, const float *d_cepstrum, int stride,
                                                             float *d_lpc, int n_rows, void *cuda_stream);

/* ------------------------------------------------------------------------------------
 * 3. Feature analysis (the step before the synthesis path)
 * ------------------------------------------------------------------------------------ */
/* The reference's per-frame analysis entry points (include/lpcnet.h:98-158, src/lpcnet_enc.c:467-486, 913-933):
 * 160 samples of 16 kHz PCM -> NB_TOTAL_FEATURES (36) floats = 18 cepstral coefficients, pitch period,
 * pitch correlation, 16 LPC coefficients.  One stream; the state lives on the GPU (device 0). */
typedef struct LPCNetEncState LPCNetEncState;
LPCNET_B200_EXPORT int lpcnet_encoder_get_size(void);
LPCNET_B200_EXPORT int lpcnet_encoder_init(LPCNetEncState *st);
LPCNET_B200_EXPORT LPCNetEncState *lpcnet_encoder_create(void);
LPCNET_B200_EXPORT void lpcnet_encoder_destroy(LPCNetEncState *st);
LPCNET_B200_EXPORT int lpcnet_compute_single_frame_features(LPCNetEncState *st, const short *pcm, float *features);
LPCNET_B200_EXPORT int lpcnet_compute_single_frame_features_float(LPCNetEncState *st, const float *pcm,
                                                                  float *features);

/* Batched form: n_streams independent analysis states, any number of frames per call (the state carries over,
 * so one call of n frames equals n calls of one frame).  pcm: [n_streams][n_frames*160]; features:
 * [n_streams][n_frames][36].  Returns 0, or -1 with lpcnet_b200_last_error() set.  No CPU fallback. */
typedef struct LPCNetFeatureBatch LPCNetFeatureBatch;
LPCNET_B200_EXPORT LPCNetFeatureBatch *lpcnet_feature_batch_create(int n_streams, int device);
LPCNET_B200_EXPORT void lpcnet_feature_batch_destroy(LPCNetFeatureBatch *b);
LPCNET_B200_EXPORT int lpcnet_feature_batch_n_streams(const LPCNetFeatureBatch *b);
LPCNET_B200_EXPORT long long lpcnet_feature_batch_launch_count(const LPCNetFeatureBatch *b);
/* lpcnet_encoder_init for one stream, or for all of them (stream = -1). */
LPCNET_B200_EXPORT int lpcnet_feature_batch_reset(LPCNetFeatureBatch *b, int stream);
LPCNET_B200_EXPORT int lpcnet_feature_batch_compute(LPCNetFeatureBatch *b, const short *pcm, int n_frames,
                                                    float *features);
LPCNET_B200_EXPORT int lpcnet_feature_batch_compute_float(LPCNetFeatureBatch *b, const float *pcm, int n_frames,
                                                          float *features);
/* Device pointers (exactly one of d_pcm16 / d_pcmf non-NULL); enqueued on cuda_stream (a cudaStream_t, NULL = the
 * batch's own stream) without synchronising. */
LPCNET_B200_EXPORT int lpcnet_feature_batch_compute_device(LPCNetFeatureBatch *b, const short *d_pcm16,
                                                           const float *d_pcmf, int n_frames, float *d_features,
                                                           void *cuda_stream);

/* 1.6 kb/s encoder, the wire format's other half.  The reference's entry points
 * (include/lpcnet.h:98-131, src/lpcnet_enc.c:882-911): 640 samples -> LPCNET_COMPRESSED_SIZE (8) bytes
 * (lpcnet_encode), or -> 4 x 36 unquantised features with the superframe pitch tracker (lpcnet_compute_features;
 * features = float[4][36]).  The reference links its VQ tables (ceps_codebooks.c) into the library; here they are
 * given once per process (lpcnet_b200_set_codebooks: ceps_codebook1|2|3 [1024][17] then ceps_codebook_diff4
 * [4096][18], 125952 floats, copied; encoder states created afterwards use them) or per batch. */
LPCNET_B200_EXPORT int lpcnet_b200_set_codebooks(const float *codebooks, int n_floats);
LPCNET_B200_EXPORT int lpcnet_encode(LPCNetEncState *st, const short *pcm, unsigned char *buf);
LPCNET_B200_EXPORT int lpcnet_compute_features(LPCNetEncState *st, const short *pcm, float *features);
/* Batched: pcm [n_streams][n_packets*640] -> packets [n_streams][n_packets][8]; quantised_features (nullable)
 * [n_streams][n_packets*4][36] receives the encoder's quantised rows, entries [0..20) (= what lpcnet_decode's
 * decode_packet rebuilds from the packet; entries [20..36) hold the LPC of the unquantised cepstrum). */
LPCNET_B200_EXPORT int lpcnet_feature_batch_set_codebooks(LPCNetFeatureBatch *b, const float *codebooks, int n_floats);
LPCNET_B200_EXPORT int lpcnet_feature_batch_encode(LPCNetFeatureBatch *b, const short *pcm, int n_packets,
                                                   unsigned char *packets, float *quantised_features);
/* lpcnet_compute_features batched: features [n_streams][n_packets*4][36]. */
LPCNET_B200_EXPORT int lpcnet_feature_batch_compute_features(LPCNetFeatureBatch *b, const short *pcm, int n_packets,
                                                             float *features);

/* ------------------------------------------------------------------------------------
 * 4. Packet-loss concealment: non-causal, causal and codec modes (BASELINE config 4)
 * ------------------------------------------------------------------------------------ */
/* The reference's PLC entry points (include/lpcnet.h:190-215, src/lpcnet_plc.c:44-109, 493-503).  options =
 * LPCNET_PLC_CAUSAL (0), LPCNET_PLC_NONCAUSAL (1) or LPCNET_PLC_CODEC (2), optionally | LPCNET_PLC_DC_FILTER (4); 3
 * returns -1.  The blob must carry the plc_* arrays next to the LPCNet model.  The non-causal mode runs the model with
 * FEATURES_DELAY 0, as src/lpcnet_plc.c:357-361 demands, and its output lags the input by 80 samples; the causal and
 * codec modes run it with the delay of lpcnet_b200_set_model_constants (default 2).  pcm: 160 samples in (update) /
 * out (both). */
typedef struct LPCNetPLCState LPCNetPLCState;
LPCNET_B200_EXPORT int lpcnet_plc_get_size(void);
LPCNET_B200_EXPORT int lpcnet_plc_init(LPCNetPLCState *st, int options);
LPCNET_B200_EXPORT LPCNetPLCState *lpcnet_plc_create(int options);
LPCNET_B200_EXPORT void lpcnet_plc_destroy(LPCNetPLCState *st);
LPCNET_B200_EXPORT int lpcnet_plc_load_model(LPCNetPLCState *st, const unsigned char *data, int len);
/* lpcnet_plc_reset (src/lpcnet_plc.c:46-60): back to the state after lpcnet_plc_init + lpcnet_plc_load_model. */
LPCNET_B200_EXPORT void lpcnet_plc_reset(LPCNetPLCState *st);
/* The FEC queue (include/lpcnet.h:213-215, src/lpcnet_plc.c:111-132): read by the causal and codec modes
 * (get_fec_or_pred); in the non-causal mode -- as in the reference -- it does not change the output. */
LPCNET_B200_EXPORT void lpcnet_plc_fec_add(LPCNetPLCState *st, const float *features);
LPCNET_B200_EXPORT void lpcnet_plc_fec_clear(LPCNetPLCState *st);
LPCNET_B200_EXPORT int lpcnet_plc_update(LPCNetPLCState *st, short *pcm);
LPCNET_B200_EXPORT int lpcnet_plc_conceal(LPCNetPLCState *st, short *pcm);

/* Batched form: n_streams independent PLC states, each with its own loss pattern.  lost: [n_streams][n_frames]
 * (non-zero = the 10 ms frame is lost -> lpcnet_plc_conceal, else lpcnet_plc_update); pcm: [n_streams][n_frames*160]
 * in/out.  Streams are seeded like lpcnet_batch_set_seed. */
typedef struct LPCNetPLCBatch LPCNetPLCBatch;
LPCNET_B200_EXPORT LPCNetPLCBatch *lpcnet_plc_batch_create(const unsigned char *blob, int len, int n_streams, int device,
                                                           int options, float lpc_gamma);
LPCNET_B200_EXPORT void lpcnet_plc_batch_destroy(LPCNetPLCBatch *b);
LPCNET_B200_EXPORT int lpcnet_plc_batch_n_streams(const LPCNetPLCBatch *b);
LPCNET_B200_EXPORT int lpcnet_plc_batch_set_seed(LPCNetPLCBatch *b, int stream, const unsigned char *data, int n);
LPCNET_B200_EXPORT long long lpcnet_plc_batch_launch_count(const LPCNetPLCBatch *b);
LPCNET_B200_EXPORT int lpcnet_plc_batch_reset(LPCNetPLCBatch *b);
/* lpcnet_plc_fec_add / _fec_clear for one stream of the batch (features: 20 floats, NULL = "skip one"); the vectors
 * a call of lpcnet_plc_batch_process consumes must have been added before it. */
LPCNET_B200_EXPORT int lpcnet_plc_batch_fec_add(LPCNetPLCBatch *b, int stream, const float *features);
LPCNET_B200_EXPORT int lpcnet_plc_batch_fec_clear(LPCNetPLCBatch *b, int stream);
LPCNET_B200_EXPORT int lpcnet_plc_batch_process(LPCNetPLCBatch *b, const unsigned char *lost, short *pcm, int n_frames);
/* Frame-rate state of one stream for step-by-step checks: st->features[0..20), the predictor's two GRU states
 * (256 each), {dc_mem, syn_dc}; any pointer may be NULL. */
LPCNET_B200_EXPORT int lpcnet_plc_batch_peek(LPCNetPLCBatch *b, int stream, float *features20, float *gru1, float *gru2,
                                             double *dc2);
/* The last processed frame's burg_cepstral_analysis output and analysis features of one stream (good frames). */
LPCNET_B200_EXPORT int lpcnet_plc_batch_peek_frame(LPCNetPLCBatch *b, int stream, float *burg36, float *enc36);

/* ------------------------------------------------------------------------------------
 * 5. DRED: the RDOVAE decoder that turns received latent vectors into FEC features
 * ------------------------------------------------------------------------------------ */
/* DRED_rdovae_decode_all (include/dred_rdovae.h:40, src/dred_rdovae.c:37-51 over src/dred_rdovae_dec.c:37-98): the
 * 24-value initial state and nb_latents latent vectors of 80 floats -> nb_latents x 4 feature vectors of 20 floats
 * (each group of four in the reference's reverse order), ready for lpcnet_plc_fec_add.  The reference's RDOVAEDec model
 * struct is replaced by a handle created from a weights blob that carries the decoder arrays (dec_dense1..8, dec_final,
 * state1..3 in the layout of torch/rdovae/export_rdovae_weights.py; write_lpcnet_weights.c writes them into the same
 * blob as the LPCNet model).  Sizes are those of the reference's default export: cond_size 256, latent 80, state 24. */
typedef struct LPCNetDredDecoder LPCNetDredDecoder;
LPCNET_B200_EXPORT LPCNetDredDecoder *lpcnet_dred_decoder_create(const unsigned char *blob, int len, int device);
LPCNET_B200_EXPORT void lpcnet_dred_decoder_destroy(LPCNetDredDecoder *d);
LPCNET_B200_EXPORT long long lpcnet_dred_decoder_launch_count(const LPCNetDredDecoder *d);
LPCNET_B200_EXPORT int lpcnet_dred_decode_all(LPCNetDredDecoder *d, float *features, const float *state,
                                              const float *latents, int nb_latents);
/* n_streams independent packets per call: states [n][24], latents [n][nb_latents][80] -> features [n][nb_latents][80] */
LPCNET_B200_EXPORT int lpcnet_dred_decode_batch(LPCNetDredDecoder *d, int n_streams, float *features, const float *states,
                                                const float *latents, int nb_latents);

/* The reference's tanh/sigmoid multiply by `rcpps` (src/vec_avx.h:408,437), a CPU
 * hardware estimate that is exactly a 2048-entry table over the top 11 mantissa bits.
 * Default: the Intel table model (oracle/orc_prim.c).  A process-wide override takes the
 * 2048 binary32 bit patterns of rcpps(1 + i/2048); applies to batches created later. */
LPCNET_B200_EXPORT void lpcnet_b200_set_rcp_table(const uint32_t *table2048);
LPCNET_B200_EXPORT void lpcnet_b200_default_rcp_table(uint32_t *table2048);

/* FEATURES_DELAY / LPC_GAMMA used by the drop-in lpcnet_load_model (compile-time constants of
 * the reference's generated nnet_data.h; default 2 and 1.0). */
LPCNET_B200_EXPORT void lpcnet_b200_set_model_constants(int features_delay, float lpc_gamma);
/* The reference's -DEND2END build (src/lpcnet.c:57-78, 108: the frame's LPC are rc2lpc of the first 16 outputs of
 * feature_dense2 instead of lpc_from_cepstrum through the FEATURES_DELAY line) as a run-time switch: process-wide
 * default for states / batches created later (0), and per batch. */
LPCNET_B200_EXPORT void lpcnet_b200_set_end2end(int on);
LPCNET_B200_EXPORT int lpcnet_batch_set_end2end(LPCNetBatch *b, int on);

/* Last error message of the calling thread ("" if none); sticky until read. */
LPCNET_B200_EXPORT const char *lpcnet_b200_last_error(void);
LPCNET_B200_EXPORT int lpcnet_b200_device_count(void);

#ifdef __cplusplus
}
#endif
#endif
