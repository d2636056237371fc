// Internal structures shared by the host shim and the CUDA kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>

namespace lpcb {

// Default model dimensions (training_tf2/lpcnet.py:312-325).  The blob loader
// checks every array against them, exactly like the reference's generated init function.
constexpr int kN = 384;        // GRU_A_STATE_SIZE
constexpr int kNB = 16;        // GRU_B_STATE_SIZE
constexpr int kC = 128;        // FEATURE_DENSE2_OUT_SIZE
constexpr int kPitchEmb = 64;  // EMBED_PITCH_OUT_SIZE
constexpr int kFeat = 20;      // NB_FEATURES
constexpr int kConvIn = kFeat + kPitchEmb;  // 84
constexpr int kLpcOrder = 16;
constexpr int kFrame = 160;
constexpr int kMaxDelay = 2;
constexpr int kRowsA = 3 * kN;   // 1152
constexpr int kRowsB = 3 * kNB;  // 48

// u8 activation rows in shared memory, word stride = 4 (mod 32): in a 4-byte gather of an MMA A
// fragment lane (g, t) reads word 100*g + w[t], i.e. bank 4*g + w[t]; the eight streams of one t
// cover the banks of one residue class mod 4, so a gather is conflict-free when its four block
// positions have distinct residues -- model.cu orders every chunk's blocks that way.
constexpr int kXaStride = kN / 4 + 4;    // words per stream (100)
constexpr int kXbStride = kNB / 4;       // 4: eight streams x four words = 32 distinct banks
constexpr int kFcStride = 36;            // dual_fc node stride in shared memory: 32 weights padded so that 8 lanes' 16-byte reads hit disjoint banks

// GRU-B input product (48 x 384) in the sample kernel: warp w of the 16 forms the K slice of its own 24 units (unit
// groups w, w+16, w+32) straight from the state bytes it has just produced, one mma.m16n8k32 per 8-row group with the
// K order [ug w: 2t 2t+1 | ug w+16: 2t 2t+1] (k = 4t..4t+3) and [ug w+32: 2t 2t+1 | 0 0] (k = 16+4t..), t = lane % 4;
// DevModel::b_in_frag holds the matching B fragments [warp][row group][lane] (absent blocks are zeros).
constexpr int kBinFragWords = 16 * 6 * 32 * 2;
// (the block-list form of the same matrix, DevModel::b_in, keeps these packing parameters; the kernel no longer reads it)
constexpr int kPartsB = 2;
constexpr int kPadB = 6;

// Column order of the GRU-A input rows (embedding tables, frame conditioning) in device memory: the six
// values lane t of a quad needs for unit group ug share one 32-byte sector, [ug][t][z0 z1 r0 r1 h0 h1 - -]:
// one 256-bit load per lane fetches them, a warp request covers exactly one 128-byte line per stream, and a
// row pointer per table serves all unit groups at immediate offsets.  Rows are kRowsAP floats.
// col = gate*384 + unit  ->  storage index.
constexpr int kRowsAP = kN / 8 * 32;   // 1536
__host__ __device__ inline int gru_a_col_index(int col) {
  const int gate = col / kN, unit = col % kN;
  return ((unit >> 3) * 4 + ((unit & 7) >> 1)) * 8 + gate * 2 + (unit & 1);
}

// Block-sparse int8 matrix re-tiled for mma.sync.m16n8k32 B fragments.
// A "chunk" is 8 blocks (K = 32) of one 8-row group; its 64 weight words are stored [lane][2] so that a
// lane fetches both B registers with one 8-byte load.
struct PackedSparse {
  int n_groups = 0;   // rows / 8
  int n_chunks = 0;
  int n_blocks = 0;   // real (un-padded) blocks
  std::vector<uint32_t> chunk_start;  // [n_groups + 1]   (slotted GRU-A: [unit groups + 1])
  std::vector<uint8_t> pos;           // [n_chunks][8]  column position / 4
  std::vector<uint32_t> w;            // [n_chunks][32 lanes][2] words, fragment order
  // GRU-A only ("slotted" layout): unit group ug owns the fixed chunks [ug*ST, (ug+1)*ST), ST = slots[0]+
  // slots[1]+slots[2], gate-major inside; blocks that do not fit go to overflow chunks after the fixed region,
  // one list per unit group (chunk_start[ug] .. chunk_start[ug+1]) with the gate of each chunk in its offset words.
  // off: per chunk, per lane%4, the two gather byte offsets and the gate: lo | hi << 16 | gate << 30.
  // off_fixed: the same words of the fixed chunks regrouped [ug][lane%4][8] (ST <= 8 used) for 16-byte loads.
  int slots[3] = {0, 0, 0};
  std::vector<uint32_t> sat_start;    // saturating pairs (DevSat): [lists + 1]
  std::vector<uint2> sat;
  std::vector<uint32_t> off;          // [n_chunks][4]
  std::vector<uint32_t> off_fixed;    // [unit groups][4][8]
};

// vpmaddubsw (src/vec_avx.h:814-818) adds the two u8 x s8 products of an adjacent input pair in a SATURATING s16;
// integer MMA / dp4a sums cannot saturate.  A pair can only saturate when both weights have the same sign and
// 255*(|w0|+|w1|) > 32767; such pairs are taken out of the tensor-core operands at load (weights zeroed there) and
// listed here, and the kernels add sat16(u0*w0 + u1*w1) for them -- exact for every int8 model, free when the list
// is empty (clipped models).  entry.x = column | row_in_group << 16 | gate << 20, entry.y = w0 & 255 | (w1 & 255) << 8.
struct DevSat {
  int n;                  // entries in total (0: the kernels skip the pass)
  const uint32_t *start;  // [lists + 1]
  const uint2 *e;
};
__host__ __device__ inline int sat16_pair(uint2 f, const unsigned char *x) {
  const int col = f.x & 0xffff;
  const int w0 = (int)(signed char)(f.y & 0xffu), w1 = (int)(signed char)((f.y >> 8) & 0xffu);
  const int p = (int)x[col] * w0 + (int)x[col + 1] * w1;
  return p > 32767 ? 32767 : (p < -32768 ? -32768 : p);
}

struct DevSparse {
  DevSat sat;                   // lists per unit group (GRU-A) or per 8-row group
  int n_groups, n_chunks;
  const uint32_t *chunk_start;  // device
  const uint2 *pos;             // device, [n_chunks]
  const uint32_t *w;            // device, [n_chunks][64]
  const uint32_t *off;          // device, [n_chunks][4]
  const uint32_t *off_fixed;    // device, [unit groups][4][8] (GRU-A only)
  int slots[3];                 // fixed chunk slots per gate per unit group (0,0,0 = all chunks in the overflow lists)
};

// PLC predictor (src/lpcnet_plc.c:135-145; plc_data.c): dense 57->128, two int8 GRUs of 256 units, dense 256->20.
// GRU matrices re-tiled as words of four consecutive inputs, [input/4][3*256 rows]: thread = row, coalesced.
constexpr int kPlcIn = 57, kPlcCond = 128, kPlcUnits = 256, kPlcRows = 3 * kPlcUnits;
struct DevPlc {
  DevSat sat_g1_in, sat_g1_rec, sat_g2_in, sat_g2_rec;   // per output row (lists = 768), entry.x = column
  const float *dense1_w, *dense1_b;     // [57][128], [128]
  const float *out_w, *out_b;           // [256][20], [20]
  const uint32_t *g1_in, *g1_rec;       // [128/4][768], [256/4][768]
  const uint32_t *g2_in, *g2_rec;       // [256/4][768], [256/4][768]
  const float *g1_subias, *g2_subias;   // [2][768]
};

// RDOVAE decoder of DRED (src/dred_rdovae_dec.c): cond_size = cond_size2 = 256, i.e. the PLC predictor's GRU geometry
constexpr int kDredLatent = 80, kDredState = 24, kDredOut = 80;
struct DevDred {
  DevSat sat_in[3], sat_rec[3];             // per output row
  const float *state_w[3], *state_b[3];     // state1..3: [24][256], [256]
  const float *dense_w[5], *dense_b[5];     // dec_dense1 [80][256]; dec_dense3/5/7/8 [256][256]
  const float *final_w, *final_b;           // dec_final [2048][80], [80]
  const uint32_t *g_in[3], *g_rec[3];       // dec_dense2/4/6: [256/4][768] words each
  const float *g_subias[3];                 // [2][768]
};

// Device-side model: pointers into one device allocation.
struct DevModel {
  // 100 Hz frame network (fp32, blob layout)
  const float *embed_pitch;                 // [256][64]
  const float *conv1_w, *conv1_b;           // [3*84][128], [128]
  const float *conv2_w, *conv2_b;           // [3*128][128], [128]
  const float *dense1_w, *dense1_b, *dense2_w, *dense2_b;
  const float *a_feat_w, *a_feat_b;         // [128][1152], [1152]
  const float *b_feat_w, *b_feat_b;         // [128][48], [48]
  // 16 kHz sample network
  const float *e_sig, *e_pred, *e_exc;      // [256][kRowsAP], columns in gru_a_col_index order
  const float *a_subias_rec;                // [1152] = &sparse_gru_a_subias[3N]
  const float *a_diag;                      // [1152]
  const float *a_sd_packed;                 // [gate*48 + ug][lane%4][sub0 sub1 diag0 diag1]: the kernel's shared-memory image
  const float *fc_w_padded;                 // dual_fc weights at node stride kFcStride (36 floats): shared-memory image
  const float *fc_bf;                       // dual_fc bias [2][256] | factor [2][256]: shared-memory image
  DevSparse a_rec;                          // 1152 x 384 block-sparse
  const float *b_subias;                    // [96]
  DevSparse b_in;                           // 48 x 384 (sparse stream format, dense by default): block count, DevSat lists
  const uint2 *b_in_frag;                   // the same matrix as register-operand B fragments [16][6][32] (see kBinFragWords)
  DevSparse b_rec;                          // 48 x 16 dense
  const float *fc_w, *fc_bias, *fc_factor;  // [256][2][16], [2][256], [2][256]
  const uint32_t *rcp_tab;                  // [2048]
  int small_acc;                            // 1: every int32 sum of the sample network stays inside (-2^22, 2^22) (proven at load)
  int has_sat;                              // 1: some int8 pair can saturate vpmaddubsw's s16 (DevSat lists are not empty)
  int rcp_arith;                            // 1: the table equals mid-point MUFU.RCP rounded to 12 bits (checked on this GPU)
  const float *ulaw2lin_tab;                // [256]
  const float *logit_tab;                   // [256]
  // DSP tables
  const float *fft_tw;                      // [320][2]
  const int16_t *fft_bitrev;                // [320]
  const float *dct;                         // [18*18]
  const float *codebooks;                   // cb1|cb2|cb3|diff4 or NULL
  int features_delay;
  float lpc_gamma;
  int end2end;               // the reference's -DEND2END build: LPC = rc2lpc(condition[0..16)) (src/lpcnet.c:57-78, 108)
  DevPlc plc;                               // valid when Model::has_plc
  DevDred dred;                             // valid when Model::has_dred
};

// Per-stream mutable state, structure-of-arrays in HBM ([field][stream] where noted).
struct DevState {
  float *conv1_state;   // [B][168]
  float *conv2_state;   // [B][256]
  float *gru_a;         // [B][384]
  float *gru_b;         // [B][16]
  float *last_sig;      // [B][16]
  float *old_lpc;       // [B][2][16]
  float *lpc;           // [B][16]      (current frame)
  float *cond_a;        // [B][kRowsAP]  (current frame), gru_a_col_index order
  float *cond_b;        // [B][48]
  float *deemph;        // [B]
  int *last_exc;        // [B]
  int *frame_count;     // [B]
  uint32_t *rng;        // [B][4]
  float *vq_mem;        // [B][18]
};

struct Model {
  DevModel d{};
  void *dev_blob = nullptr;
  void *dev_codebooks = nullptr;   // tables installed after load (install_codebooks); NULL when they came with the blob
  size_t dev_bytes = 0;
  int nblocks_a = 0, nblocks_b = 0;
  bool has_codebooks = false;
  bool has_plc = false;   // the blob carried the plc_* arrays
  bool has_dred = false;  // ... the RDOVAE decoder arrays (dec_*, state1..3)
};

// Loads and validates a DNNw blob (src/parse_lpcnet_weights.c) and builds the device model.
// Returns false and fills `err` on failure.
bool load_model(Model &m, const unsigned char *blob, int len, int features_delay, float lpc_gamma,
                const uint32_t *rcp_tab, std::string &err);
void free_model(Model &m);
bool install_codebooks(Model &m, const float *cb, int n, std::string &err);

// Latches the thread's error text (lpcnet_b200_last_error) and returns -1.
int set_last_error(const std::string &msg);

void default_rcp_table(uint32_t *tab);
void host_dsp_tables(std::vector<float> &tw, std::vector<int16_t> &br, std::vector<float> &dct);

// Per-chunk scratch: conditioning for F frames.
struct FrameBuffers {
  int cap_frames = 0;
  int cap_streams = 0;       // the arrays below are laid out for batches of at most this many streams
  float *in_vec = nullptr;   // [B][F+2][84]   frame-network input rows incl. 2 history rows
  float *c1 = nullptr;       // [B][F+2][128]  conv1 outputs incl. 2 history rows
  float *cond_a = nullptr;   // [F][B][kRowsAP], gru_a_col_index order
  float *cond_b = nullptr;   // [F][B][48]
  float *lpc = nullptr;      // [F][B][16]
  uint8_t *active = nullptr; // [F][B]
  float *fft_scratch = nullptr;  // [2][320][2][B*F]  (ping-pong not needed: in-place) -> [320][2][B*F]
};

struct SampleArgs {
  DevModel m;
  DevState st;
  int n_streams;
  int stream_begin = 0;      // the launch advances streams [stream_begin, stream_end); stream_end 0 = n_streams
  int stream_end = 0;
  int n_frames;              // frames in this chunk
  int spf;                   // samples per frame (<=160)
  int frame0;                // index of the chunk's first frame within the call
  int total_frames;          // frames in the whole call (pcm row length = total_frames*spf)
  const float *cond_a;       // [F][B][kRowsAP], gru_a_col_index order
  const float *cond_b;       // [F][B][48]
  const float *lpc;          // [F][B][16]
  const uint8_t *active;     // [F][B]
  short *pcm;                // [B][total_frames*spf]
  const short *forced;       // nullable, same layout
  int forced_by_flag;        // 0: `forced` (if given) teacher-forces every active stream; 1: only streams whose
                             //    active flag has bit 1 set (PLC: forced and free-running streams in one launch)
  void *trace;               // nullable LPCNetB200Trace [B][total_frames*spf]
  float *logits;             // nullable [B][total_frames*spf][256]
  unsigned long long *phase_clk;   // development builds (-DLPCB_PHASE_CLOCKS) only: per-phase cycle sums of CTA 0
  // 1.0f, 1.0f -- as a kernel argument, i.e. a value ptxas cannot see: fma.rn.f32x2(p, one, c) is the exactly rounded
  // packed sum p + c, and unlike add.rn.f32x2 it is not contracted with the packed multiply that produced p
  // (ptxas fuses mul.rn.f32x2 -> add.rn.f32x2 into FFMA2, which would drop the product's rounding)
  float2 one2 = {1.f, 1.f};
};

// tile_streams 8 / 16 / 32, or 0: 16-stream tiles in whole waves of one tile per SM, and the remainder in 8-stream
// tiles when it fits one per SM (lower step latency than a part-filled wave of 16-stream tiles), else in 16s.
cudaError_t launch_sample_kernel(const SampleArgs &a, int tile_streams, int n_sms, cudaStream_t s,
                                 int *n_launches);
size_t sample_kernel_smem(int tile_streams, const DevModel &m);

struct FrameArgs {
  DevModel m;
  DevState st;
  FrameBuffers fb;
  int n_streams, n_frames;
  const float *features;     // device [B][total_frames][stride]
  int feature_stride;
  int frame0, total_frames;
  const uint8_t *stream_mask;  // nullable [B]: streams with 0 keep their frame-rate state and their conditioning rows
  const int *rows = nullptr;   // nullable device list of n_rows stream indices: only these streams run (instead of a mask:
  int n_rows = 0;              //   the kernels are launched over the list, not over the batch); layouts stay those of the batch
  int state_only = 0;          // run_frame_network_flush (lpcnet.c:134-144): only the frame-rate state moves
};
// fork (optional): an auxiliary CUDA stream and two events.  The cepstrum -> LPC kernel depends on the features only,
// not on the conv / dense chain, and both are latency-bound at small sizes: with a fork it runs on `aux` beside
// conv1 + cond (joined before the finalize kernel) instead of behind them.
struct FrameFork { cudaStream_t aux = nullptr; cudaEvent_t begin = nullptr, done = nullptr; };
cudaError_t launch_frame_network(const FrameArgs &a, cudaStream_t s, int *n_launches, const FrameFork *fork = nullptr);

struct DecodeArgs {
  DevModel m;
  DevState st;
  int n_streams, n_packets;
  const unsigned char *packets;  // device [B][n_packets][8]
  float *features;               // device [B][n_packets*4][36]
};
cudaError_t launch_decode_packets(const DecodeArgs &a, cudaStream_t s, int *n_launches);

}  // namespace lpcb

// Hooks the PLC driver (plc.cu) uses on the batched engine (api.cu) and the feature analysis (analysis.cu).
struct LPCNetBatch;
struct LPCNetFeatureBatch;
namespace lpcb {
struct BatchView {
  DevState st;
  const Model *model;
  cudaStream_t stream;
  int n_streams, device;
};
BatchView batch_view(LPCNetBatch *b);
int default_features_delay();   // lpcnet_b200_set_model_constants (default 2)
float default_lpc_gamma();      // ... (default 1.0)
int current_device();           // the calling thread's current CUDA device (0 if there is none to ask)
// the process-wide VQ tables of lpcnet_b200_set_codebooks (analysis.cu), empty if none were set
std::vector<float> default_codebooks();
// ... and the ceps_codebook1/2/3 + ceps_codebook_diff4 arrays of $LPCNET_B200_WEIGHTS or ./weights_blob.bin (the file
// that stands in for the reference's compiled-in model and tables), empty if there is no such file or no such arrays
std::vector<float> file_codebooks();
// run_frame_network for one frame of the streams with mask != 0 (features: device [B][stride]); the conditioning
// stays in the batch's frame buffers for the sample segments that follow.
// state_only: run_frame_network_flush -- the outputs go to a scratch set, the streams' conditioning stays.
// d_rows / n_rows (optional): the device list of the streams the mask selects -- the kernels then run over the list
// (the PLC's small classes: a tenth of the batch) instead of testing the mask for every stream.
int batch_frame_network_masked(LPCNetBatch *b, const float *d_features, int stride, const uint8_t *d_mask,
                               bool state_only = false, const int *d_rows = nullptr, int n_rows = 0,
                               cudaStream_t side = nullptr);
// n sample steps (lpcnet_synthesize_tail_impl) on the conditioning of the last frame network.  flags [B]: 0 the
// stream does not move, 1 free-running, 3 teacher-forced on d_forced [B][n]; d_out [B][n].
// d_perm (nullable): device list of the n_act streams whose flag is not 0, ascending; when they are at most half of
// the batch they are advanced as a dense batch of their own (gather, sample kernel, scatter).
// side (both functions): run on this CUDA stream instead of the batch's, with a gather buffer of its own, so that a
// segment of other streams can run on the batch's stream at the same time (the caller orders the two with events;
// the frame buffers are rows per stream, and the sample kernel stores the state of the streams it moved only).
int batch_sample_segment(LPCNetBatch *b, int n, const uint8_t *d_flags, const short *d_forced, short *d_out,
                         const int *d_perm = nullptr, int n_act = 0, cudaStream_t side = nullptr, int tile_override = 0);
// one frame of analysis for the streams with mask != 0: d_pcm16 [B][160] -> d_features [B][36] (others untouched)
// d_rows / n_rows (optional): the device list of those streams -- the kernels' grids then cover the list only
int feature_batch_frame_masked(LPCNetFeatureBatch *b, const short *d_pcm16, const uint8_t *d_mask, float *d_features,
                               cudaStream_t stream, const int *d_rows = nullptr, int n_rows = 0);

}  // namespace lpcb
