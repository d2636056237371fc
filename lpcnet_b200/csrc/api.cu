// C-ABI host shim: the reference's lpcnet.h entry points for the synthesis path and the batched
// multi-stream API, over the CUDA kernels in frame.cu / sample.cu.  No CPU fallback anywhere.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/lpcnet_b200.h"
#include "engine.h"

using namespace lpcb;

namespace {
thread_local std::string g_err;
}
int lpcb::set_last_error(const std::string &msg) {
  g_err = msg;
  return -1;
}

namespace {

#ifdef LPCB_PHASE_CLOCKS
unsigned long long *g_phase_clk = nullptr;
#endif
thread_local std::string g_err_ret;
std::mutex g_cfg_mu;
bool g_have_rcp = false;
uint32_t g_rcp[2048];
int g_default_delay = 2;
float g_default_gamma = 1.0f;
int g_default_end2end = 0;

int fail(const std::string &msg) { return lpcb::set_last_error(msg); }
int fail_cuda(const char *what, cudaError_t e) { return fail(std::string(what) + ": " + cudaGetErrorString(e)); }

#define CK(call)                                        \
  do {                                                  \
    cudaError_t e_ = (call);                            \
    if (e_ != cudaSuccess) return fail_cuda(#call, e_); \
  } while (0)

void lpcnet_kiss99_srand_host(uint32_t st[4], const unsigned char *d, int n) {
  // kiss99_srand (src/kiss99.c:32-57)
  auto next = [&]() {
    st[0] = 36969u * (st[0] & 0xFFFF) + (st[0] >> 16);
    st[1] = 18000u * (st[1] & 0xFFFF) + (st[1] >> 16);
    st[2] ^= st[2] << 13; st[2] ^= st[2] >> 17; st[2] ^= st[2] << 5;
    st[3] = 69069u * st[3] + 1234567u;
  };
  st[0] = 362436069u; st[1] = 521288629u; st[2] = 123456789u; st[3] = 380116160u;
  int i;
  for (i = 3; i < n; i += 4) {
    st[0] ^= d[i - 3]; st[1] ^= d[i - 2]; st[2] ^= d[i - 1]; st[3] ^= d[i];
    next();
  }
  if (i - 3 < n) st[0] ^= d[i - 3];
  if (i - 2 < n) st[1] ^= d[i - 2];
  if (i - 1 < n) st[2] ^= d[i - 1];
  if (st[0] == 0 || st[0] == 0x9068FFFFu) st[0]++;
  if (st[1] == 0 || st[1] == 0x464FFFFFu) st[1]++;
  if (st[2] == 0) st[2]++;
}

}  // namespace

struct LPCNetBatch {
  int device = 0;
  int n_streams = 0;
  int n_sms = 0;
  int tile = 0;  // 0 = auto
  int max_tile = 32;
  Model model;
  DevState st{};
  void *st_alloc = nullptr;
  FrameBuffers fb;                // frame-network outputs of one chunk ...
  void *fb_alloc = nullptr;
  FrameBuffers fb2;               // ... and of the next one: the frame network of chunk k+1 runs on its own stream
  void *fb2_alloc = nullptr;      //     beside the sample kernel of chunk k (it depends only on frame-rate state)
  size_t fb_rows = 0, fb2_rows = 0;   // stream-frames the two buffer sets hold
  cudaStream_t fstream = nullptr;
  cudaStream_t cstream = nullptr;                       // device->host copies of finished chunks (pinned destination)
  cudaEvent_t ev_chunk = nullptr, ev_copied = nullptr;
  cudaStream_t hstream = nullptr;                       // host->device copies of the chunks' feature rows (pinned source)
  cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_h2d_entry = nullptr;
  cudaEvent_t ev_entry = nullptr, ev_frames[2] = {nullptr, nullptr}, ev_samples[2] = {nullptr, nullptr};
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  long long launches = 0;
  bool profiling = false;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_sample, ev_frame;
  std::vector<int> ev_sample_steps, last_launch_steps;   // sample steps of every profiled sample-kernel launch
  std::vector<float> last_launch_ms;                     // their durations, as of the last read_profile
  // grow-only staging for the host-buffer entry points
  void *d_feat = nullptr; size_t d_feat_bytes = 0;
  void *d_pcm = nullptr; size_t d_pcm_bytes = 0;
  void *d_forced = nullptr; size_t d_forced_bytes = 0;
  void *d_trace = nullptr; size_t d_trace_bytes = 0;
  void *d_logits = nullptr; size_t d_logits_bytes = 0;
  void *d_packets = nullptr; size_t d_packets_bytes = 0;
  // compact copy of the streams one PLC sample segment moves (batch_sample_segment with a stream list)
  void *d_compact = nullptr; size_t d_compact_bytes = 0;
  void *d_compact2 = nullptr; size_t d_compact2_bytes = 0;   // ... of a segment on a side stream (beside one on `stream`)
  std::vector<std::pair<cudaStream_t, FrameFork>> forks;   // per CUDA stream that launches frame networks (fork_for)
  size_t l2_window_max = 0;       // > 0: L2's persisting carve-out in bytes; a chunk's conditioning rows are kept in it
  size_t l2_window_bytes = 0;     // the device's largest access-policy window
};

namespace {

// One PLC sample segment usually moves a small part of the batch (the streams that come out of a loss burst).  The
// sample kernel walks whole tiles in lock step, so those streams are gathered into a dense batch of their own
// (state, conditioning, flags, teacher-forcing samples), advanced there, and scattered back.
struct CompactView {
  DevState st;            // gru_a, gru_b, last_sig, deemph, last_exc, rng of the listed streams
  float *cond_a, *cond_b, *lpc;
  uint8_t *flags;
  short *forced, *out;    // [n_act][n]
};

__global__ void compact_gather_kernel(CompactView c, DevState st, const float *cond_a, const float *cond_b, const float *lpc,
                                      const uint8_t *flags, const short *forced, const int *perm, int n) {
  const int j = blockIdx.x, t = threadIdx.x, s = perm[j];
  for (int i = t; i < kRowsAP / 4; i += blockDim.x)
    reinterpret_cast<float4 *>(c.cond_a + (size_t)j * kRowsAP)[i] = reinterpret_cast<const float4 *>(cond_a + (size_t)s * kRowsAP)[i];
  for (int i = t; i < kN; i += blockDim.x) c.st.gru_a[(size_t)j * kN + i] = st.gru_a[(size_t)s * kN + i];
  if (t < kRowsB) c.cond_b[(size_t)j * kRowsB + t] = cond_b[(size_t)s * kRowsB + t];
  if (t < kNB) c.st.gru_b[(size_t)j * kNB + t] = st.gru_b[(size_t)s * kNB + t];
  if (t < 16) {
    c.st.last_sig[(size_t)j * 16 + t] = st.last_sig[(size_t)s * 16 + t];
    c.lpc[(size_t)j * 16 + t] = lpc[(size_t)s * 16 + t];
  }
  if (t < 4) c.st.rng[(size_t)j * 4 + t] = st.rng[(size_t)s * 4 + t];
  if (t == 0) { c.st.deemph[j] = st.deemph[s]; c.st.last_exc[j] = st.last_exc[s]; c.flags[j] = flags[s]; }
  if (forced) for (int i = t; i < n; i += blockDim.x) c.forced[(size_t)j * n + i] = forced[(size_t)s * n + i];
}

__global__ void compact_scatter_kernel(CompactView c, DevState st, short *out, const int *perm, int n) {
  const int j = blockIdx.x, t = threadIdx.x, s = perm[j];
  for (int i = t; i < kN; i += blockDim.x) st.gru_a[(size_t)s * kN + i] = c.st.gru_a[(size_t)j * kN + i];
  if (t < kNB) st.gru_b[(size_t)s * kNB + t] = c.st.gru_b[(size_t)j * kNB + t];
  if (t < 16) st.last_sig[(size_t)s * 16 + t] = c.st.last_sig[(size_t)j * 16 + t];
  if (t < 4) st.rng[(size_t)s * 4 + t] = c.st.rng[(size_t)j * 4 + t];
  if (t == 0) { st.deemph[s] = c.st.deemph[j]; st.last_exc[s] = c.st.last_exc[j]; }
  for (int i = t; i < n; i += blockDim.x) out[(size_t)s * n + i] = c.out[(size_t)j * n + i];
}

int grow(void **p, size_t *have, size_t want) {
  if (*have >= want) return 0;
  if (*p) cudaFree(*p);
  *p = nullptr; *have = 0;
  CK(cudaMalloc(p, want));
  *have = want;
  return 0;
}

int alloc_state(LPCNetBatch *b) {
  const size_t B = b->n_streams;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
  const size_t o_c1 = take(B * 168 * 4), o_c2 = take(B * 256 * 4), o_a = take(B * kN * 4), o_b = take(B * kNB * 4);
  const size_t o_ls = take(B * 16 * 4), o_ol = take(B * kMaxDelay * 16 * 4), o_lpc = take(B * 16 * 4);
  const size_t o_ca = take(B * kRowsAP * 4), o_cb = take(B * kRowsB * 4), o_de = take(B * 4), o_le = take(B * 4);
  const size_t o_fc = take(B * 4), o_rng = take(B * 16), o_vq = take(B * 18 * 4);
  CK(cudaMalloc(&b->st_alloc, off));
  CK(cudaMemset(b->st_alloc, 0, off));
  char *p = static_cast<char *>(b->st_alloc);
  b->st.conv1_state = (float *)(p + o_c1); b->st.conv2_state = (float *)(p + o_c2);
  b->st.gru_a = (float *)(p + o_a); b->st.gru_b = (float *)(p + o_b); b->st.last_sig = (float *)(p + o_ls);
  b->st.old_lpc = (float *)(p + o_ol); b->st.lpc = (float *)(p + o_lpc); b->st.cond_a = (float *)(p + o_ca);
  b->st.cond_b = (float *)(p + o_cb); b->st.deemph = (float *)(p + o_de); b->st.last_exc = (int *)(p + o_le);
  b->st.frame_count = (int *)(p + o_fc); b->st.rng = (uint32_t *)(p + o_rng); b->st.vq_mem = (float *)(p + o_vq);
  return 0;
}

// A call is walked in chunks of at most 32 frames (the sample kernel keeps a tile's state in registers / shared
// memory for a whole chunk, the frame network evaluates the chunk's frames side by side).  The conditioning of a
// chunk is 6400 B per stream-frame; instead of shrinking the chunk when the batch is large (1 frame per launch at
// 65 536 streams: the tile state went through HBM every 160 steps and the frame network was never batched over
// frames), the batch is walked in GROUPS of streams that each get full 32-frame chunks: a group is a whole number
// of 16-stream-tile waves (n_sms * 16 streams) and its chunk buffer stays under kChunkBytes.
constexpr long long kChunkBytes = 1ll << 30;
constexpr int kChunkFrames = 32;

// Round 2: the chunk is sized for L2 instead.  The frame network writes a chunk's GRU-A conditioning (6 KB per
// stream-frame) and the sample kernel reads it back; at 32 frames x 1024 streams that is 201 MB per buffer and every
// byte made the round trip through HBM (ncu: 53 + 91 B per output sample of DRAM traffic for the two kernels against
// 2.5 algorithmic).  With both buffer sets of the pipeline inside L2's persisting carve-out (cudaLimitPersistingL2-
// CacheSize, an access-policy window on the writing and on the reading CUDA stream) the rows never leave the chip:
// 5.1 + 2.4 B per sample measured at 4 frames x 1024 streams, for 0.5 % of throughput (8x the launches).
// Only where it is (nearly) free: batches whose two buffer sets of >= 2 frames fit two thirds of the carve-out (2140 streams
// on a B200) -- the latency regime, where the frame network of the next chunk runs on the SMs the sample kernel leaves
// idle.  A saturated batch has no idle SM: its frame network runs between sample-kernel launches, and short chunks
// cost 5.7 % there (measured at 4736 and 65 536 streams, one-wave groups of 2 frames), so those keep the 32-frame
// chunks and their 140 B per sample of HBM traffic (50 GB/s, 0.8 % of the bandwidth).  Calls shorter than
// kL2MinFrames (frame-by-frame streaming) keep one chunk per call.
// LPCNET_B200_L2_PERSIST_MB sets the carve-out (default: the device's maximum, 79 MB of 126); 0 = the round-1 chunking.
constexpr int kL2MinFrames = 8;
constexpr long long kCondRowBytes = (long long)kRowsAP * 4;

long long l2_budget(const LPCNetBatch *b) { return (long long)(b->l2_window_max / 3 * 2); }   // (measured: fuller evicts)
bool l2_sized(const LPCNetBatch *b, int n_frames) {
  return b->l2_window_max > 0 && n_frames >= kL2MinFrames && b->n_streams <= l2_budget(b) / (4 * kCondRowBytes);
}

int l2_window(const LPCNetBatch *b, cudaStream_t st, const void *base, size_t bytes) {
  if (!b->l2_window_max) return 0;
  cudaStreamAttrValue v{};
  v.accessPolicyWindow.base_ptr = const_cast<void *>(base);
  v.accessPolicyWindow.num_bytes = bytes < b->l2_window_bytes ? bytes : b->l2_window_bytes;
  // both buffer sets share the carve-out; a window that is larger (short calls on large batches) persists in part
  v.accessPolicyWindow.hitRatio = 2 * bytes <= b->l2_window_max ? 1.f : (float)b->l2_window_max / (float)(2 * bytes);
  v.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
  v.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
  CK(cudaStreamSetAttribute(st, cudaStreamAttributeAccessPolicyWindow, &v));
  return 0;
}

// streams per group for a call of n_frames frames (n_streams = no grouping)
int group_streams(const LPCNetBatch *b, int n_frames) {
  const int wave = b->n_sms * 16;
  if (l2_sized(b, n_frames) && !getenv("LPCNET_B200_CHUNK_BYTES")) return b->n_streams;
  const int Fc = n_frames < kChunkFrames ? n_frames : kChunkFrames;
  long long bytes = kChunkBytes;
  if (const char *e = getenv("LPCNET_B200_CHUNK_BYTES")) bytes = atoll(e) > 0 ? atoll(e) : bytes;   // tests: small groups
  const long long cap = bytes / ((long long)Fc * 6400);
  if (cap >= b->n_streams) return b->n_streams;
  long long g = cap / wave * wave;
  if (g < wave) g = wave;
  return (int)g;
}

// frames per chunk for groups of G streams
int chunk_frames(const LPCNetBatch *b, int n_frames, int G) {
  int fc = kChunkFrames;
  if (l2_sized(b, n_frames) && !getenv("LPCNET_B200_CHUNK_BYTES")) {
    const long long fit = l2_budget(b) / (2 * kCondRowBytes * G);
    fc = (int)(fit < 1 ? 1 : fit > kChunkFrames ? kChunkFrames : fit);
  }
  if (const char *e = getenv("LPCNET_B200_CHUNK_FRAMES")) fc = atoi(e) > 0 ? atoi(e) : fc;
  return n_frames < fc ? n_frames : fc;
}

// the stream state of streams [s0, ...) as a batch of its own: every array is [stream][field]
DevState state_view(const DevState &S, size_t s0) {
  DevState v = S;
  v.conv1_state += s0 * 168; v.conv2_state += s0 * 256; v.gru_a += s0 * kN; v.gru_b += s0 * kNB;
  v.last_sig += s0 * 16; v.old_lpc += s0 * kMaxDelay * 16; v.lpc += s0 * 16; v.cond_a += s0 * kRowsAP;
  v.cond_b += s0 * kRowsB; v.deemph += s0; v.last_exc += s0; v.frame_count += s0; v.rng += s0 * 4; v.vq_mem += s0 * 18;
  return v;
}

// buffers for F frames of `streams` streams (0 = the whole batch); they only grow
int ensure_frame_buffers(LPCNetBatch *b, int F, bool second = false, int streams = 0) {
  FrameBuffers &fb = second ? b->fb2 : b->fb;
  void *&alloc = second ? b->fb2_alloc : b->fb_alloc;
  size_t &have = second ? b->fb2_rows : b->fb_rows;
  const size_t B = streams > 0 ? streams : b->n_streams;
  // every array is a product of the stream count with F or F + 2: both products bound all of them
  if (have >= B * (size_t)F && (size_t)fb.cap_streams * (fb.cap_frames + 2) >= B * (size_t)(F + 2)) return 0;
  if (alloc) cudaFree(alloc);
  alloc = nullptr; fb.cap_frames = 0; fb.cap_streams = 0; have = 0;
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
  const size_t o_iv = take(B * (F + 2) * kConvIn * 4), o_c1 = take(B * (F + 2) * kC * 4);
  const size_t o_ca = take((size_t)F * B * kRowsAP * 4), o_cb = take((size_t)F * B * kRowsB * 4);
  const size_t o_lpc = take((size_t)F * B * 16 * 4), o_act = take((size_t)F * B);
  const size_t o_fft = take(16);       // (the transforms' scratch lives in shared memory now: dsp.cuh block_scratch)
  CK(cudaMalloc(&alloc, off));
  char *p = static_cast<char *>(alloc);
  fb.in_vec = (float *)(p + o_iv); fb.c1 = (float *)(p + o_c1); fb.cond_a = (float *)(p + o_ca);
  fb.cond_b = (float *)(p + o_cb); fb.lpc = (float *)(p + o_lpc); fb.active = (uint8_t *)(p + o_act);
  fb.fft_scratch = (float *)(p + o_fft);
  fb.cap_frames = F; fb.cap_streams = (int)B; have = B * (size_t)F;
  return 0;
}

int auto_tile(const LPCNetBatch *b) {
  if (b->tile) return b->tile;
  const int B = b->n_streams, sms = b->n_sms;
  // 16 streams per CTA is the throughput optimum (2.97e8 samples/s at 2368+ streams against 2.91e8 for 32 and
  // 2.45e8 for 8, measured); batches that fit one 8-stream tile per SM get the lowest step latency from 8
  int t = 16;
  if (B <= 8 * sms) t = 8;
  return t < b->max_tile ? t : b->max_tile;
}

// the auxiliary stream + events of the frame networks launched on `st` (engine.h FrameFork); nullptr = no fork
const FrameFork *fork_for(LPCNetBatch *b, cudaStream_t st) {
  if (getenv("LPCNET_B200_NO_FORK")) return nullptr;
  for (auto &f : b->forks) if (f.first == st) return &f.second;
  FrameFork f;
  int lo = 0, hi = 0;
  if (cudaDeviceGetStreamPriorityRange(&lo, &hi) != cudaSuccess ||
      cudaStreamCreateWithPriority(&f.aux, cudaStreamNonBlocking, hi) != cudaSuccess ||
      cudaEventCreateWithFlags(&f.begin, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&f.done, cudaEventDisableTiming) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  b->forks.emplace_back(st, f);
  return &b->forks.back().second;
}

// frame network for frames [f0, f0+F) of the call, streams [s0, s0+G) of the batch, on stream st; results left in fb
// (laid out [F][G][...]).  d_features points at the whole batch's rows.
int run_frames(LPCNetBatch *b, const float *d_features, int stride, int f0, int F, int total_frames,
               const FrameBuffers &fb, cudaStream_t st, int s0 = 0, int G = 0) {
  if (G <= 0) G = b->n_streams - s0;
  FrameArgs fa{};
  fa.m = b->model.d; fa.st = state_view(b->st, s0); fa.fb = fb; fa.n_streams = G; fa.n_frames = F;
  fa.features = d_features + (size_t)s0 * total_frames * stride; fa.feature_stride = stride; fa.frame0 = f0; fa.total_frames = total_frames;
  int nl = 0;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (b->profiling) { CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventRecord(e0, st)); }
  CK(launch_frame_network(fa, st, &nl, fork_for(b, st)));
  if (b->profiling) { CK(cudaEventRecord(e1, st)); b->ev_frame.emplace_back(e0, e1); }
  b->launches += nl;
  // keep the last frame's conditioning in the stream state (lpcnet_private.h:43-47)
  const size_t B = G;
  CK(cudaMemcpyAsync(fa.st.cond_a, fb.cond_a + (size_t)(F - 1) * B * kRowsAP, B * kRowsAP * 4, cudaMemcpyDeviceToDevice, st));
  CK(cudaMemcpyAsync(fa.st.cond_b, fb.cond_b + (size_t)(F - 1) * B * kRowsB, B * kRowsB * 4, cudaMemcpyDeviceToDevice, st));
  return 0;
}
int run_frames(LPCNetBatch *b, const float *d_features, int stride, int f0, int F, int total_frames) {
  return run_frames(b, d_features, stride, f0, F, total_frames, b->fb, b->stream);
}

bool is_pinned_host(const void *p) {
  cudaPointerAttributes at{};
  if (p && cudaPointerGetAttributes(&at, p) == cudaSuccess) return at.type == cudaMemoryTypeHost;
  cudaGetLastError();
  return false;
}

// h_pcm (optional, pinned host memory [B][n_frames * spf]): every chunk's samples are copied out on a third stream
// as soon as its sample kernel is done, beside the next chunk's kernel; the caller's stream ends after the last copy.
// h_features (optional, pinned host memory [B][n_frames][stride], with d_features the device buffer of that shape to
// fill): every chunk's feature rows are copied in on a fourth stream just ahead of its frame network, so that the call
// starts computing after the first chunk's rows instead of after the whole upload.
int synth_device(LPCNetBatch *b, const float *d_features, int stride, int n_frames, int spf, short *d_pcm,
                 const short *d_forced, LPCNetB200Trace *d_trace, float *d_logits, short *h_pcm = nullptr,
                 const float *h_features = nullptr) {
  if (!b) return fail("null batch");
  if (n_frames <= 0) return 0;
  if (spf <= 0 || spf > kFrame) return fail("samples_per_frame must be 1..160");
  if (stride < kFeat) return fail("feature_stride must be >= 20");
  CK(cudaSetDevice(b->device));
  // large batches: groups of streams, each walked in full chunks (see group_streams); trace / logit runs stay whole
  const int G = (d_trace || d_logits) ? b->n_streams : group_streams(b, n_frames);
  const int Fc = chunk_frames(b, n_frames, G);
  const int n_groups = (b->n_streams + G - 1) / G;
  if (ensure_frame_buffers(b, Fc, false, G)) return -1;
  // no explicit tile: whole waves of 16-stream tiles, the remainder in 8-stream tiles when that is faster (0 = mixed)
  const int tile = (b->tile == 0 && b->max_tile >= 16 && !getenv("LPCNET_B200_NO_MIX")) ? 0 : auto_tile(b);
  // Several chunks: the frame network of chunk k+1 runs on a second stream while the sample kernel works on
  // chunk k (two buffer sets).  It touches only frame-rate state (conv histories, frame counter, LPC delay
  // line); the sample kernel only sample-rate state.  Everything is ordered after the caller's stream at
  // entry, and the caller's stream ends after all of it.  The (group, chunk) pairs form one sequence of work items.
  const bool pipe = (n_frames > Fc || n_groups > 1) && !getenv("LPCNET_B200_NO_PIPELINE");
  if (pipe) {
    if (ensure_frame_buffers(b, Fc, true, G)) return -1;
    if (!b->fstream) {
      CK(cudaStreamCreateWithFlags(&b->fstream, cudaStreamNonBlocking));
      CK(cudaEventCreateWithFlags(&b->ev_entry, cudaEventDisableTiming));
      for (int k = 0; k < 2; k++) {
        CK(cudaEventCreateWithFlags(&b->ev_frames[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&b->ev_samples[k], cudaEventDisableTiming));
      }
    }
    CK(cudaEventRecord(b->ev_entry, b->stream));
    CK(cudaStreamWaitEvent(b->fstream, b->ev_entry, 0));
  }
  const size_t row_len = (size_t)n_frames * spf;
  if (h_features) {
    if (!b->hstream) {
      CK(cudaStreamCreateWithFlags(&b->hstream, cudaStreamNonBlocking));
      for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&b->ev_h2d[k], cudaEventDisableTiming));
      CK(cudaEventCreateWithFlags(&b->ev_h2d_entry, cudaEventDisableTiming));
    }
    CK(cudaEventRecord(b->ev_h2d_entry, b->stream));      // the feature buffer's last readers are ordered before it
    CK(cudaStreamWaitEvent(b->hstream, b->ev_h2d_entry, 0));
  }
  int item = 0;
  for (int s0 = 0; s0 < b->n_streams; s0 += G) {
    const int Gs = (b->n_streams - s0) < G ? (b->n_streams - s0) : G;
    for (int f0 = 0; f0 < n_frames; f0 += Fc, item++) {
      const int F = (n_frames - f0) < Fc ? (n_frames - f0) : Fc;
      const int buf = pipe ? (item & 1) : 0;
      const FrameBuffers &fbk = buf ? b->fb2 : b->fb;
      if (h_features) {
        // rows [s0, s0 + Gs), frames [f0, f0 + F) of [B][n_frames][stride]: a 2-D copy
        const size_t pitch = (size_t)n_frames * stride * 4, off = ((size_t)s0 * n_frames + f0) * stride;
        if (F == n_frames)      // whole rows: one contiguous block (a 2-D copy of short rows pays per row: 0.8 ms for
          CK(cudaMemcpyAsync(const_cast<float *>(d_features) + off, h_features + off, pitch * Gs,   // 12 500 rows of 144 B)
                             cudaMemcpyHostToDevice, b->hstream));
        else
          CK(cudaMemcpy2DAsync(const_cast<float *>(d_features) + off, pitch, h_features + off, pitch, (size_t)F * stride * 4, Gs,
                               cudaMemcpyHostToDevice, b->hstream));
        CK(cudaEventRecord(b->ev_h2d[item & 1], b->hstream));
        CK(cudaStreamWaitEvent(pipe ? b->fstream : b->stream, b->ev_h2d[item & 1], 0));
      }
      if (pipe) {
        if (item >= 2) CK(cudaStreamWaitEvent(b->fstream, b->ev_samples[buf], 0));   // the buffers are free again
        if (l2_window(b, b->fstream, fbk.cond_a, (size_t)F * Gs * kRowsAP * 4)) return -1;
        if (run_frames(b, d_features, stride, f0, F, n_frames, fbk, b->fstream, s0, Gs)) return -1;
        CK(cudaEventRecord(b->ev_frames[buf], b->fstream));
        CK(cudaStreamWaitEvent(b->stream, b->ev_frames[buf], 0));
      } else if (run_frames(b, d_features, stride, f0, F, n_frames, b->fb, b->stream, s0, Gs)) {
        return -1;
      }
      SampleArgs sa{};
      sa.m = b->model.d; sa.st = state_view(b->st, s0); sa.n_streams = Gs; sa.n_frames = F; sa.spf = spf;
      sa.frame0 = f0; sa.total_frames = n_frames;
      sa.cond_a = fbk.cond_a; sa.cond_b = fbk.cond_b; sa.lpc = fbk.lpc; sa.active = fbk.active;
      sa.pcm = d_pcm + (size_t)s0 * row_len; sa.forced = d_forced ? d_forced + (size_t)s0 * row_len : nullptr;
      sa.trace = d_trace; sa.logits = d_logits;     // whole-batch runs only (G = n_streams above)
#ifdef LPCB_PHASE_CLOCKS
      if (!g_phase_clk) { CK(cudaMalloc(&g_phase_clk, 16 * 8)); CK(cudaMemset(g_phase_clk, 0, 16 * 8)); }
      sa.phase_clk = g_phase_clk;
#endif
      int nl = 0;
      cudaEvent_t e0 = nullptr, e1 = nullptr;
      if (b->profiling) { CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1)); CK(cudaEventRecord(e0, b->stream)); }
      if (l2_window(b, b->stream, fbk.cond_a, (size_t)F * Gs * kRowsAP * 4)) return -1;
      CK(launch_sample_kernel(sa, tile, b->n_sms, b->stream, &nl));
      if (b->profiling) { CK(cudaEventRecord(e1, b->stream)); b->ev_sample.emplace_back(e0, e1); b->ev_sample_steps.push_back(F * spf); }
      if (pipe) CK(cudaEventRecord(b->ev_samples[buf], b->stream));
      b->launches += nl;
      if (h_pcm) {
        if (!b->cstream) {
          CK(cudaStreamCreateWithFlags(&b->cstream, cudaStreamNonBlocking));
          CK(cudaEventCreateWithFlags(&b->ev_chunk, cudaEventDisableTiming));
          CK(cudaEventCreateWithFlags(&b->ev_copied, cudaEventDisableTiming));
        }
        const size_t row = row_len * sizeof(short), off = (size_t)s0 * row_len + (size_t)f0 * spf;
        CK(cudaEventRecord(b->ev_chunk, b->stream));
        CK(cudaStreamWaitEvent(b->cstream, b->ev_chunk, 0));
        CK(cudaMemcpy2DAsync(h_pcm + off, row, d_pcm + off, row, (size_t)F * spf * sizeof(short), Gs,
                             cudaMemcpyDeviceToHost, b->cstream));
      }
    }
  }
  if (h_pcm) {
    CK(cudaEventRecord(b->ev_copied, b->cstream));
    CK(cudaStreamWaitEvent(b->stream, b->ev_copied, 0));
  }
  return 0;
}

}  // namespace

// ---- hooks for the PLC driver (plc.cu) ----
int lpcb::default_features_delay() {
  std::lock_guard<std::mutex> lk(g_cfg_mu);
  return g_default_delay;
}

float lpcb::default_lpc_gamma() {
  std::lock_guard<std::mutex> lk(g_cfg_mu);
  return g_default_gamma;
}
int lpcb::current_device() {
  int d = 0;
  if (cudaGetDevice(&d) != cudaSuccess) { cudaGetLastError(); d = 0; }
  return d;
}

std::vector<float> lpcb::file_codebooks() {
  const char *path = getenv("LPCNET_B200_WEIGHTS");
  if (!path || !*path) path = "weights_blob.bin";
  std::vector<float> out;
  FILE *f = fopen(path, "rb");
  if (!f) return out;
  std::vector<unsigned char> data;
  unsigned char buf[65536];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) data.insert(data.end(), buf, buf + n);
  fclose(f);
  // DNNw records (src/nnet.h:41-61): 64-byte head {"DNNw", version, type, size, block_size, name[44]} + payload
  static const char *names[4] = {"ceps_codebook1", "ceps_codebook2", "ceps_codebook3", "ceps_codebook_diff4"};
  const size_t want[4] = {1024 * 17 * 4, 1024 * 17 * 4, 1024 * 17 * 4, 4096 * 18 * 4};
  const unsigned char *found[4] = {nullptr, nullptr, nullptr, nullptr};
  for (size_t off = 0; off + 64 <= data.size();) {
    int32_t size, block;
    memcpy(&size, &data[off + 12], 4);
    memcpy(&block, &data[off + 16], 4);
    if (size <= 0 || block < size || (size_t)block > data.size() - off - 64 || data[off + 63] != 0) break;
    for (int k = 0; k < 4; k++)
      if (!strcmp(reinterpret_cast<const char *>(&data[off + 20]), names[k]) && (size_t)size == want[k]) found[k] = &data[off + 64];
    off += 64 + (size_t)block;
  }
  if (!found[0] || !found[1] || !found[2] || !found[3]) return out;
  out.resize((want[0] + want[1] + want[2] + want[3]) / 4);
  size_t o = 0;
  for (int k = 0; k < 4; k++) { memcpy(reinterpret_cast<char *>(out.data()) + o, found[k], want[k]); o += want[k]; }
  return out;
}

BatchView lpcb::batch_view(LPCNetBatch *b) { return BatchView{b->st, &b->model, b->stream, b->n_streams, b->device}; }

int lpcb::batch_frame_network_masked(LPCNetBatch *b, const float *d_features, int stride, const uint8_t *d_mask,
                                     bool state_only, const int *d_rows, int n_rows, cudaStream_t side) {
  CK(cudaSetDevice(b->device));
  if (ensure_frame_buffers(b, 1) || (state_only && ensure_frame_buffers(b, 1, true))) return -1;
  FrameArgs fa{};
  fa.m = b->model.d; fa.st = b->st; fa.fb = state_only ? b->fb2 : b->fb; fa.n_streams = b->n_streams; fa.n_frames = 1;
  fa.state_only = state_only ? 1 : 0;
  fa.features = d_features; fa.feature_stride = stride; fa.frame0 = 0; fa.total_frames = 1;
  fa.stream_mask = d_mask;
  if (d_rows && !getenv("LPCNET_B200_NO_COMPACT")) { fa.rows = d_rows; fa.n_rows = n_rows; fa.stream_mask = nullptr; }
  int nl = 0;
  CK(launch_frame_network(fa, side ? side : b->stream, &nl, fork_for(b, side ? side : b->stream)));
  b->launches += nl;
  return 0;
}

int lpcb::batch_sample_segment(LPCNetBatch *b, int n, const uint8_t *d_flags, const short *d_forced, short *d_out,
                               const int *d_perm, int n_act, cudaStream_t side, int tile_override) {
  cudaStream_t cs = side ? side : b->stream;
  if (n <= 0 || n > kFrame) return fail("sample segment must be 1..160 steps");
  CK(cudaSetDevice(b->device));
  if (d_perm && n_act == 0) return 0;
  if (d_perm && 2 * n_act <= b->n_streams && !getenv("LPCNET_B200_NO_COMPACT")) {
    const size_t A = n_act;
    size_t off = 0;
    auto take = [&](size_t bytes) { size_t r = off; off += (bytes + 255) & ~size_t(255); return r; };
    const size_t o_ca = take(A * kRowsAP * 4), o_ga = take(A * kN * 4), o_cb = take(A * kRowsB * 4), o_gb = take(A * kNB * 4),
                 o_ls = take(A * 16 * 4), o_lpc = take(A * 16 * 4), o_rng = take(A * 16), o_de = take(A * 4), o_le = take(A * 4),
                 o_fl = take(A), o_fo = take(A * kFrame * 2), o_out = take(A * kFrame * 2);
    if (side ? grow(&b->d_compact2, &b->d_compact2_bytes, off) : grow(&b->d_compact, &b->d_compact_bytes, off)) return -1;
    char *p = static_cast<char *>(side ? b->d_compact2 : b->d_compact);
    CompactView c{};
    c.cond_a = (float *)(p + o_ca); c.st.gru_a = (float *)(p + o_ga); c.cond_b = (float *)(p + o_cb); c.st.gru_b = (float *)(p + o_gb);
    c.st.last_sig = (float *)(p + o_ls); c.lpc = (float *)(p + o_lpc); c.st.rng = (uint32_t *)(p + o_rng);
    c.st.deemph = (float *)(p + o_de); c.st.last_exc = (int *)(p + o_le); c.flags = (uint8_t *)(p + o_fl);
    c.forced = (short *)(p + o_fo); c.out = (short *)(p + o_out);
    compact_gather_kernel<<<n_act, 128, 0, cs>>>(c, b->st, b->fb.cond_a, b->fb.cond_b, b->fb.lpc, d_flags, d_forced, d_perm, n);
    SampleArgs sa{};
    sa.m = b->model.d; sa.st = c.st; sa.n_streams = n_act; sa.n_frames = 1; sa.spf = n;
    sa.frame0 = 0; sa.total_frames = 1;
    sa.cond_a = c.cond_a; sa.cond_b = c.cond_b; sa.lpc = c.lpc; sa.active = c.flags;
    sa.pcm = c.out; sa.forced = d_forced ? c.forced : nullptr; sa.forced_by_flag = 1;
    int nl = 0;
    int tile = tile_override > 0 ? tile_override : b->tile ? b->tile : (n_act <= 8 * b->n_sms ? 8 : 16);
    tile = tile < b->max_tile ? tile : b->max_tile;
    CK(launch_sample_kernel(sa, tile, b->n_sms, cs, &nl));
    compact_scatter_kernel<<<n_act, 128, 0, cs>>>(c, b->st, d_out, d_perm, n);
    CK(cudaGetLastError());
    b->launches += nl + 2;
    return 0;
  }
  SampleArgs sa{};
  sa.m = b->model.d; sa.st = b->st; sa.n_streams = b->n_streams; sa.n_frames = 1; sa.spf = n;
  sa.frame0 = 0; sa.total_frames = 1;
  sa.cond_a = b->fb.cond_a; sa.cond_b = b->fb.cond_b; sa.lpc = b->fb.lpc; sa.active = d_flags;
  sa.pcm = d_out; sa.forced = d_forced; sa.forced_by_flag = 1;
  int nl = 0;
  int tile = tile_override > 0 ? tile_override : auto_tile(b);
  tile = tile < b->max_tile ? tile : b->max_tile;
  CK(launch_sample_kernel(sa, tile, b->n_sms, cs, &nl));
  b->launches += nl;
  return 0;
}

extern "C" {

#ifdef LPCB_PHASE_CLOCKS
// development build only (tools/phase_clocks.py): cycle sums per phase of the sample kernel's CTA 0, then cleared
__attribute__((visibility("default"))) int lpcnet_b200_read_phase_clocks(unsigned long long *out) {
  if (!g_phase_clk) return -1;
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpy(out, g_phase_clk, 16 * 8, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  cudaMemset(g_phase_clk, 0, 16 * 8);
  return 0;
}
#endif

const char *lpcnet_b200_last_error(void) {
  g_err_ret = g_err;
  g_err.clear();
  return g_err_ret.c_str();
}

int lpcnet_b200_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

void lpcnet_b200_set_rcp_table(const uint32_t *t) {
  std::lock_guard<std::mutex> lk(g_cfg_mu);
  if (t) { memcpy(g_rcp, t, sizeof(g_rcp)); g_have_rcp = true; }
  else g_have_rcp = false;
}
void lpcnet_b200_default_rcp_table(uint32_t *t) { default_rcp_table(t); }

void lpcnet_b200_set_model_constants(int features_delay, float lpc_gamma) {
  std::lock_guard<std::mutex> lk(g_cfg_mu);
  g_default_delay = features_delay;
  g_default_gamma = lpc_gamma;
}

void lpcnet_b200_set_end2end(int on) {
  std::lock_guard<std::mutex> lk(g_cfg_mu);
  g_default_end2end = on ? 1 : 0;
}

int lpcnet_batch_set_end2end(LPCNetBatch *b, int on) {
  if (!b) return fail("null batch");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  b->model.d.end2end = on ? 1 : 0;
  return 0;
}

LPCNetBatch *lpcnet_batch_create(const unsigned char *blob, int len, int n_streams, int device, int features_delay,
                                 float lpc_gamma) {
  if (n_streams <= 0) { fail("n_streams must be positive"); return nullptr; }
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0) {
    cudaGetLastError();
    fail("no CUDA device available (liblpcnet_b200 has no CPU fallback)");
    return nullptr;
  }
  if (device < 0) { if (cudaGetDevice(&device) != cudaSuccess) device = 0; }
  if (device >= ndev) { fail("device index out of range"); return nullptr; }
  if (cudaSetDevice(device) != cudaSuccess) { fail("cudaSetDevice failed"); return nullptr; }
  LPCNetBatch *b = new (std::nothrow) LPCNetBatch();
  if (!b) { fail("out of memory"); return nullptr; }
  b->device = device;
  b->n_streams = n_streams;
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  b->n_sms = prop.multiProcessorCount;
  {
    // L2 carve-out for the conditioning rows (see group_streams): what the device allows, unless the caller says less
    size_t want = (size_t)prop.persistingL2CacheMaxSize;
    if (const char *e = getenv("LPCNET_B200_L2_PERSIST_MB")) want = (size_t)(atoi(e) > 0 ? atoi(e) : 0) << 20;
    if (want > (size_t)prop.persistingL2CacheMaxSize) want = prop.persistingL2CacheMaxSize;
    if (want && prop.accessPolicyMaxWindowSize > 0 && cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want) == cudaSuccess) {
      b->l2_window_max = want;
      b->l2_window_bytes = prop.accessPolicyMaxWindowSize;
    } else {
      cudaGetLastError();
    }
    if (getenv("LPCNET_B200_DEBUG"))
      fprintf(stderr, "L2 %d MB, persisting max %d MB, window max %d MB, carve-out %zu MB\n", prop.l2CacheSize >> 20,
              prop.persistingL2CacheMaxSize >> 20, prop.accessPolicyMaxWindowSize >> 20, b->l2_window_max >> 20);
  }
  uint32_t rcp[2048];
  bool have;
  {
    std::lock_guard<std::mutex> lk(g_cfg_mu);
    have = g_have_rcp;
    if (have) memcpy(rcp, g_rcp, sizeof(rcp));
  }
  std::string err;
  if (!load_model(b->model, blob, len, features_delay, lpc_gamma, have ? rcp : nullptr, err)) {
    fail(err);
    delete b;
    return nullptr;
  }
  // largest streams-per-CTA tile whose resident weights + per-stream buffers fit this GPU's shared memory
  { std::lock_guard<std::mutex> lk(g_cfg_mu); b->model.d.end2end = g_default_end2end; }
  b->max_tile = 0;
  for (int t : {32, 16, 8})
    if (!b->max_tile && sample_kernel_smem(t, b->model.d) <= (size_t)prop.sharedMemPerBlockOptin) b->max_tile = t;
  if (!b->max_tile) {
    fail("model does not fit the shared-memory-resident kernel on this GPU");
    free_model(b->model); delete b;
    return nullptr;
  }
  if (alloc_state(b) != 0) { free_model(b->model); delete b; return nullptr; }
  if (cudaStreamCreateWithFlags(&b->own_stream, cudaStreamNonBlocking) != cudaSuccess) {
    fail("cudaStreamCreate failed");
    lpcnet_batch_destroy(b);
    return nullptr;
  }
  b->stream = b->own_stream;
  if (lpcnet_batch_reset(b, -1) != 0) { lpcnet_batch_destroy(b); return nullptr; }
  return b;
}

void lpcnet_batch_destroy(LPCNetBatch *b) {
  if (!b) return;
  cudaSetDevice(b->device);
  if (b->stream) cudaStreamSynchronize(b->stream);
  if (b->l2_window_max) {                      // hand the conditioning rows' lines back to normal caching
    cudaStreamAttrValue v{};                   // (num_bytes 0 disables the window; the stream may be the caller's)
    if (b->stream) cudaStreamSetAttribute(b->stream, cudaStreamAttributeAccessPolicyWindow, &v);
    cudaCtxResetPersistingL2Cache();
    cudaGetLastError();
  }
  free_model(b->model);
  if (b->fstream) {
    cudaStreamSynchronize(b->fstream);
    cudaStreamDestroy(b->fstream);
    cudaEventDestroy(b->ev_entry);
    for (int k = 0; k < 2; k++) { cudaEventDestroy(b->ev_frames[k]); cudaEventDestroy(b->ev_samples[k]); }
  }
  for (auto &f : b->forks) { cudaStreamSynchronize(f.second.aux); cudaStreamDestroy(f.second.aux); cudaEventDestroy(f.second.begin); cudaEventDestroy(f.second.done); }
  void *ptrs[] = {b->st_alloc, b->fb_alloc, b->fb2_alloc, b->d_feat, b->d_pcm, b->d_forced, b->d_trace, b->d_logits, b->d_packets, b->d_compact, b->d_compact2};
  for (void *p : ptrs) if (p) cudaFree(p);
  if (b->hstream) { cudaStreamSynchronize(b->hstream); cudaStreamDestroy(b->hstream); cudaEventDestroy(b->ev_h2d[0]); cudaEventDestroy(b->ev_h2d[1]); cudaEventDestroy(b->ev_h2d_entry); }
  if (b->cstream) { cudaStreamSynchronize(b->cstream); cudaStreamDestroy(b->cstream); cudaEventDestroy(b->ev_chunk); cudaEventDestroy(b->ev_copied); }
  if (b->own_stream) cudaStreamDestroy(b->own_stream);
  delete b;
}

int lpcnet_batch_n_streams(const LPCNetBatch *b) { return b ? b->n_streams : 0; }
int lpcnet_batch_nblocks(const LPCNetBatch *b, int which) {
  if (!b) return 0;
  switch (which) {
    case 0: return b->model.nblocks_a;
    case 1: return b->model.nblocks_b;
    case 2: return b->model.d.rcp_arith;               // 1: reciprocal estimate computed, 0: table lookup
    case 3: return b->model.d.a_rec.slots[2] != 0;     // 1: GRU-A runs the unrolled fixed-slot path
    case 4: return b->model.d.a_rec.n_chunks;
    case 7: return b->model.d.small_acc;               // 1: conversion-free seeds (int32 sums proven inside +-2^22)
    case 6: return b->model.d.a_rec.sat.n + b->model.d.b_in.sat.n + b->model.d.b_rec.sat.n;   // int8 pairs on the exact s16 pass
    case 5: return auto_tile(b);                        // streams per CTA the next synthesize call will use
    case 8: case 16: case 32: return (int)sample_kernel_smem(which, b->model.d);   // bytes of shared memory per CTA
    default: return 0;
  }
}
long long lpcnet_batch_launch_count(const LPCNetBatch *b) { return b ? b->launches : 0; }

int lpcnet_batch_set_cuda_stream(LPCNetBatch *b, void *s) {
  if (!b) return fail("null batch");
  b->stream = s ? static_cast<cudaStream_t>(s) : b->own_stream;
  return 0;
}

int lpcnet_batch_set_tile(LPCNetBatch *b, int t) {
  if (!b) return fail("null batch");
  if (t != 0 && t != 8 && t != 16 && t != 32 && t != 28) return fail("tile must be 0, 8, 16, 32 or 28 (two 8-stream tiles per CTA)");
  if ((t == 28 ? 16 : t) > b->max_tile) return fail("tile does not fit in shared memory for this model");
  b->tile = t;
  return 0;
}

int lpcnet_batch_set_profiling(LPCNetBatch *b, int on) {
  if (!b) return fail("null batch");
  b->profiling = on != 0;
  return 0;
}

// Sums (and clears) the CUDA-event durations recorded around the sample-kernel launches and the
// frame-network launch groups since the last read.  Synchronises the batch's stream.
int lpcnet_batch_read_profile(LPCNetBatch *b, double *sample_ms, int *sample_launches, double *frame_ms,
                              int *frame_groups) {
  if (!b) return fail("null batch");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  b->last_launch_ms.clear();
  b->last_launch_steps.swap(b->ev_sample_steps);
  b->ev_sample_steps.clear();
  auto drain = [&](std::vector<std::pair<cudaEvent_t, cudaEvent_t>> &v, double *ms, int *n) -> int {
    double tot = 0;
    for (auto &p : v) {
      float t = 0;
      CK(cudaEventElapsedTime(&t, p.first, p.second));
      tot += t;
      if (&v == &b->ev_sample) b->last_launch_ms.push_back(t);
      cudaEventDestroy(p.first); cudaEventDestroy(p.second);
    }
    if (ms) *ms = tot;
    if (n) *n = (int)v.size();
    v.clear();
    return 0;
  };
  if (drain(b->ev_sample, sample_ms, sample_launches)) return -1;
  if (drain(b->ev_frame, frame_ms, frame_groups)) return -1;
  return 0;
}

int lpcnet_batch_profile_launches(LPCNetBatch *b, float *ms, int *steps, int max) {
  if (!b) return fail("null batch");
  const int n = (int)b->last_launch_ms.size();
  for (int i = 0; i < n && i < max; i++) {
    if (ms) ms[i] = b->last_launch_ms[i];
    if (steps) steps[i] = i < (int)b->last_launch_steps.size() ? b->last_launch_steps[i] : 0;
  }
  return n;
}

int lpcnet_batch_sync(LPCNetBatch *b) {
  if (!b) return fail("null batch");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  return 0;
}

// lpcnet_reset (src/lpcnet.c:174-182)
int lpcnet_batch_reset(LPCNetBatch *b, int stream) {
  if (!b) return fail("null batch");
  if (stream >= b->n_streams) return fail("stream index out of range");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  uint32_t seed[4];
  lpcnet_kiss99_srand_host(seed, reinterpret_cast<const unsigned char *>("LPCNet"), 6);
  const int s0 = stream < 0 ? 0 : stream, n = stream < 0 ? b->n_streams : 1;
  const DevState &S = b->st;
  CK(cudaMemset(S.conv1_state + (size_t)s0 * 168, 0, (size_t)n * 168 * 4));
  CK(cudaMemset(S.conv2_state + (size_t)s0 * 256, 0, (size_t)n * 256 * 4));
  CK(cudaMemset(S.gru_a + (size_t)s0 * kN, 0, (size_t)n * kN * 4));
  CK(cudaMemset(S.gru_b + (size_t)s0 * kNB, 0, (size_t)n * kNB * 4));
  CK(cudaMemset(S.last_sig + (size_t)s0 * 16, 0, (size_t)n * 16 * 4));
  CK(cudaMemset(S.old_lpc + (size_t)s0 * kMaxDelay * 16, 0, (size_t)n * kMaxDelay * 16 * 4));
  CK(cudaMemset(S.lpc + (size_t)s0 * 16, 0, (size_t)n * 16 * 4));
  CK(cudaMemset(S.cond_a + (size_t)s0 * kRowsAP, 0, (size_t)n * kRowsAP * 4));
  CK(cudaMemset(S.cond_b + (size_t)s0 * kRowsB, 0, (size_t)n * kRowsB * 4));
  CK(cudaMemset(S.deemph + s0, 0, (size_t)n * 4));
  CK(cudaMemset(S.frame_count + s0, 0, (size_t)n * 4));
  CK(cudaMemset(S.vq_mem + (size_t)s0 * 18, 0, (size_t)n * 18 * 4));
  std::vector<int> le(n, 128);  // lin2ulaw(0) = 128
  std::vector<uint32_t> rng((size_t)n * 4);
  for (int i = 0; i < n; i++) memcpy(&rng[(size_t)i * 4], seed, 16);
  CK(cudaMemcpy(S.last_exc + s0, le.data(), (size_t)n * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.rng + (size_t)s0 * 4, rng.data(), (size_t)n * 16, cudaMemcpyHostToDevice));
  return 0;
}

int lpcnet_batch_set_seed(LPCNetBatch *b, int stream, const unsigned char *data, int n) {
  if (!b) return fail("null batch");
  if (stream < 0 || stream >= b->n_streams) return fail("stream index out of range");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  uint32_t st[4];
  lpcnet_kiss99_srand_host(st, data, n);
  CK(cudaMemcpy(b->st.rng + (size_t)stream * 4, st, 16, cudaMemcpyHostToDevice));
  return 0;
}

int lpcnet_batch_set_codebooks(LPCNetBatch *b, const float *cb, int n) {
  if (!b) return fail("null batch");
  CK(cudaSetDevice(b->device));
  std::string err;
  if (!install_codebooks(b->model, cb, n, err)) return fail(err);
  return 0;
}

int lpcnet_batch_synthesize_device(LPCNetBatch *b, const float *d_features, int stride, int n_frames, int spf,
                                   short *d_pcm, const short *d_forced, LPCNetB200Trace *d_trace, float *d_logits) {
  return synth_device(b, d_features, stride, n_frames, spf, d_pcm, d_forced, d_trace, d_logits);
}

int lpcnet_batch_synthesize(LPCNetBatch *b, const float *features, int stride, int n_frames, short *pcm,
                            const short *forced, LPCNetB200Trace *trace, float *logits) {
  if (!b) return fail("null batch");
  if (n_frames <= 0) return 0;
  if (!features || !pcm) return fail("null buffer");
  CK(cudaSetDevice(b->device));
  const size_t B = b->n_streams, T = (size_t)n_frames * kFrame;
  const size_t fbytes = B * n_frames * (size_t)stride * 4, pbytes = B * T * 2;
  if (grow(&b->d_feat, &b->d_feat_bytes, fbytes) || grow(&b->d_pcm, &b->d_pcm_bytes, pbytes)) return -1;
  // pinned features go in chunk by chunk beside the computation (synth_device); pageable ones in one staged copy
  const bool feat_pinned = is_pinned_host(features) && !getenv("LPCNET_B200_NO_H2D_PIPELINE");
  if (!feat_pinned) CK(cudaMemcpyAsync(b->d_feat, features, fbytes, cudaMemcpyHostToDevice, b->stream));
  if (forced) {
    if (grow(&b->d_forced, &b->d_forced_bytes, pbytes)) return -1;
    CK(cudaMemcpyAsync(b->d_forced, forced, pbytes, cudaMemcpyHostToDevice, b->stream));
  }
  if (trace && grow(&b->d_trace, &b->d_trace_bytes, B * T * sizeof(LPCNetB200Trace))) return -1;
  // a pinned destination takes the finished chunks while the next ones are computed (pageable memory would make the
  // "asynchronous" copy block the host inside the chunk loop)
  const bool pinned = is_pinned_host(pcm);
  if (logits && grow(&b->d_logits, &b->d_logits_bytes, B * T * 256 * 4)) return -1;
  if (synth_device(b, static_cast<const float *>(b->d_feat), stride, n_frames, kFrame, static_cast<short *>(b->d_pcm),
                   forced ? static_cast<const short *>(b->d_forced) : nullptr,
                   trace ? static_cast<LPCNetB200Trace *>(b->d_trace) : nullptr,
                   logits ? static_cast<float *>(b->d_logits) : nullptr, pinned ? pcm : nullptr,
                   feat_pinned ? features : nullptr))
    return -1;
  if (!pinned) CK(cudaMemcpyAsync(pcm, b->d_pcm, pbytes, cudaMemcpyDeviceToHost, b->stream));
  if (trace) CK(cudaMemcpyAsync(trace, b->d_trace, B * T * sizeof(LPCNetB200Trace), cudaMemcpyDeviceToHost, b->stream));
  if (logits) CK(cudaMemcpyAsync(logits, b->d_logits, B * T * 256 * 4, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return 0;
}

int lpcnet_batch_frame_network(LPCNetBatch *b, const float *features, int stride, int n_frames, float *ga, float *gb,
                               float *lpc) {
  if (!b) return fail("null batch");
  if (n_frames <= 0) return 0;
  if (stride < kFeat) return fail("feature_stride must be >= 20");
  CK(cudaSetDevice(b->device));
  const size_t B = b->n_streams;
  const size_t fbytes = B * n_frames * (size_t)stride * 4;
  if (grow(&b->d_feat, &b->d_feat_bytes, fbytes)) return -1;
  CK(cudaMemcpyAsync(b->d_feat, features, fbytes, cudaMemcpyHostToDevice, b->stream));
  long long cap = kChunkBytes / ((long long)b->n_streams * 6400);      // a test entry point: whole batch, smaller chunks
  const int Fc0 = n_frames < kChunkFrames ? n_frames : kChunkFrames;
  const int Fc = (int)(cap < 1 ? 1 : (cap < Fc0 ? cap : Fc0));
  if (ensure_frame_buffers(b, Fc)) return -1;
  for (int f0 = 0; f0 < n_frames; f0 += Fc) {
    const int F = (n_frames - f0) < Fc ? (n_frames - f0) : Fc;
    if (run_frames(b, static_cast<const float *>(b->d_feat), stride, f0, F, n_frames)) return -1;
    CK(cudaStreamSynchronize(b->stream));
    // device [F][B][W] -> host [B][n_frames][W]
    auto fetch = [&](float *dst, const float *src, int W) -> int {
      if (!dst) return 0;
      for (int f = 0; f < F; f++)
        CK(cudaMemcpy2D(dst + (size_t)(f0 + f) * W, (size_t)n_frames * W * 4, src + (size_t)f * B * W, (size_t)W * 4,
                        (size_t)W * 4, B, cudaMemcpyDeviceToHost));
      return 0;
    };
    if (fetch(gb, b->fb.cond_b, kRowsB) || fetch(lpc, b->fb.lpc, 16)) return -1;
    if (ga) {
      // device rows are [F][B][kRowsAP] in gru_a_col_index order; the caller gets the reference's column order
      std::vector<float> tmp((size_t)F * B * kRowsAP);
      CK(cudaMemcpy(tmp.data(), b->fb.cond_a, tmp.size() * 4, cudaMemcpyDeviceToHost));
      for (int f = 0; f < F; f++)
        for (size_t s = 0; s < (size_t)B; s++) {
          const float *src = tmp.data() + ((size_t)f * B + s) * kRowsAP;
          float *row = ga + (s * n_frames + f0 + f) * kRowsA;
          for (int col = 0; col < kRowsA; col++) row[col] = src[gru_a_col_index(col)];
        }
    }
  }
  return 0;
}

int lpcnet_batch_decode(LPCNetBatch *b, const unsigned char *packets, int n_packets, short *pcm, float *features_out) {
  if (!b) return fail("null batch");
  if (n_packets <= 0) return 0;
  if (!b->model.has_codebooks) return fail("no VQ codebooks: put ceps_codebook* arrays in the blob or call lpcnet_batch_set_codebooks");
  CK(cudaSetDevice(b->device));
  const size_t B = b->n_streams;
  const int n_frames = 4 * n_packets;
  const size_t pk_bytes = B * n_packets * 8, fbytes = B * n_frames * 36 * 4, pbytes = B * (size_t)n_frames * kFrame * 2;
  if (grow(&b->d_packets, &b->d_packets_bytes, pk_bytes) || grow(&b->d_feat, &b->d_feat_bytes, fbytes) ||
      grow(&b->d_pcm, &b->d_pcm_bytes, pbytes))
    return -1;
  CK(cudaMemcpyAsync(b->d_packets, packets, pk_bytes, cudaMemcpyHostToDevice, b->stream));
  DecodeArgs da{};
  da.m = b->model.d; da.st = b->st; da.n_streams = b->n_streams; da.n_packets = n_packets;
  da.packets = static_cast<const unsigned char *>(b->d_packets); da.features = static_cast<float *>(b->d_feat);
  int nl = 0;
  CK(launch_decode_packets(da, b->stream, &nl));
  b->launches += nl;
  const bool pinned = is_pinned_host(pcm);     // finished chunks leave beside the next chunk's kernel
  if (synth_device(b, static_cast<const float *>(b->d_feat), 36, n_frames, kFrame, static_cast<short *>(b->d_pcm), nullptr,
                   nullptr, nullptr, pinned ? pcm : nullptr))
    return -1;
  if (pcm && !pinned) CK(cudaMemcpyAsync(pcm, b->d_pcm, pbytes, cudaMemcpyDeviceToHost, b->stream));
  if (features_out) CK(cudaMemcpyAsync(features_out, b->d_feat, fbytes, cudaMemcpyDeviceToHost, b->stream));
  CK(cudaStreamSynchronize(b->stream));
  return 0;
}

int lpcnet_batch_get_state(LPCNetBatch *b, int s, LPCNetB200StreamState *o) {
  if (!b || !o) return fail("null argument");
  if (s < 0 || s >= b->n_streams) return fail("stream index out of range");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  const DevState &S = b->st;
  memset(o, 0, sizeof(*o));
  CK(cudaMemcpy(o->conv1_state, S.conv1_state + (size_t)s * 168, 168 * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->conv2_state, S.conv2_state + (size_t)s * 256, 256 * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->gru_a_state, S.gru_a + (size_t)s * kN, kN * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->gru_b_state, S.gru_b + (size_t)s * kNB, kNB * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->last_sig, S.last_sig + (size_t)s * 16, 64, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->lpc, S.lpc + (size_t)s * 16, 64, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->old_lpc, S.old_lpc + (size_t)s * kMaxDelay * 16, (size_t)b->model.d.features_delay * 64, cudaMemcpyDeviceToHost));
  {
    float tmp[kRowsAP];
    CK(cudaMemcpy(tmp, S.cond_a + (size_t)s * kRowsAP, kRowsAP * 4, cudaMemcpyDeviceToHost));
    for (int col = 0; col < kRowsA; col++) o->gru_a_condition[col] = tmp[gru_a_col_index(col)];
  }
  CK(cudaMemcpy(o->gru_b_condition, S.cond_b + (size_t)s * kRowsB, kRowsB * 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&o->deemph_mem, S.deemph + s, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&o->last_exc, S.last_exc + s, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(&o->frame_count, S.frame_count + s, 4, cudaMemcpyDeviceToHost));
  CK(cudaMemcpy(o->rng, S.rng + (size_t)s * 4, 16, cudaMemcpyDeviceToHost));
  return 0;
}

int lpcnet_batch_set_state(LPCNetBatch *b, int s, const LPCNetB200StreamState *in) {
  if (!b || !in) return fail("null argument");
  if (s < 0 || s >= b->n_streams) return fail("stream index out of range");
  // a GRU state is a convex combination of values in [-1, 1]; the kernels' byte quantiser relies on it
  for (int i = 0; i < kN; i++)
    if (!(in->gru_a_state[i] >= -1.f && in->gru_a_state[i] <= 1.f)) return fail("gru_a_state outside [-1, 1]");
  for (int i = 0; i < kNB; i++)
    if (!(in->gru_b_state[i] >= -1.f && in->gru_b_state[i] <= 1.f)) return fail("gru_b_state outside [-1, 1]");
  CK(cudaSetDevice(b->device));
  CK(cudaStreamSynchronize(b->stream));
  const DevState &S = b->st;
  CK(cudaMemcpy(S.conv1_state + (size_t)s * 168, in->conv1_state, 168 * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.conv2_state + (size_t)s * 256, in->conv2_state, 256 * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.gru_a + (size_t)s * kN, in->gru_a_state, kN * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.gru_b + (size_t)s * kNB, in->gru_b_state, kNB * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.last_sig + (size_t)s * 16, in->last_sig, 64, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.lpc + (size_t)s * 16, in->lpc, 64, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.old_lpc + (size_t)s * kMaxDelay * 16, in->old_lpc, (size_t)b->model.d.features_delay * 64, cudaMemcpyHostToDevice));
  {
    float tmp[kRowsAP] = {0.f};
    for (int col = 0; col < kRowsA; col++) tmp[gru_a_col_index(col)] = in->gru_a_condition[col];
    CK(cudaMemcpy(S.cond_a + (size_t)s * kRowsAP, tmp, kRowsAP * 4, cudaMemcpyHostToDevice));
  }
  CK(cudaMemcpy(S.cond_b + (size_t)s * kRowsB, in->gru_b_condition, kRowsB * 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.deemph + s, &in->deemph_mem, 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.last_exc + s, &in->last_exc, 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.frame_count + s, &in->frame_count, 4, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(S.rng + (size_t)s * 4, in->rng, 16, cudaMemcpyHostToDevice));
  return 0;
}

/* ---------------------------------------------------------------------------------------------
 * Drop-in lpcnet.h symbols: a one-stream batch behind the reference's opaque state.
 * ------------------------------------------------------------------------------------------- */
// The state is a handle to device-resident data, not the relocatable POD of the reference (lpcnet_private.h:28-53): a
// copy by value (memcpy / struct assignment into other memory) would alias one device slot from two places.  `self`
// makes that detectable: every entry point refuses a state whose address is not the one lpcnet_init wrote.
struct LPCNetState {
  uint32_t magic;
  LPCNetState *self;
  LPCNetBatch *batch;
  float *pinned_feat;   // 36 floats
  short *pinned_pcm;    // 160 samples
};
struct LPCNetDecState {
  LPCNetState lpcnet_state;
};
static const uint32_t kMagic = 0x4C50434Eu;

int lpcnet_get_size(void) { return (int)sizeof(LPCNetState); }

int lpcnet_init(LPCNetState *st) {
  memset(st, 0, sizeof(*st));
  st->magic = kMagic;
  st->self = st;
  return 0;
}

LPCNetState *lpcnet_create(void) {
  LPCNetState *st = static_cast<LPCNetState *>(calloc(1, sizeof(LPCNetState)));
  if (st) lpcnet_init(st);
  return st;
}

static bool state_ok(const LPCNetState *st, const char *who) {
  if (!st) { fail(std::string(who) + ": null state"); return false; }
  if (st->magic != kMagic) { fail(std::string(who) + ": state not initialised (lpcnet_init / lpcnet_create)"); return false; }
  if (st->self != st) {
    fail(std::string(who) + ": this LPCNetState was copied by value; it is a handle to device-resident state and cannot be relocated (INTEGRATION.md)");
    return false;
  }
  return true;
}

static void state_release(LPCNetState *st) {
  if (st->batch) lpcnet_batch_destroy(st->batch);
  if (st->pinned_feat) cudaFreeHost(st->pinned_feat);
  if (st->pinned_pcm) cudaFreeHost(st->pinned_pcm);
  st->batch = nullptr; st->pinned_feat = nullptr; st->pinned_pcm = nullptr;
}

void lpcnet_destroy(LPCNetState *st) {
  if (!st) return;
  if (st->magic == kMagic && st->self == st) state_release(st);   // a by-value copy does not own the device state
  free(st);
}

int lpcnet_load_model(LPCNetState *st, const unsigned char *data, int len) {
  if (!state_ok(st, "lpcnet_load_model")) return -1;
  state_release(st);
  int delay; float gamma;
  { std::lock_guard<std::mutex> lk(g_cfg_mu); delay = g_default_delay; gamma = g_default_gamma; }
  st->batch = lpcnet_batch_create(data, len, 1, -1, delay, gamma);
  if (!st->batch) return -1;
  if (cudaMallocHost(&st->pinned_feat, 36 * 4) != cudaSuccess || cudaMallocHost(&st->pinned_pcm, 640 * 2) != cudaSuccess) {
    state_release(st);
    return fail("cudaMallocHost failed"), -1;
  }
  return 0;
}

// The reference's lpcnet_create() returns a state with the COMPILED-IN model ready (src/lpcnet.c:184-200, unless it was
// built -DUSE_WEIGHTS_FILE); callers such as `lpcnet_demo -decode` (src/lpcnet_demo.c:176-189) never load one.  The
// stand-in for the compiled-in model is a blob file: $LPCNET_B200_WEIGHTS, else ./weights_blob.bin (the name the
// reference's own USE_WEIGHTS_FILE build reads, src/lpcnet_demo.c:113), loaded on first use.  -1 if there is none.
static int ensure_default_model(LPCNetState *st, bool decoder) {
  if (st->batch) return 0;
  const char *path = getenv("LPCNET_B200_WEIGHTS");
  if (!path || !*path) path = "weights_blob.bin";
  FILE *f = fopen(path, "rb");
  if (!f) return fail(std::string("no model loaded (call lpcnet_load_model first, or provide ") + path + ")");
  std::vector<unsigned char> data;
  unsigned char buf[65536];
  size_t n;
  while ((n = fread(buf, 1, sizeof(buf), f)) > 0) data.insert(data.end(), buf, buf + n);
  fclose(f);
  if (data.empty() || data.size() > 0x7fffffffu) return fail(std::string("cannot use ") + path + " as a model blob");
  return decoder ? lpcnet_decoder_load_model(reinterpret_cast<LPCNetDecState *>(st), data.data(), (int)data.size())
                 : lpcnet_load_model(st, data.data(), (int)data.size());
}

void lpcnet_reset(LPCNetState *st) {
  if (st && st->batch && state_ok(st, "lpcnet_reset")) lpcnet_batch_reset(st->batch, -1);
}

void lpcnet_synthesize(LPCNetState *st, const float *features, short *output, int N) {
  if (!state_ok(st, "lpcnet_synthesize") || (!st->batch && ensure_default_model(st, false))) return;
  if (N <= 0) return;
  if (N > kFrame) { fail("lpcnet_synthesize: N > 160"); return; }
  LPCNetBatch *b = st->batch;
  if (cudaSetDevice(b->device) != cudaSuccess) { fail("cudaSetDevice failed"); return; }
  memcpy(st->pinned_feat, features, kFeat * 4);
  if (grow(&b->d_feat, &b->d_feat_bytes, 36 * 4) || grow(&b->d_pcm, &b->d_pcm_bytes, kFrame * 2)) return;
  if (cudaMemcpyAsync(b->d_feat, st->pinned_feat, kFeat * 4, cudaMemcpyHostToDevice, b->stream) != cudaSuccess) { fail("H2D failed"); return; }
  if (synth_device(b, static_cast<const float *>(b->d_feat), kFeat, 1, N, static_cast<short *>(b->d_pcm), nullptr, nullptr, nullptr)) return;
  if (cudaMemcpyAsync(st->pinned_pcm, b->d_pcm, (size_t)N * 2, cudaMemcpyDeviceToHost, b->stream) != cudaSuccess) { fail("D2H failed"); return; }
  cudaError_t e = cudaStreamSynchronize(b->stream);
  if (e != cudaSuccess) { fail_cuda("lpcnet_synthesize", e); return; }
  memcpy(output, st->pinned_pcm, (size_t)N * 2);
}

int lpcnet_decoder_get_size(void) { return (int)sizeof(LPCNetDecState); }
int lpcnet_decoder_init(LPCNetDecState *st) { return lpcnet_init(&st->lpcnet_state); }
LPCNetDecState *lpcnet_decoder_create(void) {
  LPCNetDecState *st = static_cast<LPCNetDecState *>(calloc(1, sizeof(LPCNetDecState)));
  if (st) lpcnet_decoder_init(st);
  return st;
}
void lpcnet_decoder_destroy(LPCNetDecState *st) {
  if (!st) return;
  if (st->lpcnet_state.magic == kMagic && st->lpcnet_state.self == &st->lpcnet_state) state_release(&st->lpcnet_state);
  free(st);
}
int lpcnet_decoder_load_model(LPCNetDecState *st, const unsigned char *data, int len) {
  if (lpcnet_load_model(&st->lpcnet_state, data, len)) return -1;
  // the reference links its VQ tables into the library (ceps_codebooks.c); here they come with the blob or, for an
  // opaque decoder state that cannot reach lpcnet_batch_set_codebooks, from lpcnet_b200_set_codebooks
  LPCNetBatch *b = st->lpcnet_state.batch;
  if (!b->model.has_codebooks) {
    std::vector<float> cb = default_codebooks();
    if (cb.empty()) cb = file_codebooks();
    if (!cb.empty() && lpcnet_batch_set_codebooks(b, cb.data(), (int)cb.size())) return -1;
  }
  return 0;
}
int lpcnet_decode(LPCNetDecState *st, const unsigned char *buf, short *pcm) {
  if (!st || !state_ok(&st->lpcnet_state, "lpcnet_decode")) return st ? -1 : fail("lpcnet_decode: null state");
  if (!st->lpcnet_state.batch && ensure_default_model(&st->lpcnet_state, true)) return -1;
  return lpcnet_batch_decode(st->lpcnet_state.batch, buf, 1, pcm, nullptr);
}

}  // extern "C"
