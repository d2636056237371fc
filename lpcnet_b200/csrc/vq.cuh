// The quantisers of the 1.6 kb/s encoder (process_superframe with quantize = encode = 1, src/lpcnet_enc.c:701-733):
// c0 + 3-stage M-best VQ of the packet's last cepstrum (quantize_3stage_mbest, 134-248), the 13-bit predictive VQ of
// the second one (quantize_diff / find_nearest_multi, 250-321), the interpolation search (323-400), and the bit
// packer (441-463).  Included by analysis.cu (inside its namespace).
//
// One warp per packet, eight packets per CTA.  A codebook (or a 1024-entry tile of the 4096-entry one) is staged in
// shared memory once per CTA with coalesced loads and scanned by every warp; rows are read at an odd word stride
// (17, or 18 padded to 19), so a warp's 32 rows sit in 32 different banks.  Every distance is the reference's
// j-ascending float sum (no FMA: -fmad=false).  The reference scans entries in index order and accepts on a strict
// '<', so its M best are the M smallest (distance, index) pairs in lexicographic order: a 64-bit key
// (distance bits << 32 | index) turns the whole search into integer minima (distances are non-negative, so their
// bit patterns order like the floats; NaN and +inf sort above the 1e15 sentinel and are never accepted, as there).
#pragma once

struct VqArgs {
  const float *cb;          // cb1[1024][17] | cb2 | cb3 | diff4[4096][18]
  float *features;          // [B][4P][36] analysis rows in, quantised rows [0..20) out
  int *ids;                 // [B][P][kIds]
  float *vq_mem;            // [B][18] stream state: the previous packet's quantised last cepstrum
  float *left0;             // [B][18] vq_mem as it was when the call started
  unsigned char *packets;   // [B][P][8]
  int n_streams, n_packets;
};

constexpr int kVqWarps = 8, kVqThreads = kVqWarps * 32, kNb1 = 17, kCbN = 1024, kSurv = 5;
constexpr int kDiffStride = 19, kDiffTile = 1024;
constexpr size_t kVqEndSmem = (size_t)kCbN * kNb1 * 4;
constexpr size_t kVqMidSmem = (size_t)kDiffTile * kDiffStride * 4 + (size_t)kVqWarps * 6 * kBands * 4;
constexpr unsigned long long kVqSentinel = ((unsigned long long)0x58635FA9u << 32);   // (1e15f, index 0)

__device__ __forceinline__ unsigned long long vq_key(float d, int i) {
  return ((unsigned long long)__float_as_uint(d) << 32) | (unsigned)i;
}
__device__ __forceinline__ float key_dist(unsigned long long k) { return __uint_as_float((unsigned)(k >> 32)); }
__device__ __forceinline__ int key_index(unsigned long long k) { return (int)(unsigned)(k & 0xffffffffu); }

__device__ __forceinline__ unsigned long long warp_min_u64(unsigned long long k) {
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, k, o);
    k = other < k ? other : k;
  }
  return k;
}

// vq_quantize_mbest (lpcnet_enc.c:53-78) of x against the 1024 x 17 codebook in shared memory; res = the five best
// keys in order, the same in every lane
__device__ __forceinline__ void scan_mbest(const float *cbs, const float (&x)[kNb1], int lane, unsigned long long (&res)[kSurv]) {
  unsigned long long k[kSurv];
#pragma unroll
  for (int m = 0; m < kSurv; m++) k[m] = kVqSentinel;
  for (int i = lane; i < kCbN; i += 32) {
    const float *c = cbs + i * kNb1;
    float d = 0;
#pragma unroll
    for (int j = 0; j < kNb1; j++) {
      const float e = x[j] - c[j];
      d += e * e;
    }
    const unsigned long long key = vq_key(d, i);
    if (key < k[kSurv - 1]) {
      k[kSurv - 1] = key;
#pragma unroll
      for (int m = kSurv - 1; m > 0; m--)
        if (k[m] < k[m - 1]) { const unsigned long long tmp = k[m]; k[m] = k[m - 1]; k[m - 1] = tmp; }
    }
  }
#pragma unroll
  for (int m = 0; m < kSurv; m++) {
    const unsigned long long mn = warp_min_u64(k[0]);
    res[m] = mn;
    if (k[0] == mn) {
#pragma unroll
      for (int q = 0; q < kSurv - 1; q++) k[q] = k[q + 1];
      k[kSurv - 1] = ~0ull;
    }
  }
}

// the survivor merge of quantize_3stage_mbest (lpcnet_enc.c:160-178, 200-220), `nlead` leading indices per path
__device__ __forceinline__ void merge_survivors(float (&gd)[kSurv], int (&g)[kSurv][3], const unsigned long long (&cur)[kSurv],
                                                const int *lead, int nlead) {
  int m = 0;
  for (int pos = 0; pos < kSurv; pos++) {
    if (m < kSurv && key_dist(cur[m]) < gd[pos]) {
      for (int j = kSurv - 1; j >= pos + 1; j--) {
        gd[j] = gd[j - 1];
        for (int c = 0; c <= nlead; c++) g[j][c] = g[j - 1][c];
      }
      gd[pos] = key_dist(cur[m]);
      for (int c = 0; c < nlead; c++) g[pos][c] = lead[c];
      g[pos][nlead] = key_index(cur[m]);
      m++;
    }
  }
}

__device__ __forceinline__ void stage_rows(float *dst, const float *src, int n_floats, int tid) {
  for (int i = tid; i < n_floats; i += kVqThreads) dst[i] = __ldg(src + i);
}

// c0 and the 3-stage M-best VQ of the packet's fourth cepstrum; writes the quantised row in place
__global__ void __launch_bounds__(kVqThreads) vq_end_kernel(VqArgs A) {
  extern __shared__ float cbs[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long unit = (long long)blockIdx.x * kVqWarps + warp;
  const bool valid = unit < (long long)A.n_streams * A.n_packets;
  const int s = valid ? (int)(unit / A.n_packets) : 0, p = valid ? (int)(unit % A.n_packets) : 0;
  const float *cb1 = A.cb, *cb2 = cb1 + kCbN * kNb1, *cb3 = cb2 + kCbN * kNb1;
  float *F3 = A.features + ((size_t)s * A.n_packets * 4 + 4 * p + 3) * kNF;
  float x[kNb1];
#pragma unroll
  for (int j = 0; j < kNb1; j++) x[j] = F3[1 + j];
  if (valid && p == 0 && lane < kBands) A.left0[(size_t)s * kBands + lane] = A.vq_mem[(size_t)s * kBands + lane];

  unsigned long long r1[kSurv], cur[kSurv];
  float gd[kSurv];
  int g2[kSurv][3], g3[kSurv][3];
  stage_rows(cbs, cb1, kCbN * kNb1, tid);
  __syncthreads();
  if (valid) scan_mbest(cbs, x, lane, r1);
  __syncthreads();
  stage_rows(cbs, cb2, kCbN * kNb1, tid);
  __syncthreads();
  if (valid) {
    for (int k = 0; k < kSurv; k++) {
      const int i1 = key_index(r1[k]);
      float diff[kNb1];
#pragma unroll
      for (int j = 0; j < kNb1; j++) diff[j] = x[j] - __ldg(cb1 + i1 * kNb1 + j);
      scan_mbest(cbs, diff, lane, cur);
      if (k == 0) {
        for (int m = 0; m < kSurv; m++) { g2[m][0] = i1; g2[m][1] = key_index(cur[m]); gd[m] = key_dist(cur[m]); }
      } else if (key_dist(cur[0]) < gd[kSurv - 1]) {
        merge_survivors(gd, g2, cur, &i1, 1);
      }
    }
  }
  __syncthreads();
  stage_rows(cbs, cb3, kCbN * kNb1, tid);
  __syncthreads();
  if (!valid) return;
  for (int k = 0; k < kSurv; k++) {
    const int lead[2] = {g2[k][0], g2[k][1]};
    float diff[kNb1];
#pragma unroll
    for (int j = 0; j < kNb1; j++) diff[j] = x[j] - __ldg(cb1 + lead[0] * kNb1 + j) - __ldg(cb2 + lead[1] * kNb1 + j);
    scan_mbest(cbs, diff, lane, cur);
    if (k == 0) {
      for (int m = 0; m < kSurv; m++) {
        g3[m][0] = lead[0]; g3[m][1] = lead[1]; g3[m][2] = key_index(cur[m]);
        gd[m] = key_dist(cur[m]);
      }
    } else if (key_dist(cur[0]) < gd[kSurv - 1]) {
      merge_survivors(gd, g3, cur, lead, 2);
    }
  }
  const int e0 = g3[0][0], e1 = g3[0][1], e2 = g3[0][2];
  int *id = A.ids + ((size_t)s * A.n_packets + p) * kIds;
  if (lane < kNb1)
    F3[1 + lane] = __ldg(cb1 + e0 * kNb1 + lane) + __ldg(cb2 + e1 * kNb1 + lane) + cbs[e2 * kNb1 + lane];
  if (lane == 0) {
    int c0 = cvt_x86(floor(.5 + (double)(F3[0] * 4.f)));       // lpcnet_enc.c:704-706
    c0 = max(-64, min(63, c0));
    F3[0] = (float)c0 / 4.f;
    id[0] = c0 + 64; id[4] = e0; id[5] = e1; id[6] = e2;
  }
}

// quantize_diff of the second cepstrum against the 4096-entry signed codebook, double_interp_search,
// perform_double_interp and the bit packer; runs after vq_end_kernel has quantised every packet's last row
__global__ void __launch_bounds__(kVqThreads) vq_mid_kernel(VqArgs A) {
  extern __shared__ float sm[];
  float *tile = sm;                                              // [kDiffTile][19]
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  float *W = sm + kDiffTile * kDiffStride + warp * 6 * kBands;   // left | F0 | F1 | F2 | F3 | pred(.5(l+r))
  float *left = W, *F0s = W + kBands, *F1s = W + 2 * kBands, *F2s = W + 3 * kBands, *F3s = W + 4 * kBands, *avg = W + 5 * kBands;
  const long long unit = (long long)blockIdx.x * kVqWarps + warp;
  const bool valid = unit < (long long)A.n_streams * A.n_packets;
  const int s = valid ? (int)(unit / A.n_packets) : 0, p = valid ? (int)(unit % A.n_packets) : 0;
  const float *diff4 = A.cb + 3 * kCbN * kNb1;
  float *F = A.features + ((size_t)s * A.n_packets * 4 + 4 * p) * kNF;
  if (lane < kBands) {
    left[lane] = p == 0 ? A.left0[(size_t)s * kBands + lane] : F[-kNF + lane];
    F0s[lane] = F[lane]; F1s[lane] = F[kNF + lane]; F2s[lane] = F[2 * kNF + lane]; F3s[lane] = F[3 * kNF + lane];
  }
  __syncwarp();
  if (lane < kBands) avg[lane] = .5f * (left[lane] + F3s[lane]);
  __syncwarp();
  // this lane's entries all have (i & MULTI_MASK) = lane & 3: one of the four prediction modes (lpcnet_enc.c:302-306)
  const int mode = lane & 3;
  float tg[kBands];
#pragma unroll
  for (int j = 0; j < kBands; j++) {
    const float pr = mode < 2 ? avg[j] : (mode == 2 ? left[j] : F3s[j]);
    tg[j] = F1s[j] - pr;
  }
  unsigned long long best = kVqSentinel;
  for (int t0 = 0; t0 < 4096; t0 += kDiffTile) {
    __syncthreads();
    for (int i = tid; i < kDiffTile * kBands; i += kVqThreads)
      tile[(i / kBands) * kDiffStride + i % kBands] = __ldg(diff4 + (size_t)t0 * kBands + i);
    __syncthreads();
    if (valid) {
      for (int i = lane; i < kDiffTile; i += 32) {
        const float *c = tile + i * kDiffStride;
        float dm = 0, dp = 0;
#pragma unroll
        for (int j = 0; j < kBands; j++) {
          const float em = tg[j] - c[j], ep = tg[j] + c[j];
          dm += em * em;
          dp += ep * ep;
        }
        const unsigned long long km = vq_key(dm, t0 + i), kp = vq_key(dp, t0 + i + 4096);
        best = km < best ? km : best;
        best = kp < best ? kp : best;
      }
    }
  }
  if (!valid) return;
  best = warp_min_u64(best);
  const int vq_mid = key_index(best);            // sentinel: nothing below 1e15 -> entry 0, as in the reference
  const float sgn = vq_mid >= 4096 ? -1.f : 1.f;
  const int idm = vq_mid & 4095;
  if (lane < kBands) {
    const int m = idm & 3;
    const float pr = m < 2 ? avg[lane] : (m == 2 ? left[lane] : F3s[lane]);
    F1s[lane] = pr + sgn * __ldg(diff4 + idm * kBands + lane);
  }
  __syncwarp();
  // double_interp_search (lpcnet_enc.c:378-400) over interp_search (323-341)
  int interp_id = 0;
  if (lane == 0) {
    float dist[2][3];
    for (int h = 0; h < 2; h++) {
      const float *x = h ? F2s : F0s, *l = h ? F1s : left, *r = h ? F3s : F1s;
      for (int k = 0; k < 3; k++) {
        float d = 0;
        for (int i = 0; i < kBands; i++) {
          const float pr = k == 0 ? .5f * (l[i] + r[i]) : (k == 1 ? l[i] : r[i]);
          const float e = x[i] - pr;
          d += e * e;
        }
        dist[h][k] = d;
      }
    }
    float min_dist = 1e15f;
    int best_id = 0;
    for (int i = 0; i < 3; i++)
      for (int j = 0; j < 3; j++) {
        const int id = 3 * i + j;
        const float d = dist[0][i] + dist[1][j];
        if (d < min_dist && id != 7) { min_dist = d; best_id = id; }
      }
    interp_id = best_id - (best_id >= 7);
  }
  interp_id = __shfl_sync(0xffffffffu, interp_id, 0);
  // perform_double_interp (common.c:57-65)
  const int full = interp_id + (interp_id >= 7), id0 = full / 3, id1 = full % 3;
  if (lane < kBands) {
    const float l0 = left[lane], q1 = F1s[lane], q3 = F3s[lane];
    F[lane] = id0 == 0 ? .5f * (l0 + q1) : (id0 == 1 ? l0 : q1);
    F[kNF + lane] = q1;
    F[2 * kNF + lane] = id1 == 0 ? .5f * (q1 + q3) : (id1 == 1 ? q1 : q3);
    if (p == A.n_packets - 1) A.vq_mem[(size_t)s * kBands + lane] = q3;
  }
  if (lane == 0) {
    // bits_pack (lpcnet_enc.c:441-463): MSB first, 7+6+3+2+10+10+10+13+3 = 64 bits
    int *id = A.ids + ((size_t)s * A.n_packets + p) * kIds;
    id[7] = vq_mid; id[8] = interp_id;
    const int width[9] = {7, 6, 3, 2, 10, 10, 10, 13, 3};
    unsigned long long v = 0;
    for (int k = 0; k < 9; k++) v = (v << width[k]) | ((unsigned long long)(unsigned)id[k] & ((1ull << width[k]) - 1));
    unsigned char *out = A.packets + ((size_t)s * A.n_packets + p) * 8;
    for (int k = 0; k < 8; k++) out[k] = (unsigned char)(v >> (56 - 8 * k));
  }
}
