// The 16 kHz sample network as one persistent kernel: run_sample_network + the synthesis loop
// of the reference (src/lpcnet.c:146-167, 235-271; src/nnet.c:163-214, 326-372, 410-448, 484-491;
// src/vec_avx.h:321-336, 690-858) for a tile of NS = MT*VR independent streams per CTA.
//
// Layout of one CTA (512 threads, 16 warps), per sample step:
//   phase A  (all warps)  GRU-A.  Warp w owns unit groups {w, w+16, w+32} (8 units each) of every
//            16-row stream tile.  Per (unit group, tile) it gathers the three embedding rows and the
//            frame conditioning (fp32, L1/L2), seeds int32 accumulators exactly like the AVX2 kernel,
//            runs the block-sparse u8 x s8 product as mma.sync.m16n8k32 (streams = M, the 8 rows of a
//            row group = N, 8 blocks of 4 inputs = K; activations gathered from shared memory),
//            applies the Pade/rcpps gates in registers and keeps the fp32 state in registers.
//            The GRU-B input product (48x384) rides on phase A: a warp multiplies the state bytes it has just
//            produced (its own 24 units, still in registers) with its K slice of the matrix, one MMA per 8-row
//            group, and adds the partial int32 sums into shared memory (red.shared.add; integer addition is
//            associative, so the K split over the 16 warps is exact).  The recurrent 48x16 product runs on six
//            warps at the start of phase A.  No separate product phase, two barriers per step.
//   phase B2+T (16 lanes per stream) GRU-B gates, then the 8-level dual-FC tree sampler evaluated
//            speculatively in two rounds of 15 nodes (identical decisions to the sequential descent),
//            kiss99, mu-law, LPC prediction, de-emphasis and the PCM store.
// Weights, index lists and tables stay resident in shared memory for the life of the CTA.
#pragma once
#include <stdio.h>

#include "../../include/lpcnet_b200.h"
#include "bitexact.cuh"
#include "engine.h"
#include "tma.cuh"


namespace lpcb {

constexpr int kThreads = 512;
constexpr int kWarps = kThreads / 32;
constexpr int kUGroups = kN / 8;           // 48 unit groups
constexpr int kRGroupsB = kRowsB / 8;      // 6

struct SmemLayout {
  // byte offsets
  unsigned wA, posA, offF, csA, binF, wBr, posBr, csBr;
  unsigned sdA, subB, u2l, logit;
  unsigned xa, xb, bst, condb, part;
  unsigned usig, upred, lexc, deemph, rng, act, ix, tst, accx;
  unsigned bar, lpcb, condb1;  // mbarriers [1 + 2*G]; landing buffers of the bulk copies: LPC rows [G][2][NS][16], GRU-B conditioning [G][2][NS][48]
  unsigned fcw, fcb;   // dual_fc weights (node stride kFcStride floats) and bias|factor, when resident
  unsigned total;
};


__host__ __device__ inline bool fc_resident(int NS) { return NS <= 32; }

// G: tiles per CTA (1, or 2 as two groups of 8 warps: the per-tile buffers then exist once per group)
__host__ __device__ inline SmemLayout make_layout(int NS, int MT, int chA, bool fixed, int chBr, int G = 1) {
  SmemLayout L;
  unsigned o = 0;
  auto take = [&](unsigned bytes) { unsigned r = o; o += (bytes + 15u) & ~15u; return r; };
  if (NS == 8 && G == 1) {
    // 8 streams per CTA: the arrays whose size depends on the tile shape only come first, so their offsets fold to
    // constants (immediate operands instead of live registers or re-computed sums); the model-sized weight and
    // offset lists follow: 4.17 -> 4.13 us per step.  The larger tiles measured 2 % slower in this order
    // (register allocation) and keep the lists in front.
    L.offF = take(fixed ? kN / 8 * 32 * 4 : 0); L.csA = take((kN / 8 + 1) * 4);
    L.binF = take(kBinFragWords * 4);
    L.csBr = take((kRGroupsB + 1) * 4);
    L.sdA = take(kRowsA * 8); L.subB = take(2 * kRowsB * 4);
    L.u2l = take(256 * 4); L.logit = take(256 * 4);
    L.xa = take(2 * NS * kXaStride * 4);
    L.xb = take(NS * kXbStride * 4);
    L.bst = take(NS * 20 * 4);
    L.condb = take(NS * kRowsB * 4);
    L.part = take(MT * 2 * kRGroupsB * 128 * 4);
    L.usig = take(NS * 4); L.upred = take(NS * 4); L.lexc = take(NS * 4);
    L.deemph = take(NS * 4); L.rng = take(4 * NS * 4); L.act = take(NS * 4); L.ix = take(NS * 4);
    L.tst = 0;
    L.accx = take(12 * 3 * 32 * 8);   // GRU-A sums computed ahead for the sampler warps' unit groups
    L.fcb = take(4 * 256 * 4);
    L.fcw = take(256 * kFcStride * 4);
    L.wA = take(chA * 256); L.posA = take(chA * 16);
    L.wBr = take(chBr * 256); L.posBr = take(chBr * 16);
    // (after everything else, so that the arrays of the step loop keep the offsets they were tuned with)
    L.condb1 = take(2 * NS * kRowsB * 4); L.lpcb = take(2 * NS * 16 * 4); L.bar = take((1 + 2) * 8);
    L.total = o;
    return L;
  }
  L.wA = take(chA * 256); L.posA = take(chA * 16); L.offF = take(fixed ? kN / 8 * 32 * 4 : 0); L.csA = take((kN / 8 + 1) * 4);
  L.binF = take(kBinFragWords * 4);
  L.wBr = take(chBr * 256); L.posBr = take(chBr * 16); L.csBr = take((kRGroupsB + 1) * 4);
  L.sdA = take(kRowsA * 8); L.subB = take(2 * kRowsB * 4);
  L.u2l = take(256 * 4); L.logit = take(256 * 4);
  L.xa = take(G * 2 * NS * kXaStride * 4);
  L.xb = take(G * NS * kXbStride * 4);
  L.bst = take(G * NS * 20 * 4);
  L.condb = take(G * NS * kRowsB * 4);
  L.part = take(G * MT * 2 * kRGroupsB * 128 * 4);
  L.usig = take(G * NS * 4); L.upred = take(G * NS * 4); L.lexc = take(G * NS * 4);
  L.deemph = take(G * NS * 4); L.rng = take(G * 4 * NS * 4); L.act = take(G * NS * 4); L.ix = take(G * NS * 4);
  L.tst = NS >= 32 ? take(3 * kThreads * 4) : 0;
  L.accx = NS == 8 && G == 1 ? take(12 * 3 * 32 * 8) : 0;   // GRU-A sums computed ahead for the sampler warps' unit groups
  L.fcb = take(4 * 256 * 4);
  L.fcw = fc_resident(NS) ? take(256 * kFcStride * 4) : 0;
  L.condb1 = take(G * 2 * NS * kRowsB * 4); L.lpcb = take(G * 2 * NS * 16 * 4); L.bar = take((1 + 2 * G) * 8);
  L.total = o;
  return L;
}

__device__ __forceinline__ float2 ldg2(const float *p) { return __ldg(reinterpret_cast<const float2 *>(p)); }

// the six GRU-A input values of a lane for one unit group: p points at the lane's 32-byte sector
// [z0 z1 r0 r1 h0 h1 - -] of a row in gru_a_col_index sector order (engine.h)
__device__ __forceinline__ void ldg_zrh(const float *p, float2 &z, float2 &r, float2 &h) {
  // one 256-bit load (sm_100: LDG.E.256): a warp request then covers exactly one 128-byte line per stream,
  // 8 L1 wavefronts for all six values instead of 8 + 8 with a 16-byte and an 8-byte load
  float pad0, pad1;
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(z.x), "=f"(z.y), "=f"(r.x), "=f"(r.y), "=f"(h.x), "=f"(h.y), "=f"(pad0), "=f"(pad1) : "l"(p));
}
template <bool ARITH> struct RcpSel {
  typedef const uint32_t *type;
  static __device__ __forceinline__ type make(const uint32_t *tab) { return tab; }
};
template <> struct RcpSel<true> {
  typedef RcpArith type;
  static __device__ __forceinline__ type make(const uint32_t *) { return RcpArith{}; }
};

// one dual-FC node logit (src/nnet.c:192-202); bv = GRU-B state.  w: the node's 32 weights
// [channel][input], bf: bias|factor as [bias c0 | bias c1 | factor c0 | factor c1][256].
template <bool SMEM, class R>
__device__ __forceinline__ float node_logit(const float *__restrict__ wbase, int wstride, const float *__restrict__ bf,
                                            int i, const float (&bv)[16], R rcp) {
  const float4 *w4 = reinterpret_cast<const float4 *>(wbase + (size_t)i * wstride);
  float w[32];
#pragma unroll
  for (int k = 0; k < 8; k++) {
    const float4 v = SMEM ? w4[k] : __ldg(w4 + k);
    w[4 * k] = v.x; w[4 * k + 1] = v.y; w[4 * k + 2] = v.z; w[4 * k + 3] = v.w;
  }
  float sum1 = bf[i], sum2 = bf[256 + i];       // bias|factor are always resident
  const float f1 = bf[512 + i], f2 = bf[768 + i];
  // products two at a time (FMUL2), sums in the reference's order: each product is rounded before its add
#pragma unroll
  for (int j = 0; j < 16; j += 2) {
    const float2 b2 = make_float2(bv[j], bv[j + 1]);
    const float2 p1 = f2_mul(make_float2(w[j], w[j + 1]), b2), p2 = f2_mul(make_float2(w[16 + j], w[17 + j]), b2);
    sum1 = sum1 + p1.x; sum2 = sum2 + p2.x;
    sum1 = sum1 + p1.y; sum2 = sum2 + p2.y;
  }
  sum1 = f1 * tanh_ref(sum1, rcp);
  sum2 = f2 * tanh_ref(sum2, rcp);
  return sum1 + sum2;
}

__device__ __forceinline__ uint32_t fnv1a_bytes(const unsigned char *p, int n) {
  uint32_t h = 2166136261u;
  for (int i = 0; i < n; i++) { h ^= p[i]; h *= 16777619u; }
  return h;
}

// development builds: thread 0 of CTA 0 accumulates the cycles between phase marks
#ifdef LPCB_PHASE_CLOCKS
#define LPCB_MARK(k) do { if (blockIdx.x == 0 && threadIdx.x == 0) { const long long now_ = clock64(); clk_[k] += now_ - last_; last_ = now_; } } while (0)
#else
#define LPCB_MARK(k) do { } while (0)
#endif

// SUB: the launch advances a sub-range of the batch (SampleArgs::stream_begin / stream_end).  A template parameter
// because carrying the two bounds in the whole-batch kernel cost 3.5 % at 8 streams per CTA (4.18 -> 4.33 us per
// step, A/B on one B200): the whole-batch instantiation keeps the constants.
// SAT: the model holds int8 pairs that can saturate vpmaddubsw's s16 (engine.h DevSat); their weights are zero in the
// MMA operands and the saturated pair sums are added by a scalar pass after each product.  Built only on the generic
// variants (FIX = RA = false), so the instantiations clipped models run are unchanged.
// G = 2: two independent 8-stream tiles per CTA, each advanced by its own group of 8 warps on its own named barrier,
// sharing the one resident copy of the weights (two CTAs per SM cannot fit: 184 KB of shared memory each).  The
// groups run out of phase, so the serial sampler phase of one tile (2 busy warps of its 8) hides under the gate
// phase of the other: the throughput schedule for batches of more than one 8-stream tile per SM.
template <int MT, int VR, bool DEBUG, bool FIX, bool RA, bool SUB = false, bool SAT = false, int G = 1>
__global__ void __launch_bounds__(kThreads, 1) sample_kernel(const SampleArgs A) {
  static_assert(!SAT || (!FIX && !RA), "the s16 pass is built on the generic variants only");
  static_assert(G == 1 || (G == 2 && MT == 1 && VR == 8 && !DEBUG), "two groups: 8-stream tiles, production kernel");
  constexpr int WG = kWarps / G, TG = kThreads / G;   // warps / threads that advance one tile
  constexpr int UW = kUGroups / WG;                   // unit groups per warp (3, or 6 with two groups)
  constexpr int US = UW / 3;                          // K slices of the GRU-B input product per warp
  constexpr int NS = MT * VR;
  constexpr int NV = (VR == 16) ? 4 : 2;   // accumulator registers in use per thread
  constexpr bool CREG = (MT == 1 && VR == 8 && G == 1);   // frame conditioning of GRU-A held in registers (18 per thread)
  constexpr bool FCS = NS <= 32;            // dual_fc resident in shared memory (fc_resident)
  // 32 streams per CTA: phase A has no registers to spare, so the three phase-T values each thread carries
  // from step to step wait in shared memory while phase A runs (measured +2 %; -2 % for the smaller tiles)
  constexpr bool TST = NS >= 32;
  // 8 streams per CTA: while four warps run phase T of step i, the other twelve already form the recurrent GRU-A
  // sums of step i+1 (they need only the state of step i): their own three unit groups into registers and one
  // of the sampler warps' twelve into shared memory.  Phase A of step i+1 then starts at the gates.
  constexpr bool OVL = (MT == 1 && VR == 8 && G == 1);
  constexpr bool TMAW = (NS == 16 && G == 1);   // weights staged by bulk copies (see the prologue)
  // per-frame conditioning fetched one frame ahead by bulk copies: everywhere but at 16 streams per CTA, where the same
  // A/B gives 6.44 us per step without it and 6.59 with it (8 streams per CTA: 3.58 either way)
  constexpr bool TMAF = !(NS == 16 && G == 1);
  extern __shared__ __align__(16) unsigned char smem[];
  const DevModel &M = A.m;
  const SmemLayout L = make_layout(NS, MT, M.a_rec.n_chunks, FIX, M.b_rec.n_chunks, G);

  uint32_t *s_wA = reinterpret_cast<uint32_t *>(smem + L.wA);
  uint32_t *s_offA = reinterpret_cast<uint32_t *>(smem + L.posA);   // [chunk][4] packed gather offsets
  uint32_t *s_offF = reinterpret_cast<uint32_t *>(smem + L.offF);   // [ug][lane%4][8] offsets of the fixed chunks
  uint32_t *s_csA = reinterpret_cast<uint32_t *>(smem + L.csA);     // [ug + 1] overflow chunk lists
  uint32_t *s_binF = reinterpret_cast<uint32_t *>(smem + L.binF);     // GRU-B input B fragments [warp][row group][lane][2]
  uint32_t *s_wBr = reinterpret_cast<uint32_t *>(smem + L.wBr);
  uint32_t *s_offBr = reinterpret_cast<uint32_t *>(smem + L.posBr);
  uint32_t *s_csBr = reinterpret_cast<uint32_t *>(smem + L.csBr);
  float *s_sdA = reinterpret_cast<float *>(smem + L.sdA);   // [gate][ug][lane%4][sub0 sub1 diag0 diag1]
  float *s_subB = reinterpret_cast<float *>(smem + L.subB);
  // rcpps: arithmetic, or (trace / logit builds, models on the s16 pass, GPUs where the arithmetic form failed its
  // check) the 2048-entry table read through L1 -- not worth 8 KB of shared memory any more
  const typename RcpSel<RA>::type s_rcp = RcpSel<RA>::make(M.rcp_tab);
  float *s_u2l = reinterpret_cast<float *>(smem + L.u2l);
  float *s_logit = reinterpret_cast<float *>(smem + L.logit);
  uint32_t *s_xa = reinterpret_cast<uint32_t *>(smem + L.xa);
  uint32_t *s_xb = reinterpret_cast<uint32_t *>(smem + L.xb);
  float *s_bst = reinterpret_cast<float *>(smem + L.bst);
  float *s_condb = reinterpret_cast<float *>(smem + L.condb);   // [stream][48] of the frame in progress (fixed offset: phase B2 reads it every step)
  float *s_condb1 = reinterpret_cast<float *>(smem + L.condb1); // [frame & 1][stream][48]: where the bulk copies land
  float *s_lpcb = reinterpret_cast<float *>(smem + L.lpcb);     // [frame & 1][stream][16]
  uint64_t *s_bar = reinterpret_cast<uint64_t *>(smem + L.bar);  // [0] weights; [1 + 2*group + (frame & 1)] conditioning
  int32_t *s_part = reinterpret_cast<int32_t *>(smem + L.part);
  int *s_usig = reinterpret_cast<int *>(smem + L.usig);
  int *s_upred = reinterpret_cast<int *>(smem + L.upred);
  int *s_lexc = reinterpret_cast<int *>(smem + L.lexc);
  float *s_deemph = reinterpret_cast<float *>(smem + L.deemph);
  uint32_t *s_rng = reinterpret_cast<uint32_t *>(smem + L.rng);
  int *s_act = reinterpret_cast<int *>(smem + L.act);
  uint32_t *s_ix = reinterpret_cast<uint32_t *>(smem + L.ix);
  float *s_tst = reinterpret_cast<float *>(smem + L.tst);
  int2 *s_accx = reinterpret_cast<int2 *>(smem + L.accx);   // [job of a sampler warp: u*4 + warp][gate][lane]   // u_sig | u_pred << 8 | last_exc << 16 | active << 24
  // dual_fc weights: resident copy (padded node stride) or the blob array in global memory; bias|factor resident
  const float *fc_w = FCS ? reinterpret_cast<const float *>(smem + L.fcw) : M.fc_w;
  const int fc_stride = FCS ? kFcStride : 32;
  const float *fc_bf = reinterpret_cast<const float *>(smem + L.fcb);

  // tid / warp: position within the group that advances this thread's tile (the whole CTA when G = 1)
  const int tid0 = threadIdx.x, grp_id = G == 1 ? 0 : tid0 / TG;
  const int tid = tid0 - grp_id * TG, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  if (G > 1) {
    s_xa += grp_id * 2 * NS * kXaStride; s_xb += grp_id * NS * kXbStride; s_bst += grp_id * NS * 20;
    s_condb += grp_id * NS * kRowsB; s_condb1 += grp_id * 2 * NS * kRowsB; s_lpcb += grp_id * 2 * NS * 16; s_part += grp_id * MT * 2 * kRGroupsB * 128;
    s_usig += grp_id * NS; s_upred += grp_id * NS; s_lexc += grp_id * NS; s_deemph += grp_id * NS;
    s_rng += grp_id * 4 * NS; s_act += grp_id * NS; s_ix += grp_id * NS;
  }
  auto gsync = [&]() {
    if (G == 1) __syncthreads();
    else asm volatile("bar.sync %0, %1;" ::"r"(1 + grp_id), "n"(TG) : "memory");
  };
  const int B = A.n_streams;                                   // row stride of the per-frame arrays
  const int Bend = SUB && A.stream_end ? A.stream_end : B;      // this launch advances streams [sbegin, Bend)
  const int sbegin = SUB ? A.stream_begin : 0;
  const int row_len = A.total_frames * A.spf;

  // sparse_sgemv_accum8x4 (src/vec_avx.h:790-858) of unit group ug for the M tile at xrow0 on tensor cores: the
  // three gates' int32 sums without their seeds (the sum is exact in any order; the rne(16256*t) seed of the AVX2
  // kernel is added afterwards).
  auto gru_a_product = [&](int ug, const uint32_t *xrow0, int32_t (&acc)[3][4]) {
#pragma unroll
    for (int gate = 0; gate < 3; gate++) acc[gate][0] = acc[gate][1] = acc[gate][2] = acc[gate][3] = 0;
    const unsigned char *xrowb = reinterpret_cast<const unsigned char *>(xrow0 + g * kXaStride);
    auto chunk = [&](int32_t (&d)[4], uint32_t lo, uint32_t hi, uint2 bw) {
      const uint32_t a0 = *reinterpret_cast<const uint32_t *>(xrowb + lo);
      const uint32_t a2 = *reinterpret_cast<const uint32_t *>(xrowb + hi);
      uint32_t a1 = 0, a3 = 0;
      if (VR == 16) {
        a1 = *reinterpret_cast<const uint32_t *>(xrowb + 8 * kXaStride * 4 + lo);
        a3 = *reinterpret_cast<const uint32_t *>(xrowb + 8 * kXaStride * 4 + hi);
      }
      mma_u8s8(d, a0, a1, a2, a3, bw.x, bw.y);
    };
    if (FIX) {
      // fixed slots of this unit group: 1 chunk for z, 1 for r, 3 for h -- fully unrolled
      constexpr int ST = 5;
      const uint2 *wq = reinterpret_cast<const uint2 *>(s_wA) + ug * ST * 32 + lane;
      const uint4 o4 = *reinterpret_cast<const uint4 *>(s_offF + (ug * 4 + t) * 8);
      const uint32_t pk[ST] = {o4.x, o4.y, o4.z, o4.w, s_offF[(ug * 4 + t) * 8 + 4]};
      uint2 bw[ST];
#pragma unroll
      for (int sl = 0; sl < ST; sl++) bw[sl] = wq[sl * 32];
#pragma unroll
      for (int sl = 0; sl < ST; sl++) chunk(acc[sl < 2 ? sl : 2], pk[sl] & 0xffffu, pk[sl] >> 16, bw[sl]);
    }
    // overflow chunks (all chunks when the model does not fit the fixed slots); bits 30-31 of the
    // offset word name the gate
    const int c0 = s_csA[ug], c1 = s_csA[ug + 1];
#pragma unroll 1
    for (int c = c0; c < c1; c++) {   // rarely more than one trip: keep the code small (instruction cache)
      const uint32_t pk = s_offA[c * 4 + t];
      const uint2 bw = reinterpret_cast<const uint2 *>(s_wA)[c * 32 + lane];
      const uint32_t lo = pk & 0xffffu, hi = (pk >> 16) & 0x3fffu, gate = pk >> 30;
      if (gate == 0) chunk(acc[0], lo, hi, bw);
      else if (gate == 1) chunk(acc[1], lo, hi, bw);
      else chunk(acc[2], lo, hi, bw);
    }
    if (SAT && M.a_rec.sat.n) {
      // lane (g, t) holds outputs 2t, 2t+1 of the 8-row group for stream g (and g + 8): entry row r belongs to t = r / 2
      const int e1 = (int)__ldg(M.a_rec.sat.start + ug + 1);
      for (int e = (int)__ldg(M.a_rec.sat.start + ug); e < e1; e++) {
        const uint2 f = __ldg(M.a_rec.sat.e + e);
        const int r = (f.x >> 16) & 7, gate = f.x >> 20;
        if ((r >> 1) != t) continue;
        const int v0 = sat16_pair(f, xrowb), v1 = VR == 16 ? sat16_pair(f, xrowb + 8 * kXaStride * 4) : 0;
#pragma unroll
        for (int gt = 0; gt < 3; gt++)
#pragma unroll
          for (int q = 0; q < 2; q++) {
            const bool hit = gate == gt && (r & 1) == q;
            acc[gt][q] += hit ? v0 : 0;
            acc[gt][2 + q] += hit ? v1 : 0;
          }
      }
    }
  };

  // ---- one-time: weights, index lists and tables into shared memory, as bulk asynchronous copies (tma.cuh): one
  // thread arms the barrier with the byte count and issues them; everybody waits on the barrier ----
  {
    if (tid0 == 0) {
      for (int k = 0; k < 1 + 2 * G; k++) mbar_init(s_bar + k, 1);
      mbar_fence_init();
    }
    __syncthreads();
    // Staged by bulk copies where that is also the faster build: the step loop is the same source either way, but the
    // two prologues steer the register allocation of the whole kernel differently -- measured A/B on one B200, bulk
    // copies: 6.52 -> 6.44 us per step at 16 streams per CTA, 3.58 -> 3.75 at 8 (where the threads copy, as before).
    if (!TMAW) {
      auto copy_words = [&](void *dstv, const void *srcv, uint32_t bytes) {
        uint32_t *dst = (uint32_t *)dstv; const uint32_t *src = (const uint32_t *)srcv;
        for (int i = tid0; i < (int)(bytes / 4); i += kThreads) dst[i] = __ldg(src + i);
      };
      copy_words(s_wA, M.a_rec.w, (uint32_t)M.a_rec.n_chunks * 256u);
      copy_words(s_offA, M.a_rec.off, (uint32_t)M.a_rec.n_chunks * 16u);
      if (FIX) copy_words(s_offF, M.a_rec.off_fixed, kUGroups * 32u * 4u);
      copy_words(s_csA, M.a_rec.chunk_start, (kUGroups + 1) * 4u);
      copy_words(s_binF, M.b_in_frag, kBinFragWords * 4u);
      copy_words(s_wBr, M.b_rec.w, (uint32_t)M.b_rec.n_chunks * 256u);
      copy_words(s_offBr, M.b_rec.off, (uint32_t)M.b_rec.n_chunks * 16u);
      copy_words(s_csBr, M.b_rec.chunk_start, (kRGroupsB + 1) * 4u);
      copy_words(s_sdA, M.a_sd_packed, kRowsA * 8u);
      copy_words(s_subB, M.b_subias, 2u * kRowsB * 4u);
      copy_words(s_u2l, M.ulaw2lin_tab, 256u * 4u);
      copy_words(s_logit, M.logit_tab, 256u * 4u);
      copy_words(smem + L.fcb, M.fc_bf, 1024u * 4u);
      if (FCS) copy_words(smem + L.fcw, M.fc_w_padded, 256u * kFcStride * 4u);
    }
    if (TMAW && tid0 == 0) {
      struct Job { void *dst; const void *src; uint32_t bytes; };
      const Job jobs[] = {
          {s_wA, M.a_rec.w, (uint32_t)M.a_rec.n_chunks * 256u},
          {s_offA, M.a_rec.off, (uint32_t)M.a_rec.n_chunks * 16u},
          {s_offF, M.a_rec.off_fixed, FIX ? kUGroups * 32u * 4u : 0u},
          {s_csA, M.a_rec.chunk_start, (kUGroups + 1) * 4u},
          {s_binF, M.b_in_frag, kBinFragWords * 4u},
          {s_wBr, M.b_rec.w, (uint32_t)M.b_rec.n_chunks * 256u},
          {s_offBr, M.b_rec.off, (uint32_t)M.b_rec.n_chunks * 16u},
          {s_csBr, M.b_rec.chunk_start, (kRGroupsB + 1) * 4u},
          {s_sdA, M.a_sd_packed, kRowsA * 8u},
          {s_subB, M.b_subias, 2u * kRowsB * 4u},
          {s_u2l, M.ulaw2lin_tab, 256u * 4u},
          {s_logit, M.logit_tab, 256u * 4u},
          {smem + L.fcb, M.fc_bf, 1024u * 4u},
          {smem + L.fcw, M.fc_w_padded, FCS ? 256u * kFcStride * 4u : 0u},
      };
      uint32_t total = 0;
      for (const Job &j : jobs) total += (j.bytes + 15u) & ~15u;     // (the model's arrays are readable to the next 16 bytes)
      mbar_arrive_expect_tx(s_bar, total);
      for (const Job &j : jobs)
        if (j.bytes) bulk_g2s(j.dst, j.src, (j.bytes + 15u) & ~15u, s_bar);
    }
    // the sums of the GRU-B input product start at zero; phase B2 re-zeroes them
    for (int i = tid0; i < G * MT * 2 * kRGroupsB * 128; i += kThreads) (s_part - grp_id * MT * 2 * kRGroupsB * 128)[i] = 0;
    if (TMAW) mbar_wait(s_bar, 0);
  }
  __syncthreads();     // the groups go their own ways from here

#ifdef LPCB_PHASE_CLOCKS
  long long clk_[10] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, last_ = clock64();
#endif
  const int n_tiles = (Bend - sbegin + NS - 1) / NS;
  uint32_t bar_phase = 0;      // bit b: the phase the next wait on conditioning buffer b expects
  if (G > 1 && grp_id == 1) __nanosleep(1800);     // start half a step apart; the phases keep the groups apart afterwards
  for (int tile = blockIdx.x * G + grp_id; tile < n_tiles; tile += gridDim.x * G) {
    const int stream0 = sbegin + tile * NS;
    gsync();
    // ---- load the tile's stream state ----
    float2 a_reg[UW][MT][NV / 2];
    int cur = 0;
#pragma unroll
    for (int u = 0; u < UW; u++) {
      const int ug = warp + u * WG;
#pragma unroll
      for (int mt = 0; mt < MT; mt++) {
#pragma unroll
        for (int h = 0; h < NV / 2; h++) {
          const int s = mt * VR + g + h * 8;
          const int gs = min(stream0 + s, Bend - 1);
          const int unit = ug * 8 + 2 * t;
          const float2 v = *reinterpret_cast<const float2 *>(A.st.gru_a + (size_t)gs * kN + unit);
          a_reg[u][mt][h] = v;
          reinterpret_cast<unsigned char *>(s_xa)[(size_t)s * kXaStride * 4 + unit] = (unsigned char)quant_u8_ref(v.x);
          reinterpret_cast<unsigned char *>(s_xa)[(size_t)s * kXaStride * 4 + unit + 1] = (unsigned char)quant_u8_ref(v.y);
        }
      }
    }
    for (int i = tid; i < NS * 16; i += TG) {
      const int s = i >> 4, j = i & 15;
      const int gs = min(stream0 + s, Bend - 1);
      float v = A.st.gru_b[(size_t)gs * kNB + j];
      s_bst[s * 20 + j] = v;
      reinterpret_cast<unsigned char *>(s_xb)[s * kXbStride * 4 + j] = (unsigned char)quant_u8_ref(v);
    }
    // Phase T state of the 16-lane group that owns stream `grp` of the tile (NS <= 32 groups): lane j holds
    // last_sig[j] (j = 0 newest) and lpc[j]; the LPC prediction is known to all 16 lanes.
    // Sampler warps: the first NS/2 warps (one per SM sub-partition when there are four; putting the four on one
    // sub-partition, away from the product warps of OVL, measured 3 % slower).
    const int tw = 2 * warp < NS ? warp : -1;
    const bool twarp = tw >= 0;                             // both 16-lane halves own a stream (NS is even)
    const int l16 = tid & 15, grp = twarp ? 2 * tw + (lane >> 4) : NS;
    const unsigned gbase = lane & 16;
    const int tgs = min(stream0 + min(grp, NS - 1), Bend - 1);
    float t_x = A.st.last_sig[(size_t)tgs * 16 + l16], t_lpc = 0.f, t_pred = 0.f;
    int32_t accp[OVL ? 3 : 1][3][2];
    auto product_ahead = [&](const uint32_t *xa_buf) {   // the twelve non-sampler warps (OVL)
      int32_t acc[3][4];
#pragma unroll
      for (int u = 0; u < 3; u++) {
        gru_a_product(warp + u * WG, xa_buf, acc);
#pragma unroll
        for (int gate = 0; gate < 3; gate++) { accp[OVL ? u : 0][gate][0] = acc[gate][0]; accp[OVL ? u : 0][gate][1] = acc[gate][1]; }
      }
      const int e = warp - 4;                              // 0..11: job e / 4 of sampler warp e % 4
      gru_a_product((e & 3) + WG * (e >> 2), xa_buf, acc);
#pragma unroll
      for (int gate = 0; gate < 3; gate++) s_accx[(e * 3 + gate) * 32 + lane] = make_int2(acc[gate][0], acc[gate][1]);
    };
    for (int s = tid; s < NS; s += TG) {
      const int gs = min(stream0 + s, Bend - 1);
      s_deemph[s] = A.st.deemph[gs];
      s_lexc[s] = A.st.last_exc[gs];
#pragma unroll
      for (int k = 0; k < 4; k++) s_rng[k * NS + s] = A.st.rng[(size_t)gs * 4 + k];
    }
    // The frame-rate inputs of the tile -- GRU-B conditioning [stream][48] and LPC [stream][16], contiguous over the
    // tile's streams -- arrive by bulk copy into one of two buffers: frame f + 1 is fetched while frame f's 160 steps
    // run.  A last, part-filled tile copies its own rows only; the rest of the buffers is cleared once.
    const int tile_rows = min(NS, Bend - stream0);
    if (tile_rows < NS) {
      for (int i = tid; i < 2 * NS * kRowsB; i += TG) if ((i % (NS * kRowsB)) >= tile_rows * kRowsB) s_condb1[i] = 0.f;
      for (int i = tid; i < 2 * NS * 16; i += TG) if ((i % (NS * 16)) >= tile_rows * 16) s_lpcb[i] = 0.f;
    }
    auto fetch_frame = [&](int f) {   // one thread of the group
      uint64_t *bar = s_bar + 1 + 2 * grp_id + (f & 1);
      fence_proxy_async();            // the buffer's earlier reads (two frames ago) are ordered before the engine's writes
      mbar_arrive_expect_tx(bar, (uint32_t)tile_rows * (kRowsB + 16) * 4u);
      bulk_g2s(s_condb1 + (f & 1) * NS * kRowsB, A.cond_b + ((size_t)f * B + stream0) * kRowsB, (uint32_t)tile_rows * kRowsB * 4u, bar);
      bulk_g2s(s_lpcb + (f & 1) * NS * 16, A.lpc + ((size_t)f * B + stream0) * 16, (uint32_t)tile_rows * 16u * 4u, bar);
    };
    gsync();
    if (TMAF && tid == 0) fetch_frame(0);
    if (OVL && !twarp) product_ahead(s_xa + cur * NS * kXaStride);   // sums for the tile's first step

    int ever_act = 0;   // (threads < NS) was this stream active in any frame of the launch?
    for (int f = 0; f < A.n_frames; f++) {
      // ---- frame prologue: LPC, GRU-B conditioning, active flags ----
      // (every thread of the group is past frame f - 1 here: its reads of the other buffer ended before the last barrier)
      if (TMAF) {
        if (tid == 0 && f + 1 < A.n_frames) fetch_frame(f + 1);
        mbar_wait(s_bar + 1 + 2 * grp_id + (f & 1), (bar_phase >> (f & 1)) & 1u);
        bar_phase ^= 1u << (f & 1);
        t_lpc = s_lpcb[(f & 1) * NS * 16 + min(grp, NS - 1) * 16 + l16];
        for (int i = tid; i < NS * kRowsB; i += TG) s_condb[i] = s_condb1[(f & 1) * NS * kRowsB + i];   // (ordered by the barrier below)
      } else {
        t_lpc = A.lpc[((size_t)f * B + tgs) * 16 + l16];
        for (int i = tid; i < NS * kRowsB; i += TG) {
          const int s = i / kRowsB, j = i % kRowsB;
          const int gs = min(stream0 + s, Bend - 1);
          s_condb[i] = A.cond_b[((size_t)f * B + gs) * kRowsB + j];
        }
      }
      int my_act = 0;
      if (tid < NS) {
        const int gs = stream0 + tid;
        my_act = (gs < Bend) ? (int)A.active[(size_t)f * B + gs] : 0;
        s_act[tid] = my_act;
        ever_act |= my_act;
      }
      int any_act;
      if (G == 1) {
        any_act = __syncthreads_or(my_act);
      } else {
        gsync();
        any_act = 0;
#pragma unroll
        for (int k = 0; k < NS; k++) any_act |= s_act[k];
      }
      const size_t out_base = (size_t)(A.frame0 + f) * A.spf;
      if (!any_act) {
        // warm-up frames: N zeros, nothing else moves (src/lpcnet.c:239-243)
        {
          for (int i = tid; i < NS * A.spf; i += TG) {
            const int s = i / A.spf, k = i % A.spf;
            if (stream0 + s < Bend) {
              A.pcm[(size_t)(stream0 + s) * row_len + out_base + k] = 0;
              if (DEBUG && A.trace) {
                LPCNetB200Trace tr = {0, 0, 0, 0, 0, 0, 0, 0};
                reinterpret_cast<LPCNetB200Trace *>(A.trace)[(size_t)(stream0 + s) * row_len + out_base + k] = tr;
              }
            }
          }
          if (DEBUG && A.logits) {
            for (size_t i = tid; i < (size_t)NS * A.spf * 256; i += TG) {
              const int s = (int)(i / ((size_t)A.spf * 256));
              const size_t r = i % ((size_t)A.spf * 256);
              if (stream0 + s < Bend) A.logits[((size_t)(stream0 + s) * row_len + out_base) * 256 + r] = 0.f;
            }
          }
        }
        continue;
      }
      // LPC prediction + mu-law indices of the next step (src/lpcnet.c:252-254), by the 16 lanes of a stream:
      // lane j forms last_sig[j]*lpc[j]; the sum runs in the reference's order j = 0..15 over shuffles.
      auto predict = [&]() {
        const float p = t_x * t_lpc;
        float pred = 0.f;
#pragma unroll
        for (int j = 0; j < 16; j++) pred = pred - __shfl_sync(0xffffffffu, p, gbase + j);
        t_pred = pred;
        const int us = lin2ulaw_ref(__shfl_sync(0xffffffffu, t_x, gbase)), up = lin2ulaw_ref(pred);
        if (l16 == 0 && grp < NS) {
          if (DEBUG) { s_usig[grp] = us; s_upred[grp] = up; }
          s_ix[grp] = (uint32_t)us | (uint32_t)up << 8 | (uint32_t)s_lexc[grp] << 16 | (uint32_t)s_act[grp] << 24;
        }
      };
      if (twarp) predict();
      if (TST) { s_tst[tid] = t_x; s_tst[TG + tid] = t_lpc; s_tst[2 * TG + tid] = t_pred; }
      gsync();

      const float *condA_f = A.cond_a + (size_t)f * B * kRowsAP;
      float2 cond_reg[UW][3];
      if (CREG) {
        const int gs = min(stream0 + g, Bend - 1);
#pragma unroll
        for (int u = 0; u < 3; u++)
          ldg_zrh(condA_f + (size_t)gs * kRowsAP + ((warp + u * WG) * 4 + t) * 8, cond_reg[u][0], cond_reg[u][1], cond_reg[u][2]);
      }

      LPCB_MARK(0);   // frame prologue and everything outside the step loop
      for (int step = 0; step < A.spf; step++) {
        uint32_t *xa_cur = s_xa + cur * NS * kXaStride;
        uint32_t *xa_new = s_xa + (cur ^ 1) * NS * kXaStride;
        // ================= phase A: GRU-A =================
        // per-stream indices of this step (the same for the three unit groups of the warp)
        // rp: row pointers (this lane's sector of unit group `warp`) of the three embedding rows, held for the
        // step when there are registers for them; a unit group further on is an immediate offset
        constexpr bool PH = MT * (NV / 2) <= 2;
        const int lane_off = (warp * 4 + t) * 8;
        uint32_t ixw[MT][NV / 2];
        const float *rp[PH ? MT : 1][PH ? NV / 2 : 1][3];
#pragma unroll
        for (int mt = 0; mt < MT; mt++)
#pragma unroll
          for (int h = 0; h < NV / 2; h++) {
            const uint32_t w = s_ix[mt * VR + g + 8 * h];
            ixw[mt][h] = w;
            if (PH) {
              rp[mt][h][0] = M.e_sig + (w & 0xffu) * kRowsAP + lane_off;
              rp[mt][h][1] = M.e_pred + ((w >> 8) & 0xffu) * kRowsAP + lane_off;
              rp[mt][h][2] = M.e_exc + ((w >> 16) & 0xffu) * kRowsAP + lane_off;
            }
          }
        // GRU-B recurrent product (48 x 16 on the GRU-B state of the previous step, src/nnet.c:355-361): one row group
        // per warp on warps 10..15 (never sampler warps at 8 or 16 streams per CTA); phase B2 reads the sums after
        // the barrier
        if (warp >= WG - kRGroupsB) {
          const int rg = warp - (WG - kRGroupsB);
          int32_t acc[MT][4];
#pragma unroll
          for (int mt = 0; mt < MT; mt++) acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0;
          const unsigned char *xb0 = reinterpret_cast<const unsigned char *>(s_xb + g * kXbStride);
          for (int c = s_csBr[rg]; c < (int)s_csBr[rg + 1]; c++) {
            const uint32_t pk = s_offBr[c * 4 + t];
            const uint2 bw = reinterpret_cast<const uint2 *>(s_wBr)[c * 32 + lane];
            const uint32_t lo = pk & 0xffffu, hi2 = pk >> 16;
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
              const unsigned char *xr = xb0 + mt * VR * kXbStride * 4;
              const uint32_t a0 = *reinterpret_cast<const uint32_t *>(xr + lo), a2 = *reinterpret_cast<const uint32_t *>(xr + hi2);
              uint32_t a1 = 0, a3 = 0;
              if (VR == 16) {
                a1 = *reinterpret_cast<const uint32_t *>(xr + 8 * kXbStride * 4 + lo);
                a3 = *reinterpret_cast<const uint32_t *>(xr + 8 * kXbStride * 4 + hi2);
              }
              mma_u8s8(acc[mt], a0, a1, a2, a3, bw.x, bw.y);
            }
          }
          if (SAT && M.b_rec.sat.n) {
            const int e1 = (int)__ldg(M.b_rec.sat.start + rg + 1);
            for (int e = (int)__ldg(M.b_rec.sat.start + rg); e < e1; e++) {
              const uint2 f = __ldg(M.b_rec.sat.e + e);
              const int r = (f.x >> 16) & 7;
              if ((r >> 1) != t) continue;
#pragma unroll
              for (int mt = 0; mt < MT; mt++) {
                const unsigned char *xr = xb0 + mt * VR * kXbStride * 4;
                const int v0 = sat16_pair(f, xr), v1 = VR == 16 ? sat16_pair(f, xr + 8 * kXbStride * 4) : 0;
#pragma unroll
                for (int q = 0; q < 2; q++) {
                  acc[mt][q] += (r & 1) == q ? v0 : 0;
                  acc[mt][2 + q] += (r & 1) == q ? v1 : 0;
                }
              }
            }
          }
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            int32_t *dst = s_part + ((mt * 2 + 1) * kRGroupsB + rg) * 128;
            *reinterpret_cast<int2 *>(dst + g * 8 + 2 * t) = make_int2(acc[mt][0], acc[mt][1]);
            if (VR == 16) *reinterpret_cast<int2 *>(dst + (g + 8) * 8 + 2 * t) = make_int2(acc[mt][2], acc[mt][3]);
          }
        }
        // the new state bytes of this lane, as A fragments of the GRU-B input product (engine.h kBinFragWords):
        // qa0 = [ug w: 2t 2t+1 | ug w+16: 2t 2t+1], qa2 = [ug w+32: 2t 2t+1 | 0 0], per stream row
        // (with two groups a warp owns two such K slices: unit groups w + 8*set + 16*slot)
        uint32_t qa0[US][MT][NV / 2], qa2[US][MT][NV / 2];
        // p + c for a p that comes out of a packed multiply.  ptxas contracts mul.rn.f32x2 -> add.rn.f32x2 into FFMA2
        // (it never contracts the scalar .rn forms), which would drop the product's rounding; an fma with a one it
        // cannot see (SampleArgs::one2, a kernel argument) is the exactly rounded packed sum and stays apart.  Worth
        // 0.6 % at 16 and 1.6 % at 32 streams per CTA; at 8 the two extra registers cost 2 %, so the sums stay scalar.
        auto padd = [&](float2 p, float2 c) {
          if (NS > 8) return f2_fma(p, A.one2, c);
          return make_float2(p.x + c.x, p.y + c.y);
        };
#pragma unroll
        for (int u = 0; u < UW; u++) {
          const int ug = warp + u * WG;
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            const int sA = mt * VR + g;
            // (1) issue the gathers first: compute_gru_a_input (src/nnet.c:484-491) reads the frame
            //     conditioning and one row of each embedding table; none of it feeds the recurrent product
            float2 cv[3][NV / 2], e1[3][NV / 2], e2[3][NV / 2], e3[3][NV / 2];
#pragma unroll
            for (int h = 0; h < NV / 2; h++) {
              const int gs = min(stream0 + sA + 8 * h, Bend - 1);
              if (CREG) { cv[0][h] = cond_reg[u][0]; cv[1][h] = cond_reg[u][1]; cv[2][h] = cond_reg[u][2]; }
              else ldg_zrh(condA_f + (size_t)gs * kRowsAP + lane_off + u * (WG * 32), cv[0][h], cv[1][h], cv[2][h]);
              if (PH) {
                ldg_zrh(rp[mt][h][0] + u * (WG * 32), e1[0][h], e1[1][h], e1[2][h]);
                ldg_zrh(rp[mt][h][1] + u * (WG * 32), e2[0][h], e2[1][h], e2[2][h]);
                ldg_zrh(rp[mt][h][2] + u * (WG * 32), e3[0][h], e3[1][h], e3[2][h]);
              } else {
                const uint32_t w = ixw[mt][h];
                const int off = lane_off + u * (WG * 32);
                ldg_zrh(M.e_sig + (w & 0xffu) * kRowsAP + off, e1[0][h], e1[1][h], e1[2][h]);
                ldg_zrh(M.e_pred + ((w >> 8) & 0xffu) * kRowsAP + off, e2[0][h], e2[1][h], e2[2][h]);
                ldg_zrh(M.e_exc + ((w >> 16) & 0xffu) * kRowsAP + off, e3[0][h], e3[1][h], e3[2][h]);
              }
            }
            // (2) the recurrent product: formed here, or already at hand from the previous step's phase T
            int32_t acc[3][4];
            if (!OVL) {
              gru_a_product(ug, xa_cur + mt * VR * kXaStride, acc);
            } else if (!twarp) {
#pragma unroll
              for (int gate = 0; gate < 3; gate++) { acc[gate][0] = accp[u][gate][0]; acc[gate][1] = accp[u][gate][1]; acc[gate][2] = acc[gate][3] = 0; }
            } else {
#pragma unroll
              for (int gate = 0; gate < 3; gate++) {
                const int2 v = s_accx[((u * 4 + tw) * 3 + gate) * 32 + lane];
                acc[gate][0] = v.x; acc[gate][1] = v.y; acc[gate][2] = acc[gate][3] = 0;
              }
            }
            // (3) seeds and gates: compute_sparse_gru (src/nnet.c:431-447), two units per packed operation
            float2 gh[NV / 2];
#pragma unroll
            for (int gate = 0; gate < 3; gate++) {
              const float4 sd4 = *reinterpret_cast<const float4 *>(s_sdA + ((gate * kUGroups + ug) * 4 + t) * 4);
              const float2 sub = make_float2(sd4.x, sd4.y), dg = make_float2(sd4.z, sd4.w);
#pragma unroll
              for (int h = 0; h < NV / 2; h++) {
                const float2 gx = f2_add(f2_add(f2_add(cv[gate][h], e1[gate][h]), e2[gate][h]), e3[gate][h]);
                // bias + diag*state: a packed product feeding a sum goes through padd (above)
                const float2 a2 = a_reg[u][mt][h];
                const float2 da = f2_mul(dg, a2);
                float2 tx = padd(da, sub);                                // = sub + da
                if (gate < 2) tx = f2_add(tx, gx);                        // (+ input for z, r)
                else gh[h] = gx;
                const float2 sd = f2_mul(tx, f2_set(LPCB_SCALE));
                if (RA) {   // (RA variants run only on models with small sums: bitexact.cuh acc_seed_small)
                  acc[gate][2 * h] = (int32_t)((uint32_t)acc[gate][2 * h] + __float_as_uint(__fadd_rn(sd.x, LPCB_MAGIC)));
                  acc[gate][2 * h + 1] = (int32_t)((uint32_t)acc[gate][2 * h + 1] + __float_as_uint(__fadd_rn(sd.y, LPCB_MAGIC)));
                } else {
                  acc[gate][2 * h] = (int32_t)((uint32_t)acc[gate][2 * h] + (uint32_t)__float2int_rn(sd.x));
                  acc[gate][2 * h + 1] = (int32_t)((uint32_t)acc[gate][2 * h + 1] + (uint32_t)__float2int_rn(sd.y));
                }
              }
            }
#pragma unroll
            for (int h = 0; h < NV / 2; h++) {
              const float2 k1 = f2_set(LPCB_SCALE_1);
              auto sum_f = [&](int gate) {   // the gate's two int32 sums as floats
                if (RA) return f2_add(make_float2(__uint_as_float((uint32_t)acc[gate][2 * h]), __uint_as_float((uint32_t)acc[gate][2 * h + 1])), f2_set(-LPCB_MAGIC));
                return make_float2((float)acc[gate][2 * h], (float)acc[gate][2 * h + 1]);
              };
              const float2 z = sigmoid2_ref(f2_mul(sum_f(0), k1), s_rcp);
              const float2 r = sigmoid2_ref(f2_mul(sum_f(1), k1), s_rcp);
              float2 hh = f2_mul(sum_f(2), k1);
              const float2 hr = f2_mul(hh, r);
              hh = tanh2_ref(padd(hr, gh[h]), s_rcp);                       // = hr + gh
              const float2 a_old = a_reg[u][mt][h];
              // z*state + (1-z)*h, summed the same way
              const float2 za = f2_mul(z, a_old), zh = f2_mul(f2_add(f2_set(1.f), make_float2(-z.x, -z.y)), hh);
              const float2 a_nw = padd(zh, za);                             // = za + zh
              unsigned short *dst = reinterpret_cast<unsigned short *>(xa_new) + (((sA + 8 * h) * kXaStride * 4 + ug * 8 + 2 * t) >> 1);
              uint32_t pv;
              if (ixw[mt][h] >> 24) {
                a_reg[u][mt][h] = a_nw;
                pv = quant_u8_unit(a_nw.x) | (quant_u8_unit(a_nw.y) << 8);
              } else {
                // an inactive stream keeps its state and the byte pair it already has
                pv = reinterpret_cast<const unsigned short *>(xa_cur)[((sA + 8 * h) * kXaStride * 4 + ug * 8 + 2 * t) >> 1];
              }
              *dst = (unsigned short)pv;
              if (u / US == 0) qa0[u % US][mt][h] = pv;
              else if (u / US == 1) qa0[u % US][mt][h] |= pv << 16;
              else qa2[u % US][mt][h] = pv;
            }
          }
        }
        // GRU-B input product (src/nnet.c:353) of this warp's 24 units, operands from registers; the 16 warps' partial
        // sums meet in shared memory
        {
          // fragment set w + WG*set of engine.h kBinFragWords = unit groups {w, w+16, w+32} + 8*set
          const uint2 *bf = reinterpret_cast<const uint2 *>(s_binF) + warp * kRGroupsB * 32 + lane;
          constexpr int NJ = kRGroupsB * US;              // (row group, K slice) products of this warp
          constexpr int PF = NS == 8 ? 6 : 2;             // B fragments fetched ahead (registers are short above 8 streams)
          auto frag = [&](int j) { return bf[((j % US) * WG * kRGroupsB + j / US) * 32]; };
          uint2 bw[PF];
#pragma unroll
          for (int k = 0; k < PF; k++) bw[k] = frag(k);
#pragma unroll
          for (int rg = 0; rg < kRGroupsB; rg++) {
            int32_t c[MT][4];
#pragma unroll
            for (int mt = 0; mt < MT; mt++) c[mt][0] = c[mt][1] = c[mt][2] = c[mt][3] = 0;
#pragma unroll
            for (int set = 0; set < US; set++) {
              const int j = rg * US + set;
              const uint2 bcur = bw[j % PF];
              if (j + PF < NJ) bw[j % PF] = frag(j + PF);
#pragma unroll
              for (int mt = 0; mt < MT; mt++)
                mma_u8s8(c[mt], qa0[set][mt][0], NV == 4 ? qa0[set][mt][NV / 2 - 1] : 0u, qa2[set][mt][0],
                         NV == 4 ? qa2[set][mt][NV / 2 - 1] : 0u, bcur.x, bcur.y);
            }
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
              int32_t *dst = s_part + ((mt * 2 + 0) * kRGroupsB + rg) * 128;
              atomicAdd(dst + g * 8 + 2 * t, c[mt][0]);
              atomicAdd(dst + g * 8 + 2 * t + 1, c[mt][1]);
              if (VR == 16) {
                atomicAdd(dst + (g + 8) * 8 + 2 * t, c[mt][2]);
                atomicAdd(dst + (g + 8) * 8 + 2 * t + 1, c[mt][3]);
              }
            }
          }
        }
        LPCB_MARK(1);   // phase A (this warp)
        gsync();
        LPCB_MARK(2);   // wait for the other warps
        // ================= phase B2 + T: GRU-B gates, sampling, signal update =================
        {
          if (twarp) {
            // pairs of streams share a warp; NS is even so both halves are valid together
            const int s = grp;
            if (TST) { t_x = s_tst[tid]; t_lpc = s_tst[TG + tid]; t_pred = s_tst[2 * TG + tid]; }
            const int mt = s / VR, row = s % VR;
            const int act = s_act[s];
            // the step's two kiss99 draws, by every lane of the stream from broadcast loads (no branch, no shuffle;
            // the integer work fills the latency of the loads below)
            uint32_t r0, r1;
            {
              uint32_t zz = s_rng[s], ww = s_rng[NS + s], js = s_rng[2 * NS + s], jc = s_rng[3 * NS + s];
              r0 = kiss99_next(zz, ww, js, jc);
              r1 = kiss99_next(zz, ww, js, jc);
              __syncwarp();
              if (l16 == 0 && act) { s_rng[s] = zz; s_rng[NS + s] = ww; s_rng[2 * NS + s] = js; s_rng[3 * NS + s] = jc; }
            }
            // compute_gruB (src/nnet.c:326-372)
            float zrh[3], rec[3];
#pragma unroll
            for (int gate = 0; gate < 3; gate++) {
              const int i = gate * kNB + l16;
              const int rg = gate * 2 + (l16 >> 3);
              const int e = row * 8 + (l16 & 7);
              int32_t ai = RA ? acc_seed_small(s_subB[i] + s_condb[s * kRowsB + i]) : acc_seed(s_subB[i] + s_condb[s * kRowsB + i]);
              int32_t *pin = s_part + ((mt * 2 + 0) * kRGroupsB + rg) * 128 + e;
              ai += *pin;
              *pin = 0;                                   // the next step's partial sums add up from zero
              if (SAT && M.b_in.sat.n) {
                // pairs on the exact s16 pass: this lane's row is row l16 % 8 of row group rg
                const unsigned char *xr = reinterpret_cast<const unsigned char *>(xa_new + s * kXaStride);
                const int e1 = (int)__ldg(M.b_in.sat.start + rg + 1);
                for (int q = (int)__ldg(M.b_in.sat.start + rg); q < e1; q++) {
                  const uint2 f = __ldg(M.b_in.sat.e + q);
                  if (((f.x >> 16) & 7) == (uint32_t)(l16 & 7)) ai += sat16_pair(f, xr);
                }
              }
              zrh[gate] = RA ? acc_out_small(ai) : acc_out(ai);
              const int32_t ar = (RA ? acc_seed_small(s_subB[kRowsB + i]) : acc_seed(s_subB[kRowsB + i])) +
                                 s_part[((mt * 2 + 1) * kRGroupsB + rg) * 128 + e];
              rec[gate] = RA ? acc_out_small(ar) : acc_out(ar);
            }
            const float z = sigmoid_ref(zrh[0] + rec[0], s_rcp);
            const float r = sigmoid_ref(zrh[1] + rec[1], s_rcp);
            const float hh = tanh_ref(zrh[2] + rec[2] * r, s_rcp);
            const float b_old = s_bst[s * 20 + l16];
            const float b_new = z * b_old + (1.f - z) * hh;
            __syncwarp();
            if (act) {
              s_bst[s * 20 + l16] = b_new;
              reinterpret_cast<unsigned char *>(s_xb)[s * kXbStride * 4 + l16] = (unsigned char)quant_u8_fast(b_new);
            }
            __syncwarp();
            LPCB_MARK(5);   // GRU-B gates
            // the product warps start now: holding their shared-memory burst back until the GRU-B gates are through
            // keeps it off the sampler's most latency-bound stretch (measured 3.67 -> 3.63 us per step; releasing
            // after the first sampler round gives the same, after the second 3.81)
            if (OVL) asm volatile("bar.arrive 1, %0;" ::"n"(kThreads) : "memory");
            // sample_mdense (src/nnet.c:163-214)
            float bv[16];
#pragma unroll
            for (int k = 0; k < 4; k++) {
              const float4 v = *reinterpret_cast<const float4 *>(s_bst + s * 20 + 4 * k);
              bv[4 * k] = v.x; bv[4 * k + 1] = v.y; bv[4 * k + 2] = v.z; bv[4 * k + 3] = v.w;
            }
            const int m = l16 ? l16 : 1;              // heap index within a depth-4 subtree (lane 0 idles on node 1)
            const int d = 31 - __clz(m);
            // round 1: levels 0..3, nodes 1..15
            int bit = s_logit[(r0 >> (8 * d)) & 0xffu] < node_logit<FCS>(fc_w, fc_stride, fc_bf, m, bv, s_rcp);
            int n = 1;
#pragma unroll
            for (int lvl = 0; lvl < 4; lvl++) n = 2 * n + __shfl_sync(0xffffffffu, bit, gbase + n);
            LPCB_MARK(6);   // sampler round 1
            // round 2: levels 4..7 below node n
            const int node = (n << d) | (m - (1 << d));
            bit = s_logit[(r1 >> (8 * d)) & 0xffu] < node_logit<FCS>(fc_w, fc_stride, fc_bf, node, bv, s_rcp);
            int mm = 1;
#pragma unroll
            for (int lvl = 0; lvl < 4; lvl++) mm = 2 * mm + __shfl_sync(0xffffffffu, bit, gbase + mm);
            int exc = ((n - 16) << 4) | (mm - 16);
            if (DEBUG && A.logits && stream0 + s < Bend) {
              float *lg = A.logits + ((size_t)(stream0 + s) * row_len + out_base + step) * 256;
              for (int k = 0; k < 16; k++) {
                const int nd = k * 16 + l16;
                lg[nd] = (nd == 0 || !act) ? 0.f : node_logit<FCS>(fc_w, fc_stride, fc_bf, nd, bv, s_rcp);
              }
            }
            LPCB_MARK(7);   // sampler round 2
            // signal update (src/lpcnet.c:256-269): every lane of the stream follows the scalar recurrence (same
            // values in all 16 lanes), the history shifts by one lane, lane 0 stores.  Shuffles stay outside the
            // per-stream conditions: the two streams of a warp may differ in `act`.
            {
              const bool valid = stream0 + s < Bend;
              const size_t oidx = (size_t)min(stream0 + s, Bend - 1) * row_len + out_base + step;
              const float dm = s_deemph[s];
              float pcm;
              int forced_v = 0;
              const bool teach = A.forced && (!A.forced_by_flag || (act & 2));
              if (teach) {
                forced_v = A.forced[oidx];
                const float fo = (float)forced_v;
                exc = lin2ulaw_ref(fo - 0.85f * dm - t_pred);
                pcm = fo - 0.85f * dm;
              } else {
                pcm = t_pred + s_u2l[exc];
              }
              const float x_up = __shfl_up_sync(0xffffffffu, t_x, 1, 16);
              if (act) t_x = l16 ? x_up : pcm;
              float out = pcm + 0.85f * dm;
              const float dm_new = out;
              if (out < -32767.f) out = -32767.f;
              if (out > 32767.f) out = 32767.f;
              const int outv = teach ? forced_v : round_half_up_ref(out);
              if (l16 == 0 && valid) {
                if (act) {
                  s_lexc[s] = exc;
                  s_deemph[s] = dm_new;
                  A.pcm[oidx] = (short)outv;
                } else {
                  A.pcm[oidx] = 0;
                }
                if (DEBUG && A.trace) {
                  LPCNetB200Trace tr = {0, 0, 0, 0, 0, 0, 0, 0};
                  if (act) {
                    tr.exc = exc; tr.u_sig = s_usig[s]; tr.u_pred = s_upred[s];
                    tr.rng_z = s_rng[s]; tr.rng_w = s_rng[NS + s];
                    tr.xa_hash = fnv1a_bytes(reinterpret_cast<const unsigned char *>(xa_new + s * kXaStride), kN);
                    tr.xb_hash = fnv1a_bytes(reinterpret_cast<const unsigned char *>(s_xb + s * kXbStride), kNB);
                    tr.pcm = outv;
                  }
                  reinterpret_cast<LPCNetB200Trace *>(A.trace)[oidx] = tr;
                }
              }
              __syncwarp();
              if (step + 1 < A.spf) predict();
              if (TST) { s_tst[tid] = t_x; s_tst[2 * TG + tid] = t_pred; }
            }
          } else if (OVL) {
            asm volatile("bar.sync 1, %0;" ::"n"(kThreads) : "memory");   // released by the sampler warps after the GRU-B gates
            product_ahead(xa_new);   // the next step's recurrent sums, while warps 0-3 sample
          }
        }
        LPCB_MARK(8);   // signal update
        gsync();
        LPCB_MARK(9);
        cur ^= 1;
      }  // step
    }    // frame

    // ---- store the tile's stream state ----
    // Only of the streams the launch moved: a stream that stayed inactive may be advanced by another launch on
    // another CUDA stream at the same time (the PLC's side chain, plc.cu), whose stores these would overwrite.
    if (tid < NS) s_act[tid] = ever_act;
    gsync();
#pragma unroll
    for (int u = 0; u < UW; u++) {
      const int ug = warp + u * WG;
#pragma unroll
      for (int mt = 0; mt < MT; mt++)
#pragma unroll
        for (int h = 0; h < NV / 2; h++) {
          const int s = mt * VR + g + h * 8;
          if (stream0 + s < Bend && s_act[s]) *reinterpret_cast<float2 *>(A.st.gru_a + (size_t)(stream0 + s) * kN + ug * 8 + 2 * t) = a_reg[u][mt][h];
        }
    }
    for (int i = tid; i < NS * 16; i += TG) {
      const int s = i >> 4, j = i & 15;
      if (stream0 + s < Bend && s_act[s]) {
        A.st.gru_b[(size_t)(stream0 + s) * kNB + j] = s_bst[s * 20 + j];
      }
    }
    if (grp < NS && stream0 + grp < Bend && s_act[grp]) A.st.last_sig[(size_t)(stream0 + grp) * 16 + l16] = t_x;
    for (int s = tid; s < NS; s += TG) {
      if (stream0 + s < Bend && s_act[s]) {
        A.st.deemph[stream0 + s] = s_deemph[s];
        A.st.last_exc[stream0 + s] = s_lexc[s];
#pragma unroll
        for (int k = 0; k < 4; k++) A.st.rng[(size_t)(stream0 + s) * 4 + k] = s_rng[k * NS + s];
      }
    }
  }  // tile
#ifdef LPCB_PHASE_CLOCKS
  if (blockIdx.x == 0 && threadIdx.x == 0 && A.phase_clk)
    for (int k = 0; k < 10; k++) atomicAdd(A.phase_clk + k, (unsigned long long)clk_[k]);
#endif
}

template <int MT, int VR>
static cudaError_t launch_tile(const SampleArgs &a, int n_sms, cudaStream_t s) {
  const int NS = MT * VR;
  const size_t smem = make_layout(NS, MT, a.m.a_rec.n_chunks, a.m.a_rec.slots[2] != 0 && !a.m.has_sat, a.m.b_rec.n_chunks).total;
  const bool dbg = a.trace || a.logits;
  const bool sat = a.m.has_sat != 0;                // exact s16 pass: generic chunk lists, rcpps table (model.cu)
  const bool fix = a.m.a_rec.slots[2] != 0 && !sat;
  // Arithmetic rcpps when it was verified against the table on this GPU (model.cu): no shared-memory lookups,
  // 2-4 % faster at every tile size (measured); the same variants seed and read the int32 sums without conversion
  // instructions, which needs the model's sums to stay inside (-2^22, 2^22) (DevModel::small_acc).  The trace/logit
  // variants always use the table and the conversions.
  const bool ra = a.m.rcp_arith != 0 && a.m.small_acc != 0 && !sat;
  const bool sub = a.stream_begin != 0 || a.stream_end != 0;
  if (sub && dbg) return cudaErrorInvalidValue;     // the trace / logit builds advance whole batches only
  void (*kern)(const SampleArgs) =
      sat ? (dbg ? sample_kernel<MT, VR, true, false, false, false, true>
                 : sub ? sample_kernel<MT, VR, false, false, false, true, true> : sample_kernel<MT, VR, false, false, false, false, true>)
      : dbg ? (fix ? sample_kernel<MT, VR, true, true, false> : sample_kernel<MT, VR, true, false, false>)
      : sub ? (ra ? (fix ? sample_kernel<MT, VR, false, true, true, true> : sample_kernel<MT, VR, false, false, true, true>)
                  : (fix ? sample_kernel<MT, VR, false, true, false, true> : sample_kernel<MT, VR, false, false, false, true>))
      : ra ? (fix ? sample_kernel<MT, VR, false, true, true> : sample_kernel<MT, VR, false, false, true>)
           : (fix ? sample_kernel<MT, VR, false, true, false> : sample_kernel<MT, VR, false, false, false>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int n_tiles = ((a.stream_end ? a.stream_end : a.n_streams) - a.stream_begin + NS - 1) / NS;
  const int grid = n_tiles < n_sms ? n_tiles : n_sms;
  kern<<<grid, kThreads, smem, s>>>(a);
  return cudaGetLastError();
}

// Two 8-stream tiles per CTA (G = 2): production kernel only (no trace / logits, no s16 pass).
static cudaError_t launch_tile_pairs(const SampleArgs &a, int n_sms, cudaStream_t s) {
  const size_t smem = make_layout(8, 1, a.m.a_rec.n_chunks, a.m.a_rec.slots[2] != 0, a.m.b_rec.n_chunks, 2).total;
  if (a.trace || a.logits || a.m.has_sat) return cudaErrorInvalidValue;
  const bool fix = a.m.a_rec.slots[2] != 0, ra = a.m.rcp_arith != 0 && a.m.small_acc != 0;
  const bool sub = a.stream_begin != 0 || a.stream_end != 0;
  void (*kern)(const SampleArgs) =
      sub ? (ra ? (fix ? sample_kernel<1, 8, false, true, true, true, false, 2> : sample_kernel<1, 8, false, false, true, true, false, 2>)
                : (fix ? sample_kernel<1, 8, false, true, false, true, false, 2> : sample_kernel<1, 8, false, false, false, true, false, 2>))
      : ra ? (fix ? sample_kernel<1, 8, false, true, true, false, false, 2> : sample_kernel<1, 8, false, false, true, false, false, 2>)
           : (fix ? sample_kernel<1, 8, false, true, false, false, false, 2> : sample_kernel<1, 8, false, false, false, false, false, 2>);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  const int n_tiles = ((a.stream_end ? a.stream_end : a.n_streams) - a.stream_begin + 7) / 8;
  const int ctas = (n_tiles + 1) / 2, grid = ctas < n_sms ? ctas : n_sms;
  kern<<<grid, kThreads, smem, s>>>(a);
  return cudaGetLastError();
}

}  // namespace lpcb
