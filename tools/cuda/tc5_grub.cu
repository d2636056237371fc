// Decision by measurement: GRU-B's dense recurrent input product
// [streams x 384 u8 state bytes] x [384 x 48 s8 weights] -> s32 as ONE tcgen05.mma kind::i8 chain with M = 128 streams
// per CTA, against the mma.sync m16n8k32 form the sample kernel uses (csrc/sample_kernel.cuh: 6 register-operand IMMA
// per warp and step inside phase A, partial sums added in shared memory).
//
// The product sits on the serial path of every sample step (its inputs are the step's new GRU-A state, its outputs
// feed GRU-B's gates and the sampler), so what matters is the ROUND TRIP of one step, not throughput:
//   tc5:  [state bytes -> shared memory in the UMMA canonical K-major layout] -> fence.proxy.async -> barrier ->
//         12 x tcgen05.mma (K = 32 each) by one thread -> tcgen05.commit -> mbarrier wait -> tcgen05.ld 48 columns
//   imma: [state bytes in shared memory, A-fragment order] -> barrier -> 16 warps x 6 IMMA on a 24-wide K slice ->
//         red.shared of the partial sums -> barrier -> read 48 sums per stream
// Both variants are checked against a host computation of the first step, then timed as a dependent chain (each
// step's state bytes are a function of the previous step's sums).  Prints cycles per step.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o tc5_grub tools/cuda/tc5_grub.cu && ./tc5_grub
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

constexpr int K = 384, N = 48, M = 128;

// the next step's state, four bytes at a time, as a function of this step's sums (a stand-in for the gate arithmetic)
__host__ __device__ inline uint32_t next_word(int32_t sum, int i) { return ((uint32_t)sum >> 3) * 2654435761u + 40503u * (uint32_t)i; }

// ---------------------------------------------------------------- tcgen05 ------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

// K-major, no swizzle (cute make_umma_desc<Major::K>, LayoutType::INTERLEAVE): core matrix = 8 rows x 16 bytes,
// contiguous (128 B); SBO = byte step between 8-row groups, LBO = byte step between the 16-byte K chunks.
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
// instruction descriptor (cute UMMA::InstrDescriptor): c = s32, a = u8, b = s8, both K-major, N = 48, M = 128
constexpr uint32_t kIdesc = (2u << 4) | (0u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);

__device__ __forceinline__ void umma_i8(uint32_t tmem_c, uint64_t da, uint64_t db, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}\n" ::"r"(tmem_c),
      "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate), "r"(0u)
      : "memory");
}

template <bool RESTAGE>
__global__ void __launch_bounds__(128, 1) tc5_kernel(const uint8_t *state0, const int8_t *w, int32_t *first_out, long long *cycles,
                                                    uint32_t *chk, int steps) {
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t *sA = smem;                  // [24 k-chunks][16 row groups][8 rows][16 B]
  uint8_t *sB = smem + M * K;          // [24 k-chunks][ 6 row groups][8 rows][16 B]
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;
  // operands into the canonical layout
  for (int i = tid; i < M * K / 16; i += 128) {
    const int m = i / (K / 16), kc = i % (K / 16);
    *reinterpret_cast<uint4 *>(sA + ((kc * (M / 8) + m / 8) * 8 + (m % 8)) * 16) = *reinterpret_cast<const uint4 *>(state0 + (size_t)m * K + kc * 16);
  }
  for (int i = tid; i < N * K / 16; i += 128) {
    const int n = i / (K / 16), kc = i % (K / 16);
    *reinterpret_cast<uint4 *>(sB + ((kc * (N / 8) + n / 8) * 8 + (n % 8)) * 16) = *reinterpret_cast<const uint4 *>(w + (size_t)n * K + kc * 16);
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base_s)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_base_s;
  const uint32_t a0 = smem_u32(sA), b0 = smem_u32(sB);
  constexpr uint32_t A_LBO = (M / 8) * 128, B_LBO = (N / 8) * 128, SBO = 128;
  uint32_t phase = 0, acc_chk = 0;
  int32_t r[N];
  long long t0 = 0;
  for (int s = 0; s <= steps; s++) {
    if (s == 1) t0 = clock64();
    if (tid == 0) {
#pragma unroll
      for (int j = 0; j < K / 32; j++)
        umma_i8(tmem, umma_desc(a0 + j * 2 * A_LBO, A_LBO, SBO), umma_desc(b0 + j * 2 * B_LBO, B_LBO, SBO), j ? 1u : 0u);
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
    }
    // wait for the MMAs (bounded: a wrong descriptor must not hang the box)
    {
      uint32_t done = 0;
      for (int spin = 0; spin < (1 << 22) && !done; spin++)
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(smem_u32(&bar)), "r"(phase) : "memory");
      if (!done) { if (tid == 0) chk[1] = 0xDEAD0000u + s; return; }
      phase ^= 1;
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll
    for (int c = 0; c < N; c += 16)
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                   : "=r"(r[c + 0]), "=r"(r[c + 1]), "=r"(r[c + 2]), "=r"(r[c + 3]), "=r"(r[c + 4]), "=r"(r[c + 5]), "=r"(r[c + 6]),
                     "=r"(r[c + 7]), "=r"(r[c + 8]), "=r"(r[c + 9]), "=r"(r[c + 10]), "=r"(r[c + 11]), "=r"(r[c + 12]),
                     "=r"(r[c + 13]), "=r"(r[c + 14]), "=r"(r[c + 15])
                   : "r"(taddr + c));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    if (s == 0)
      for (int n = 0; n < N; n++) first_out[tid * N + n] = r[n];
    // the next step's state: every byte of the row a function of this step's sums (stream tid = row tid)
    if (RESTAGE) {
#pragma unroll
      for (int kc = 0; kc < K / 16; kc++) {
        uint32_t v[4];
#pragma unroll
        for (int q = 0; q < 4; q++) v[q] = next_word(r[(kc * 4 + q) % N], kc * 4 + q);
        *reinterpret_cast<uint4 *>(sA + ((kc * (M / 8) + tid / 8) * 8 + (tid % 8)) * 16) = make_uint4(v[0], v[1], v[2], v[3]);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    } else {
      for (int n = 0; n < N; n++) acc_chk += (uint32_t)r[n];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  const long long t1 = clock64();
  for (int n = 0; n < N; n++) acc_chk += (uint32_t)r[n] * (n + 1);
  atomicAdd(&chk[0], acc_chk);
  if (tid == 0) cycles[0] = t1 - t0;
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}

// ---------------------------------------------------------------- mma.sync -----------------------------------------
// NS streams per CTA (8 or 16, as the sample kernel's tiles), 16 warps: warp w takes the K slice [24 w, 24 w + 24)
// -- padded to 32 with zero weights -- for all 48 outputs (6 n-tiles), and adds its partial sums into shared memory.
__device__ __forceinline__ void imma(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
               : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

template <int NS>
__global__ void __launch_bounds__(512, 1) imma_kernel(const uint8_t *state0, const int8_t *w, int32_t *first_out, long long *cycles,
                                                     uint32_t *chk, int steps) {
  __shared__ __align__(16) uint8_t sx[16][K + 16];       // state bytes, row = stream (rows >= NS are zero)
  __shared__ int32_t part[16][N + 1];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  for (int i = tid; i < 16 * (K + 16); i += 512) (&sx[0][0])[i] = 0;
  __syncthreads();
  for (int i = tid; i < NS * K; i += 512) sx[i / K][i % K] = state0[i];
  // this warp's B fragments: n-tile j, k = 24 warp + (4 t .. 4 t + 3) and + 16; zero beyond the 24-wide slice
  uint32_t bf[6][2];
  for (int j = 0; j < 6; j++)
    for (int h = 0; h < 2; h++) {
      uint32_t x = 0;
      for (int e = 0; e < 4; e++) {
        const int kk = h * 16 + 4 * t + e;
        const int8_t v = kk < 24 ? w[(size_t)(j * 8 + g) * K + 24 * warp + kk] : (int8_t)0;
        x |= (uint32_t)(uint8_t)v << (8 * e);
      }
      bf[j][h] = x;
    }
  for (int i = tid; i < 16 * (N + 1); i += 512) (&part[0][0])[i] = 0;
  __syncthreads();
  uint32_t acc_chk = 0;
  long long t0 = 0;
  for (int s = 0; s <= steps; s++) {
    if (s == 1) t0 = clock64();
    // A fragments of the slice: rows g and g + 8, k = 24 warp + 4 t (+16)
    const uint8_t *r0 = &sx[g][24 * warp], *r1 = &sx[g + 8][24 * warp];
    const uint32_t a0 = *reinterpret_cast<const uint32_t *>(r0 + 4 * t), a1 = *reinterpret_cast<const uint32_t *>(r1 + 4 * t);
    const uint32_t a2 = *reinterpret_cast<const uint32_t *>(r0 + 16 + 4 * t), a3 = *reinterpret_cast<const uint32_t *>(r1 + 16 + 4 * t);
#pragma unroll
    for (int j = 0; j < 6; j++) {
      int c[4] = {0, 0, 0, 0};
      imma(c, a0, a1, a2, a3, bf[j][0], bf[j][1]);
      atomicAdd(&part[g][j * 8 + 2 * t], c[0]);
      atomicAdd(&part[g][j * 8 + 2 * t + 1], c[1]);
      if (NS > 8) { atomicAdd(&part[g + 8][j * 8 + 2 * t], c[2]); atomicAdd(&part[g + 8][j * 8 + 2 * t + 1], c[3]); }
    }
    __syncthreads();
    // every stream's 48 sums -> its next state bytes (thread = (stream, 12-byte piece)), sums cleared for the next step
    if (s == 0) for (int i = tid; i < NS * N; i += 512) first_out[i] = part[i / N][i % N];
    if (tid < NS * 32) {
      const int st = tid >> 5, p = tid & 31;             // 32 pieces of 3 words per stream
#pragma unroll
      for (int e = 0; e < 3; e++) { const int i = p * 3 + e; *reinterpret_cast<uint32_t *>(&sx[st][4 * i]) = next_word(part[st][i % N], i); }
      acc_chk += (uint32_t)part[st][p % N];
    }
    __syncthreads();
    for (int i = tid; i < NS * N; i += 512) part[i / N][i % N] = 0;
    __syncthreads();
  }
  const long long t1 = clock64();
  atomicAdd(&chk[0], acc_chk);
  if (tid == 0) cycles[0] = t1 - t0;
}

// ---------------------------------------------------------------- host ---------------------------------------------
int main() {
  std::vector<uint8_t> st(M * K);
  std::vector<int8_t> w(N * K);
  uint32_t x = 12345;
  auto rnd = [&]() { x = x * 1664525u + 1013904223u; return x >> 8; };
  for (auto &v : st) v = (uint8_t)(rnd() & 255);
  for (auto &v : w) v = (int8_t)((int)(rnd() % 81) - 40);
  std::vector<int32_t> want(M * N);
  for (int m = 0; m < M; m++)
    for (int n = 0; n < N; n++) {
      int32_t a = 0;
      for (int k = 0; k < K; k++) a += (int32_t)st[m * K + k] * (int32_t)w[n * K + k];
      want[m * N + n] = a;
    }
  uint8_t *d_st; int8_t *d_w; int32_t *d_out; long long *d_cyc; uint32_t *d_chk;
  cudaMalloc(&d_st, st.size()); cudaMalloc(&d_w, w.size()); cudaMalloc(&d_out, M * N * 4); cudaMalloc(&d_cyc, 8); cudaMalloc(&d_chk, 8);
  cudaMemcpy(d_st, st.data(), st.size(), cudaMemcpyHostToDevice);
  cudaMemcpy(d_w, w.data(), w.size(), cudaMemcpyHostToDevice);
  const int steps = 2000;
  std::vector<int32_t> got(M * N);
  auto report = [&](const char *name, int rows) {
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("%s: CUDA error %s\n", name, cudaGetErrorString(e)); exit(1); }
    long long cyc; uint32_t chk[2];
    cudaMemcpy(&cyc, d_cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(chk, d_chk, 8, cudaMemcpyDeviceToHost);
    cudaMemcpy(got.data(), d_out, M * N * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int i = 0; i < rows * N; i++) bad += got[i] != want[i];
    printf("%-44s first step %s (%d of %d sums differ)  %8.1f cycles per step  [chk %08x%s]\n", name, bad ? "WRONG" : "exact", bad,
           rows * N, (double)cyc / steps, chk[0], (chk[1] >> 16) == 0xDEAD ? ", MBARRIER TIMEOUT" : "");
    cudaMemset(d_chk, 0, 8); cudaMemset(d_out, 0, M * N * 4);
  };
  cudaMemset(d_chk, 0, 8);
  const int smem = M * K + N * K + 1024;
  cudaFuncSetAttribute(tc5_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(tc5_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  for (int rep = 0; rep < 2; rep++) {
    tc5_kernel<false><<<1, 128, smem>>>(d_st, d_w, d_out, d_cyc, d_chk, steps);
    report("tcgen05 M=128: mma x12 + commit + wait + ld", M);
    tc5_kernel<true><<<1, 128, smem>>>(d_st, d_w, d_out, d_cyc, d_chk, steps);
    report("tcgen05 M=128: + state bytes restaged", M);
    imma_kernel<8><<<1, 512>>>(d_st, d_w, d_out, d_cyc, d_chk, steps);
    report("mma.sync, 8 streams, 16 warps x 6 IMMA + red", 8);
    imma_kernel<16><<<1, 512>>>(d_st, d_w, d_out, d_cyc, d_chk, steps);
    report("mma.sync, 16 streams, 16 warps x 6 IMMA + red", 16);
  }
  return 0;
}
