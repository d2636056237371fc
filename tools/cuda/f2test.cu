// GPU check: packed f32x2 arithmetic and the pair activations equal their scalar forms bit for bit.
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "../../lpcnet_b200/csrc/bitexact.cuh"
#include "../../lpcnet_b200/csrc/engine.h"
using namespace lpcb;
__global__ void k(const float *a, const float *b, const float *c, const uint32_t *tab, uint32_t *out, int n) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float2 A = make_float2(a[i], a[(i * 7 + 3) % n]), B = make_float2(b[i], b[(i * 5 + 1) % n]), C = make_float2(c[i], c[(i * 3 + 2) % n]);
  uint32_t bad = 0;
  float2 r = f2_fma(A, B, C);
  if (__float_as_uint(r.x) != __float_as_uint(fmaf(A.x, B.x, C.x)) || __float_as_uint(r.y) != __float_as_uint(fmaf(A.y, B.y, C.y))) bad |= 1;
  r = f2_add(A, B);
  if (__float_as_uint(r.x) != __float_as_uint(A.x + B.x) || __float_as_uint(r.y) != __float_as_uint(A.y + B.y)) bad |= 2;
  r = f2_mul(A, B);
  if (__float_as_uint(r.x) != __float_as_uint(A.x * B.x) || __float_as_uint(r.y) != __float_as_uint(A.y * B.y)) bad |= 4;
  r = sigmoid2_ref(A, tab);
  if (__float_as_uint(r.x) != __float_as_uint(sigmoid_ref(A.x, tab)) || __float_as_uint(r.y) != __float_as_uint(sigmoid_ref(A.y, tab))) bad |= 8;
  r = tanh2_ref(A, tab);
  if (__float_as_uint(r.x) != __float_as_uint(tanh_ref(A.x, tab)) || __float_as_uint(r.y) != __float_as_uint(tanh_ref(A.y, tab))) bad |= 16;
  r = f2_fma(A, f2_set(-1.f), f2_set(1.f));
  if (__float_as_uint(r.x) != __float_as_uint(1.f - A.x) || __float_as_uint(r.y) != __float_as_uint(1.f - A.y)) bad |= 32;
  out[i] = bad;
}
int main() {
  const int n = 1 << 20;
  std::vector<float> a(n), b(n), c(n);
  srand(1);
  auto rnd = [](int mode) {
    float u = (float)rand() / RAND_MAX * 2 - 1;
    switch (mode) { case 0: return u; case 1: return u * 20.f; case 2: return u * 1e-20f; case 3: return u * 1e-38f; case 4: return u * 1e-42f; default: return u * 1e4f; }
  };
  for (int i = 0; i < n; i++) { a[i] = rnd(i % 6); b[i] = rnd((i / 6) % 6); c[i] = rnd((i / 36) % 6); }
  a[0] = 0.f; a[1] = -0.f; b[0] = -0.f; c[0] = 0.f; c[1] = -0.f;
  uint32_t tab[2048]; default_rcp_table(tab);
  float *da, *db, *dc; uint32_t *dt, *dout;
  cudaMalloc(&da, n * 4); cudaMalloc(&db, n * 4); cudaMalloc(&dc, n * 4); cudaMalloc(&dt, 8192); cudaMalloc(&dout, n * 4);
  cudaMemcpy(da, a.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(db, b.data(), n * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dc, c.data(), n * 4, cudaMemcpyHostToDevice); cudaMemcpy(dt, tab, 8192, cudaMemcpyHostToDevice);
  k<<<n / 256, 256>>>(da, db, dc, dt, dout, n);
  std::vector<uint32_t> out(n);
  if (cudaMemcpy(out.data(), dout, n * 4, cudaMemcpyDeviceToHost) != cudaSuccess) { printf("cuda error\n"); return 1; }
  int cnt[6] = {0};
  for (int i = 0; i < n; i++) for (int k2 = 0; k2 < 6; k2++) if (out[i] >> k2 & 1) { if (cnt[k2]++ < 3) printf("bad op %d at %d: a=%g b=%g c=%g\n", k2, i, a[i], b[i], c[i]); }
  printf("mismatches fma %d add %d mul %d sigmoid %d tanh %d oneminus %d\n", cnt[0], cnt[1], cnt[2], cnt[3], cnt[4], cnt[5]);
  return 0;
}
