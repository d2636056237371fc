// The RDOVAE decoder of DRED: DRED_rdovae_decode_all = dred_rdovae_dec_init_states +
// dred_rdovae_decode_qframe per latent vector (src/dred_rdovae.c:37-51, src/dred_rdovae_dec.c:37-98).  Included by
// plc.cu inside its namespace (it shares the predictor's int8 GRU code: the decoder's GRUs have the same 256-unit
// geometry).  One block of 256 threads per stream; the latent vectors of a stream are sequential (three GRU states),
// streams are independent.  Dense layers accumulate with fmaf in input order from the bias (sgemv_accum16: every
// layer has a multiple of 16 rows), the GRUs are compute_gruB on u8 x s8 products.
#pragma once

struct DredArgs {
  DevDred net;
  const uint32_t *rcp_tab;
  const float *states;     // [B][24]
  const float *latents;    // [B][n][80]
  float *features;         // [B][n][80]  four feature vectors per latent, in the reference's (reverse) order
  int n_latents;
};

__device__ __forceinline__ void dred_dense_tanh(const float *__restrict__ w, const float *__restrict__ b, const float *x, int nin,
                                                float *out, const uint32_t *rcp, int t) {
  float acc = __ldg(b + t);
  for (int j = 0; j < nin; j++) acc = fmaf(__ldg(w + (size_t)j * kPlcUnits + t), x[j], acc);
  out[t] = tanh_ref(acc, rcp);
}

__global__ void __launch_bounds__(256) dred_decode_kernel(DredArgs A) {
  const int s = blockIdx.x, t = threadIdx.x;
  __shared__ float buf[8 * kPlcUnits];          // the eight layer outputs of a step, concatenated (dec_final's input)
  __shared__ float g[3][kPlcUnits];             // dense2 / dense4 / dense6 GRU states
  __shared__ float zin[kDredLatent];
  __shared__ uint32_t xq_in[kPlcUnits / 4], xq_st[kPlcUnits / 4];
  const DevDred &D = A.net;
  // dred_rdovae_dec_init_states: three dense layers of the 24-value initial state
  if (t < kDredState) zin[t] = A.states[(size_t)s * kDredState + t];
  __syncthreads();
  for (int k = 0; k < 3; k++) dred_dense_tanh(D.state_w[k], D.state_b[k], zin, kDredState, g[k], A.rcp_tab, t);
  __syncthreads();
  for (int i = 0; i < A.n_latents; i++) {
    if (t < kDredLatent) zin[t] = A.latents[((size_t)s * A.n_latents + i) * kDredLatent + t];
    __syncthreads();
    dred_dense_tanh(D.dense_w[0], D.dense_b[0], zin, kDredLatent, buf, A.rcp_tab, t);
    __syncthreads();
    for (int k = 0; k < 3; k++) {
      // compute_gruB (zero conditioning) on the previous dense output; the new state is the next layer's input
      plc_gru(D.g_in[k], D.g_rec[k], D.g_subias[k], buf + 2 * k * kPlcUnits, kPlcUnits, g[k], xq_in, xq_st,
              buf + (2 * k + 1) * kPlcUnits, A.rcp_tab, t, D.sat_in[k], D.sat_rec[k]);
      dred_dense_tanh(D.dense_w[k + 1], D.dense_b[k + 1], buf + (2 * k + 1) * kPlcUnits, kPlcUnits,
                      buf + (2 * k + 2) * kPlcUnits, A.rcp_tab, t);
      __syncthreads();
    }
    dred_dense_tanh(D.dense_w[4], D.dense_b[4], buf + 6 * kPlcUnits, kPlcUnits, buf + 7 * kPlcUnits, A.rcp_tab, t);
    __syncthreads();
    if (t < kDredOut) {                          // dec_final: 2048 -> 80, linear
      float acc = __ldg(D.final_b + t);
      for (int j = 0; j < 8 * kPlcUnits; j++) acc = fmaf(__ldg(D.final_w + (size_t)j * kDredOut + t), buf[j], acc);
      A.features[((size_t)s * A.n_latents + i) * kDredOut + t] = acc;
    }
    __syncthreads();
  }
}
