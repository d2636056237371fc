// Device side of the causal and codec PLC modes (lpcnet_plc_update_causal / lpcnet_plc_conceal_causal,
// src/lpcnet_plc.c:185-340, with get_fec_or_pred 147-166 and the deferred frame network, src/lpcnet.c:122-144).
// Included by plc.cu inside its namespace.  Every integer of the reference's control state (pcm_fill, skip_analysis,
// blend, loss_count, the FEC queue positions, feature_buffer_fill) is a function of the loss pattern and of the
// lpcnet_plc_fec_add calls, so the host plans each frame (plc_plan_causal in plc.cu) and these kernels only move
// data for the streams a control row selects.
#pragma once

struct CausalState {       // per stream, device
  short *cpcm;             // [B][kCpcm]  st->pcm (PLC_BUF_SIZE + FRAME_SIZE samples)
  float *pcopy;            // [3][B][512] plc_copy[k] = (gru1 | gru2)
  float *fbuf;             // [B][4][20]  LPCNetState::feature_buffer
  float *fecst;            // [4][B][20]  the FEC feature vectors this frame's predictions consume (staged by the host)
};
constexpr int kCpcm = kMaxDelay * kFrame + kTOff + kFrame;   // 560

// RNN_MOVE(&plc_copy[1], &plc_copy[0], FEATURES_DELAY); plc_copy[0] = plc_net   (lpcnet_plc.c:305-306, 313-314)
__global__ void c_push_copy_kernel(PlcState st, CausalState cs, const uint8_t *mask, int B, int D) {
  const int s = blockIdx.x, t = threadIdx.x;
  if (!mask[s]) return;
  for (int i = t; i < 2 * kPlcUnits; i += blockDim.x) {
    for (int k = D; k >= 1; k--) cs.pcopy[((size_t)k * B + s) * 512 + i] = cs.pcopy[((size_t)(k - 1) * B + s) * 512 + i];
    cs.pcopy[(size_t)s * 512 + i] = i < kPlcUnits ? st.gru1[(size_t)s * kPlcUnits + i] : st.gru2[(size_t)s * kPlcUnits + i - kPlcUnits];
  }
}

// st->plc_net = st->plc_copy[k]   (lpcnet_plc.c:218, 236)
__global__ void c_set_net_kernel(PlcState st, CausalState cs, const uint8_t *mask, int B, int k) {
  const int s = blockIdx.x, t = threadIdx.x;
  if (!mask[s]) return;
  for (int i = t; i < 2 * kPlcUnits; i += blockDim.x) {
    const float v = cs.pcopy[((size_t)k * B + s) * 512 + i];
    if (i < kPlcUnits) st.gru1[(size_t)s * kPlcUnits + i] = v; else st.gru2[(size_t)s * kPlcUnits + i - kPlcUnits] = v;
  }
}

// run_frame_network_deferred (lpcnet.c:122-133): code 0 = no push, else (slot + 1) | 8 if the buffer shifts first
__global__ void c_fb_push_kernel(CausalState cs, const uint8_t *code, const float *src, int src_stride, int B) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= B * 20) return;
  const int s = i / 20, k = i % 20, c = code[s];
  if (!c) return;
  float *fb = cs.fbuf + (size_t)s * 80;
  if (c & 8) { fb[k] = fb[20 + k]; fb[20 + k] = fb[40 + k]; fb[40 + k] = fb[60 + k]; }
  fb[((c & 7) - 1) * 20 + k] = src[(size_t)s * src_stride + k];
}

// st->pcm moves.  op per stream: 0 none; 1 dst[off..off+n) = src (append / tail copy); 2 shift left by 160 keeping
// `keep` samples (RNN_MOVE(st->pcm, &st->pcm[FRAME_SIZE], PLC_BUF_SIZE)); one block per stream
__global__ void c_pcm_store_kernel(CausalState cs, const uint8_t *mask, const uint16_t *off, int fixed_off, const short *src,
                                   int src_stride, int src_off, int n) {
  const int s = blockIdx.x;
  if (!mask[s]) return;
  const int o = off ? off[s] : fixed_off;
  for (int i = threadIdx.x; i < n; i += blockDim.x) cs.cpcm[(size_t)s * kCpcm + o + i] = src[(size_t)s * src_stride + src_off + i];
}
__global__ void c_pcm_shift_kernel(CausalState cs, const uint8_t *mask, int keep) {
  const int s = blockIdx.x;
  if (!mask[s]) return;
  __shared__ short tmp[kCpcm];
  for (int i = threadIdx.x; i < keep; i += blockDim.x) tmp[i] = cs.cpcm[(size_t)s * kCpcm + kFrame + i];
  __syncthreads();
  for (int i = threadIdx.x; i < keep; i += blockDim.x) cs.cpcm[(size_t)s * kCpcm + i] = tmp[i];
}
// teacher-forcing input of a catch-up segment: segf[s][0..n) = st->pcm[0..n)
__global__ void c_pcm_load_kernel(CausalState cs, const uint8_t *mask, short *dst, int n) {
  const int s = blockIdx.x;
  if (!mask[s]) return;
  for (int i = threadIdx.x; i < n; i += blockDim.x) dst[(size_t)s * n + i] = cs.cpcm[(size_t)s * kCpcm + i];
}

// compute_plc_pred for the causal modes.  mode per stream: 0 skip; 1 zeros -> features; 2 [burg | 0 | 1] ->
// features; 3 [burg | analysis features | 1] -> features; 4 FEC: input [0 | fec | -1], the output is discarded and
// features = fec (get_fec_or_pred, lpcnet_plc.c:147-166).  +8: then the loss attenuation of features[0] with the
// loss count in lossc (lpcnet_plc.c:318-319).
__global__ void __launch_bounds__(256) c_pred_kernel(PlcArgs A, const uint8_t *mode_row, const float *fec) {
  const int s = blockIdx.x, t = threadIdx.x;
  const int code = mode_row[s], mode = code & 7;
  if (!mode) return;
  __shared__ float s_in[kPlcIn + 3], s_d1[kPlcCond], s_g1[kPlcUnits], s_g2[kPlcUnits];
  __shared__ uint32_t xq_in[kPlcUnits / 4], xq_st[kPlcUnits / 4];
  if (t < kPlcIn) {
    float v = 0.f;
    if (mode == 2 || mode == 3) {
      if (t < 2 * kBands) v = A.w.burg[(size_t)s * 2 * kBands + t];
      else if (mode == 3 && t < 2 * kBands + 20) v = A.w.encf[(size_t)s * 36 + t - 2 * kBands];
      else if (t == kPlcIn - 1) v = 1.f;
    } else if (mode == 4) {
      if (t >= 2 * kBands && t < 2 * kBands + 20) v = fec[(size_t)s * 20 + t - 2 * kBands];
      else if (t == kPlcIn - 1) v = -1.f;
    }
    s_in[t] = v;
  }
  __syncthreads();
  if (t < kPlcCond) {
    float acc = __ldg(A.net.dense1_b + t);
    for (int j = 0; j < kPlcIn; j++) acc = fmaf(__ldg(A.net.dense1_w + j * kPlcCond + t), s_in[j], acc);
    s_d1[t] = tanh_ref(acc, A.rcp_tab);
  }
  __syncthreads();
  plc_gru(A.net.g1_in, A.net.g1_rec, A.net.g1_subias, s_d1, kPlcCond, A.st.gru1 + (size_t)s * kPlcUnits, xq_in, xq_st,
          s_g1, A.rcp_tab, t, A.net.sat_g1_in, A.net.sat_g1_rec);
  plc_gru(A.net.g2_in, A.net.g2_rec, A.net.g2_subias, s_g1, kPlcUnits, A.st.gru2 + (size_t)s * kPlcUnits, xq_in, xq_st,
          s_g2, A.rcp_tab, t, A.net.sat_g2_in, A.net.sat_g2_rec);
  if (t < 20) {
    float acc;
    if (mode == 4) {
      acc = fec[(size_t)s * 20 + t];
    } else {
      acc = __ldg(A.net.out_b + t);
      for (int j = 0; j < kPlcUnits; j++) acc += __ldg(A.net.out_w + j * 20 + t) * s_g2[j];
      if (t == 19) acc = fminf(.5f, acc + .1f);
    }
    if (t == 0 && (code & 8)) {
      const int lc = A.lossc[s];
      const float att[10] = {0, 0, -.2f, -.2f, -.4f, -.4f, -.8f, -.8f, -1.6f, -1.6f};
      float v;
      if (lc >= 10) v = acc + att[9] - (float)(2 * (lc - 9));
      else v = acc + att[lc];
      acc = -10.f > v ? -10.f : v;
    }
    A.st.features[(size_t)s * 20 + t] = acc;
  }
}

// fnfeat[s] = st->features for the streams of a mask
__global__ void c_feat_kernel(PlcArgs A, const uint8_t *mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_streams * 20) return;
  if (mask[i / 20]) A.w.fnfeat[i] = A.st.features[i];
}

// update streams, causal DC removal (lpcnet_plc.c:192-204): the frame -> w.in, the removed offset -> w.lp, delta
__global__ void __launch_bounds__(kWpsThreads) c_pre_kernel(PlcArgs A, const uint8_t *mask) {
  __shared__ short xs[kWpsThreads / 32][kFrame], ls[kWpsThreads / 32][kFrame];
  int lane, w;
  const int s = wps_stream(lane, w);                 // one warp per stream, see plc_pre_kernel
  if (s >= A.n_streams || !mask[s]) return;
  const short *x = A.pcm_io + ((size_t)s * A.total_frames + A.frame) * kFrame;
  short *in = A.w.in + (size_t)s * kFrame, *lp = A.w.lp + (size_t)s * kFrame;
  for (int i = lane; i < kFrame; i += 32) xs[w][i] = x[i];
  __syncwarp();
  if (lane == 0) {
    A.w.delta[s] = 0;
    if (A.remove_dc) {
      double dc = A.st.dc_mem[s] + A.st.syn_dc[s];
      A.w.delta[s] = (int)A.st.syn_dc[s];
      A.st.syn_dc[s] = 0;
      for (int i = 0; i < kFrame; i++) {
        ls[w][i] = (short)(int)floor(.5 + dc);
        dc += DC_CONST * ((double)xs[w][i] - dc);
      }
      A.st.dc_mem[s] = dc;
    }
  }
  __syncwarp();
  for (int i = lane; i < kFrame; i += 32) {
    const short l = A.remove_dc ? ls[w][i] : (short)0;
    lp[i] = l; in[i] = (short)(xs[w][i] - l);
  }
}

// the cross-fade of the first good frame (lpcnet_plc.c:225-229): w over FRAME_SIZE - TRAINING_OFFSET = 80 samples;
// the faded samples are also the teacher-forcing input of the segment that follows
__global__ void c_crossfade_kernel(PlcArgs A, const uint8_t *mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= A.n_streams * kTOff) return;
  const int s = i / kTOff, k = i % kTOff;
  if (!mask[s]) return;
  const float w = A.fade[k];
  short *in = A.w.in + (size_t)s * kFrame;
  const short tmp = A.w.sego[(size_t)s * kTOff + k];
  const short v = (short)(int)floor(.5 + (double)(w * (float)in[k]) + (double)((1.f - w) * (float)((int)tmp - A.w.delta[s])));
  in[k] = v;
  A.w.segf[(size_t)s * kTOff + k] = v;
}

// end of a frame: update streams restore the DC offset (lpcnet_plc.c:281-285); conceal streams assemble the two
// 80-sample segments, hand them to the analysis (w.in) and add the DC estimate (lpcnet_plc.c:332-338)
__global__ void __launch_bounds__(kWpsThreads) c_finish_kernel(PlcArgs A, const uint8_t *upd, const uint8_t *con, int stage) {
  __shared__ short xs[kWpsThreads / 32][kFrame];
  int lane, w;
  const int s = wps_stream(lane, w);
  if (s >= A.n_streams) return;
  short *io = A.pcm_io + ((size_t)s * A.total_frames + A.frame) * kFrame;
  short *in = A.w.in + (size_t)s * kFrame;
  if (stage == 0) {                       // conceal: before the analysis
    if (!con[s]) return;
    for (int i = lane; i < kFrame; i += 32) {
      const short v = i < kTOff ? A.w.sego[(size_t)s * kTOff + i] : A.w.sego2[(size_t)s * kTOff + i - kTOff];
      in[i] = v; xs[w][i] = v;
    }
    __syncwarp();
    int dc = 0;
    if (A.remove_dc) {
      dc = (int)floor(.5 + A.st.dc_mem[s]);
      if (lane == 0) {
        double sd = A.st.syn_dc[s];
        for (int i = 0; i < kFrame; i++) sd += DC_CONST * ((double)xs[w][i] - sd);
        A.st.syn_dc[s] = sd;
      }
    }
    for (int i = lane; i < kFrame; i += 32) io[i] = (short)(xs[w][i] + dc);
  } else {                                // update: after everything
    if (!upd[s]) return;
    const short *lp = A.w.lp + (size_t)s * kFrame;
    for (int i = lane; i < kFrame; i += 32) io[i] = (short)(in[i] + lp[i]);
  }
}
