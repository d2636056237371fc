// Weight-blob loader: validates a "DNNw" blob with the rules of the reference parser
// (src/parse_lpcnet_weights.c:36-113), looks the LPCNet arrays up by name and exact size
// (the call list of the generated nnet_data.c, training_tf2/dump_lpcnet.py),
// re-tiles the block-sparse int8 matrices into mma.sync B-fragment order and uploads
// everything to the GPU in one allocation.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <map>

#include "engine.h"
#include "bitexact.cuh"

namespace lpcb {

namespace {

struct Arr {
  const unsigned char *data;
  int size;
};

// parse_record / parse_weights (parse_lpcnet_weights.c:36-77): 64-byte head
// {char[4], int version, int type, int size, int block_size, char name[44]}.
bool parse_records(const unsigned char *d, int len, std::map<std::string, Arr> &out, std::string &err) {
  while (len > 0) {
    if (len < 64) { err = "truncated record head"; return false; }
    int32_t size, block;
    memcpy(&size, d + 12, 4);
    memcpy(&block, d + 16, 4);
    const char *name = reinterpret_cast<const char *>(d) + 20;
    if (block < size || block > len - 64 || name[43] != 0 || size <= 0) {
      err = "malformed weight record";
      return false;
    }
    if (!out.count(name)) out[name] = Arr{d + 64, size};  // first match wins (find_array_entry)
    d += 64 + block;
    len -= 64 + block;
  }
  return true;
}

struct Builder {
  std::vector<unsigned char> buf;
  size_t add(const void *p, size_t bytes) {
    size_t off = (buf.size() + 255) & ~size_t(255);
    buf.resize(off + ((bytes + 15) & ~size_t(15)));   // bulk copies move whole 16-byte blocks: the tail is readable zeros
    if (bytes) memcpy(buf.data() + off, p, bytes);
    return off;
  }
};

// find_idx_check (parse_lpcnet_weights.c:90-113) + re-tiling.  `idx` NULL means a dense
// dotp-layout matrix [rows/8][cols/4][8][4] (dump_lpcnet.py:56-60).
bool pack_sparse(PackedSparse &ps, const int32_t *idx, int idx_len, const int8_t *w, int w_bytes, int rows,
                 int cols, std::string &err, const char *what, bool slotted = false, int parts = 1, int pad_to = 1,
                 bool no_slots = false) {
  ps.n_groups = rows / 8;
  ps.chunk_start.assign(ps.n_groups + 1, 0);
  std::vector<std::vector<int>> gpos(ps.n_groups);
  int total = 0;
  if (idx) {
    int remain = idx_len, grp = 0;
    const int32_t *p = idx;
    while (remain > 0) {
      int nb = *p++;
      if (nb < 0 || nb > remain - 1 || grp >= ps.n_groups) { err = std::string(what) + ": bad block index stream"; return false; }
      for (int k = 0; k < nb; k++) {
        int pos = *p++;
        if (pos + 3 >= cols || (pos & 3) || pos < 0) { err = std::string(what) + ": bad block position"; return false; }
        gpos[grp].push_back(pos);
      }
      remain -= nb + 1;
      total += nb;
      grp++;
    }
    if (grp != ps.n_groups) { err = std::string(what) + ": block index stream does not cover all rows"; return false; }
  } else {
    for (int g = 0; g < ps.n_groups; g++)
      for (int pos = 0; pos < cols; pos += 4) gpos[g].push_back(pos);
    total = ps.n_groups * (cols / 4);
  }
  if (w_bytes != 32 * total) { err = std::string(what) + ": weight array size does not match the block count"; return false; }
  // The AVX2 kernel sums adjacent input pairs in a saturating s16 (vpmaddubsw, src/vec_avx.h:814-818); integer
  // tensor-core accumulation cannot saturate.  Pairs that can saturate (same sign, 255*(|q0|+|q1|) > 32767 -- the
  // training-time WeightClip bounds adjacent OUTPUTS, not these adjacent inputs, so trained models may hold some) are
  // zeroed in the MMA operands and listed for the kernels' exact s16 pass (engine.h DevSat).
  std::vector<int8_t> wz(w, w + w_bytes);
  struct SatRaw { int grp, r, col; uint32_t ww; };
  std::vector<SatRaw> sat_raw;
  {
    int blk = 0;
    for (int g = 0; g < ps.n_groups; g++)
      for (size_t b = 0; b < gpos[g].size(); b++, blk++)
        for (int r = 0; r < 8; r++)
          for (int pr = 0; pr < 2; pr++) {
            const int i = blk * 32 + r * 4 + 2 * pr, q0 = w[i], q1 = w[i + 1];
            if ((q0 > 0) == (q1 > 0) && q0 != 0 && q1 != 0 && 255 * (abs(q0) + abs(q1)) > 32767) {
              sat_raw.push_back(SatRaw{g, r, gpos[g][b] + 2 * pr, (uint32_t)(uint8_t)q0 | (uint32_t)(uint8_t)q1 << 8});
              wz[i] = wz[i + 1] = 0;
            }
          }
  }
  w = wz.data();      // everything below packs the copy
  {
    // lists per unit group (GRU-A: rows are gate-major, group = gate * nug + ug) or per row group
    const int nug = slotted ? ps.n_groups / 3 : ps.n_groups;
    std::vector<std::vector<uint2>> lists(nug);
    for (const SatRaw &e : sat_raw) {
      const int list = slotted ? e.grp % nug : e.grp, gate = slotted ? e.grp / nug : 0;
      lists[list].push_back(make_uint2((uint32_t)e.col | (uint32_t)e.r << 16 | (uint32_t)gate << 20, e.ww));
    }
    ps.sat_start.assign(nug + 1, 0);
    ps.sat.clear();
    for (int l = 0; l < nug; l++) {
      ps.sat_start[l] = (uint32_t)ps.sat.size();
      ps.sat.insert(ps.sat.end(), lists[l].begin(), lists[l].end());
    }
    ps.sat_start[nug] = (uint32_t)ps.sat.size();
  }
  ps.n_blocks = total;
  // weight pointer of every block, in stream order
  std::vector<std::vector<const int8_t *>> gw(ps.n_groups);
  {
    const int8_t *wp = w;
    for (int g = 0; g < ps.n_groups; g++)
      for (size_t b = 0; b < gpos[g].size(); b++, wp += 32) gw[g].push_back(wp);
  }
  // Order every row group's blocks so that each run of four (one 4-byte gather per lane%4, engine.h
  // kXaStride) has four distinct positions mod 4 wherever the group allows it; the int32 sum does not
  // depend on the order.  Unused slots repeat a position of their own run (same word: a broadcast,
  // not a conflict) with zero weights.
  struct Slot { int pos; const int8_t *w; };
  std::vector<std::vector<Slot>> gslots(ps.n_groups);
  for (int g = 0; g < ps.n_groups; g++) {
    const int n = (int)gpos[g].size(), Q = 2 * ((n + 7) / 8);
    std::vector<std::vector<Slot>> quad(Q);
    std::vector<Slot> bucket[4], left;
    for (int b = 0; b < n; b++) bucket[(gpos[g][b] / 4) & 3].push_back(Slot{gpos[g][b], gw[g][b]});
    for (int r = 0; r < 4; r++)
      for (size_t k = 0; k < bucket[r].size(); k++) {
        if ((int)k < Q) quad[k].push_back(bucket[r][k]);
        else left.push_back(bucket[r][k]);
      }
    // fullest runs first, so the sparse ones (which absorb the leftovers) end up last
    std::stable_sort(quad.begin(), quad.end(), [](const std::vector<Slot> &a, const std::vector<Slot> &b) { return a.size() > b.size(); });
    for (int q = Q - 1; q >= 0 && !left.empty(); q--)
      while (quad[q].size() < 4 && !left.empty()) { quad[q].push_back(left.back()); left.pop_back(); }
    for (int q = 0; q < Q; q++) {
      const int fill = quad[q].empty() ? 0 : quad[q][0].pos;
      for (int k = 0; k < 4; k++) gslots[g].push_back(k < (int)quad[q].size() ? quad[q][k] : Slot{fill, nullptr});
    }
  }
  auto emit = [&](int c, int slot, const Slot &sl) {
    ps.pos[(size_t)c * 8 + slot] = (uint8_t)(sl.pos / 4);
    if (!sl.w) return;
    // B fragment of mma.m16n8k32: register slot/4 of lane = row*4 + slot%4 holds the four input
    // weights of output row `row` of block `slot`; the two registers of a lane are adjacent.
    for (int r = 0; r < 8; r++) {
      uint32_t word;
      memcpy(&word, sl.w + r * 4, 4);
      ps.w[(size_t)c * 64 + (r * 4 + (slot % 4)) * 2 + slot / 4] = word;
    }
  };
  if (!slotted) {
    // every row group's chunks split into `parts` K ranges (one kernel job each), each range padded with
    // empty chunks to a multiple of `pad_to` so that the kernel's unrolled loops need no tail
    int nchunks = 0;
    ps.chunk_start.assign((size_t)ps.n_groups * parts + 1, 0);
    std::vector<int> first(ps.n_groups * parts), count(ps.n_groups * parts);
    for (int g = 0; g < ps.n_groups; g++) {
      const int n = (int)gslots[g].size() / 8;
      for (int p = 0; p < parts; p++) {
        const int lo = n * p / parts, hi = n * (p + 1) / parts;
        ps.chunk_start[g * parts + p] = nchunks;
        first[g * parts + p] = lo; count[g * parts + p] = hi - lo;
        nchunks += (hi - lo + pad_to - 1) / pad_to * pad_to;
      }
    }
    ps.chunk_start[(size_t)ps.n_groups * parts] = nchunks;
    ps.n_chunks = nchunks;
    ps.pos.assign((size_t)nchunks * 8, 0);
    ps.w.assign((size_t)nchunks * 64, 0);
    for (int g = 0; g < ps.n_groups; g++)
      for (int p = 0; p < parts; p++)
        for (int k = 0; k < count[g * parts + p] * 8; k++)
          emit(ps.chunk_start[g * parts + p] + k / 8, k % 8, gslots[g][(size_t)first[g * parts + p] * 8 + k]);
    ps.off.assign((size_t)nchunks * 4, 0);
    for (int c = 0; c < nchunks; c++)
      for (int t = 0; t < 4; t++)
        ps.off[(size_t)c * 4 + t] = (uint32_t)ps.pos[(size_t)c * 8 + t] * 4u | ((uint32_t)ps.pos[(size_t)c * 8 + 4 + t] * 4u) << 16;
  } else {
    // rows are gate-major: group index = gate*(n_groups/3) + ug
    const int nug = ps.n_groups / 3;
    for (int gate = 0; gate < 3; gate++) {
      long long blocks = 0;
      for (int ug = 0; ug < nug; ug++) blocks += (long long)gpos[gate * nug + ug].size();
      const double mean = (double)blocks / nug;
      int sl = (int)((mean + 7.999) / 8.0);         // chunks that hold the average row group
      ps.slots[gate] = sl < 1 ? 1 : (sl > 3 ? 3 : sl);
    }
    // the kernel's unrolled fast path is built for (1,1,3); anything else runs the generic lists
    if (!(ps.slots[0] == 1 && ps.slots[1] == 1 && ps.slots[2] == 3) || getenv("LPCNET_B200_NO_SLOTS") || no_slots)
      ps.slots[0] = ps.slots[1] = ps.slots[2] = 0;
    const int ST = ps.slots[0] + ps.slots[1] + ps.slots[2];
    int nchunks = nug * ST;
    ps.chunk_start.assign(nug + 1, 0);
    for (int ug = 0; ug < nug; ug++) {
      ps.chunk_start[ug] = nchunks;
      for (int gate = 0; gate < 3; gate++) {
        const int over = (int)gslots[gate * nug + ug].size() - 8 * ps.slots[gate];
        if (over > 0) nchunks += over / 8;
      }
    }
    ps.chunk_start[nug] = nchunks;
    ps.n_chunks = nchunks;
    ps.pos.assign((size_t)nchunks * 8, 0);
    ps.w.assign((size_t)nchunks * 64, 0);
    std::vector<int> gate_of(nchunks, 0);
    for (int ug = 0; ug < nug; ug++) {
      int slot0 = 0, oc = ps.chunk_start[ug];
      for (int gate = 0; gate < 3; gate++) {
        const int g = gate * nug + ug, cap = 8 * ps.slots[gate];
        for (int b = 0; b < (int)gslots[g].size(); b++) {
          const int c = b < cap ? ug * ST + slot0 + b / 8 : oc + (b - cap) / 8;
          emit(c, b % 8, gslots[g][b]);
          gate_of[c] = gate;
        }
        if ((int)gslots[g].size() > cap) oc += ((int)gslots[g].size() - cap) / 8;
        slot0 += ps.slots[gate];
      }
    }
    ps.off.assign((size_t)nchunks * 4, 0);
    for (int c = 0; c < nchunks; c++)
      for (int t = 0; t < 4; t++)
        ps.off[(size_t)c * 4 + t] = (uint32_t)ps.pos[(size_t)c * 8 + t] * 4u | ((uint32_t)ps.pos[(size_t)c * 8 + 4 + t] * 4u) << 16 |
                                    (uint32_t)gate_of[c] << 30;
    ps.off_fixed.assign((size_t)nug * 32, 0);
    for (int ug = 0; ug < nug; ug++)
      for (int sl = 0; sl < ST; sl++)
        for (int t = 0; t < 4; t++) ps.off_fixed[((size_t)ug * 4 + t) * 8 + sl] = ps.off[((size_t)ug * ST + sl) * 4 + t] & 0x3fffffffu;
  }
  return true;
}

}  // namespace

// Model of the Intel `rcpps` estimate: entry i = 1/midpoint of bucket i rounded to 2^-13
// (measured; oracle/orc_prim.c).
void default_rcp_table(uint32_t *tab) {
  for (int i = 0; i < 2048; i++) {
    uint32_t k = 4097u + 2u * (uint32_t)i;
    uint32_t q = (uint32_t)(((1ull << 26) / k + 1) >> 1);
    tab[i] = 0x3f000000u + ((q - 4096u) << 11);
  }
}

// every mantissa bucket (both ends of it) x every exponent the activations can produce, and far beyond
__global__ void rcp_check_kernel(const uint32_t *tab, int *mismatch) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;   // bucket*2 + end
  const uint32_t mant = ((i >> 1) << 12) | ((i & 1) ? 0xfffu : 0u);
  int bad = 0;
  for (uint32_t e = 8; e <= 246; e++) {
    const float x = __uint_as_float((e << 23) | mant);
    bad |= __float_as_uint(rcp_estimate(x, tab)) != __float_as_uint(rcp_estimate(x, RcpArith{}));
  }
  if (bad) atomicOr(mismatch, 1);
}

void free_model(Model &m) {
  if (m.dev_blob) cudaFree(m.dev_blob);
  if (m.dev_codebooks) cudaFree(m.dev_codebooks);
  m.dev_blob = nullptr;
  m.dev_codebooks = nullptr;
}

// kiss_fft tables for nfft = 320 (src/kiss_fft.c:409-424, 312-345; factors {5,64,4,16,4,4,4,1}
// src/lpcnet_tables.c:197-205) and the 18x18 DCT (src/dump_lpcnet_tables.c:83-99).
void host_dsp_tables(std::vector<float> &tw, std::vector<int16_t> &br, std::vector<float> &dct) {
  tw.assign(640, 0.f);
  for (int i = 0; i < 320; i++) {
    const double pi = 3.14159265358979323846264338327;
    double phase = (-2 * pi / 320) * i;
    tw[2 * i] = (float)cos(phase);
    tw[2 * i + 1] = (float)sin(phase);
  }
  static const int factors[8] = {5, 64, 4, 16, 4, 4, 4, 1};
  br.assign(320, 0);
  struct Rec {
    static void go(int Fout, int16_t *f, int fstride, const int *fac) {
      int p = *fac++, mm = *fac++;
      if (mm == 1) {
        for (int j = 0; j < p; j++) { *f = (int16_t)(Fout + j); f += fstride; }
      } else {
        for (int j = 0; j < p; j++) { go(Fout, f, fstride * p, fac); f += fstride; Fout += mm; }
      }
    }
  };
  Rec::go(0, br.data(), 1, factors);
  dct.assign(18 * 18, 0.f);
  for (int i = 0; i < 18; i++)
    for (int j = 0; j < 18; j++) {
      float v = (float)cos((i + .5) * j * M_PI / 18);
      if (j == 0) v = (float)((double)v * sqrt(.5));
      dct[i * 18 + j] = v;
    }
}

bool load_model(Model &m, const unsigned char *blob, int len, int features_delay, float lpc_gamma,
                const uint32_t *rcp_tab, std::string &err) {
  if (features_delay < 0 || features_delay > kMaxDelay) { err = "features_delay must be 0..2"; return false; }
  if (!blob || len <= 0) { err = "empty blob"; return false; }
  std::map<std::string, Arr> A;
  if (!parse_records(blob, len, A, err)) return false;
  // find_array_check (parse_lpcnet_weights.c:84-88): name and exact byte size.
  auto need = [&](const char *name, int bytes) -> const unsigned char * {
    auto it = A.find(name);
    if (it == A.end()) { if (err.empty()) err = std::string("missing array ") + name; return nullptr; }
    if (it->second.size != bytes) { if (err.empty()) err = std::string("wrong size for array ") + name; return nullptr; }
    return it->second.data;
  };
  Builder B;
  struct Fix { size_t off; const void **dst; };
  std::vector<Fix> fixes;
  auto put = [&](const void **dst, const void *src, size_t bytes) { fixes.push_back(Fix{B.add(src, bytes), dst}); };
  auto put_f = [&](const float **dst, const char *name, int n) -> bool {
    const unsigned char *p = need(name, n * 4);
    if (!p) return false;
    put(reinterpret_cast<const void **>(dst), p, (size_t)n * 4);
    return true;
  };
  DevModel &d = m.d;
  bool ok = true;
  // embedding rows are stored in the column order the sample kernel's 16-byte gathers want
  auto put_embed = [&](const float **dst, const char *name) -> bool {
    const unsigned char *p = need(name, 256 * kRowsA * 4);
    if (!p) return false;
    std::vector<float> t((size_t)256 * kRowsAP, 0.f);
    for (int v = 0; v < 256; v++)
      for (int col = 0; col < kRowsA; col++)
        memcpy(&t[(size_t)v * kRowsAP + gru_a_col_index(col)], p + ((size_t)v * kRowsA + col) * 4, 4);
    put(reinterpret_cast<const void **>(dst), t.data(), t.size() * 4);
    return true;
  };
  ok &= put_embed(&d.e_sig, "gru_a_embed_sig_weights");
  ok &= put_embed(&d.e_pred, "gru_a_embed_pred_weights");
  ok &= put_embed(&d.e_exc, "gru_a_embed_exc_weights");
  ok &= put_f(&d.a_feat_b, "gru_a_dense_feature_bias", kRowsA);
  ok &= put_f(&d.a_feat_w, "gru_a_dense_feature_weights", kC * kRowsA);
  ok &= put_f(&d.b_feat_b, "gru_b_dense_feature_bias", kRowsB);
  ok &= put_f(&d.b_feat_w, "gru_b_dense_feature_weights", kC * kRowsB);
  ok &= put_f(&d.conv1_b, "feature_conv1_bias", kC);
  ok &= put_f(&d.conv1_w, "feature_conv1_weights", 3 * kConvIn * kC);
  ok &= put_f(&d.conv2_b, "feature_conv2_bias", kC);
  ok &= put_f(&d.conv2_w, "feature_conv2_weights", 3 * kC * kC);
  ok &= put_f(&d.embed_pitch, "embed_pitch_weights", 256 * kPitchEmb);
  ok &= need("embed_sig_weights", 256 * 128 * 4) != nullptr;  // loaded by the reference, unused at inference
  ok &= put_f(&d.dense1_b, "feature_dense1_bias", kC);
  ok &= put_f(&d.dense1_w, "feature_dense1_weights", kC * kC);
  ok &= put_f(&d.dense2_b, "feature_dense2_bias", kC);
  ok &= put_f(&d.dense2_w, "feature_dense2_weights", kC * kC);
  ok &= put_f(&d.fc_bias, "dual_fc_bias", 2 * 256);
  ok &= put_f(&d.fc_w, "dual_fc_weights", 256 * 2 * kNB);
  ok &= put_f(&d.fc_factor, "dual_fc_factor", 2 * 256);
  ok &= put_f(&d.a_diag, "sparse_gru_a_recurrent_weights_diag", kRowsA);
  ok &= put_f(&d.b_subias, "gru_b_subias", 2 * kRowsB);
  ok &= need("gru_b_bias", 2 * kRowsB * 4) != nullptr;
  ok &= need("sparse_gru_a_bias", 2 * kRowsA * 4) != nullptr;
  const unsigned char *sub_a = need("sparse_gru_a_subias", 2 * kRowsA * 4);
  ok &= sub_a != nullptr;
  if (!ok) return false;
  put(reinterpret_cast<const void **>(&d.a_subias_rec), sub_a + kRowsA * 4, kRowsA * 4);
  // the sample kernel's shared-memory images of three arrays, laid out here so that the kernel stages them with plain
  // bulk copies: [gate*48 + ug][lane%4][sub0 sub1 diag0 diag1]; dual_fc weights at node stride kFcStride; bias | factor
  {
    const float *sub = reinterpret_cast<const float *>(sub_a) + kRowsA, *diag = reinterpret_cast<const float *>(A["sparse_gru_a_recurrent_weights_diag"].data);
    std::vector<float> sd((size_t)kRowsA * 2);
    for (int i = 0; i < kRowsA; i++) {
      const int k = (i >> 3) * 16 + ((i & 7) >> 1) * 4 + (i & 1);
      sd[k] = sub[i];
      sd[k + 2] = diag[i];
    }
    put(reinterpret_cast<const void **>(&d.a_sd_packed), sd.data(), sd.size() * 4);
    const float *fw = reinterpret_cast<const float *>(A["dual_fc_weights"].data);
    std::vector<float> fp((size_t)256 * kFcStride, 0.f);
    for (int i = 0; i < 256 * 32; i++) fp[(size_t)(i >> 5) * kFcStride + (i & 31)] = fw[i];
    put(reinterpret_cast<const void **>(&d.fc_w_padded), fp.data(), fp.size() * 4);
    std::vector<float> fb(1024);
    memcpy(fb.data(), A["dual_fc_bias"].data, 512 * 4);
    memcpy(fb.data() + 512, A["dual_fc_factor"].data, 512 * 4);
    put(reinterpret_cast<const void **>(&d.fc_bf), fb.data(), fb.size() * 4);
  }

  // The kernels seed their int32 accumulators with cvt.rni(16256*t) and skip the out-of-range fix-up of
  // cvtps2dq (-> 0x80000000): prove here that no pre-activation t can reach 2^31/16256 for states in [-1,1].
  double ta_bound = 0, tb_bound = 0;
  {
    auto amax = [](const unsigned char *p, size_t n) {
      const float *f = reinterpret_cast<const float *>(p);
      double m = 0;
      for (size_t i = 0; i < n; i++) { double v = fabs((double)f[i]); if (!(v <= m)) m = v; }
      return m;
    };
    auto cond_bound = [&](const char *bn, const char *wn, int nout) {   // |bias| + sum_j |W[j][i]| (inputs are tanh outputs)
      const float *b = reinterpret_cast<const float *>(A[bn].data), *w = reinterpret_cast<const float *>(A[wn].data);
      double m = 0;
      for (int i = 0; i < nout; i++) {
        double v = fabs((double)b[i]);
        for (int j = 0; j < kC; j++) v += fabs((double)w[(size_t)j * nout + i]);
        if (!(v <= m)) m = v;
      }
      return m;
    };
    const double ta = amax(sub_a, 2 * kRowsA) + amax(A["sparse_gru_a_recurrent_weights_diag"].data, kRowsA) +
                      amax(A["gru_a_embed_sig_weights"].data, 256 * kRowsA) + amax(A["gru_a_embed_pred_weights"].data, 256 * kRowsA) +
                      amax(A["gru_a_embed_exc_weights"].data, 256 * kRowsA) +
                      cond_bound("gru_a_dense_feature_bias", "gru_a_dense_feature_weights", kRowsA);
    const double tb = amax(A["gru_b_subias"].data, 2 * kRowsB) + cond_bound("gru_b_dense_feature_bias", "gru_b_dense_feature_weights", kRowsB);
    const double lim = 2147483648.0 / 16256.0 * 0.99;
    if (!(ta < lim) || !(tb < lim)) { err = "model pre-activations can exceed the int32 accumulator seed range (or contain NaN)"; return false; }
    ta_bound = ta; tb_bound = tb;
  }

  // int8 matrices
  PackedSparse pa, pb, pr;
  {
    auto ia = A.find("sparse_gru_a_recurrent_weights_idx"), wa = A.find("sparse_gru_a_recurrent_weights");
    if (ia == A.end() || wa == A.end()) { err = "missing GRU-A recurrent arrays"; return false; }
    if (!pack_sparse(pa, reinterpret_cast<const int32_t *>(ia->second.data), ia->second.size / 4,
                     reinterpret_cast<const int8_t *>(wa->second.data), wa->second.size, kRowsA, kN, err, "sparse_gru_a", true))
      return false;
    auto ib = A.find("gru_b_weights_idx"), wb = A.find("gru_b_weights");
    if (ib == A.end() || wb == A.end()) { err = "missing GRU-B input arrays"; return false; }
    if (!pack_sparse(pb, reinterpret_cast<const int32_t *>(ib->second.data), ib->second.size / 4,
                     reinterpret_cast<const int8_t *>(wb->second.data), wb->second.size, kRowsB, kN, err, "gru_b", false, kPartsB, kPadB))
      return false;
    // dense 48 x 384 copy of the GRU-B input matrix (absent blocks and the pairs on the s16 pass are zeros), cut into
    // the B fragments of the kernel's register-operand product (engine.h kBinFragWords)
    {
      std::vector<int8_t> W((size_t)kRowsB * kN, 0);
      const int32_t *ip = reinterpret_cast<const int32_t *>(ib->second.data);
      const int8_t *wp = reinterpret_cast<const int8_t *>(wb->second.data);
      for (int rg = 0; rg < kRowsB / 8; rg++) {
        const int nb = *ip++;
        for (int k = 0; k < nb; k++, wp += 32) {
          const int pos = *ip++;
          for (int r = 0; r < 8; r++)
            for (int c = 0; c < 4; c++) W[(size_t)(rg * 8 + r) * kN + pos + c] = wp[r * 4 + c];
        }
      }
      for (int rg = 0; rg < kRowsB / 8; rg++)
        for (uint32_t e = pb.sat_start[rg]; e < pb.sat_start[rg + 1]; e++) {
          const int col = pb.sat[e].x & 0xffff, r = (pb.sat[e].x >> 16) & 7;
          W[(size_t)(rg * 8 + r) * kN + col] = W[(size_t)(rg * 8 + r) * kN + col + 1] = 0;
        }
      std::vector<uint32_t> frag(kBinFragWords, 0u);
      for (int w = 0; w < 16; w++)
        for (int rg = 0; rg < 6; rg++)
          for (int lane = 0; lane < 32; lane++) {
            const int n = lane >> 2, t = lane & 3;
            const int8_t *row = &W[(size_t)(rg * 8 + n) * kN];
            auto byte = [&](int ug, int j) { return (uint32_t)(uint8_t)row[ug * 8 + 2 * t + j]; };
            const uint32_t b0 = byte(w, 0) | byte(w, 1) << 8 | byte(w + 16, 0) << 16 | byte(w + 16, 1) << 24;
            const uint32_t b1 = byte(w + 32, 0) | byte(w + 32, 1) << 8;
            frag[((size_t)(w * 6 + rg) * 32 + lane) * 2] = b0;
            frag[((size_t)(w * 6 + rg) * 32 + lane) * 2 + 1] = b1;
          }
      put(reinterpret_cast<const void **>(&d.b_in_frag), frag.data(), frag.size() * 4);
    }
    const unsigned char *wr = need("gru_b_recurrent_weights", kRowsB * kNB);
    if (!wr) return false;
    if (!pack_sparse(pr, nullptr, 0, reinterpret_cast<const int8_t *>(wr), kRowsB * kNB, kRowsB, kNB, err, "gru_b_recurrent"))
      return false;
  }
  // The fast kernel variants seed and read their int32 sums through float bit patterns (bitexact.cuh acc_seed_small),
  // which is exact while |seed + sum of products| < 2^22: bound it per row (u8 activations are at most 255).
  {
    auto row_l1 = [](const int32_t *idx, const int8_t *w, int n_groups, int dense_blocks) {   // max over rows of sum |w|
      double worst = 0;
      for (int g = 0; g < n_groups; g++) {
        const int nb = idx ? *idx++ : dense_blocks;
        double l1[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int k = 0; k < nb; k++, w += 32) {
          if (idx) idx++;
          for (int r = 0; r < 8; r++)
            for (int c = 0; c < 4; c++) l1[r] += abs((int)w[r * 4 + c]);
        }
        for (int r = 0; r < 8; r++) worst = l1[r] > worst ? l1[r] : worst;
      }
      return worst;
    };
    auto ia = A.find("sparse_gru_a_recurrent_weights_idx"), wa = A.find("sparse_gru_a_recurrent_weights");
    auto ib = A.find("gru_b_weights_idx"), wb = A.find("gru_b_weights");
    const double la = row_l1(reinterpret_cast<const int32_t *>(ia->second.data), reinterpret_cast<const int8_t *>(wa->second.data), kRowsA / 8, 0);
    const double lb = row_l1(reinterpret_cast<const int32_t *>(ib->second.data), reinterpret_cast<const int8_t *>(wb->second.data), kRowsB / 8, 0);
    const double lr = row_l1(nullptr, reinterpret_cast<const int8_t *>(A["gru_b_recurrent_weights"].data), kRowsB / 8, kNB / 4);
    const double lim22 = 4194304.0 * 0.98;
    d.small_acc = (ta_bound * 16256.0 + 255.0 * la < lim22) && (tb_bound * 16256.0 + 255.0 * (lb > lr ? lb : lr) < lim22) ? 1 : 0;
  }
  // models with pairs that can saturate vpmaddubsw run the kernel variant with the exact s16 pass, which is built on
  // the generic chunk lists and the rcpps table (sample_kernel.cuh launch_tile): re-tile GRU-A without the fixed slots
  d.has_sat = (pa.sat.size() + pb.sat.size() + pr.sat.size()) > 0 ? 1 : 0;
  if (d.has_sat && pa.slots[2] != 0) {
    auto ia = A.find("sparse_gru_a_recurrent_weights_idx"), wa = A.find("sparse_gru_a_recurrent_weights");
    pa = PackedSparse();
    if (!pack_sparse(pa, reinterpret_cast<const int32_t *>(ia->second.data), ia->second.size / 4,
                     reinterpret_cast<const int8_t *>(wa->second.data), wa->second.size, kRowsA, kN, err, "sparse_gru_a", true, 1, 1, true))
      return false;
  }
  m.nblocks_a = pa.n_blocks;
  m.nblocks_b = pb.n_blocks;
  auto put_sparse = [&](DevSparse &ds, const PackedSparse &ps) {
    ds.n_groups = ps.n_groups;
    ds.n_chunks = ps.n_chunks;
    put(reinterpret_cast<const void **>(&ds.chunk_start), ps.chunk_start.data(), ps.chunk_start.size() * 4);
    put(reinterpret_cast<const void **>(&ds.pos), ps.pos.data(), ps.pos.size());
    put(reinterpret_cast<const void **>(&ds.w), ps.w.data(), ps.w.size() * 4);
    ds.off = ds.off_fixed = nullptr;
    ds.sat.n = (int)ps.sat.size();
    ds.sat.start = nullptr; ds.sat.e = nullptr;
    put(reinterpret_cast<const void **>(&ds.sat.start), ps.sat_start.data(), ps.sat_start.size() * 4);
    if (!ps.sat.empty()) put(reinterpret_cast<const void **>(&ds.sat.e), ps.sat.data(), ps.sat.size() * 8);
    for (int k = 0; k < 3; k++) ds.slots[k] = ps.slots[k];
    if (!ps.off.empty()) put(reinterpret_cast<const void **>(&ds.off), ps.off.data(), ps.off.size() * 4);
    if (!ps.off_fixed.empty()) put(reinterpret_cast<const void **>(&ds.off_fixed), ps.off_fixed.data(), ps.off_fixed.size() * 4);
  };
  put_sparse(d.a_rec, pa);
  put_sparse(d.b_in, pb);
  put_sparse(d.b_rec, pr);

  // tables
  std::vector<uint32_t> rcp(2048);
  if (rcp_tab) memcpy(rcp.data(), rcp_tab, 2048 * 4);
  else default_rcp_table(rcp.data());
  put(reinterpret_cast<const void **>(&d.rcp_tab), rcp.data(), 2048 * 4);
  float u2l[256], logit[256];
  for (int i = 0; i < 256; i++) {
    // ulaw2lin (src/common.h:37-45): double exp, then float
    float u = (float)i - 128.f;
    float s = u >= 0.f ? 1.f : -1.f;
    float scale_1 = 32768.f / 255.f;
    u = fabsf(u);
    u2l[i] = (float)((double)(s * scale_1) * (exp((double)u / 128. * (double)5.5451774445f) - 1));
    // sampling_logit_table (src/lpcnet.c:188-191)
    float prob = .025f + .95f * (float)i / 255.f;
    logit[i] = (float)(-log((double)(((float)1 - prob) / prob)));
  }
  put(reinterpret_cast<const void **>(&d.ulaw2lin_tab), u2l, sizeof(u2l));
  put(reinterpret_cast<const void **>(&d.logit_tab), logit, sizeof(logit));
  {
    std::vector<float> tw, dct;
    std::vector<int16_t> br;
    host_dsp_tables(tw, br, dct);
    put(reinterpret_cast<const void **>(&d.fft_tw), tw.data(), tw.size() * 4);
    put(reinterpret_cast<const void **>(&d.fft_bitrev), br.data(), br.size() * 2);
    put(reinterpret_cast<const void **>(&d.dct), dct.data(), dct.size() * 4);
  }
  // optional VQ codebooks carried in the blob under their C names (src/lpcnet_private.h:109-112)
  d.codebooks = nullptr;
  {
    auto c1 = A.find("ceps_codebook1"), c2 = A.find("ceps_codebook2"), c3 = A.find("ceps_codebook3"),
         c4 = A.find("ceps_codebook_diff4");
    if (c1 != A.end() && c2 != A.end() && c3 != A.end() && c4 != A.end() && c1->second.size == 1024 * 17 * 4 &&
        c2->second.size == 1024 * 17 * 4 && c3->second.size == 1024 * 17 * 4 && c4->second.size == 4096 * 18 * 4) {
      std::vector<float> cb(3 * 1024 * 17 + 4096 * 18);
      memcpy(cb.data(), c1->second.data, 1024 * 17 * 4);
      memcpy(cb.data() + 1024 * 17, c2->second.data, 1024 * 17 * 4);
      memcpy(cb.data() + 2 * 1024 * 17, c3->second.data, 1024 * 17 * 4);
      memcpy(cb.data() + 3 * 1024 * 17, c4->second.data, 4096 * 18 * 4);
      put(reinterpret_cast<const void **>(&d.codebooks), cb.data(), cb.size() * 4);
      m.has_codebooks = true;
    }
  }
  // [input/4][row] words from the block-sparse stream (absent blocks = 0) or the dense 8x4 layout
  auto retile = [&](const uint32_t **dst, DevSat *sat, const char *wname, const char *iname, int cols) -> bool {
    auto wi = A.find(wname);
    if (wi == A.end()) { if (err.empty()) err = std::string("missing array ") + wname; return false; }
    const int8_t *w = reinterpret_cast<const int8_t *>(wi->second.data);
    const int wb = wi->second.size;
    std::vector<uint32_t> t((size_t)(cols / 4) * kPlcRows, 0u);
    // pairs that can saturate vpmaddubsw's s16 leave the dp4a words and go to per-row lists (engine.h DevSat)
    std::vector<std::vector<uint2>> rows_sat(kPlcRows);
    auto place = [&](int grp, int pos, const int8_t *blk) {
      for (int r = 0; r < 8; r++) {
        int8_t q[4];
        memcpy(q, blk + 4 * r, 4);
        for (int pr = 0; pr < 2; pr++) {
          const int q0 = q[2 * pr], q1 = q[2 * pr + 1];
          if ((q0 > 0) == (q1 > 0) && q0 != 0 && q1 != 0 && 255 * (abs(q0) + abs(q1)) > 32767) {
            rows_sat[grp * 8 + r].push_back(make_uint2((uint32_t)(pos + 2 * pr), (uint32_t)(uint8_t)q0 | (uint32_t)(uint8_t)q1 << 8));
            q[2 * pr] = q[2 * pr + 1] = 0;
          }
        }
        memcpy(&t[(size_t)(pos / 4) * kPlcRows + grp * 8 + r], q, 4);
      }
    };
    if (iname) {
      auto ii = A.find(iname);
      if (ii == A.end() || (ii->second.size & 3)) { if (err.empty()) err = std::string("missing array ") + iname; return false; }
      const int32_t *p = reinterpret_cast<const int32_t *>(ii->second.data);
      int remain = ii->second.size / 4, grp = 0, total = 0;
      while (remain > 0) {
        const int nb = *p++;
        if (nb < 0 || nb > remain - 1 || grp >= kPlcRows / 8 || (long long)nb > (long long)wb / 32 - total) { err = std::string(iname) + ": bad block index stream"; return false; }
        for (int k = 0; k < nb; k++) {
          const int pos = *p++;
          if (pos < 0 || (pos & 3) || pos + 3 >= cols) { err = std::string(iname) + ": bad block position"; return false; }
          place(grp, pos, w + (size_t)(total + k) * 32);
        }
        remain -= nb + 1; total += nb; grp++;
      }
      if (grp != kPlcRows / 8 || total * 32 != wb) { err = std::string(iname) + ": block index stream does not match the weights"; return false; }
    } else {
      if (wb != kPlcRows * cols) { err = std::string("wrong size for array ") + wname; return false; }
      for (int grp = 0; grp < kPlcRows / 8; grp++)
        for (int pos = 0; pos < cols; pos += 4) place(grp, pos, w + ((size_t)grp * (cols / 4) + pos / 4) * 32);
    }
    put(reinterpret_cast<const void **>(dst), t.data(), t.size() * 4);
    std::vector<uint32_t> start(kPlcRows + 1, 0);
    std::vector<uint2> flat;
    for (int r = 0; r < kPlcRows; r++) { start[r] = (uint32_t)flat.size(); flat.insert(flat.end(), rows_sat[r].begin(), rows_sat[r].end()); }
    start[kPlcRows] = (uint32_t)flat.size();
    sat->n = (int)flat.size(); sat->start = nullptr; sat->e = nullptr;
    put(reinterpret_cast<const void **>(&sat->start), start.data(), start.size() * 4);
    if (!flat.empty()) put(reinterpret_cast<const void **>(&sat->e), flat.data(), flat.size() * 8);
    return true;
  };
  // optional PLC predictor arrays (init_plc_model, training_tf2/dump_plc.py:319-330): all or none
  if (A.find("plc_dense1_bias") != A.end()) {
    DevPlc &P = d.plc;
    bool pk = true;
    pk &= put_f(&P.dense1_b, "plc_dense1_bias", kPlcCond);
    pk &= put_f(&P.dense1_w, "plc_dense1_weights", kPlcIn * kPlcCond);
    pk &= put_f(&P.out_b, "plc_out_bias", 20);
    pk &= put_f(&P.out_w, "plc_out_weights", kPlcUnits * 20);
    pk &= put_f(&P.g1_subias, "plc_gru1_subias", 2 * kPlcRows);
    pk &= put_f(&P.g2_subias, "plc_gru2_subias", 2 * kPlcRows);
    pk &= need("plc_gru1_bias", 2 * kPlcRows * 4) != nullptr;
    pk &= need("plc_gru2_bias", 2 * kPlcRows * 4) != nullptr;
    pk = pk && retile(&P.g1_in, &P.sat_g1_in, "plc_gru1_weights", "plc_gru1_weights_idx", kPlcCond);
    pk = pk && retile(&P.g1_rec, &P.sat_g1_rec, "plc_gru1_recurrent_weights", nullptr, kPlcUnits);
    pk = pk && retile(&P.g2_in, &P.sat_g2_in, "plc_gru2_weights", "plc_gru2_weights_idx", kPlcUnits);
    pk = pk && retile(&P.g2_rec, &P.sat_g2_rec, "plc_gru2_recurrent_weights", nullptr, kPlcUnits);
    if (!pk) return false;
    m.has_plc = true;
  }
  // optional RDOVAE decoder of DRED (init_rdovaedec, torch/rdovae/export_rdovae_weights.py:178-222): all or none
  if (A.find("dec_dense1_bias") != A.end()) {
    DevDred &R = d.dred;
    bool pk = true;
    static const char *dn[5] = {"dec_dense1", "dec_dense3", "dec_dense5", "dec_dense7", "dec_dense8"};
    static const char *gn[3] = {"dec_dense2", "dec_dense4", "dec_dense6"};
    for (int k = 0; k < 3; k++) {
      const std::string n = "state" + std::to_string(k + 1);
      pk &= put_f(&R.state_b[k], (n + "_bias").c_str(), kPlcUnits);
      pk &= put_f(&R.state_w[k], (n + "_weights").c_str(), kDredState * kPlcUnits);
    }
    for (int k = 0; k < 5; k++) {
      pk &= put_f(&R.dense_b[k], (std::string(dn[k]) + "_bias").c_str(), kPlcUnits);
      pk &= put_f(&R.dense_w[k], (std::string(dn[k]) + "_weights").c_str(), (k == 0 ? kDredLatent : kPlcUnits) * kPlcUnits);
    }
    pk &= put_f(&R.final_b, "dec_final_bias", kDredOut);
    pk &= put_f(&R.final_w, "dec_final_weights", 8 * kPlcUnits * kDredOut);
    for (int k = 0; k < 3 && pk; k++) {
      const std::string n = gn[k];
      pk &= put_f(&R.g_subias[k], (n + "_subias").c_str(), 2 * kPlcRows);
      pk &= need((n + "_bias").c_str(), 2 * kPlcRows * 4) != nullptr;
      pk = pk && retile(&R.g_in[k], &R.sat_in[k], (n + "_weights").c_str(), (n + "_weights_idx").c_str(), kPlcUnits);
      pk = pk && retile(&R.g_rec[k], &R.sat_rec[k], (n + "_recurrent_weights").c_str(), nullptr, kPlcUnits);
    }
    if (!pk) return false;
    m.has_dred = true;
  }
  d.features_delay = features_delay;
  d.lpc_gamma = lpc_gamma;
  d.end2end = 0;                       // lpcnet_batch_set_end2end / lpcnet_b200_set_end2end (api.cu)

  cudaError_t e = cudaMalloc(&m.dev_blob, B.buf.size());
  if (e != cudaSuccess) { err = std::string("cudaMalloc(model): ") + cudaGetErrorString(e); return false; }
  m.dev_bytes = B.buf.size();
  e = cudaMemcpy(m.dev_blob, B.buf.data(), B.buf.size(), cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { err = std::string("cudaMemcpy(model): ") + cudaGetErrorString(e); free_model(m); return false; }
  for (auto &f : fixes) *f.dst = static_cast<const unsigned char *>(m.dev_blob) + f.off;
  // The sample kernel may replace the table lookup by MUFU.RCP of the bucket mid-point rounded to 12
  // mantissa bits -- only if that reproduces this table on this GPU for every bucket and exponent.
  d.rcp_arith = 0;
  if (!getenv("LPCNET_B200_RCP_TABLE")) {
    int *flag = nullptr;
    if (cudaMalloc(&flag, 4) == cudaSuccess) {
      cudaMemset(flag, 0, 4);
      rcp_check_kernel<<<2048 * 2 / 256, 256>>>(d.rcp_tab, flag);
      int bad = 1;
      if (cudaMemcpy(&bad, flag, 4, cudaMemcpyDeviceToHost) == cudaSuccess) d.rcp_arith = bad == 0;
      cudaFree(flag);
    }
  }
  return true;
}

bool install_codebooks(Model &m, const float *cb, int n, std::string &err) {
  const int want = 3 * 1024 * 17 + 4096 * 18;
  if (n != want) { err = "codebooks: expected 3*1024*17 + 4096*18 floats"; return false; }
  float *dcb = nullptr;
  cudaError_t e = cudaMalloc(&dcb, (size_t)want * 4);
  if (e == cudaSuccess) e = cudaMemcpy(dcb, cb, (size_t)want * 4, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) { err = std::string("codebooks: ") + cudaGetErrorString(e); return false; }
  if (m.dev_codebooks) cudaFree(m.dev_codebooks);   // a re-install replaces the earlier tables
  m.dev_codebooks = dcb;                            // freed with the model (free_model)
  m.d.codebooks = dcb;
  m.has_codebooks = true;
  return true;
}

}  // namespace lpcb
