// 12 Hz ENCODE path: weight packing, per-(B, T) launch plans and the C entry points.
//
// Reference: Qwen3TTSTokenizerV2Model.encode (modeling :960-990) = MimiModel.encode of the `transformers` package (not vendored; line
// numbers are transformers/models/mimi/modeling_mimi.py of this image) as instantiated by Qwen3TTSTokenizerV2Encoder (modeling :898-907):
//   MimiEncoder :454-496, MimiResnetBlock :412-451, MimiConv1d :214-352, MimiTransformerLayer :926-993, down-sampling conv :1421-1431,
//   MimiSplitResidualVectorQuantizer.encode :1296-1350 (rvq_encode.cu).
//
// Design.  Activations are fp32 [B][L][C] in HBM.  Every conv / linear is the implicit-GEMM tensor-core kernel of conv_gemm.cu with
// split-precision operands (gemm_params.h, k_mult 3): A rows are fp16 [hi | lo] written by the SIMT kernel that applies the
// activation in front of the layer (encode_kernels.cu), weights are packed [hi | hi | lo] at load.  The accumulators are stored as
// fp32 with bias, layer scale and the fp32 residual fused into the epilogue.
//   * a k-tap stride-1 causal conv: taps at rows -(k-1) .. 0, rows < 0 read as zero by TMA (the reference's constant left padding);
//   * a stride-r, kernel-2r causal conv: the activation kernel regroups r input rows into one super-row [hi(r x C) | lo(r x C)], so
//     the conv is a two-tap GEMM over super-rows q-1 and q (left padding r = one zero super-row; the reference's right padding up to
//     a whole frame = the zero tail of the last super-row);
//   * the down-sampling conv (stride 2, kernel 4, pad_mode "replicate"): same, with the replicated rows materialised by the activation
//     kernel (one leading super-row) and taps at rows q and q+1.
#include <cmath>
#include <cstring>
#include <string>

#include "encode_kernels.h"
#include "engine.h"
#include "launch.cuh"

namespace q3 {

namespace {

struct EncFin {
  Engine* e;
  int rc = 0;
  cudaStream_t s = 0;

  const RawTensor* find(const std::string& key) const {
    auto it = e->raw.find(key);
    return it == e->raw.end() ? nullptr : &it->second;
  }
  const RawTensor* get(const std::string& key, std::initializer_list<int64_t> shape) {
    if (rc) return nullptr;
    const RawTensor* t = find(key);
    if (!t) { rc = e->fail(Q3TTS_BAD_KEY, "missing tensor: " + key); return nullptr; }
    if (t->shape != std::vector<int64_t>(shape)) { rc = e->fail(Q3TTS_BAD_KEY, "bad shape for " + key); return nullptr; }
    return t;
  }
  void* alloc(size_t bytes) {
    if (rc) return nullptr;
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { cudaGetLastError(); rc = e->fail(Q3TTS_OOM, "cudaMalloc failed while packing encoder weights"); return nullptr; }
    e->owned.push_back(p);
    return p;
  }
  void check(cudaError_t err, const char* what) {
    if (!rc && err != cudaSuccess) rc = e->fail(Q3TTS_CUDA_ERROR, std::string(what) + ": " + cudaGetErrorString(err));
  }
  // The tensor cores add into their fp32 accumulators with truncation, so the error of a dot product grows linearly with the
  // number of MMA instructions chained on one accumulator (measured: latents 7e-5 off with chains of up to 1536 instructions, 1.2e-5 with chains
  // of at most 96, against 3e-6 for the same operands summed with fmaf).  A layer whose K' = taps * 3 * kreal exceeds kMaxChain is therefore cut into K
  // groups of cg input values: one launch per group, each adding its partial sum to the fp32 output in the epilogue (a rounded add).
  static constexpr int kMaxChain = 2048;   // K' per launch = 128 MMA instructions of K = 16
  static int group_size(int kreal, int taps, int kb) {
    if (taps * 3 * kreal <= kMaxChain) return 0;
    int best = kb;
    for (int cg = kb; cg < kreal; cg += kb)
      if (kreal % cg == 0 && taps * 3 * cg <= kMaxChain) best = cg;
    return best;
  }
  // k_mult 3: K the MMAs walk = 3 x real input width, A rows are 2 x real width
  void finish(ConvW& w, int kreal) {
    w.k_mult = 3; w.cin = 3 * kreal; w.a_cols = 2 * kreal;
    w.bn = 0;
    for (int bn = 256; bn >= 16; bn -= 16) if (w.n_total % bn == 0) { w.bn = bn; break; }
    if (!rc && (w.bn == 0 || kreal % w.kb != 0 || w.n_total % 16 != 0)) rc = e->fail(Q3TTS_BAD_CONFIG, "encoder layer shape not tileable");
  }
  static int kblock(int kreal) { return (kreal % 64 == 0) ? 64 : 32; }
  ConvW conv(const std::string& prefix, int cout, int cin, int k, bool has_bias) {
    ConvW w;
    const RawTensor* t = get(prefix + ".weight", {cout, cin, k});
    const RawTensor* b = has_bias ? get(prefix + ".bias", {cout}) : nullptr;
    if (rc) return w;
    w.taps = k; w.n_total = cout; w.ncol_mod = cout;
    for (int j = 0; j < k; ++j) w.tap_off[j] = -(k - 1 - j);
    w.bias = b ? b->d : nullptr;
    w.kb = kblock(cin); w.cg = group_size(cin, k, w.kb);
    w.w = alloc((size_t)cout * cin * k * 2 * 3);
    if (!rc) check(pack_conv_weight(t->d, w.w, cout, cin, k, false, s, 3, w.cg), "pack_conv_weight");
    finish(w, cin);
    return w;
  }
  // stride-r, kernel-2r conv as two taps over super-rows; first_off = row offset of the first tap (-1: zero padding by TMA, 0: the
  // activation kernel wrote a leading super-row)
  ConvW strided(const std::string& prefix, int cout, int c, int r, bool has_bias, int first_off) {
    ConvW w;
    const RawTensor* t = get(prefix + ".weight", {cout, c, 2 * r});
    const RawTensor* b = has_bias ? get(prefix + ".bias", {cout}) : nullptr;
    if (rc) return w;
    w.taps = 2; w.n_total = cout; w.ncol_mod = cout;
    w.tap_off[0] = first_off; w.tap_off[1] = first_off + 1;
    w.bias = b ? b->d : nullptr;
    w.kb = kblock(r * c); w.cg = group_size(r * c, 2, w.kb);
    if (first_off < 0 && 2 * 3 * c <= kMaxChain && c % w.kb == 0) {
      // one K group per input row of the super-row: the operand is then simply every input row as its own [hi | lo] pair - what a
      // producing GEMM's epilogue can write directly (ACT_ELU_SPLIT) - and a launch walks as many groups as the chain limit allows
      w.cg = c; w.gl = kMaxChain / (2 * 3 * c);
    }
    w.w = alloc((size_t)2 * cout * r * c * 2 * 3);
    if (!rc) check(pack_strided_conv_weight(t->d, w.w, cout, c, r, false, s, 3, w.cg), "pack_strided_conv_weight");
    finish(w, r * c);
    return w;
  }
  // rows of up to three [n_i][k] matrices stacked (q | k | v)
  ConvW linear(std::initializer_list<std::string> keys, int n_each, int k) {
    ConvW w;
    const int n = n_each * (int)keys.size();
    w.taps = 1; w.n_total = n; w.ncol_mod = n;
    w.kb = kblock(k); w.cg = group_size(k, 1, w.kb);
    w.w = alloc((size_t)n * k * 2 * 3);
    int i = 0;
    for (const std::string& key : keys) {
      const RawTensor* t = get(key, {n_each, k});
      if (rc) return w;
      check(pack_rows(t->d, (char*)w.w + (size_t)i * n_each * k * 2 * 3, n_each, k, false, s, 3, w.cg), "pack_rows");
      ++i;
    }
    finish(w, k);
    return w;
  }
  const float* vec(const std::string& key, int n) {
    const RawTensor* t = get(key, {n});
    return t ? t->d : nullptr;
  }
};

struct EncBuilder {
  Engine* e;
  Plan* plan;
  uint8_t* base;
  size_t off = 0;
  bool dry;
  int rc = 0;

  void* alloc(size_t bytes) {
    off = (off + 1023) & ~(size_t)1023;
    void* p = dry ? nullptr : base + off;
    off += bytes;
    return p;
  }
  // A [batches][rows][2 kreal] fp16 -> fp32 out [batches][rows_valid][n] (+ bias, * scale, + fp32 residual).  A layer with K groups
  // (ConvW::cg) runs as one launch per `gl` groups: a launch reads A columns [2 g0 cg, 2 (g0 + ng) cg) and weight columns
  // [3 g0 cg, 3 (g0 + ng) cg) of the full-pitch rows, and from the second launch on adds to what the earlier ones stored.
  // pair_out: the (last) launch also / instead writes ELU(result) as fp16 [hi | lo] rows [batches][rows_valid][2 n] (ACT_ELU_SPLIT);
  // `out` may then be null when the fp32 result itself has no reader (single-launch layers only).
  void gemm(const ConvW& w, const void* in, int batches, int rows, int rows_valid, float* out, const float* res = nullptr, const float* scale = nullptr,
            void* pair_out = nullptr, int pair_act = ACT_ELU_SPLIT, int pair_group = 0) {
    const int kreal = w.a_cols / 2;
    const int cg = w.cg > 0 ? w.cg : kreal, G = kreal / cg, gl = w.gl > 0 ? w.gl : 1;
    for (int g0 = 0; g0 < G; g0 += gl) {
      const int ng = G - g0 < gl ? G - g0 : gl;
      const bool last = g0 + ng >= G;
      plan->ops.emplace_back();
      Op& op = plan->ops.back();
      op.type = OP_GEMM;
      GemmParams& p = op.g.p;
      memset(&p, 0, sizeof(p));
      p.batches = batches; p.rows = rows; p.rows_valid = rows_valid;
      p.cin = 3 * cg * ng; p.ntaps = w.taps;
      for (int i = 0; i < 8; ++i) p.tap_off[i] = w.tap_off[i];
      p.n_total = w.n_total; p.bn = w.bn; p.kb = w.kb; p.ncol_mod = w.ncol_mod;
      p.bias = g0 == 0 ? w.bias : nullptr; p.bf16 = 0;   // the encoder's operands are always fp16 pairs
      p.a_cols = 2 * cg * ng; p.k_mult = 3; p.kgroup = ng > 1 ? cg : 0;
      p.a_row_stride = 2 * kreal; p.w_row_stride = 3 * kreal;
      p.a_ptr = (const char*)in + (size_t)g0 * 2 * cg * 2; p.w_ptr = (const char*)w.w + (size_t)g0 * 3 * cg * 2;
      p.out_row_stride = w.n_total;
      p.out_batch_stride = (long long)rows_valid * w.n_total;
      p.out_f32 = out; p.scale = scale;
      const float* r = g0 == 0 ? res : out;
      if (r) { p.res = r; p.res_f32 = 1; }
      if (pair_out && last) {
        p.act = pair_act;
        p.out_raw = pair_out; p.out_snake = pair_out;   // one tensor; out_snake only selects the two-output epilogue paths
        p.pair_row_stride = 2 * w.n_total; p.pair_batch_stride = (long long)rows_valid * 2 * w.n_total;
        p.pair_group = pair_group > 0 && pair_group < w.n_total ? pair_group : 0;
      }
      if (!out && (G > gl || !pair_out)) rc = e->fail(Q3TTS_BAD_CONFIG, "encoder plan: a layer without an fp32 output must be a single launch with a pair output");
      op.g.retile = 1;   // few rows (the 25 Hz stages): shrink the N tile until the grid covers the SMs
    }
  }
  // `next`: the layer that reads the result (its K group size decides the layout)
  Op& act(const ConvW& next, const float* in, void* out, int B, int L, int C, int act_id, int r = 1, int lead = 0, int replicate = 0) {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_ENC_ACT; op.in = in; op.out = out; op.B = B; op.L = L; op.C = C; op.aux0 = act_id; op.aux1 = r; op.split = lead | (replicate << 8);
    op.kgroup = next.cg;
    return op;
  }
};

inline int cdiv(long long a, int b) { return (int)((a + b - 1) / b); }

int build(Engine* e, Plan* plan, int B, int T, bool dry, size_t* bytes) {
  EncBuilder b{e, plan, e->enc_ws, 0, dry};
  plan->B = B; plan->T = T; plan->ops.clear();
  const int C0 = e->enc_c0, H = e->enc_hidden, I = e->enc_inter;
  std::vector<size_t> xbytes;
  {
    int L = T, C = C0;
    xbytes.push_back((size_t)B * L * C * 4);
    for (const EncStageW& S : e->enc_stages) { L = cdiv(L, S.r); C *= 2; xbytes.push_back((size_t)B * L * C * 4); }
  }
  // The fp32 main-path tensors X0, X1, ... persist (~460 B per sample in all); a stage's operands live in a scratch region every stage
  // reuses (~512 B per sample) - except an operand that the PREVIOUS stage's strided conv writes from its epilogue while it still reads
  // its own scratch: that one is persistent too.
  const size_t ns_ = e->enc_stages.size();
  std::vector<char> pre(ns_ + 1, 0);   // pre[s]: stage s gets ELU(X_s) [hi | lo] from its producer's epilogue
  pre[0] = 1;                          // the first conv writes it
  for (size_t s = 1; s < ns_; ++s) pre[s] = e->enc_stages[s].c3.cg == 0 ? 1 : 0;
  std::vector<float*> Xs;
  std::vector<void*> A1pre(ns_ + 1, nullptr);
  {
    int L = T, C = C0;
    for (size_t s = 0; s <= ns_; ++s) {
      Xs.push_back((float*)b.alloc(xbytes[s]));
      if (s < ns_ && s >= 1 && pre[s]) A1pre[s] = b.alloc((size_t)B * L * 2 * C * 2);
      if (s < ns_) { L = cdiv(L, e->enc_stages[s].r); C *= 2; }
    }
  }
  const size_t scratch0 = b.off;
  size_t scratch_max = b.off;
  int L = T, C = C0;
  float* X = Xs[0];
  size_t si = 0;
  for (const EncStageW& S : e->enc_stages) {
    b.off = scratch0;
    const int r = S.r, Ln = cdiv(L, r);
    // per-row operand layout of the strided conv (one K group per input row): the k1 conv's epilogue can then write it directly
    const bool rowwise = S.down.cg == C && S.c1.cg == 0;
    const size_t a1 = (size_t)B * L * 2 * C * 2, a3 = (size_t)B * Ln * 2 * r * C * 2;
    void* A1 = A1pre[si] ? A1pre[si] : b.alloc(a1);
    void* A3 = b.alloc(a3);                                      // the strided conv's operand (super-rows of r input rows)
    const bool fused = S.c3.cg == 0 && S.c1.cg == 0 && (C / 2) % 32 == 0;   // ELU(conv_k3) leaves the epilogue as the k1 conv's operand
    float* Hh = fused ? nullptr : (float*)b.alloc((size_t)B * L * (C / 2) * 4);
    void* A2 = b.alloc((size_t)B * L * C * 2);
    if (si == 0) { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_CONV0; op.out = X; op.out2 = A1; op.B = B; op.L = L; op.C = C; }
    else if (!pre[si]) b.act(S.c3, X, A1, B, L, C, 1);
    if (fused) {
      b.gemm(S.c3, A1, B, L, L, nullptr, nullptr, nullptr, A2);
    } else {
      b.gemm(S.c3, A1, B, L, L, Hh);
      b.act(S.c1, Hh, A2, B, L, C / 2, 1);
    }
    if (rowwise && L % r == 0) {
      // x + conv_k1(...) is only ever read through the ELU in front of the strided conv: it leaves the k1 conv's epilogue as that
      // conv's operand (rows of [hi | lo] = super-rows of r such rows) and is never stored as fp32
      b.gemm(S.c1, A2, 1, B * L, B * L, nullptr, X, nullptr, A3);
    } else {
      b.gemm(S.c1, A2, 1, B * L, B * L, X, X);                  // x += conv_k1(...): residual read and written in place, fp32
      b.act(S.down, X, A3, B, L, C, 1, r);
    }
    float* Xn = Xs[si + 1];
    b.gemm(S.down, A3, B, Ln, Ln, Xn, nullptr, nullptr, si + 1 < ns_ && pre[si + 1] ? A1pre[si + 1] : nullptr);
    if (b.off > scratch_max) scratch_max = b.off;
    X = Xn; L = Ln; C = 2 * C; ++si;
  }
  b.off = scratch0;
  // ---- ELU, conv k3 -> Y [B][L][H]: the transformer's residual stream
  const int M = B * L;
  float* Y = (float*)b.alloc((size_t)M * H * 4);
  {
    void* A = b.alloc((size_t)M * 2 * C * 2);
    b.act(e->enc_last, X, A, B, L, C, 1);
    b.gemm(e->enc_last, A, B, L, L, Y);
  }
  void* Hn = b.alloc((size_t)M * 2 * H * 2);
  float* QKV = (float*)b.alloc((size_t)M * 3 * H * 4);
  void* AO = b.alloc((size_t)M * 2 * H * 2);
  void* QKVS = b.alloc((size_t)M * 6 * H * 2);               // q, k (RoPE applied), v as fp16 pairs: the attention kernel's operands
  const bool mlp_fused = !e->enc_layers.empty() && e->enc_layers[0].fc1.cg == 0;   // GELU(fc1) leaves the epilogue as fc2's operand
  float* F1 = mlp_fused ? nullptr : (float*)b.alloc((size_t)M * I * 4);
  void* G = b.alloc((size_t)M * 2 * I * 2);
  for (const EncLayerW& W : e->enc_layers) {
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_LN; op.in = Y; op.out = Hn; op.w0 = W.ln1w; op.w1 = W.ln1b; op.B = M; op.C = H; }
    b.gemm(W.qkv, Hn, 1, M, M, QKV);
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_ATTN; op.in = QKV; op.out = AO; op.out2 = QKVS; op.B = B; op.L = L; op.C = H; }
    b.gemm(W.o, AO, 1, M, M, Y, Y, W.ls1);
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_LN; op.in = Y; op.out = Hn; op.w0 = W.ln2w; op.w1 = W.ln2b; op.B = M; op.C = H; }
    if (mlp_fused) {
      b.gemm(W.fc1, Hn, 1, M, M, nullptr, nullptr, nullptr, G, ACT_GELU_SPLIT, W.fc2.cg);
    } else {
      b.gemm(W.fc1, Hn, 1, M, M, F1);
      b.act(W.fc2, F1, G, 1, M, I, 2);
    }
    b.gemm(W.fc2, G, 1, M, M, Y, Y, W.ls2);
  }
  // ---- down-sampling conv (stride 2, kernel 4, replicate padding, no bias) -> latents [B][F][H]
  const int F = cdiv(L, 2), MF = B * F;
  void* AD = b.alloc((size_t)B * (F + 1) * 4 * H * 2);
  float* LAT = (float*)b.alloc((size_t)MF * H * 4);
  b.act(e->enc_down, Y, AD, B, L, H, 0, 2, 1, 1);
  b.gemm(e->enc_down, AD, B, F + 1, F, LAT);
  { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_RVQ; op.in = LAT; op.B = B; op.L = F; op.C = H; op.tag = "latents"; }
  // ---- RVQ search (MimiSplitResidualVectorQuantizer.encode :1296-1350) on the tensor cores: per codebook one GEMM residual x codes
  // (fp16 hi + lo operands: K' = 3 * 256) and one SIMT pass that takes the argmin and updates the fp32 residual
  const int dq = e->cfg.codebook_dim / 2, NC = e->cfg.codebook_size, K = e->cfg.num_quantizers;
  void* AL = b.alloc((size_t)MF * 2 * H * 2);
  float* R = (float*)b.alloc((size_t)MF * 2 * dq * 4);          // [semantic residual | acoustic residual]
  void* AR = b.alloc((size_t)MF * 2 * dq * 2);
  float* SC = (float*)b.alloc((size_t)MF * NC * 4);
  b.act(e->enc_qin, LAT, AL, 1, MF, H, 0);
  b.gemm(e->enc_qin, AL, 1, MF, MF, R);
  for (int k = 0; k < K; ++k) {
    if (k <= 1) { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ENC_SPLITCOLS; op.in = R; op.out = AR; op.B = MF; op.C = dq; op.aux0 = k * dq; op.aux1 = 2 * dq; }
    b.gemm(e->enc_cb[k], AR, 1, MF, MF, SC);
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_ENC_ARGMIN; op.w0 = SC; op.w1 = R; op.out = (k == 0 || k == K - 1) ? nullptr : AR; op.B = B; op.L = F; op.C = NC; op.aux0 = k; op.aux1 = k == 0 ? 0 : dq;
  }
  if (b.off > scratch_max) scratch_max = b.off;
  plan->out_len = F;
  if (!dry) {
    for (Op& op : plan->ops)
      if (op.type == OP_GEMM && gemm_make_epilogue_maps(&op.g)) b.rc = e->fail(Q3TTS_CUDA_ERROR, "epilogue tensor map failed (encoder)");
  }
  if (bytes) *bytes = scratch_max + 1024;
  return b.rc;
}

int get_plan(Engine* e, int B, int T, Plan** out) {
  auto key = std::make_pair(B, T);
  auto it = e->enc_plans.find(key);
  if (it != e->enc_plans.end()) { *out = &it->second; return 0; }
  Plan tmp;
  size_t need = 0;
  build(e, &tmp, B, T, true, &need);
  if (need > e->enc_ws_bytes) {
    cudaDeviceSynchronize();
    if (e->enc_ws) cudaFree(e->enc_ws);
    e->enc_ws = nullptr; e->enc_ws_bytes = 0;
    e->enc_plans.clear();  // tensor maps point into the old workspace
    if (cudaMalloc((void**)&e->enc_ws, need) != cudaSuccess) { cudaGetLastError(); return e->fail(Q3TTS_OOM, "encoder workspace cudaMalloc failed"); }
    e->enc_ws_bytes = need;
  }
  if (e->enc_plans.size() >= 16) e->enc_plans.clear();  // ragged batches: every distinct (B, T) is a plan; keep the cache small
  Plan& p = e->enc_plans[key];
  const int rc = build(e, &p, B, T, false, nullptr);
  if (rc) { e->enc_plans.erase(key); return rc; }
  *out = &p;
  return 0;
}

// codes: [B][Q][F] (codes_tq = 0) or [B][F][Q]; latents (optional) [B][H][F]
int run(Engine* e, Plan* plan, const float* wav, long long wav_stride, long long* codes, int codes_tq, float* latents, cudaStream_t s) {
  size_t pi = e->profile ? prof_begin(e, plan->ops.size()) : 0;
  // programmatic dependent launch between consecutive tensor-core kernels (k3 -> k1 -> strided conv, the K-group launches): the next
  // kernel's prologue overlaps the previous one's tail; off while profiling (the event pairs would time the overlap)
  PdlScope pdl(!e->profile);
  bool searching = false;
  for (Op& op : plan->ops) {
    cudaError_t err = cudaSuccess;
    if (searching && !codes) break;   // latents only: the search is not run
    if (op.type == OP_ENC_RVQ && !latents) { searching = true; continue; }   // the parity tap is a no-op unless asked for
    if (e->profile) cudaEventRecord(e->prof_events[2 * pi], s);
    if (op.type == OP_ENC_RVQ) searching = true;
    switch (op.type) {
      case OP_ENC_CONV0: err = enc_conv0(wav, wav_stride, e->enc_conv0_w, e->enc_conv0_b, (float*)op.out, op.out2, op.B, op.L, op.C, s); break;
      case OP_ENC_ACT: err = enc_act_split((const float*)op.in, op.out, op.B, op.L, op.C, op.aux0, op.aux1, op.split & 0xff, op.split >> 8, op.kgroup, s); break;
      case OP_GEMM: err = launch_conv_gemm(op.g, e->engine, s); break;
      case OP_ENC_LN: err = enc_layernorm_split((const float*)op.in, op.w0, op.w1, op.out, op.B, op.C, e->enc_cfg.norm_eps, s); break;
      case OP_ENC_ATTN:
        if (e->engine == ENGINE_TC) {   // tensor-core form: rotate + split once, then mma.sync attention over fp16 pairs
          err = enc_rope_split((const float*)op.in, e->enc_rope_cos, e->enc_rope_sin, op.out2, (long long)op.B * op.L, op.L, e->enc_cfg.num_attention_heads, s);
          if (err == cudaSuccess) err = enc_attention_mma(op.out2, op.out, op.B, op.L, e->enc_cfg.num_attention_heads, 64, e->enc_cfg.sliding_window, s);
          ++e->launches;
        } else {                        // the cross-check engine (Q3TTS_ENGINE_SIMT): fp32 SIMT attention
          err = enc_attention((const float*)op.in, e->enc_rope_cos, e->enc_rope_sin, op.out, op.B, op.L, e->enc_cfg.num_attention_heads, 64,
                              e->enc_cfg.sliding_window, s);
        }
        break;
      case OP_ENC_RVQ:   // parity tap: the quantiser's input in the reference's layout
        if (latents) err = enc_to_channels_first((const float*)op.in, latents, op.B, op.L, op.C, s);
        break;
      case OP_ENC_SPLITCOLS: err = enc_split_cols((const float*)op.in, op.aux1, op.aux0, op.out, op.B, op.C, s); break;
      case OP_ENC_ARGMIN: {
        const int K = e->cfg.num_quantizers, dq = e->cfg.codebook_dim / 2, k = op.aux0;
        // codes [B][K][F] (stride between frames 1, between items K * F, offset k * F) or [B][F][K] (K, F * K, k)
        err = enc_argmin_update(op.w0, e->enc_norms + (size_t)k * op.C, e->enc_tables + (size_t)k * op.C * dq, const_cast<float*>(op.w1) + op.aux1, 2 * dq, op.out,
                                codes, codes_tq ? K : 1, (long long)K * op.L, codes_tq ? k : (long long)k * op.L, op.B * op.L, op.L, op.C, dq, s);
        break;
      }
      default: err = cudaErrorInvalidValue; break;
    }
    ++e->launches;
    if (e->profile) {
      cudaEventRecord(e->prof_events[2 * pi + 1], s);
      prof_account(e, op);
      ++pi;
    }
    if (err != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("encode launch failed: ") + cudaGetErrorString(err));
  }
  return Q3TTS_OK;
}

}  // namespace

int encoder_front_finalize(Engine* e) {
  const std::string P = "encoder.encoder.layers.";
  EncFin f{e};
  e->enc_front_ready = false;
  const RawTensor* w0 = f.find(P + "0.conv.weight");
  if (!w0) return Q3TTS_OK;   // no encoder in this state dict: decode-only handle
  if (w0->shape.size() != 3 || w0->shape[1] != 1 || w0->shape[2] != 7 || w0->shape[0] % 32) return e->fail(Q3TTS_BAD_KEY, "bad shape for " + P + "0.conv.weight");
  const q3tts_encoder_config& ec = e->enc_cfg;
  e->enc_c0 = (int)w0->shape[0];
  {  // own copy: engine_finalize frees the raw >= 2-D tensors once everything is packed
    float* w = (float*)f.alloc((size_t)w0->numel * sizeof(float));
    if (f.rc) return f.rc;
    f.check(cudaMemcpy(w, w0->d, (size_t)w0->numel * sizeof(float), cudaMemcpyDeviceToDevice), "conv0 weight copy");
    e->enc_conv0_w = w;
  }
  e->enc_conv0_b = f.vec(P + "0.conv.bias", e->enc_c0);
  e->enc_stages.clear(); e->enc_layers.clear();
  // stages: layers i (ResnetBlock), i + 1 (ELU), i + 2 (strided conv), i = 1, 4, 7, ...; then ELU and the last conv
  int C = e->enc_c0, i = 1;
  while (!f.rc && f.find(P + std::to_string(i) + ".block.1.conv.weight")) {
    EncStageW S;
    const RawTensor* dw = f.find(P + std::to_string(i + 2) + ".conv.weight");
    if (!dw || dw->shape.size() != 3 || dw->shape[0] != 2 * C || dw->shape[1] != C || dw->shape[2] % 2) return e->fail(Q3TTS_BAD_KEY, "bad strided conv after encoder block " + std::to_string(i));
    S.C = C; S.r = (int)dw->shape[2] / 2;
    S.c3 = f.conv(P + std::to_string(i) + ".block.1.conv", C / 2, C, 3, true);
    S.c1 = f.conv(P + std::to_string(i) + ".block.3.conv", C, C / 2, 1, true);
    S.down = f.strided(P + std::to_string(i + 2) + ".conv", 2 * C, C, S.r, true, -1);
    e->enc_stages.push_back(S);
    C *= 2; i += 3;
  }
  if (f.rc) return f.rc;
  const RawTensor* lw = f.find(P + std::to_string(i + 1) + ".conv.weight");
  if (!lw || lw->shape.size() != 3 || lw->shape[1] != C || lw->shape[2] != 3) return e->fail(Q3TTS_BAD_KEY, "missing or mis-shaped last encoder conv");
  const int H = (int)lw->shape[0];
  e->enc_hidden = H;
  if (H != e->cfg.codebook_dim) return e->fail(Q3TTS_BAD_CONFIG, "encoder hidden size must equal codebook_dim (the quantiser's input width)");
  if (H != ec.num_attention_heads * 64 || H % 128) return e->fail(Q3TTS_BAD_CONFIG, "encoder attention is specialised for head_dim 64");
  e->enc_last = f.conv(P + std::to_string(i + 1) + ".conv", H, C, 3, true);
  const std::string TP = "encoder.encoder_transformer.layers.";
  for (int l = 0; !f.rc && f.find(TP + std::to_string(l) + ".self_attn.q_proj.weight"); ++l) {
    const std::string p = TP + std::to_string(l) + ".";
    EncLayerW W;
    const RawTensor* fc1 = f.find(p + "mlp.fc1.weight");
    if (!fc1 || fc1->shape.size() != 2) return e->fail(Q3TTS_BAD_KEY, "missing " + p + "mlp.fc1.weight");
    const int I = (int)fc1->shape[0];
    if (l == 0) e->enc_inter = I;
    if (I != e->enc_inter) return e->fail(Q3TTS_BAD_CONFIG, "encoder layers differ in intermediate size");
    W.qkv = f.linear({p + "self_attn.q_proj.weight", p + "self_attn.k_proj.weight", p + "self_attn.v_proj.weight"}, H, H);
    W.o = f.linear({p + "self_attn.o_proj.weight"}, H, H);
    W.fc1 = f.linear({p + "mlp.fc1.weight"}, I, H);
    W.fc2 = f.linear({p + "mlp.fc2.weight"}, H, I);
    W.ln1w = f.vec(p + "input_layernorm.weight", H); W.ln1b = f.vec(p + "input_layernorm.bias", H);
    W.ln2w = f.vec(p + "post_attention_layernorm.weight", H); W.ln2b = f.vec(p + "post_attention_layernorm.bias", H);
    W.ls1 = f.vec(p + "self_attn_layer_scale.scale", H); W.ls2 = f.vec(p + "mlp_layer_scale.scale", H);
    e->enc_layers.push_back(W);
  }
  e->enc_down = f.strided("encoder.downsample.conv", H, H, 2, false, 0);
  if (f.rc) return f.rc;
  // quantiser as GEMM operands: the two input projections stacked (H -> 2 dq), and every codebook [codes][dq] (rvq_encode.cu packed
  // embed = embed_sum / clamp(cluster_usage) in fp32 just before this call)
  if (!e->enc_ready) return e->fail(Q3TTS_BAD_KEY, "the encoder front needs the encoder.quantizer.* tensors too");
  {
    const int dq = e->cfg.codebook_dim / 2, NC = e->cfg.codebook_size, K = e->cfg.num_quantizers;
    auto operand = [&](const float* src, int n, int kdim) {
      ConvW w;
      w.taps = 1; w.n_total = n; w.ncol_mod = n;
      w.kb = EncFin::kblock(kdim); w.cg = EncFin::group_size(kdim, 1, w.kb);
      w.w = f.alloc((size_t)n * kdim * 2 * 3);
      if (!f.rc) f.check(pack_rows(src, w.w, n, kdim, false, f.s, 3, w.cg), "pack_rows (quantiser)");
      f.finish(w, kdim);
      return w;
    };
    e->enc_qin = operand(e->enc_win, 2 * dq, H);
    e->enc_cb.clear();
    for (int k = 0; k < K && !f.rc; ++k) e->enc_cb.push_back(operand(e->enc_tables + (size_t)k * NC * dq, NC, dq));
    if (NC % 128 || dq % 4) return e->fail(Q3TTS_BAD_CONFIG, "quantiser shape not supported by the search kernels");
  }
  if (f.rc) return f.rc;
  // RoPE tables with the reference's float32 arithmetic: inv_freq = 1 / theta^(2i / 64) evaluated in float32 (:550-553)
  float inv[32];
  for (int k = 0; k < 32; ++k) inv[k] = 1.0f / (float)std::pow((double)ec.rope_theta, (double)((float)(2 * k) / 64.0f));
  float* d_inv = (float*)f.alloc(sizeof(inv));
  e->enc_rope_cos = (float*)f.alloc((size_t)ec.max_position_embeddings * 32 * 4);
  e->enc_rope_sin = (float*)f.alloc((size_t)ec.max_position_embeddings * 32 * 4);
  if (f.rc) return f.rc;
  f.check(cudaMemcpy(d_inv, inv, sizeof(inv), cudaMemcpyHostToDevice), "rope inv_freq copy");
  f.check(enc_rope_table(e->enc_rope_cos, e->enc_rope_sin, d_inv, ec.max_position_embeddings, 64, f.s), "enc_rope_table");
  f.check(enc_attention_setup(), "enc_attention_setup");
  f.check(cudaDeviceSynchronize(), "encoder weight packing");
  if (f.rc) return f.rc;
  e->enc_front_ready = true;
  return Q3TTS_OK;
}

void encoder_free(Engine* e) {
  if (e->enc_ws) cudaFree(e->enc_ws);
  if (e->enc_wav_stage) cudaFree(e->enc_wav_stage);
  if (e->enc_codes_stage) cudaFree(e->enc_codes_stage);
  e->enc_ws = nullptr; e->enc_wav_stage = nullptr; e->enc_codes_stage = nullptr;
  e->enc_plans.clear();
}

static long long enc_frames(const Engine& e, long long samples) {
  long long n = samples;
  for (const EncStageW& S : e.enc_stages) n = (n + S.r - 1) / S.r;
  return (n + 1) / 2;
}

// wav [B][wav_stride] -> codes ([B][Q][F] or [B][F][Q]) in sub-batches of at most enc_max_samples samples
static int encode_batched(Engine* e, const float* wav, int B, long long T, long long wav_stride, long long* codes, int codes_tq, float* latents, cudaStream_t s) {
  const long long F = enc_frames(*e, T);
  long long fr = T;
  for (const EncStageW& S : e->enc_stages) fr = (fr + S.r - 1) / S.r;
  if (fr > e->enc_cfg.max_position_embeddings) return e->fail(Q3TTS_BAD_SHAPE, "encode: item longer than max_position_embeddings frames at 25 Hz");
  long long per_ll = e->enc_max_samples / T;
  if (per_ll > 32768) per_ll = 32768;   // the first conv's grid carries the item index in its y dimension
  int per = per_ll < 1 ? 1 : (int)per_ll;
  const int Q = e->cfg.num_quantizers;
  for (int b0 = 0; b0 < B; b0 += per) {
    const int nb = B - b0 < per ? B - b0 : per;
    Plan* plan = nullptr;
    int rc = get_plan(e, nb, (int)T, &plan);
    if (rc) return rc;
    rc = run(e, plan, wav + (long long)b0 * wav_stride, wav_stride, codes ? codes + (long long)b0 * Q * F : nullptr, codes_tq,
             latents ? latents + (long long)b0 * e->enc_hidden * F : nullptr, s);
    if (rc) return rc;
  }
  return Q3TTS_OK;
}

}  // namespace q3

extern "C" {

int q3tts_set_encoder_config(q3tts_handle* h, const q3tts_encoder_config* cfg) {
  if (!h || !cfg) return Q3TTS_BAD_CONFIG;
  q3::Engine& e = h->e;
  if (e.finalized) return e.fail(Q3TTS_BAD_CONFIG, "set_encoder_config: call before q3tts_finalize_weights");
  if (cfg->num_attention_heads < 1 || cfg->sliding_window < 1 || cfg->max_position_embeddings < 1 || !(cfg->rope_theta > 0.f) || !(cfg->norm_eps > 0.f))
    return e.fail(Q3TTS_BAD_CONFIG, "set_encoder_config: values out of range");
  e.enc_cfg = *cfg;
  return Q3TTS_OK;
}

int64_t q3tts_encode_num_frames(q3tts_handle* h, int64_t samples) {
  if (!h || samples < 0) return Q3TTS_BAD_CONFIG;
  if (!h->e.enc_front_ready) return h->e.fail(Q3TTS_NOT_READY, "encode: the encoder weights (encoder.*) are not loaded");
  return q3::enc_frames(h->e, samples);
}

int q3tts_encode(q3tts_handle* h, const float* wav, int B, int64_t T, int64_t wav_stride, int64_t* codes, float* latents, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  q3::Engine& e = h->e;
  if (!e.finalized || !e.enc_front_ready || !e.enc_ready)
    return e.fail(Q3TTS_NOT_READY, "encode: the encoder weights (encoder.encoder.*, encoder.encoder_transformer.*, encoder.downsample.*, encoder.quantizer.*) are not loaded");
  if (!wav || (!codes && !latents) || B < 1 || T < 1 || T > 0x7fffffffll / 64 || wav_stride < T) return e.fail(Q3TTS_BAD_SHAPE, "encode: wav must be [B >= 1, stride >= T >= 1]");
  q3::DeviceGuard g(e.device);
  return q3::encode_batched(&e, wav, B, T, wav_stride, (long long*)codes, 0, latents, (cudaStream_t)stream);
}

int q3tts_encode_host(q3tts_handle* h, const float* wav_host, int B, int64_t T, int64_t wav_stride, const int64_t* valid_samples, int64_t* codes_host,
                      int64_t* lengths, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  q3::Engine& e = h->e;
  if (!e.finalized || !e.enc_front_ready || !e.enc_ready) return e.fail(Q3TTS_NOT_READY, "encode_host: the encoder weights (encoder.*) are not loaded");
  if (!wav_host || !codes_host || B < 1 || T < 1 || T > 0x7fffffffll / 64 || wav_stride < T) return e.fail(Q3TTS_BAD_SHAPE, "encode_host: wav must be [B >= 1, stride >= T >= 1]");
  if (valid_samples)
    for (int b = 0; b < B; ++b)
      if (valid_samples[b] < 0 || valid_samples[b] > T) return e.fail(Q3TTS_BAD_SHAPE, "encode_host: valid_samples[b] must be in [0, T]");
  q3::DeviceGuard g(e.device);
  cudaStream_t s = (cudaStream_t)stream;
  const long long F = q3::enc_frames(e, T);
  const int Q = e.cfg.num_quantizers;
  const size_t nw = (size_t)B * T, nc = (size_t)B * F * Q;
  // Page-locked waveform buffer: the first conv reads it in place over PCIe (unified addressing gives the device an alias of the host
  // pointer; the kernel stages every sample into shared memory once, so each is fetched ~1.05 times) - no separate host-to-device copy.
  const float* wav_dev = nullptr;
  long long wav_dev_stride = T;
  {
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, wav_host) == cudaSuccess && at.type == cudaMemoryTypeHost && at.devicePointer) {
      wav_dev = (const float*)at.devicePointer; wav_dev_stride = wav_stride;
    } else {
      cudaGetLastError();
    }
  }
  if (!wav_dev && nw > e.enc_wav_stage_elems) {
    cudaStreamSynchronize(s);
    if (e.enc_wav_stage) cudaFree(e.enc_wav_stage);
    e.enc_wav_stage = nullptr; e.enc_wav_stage_elems = 0;
    if (cudaMalloc((void**)&e.enc_wav_stage, nw * sizeof(float)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "encode_host: staging cudaMalloc failed"); }
    e.enc_wav_stage_elems = nw;
  }
  if (nc > e.enc_codes_stage_elems) {
    cudaStreamSynchronize(s);
    if (e.enc_codes_stage) cudaFree(e.enc_codes_stage);
    e.enc_codes_stage = nullptr; e.enc_codes_stage_elems = 0;
    if (cudaMalloc((void**)&e.enc_codes_stage, nc * sizeof(long long)) != cudaSuccess) { cudaGetLastError(); return e.fail(Q3TTS_OOM, "encode_host: staging cudaMalloc failed"); }
    e.enc_codes_stage_elems = nc;
  }
  cudaError_t err = cudaSuccess;
  if (!wav_dev) {
    err = cudaMemcpy2DAsync(e.enc_wav_stage, (size_t)T * sizeof(float), wav_host, (size_t)wav_stride * sizeof(float), (size_t)T * sizeof(float), B,
                            cudaMemcpyHostToDevice, s);
    if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("encode_host: H2D failed: ") + cudaGetErrorString(err));
    wav_dev = e.enc_wav_stage;
  }
  const int rc = q3::encode_batched(&e, wav_dev, B, T, wav_dev_stride, e.enc_codes_stage, 1, nullptr, s);
  if (rc) return rc;
  err = cudaMemcpyAsync(codes_host, e.enc_codes_stage, nc * sizeof(long long), cudaMemcpyDeviceToHost, s);
  if (err == cudaSuccess) err = cudaStreamSynchronize(s);
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("encode_host: ") + cudaGetErrorString(err));
  if (lengths)
    for (int b = 0; b < B; ++b) {
      const long long n = valid_samples ? valid_samples[b] : T;
      lengths[b] = (n + e.total_up - 1) / e.total_up;   // ceil(mask.sum() / encode_downsample_rate), modeling :983
    }
  return Q3TTS_OK;
}

}  // extern "C"
