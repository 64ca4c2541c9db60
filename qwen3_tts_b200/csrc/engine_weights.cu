// Load-time repacking: reference state_dict tensors (fp32) -> tensor-core operand layouts.
#include <cstdio>

#include "engine.h"
#include "kernels.h"

namespace q3 {

namespace {

struct Fin {
  Engine* e;
  int rc = 0;
  cudaStream_t s = 0;

  const RawTensor* get(const std::string& key, std::initializer_list<int64_t> shape) {
    if (rc) return nullptr;
    auto it = e->raw.find(key);
    if (it == e->raw.end()) { rc = e->fail(Q3TTS_BAD_KEY, "missing tensor: " + key); return nullptr; }
    std::vector<int64_t> want(shape);
    if (it->second.shape != want) {
      std::string got;
      for (auto v : it->second.shape) got += std::to_string(v) + ",";
      rc = e->fail(Q3TTS_BAD_KEY, "bad shape for " + key + ": got [" + got + "]");
      return nullptr;
    }
    return &it->second;
  }
  void* alloc(size_t bytes) {
    if (rc) return nullptr;
    void* p = nullptr;
    if (cudaMalloc(&p, bytes) != cudaSuccess) { rc = e->fail(Q3TTS_OOM, "cudaMalloc failed while packing weights"); return nullptr; }
    e->owned.push_back(p);
    return p;
  }
  void check(cudaError_t err, const char* what) {
    if (!rc && err != cudaSuccess) rc = e->fail(Q3TTS_CUDA_ERROR, std::string(what) + ": " + cudaGetErrorString(err));
  }
  // N tile: the epilogue moves 16-bit tensors in 32-column boxes (64 accumulator columns for the GLU epilogue)
  static int choose_bn(int n, int step) {
    for (int bn = 256 / step * step; bn >= step; bn -= step)
      if (n % bn == 0) return bn;
    return 0;
  }
  // k_mult / a_cols of a layer with `c` input channels for split factor sp (1 plain, 2 weights only, 3 weights and activations)
  static void split_fields(ConvW& w, int c, int sp) {
    w.k_mult = sp; w.cin = sp * c; w.a_cols = sp == 1 ? 0 : (sp == 2 ? c : 2 * c);
  }
  void tiles(ConvW& w, int step = 32) {
    w.bn = choose_bn(w.n_total, step);
    w.kb = ((w.cin / w.k_mult) % 64 == 0) ? 64 : 32;  // a K block never straddles the hi / lo parts
    if (!rc && (w.bn == 0 || (w.cin / w.k_mult) % w.kb != 0 || w.ncol_mod % 16 != 0))
      rc = e->fail(Q3TTS_BAD_CONFIG, "layer shape not tileable: n=" + std::to_string(w.n_total) + " cin=" + std::to_string(w.cin));
  }
  // Conv1d [cout][cin][k], dilation d, causal
  ConvW conv(const std::string& prefix, int cout, int cin, int k, int dil, bool has_bias = true, int sp = 1) {
    ConvW w;
    const RawTensor* t = get(prefix + ".weight", {cout, cin, k});
    const RawTensor* b = has_bias ? get(prefix + ".bias", {cout}) : nullptr;
    if (rc) return w;
    w.taps = k; w.n_total = cout; w.ncol_mod = cout;
    split_fields(w, cin, sp);
    for (int j = 0; j < k; ++j) w.tap_off[j] = -(k - 1 - j) * dil;
    w.bias = b ? b->d : nullptr;
    w.w = alloc((size_t)cout * cin * k * 2 * sp);
    if (!rc) check(pack_conv_weight(t->d, w.w, cout, cin, k, e->bf16, s, sp), "pack_conv_weight");
    tiles(w);
    return w;
  }
  // Linear [n][k]
  ConvW linear(const std::string& prefix, int n, int k, bool has_bias, int sp = 1) {
    ConvW w;
    const RawTensor* t = get(prefix + ".weight", {n, k});
    const RawTensor* b = has_bias ? get(prefix + ".bias", {n}) : nullptr;
    if (rc) return w;
    w.taps = 1; w.n_total = n; w.ncol_mod = n;
    split_fields(w, k, sp);
    w.bias = b ? b->d : nullptr;
    w.w = alloc((size_t)n * k * 2 * sp);
    if (!rc) check(pack_rows(t->d, w.w, n, k, e->bf16, s, sp), "pack_rows");
    tiles(w);
    return w;
  }
  // ConvTranspose1d [cin][cout][k], stride st (k == 2*st or k == st)
  ConvW tconv(const std::string& prefix, int cin, int cout, int k, int st, int sp = 1) {
    ConvW w;
    const RawTensor* t = get(prefix + ".weight", {cin, cout, k});
    const RawTensor* b = get(prefix + ".bias", {cout});
    if (rc) return w;
    if (k != st && k != 2 * st) { rc = e->fail(Q3TTS_BAD_CONFIG, "transposed conv needs k == stride or k == 2*stride"); return w; }
    w.taps = k / st; w.n_total = st * cout; w.ncol_mod = cout;
    split_fields(w, cin, sp);
    w.tap_off[0] = 0; w.tap_off[1] = 1;
    w.bias = b->d;
    w.out_rows_per_row = st;
    w.w = alloc((size_t)cin * cout * k * 2 * sp);
    if (!rc) check(pack_transconv_weight(t->d, w.w, cin, cout, k, st, e->bf16, s, sp), "pack_transconv_weight");
    tiles(w);
    return w;
  }
  SnakeP snake(const std::string& prefix, int c) {
    SnakeP p;
    const RawTensor* a = get(prefix + ".alpha", {c});
    const RawTensor* b = get(prefix + ".beta", {c});
    if (rc) return p;
    float* buf = (float*)alloc((size_t)2 * c * sizeof(float));
    if (rc) return p;
    check(pack_snake(a->d, b->d, buf, buf + c, c, s), "pack_snake");
    p.a = buf; p.ib = buf + c;
    return p;
  }
  const float* vec(const std::string& key, int n) {
    const RawTensor* t = get(key, {n});
    return t ? t->d : nullptr;
  }
};

}  // namespace

int engine_finalize(Engine* e, int mode) {
  if (mode < 0 || mode > Q3TTS_MODE_BF16_SPLIT_ALL) return e->fail(Q3TTS_BAD_CONFIG, "unknown compute mode");
  const int variant = mode >> 1;  // 0 default (main path split), 1 plain, 2 every early-stage GEMM split
  e->split_level = variant == 0 ? 1 : (variant == 1 ? 0 : 2);
  const int sp_main = e->split_level >= 1 ? 3 : 1;   // weights + activations hi/lo on the main signal path
  const int sp_xf_a = e->split_level >= 2 ? 3 : 1;   // transformer GEMMs fed by an RMSNorm (its output can be written hi/lo)
  const int sp_xf_w = e->split_level >= 2 ? 2 : 1;   // GEMMs whose A comes from a tensor-core epilogue in 16 bit: weights only
  const q3tts_config& c = e->cfg;
  if (c.head_dim != 64) return e->fail(Q3TTS_BAD_CONFIG, "attention kernel is specialised for head_dim 64");
  if (c.num_key_value_heads != c.num_attention_heads) return e->fail(Q3TTS_BAD_CONFIG, "GQA (kv heads != heads) not supported");
  if (c.latent_dim != 1024 || c.hidden_size != c.num_attention_heads * c.head_dim)
    return e->fail(Q3TTS_BAD_CONFIG, "latent_dim must be 1024 and hidden_size = heads*head_dim");
  if (c.codebook_dim % 4 != 0 || c.codebook_dim > 4096) return e->fail(Q3TTS_BAD_CONFIG, "codebook_dim must be a multiple of 4");
  for (void* p : e->owned) cudaFree(p);
  e->owned.clear();
  e->plans.clear();
  e->finalized = false;
  e->bf16 = (mode & 1) != 0;
  // kernel attributes (shared-memory opt-ins) are per device: set them for THIS handle's device every time, so a second handle on
  // another GPU of the same process launches with the same limits as the first
  if (gemm_setup(e->device)) return e->fail(Q3TTS_CUDA_ERROR, "gemm_setup failed (cuTensorMapEncodeTiled unavailable or smem opt-in rejected)");
  if (unit_setup() || tail_setup() || attention_setup() != cudaSuccess || stream_attention_setup() != cudaSuccess || static_scratch_setup(e->device)) {
    cudaGetLastError();
    return e->fail(Q3TTS_CUDA_ERROR, "per-device kernel set-up failed (shared-memory opt-in or static scratch address)");
  }

  if (!e->d_bad_code) {
    if (cudaMalloc((void**)&e->d_bad_code, sizeof(int)) != cudaSuccess || cudaMemset(e->d_bad_code, 0, sizeof(int)) != cudaSuccess ||
        cudaMallocHost((void**)&e->h_bad_code, sizeof(int)) != cudaSuccess) {
      cudaGetLastError();
      return e->fail(Q3TTS_OOM, "finalize: code-check flag allocation failed");
    }
    *e->h_bad_code = 0;
  }
  Fin f{e};
  const int H = c.hidden_size, Lat = c.latent_dim, I = c.intermediate_size, D = c.codebook_dim, dq = D / 2, Q = c.num_quantizers;
  // ---- RVQ: fold output_proj into the codebooks (fp32)
  e->rvq_tables = (float*)f.alloc((size_t)Q * c.codebook_size * D * sizeof(float));
  for (int k = 0; k < Q && !f.rc; ++k) {
    const std::string name = k == 0 ? "rvq_first" : "rvq_rest";
    const int li = k == 0 ? 0 : k - 1;
    const std::string cb = "quantizer." + name + ".vq.layers." + std::to_string(li) + "._codebook.";
    const RawTensor* es = f.get(cb + "embedding_sum", {c.codebook_size, dq});
    const RawTensor* cu = f.get(cb + "cluster_usage", {c.codebook_size});
    const RawTensor* wo = f.get("quantizer." + name + ".output_proj.weight", {D, dq, 1});
    if (f.rc) break;
    f.check(fold_codebook(es->d, cu->d, wo->d, e->rvq_tables + (size_t)k * c.codebook_size * D, c.codebook_size, dq, D, f.s), "fold_codebook");
  }
  // ---- pre_conv + transformer
  e->pre_conv = f.conv("pre_conv.conv", Lat, D, 3, 1, true, sp_main);
  e->in_proj = f.linear("pre_transformer.input_proj", H, Lat, true, sp_main);
  e->out_proj = f.linear("pre_transformer.output_proj", Lat, H, true, sp_main);
  e->final_norm = f.vec("pre_transformer.norm.weight", H);
  e->layers.assign(c.num_hidden_layers, LayerW());
  for (int l = 0; l < c.num_hidden_layers && !f.rc; ++l) {
    const std::string p = "pre_transformer.layers." + std::to_string(l) + ".";
    LayerW& L = e->layers[l];
    const RawTensor* q = f.get(p + "self_attn.q_proj.weight", {H, H});
    const RawTensor* k = f.get(p + "self_attn.k_proj.weight", {H, H});
    const RawTensor* v = f.get(p + "self_attn.v_proj.weight", {H, H});
    if (f.rc) break;
    L.qkv.taps = 1; L.qkv.n_total = 3 * H; L.qkv.ncol_mod = 3 * H;
    Fin::split_fields(L.qkv, H, sp_xf_a);
    const size_t part = (size_t)H * H * 2 * sp_xf_a;
    L.qkv.w = f.alloc(3 * part);
    if (f.rc) break;
    f.check(pack_rows(q->d, L.qkv.w, H, H, e->bf16, f.s, sp_xf_a), "pack q");
    f.check(pack_rows(k->d, (uint8_t*)L.qkv.w + part, H, H, e->bf16, f.s, sp_xf_a), "pack k");
    f.check(pack_rows(v->d, (uint8_t*)L.qkv.w + 2 * part, H, H, e->bf16, f.s, sp_xf_a), "pack v");
    f.tiles(L.qkv);
    L.o = f.linear(p + "self_attn.o_proj", H, H, false, sp_xf_w);
    const RawTensor* g = f.get(p + "mlp.gate_proj.weight", {I, H});
    const RawTensor* u = f.get(p + "mlp.up_proj.weight", {I, H});
    if (f.rc) break;
    L.gateup.taps = 1; L.gateup.n_total = 2 * I; L.gateup.ncol_mod = 2 * I;
    Fin::split_fields(L.gateup, H, sp_xf_a);
    L.gateup.w = f.alloc((size_t)2 * I * H * 2 * sp_xf_a);
    if (f.rc) break;
    f.check(pack_interleave2(g->d, u->d, L.gateup.w, I, H, e->bf16, f.s, sp_xf_a), "pack gate/up");
    f.tiles(L.gateup, 64);
    L.down = f.linear(p + "mlp.down_proj", H, I, false, sp_xf_w);
    L.ln1 = f.vec(p + "input_layernorm.weight", H);
    L.ln2 = f.vec(p + "post_attention_layernorm.weight", H);
    L.ls1 = f.vec(p + "self_attn_layer_scale.scale", H);
    L.ls2 = f.vec(p + "mlp_layer_scale.scale", H);
  }
  // ---- upsample stages
  e->ups.assign(c.num_upsampling_ratios, UpsampleW());
  for (int i = 0; i < c.num_upsampling_ratios && !f.rc; ++i) {
    const std::string p = "upsample." + std::to_string(i) + ".";
    const int r = c.upsampling_ratios[i];
    UpsampleW& U = e->ups[i];
    U.tconv = f.tconv(p + "0.conv", Lat, Lat, r, r, sp_main);
    const RawTensor* dw = f.get(p + "1.dwconv.conv.weight", {Lat, 1, 7});
    if (f.rc) break;
    float* dww = (float*)f.alloc((size_t)7 * Lat * sizeof(float));
    if (f.rc) break;
    f.check(pack_dw_weight(dw->d, dww, Lat, 7, f.s), "pack_dw_weight");
    U.dww = dww;
    U.dwb = f.vec(p + "1.dwconv.conv.bias", Lat);
    U.lnw = f.vec(p + "1.norm.weight", Lat);
    U.lnb = f.vec(p + "1.norm.bias", Lat);
    U.gamma = f.vec(p + "1.gamma", Lat);
    U.pw1 = f.linear(p + "1.pwconv1", 4 * Lat, Lat, true, sp_xf_w);
    U.pw2 = f.linear(p + "1.pwconv2", Lat, 4 * Lat, true, sp_xf_w);
  }
  // ---- vocoder
  const int Dd = c.decoder_dim;
  // decoder[0] is the one big GEMM of the early stages (44 MMAC/frame): 3x its MMAs cost 1.0 ms per 64 x 10 s step for ~4 % of the noise,
  // so it takes split operands only at level 2; at level 1 it reads the fp32 main-path tensor rounded once to 16 bit
  e->dec0 = f.conv("decoder.0.conv", Dd, Lat, 7, 1, true, e->split_level >= 2 ? 3 : 1);
  e->blocks.assign(c.num_upsample_rates, BlockW());
  int total_up = 1, look = 0;
  for (int i = 0; i < c.num_upsampling_ratios; ++i) total_up *= c.upsampling_ratios[i];
  for (int i = 0; i < c.num_upsample_rates && !f.rc; ++i) {
    BlockW& Bk = e->blocks[i];
    Bk.cin = Dd >> i; Bk.cout = Dd >> (i + 1); Bk.stride = c.upsample_rates[i];
    const std::string p = "decoder." + std::to_string(i + 1) + ".block.";
    Bk.act0 = f.snake(p + "0", Bk.cin);
    Bk.tconv = f.tconv(p + "1.conv", Bk.cin, Bk.cout, 2 * Bk.stride, Bk.stride);
    const int dil[3] = {1, 3, 9};
    for (int j = 0; j < 3 && !f.rc; ++j) {
      const std::string u = p + std::to_string(j + 2) + ".";
      Bk.unit[j].act1 = f.snake(u + "act1", Bk.cout);
      Bk.unit[j].conv7 = f.conv(u + "conv1.conv", Bk.cout, Bk.cout, 7, dil[j]);
      Bk.unit[j].act2 = f.snake(u + "act2", Bk.cout);
      Bk.unit[j].conv1 = f.conv(u + "conv2.conv", Bk.cout, Bk.cout, 1, 1);
    }
    total_up *= Bk.stride;
    look = look * Bk.stride + Bk.stride;  // each k=2s transposed conv drops s samples at its own rate
  }
  e->total_up = total_up;
  e->lookahead = look;
  const int n = c.num_upsample_rates;
  const int Cl = Dd >> n;
  e->tail_c = Cl;
  e->tail_act = f.snake("decoder." + std::to_string(n + 1), Cl);
  {
    const RawTensor* tw = f.get("decoder." + std::to_string(n + 2) + ".conv.weight", {1, Cl, 7});
    const RawTensor* tb = f.get("decoder." + std::to_string(n + 2) + ".conv.bias", {1});
    if (!f.rc) {
      e->tail_w = (float*)f.alloc((size_t)7 * Cl * sizeof(float));
      if (!f.rc) f.check(pack_dw_weight(tw->d, e->tail_w, Cl, 7, f.s), "pack tail");
      if (!f.rc) f.check(cudaMemcpy(&e->tail_bias, tb->d, sizeof(float), cudaMemcpyDeviceToHost), "tail bias");
      // tensor-core form: N padded 1 -> 16 (smallest UMMA N at M = 128); only output column 0 is stored
      ConvW& t = e->tail;
      t.taps = 7; t.n_total = 16; t.cin = Cl; t.ncol_mod = 16; t.bn = 16; t.kb = (Cl % 64 == 0) ? 64 : 32;
      for (int j = 0; j < 7; ++j) t.tap_off[j] = -(6 - j);
      void* tmp = f.alloc((size_t)7 * Cl * 2);
      t.w = f.alloc((size_t)7 * 16 * Cl * 2);
      e->tail_w16 = f.alloc((size_t)16 * Cl * 2);
      float* tbias = (float*)f.alloc(16 * sizeof(float));
      if (!f.rc) {
        f.check(cudaMemset(t.w, 0, (size_t)7 * 16 * Cl * 2), "tail memset");
        f.check(cudaMemset(tbias, 0, 16 * sizeof(float)), "tail bias memset");
        f.check(pack_conv_weight(tw->d, tmp, 1, Cl, 7, e->bf16, f.s), "pack tail gemm");  // [7 taps][1][Cl]
        f.check(cudaMemset(e->tail_w16, 0, (size_t)16 * Cl * 2), "tail w16 memset");
        f.check(cudaMemcpy(e->tail_w16, tmp, (size_t)7 * Cl * 2, cudaMemcpyDeviceToDevice), "tail w16 copy");
        f.check(cudaMemcpy2D(t.w, (size_t)16 * Cl * 2, tmp, (size_t)Cl * 2, (size_t)Cl * 2, 7, cudaMemcpyDeviceToDevice), "tail scatter");
        f.check(cudaMemcpy(tbias, tb->d, sizeof(float), cudaMemcpyDeviceToDevice), "tail bias copy");
        t.bias = tbias;
        if (Cl % t.kb) f.rc = e->fail(Q3TTS_BAD_CONFIG, "tail channels not tileable");
      }
    }
  }
  // ---- RoPE table
  if (!f.rc) {
    e->rope_max = c.max_position_embeddings;
    e->rope_cos = (float*)f.alloc((size_t)e->rope_max * 32 * sizeof(float));
    e->rope_sin = (float*)f.alloc((size_t)e->rope_max * 32 * sizeof(float));
    if (!f.rc) f.check(rope_table(e->rope_cos, e->rope_sin, e->rope_max, c.head_dim, c.rope_theta, f.s), "rope_table");
  }
  if (!f.rc) f.check(cudaDeviceSynchronize(), "finalize sync");
  if (f.rc) return f.rc;
  if (int rc = encode_finalize(e)) return rc;
  if (int rc = encoder_front_finalize(e)) return rc;
  // big fp32 sources are no longer needed; keep vectors (biases, norms, scales) that kernels read directly
  for (auto it = e->raw.begin(); it != e->raw.end();) {
    if (it->second.shape.size() >= 2) {
      cudaFree(it->second.d);
      it = e->raw.erase(it);
    } else {
      ++it;
    }
  }
  e->finalized = true;
  return Q3TTS_OK;
}

}  // namespace q3
