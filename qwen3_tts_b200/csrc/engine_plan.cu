// Per-(B,T) launch plans for Qwen3TTSTokenizerV2Decoder.forward (reference modeling :868-883).
#include <cstdlib>
#include <cstring>

#include "engine.h"
#include "kernels.h"

namespace q3 {

namespace {

struct Builder {
  Engine* e;
  Plan* plan;
  uint8_t* base;
  size_t off = 0;
  bool dry;  // size pass only
  int rc = 0;

  void* alloc(size_t bytes) {
    off = (off + 1023) & ~(size_t)1023;
    void* p = dry ? nullptr : base + off;
    off += bytes;
    return p;
  }
  // GEMM: in [batches][rows][cin] -> accumulator rows_valid per batch item
  Op& gemm(const ConvW& w, const void* in, int batches, int rows, int rows_valid) {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_GEMM;
    GemmParams& p = op.g.p;
    memset(&p, 0, sizeof(p));
    p.batches = batches; p.rows = rows; p.rows_valid = rows_valid;
    p.cin = w.cin; p.ntaps = w.taps;
    for (int i = 0; i < 8; ++i) p.tap_off[i] = w.tap_off[i];
    p.n_total = w.n_total; p.bn = w.bn; p.kb = w.kb; p.ncol_mod = w.ncol_mod;
    p.bias = w.bias; p.bf16 = e->bf16 ? 1 : 0;
    p.a_cols = w.a_cols; p.k_mult = w.k_mult;
    { const char* d = Q3_ENV("Q3TTS_DBG"); p.dbg = d ? atoi(d) : 0; }
    p.a_ptr = in; p.w_ptr = w.w;
    p.out_row_stride = w.n_total;
    p.out_batch_stride = (long long)rows_valid * w.n_total;
    if (!dry) {
      if (make_tmap_3d(&op.g.tmA, in, (uint64_t)(w.a_cols ? w.a_cols : w.cin), (uint64_t)rows, (uint64_t)batches, (uint32_t)w.kb, 128, e->bf16) ||
          make_tmap_3d(&op.g.tmB, w.w, (uint64_t)w.cin, (uint64_t)w.n_total, (uint64_t)w.taps, (uint32_t)w.kb, (uint32_t)w.bn, e->bf16))
        rc = e->fail(Q3TTS_CUDA_ERROR, "cuTensorMapEncodeTiled failed");
    }
    return op;
  }
  void tag(Op& op, const char* name, const void* ptr, bool f32, int C, int L, int ld = 0) {
    op.tag = name; op.probe_ptr = ptr; op.probe_f32 = f32; op.probe_C = C; op.probe_L = L; op.probe_ld = ld;
  }
  // fp32 [batches][rows][C] -> 16-bit rows [hi | lo]: the split-precision A operand of the next GEMM
  Op& split(const float* in, void* out, int batches, int rows, int C, bool with_lo = true) {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_SPLIT; op.in = in; op.out = out; op.B = batches; op.L = rows; op.C = C; op.split = with_lo ? 1 : 0;
    op.sp_in_bs = (long long)rows * C; op.sp_out_bs = (long long)rows * (with_lo ? 2 : 1) * C;
    return op;
  }
};

int build(Engine* e, Plan* plan, int B, int T, bool dry, size_t* bytes) {
  const q3tts_config& c = e->cfg;
  Builder b{e, plan, e->ws, 0, dry};
  plan->B = B; plan->T = T; plan->ops.clear();
  const int H = c.hidden_size, Lat = c.latent_dim, I = c.intermediate_size, D = c.codebook_dim;
  const int M = B * T;
  const size_t es = 2;  // 16-bit activations

  // Split-precision plans (e->split_level >= 1, DESIGN.md "Precision"): a layer packed with k_mult 3 reads A rows [hi | lo]; such
  // tensors are written by the SIMT kernels (RVQ, RMSNorm) or made from an fp32 main-path tensor by OP_SPLIT.
  const bool sp_main = e->pre_conv.k_mult == 3;
  const bool sp_xf = !e->layers.empty() && e->layers[0].qkv.k_mult == 3;
  // ---- RVQ -> E [B,T,D] (rows [hi | lo] when pre_conv takes split operands)
  void* E = b.alloc((size_t)M * D * es * (sp_main ? 2 : 1));
  {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_RVQ; op.out = E; op.B = B; op.L = T; op.C = D; op.split = sp_main ? 1 : 0;
    b.tag(op, "quantizer", E, false, D, T, sp_main ? 2 * D : 0);
  }
  // ---- pre_conv (k=3 causal) -> P [B,T,Lat]: 16-bit, or fp32 + its [hi | lo] copy for in_proj
  void* P = b.alloc((size_t)M * Lat * (sp_main ? 4 : es));
  void* P16 = sp_main ? b.alloc((size_t)M * 2 * Lat * es) : P;
  {
    Op& op = b.gemm(e->pre_conv, E, B, T, T);
    if (sp_main) op.g.p.out_f32 = (float*)P; else op.g.p.out_raw = P;
    b.tag(op, "pre_conv", P, sp_main, Lat, T);
  }
  if (sp_main) b.split((const float*)P, P16, 1, M, Lat);
  // ---- transformer (reference modeling :500-574); rows = all tokens, residual stream in fp32
  float* X = (float*)b.alloc((size_t)M * H * 4);
  void* Hn = b.alloc((size_t)M * H * es * 2);  // RMSNorm output, [hi | lo] rows where the consumer takes split operands
  void* QKV = b.alloc((size_t)M * 3 * H * es);
  void* AO = b.alloc((size_t)M * H * es);
  void* MI = b.alloc((size_t)M * I * es);
  void* TO = b.alloc((size_t)M * Lat * (sp_main ? 4 : es));
  void* TO16 = sp_main ? b.alloc((size_t)M * 2 * Lat * es) : TO;
  {
    Op& op = b.gemm(e->in_proj, P16, 1, M, M);
    op.g.p.out_f32 = X;
  }
  {
    Op& op0 = plan->ops.back();
    b.tag(op0, "transformer.in", X, true, H, T);
  }
  for (int l = 0; l < c.num_hidden_layers; ++l) {
    const LayerW& W = e->layers[l];
    const std::string L = "transformer." + std::to_string(l) + ".";  // debug probe taps (not oracle taps)
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_RMSNORM; op.in = X; op.out = Hn; op.w0 = W.ln1; op.B = M; op.C = H; op.split = sp_xf ? 1 : 0;
      b.tag(op, (L + "ln1").c_str(), Hn, false, H, T, sp_xf ? 2 * H : 0); }
    { Op& op = b.gemm(W.qkv, Hn, 1, M, M); op.g.p.out_raw = QKV; b.tag(op, (L + "qkv").c_str(), QKV, false, 3 * H, T); }
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_ATTN; op.in = QKV; op.out = AO; op.B = B; op.L = T;
      b.tag(op, (L + "attn").c_str(), AO, false, H, T); }
    { Op& op = b.gemm(W.o, AO, 1, M, M); op.g.p.scale = W.ls1; op.g.p.res = X; op.g.p.res_f32 = 1; op.g.p.out_f32 = X;
      b.tag(op, (L + "o").c_str(), X, true, H, T); }
    { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_RMSNORM; op.in = X; op.out = Hn; op.w0 = W.ln2; op.B = M; op.C = H; op.split = sp_xf ? 1 : 0;
      b.tag(op, (L + "ln2").c_str(), Hn, false, H, T, sp_xf ? 2 * H : 0); }
    {
      Op& op = b.gemm(W.gateup, Hn, 1, M, M);
      op.g.p.act = ACT_GLU; op.g.p.out_raw = MI; op.g.p.out_row_stride = I; op.g.p.out_batch_stride = (long long)M * I;
      b.tag(op, (L + "gateup").c_str(), MI, false, I, T);
    }
    { Op& op = b.gemm(W.down, MI, 1, M, M); op.g.p.scale = W.ls2; op.g.p.res = X; op.g.p.res_f32 = 1; op.g.p.out_f32 = X;
      b.tag(op, (L + "down").c_str(), X, true, H, T); }
  }
  { plan->ops.emplace_back(); Op& op = plan->ops.back(); op.type = OP_RMSNORM; op.in = X; op.out = Hn; op.w0 = e->final_norm; op.B = M; op.C = H; op.split = sp_main ? 1 : 0; }
  {
    Op& op = b.gemm(e->out_proj, Hn, 1, M, M);
    if (sp_main) op.g.p.out_f32 = (float*)TO; else op.g.p.out_raw = TO;
    b.tag(op, "pre_transformer", TO, sp_main, Lat, T);
  }
  if (sp_main) b.split((const float*)TO, TO16, 1, M, Lat);
  // ---- upsample: [TransConv k=s ; ConvNeXt] x n (reference modeling :877-879, :226-242)
  // Split-precision plans keep the main path (transposed-conv output Uo, block output V = Uo + gamma * ...) in fp32 and hand the
  // next GEMM a [hi | lo] copy of V; the ConvNeXt's inner tensors (LayerNorm output, GELU output) stay one 16-bit value.
  const void* cur = TO16;
  int L = T;
  for (size_t i = 0; i < e->ups.size(); ++i) {
    const UpsampleW& U = e->ups[i];
    const int r = c.upsampling_ratios[i];
    const int Lo = L * r;
    void* Uo = b.alloc((size_t)B * Lo * Lat * (sp_main ? 4 : es));
    void* Nn = b.alloc((size_t)B * Lo * Lat * es);
    void* G = b.alloc((size_t)B * Lo * 4 * Lat * es);
    void* V = b.alloc((size_t)B * Lo * Lat * (sp_main ? 4 : es));
    // the 16-bit copy the next GEMM reads: [hi | lo] rows when it takes split operands (the next upsample conv always does; decoder[0]
    // only in the split-all mode), else the fp32 tensor rounded once
    const bool next_split = i + 1 < e->ups.size() ? e->ups[i + 1].tconv.k_mult == 3 : e->dec0.k_mult == 3;
    void* V16 = sp_main ? b.alloc((size_t)B * Lo * (next_split ? 2 : 1) * Lat * es) : V;
    {
      Op& op = b.gemm(U.tconv, cur, 1, B * L, B * L);  // k == stride: rows are independent, flatten the batch
      if (sp_main) op.g.p.out_f32 = (float*)Uo; else op.g.p.out_raw = Uo;
      const std::string t = "upsample." + std::to_string(i) + ".0";
      b.tag(op, t.c_str(), Uo, sp_main, Lat, Lo);
    }
    {
      plan->ops.emplace_back();
      Op& op = plan->ops.back();
      op.type = OP_DWLN; op.in = Uo; op.out = Nn; op.w0 = U.dww; op.w1 = U.dwb; op.w2 = U.lnw; op.w3 = U.lnb; op.B = B; op.L = Lo; op.C = Lat;
      op.split = sp_main ? 1 : 0;  // fp32 input
    }
    { Op& op = b.gemm(U.pw1, Nn, 1, B * Lo, B * Lo); op.g.p.act = ACT_GELU; op.g.p.out_raw = G; }
    {
      Op& op = b.gemm(U.pw2, G, 1, B * Lo, B * Lo);
      op.g.p.scale = U.gamma; op.g.p.res = Uo;
      if (sp_main) { op.g.p.res_f32 = 1; op.g.p.out_f32 = (float*)V; } else op.g.p.out_raw = V;
      const std::string t = "upsample." + std::to_string(i) + ".1";
      b.tag(op, t.c_str(), V, sp_main, Lat, Lo);
    }
    if (sp_main) b.split((const float*)V, V16, 1, B * Lo, Lat, next_split);
    cur = V16; L = Lo;
  }
  // ---- vocoder: three rotating buffers sized for the largest stage
  size_t big = (size_t)B * L * c.decoder_dim;
  {
    int Lb = L;
    for (size_t i = 0; i < e->blocks.size(); ++i) {
      Lb = (Lb - 1) * e->blocks[i].stride;
      if (Lb < 0) Lb = 0;
      const size_t n = (size_t)B * Lb * e->blocks[i].cout;
      if (n > big) big = n;
    }
  }
  void* R[3];
  for (int i = 0; i < 3; ++i) R[i] = b.alloc(big * es + 4096);
  int ia = 0, iy = 1, ifree = 2;  // ia: Snake'd conv input, iy: raw residual stream, ifree: scratch
  {
    // decoder[0]: conv7 Lat -> decoder_dim, epilogue applies block 1's input Snake (modeling :856, :645)
    Op& op = b.gemm(e->dec0, cur, B, L, L);
    op.g.p.out_snake = R[ia];
    if (!e->blocks.empty()) { op.g.p.snake_a = e->blocks[0].act0.a; op.g.p.snake_ib = e->blocks[0].act0.ib; }
    b.tag(op, "decoder.0.snake", R[ia], false, c.decoder_dim, L);
  }
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    const BlockW& K = e->blocks[i];
    const int Lo = L > 0 ? (L - 1) * K.stride : 0;
    // raw-input units (C = 96): the block's tensors travel as x only, every unit applies its input Snake in shared memory
    bool xin = e->engine == ENGINE_TC && unit_xin_available(K.cout);
    for (int j = 0; j < 3 && xin; ++j) xin = K.unit[j].conv7.taps == 7 && K.unit[j].conv1.taps == 1 && 128 + 6 * (-K.unit[j].conv7.tap_off[0] / 6) <= 256;
    {
      // transposed conv: accumulator row q -> output rows q*s .. q*s+s-1 (modeling :200-207)
      Op& op = b.gemm(K.tconv, R[ia], B, L, L - 1);
      op.g.p.out_batch_stride = (long long)Lo * K.cout;
      op.g.p.out_raw = R[iy];
      const std::string t = "decoder." + std::to_string(i + 1) + ".block.1";
      if (!xin) {
        op.g.p.out_snake = R[ifree];
        op.g.p.snake_a = K.unit[0].act1.a; op.g.p.snake_ib = K.unit[0].act1.ib;
      }
      b.tag(op, t.c_str(), R[iy], false, K.cout, Lo);
      if (!xin) std::swap(ia, ifree);
    }
    L = Lo;
    for (int j = 0; j < 3; ++j) {
      const UnitW& U = K.unit[j];
      const SnakeP& nx = j < 2 ? K.unit[j + 1].act1 : (i + 1 < e->blocks.size() ? e->blocks[i + 1].act0 : e->tail_act);
      const std::string t = "decoder." + std::to_string(i + 1) + ".block." + std::to_string(j + 2);
      if (xin) {
        // x in R[iy] -> y in R[ifree] (never in place: tiles read their neighbours' halo rows); the last unit leaves snake_next(y) in R[ia]
        UnitOp u;
        memset(&u.p, 0, sizeof(u.p));
        u.C = K.cout; u.dil = -U.conv7.tap_off[0] / 6; u.xin = 1;
        u.bias7 = U.conv7.bias; u.snake2_a = U.act2.a; u.snake2_ib = U.act2.ib;
        u.snake1_a = U.act1.a; u.snake1_ib = U.act1.ib;
        { const char* d = Q3_ENV("Q3TTS_DBG"); u.p.dbg = d ? atoi(d) : 0; }
        u.p.bf16 = e->bf16 ? 1 : 0; u.p.bias = U.conv1.bias; u.p.res = R[iy]; u.p.out_raw = R[ifree];
        if (j == 2) { u.p.out_snake = R[ia]; u.p.snake_a = nx.a; u.p.snake_ib = nx.ib; }
        u.p.w_ptr = U.conv7.w;
        int mk = 0;
        if (!dry) mk = unit_make(&u, R[iy], U.conv7.w, U.conv1.w, B, L);
        if (mk != 0) { b.rc = e->fail(Q3TTS_CUDA_ERROR, "raw-input fused unit could not be built"); continue; }
        plan->ops.emplace_back();
        Op& op = plan->ops.back();
        op.type = OP_UNIT; op.u = u; op.B = B; op.L = L; op.C = K.cout;
        op.drop_raw = j == 2;
        b.tag(op, t.c_str(), R[ifree], false, K.cout, L);
        std::swap(iy, ifree);
        continue;
      }
      {
        // fused unit: one launch, the Snake'd conv7 output stays in tensor memory (C <= 192); x is updated in place
        UnitOp u;
        memset(&u.p, 0, sizeof(u.p));
        u.C = K.cout; u.dil = -U.conv7.tap_off[0] / 6;
        u.bias7 = U.conv7.bias; u.snake2_a = U.act2.a; u.snake2_ib = U.act2.ib;
        { const char* d = Q3_ENV("Q3TTS_DBG"); u.p.dbg = d ? atoi(d) : 0; }
        { const char* tc = Q3_ENV("Q3TTS_TRACE_C"); if (tc && atoi(tc) != K.cout) u.p.dbg &= ~32; }  // timeline of one channel count only
        u.p.bf16 = e->bf16 ? 1 : 0; u.p.bias = U.conv1.bias; u.p.res = R[iy]; u.p.out_raw = R[iy]; u.p.out_snake = R[ifree];
        u.p.snake_a = nx.a; u.p.snake_ib = nx.ib;
        u.p.w_ptr = U.conv7.w;
        const bool shape_ok = e->engine == ENGINE_TC && U.conv7.taps == 7 && U.conv1.taps == 1 && unit_fusable(K.cout);
        int mk = 1;
        if (shape_ok) {
          if (dry) mk = 0;
          else mk = unit_make(&u, R[ia], U.conv7.w, U.conv1.w, B, L);
          if (mk < 0) b.rc = e->fail(Q3TTS_CUDA_ERROR, "fused unit tensor maps failed");
        }
        if (mk == 0) {
          plan->ops.emplace_back();
          Op& op = plan->ops.back();
          op.type = OP_UNIT; op.u = u; op.B = B; op.L = L; op.C = K.cout;
          op.drop_raw = j == 2;
          b.tag(op, t.c_str(), R[iy], false, K.cout, L);
          std::swap(ia, ifree);
          continue;
        }
      }
      {
        Op& op = b.gemm(U.conv7, R[ia], B, L, L);
        op.g.p.out_snake = R[ifree]; op.g.p.snake_a = U.act2.a; op.g.p.snake_ib = U.act2.ib;
        std::swap(ia, ifree);
      }
      {
        Op& op = b.gemm(U.conv1, R[ia], B, L, L);
        op.g.p.res = R[iy]; op.g.p.out_raw = R[iy];
        op.g.p.out_snake = R[ifree]; op.g.p.snake_a = nx.a; op.g.p.snake_ib = nx.ib;
        op.drop_raw = j == 2;
        b.tag(op, t.c_str(), R[iy], false, K.cout, L);
        std::swap(ia, ifree);
      }
    }
  }
  {
    // decoder[6] + clamp on the tensor cores (N padded to 16); out pointer / stride / skip are patched per call
    Op& op = b.gemm(e->tail, R[ia], B, L, L);
    op.type = OP_TAIL;
    op.g.p.act = ACT_TAIL;
    op.B = B; op.L = L; op.C = e->tail_c;
    op.tag = "wav";
    if (!dry && tail_supported(e->tail_c)) {
      if (tail_make(&op.t, R[ia], e->tail_w16, B, L, 0, e->tail_c, e->bf16) < 0) b.rc = e->fail(Q3TTS_CUDA_ERROR, "tail tensor maps failed");
      op.t.p.rows_valid = L; op.t.p.a_row0 = 0; op.t.p.bias = e->tail_bias;
    }
  }
  plan->out_len = L;
  if (!dry) {
    for (Op& op : plan->ops)
      if ((op.type == OP_GEMM || op.type == OP_TAIL) && gemm_make_epilogue_maps(&op.g)) b.rc = e->fail(Q3TTS_CUDA_ERROR, "epilogue tensor map failed");
  }
  if (bytes) *bytes = b.off + 1024;
  return b.rc;
}

}  // namespace

size_t engine_plan_bytes(const Engine* e, int B, int T) {
  Plan tmp;
  size_t bytes = 0;
  build(const_cast<Engine*>(e), &tmp, B, T, true, &bytes);
  return bytes;
}

int engine_get_plan(Engine* e, int B, int T, Plan** out) {
  auto key = std::make_pair(B, T);
  auto it = e->plans.find(key);
  if (it != e->plans.end()) { *out = &it->second; return 0; }
  const size_t need = engine_plan_bytes(e, B, T);
  if (need > e->ws_bytes) {
    cudaDeviceSynchronize();
    if (e->ws) cudaFree(e->ws);
    e->ws = nullptr; e->ws_bytes = 0;
    e->plans.clear();  // tensor maps point into the old workspace
    if (cudaMalloc((void**)&e->ws, need) != cudaSuccess) { cudaGetLastError(); return e->fail(Q3TTS_OOM, "workspace cudaMalloc failed"); }
    e->ws_bytes = need;
  }
  Plan& p = e->plans[key];
  const int rc = build(e, &p, B, T, false, nullptr);
  if (rc) { e->plans.erase(key); return rc; }
  *out = &p;
  return 0;
}

// The profile ACCUMULATES over every engine launch (all chunks and sub-batches of a step) until q3tts_get_profile reads and clears
// it, so per-kernel times come from the same launches that a timed step consists of.  Returns the index of this run's first op.
size_t prof_begin(Engine* e, size_t nops) {
  const size_t first = e->prof_type.size();
  while (e->prof_events.size() < 2 * (first + nops)) {
    cudaEvent_t ev;
    cudaEventCreate(&ev);
    e->prof_events.push_back(ev);
  }
  return first;
}
void prof_clear(Engine* e) { e->prof_type.clear(); e->prof_flops.clear(); e->prof_bytes.clear(); e->prof_dims.clear(); }

// algorithmic FLOPs and minimum HBM bytes of one op, appended to the profile of the current run
void prof_account(Engine* e, const Op& op) {
  double fl = 0.0, by = 0.0;
  int dims[4] = {op.B * (op.L > 0 ? op.L : 1), op.C, 0, 0};
  if (op.type == OP_GEMM || op.type == OP_TAIL) {
    const GemmParams& p = op.g.p;
    const double rows = (double)p.batches * p.rows_valid;
    dims[0] = p.batches * p.rows_valid; dims[1] = p.n_total; dims[2] = p.cin; dims[3] = p.ntaps;
    const int km = p.k_mult > 0 ? p.k_mult : 1;  // split-precision layers walk k_mult x the real K: the extra MMAs are not algorithmic work
    dims[2] = p.cin / km;
    fl = 2.0 * rows * p.n_total * (p.cin / km) * p.ntaps;
    by = 2.0 * ((double)p.batches * p.rows * (p.a_cols ? p.a_cols : p.cin) + (double)p.ntaps * p.n_total * p.cin);  // A once + W once
    const double outw = p.act == ACT_GLU ? p.n_total / 2 : p.n_total;
    if (p.out_raw) by += 2.0 * rows * outw;
    if (p.out_snake) by += 2.0 * rows * outw;
    if (p.out_f32) by += 4.0 * rows * outw;
    if (p.res) by += (p.res_f32 ? 4.0 : 2.0) * rows * outw;
  } else if (op.type == OP_UNIT) {
    const double rows = (double)op.B * op.L, C = op.C;
    dims[0] = op.B * op.L; dims[1] = op.C; dims[2] = op.C; dims[3] = 8;  // 7 taps + the 1x1 conv
    fl = 2.0 * rows * C * C * 8.0;
    by = 2.0 * ((op.u.xin ? 2.0 : (op.u.p.skip_raw ? 3.0 : 4.0)) * rows * C + 8.0 * C * C);  // a1 + x in, y (unless dead) + snake(y) out, weights once; raw-input form: x in, one tensor out
  } else if (op.type == OP_RVQ) {
    // the 32 MiB of tables are L2-resident after warm-up, so the HBM floor is the codes in and the embeddings out
    by = (double)op.B * op.L * (e->cfg.num_quantizers * 8.0 + 2.0 * op.C);
  } else if (op.type == OP_RMSNORM) {
    by = (double)op.B * op.C * (4.0 + 2.0);  // fp32 residual stream in, 16-bit normalised rows out (op.B = rows)
  } else if (op.type == OP_ATTN) {
    const double rows = (double)op.B * op.L, H = (double)e->cfg.num_attention_heads * e->cfg.head_dim;
    dims[1] = (int)H;
    fl = 4.0 * rows * H * e->cfg.sliding_window;  // QK^T and PV at a full window, as SURVEY 8(a) counts them
    by = 2.0 * rows * 4.0 * H;                    // q, k, v in, o out (16-bit)
  } else if (op.type == OP_SPLIT) {
    by = (4.0 + (op.split ? 4.0 : 2.0)) * (double)op.B * op.L * op.C;  // fp32 in, one or two 16-bit values out
  } else if (op.type == OP_ENC_CONV0) {
    by = (double)op.B * op.L * (4.0 + 4.0 * op.C);                       // samples in, fp32 rows out
    fl = 2.0 * 7.0 * (double)op.B * op.L * op.C;
  } else if (op.type == OP_ENC_ACT) {
    by = 8.0 * (double)op.B * op.L * op.C;                                 // fp32 in, [hi | lo] fp16 out
  } else if (op.type == OP_ENC_LN) {
    by = 8.0 * (double)op.B * op.C;                                        // op.B = rows
  } else if (op.type == OP_ENC_ATTN) {
    const double rows = (double)op.B * op.L;
    fl = 4.0 * rows * op.C * (op.L < e->enc_cfg.sliding_window ? op.L : e->enc_cfg.sliding_window);
    by = rows * op.C * (12.0 + 4.0);                                       // q, k, v fp32 in, [hi | lo] out
  } else if (op.type == OP_ENC_RVQ) {
    by = 0.0;                                                              // parity tap (only when the caller asks for latents)
  } else if (op.type == OP_ENC_SPLITCOLS) {
    by = 8.0 * (double)op.B * op.C;
  } else if (op.type == OP_ENC_ARGMIN) {
    by = (double)op.B * op.L * (4.0 * op.C + 8.0 + 3.0 * 4.0 * (e->cfg.codebook_dim / 2));   // scores in, code out, residual in / out + operand out
  } else if (op.type == OP_DWLN) {
    by = (op.split ? 4.0 + 2.0 : 2.0 * 2.0) * (double)op.B * op.L * op.C;  // one read and one write of the token tensor
  }
  // profile kernel codes (include/q3tts_decode.h q3tts_get_profile): the three implicit-GEMM kernels are told apart
  int code = (int)op.type;
  if (op.type == OP_GEMM) code = op.g.variant == 1 ? 7 : (op.g.variant == 2 ? 8 : 1);
  else if (op.type == OP_TAIL) code = 5;
  else if (op.type == OP_UNIT) code = 6;
  else if (op.type == OP_STREAM_ATTN) code = 9;
  else if (op.type == OP_STATE_LOAD || op.type == OP_STATE_SAVE || op.type == OP_POS_ADVANCE) code = 10;
  else if (op.type == OP_SPLIT) code = 11;
  e->prof_type.push_back(code);
  e->prof_flops.push_back(fl);
  e->prof_bytes.push_back(by);
  for (int d = 0; d < 4; ++d) e->prof_dims.push_back(dims[d]);
}

int engine_run(Engine* e, Plan* plan, const long long* codes, long long sb, long long sk, long long st, float* wav, long long wav_stride,
               int skip, const char* probe_stage, float* probe_out, long long probe_cap, int* pc, int* pl, cudaStream_t s,
               long long out_pos0, const long long* item_skip) {
  const q3tts_config& c = e->cfg;
  size_t pi = e->profile ? prof_begin(e, plan->ops.size()) : 0;
  static const bool sync_each = Q3_ENV("Q3TTS_SYNC") != nullptr;  // debug: serialise ops with the host
  static const bool keep_raw = Q3_ENV("Q3TTS_KEEP_RAW") != nullptr;  // debug / A-B: store the dead raw outputs too
  for (Op& op : plan->ops) {
    cudaError_t err = cudaSuccess;
    if (e->profile) cudaEventRecord(e->prof_events[2 * pi], s);
    // dead raw output of a block's last residual unit: not written (a probe of that stage still needs it)
    const bool drop = op.drop_raw && !probe_stage && !keep_raw;
    void* const saved_raw = op.g.p.out_raw;
    if (op.type == OP_UNIT) op.u.p.skip_raw = drop ? 1 : 0;
    else if (op.type == OP_GEMM && drop) op.g.p.out_raw = nullptr;
    switch (op.type) {
      case OP_RVQ:
        err = rvq_decode(codes, sb, sk, st, e->rvq_tables, op.out, nullptr, op.B, c.num_quantizers, op.L, c.codebook_size, c.codebook_dim, e->bf16, s, 0, nullptr,
                         e->d_bad_code, op.split);
        break;
      case OP_GEMM: err = launch_conv_gemm(op.g, e->engine, s); break;
      case OP_UNIT: err = launch_unit(op.u, s); break;
      case OP_RMSNORM: err = rmsnorm((const float*)op.in, op.w0, op.out, op.B, op.C, c.rms_norm_eps, e->bf16, s, op.split); break;
      case OP_SPLIT: err = split_f32((const float*)op.in, op.sp_in_bs, op.out, op.sp_out_bs, op.B, op.L, op.C, e->bf16, s, op.split); break;
      case OP_ATTN:
        err = sliding_attention(op.in, e->rope_cos, e->rope_sin, op.out, op.B, op.L, c.num_attention_heads, c.head_dim, c.sliding_window, e->bf16, s);
        break;
      case OP_DWLN: err = dwconv_layernorm(op.in, op.w0, op.w1, op.w2, op.w3, op.out, op.B, op.L, op.C, 1e-6f, e->bf16, s, 0, 0, op.split != 0); break;
      case OP_TAIL:
        if (wav) {
          if (op.t.enabled && e->engine == ENGINE_TC) {
            op.t.p.out = e->out_pcm16 ? nullptr : wav; op.t.p.out_i16 = e->out_pcm16 ? reinterpret_cast<short*>(wav) : nullptr;
            op.t.p.out_batch_stride = wav_stride; op.t.p.tail_skip = skip;
            op.t.p.out_pos0 = out_pos0; op.t.p.item_skip = item_skip;
            err = launch_tail(op.t, s);
          } else if (e->out_pcm16 || item_skip) {
            err = cudaErrorNotSupported;  // PCM16 output and the per-item prefix cut exist in the taps-as-N output conv only
          } else {
            op.g.p.out_f32 = wav + out_pos0; op.g.p.out_batch_stride = wav_stride; op.g.p.tail_skip = skip;
            err = launch_conv_gemm(op.g, e->engine, s);
          }
        }
        break;
    }
    ++e->launches;
    if (sync_each) cudaStreamSynchronize(s);
    if (e->profile) {
      cudaEventRecord(e->prof_events[2 * pi + 1], s);
      prof_account(e, op);
      ++pi;
    }
    op.g.p.out_raw = saved_raw;
    if (err != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("launch failed: ") + cudaGetErrorString(err));
    if (probe_stage && !op.tag.empty() && op.tag == probe_stage && op.type != OP_TAIL) {
      const long long n = (long long)plan->B * op.probe_C * op.probe_L;
      if (n > probe_cap) return e->fail(Q3TTS_BAD_SHAPE, "probe buffer too small");
      err = to_channels_first_f32(op.probe_ptr, op.probe_f32, probe_out, plan->B, op.probe_L, op.probe_C, e->bf16, s, op.probe_ld);
      if (err != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("probe failed: ") + cudaGetErrorString(err));
      if (pc) *pc = op.probe_C;
      if (pl) *pl = op.probe_L;
      return Q3TTS_OK;
    }
  }
  if (probe_stage) return e->fail(Q3TTS_BAD_KEY, std::string("unknown probe stage: ") + probe_stage);
  return Q3TTS_OK;
}

// forward over B items, split along the batch so that no launch exceeds the reserved frame budget
int engine_forward_batched(Engine* e, const long long* codes, long long sb, long long sk, long long st, int B, int T, float* wav,
                           long long wav_stride, int skip, cudaStream_t s, long long out_pos0, const long long* item_skip) {
  if (T > e->rope_max) return e->fail(Q3TTS_BAD_SHAPE, "T exceeds max_position_embeddings");
  int per = e->max_frames / (T > 0 ? T : 1);
  if (per < 1) per = 1;
  for (int b0 = 0; b0 < B; b0 += per) {
    const int nb = (B - b0) < per ? (B - b0) : per;
    Plan* plan = nullptr;
    int rc = engine_get_plan(e, nb, T, &plan);
    if (rc) return rc;
    // PCM16 output: `wav` is an int16 buffer, offsets are in its elements
    float* dst = e->out_pcm16 ? reinterpret_cast<float*>(reinterpret_cast<short*>(wav) + (long long)b0 * wav_stride) : wav + (long long)b0 * wav_stride;
    rc = engine_run(e, plan, codes + b0 * sb, sb, sk, st, dst, wav_stride, skip, nullptr, nullptr, 0, nullptr, nullptr, s, out_pos0,
                    item_skip ? item_skip + b0 : nullptr);
    if (rc) return rc;
  }
  return Q3TTS_OK;
}

int engine_check_codes(Engine* e, cudaStream_t s) {
  if (!e->d_bad_code || !e->h_bad_code) return Q3TTS_OK;
  if (cudaMemcpyAsync(e->h_bad_code, e->d_bad_code, sizeof(int), cudaMemcpyDeviceToHost, s) != cudaSuccess || cudaStreamSynchronize(s) != cudaSuccess)
    return e->fail(Q3TTS_CUDA_ERROR, std::string("code range check failed: ") + cudaGetErrorString(cudaGetLastError()));
  if (*e->h_bad_code == 0) return Q3TTS_OK;
  cudaMemsetAsync(e->d_bad_code, 0, sizeof(int), s);
  return e->fail(Q3TTS_CODE_OUT_OF_RANGE, "a code index is outside [0, codebook_size): the reference's F.embedding would raise here (modeling :675-678)");
}

void engine_free(Engine* e) {
  engine_stream_free(e);
  encoder_free(e);
  for (cudaEvent_t ev : e->prof_events) cudaEventDestroy(ev);
  for (void* p : e->owned) cudaFree(p);
  for (auto& kv : e->raw) cudaFree(kv.second.d);
  if (e->ws) cudaFree(e->ws);
  if (e->codes_stage) cudaFree(e->codes_stage);
  if (e->wav_stage) cudaFree(e->wav_stage);
  if (e->codes16_stage) cudaFree(e->codes16_stage);
  if (e->item_skip_stage) cudaFree(e->item_skip_stage);
  if (e->enc_tables) cudaFree(e->enc_tables);
  if (e->enc_norms) cudaFree(e->enc_norms);
  if (e->enc_win) cudaFree(e->enc_win);
  if (e->d_bad_code) cudaFree(e->d_bad_code);
  if (e->h_bad_code) cudaFreeHost(e->h_bad_code);
}

}  // namespace q3
