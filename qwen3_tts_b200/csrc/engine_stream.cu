// Streaming (stateful per-packet) decode: launch plans over history-prefixed buffers, per-slot state, q3tts_stream_step.
//
// Every stage that looks back in time keeps its input as [stream][H history rows | n new rows][C]; the history rows are loaded
// from / saved to a per-slot state block around the step, so causal convs read real history instead of zero padding, the
// transposed convs (which need input q+1 for output group q) emit with a one-input delay exactly like the reference's
// both-ends trim, and attention reads a 128-entry K/V ring per layer.  A stream's first packet ("fresh") uses the same plan
// with zero state and one output group less per transposed conv, so stream output == Decoder.forward(whole sequence).
#include <cstdlib>
#include <cstring>

#include "engine.h"
#include "kernels.h"
#include "launch.cuh"

namespace q3 {

namespace {

struct SBuilder {
  Engine* e;
  StreamPlan* plan;
  uint8_t* base;
  size_t off = 0;
  bool dry;
  int rc = 0;
  long long state_off = 0;

  void* alloc(size_t bytes) {
    off = (off + 1023) & ~(size_t)1023;
    void* p = dry ? nullptr : base + off;
    off += bytes;
    return p;
  }
  // history-prefixed buffer [S][H + n][C] of `esz`-byte elements (2: 16-bit, 4: fp32), registered for the state load / save kernels
  void* hist_buf(int S, int H, int n, int C, int esz = 2) {
    void* p = alloc((size_t)S * (H + n) * C * esz);
    StreamBufDesc d;
    d.ptr = (unsigned char*)p; d.batch_stride = (long long)(H + n) * C * esz; d.state_off = state_off; d.H = H; d.n = n; d.row_bytes = C * esz; d.pad = 0;
    plan->descs.push_back(d);
    state_off += (long long)H * C * esz;
    return p;
  }
  static void* at_row(void* p, int row, int C, int esz = 2) { return p ? (uint8_t*)p + (size_t)row * C * esz : nullptr; }

  // in: [batches][a_rows][cin] (a_rows includes a_row0 history rows), rows_valid accumulator rows per batch item
  Op& gemm(const ConvW& w, const void* in, int batches, int a_rows, int a_row0, int rows_valid) {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = OP_GEMM;
    op.g.retile = 1;
    GemmParams& p = op.g.p;
    memset(&p, 0, sizeof(p));
    p.batches = batches; p.rows = a_rows; p.a_row0 = a_row0; p.rows_valid = rows_valid;
    p.cin = w.cin; p.ntaps = w.taps;
    for (int i = 0; i < 8; ++i) p.tap_off[i] = w.tap_off[i];
    p.n_total = w.n_total; p.bn = w.bn; p.kb = w.kb; p.ncol_mod = w.ncol_mod;
    p.bias = w.bias; p.bf16 = e->bf16 ? 1 : 0;
    p.a_cols = w.a_cols; p.k_mult = w.k_mult;
    p.a_ptr = in; p.w_ptr = w.w;
    p.out_row_stride = w.n_total;
    p.out_batch_stride = (long long)rows_valid * w.n_total;
    // the A map is (re)built from p.a_ptr by gemm_make_epilogue_maps once every pointer of the plan is final
    if (!dry && make_tmap_3d(&op.g.tmB, w.w, (uint64_t)w.cin, (uint64_t)w.n_total, (uint64_t)w.taps, (uint32_t)w.kb, (uint32_t)w.bn, e->bf16))
      rc = e->fail(Q3TTS_CUDA_ERROR, "cuTensorMapEncodeTiled failed (weights)");
    return op;
  }
  Op& split(const float* in, long long in_bs, void* out, long long out_bs, int batches, int rows, int C, bool with_lo = true) {
    Op& op = simple(OP_SPLIT);
    op.in = in; op.out = out; op.B = batches; op.L = rows; op.C = C; op.sp_in_bs = in_bs; op.sp_out_bs = out_bs; op.split = with_lo ? 1 : 0;
    return op;
  }
  Op& simple(OpType t) {
    plan->ops.emplace_back();
    Op& op = plan->ops.back();
    op.type = t;
    return op;
  }
};

int build_stream(Engine* e, StreamPlan* plan, int S, int P, bool fresh, bool dry, size_t* bytes) {
  const q3tts_config& c = e->cfg;
  SBuilder b{e, plan, e->sws, 0, dry};
  plan->S = S; plan->P = P; plan->fresh = fresh; plan->ops.clear(); plan->descs.clear();
  const int H = c.hidden_size, Lat = c.latent_dim, I = c.intermediate_size, D = c.codebook_dim;
  const int M = S * P;
  const size_t es = 2;
  b.simple(OP_STATE_LOAD);
  const bool sp_main = e->pre_conv.k_mult == 3;   // split-precision main path (engine_plan.cu build(): same layout rules)
  const bool sp_xf = !e->layers.empty() && e->layers[0].qkv.k_mult == 3;
  const int Dw = sp_main ? 2 * D : D;             // width of a row of E: [hi | lo]
  // ---- RVQ -> E [S][2 | P][D]
  void* E = b.hist_buf(S, 2, P, Dw);
  { Op& op = b.simple(OP_RVQ); op.out = SBuilder::at_row(E, 2, Dw); op.B = S; op.L = P; op.C = D; op.aux0 = 2 + P; op.split = sp_main ? 1 : 0; }
  // ---- pre_conv (k = 3 causal) -> Pb [S*P][Lat]
  void* Pb = b.alloc((size_t)M * Lat * (sp_main ? 4 : es));
  void* Pb16 = sp_main ? b.alloc((size_t)M * 2 * Lat * es) : Pb;
  {
    Op& op = b.gemm(e->pre_conv, E, S, 2 + P, 2, P);
    if (sp_main) op.g.p.out_f32 = (float*)Pb; else op.g.p.out_raw = Pb;
  }
  if (sp_main) b.split((const float*)Pb, (long long)M * Lat, Pb16, (long long)M * 2 * Lat, 1, M, Lat);
  // ---- transformer over the S*P new tokens, K/V ring cache per layer
  float* X = (float*)b.alloc((size_t)M * H * 4);
  void* Hn = b.alloc((size_t)M * H * es * 2);
  void* QKV = b.alloc((size_t)M * 3 * H * es);
  void* AO = b.alloc((size_t)M * H * es);
  void* MI = b.alloc((size_t)M * I * es);
  void* TO = b.alloc((size_t)M * Lat * (sp_main ? 4 : es));
  void* TO16 = sp_main ? b.alloc((size_t)M * 2 * Lat * es) : TO;
  { Op& op = b.gemm(e->in_proj, Pb16, 1, M, 0, M); op.g.p.out_f32 = X; }
  for (int l = 0; l < c.num_hidden_layers; ++l) {
    const LayerW& W = e->layers[l];
    { Op& op = b.simple(OP_RMSNORM); op.in = X; op.out = Hn; op.w0 = W.ln1; op.B = M; op.C = H; op.split = sp_xf ? 1 : 0; }
    { Op& op = b.gemm(W.qkv, Hn, 1, M, 0, M); op.g.p.out_raw = QKV; }
    { Op& op = b.simple(OP_STREAM_ATTN); op.in = QKV; op.out = AO; op.B = S; op.L = P; op.aux0 = l; }
    { Op& op = b.gemm(W.o, AO, 1, M, 0, M); op.g.p.scale = W.ls1; op.g.p.res = X; op.g.p.res_f32 = 1; op.g.p.out_f32 = X; }
    { Op& op = b.simple(OP_RMSNORM); op.in = X; op.out = Hn; op.w0 = W.ln2; op.B = M; op.C = H; op.split = sp_xf ? 1 : 0; }
    { Op& op = b.gemm(W.gateup, Hn, 1, M, 0, M); op.g.p.act = ACT_GLU; op.g.p.out_raw = MI; op.g.p.out_row_stride = I; op.g.p.out_batch_stride = (long long)M * I; }
    { Op& op = b.gemm(W.down, MI, 1, M, 0, M); op.g.p.scale = W.ls2; op.g.p.res = X; op.g.p.res_f32 = 1; op.g.p.out_f32 = X; }
  }
  { Op& op = b.simple(OP_RMSNORM); op.in = X; op.out = Hn; op.w0 = e->final_norm; op.B = M; op.C = H; op.split = sp_main ? 1 : 0; }
  { Op& op = b.gemm(e->out_proj, Hn, 1, M, 0, M); if (sp_main) op.g.p.out_f32 = (float*)TO; else op.g.p.out_raw = TO; }
  if (sp_main) b.split((const float*)TO, (long long)M * Lat, TO16, (long long)M * 2 * Lat, 1, M, Lat);
  // ---- upsample: [TransConv k = s ; ConvNeXt]; the depthwise k7 needs 6 rows of history
  const void* cur = TO16;  // dense [S][L][Lat] ([hi | lo] rows in split-precision plans)
  const int ue = sp_main ? 4 : 2;   // element size of the main-path tensors Uo / V: fp32 in split-precision plans
  const bool sp_dec0 = e->dec0.k_mult == 3;  // decoder[0] takes split operands only in the split-all mode
  int L = P;
  void* dec_in = nullptr;  // [S][6 | L][Vw] input of decoder[0]
  for (size_t i = 0; i < e->ups.size(); ++i) {
    const UpsampleW& U = e->ups[i];
    const int r = c.upsampling_ratios[i];
    const int Lo = L * r;
    const bool last = i + 1 == e->ups.size();
    const bool next_split = sp_main && (last ? sp_dec0 : true);
    const int Vw = next_split ? 2 * Lat : Lat;  // row width of the 16-bit tensor the next GEMM reads
    void* Uo = b.hist_buf(S, 6, Lo, Lat, ue);
    void* Nn = b.alloc((size_t)S * Lo * Lat * es);
    void* G = b.alloc((size_t)S * Lo * 4 * Lat * es);
    // block output: dense fp32 + a [hi | lo] copy (split-precision), or 16-bit written straight by the epilogue; the last stage's
    // 16-bit tensor carries decoder[0]'s 6 history rows
    void* Vf = sp_main ? b.alloc((size_t)S * Lo * Lat * 4) : nullptr;
    void* V = last ? b.hist_buf(S, 6, Lo, Vw) : b.alloc((size_t)S * Lo * Vw * es);
    const long long V_bs = last ? (long long)(6 + Lo) * Vw : (long long)Lo * Vw;
    void* Vrow0 = last ? SBuilder::at_row(V, 6, Vw) : V;
    {
      Op& op = b.gemm(U.tconv, cur, S, L, 0, L);  // one accumulator row -> r output rows of the stream's own buffer
      if (sp_main) op.g.p.out_f32 = (float*)SBuilder::at_row(Uo, 6, Lat, ue); else op.g.p.out_raw = SBuilder::at_row(Uo, 6, Lat);
      op.g.p.out_batch_stride = (long long)(6 + Lo) * Lat;
    }
    { Op& op = b.simple(OP_DWLN); op.in = Uo; op.out = Nn; op.w0 = U.dww; op.w1 = U.dwb; op.w2 = U.lnw; op.w3 = U.lnb; op.B = S; op.L = Lo; op.C = Lat; op.aux0 = 6 + Lo; op.aux1 = 6;
      op.split = sp_main ? 1 : 0; }
    { Op& op = b.gemm(U.pw1, Nn, 1, S * Lo, 0, S * Lo); op.g.p.act = ACT_GELU; op.g.p.out_raw = G; }
    {
      Op& op = b.gemm(U.pw2, G, S, Lo, 0, Lo);
      op.g.p.scale = U.gamma;
      op.g.p.res = SBuilder::at_row(Uo, 6, Lat, ue); op.g.p.res_batch_stride = (long long)(6 + Lo) * Lat;
      if (sp_main) { op.g.p.res_f32 = 1; op.g.p.out_f32 = (float*)Vf; op.g.p.out_batch_stride = (long long)Lo * Lat; }
      else { op.g.p.out_raw = Vrow0; op.g.p.out_batch_stride = V_bs; }
    }
    if (sp_main) b.split((const float*)Vf, (long long)Lo * Lat, Vrow0, V_bs, S, Lo, Lat, next_split);
    cur = V; L = Lo;
    if (last) dec_in = V;
  }
  if (e->ups.empty()) return e->fail(Q3TTS_BAD_CONFIG, "streaming needs at least one upsampling ratio");
  // ---- decoder[0]: conv7 Lat -> decoder_dim, Snake of block 1 applied in the epilogue; block 1's transposed conv keeps 1 row of history
  int Cin = c.decoder_dim;
  void* Tin = b.hist_buf(S, 1, L, Cin);
  {
    Op& op = b.gemm(e->dec0, dec_in, S, 6 + L, 6, L);
    op.g.p.out_snake = SBuilder::at_row(Tin, 1, Cin);
    op.g.p.out_batch_stride = (long long)(1 + L) * Cin;
    if (!e->blocks.empty()) { op.g.p.snake_a = e->blocks[0].act0.a; op.g.p.snake_ib = e->blocks[0].act0.ib; }
  }
  void* tail_in = nullptr;
  size_t smid_max = 0;
  std::vector<std::pair<size_t, size_t>> smid_users;  // (op index, elements) of unfused units: they share one scratch buffer
  for (size_t i = 0; i < e->blocks.size(); ++i) {
    const BlockW& K = e->blocks[i];
    const int Cout = K.cout;
    // transposed conv, taps {x[q], x[q+1]}: buffer row j = input (j - 1) of this step, row 0 = last input of the previous step.
    // continuing: groups q = -1 .. L-2 (L groups, a_row0 = 0); fresh: there is no previous input, groups 0 .. L-2 (a_row0 = 1).
    const int groups = fresh ? L - 1 : L;
    const int Lo = groups > 0 ? groups * K.stride : 0;
    void* Xb = b.alloc((size_t)S * (Lo > 0 ? Lo : 1) * Cout * es);  // raw residual stream y, dense
    const int dil[3] = {1, 3, 9};
    void* A[3];
    for (int j = 0; j < 3; ++j) A[j] = b.hist_buf(S, 6 * dil[j], Lo, Cout);
    void* next = nullptr;
    int nextH = 0;
    if (i + 1 < e->blocks.size()) { next = b.hist_buf(S, 1, Lo, Cout); nextH = 1; }
    else { next = b.hist_buf(S, 6, Lo, Cout); nextH = 6; tail_in = next; }
    if (Lo > 0) {
      Op& op = b.gemm(K.tconv, Tin, S, 1 + L, fresh ? 1 : 0, groups);
      op.g.p.out_batch_stride = (long long)Lo * Cout;
      op.g.p.out_raw = Xb;
      op.g.p.out_snake = SBuilder::at_row(A[0], 6, Cout);
      op.g.p.snake_batch_stride = (long long)(6 + Lo) * Cout;
      op.g.p.snake_a = K.unit[0].act1.a; op.g.p.snake_ib = K.unit[0].act1.ib;
    }
    for (int j = 0; j < 3 && Lo > 0; ++j) {
      const UnitW& U = K.unit[j];
      const SnakeP& nx = j < 2 ? K.unit[j + 1].act1 : (i + 1 < e->blocks.size() ? e->blocks[i + 1].act0 : e->tail_act);
      const int Hc = 6 * dil[j];
      void* dst = j < 2 ? SBuilder::at_row(A[j + 1], 6 * dil[j + 1], Cout) : SBuilder::at_row(next, nextH, Cout);
      const long long dst_bs = (long long)((j < 2 ? 6 * dil[j + 1] : nextH) + Lo) * Cout;
      {
        UnitOp u;
        memset(&u.p, 0, sizeof(u.p));
        u.C = Cout; u.dil = dil[j];
        u.bias7 = U.conv7.bias; u.snake2_a = U.act2.a; u.snake2_ib = U.act2.ib;
        u.p.bf16 = e->bf16 ? 1 : 0; u.p.bias = U.conv1.bias; u.p.res = Xb; u.p.out_raw = Xb; u.p.out_snake = dst;
        u.p.snake_a = nx.a; u.p.snake_ib = nx.ib;
        u.a_rows = Hc + Lo; u.a_row0 = Hc; u.a_batch_stride = (long long)(Hc + Lo) * Cout; u.snake_batch_stride = dst_bs;
        int mk = 1;
        if (unit_fusable(Cout)) {
          mk = dry ? 0 : unit_make(&u, A[j], U.conv7.w, U.conv1.w, S, Lo);
          if (mk < 0) b.rc = e->fail(Q3TTS_CUDA_ERROR, "fused unit tensor maps failed");
        }
        // the dry pass sizes the scratch as if no unit fused, so a unit that falls back in the real pass still fits
        if (dry && (size_t)S * Lo * Cout > smid_max) smid_max = (size_t)S * Lo * Cout;
        if (mk == 0) {
          Op& op = b.simple(OP_UNIT);
          op.u = u; op.B = S; op.L = Lo; op.C = Cout;
          continue;
        }
      }
      {
        Op& op = b.gemm(U.conv7, A[j], S, Hc + Lo, Hc, Lo);
        op.g.p.snake_a = U.act2.a; op.g.p.snake_ib = U.act2.ib;  // out_snake = shared scratch, patched below
        smid_users.push_back({plan->ops.size() - 1, (size_t)S * Lo * Cout});
        if ((size_t)S * Lo * Cout > smid_max) smid_max = (size_t)S * Lo * Cout;
      }
      {
        Op& op = b.gemm(U.conv1, nullptr, S, Lo, 0, Lo);  // A = shared scratch, patched below
        op.g.p.res = Xb; op.g.p.out_raw = Xb;
        op.g.p.out_snake = dst; op.g.p.snake_batch_stride = dst_bs;
        op.g.p.snake_a = nx.a; op.g.p.snake_ib = nx.ib;
      }
    }
    Tin = next; L = Lo; Cin = Cout;
  }
  {
    void* Smid = b.alloc((smid_max ? smid_max : 1) * es);
    for (auto& u : smid_users) {
      plan->ops[u.first].g.p.out_snake = Smid;
      plan->ops[u.first + 1].g.p.a_ptr = Smid;
    }
  }
  if (L > 0) {
    Op& op = b.gemm(e->tail, tail_in, S, 6 + L, 6, L);
    op.type = OP_TAIL;
    op.g.p.act = ACT_TAIL;
    op.B = S; op.L = L; op.C = e->tail_c;
    if (!dry && tail_supported(e->tail_c)) {
      if (tail_make(&op.t, tail_in, e->tail_w16, S, 6 + L, 0, e->tail_c, e->bf16) < 0) b.rc = e->fail(Q3TTS_CUDA_ERROR, "tail tensor maps failed");
      op.t.p.rows_valid = L; op.t.p.a_row0 = 6; op.t.p.bias = e->tail_bias;
    }
  }
  b.simple(OP_STATE_SAVE);
  b.simple(OP_POS_ADVANCE);
  plan->out_len = L;
  plan->wav_tmp = (float*)b.alloc((size_t)S * (L > 0 ? L : 1) * 4);
  if (!dry) {
    if (plan->d_descs) { cudaFree(plan->d_descs); plan->d_descs = nullptr; }
    if (cudaMalloc((void**)&plan->d_descs, plan->descs.size() * sizeof(StreamBufDesc)) != cudaSuccess) { cudaGetLastError(); return e->fail(Q3TTS_OOM, "stream plan: descriptor allocation failed"); }
    for (Op& op : plan->ops)
      if ((op.type == OP_GEMM || op.type == OP_TAIL) && gemm_make_epilogue_maps(&op.g)) b.rc = e->fail(Q3TTS_CUDA_ERROR, "epilogue tensor map failed");
    if (b.off > e->sws_bytes) return e->fail(Q3TTS_CUDA_ERROR, "stream plan: arena overrun (" + std::to_string(b.off) + " > " + std::to_string(e->sws_bytes) + ")");
    const cudaError_t ce = cudaMemcpy(plan->d_descs, plan->descs.data(), plan->descs.size() * sizeof(StreamBufDesc), cudaMemcpyHostToDevice);
    if (ce != cudaSuccess) b.rc = e->fail(Q3TTS_CUDA_ERROR, std::string("stream plan: descriptor upload failed: ") + cudaGetErrorString(ce));
  }
  if (e->state_bytes == 0) e->state_bytes = (b.state_off + 255) & ~255ll;
  if (bytes) *bytes = b.off + 1024;
  return b.rc;
}

// All stream plans live in one arena; (re)allocating it drops them.  Every plan uses the arena from offset 0: plans of one
// step run one after the other, and all persistent data is in the per-slot state, not in the arena.
int get_stream_plan(Engine* e, int S, int P, bool fresh, StreamPlan** out) {
  auto key = std::make_pair(std::make_pair(S, P), fresh ? 1 : 0);
  auto it = e->splans.find(key);
  if (it != e->splans.end()) { *out = &it->second; return 0; }
  StreamPlan tmp;
  size_t need = 0;
  build_stream(e, &tmp, S, P, fresh, true, &need);
  if (need > e->sws_bytes) {
    cudaDeviceSynchronize();
    if (e->sws) cudaFree(e->sws);
    e->sws = nullptr; e->sws_bytes = 0;
    e->splans.clear();
    if (cudaMalloc((void**)&e->sws, need) != cudaSuccess) { cudaGetLastError(); return e->fail(Q3TTS_OOM, "stream workspace cudaMalloc failed"); }
    e->sws_bytes = need;
  }
  StreamPlan& p = e->splans[key];
  const int rc = build_stream(e, &p, S, P, fresh, false, nullptr);
  if (rc) { e->splans.erase(key); return rc; }
  *out = &p;
  return 0;
}

int launch_stream_plan(Engine* e, StreamPlan* plan, const long long* codes, long long sb, long long sk, long long st, const int* d_ids, const int* d_rows,
                       bool identity_rows, float* wav, long long wav_stride, cudaStream_t s);

// Runs the plan's launches directly, or as one CUDA graph once the same caller pointers have been seen twice in a row.
int run_stream_plan(Engine* e, StreamPlan* plan, const long long* codes, long long sb, long long sk, long long st, const int* d_ids, const int* d_rows,
                    bool identity_rows, float* wav, long long wav_stride, cudaStream_t s) {
  static const bool graphs_off = Q3_ENV("Q3TTS_STREAM_GRAPH") == nullptr || Q3_ENV("Q3TTS_NO_GRAPH") != nullptr || Q3_ENV("Q3TTS_SYNC") != nullptr;
  if (graphs_off || plan->graph_off || e->profile) return launch_stream_plan(e, plan, codes, sb, sk, st, d_ids, d_rows, identity_rows, wav, wav_stride, s);
  StreamPlan::GraphEntry key;
  key.codes = codes; key.sb = sb; key.sk = sk; key.st = st; key.wav = wav; key.stride = wav_stride; key.identity = identity_rows;
  for (StreamPlan::GraphEntry& g : plan->graphs)
    if (g.same(key)) {
      if (cudaGraphLaunch(g.exec, s) != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("stream graph launch failed: ") + cudaGetErrorString(cudaGetLastError()));
      e->launches += (long long)plan->ops.size() + ((!identity_rows && plan->out_len > 0) ? 1 : 0);
      return Q3TTS_OK;
    }
  if (!(plan->last_valid && plan->last.same(key))) {  // first sight of these pointers: plain launches (also the warm-up of lazy attributes)
    plan->last = key; plan->last_valid = true;
    return launch_stream_plan(e, plan, codes, sb, sk, st, d_ids, d_rows, identity_rows, wav, wav_stride, s);
  }
  if (cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal) != cudaSuccess) {
    cudaGetLastError(); plan->graph_off = true;
    return launch_stream_plan(e, plan, codes, sb, sk, st, d_ids, d_rows, identity_rows, wav, wav_stride, s);
  }
  const long long launches0 = e->launches;
  const int rc = launch_stream_plan(e, plan, codes, sb, sk, st, d_ids, d_rows, identity_rows, wav, wav_stride, s);
  cudaGraph_t graph = nullptr;
  const cudaError_t ce = cudaStreamEndCapture(s, &graph);
  cudaGraphExec_t exec = nullptr;
  if (rc != Q3TTS_OK || ce != cudaSuccess || !graph || cudaGraphInstantiate(&exec, graph, 0) != cudaSuccess) {
    cudaGetLastError();
    if (graph) cudaGraphDestroy(graph);
    plan->graph_off = true;  // this plan keeps to plain launches
    e->launches = launches0;
    return launch_stream_plan(e, plan, codes, sb, sk, st, d_ids, d_rows, identity_rows, wav, wav_stride, s);
  }
  cudaGraphDestroy(graph);
  if (plan->graphs.size() >= 4) { cudaGraphExecDestroy(plan->graphs.front().exec); plan->graphs.erase(plan->graphs.begin()); }
  key.exec = exec;
  plan->graphs.push_back(key);
  if (cudaGraphLaunch(exec, s) != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("stream graph launch failed: ") + cudaGetErrorString(cudaGetLastError()));
  return Q3TTS_OK;  // e->launches was advanced by the capture pass
}

int launch_stream_plan(Engine* e, StreamPlan* plan, const long long* codes, long long sb, long long sk, long long st, const int* d_ids, const int* d_rows,
                       bool identity_rows, float* wav, long long wav_stride, cudaStream_t s) {
  const q3tts_config& c = e->cfg;
  PdlScope pdl(!e->profile);  // programmatic dependent launch between the step's ~100 short kernels (launch.cuh); not under per-op events
  float* tail_out = identity_rows ? wav : plan->wav_tmp;
  const long long tail_stride = identity_rows ? wav_stride : plan->out_len;
  static const bool sync_each = Q3_ENV("Q3TTS_SYNC") != nullptr;  // debugging: attribute an asynchronous fault to its op
  size_t pi = e->profile ? prof_begin(e, plan->ops.size()) : 0;  // accumulates until q3tts_get_profile reads it
  for (Op& op : plan->ops) {
    cudaError_t err = cudaSuccess;
    if (e->profile) cudaEventRecord(e->prof_events[2 * pi], s);
    switch (op.type) {
      case OP_STATE_LOAD:
        err = stream_state_copy(plan->d_descs, (int)plan->descs.size(), e->stream_state, e->state_bytes, d_ids, plan->S, 0, s);
        break;
      case OP_STATE_SAVE:
        err = stream_state_copy(plan->d_descs, (int)plan->descs.size(), e->stream_state, e->state_bytes, d_ids, plan->S, 1, s);
        break;
      case OP_POS_ADVANCE: err = stream_advance_pos(e->d_slot_pos, d_ids, plan->S, plan->P, s); break;
      case OP_RVQ:
        err = rvq_decode(codes, sb, sk, st, e->rvq_tables, op.out, nullptr, op.B, c.num_quantizers, op.L, c.codebook_size, c.codebook_dim, e->bf16, s,
                         op.aux0, identity_rows ? nullptr : d_rows, e->d_bad_code, op.split);
        break;
      case OP_GEMM: err = launch_conv_gemm(op.g, ENGINE_TC, s); break;
      case OP_UNIT: err = launch_unit(op.u, s); break;
      case OP_RMSNORM: err = rmsnorm((const float*)op.in, op.w0, op.out, op.B, op.C, c.rms_norm_eps, e->bf16, s, op.split); break;
      case OP_SPLIT: err = split_f32((const float*)op.in, op.sp_in_bs, op.out, op.sp_out_bs, op.B, op.L, op.C, e->bf16, s, op.split); break;
      case OP_STREAM_ATTN:
        err = stream_attention(op.in, e->kcache[op.aux0], e->vcache[op.aux0], d_ids, e->d_slot_pos, e->rope_cos, e->rope_sin, e->rope_max, c.rope_theta,
                               op.out, op.B, op.L, c.num_attention_heads, c.head_dim, c.sliding_window, e->kv_ring, e->bf16, s);
        break;
      case OP_DWLN:
        err = dwconv_layernorm(op.in, op.w0, op.w1, op.w2, op.w3, op.out, op.B, op.L, op.C, 1e-6f, e->bf16, s, op.aux0, op.aux1, op.split != 0);
        break;
      case OP_TAIL:
        if (op.t.enabled) {
          op.t.p.out = tail_out; op.t.p.out_batch_stride = tail_stride; op.t.p.tail_skip = 0;
          err = launch_tail(op.t, s);
        } else {
          op.g.p.out_f32 = tail_out; op.g.p.out_batch_stride = tail_stride; op.g.p.tail_skip = 0;
          err = launch_conv_gemm(op.g, ENGINE_TC, s);
        }
        break;
      default: break;
    }
    ++e->launches;
    if (e->profile) {
      cudaEventRecord(e->prof_events[2 * pi + 1], s);
      prof_account(e, op);
      ++pi;
    }
    if (err == cudaSuccess && sync_each) err = cudaStreamSynchronize(s);
    if (err != cudaSuccess)
      return e->fail(Q3TTS_CUDA_ERROR, "stream launch failed (op " + std::to_string(&op - plan->ops.data()) + " type " + std::to_string((int)op.type) + " S " +
                                           std::to_string(plan->S) + " P " + std::to_string(plan->P) + (plan->fresh ? " fresh" : "") + "): " + cudaGetErrorString(err));
  }
  if (!identity_rows && plan->out_len > 0) {
    cudaError_t err = scatter_rows_f32(plan->wav_tmp, plan->out_len, wav, wav_stride, d_rows, plan->S, plan->out_len, s);
    ++e->launches;
    if (err != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, std::string("stream scatter failed: ") + cudaGetErrorString(err));
  }
  return Q3TTS_OK;
}

}  // namespace

int engine_stream_reserve(Engine* e, int max_streams, int max_frames_per_step) {
  const q3tts_config& c = e->cfg;
  if (max_frames_per_step > e->kv_ring - (c.sliding_window - 1)) return e->fail(Q3TTS_BAD_SHAPE, "stream_reserve: at most 57 frames per step (K/V ring of 128)");
  if (c.sliding_window - 1 + 1 > e->kv_ring) return e->fail(Q3TTS_BAD_CONFIG, "sliding window larger than the K/V ring");
  cudaDeviceSynchronize();
  engine_stream_free(e);
  e->max_slots = max_streams; e->stream_max_frames = max_frames_per_step;
  {  // state layout (independent of S, P): a dry build of the smallest plan
    StreamPlan tmp;
    e->state_bytes = 0;
    build_stream(e, &tmp, 1, 1, false, true, nullptr);
  }
  const size_t kv = (size_t)max_streams * e->kv_ring * c.hidden_size * 2;
  bool ok = cudaMalloc((void**)&e->stream_state, (size_t)max_streams * e->state_bytes) == cudaSuccess;
  ok = ok && cudaMemset(e->stream_state, 0, (size_t)max_streams * e->state_bytes) == cudaSuccess;
  for (int l = 0; l < c.num_hidden_layers && ok; ++l) {
    void *k = nullptr, *v = nullptr;
    ok = cudaMalloc(&k, kv) == cudaSuccess && cudaMalloc(&v, kv) == cudaSuccess;
    if (k) e->kcache.push_back(k);
    if (v) e->vcache.push_back(v);
  }
  ok = ok && cudaMalloc((void**)&e->d_slot_pos, (size_t)max_streams * sizeof(int)) == cudaSuccess;
  ok = ok && cudaMemset(e->d_slot_pos, 0, (size_t)max_streams * sizeof(int)) == cudaSuccess;
  ok = ok && cudaMalloc((void**)&e->d_slot_ids, (size_t)2 * max_streams * sizeof(int)) == cudaSuccess;
  ok = ok && cudaMalloc((void**)&e->d_row_index, (size_t)2 * max_streams * sizeof(int)) == cudaSuccess;
  ok = ok && cudaMallocHost((void**)&e->h_stage, (size_t)Engine::kStageRing * 4 * max_streams * sizeof(int)) == cudaSuccess;
  for (int i = 0; i < Engine::kStageRing && ok; ++i) {
    ok = cudaEventCreateWithFlags(&e->stage_ev[i], cudaEventDisableTiming) == cudaSuccess;
    e->stage_used[i] = false;
  }
  e->stage_next = 0;
  if (!ok) {
    cudaGetLastError();
    engine_stream_free(e);
    return e->fail(Q3TTS_OOM, "stream_reserve: allocation failed");
  }
  e->slot_frames.assign(max_streams, -1);
  // size the arena for the largest step so that no allocation happens later
  StreamPlan* p = nullptr;
  return get_stream_plan(e, max_streams, max_frames_per_step, false, &p);
}

void engine_stream_free(Engine* e) {
  if (e->stream_state) cudaFree(e->stream_state);
  for (void* p : e->kcache) cudaFree(p);
  for (void* p : e->vcache) cudaFree(p);
  if (e->d_slot_pos) cudaFree(e->d_slot_pos);
  if (e->d_slot_ids) cudaFree(e->d_slot_ids);
  if (e->d_row_index) cudaFree(e->d_row_index);
  if (e->h_stage) cudaFreeHost(e->h_stage);
  for (int i = 0; i < Engine::kStageRing; ++i) {
    if (e->stage_ev[i]) cudaEventDestroy(e->stage_ev[i]);
    e->stage_ev[i] = nullptr; e->stage_used[i] = false;
  }
  if (e->sws) cudaFree(e->sws);
  e->stream_state = nullptr; e->kcache.clear(); e->vcache.clear(); e->d_slot_pos = nullptr; e->d_slot_ids = nullptr; e->d_row_index = nullptr;
  e->h_stage = nullptr; e->sws = nullptr; e->sws_bytes = 0; e->splans.clear(); e->slot_frames.clear(); e->max_slots = 0;
}

// One step over S streams (all with P new frames).  Streams on their first packet and continuing streams run as two groups.
int engine_stream_step(Engine* e, const int32_t* stream_ids, int S, const long long* codes, long long sb, long long sk, long long st, int P, float* wav,
                       long long wav_stride, int32_t* n_valid, cudaStream_t s) {
  if (!e->max_slots) return e->fail(Q3TTS_BAD_STREAM, "stream_step: call q3tts_stream_reserve first");
  if (S < 1 || S > e->max_slots || P < 1 || P > e->stream_max_frames) return e->fail(Q3TTS_BAD_SHAPE, "stream_step: S or P outside the reserved range");
  const int ms = e->max_slots;
  // this step's block of the pinned staging ring; the only wait is for the copies that used it kStageRing steps ago
  const int blk = e->stage_next;
  if (e->stage_used[blk] && cudaEventSynchronize(e->stage_ev[blk]) != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, "stream_step: stream error");
  int* hb = e->h_stage + (size_t)blk * 4 * ms;
  int* h_ids[2] = {hb, hb + ms};
  int* h_rows[2] = {hb + 2 * ms, hb + 3 * ms};
  int n[2] = {0, 0};
  // ---- validate everything and build both plans BEFORE anything is launched or any bookkeeping advances
  for (int i = 0; i < S; ++i) {
    const int id = stream_ids[i];
    if (id < 0 || id >= ms || e->slot_frames[id] < 0) return e->fail(Q3TTS_BAD_STREAM, "stream_step: unknown or closed stream id " + std::to_string(id));
    for (int j = 0; j < i; ++j)
      if (stream_ids[j] == id) return e->fail(Q3TTS_BAD_STREAM, "stream_step: duplicate stream id " + std::to_string(id));
    const int g = e->slot_frames[id] == 0 ? 1 : 0;  // group 1 = fresh
    h_ids[g][n[g]] = id; h_rows[g][n[g]] = i; ++n[g];
  }
  StreamPlan* plans[2] = {nullptr, nullptr};
  for (int g = 0; g < 2; ++g)
    if (n[g])
      if (int rc = get_stream_plan(e, n[g], P, g == 1, &plans[g])) return rc;
  // ---- upload the ids (the two groups' id / row arrays are adjacent: two copies per step), then launch
  int* d_ids[2] = {e->d_slot_ids, e->d_slot_ids + ms};
  int* d_rows[2] = {e->d_row_index, e->d_row_index + ms};
  bool up_ok = true;
  for (int g = 0; g < 2 && up_ok; ++g) {
    if (!n[g]) continue;
    up_ok = cudaMemcpyAsync(d_ids[g], h_ids[g], n[g] * sizeof(int), cudaMemcpyHostToDevice, s) == cudaSuccess &&
            cudaMemcpyAsync(d_rows[g], h_rows[g], n[g] * sizeof(int), cudaMemcpyHostToDevice, s) == cudaSuccess;
  }
  if (!up_ok || cudaEventRecord(e->stage_ev[blk], s) != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, "stream_step: id upload failed");
  e->stage_used[blk] = true;
  e->stage_next = (blk + 1) % Engine::kStageRing;
  for (int g = 0; g < 2; ++g) {
    if (!n[g]) continue;
    StreamPlan* plan = plans[g];
    const bool identity = n[g] == S;  // one group holds every row, in order
    if (int rc = run_stream_plan(e, plan, codes, sb, sk, st, d_ids[g], d_rows[g], identity, wav, wav_stride, s)) return rc;
    for (int k = 0; k < n[g]; ++k) {
      if (n_valid) n_valid[h_rows[g][k]] = plan->out_len;
      e->slot_frames[h_ids[g][k]] += P;
    }
  }
  return Q3TTS_OK;
}

}  // namespace q3
