// Host-side engine: weight store, per-(B,T) launch plans, workspace.
#pragma once
#include <cuda_runtime.h>

#include <map>
#include <string>
#include <utility>
#include <vector>

#include "../../include/q3tts_decode.h"
#include "conv_gemm.h"
#include "kernels.h"

namespace q3 {

struct RawTensor {
  std::vector<int64_t> shape;
  float* d = nullptr;  // fp32 device copy
  long long numel = 0;
};

// One packed conv / linear operand set for the implicit-GEMM kernel.
struct ConvW {
  void* w = nullptr;  // [taps][n_total][cin] 16-bit
  int taps = 0, n_total = 0, cin = 0, ncol_mod = 0;
  int tap_off[8] = {0};
  const float* bias = nullptr;  // fp32 [ncol_mod] or null
  int bn = 0, kb = 0;
  int out_rows_per_row = 1;  // transposed conv: s output rows per accumulator row
  // split-precision operands (gemm_params.h): cin above is the K the MMAs walk = k_mult x the layer's input channels;
  // a_cols = width of the A tensor in memory (C for k_mult 2, 2C = [hi | lo] rows for k_mult 3), 0 = plain
  int k_mult = 1, a_cols = 0;
  // encode path: K group size in real input values (0 = one group): the layer runs as kreal / cg launches that accumulate into the
  // fp32 output, so that no tensor-core accumulation chain is longer than taps * 3 * cg (encode_plan.cu)
  int cg = 0;
  int gl = 1;   // K groups per launch (the kernel walks them through GemmParams.kgroup)
};

struct SnakeP { const float* a = nullptr; const float* ib = nullptr; };

struct LayerW {
  ConvW qkv, o, gateup, down;
  const float *ln1 = nullptr, *ln2 = nullptr, *ls1 = nullptr, *ls2 = nullptr;
};
struct UpsampleW {
  ConvW tconv, pw1, pw2;
  const float *dww = nullptr, *dwb = nullptr, *lnw = nullptr, *lnb = nullptr, *gamma = nullptr;
};
struct UnitW { ConvW conv7, conv1; SnakeP act1, act2; };
struct BlockW { SnakeP act0; ConvW tconv; UnitW unit[3]; int stride = 0, cin = 0, cout = 0; };

// ---- 12 Hz encoder (encode_plan.cu): HF MimiModel's encoder side as the reference instantiates it (modeling :898-907)
struct EncStageW { ConvW c3, c1, down; int C = 0, r = 0; };   // ResnetBlock (k3: C -> C/2, k1: C/2 -> C) + stride-r conv C -> 2C
struct EncLayerW {
  ConvW qkv, o, fc1, fc2;
  const float *ln1w = nullptr, *ln1b = nullptr, *ln2w = nullptr, *ln2b = nullptr, *ls1 = nullptr, *ls2 = nullptr;
};

enum OpType { OP_RVQ = 0, OP_GEMM, OP_RMSNORM, OP_ATTN, OP_DWLN, OP_TAIL, OP_UNIT, OP_STREAM_ATTN, OP_STATE_LOAD, OP_STATE_SAVE, OP_POS_ADVANCE,
              OP_SPLIT /* fp32 [B][L][C] -> 16-bit [hi | lo] rows: in, out, B, L, C, sp_in_bs / sp_out_bs */,
              // encode path (encode_plan.cu): first conv; act + split (aux0 = act, aux1 = r, split = lead rows | replicate << 8, kgroup);
              // LayerNorm -> [hi | lo]; fp32 attention; RVQ search (aux0 = 1: codes [B][T][Q])
              OP_ENC_CONV0, OP_ENC_ACT, OP_ENC_LN, OP_ENC_ATTN, OP_ENC_RVQ,
              // RVQ search on the tensor cores: split a column range of the projected latents (aux0 = first column, aux1 = row pitch);
              // argmin over a layer's scores + residual update (aux0 = codebook, aux1 = residual column offset, w0 = scores, w1 = residuals)
              OP_ENC_SPLITCOLS, OP_ENC_ARGMIN };

struct Op {
  OpType type = OP_GEMM;
  GemmOp g;
  UnitOp u;  // OP_UNIT: fused residual unit (conv7 -> Snake -> conv1 + residual)
  TailOp t;  // OP_TAIL: taps-as-N output conv (used when enabled, else the N = 16 implicit GEMM in g)
  const void* in = nullptr;
  void* out = nullptr;
  const float *w0 = nullptr, *w1 = nullptr, *w2 = nullptr, *w3 = nullptr;
  int B = 0, L = 0, C = 0;
  bool drop_raw = false;   // last residual unit of a decoder block: its raw output is dead (only the Snake'd copy is read), skipped unless probing
  int split = 0;           // OP_RVQ / OP_RMSNORM: write [hi | lo] rows;  OP_DWLN: the input tensor is fp32
  long long sp_in_bs = 0, sp_out_bs = 0;  // OP_SPLIT: elements between batch items of in / out
  int probe_ld = 0;        // row pitch of the probe buffer in elements (0 = probe_C)
  void* out2 = nullptr;    // encode path, OP_ENC_CONV0: second output (ELU(x) as [hi | lo] rows)
  int kgroup = 0;          // encode path, OP_ENC_ACT: K group size of the layer that reads the result (0 = one group)
  int aux0 = 0, aux1 = 0;  // streaming: RVQ output batch stride in rows / dwconv (input batch stride in rows, history rows) / attention layer
  // probe: the buffer that holds an oracle tap once this op has run
  std::string tag;
  const void* probe_ptr = nullptr;
  bool probe_f32 = false;
  int probe_C = 0, probe_L = 0;
};

struct Plan {
  int B = 0, T = 0;
  std::vector<Op> ops;
  int out_len = 0;  // samples per item of the forward output
};

// One launch plan of the streaming path: S streams x P new frames, `fresh` = first packet of every stream in the group.
struct StreamPlan {
  int S = 0, P = 0;
  bool fresh = false;
  std::vector<Op> ops;
  std::vector<StreamBufDesc> descs;
  StreamBufDesc* d_descs = nullptr;  // device copy, the plan's own allocation: the arena is scratch shared by every plan
  int out_len = 0;                    // samples emitted per stream
  float* wav_tmp = nullptr;           // dense [S][out_len] staging when the caller's rows are not the identity
  // CUDA graphs of this plan's launch sequence, one per set of caller pointers (a step is ~96 small launches: launch-bound at few streams).
  // A key seen twice in a row is captured; at most four are kept.
  struct GraphEntry {
    const void* codes = nullptr; long long sb = 0, sk = 0, st = 0;
    float* wav = nullptr; long long stride = 0; bool identity = false;
    cudaGraphExec_t exec = nullptr;
    bool same(const GraphEntry& o) const {
      return codes == o.codes && sb == o.sb && sk == o.sk && st == o.st && wav == o.wav && stride == o.stride && identity == o.identity;
    }
  };
  std::vector<GraphEntry> graphs;
  GraphEntry last;
  bool last_valid = false, graph_off = false;
  StreamPlan() = default;
  StreamPlan(const StreamPlan&) = delete;
  StreamPlan& operator=(const StreamPlan&) = delete;
  ~StreamPlan() {
    for (GraphEntry& g : graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
    if (d_descs) cudaFree(d_descs);
  }
};

struct Engine {
  q3tts_config cfg{};
  int device = 0;
  std::string err;
  bool finalized = false;
  bool bf16 = false;
  // Split-precision level (DESIGN.md "Precision"): 0 = every operand one 16-bit value (the round-1 scheme); 1 = the main signal
  // path of the early stages (RVQ -> pre_conv -> in_proj, final norm -> out_proj -> upsample convs -> decoder[0]) runs on hi + lo
  // operands and keeps its ConvNeXt tensors in fp32; 2 = the transformer's and ConvNeXt's own GEMMs too.
  int split_level = 1;
  int engine = ENGINE_TC;
  int total_up = 1920;
  int lookahead = 555;
  long long launches = 0;
  // code range checking (reference: F.embedding asserts on an out-of-range index, modeling :675-678): the gather kernel raises this
  // device flag; q3tts_decode_host always reads it (it synchronises anyway), the device-pointer entries read it when check_codes is on
  int* d_bad_code = nullptr;
  int* h_bad_code = nullptr;  // pinned
  bool check_codes = false;
  bool profile = false;                 // record a CUDA event pair around every op
  std::vector<cudaEvent_t> prof_events;  // 2 per op of the last profiled run
  std::vector<int> prof_type;
  std::vector<double> prof_flops;       // algorithmic FLOPs (2 * MACs) of GEMM ops, 0 otherwise
  std::vector<int> prof_dims;           // 4 per op: rows (all batch items), n_total, cin, ntaps
  std::vector<double> prof_bytes;       // minimum HBM bytes of the op (operands + outputs once)

  std::map<std::string, RawTensor> raw;
  std::vector<void*> owned;  // packed allocations

  // packed weights
  float* rvq_tables = nullptr;  // [Q][codes][D] fp32, output_proj folded in
  ConvW pre_conv, in_proj, out_proj, dec0, tail;
  std::vector<LayerW> layers;
  const float* final_norm = nullptr;
  std::vector<UpsampleW> ups;
  std::vector<BlockW> blocks;
  SnakeP tail_act;
  float* tail_w = nullptr;  // [7][C]
  float tail_bias = 0.f;
  void* tail_w16 = nullptr;  // [16][C] 16-bit, rows 0..6 = the 7 taps (taps-as-N form of the output conv)
  int tail_c = 0;
  float *rope_cos = nullptr, *rope_sin = nullptr;
  int rope_max = 0;

  // workspace
  uint8_t* ws = nullptr;
  size_t ws_bytes = 0;
  int max_frames = 8192;
  std::map<std::pair<int, int>, Plan> plans;
  long long* codes_stage = nullptr;  // device staging for decode_host
  size_t codes_stage_elems = 0;
  float* wav_stage = nullptr;
  bool out_pcm16 = false;       // transient (decode_host_pcm16): the `wav` argument of the run functions is an int16 buffer
  short* codes16_stage = nullptr;  // device staging for int16 code frames
  size_t codes16_stage_elems = 0;
  size_t wav_stage_elems = 0;
  long long* item_skip_stage = nullptr;  // device copy of decode_host's per-item skip_samples
  size_t item_skip_elems = 0;

  // ---- encoder quantiser (q3tts_rvq_encode): packed when `encoder.quantizer.*` tensors were loaded
  float* enc_tables = nullptr;  // [Q][codes][codebook_dim / 2] fp32: embed_sum / clamp(cluster_usage, 1e-5)
  float* enc_norms = nullptr;   // [Q][codes] squared norms
  float* enc_win = nullptr;     // [2][codebook_dim / 2][codebook_dim] input_proj of the semantic / acoustic group
  bool enc_ready = false;

  // ---- encoder front (q3tts_encode): packed when the `encoder.encoder.*` tensors were loaded
  q3tts_encoder_config enc_cfg{8, 250, 8000, 10000.0f, 1e-5f};   // MimiConfig defaults
  bool enc_front_ready = false;
  int enc_c0 = 0, enc_hidden = 0, enc_inter = 0;
  const float *enc_conv0_w = nullptr, *enc_conv0_b = nullptr;
  std::vector<EncStageW> enc_stages;
  ConvW enc_last, enc_down;
  ConvW enc_qin;                 // semantic | acoustic input_proj stacked: 512 -> 2 x 256
  std::vector<ConvW> enc_cb;     // per codebook: [codes][dq] as a GEMM operand (scores = residual . code)
  std::vector<EncLayerW> enc_layers;
  float *enc_rope_cos = nullptr, *enc_rope_sin = nullptr;
  uint8_t* enc_ws = nullptr;
  size_t enc_ws_bytes = 0;
  long long enc_max_samples = 16ll << 20;      // samples per launch (B * T): bounds the workspace (~1 KB per sample)
  std::map<std::pair<int, int>, Plan> enc_plans;  // (B, T samples)
  float* enc_wav_stage = nullptr;              // device staging of q3tts_encode_host
  size_t enc_wav_stage_elems = 0;
  long long* enc_codes_stage = nullptr;
  size_t enc_codes_stage_elems = 0;

  // ---- streaming state (q3tts_stream_*)
  int max_slots = 0, stream_max_frames = 0, kv_ring = 128;
  long long state_bytes = 0;             // history rows of every stage, per slot
  unsigned char* stream_state = nullptr;  // [max_slots][state_bytes]
  std::vector<void*> kcache, vcache;      // per layer: [max_slots][kv_ring][hidden] 16-bit
  int* d_slot_pos = nullptr;              // [max_slots] frames decoded so far
  int* d_slot_ids = nullptr;              // [2][max_slots] staging of the step's slot ids (continuing group, fresh group)
  int* d_row_index = nullptr;             // [2][max_slots] caller row of every group member
  // pinned host staging for the two arrays above: a ring of kStageRing blocks, each guarded by an event recorded behind its
  // H2D copies, so a step never waits for the stream (only for the copy that used the same block kStageRing steps ago)
  static constexpr int kStageRing = 4;
  int* h_stage = nullptr;                 // [kStageRing][4][max_slots]
  cudaEvent_t stage_ev[kStageRing] = {nullptr, nullptr, nullptr, nullptr};
  bool stage_used[kStageRing] = {false, false, false, false};
  int stage_next = 0;
  std::vector<int> slot_frames;           // host mirror of d_slot_pos; -1 = closed
  uint8_t* sws = nullptr;                 // streaming workspace arena
  size_t sws_bytes = 0;
  std::map<std::pair<std::pair<int, int>, int>, StreamPlan> splans;  // ((S, P), fresh)

  int fail(int code, const std::string& msg) { err = msg; return code; }
};

int engine_finalize(Engine* e, int mode);
int engine_get_plan(Engine* e, int B, int T, Plan** out);
size_t engine_plan_bytes(const Engine* e, int B, int T);
// wav: base of the output rows; this forward's samples land at column out_pos0 - item_skip[b] (item_skip: device array or null)
int engine_run(Engine* e, Plan* plan, const long long* codes, long long sb, long long sk, long long st, float* wav, long long wav_stride,
               int skip, const char* probe_stage, float* probe_out, long long probe_cap, int* pc, int* pl, cudaStream_t s,
               long long out_pos0 = 0, const long long* item_skip = nullptr);
int engine_forward_batched(Engine* e, const long long* codes, long long sb, long long sk, long long st, int B, int T, float* wav,
                           long long wav_stride, int skip, cudaStream_t s, long long out_pos0 = 0, const long long* item_skip = nullptr);
void engine_free(Engine* e);
// Reads and clears the bad-code flag; synchronises the stream.  Returns Q3TTS_CODE_OUT_OF_RANGE when a code was out of range.
int engine_check_codes(Engine* e, cudaStream_t s);
size_t prof_begin(Engine* e, size_t nops);
void prof_clear(Engine* e);
void prof_account(Engine* e, const Op& op);
int engine_stream_reserve(Engine* e, int max_streams, int max_frames_per_step);
int engine_stream_step(Engine* e, const int32_t* stream_ids, int S, const long long* codes, long long sb, long long sk, long long st, int P, float* wav,
                       long long wav_stride, int32_t* n_valid, cudaStream_t s);
void engine_stream_free(Engine* e);
int encode_finalize(Engine* e);  // rvq_encode.cu
int encoder_front_finalize(Engine* e);  // encode_plan.cu: packs encoder.encoder.* / encoder_transformer.* / downsample (no-op when absent)
void encoder_free(Engine* e);
// rvq_encode.cu: latents channels-first [B][D][T] (x_cl = 0) or channels-last [B][T][D] (x_cl = 1) -> codes [B][Q][T] or (codes_tq = 1) [B][T][Q]
cudaError_t rvq_encode_launch(Engine* e, const float* latents, int B, int T, long long* codes, int x_cl, int codes_tq, cudaStream_t s);
int static_scratch_setup(int device);  // noalloc.cu: device address of the static scratch array on the current device

struct DeviceGuard {
  int prev = -1;
  explicit DeviceGuard(int dev) {
    cudaGetDevice(&prev);
    if (prev != dev) cudaSetDevice(dev);
    else prev = -1;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace q3

struct q3tts_handle {
  q3::Engine e;
};
