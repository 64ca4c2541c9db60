// Bandwidth-bound SIMT kernels of the decode path and the load-time weight packers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace q3 {

// ---- load-time packers (fp32 reference layouts -> 16-bit K-major GEMM operands)
// Conv1d weight [Cout][Cin][k] -> [k][Cout][Cin]
// `split` (every packer below): 1 = one 16-bit value per weight; 2 / 3 = weights carried as hi + lo along K for the split-precision
// GEMMs (gemm_params.h): rows become split * cin wide, [hi | lo] or [hi | hi | lo].
// `group` (pack_conv_weight, pack_rows, pack_strided_conv_weight; 0 = the whole row): K group size in input channels - every group of a
// row is packed on its own ([hi(g) | hi(g) | lo(g)] ...), so that one launch can walk one group (encode_plan.cu: short accumulation chains)
cudaError_t pack_conv_weight(const float* src, void* dst, int cout, int cin, int k, bool bf16, cudaStream_t s, int split = 1, int group = 0);
// ConvTranspose1d weight [Cin][Cout][k], stride st.  k == 2*st: two taps, tap0 (x[q]) = W[..., p+st],
// tap1 (x[q+1]) = W[..., p]; k == st: one tap W[..., p].  dst [taps][st*Cout][Cin], row = p*Cout + co.
cudaError_t pack_transconv_weight(const float* src, void* dst, int cin, int cout, int k, int st, bool bf16, cudaStream_t s, int split = 1);
// rows of `a` and `b` ([n][k] each) interleaved: dst[2i] = a[i], dst[2i+1] = b[i]
cudaError_t pack_interleave2(const float* a, const float* b, void* dst, int n, int k, bool bf16, cudaStream_t s, int split = 1);
// [rows][k] fp32 -> [rows][split * k] 16-bit
cudaError_t pack_rows(const float* src, void* dst, long long rows, int k, bool bf16, cudaStream_t s, int split = 1, int group = 0);
// Conv1d weight [Cout][C][2r] of a stride-r causal conv -> [2][Cout][split * r * C]: two taps over super-rows of r input rows
// (encode path: tap = j / r, K index (j % r) * C + ci)
cudaError_t pack_strided_conv_weight(const float* src, void* dst, int cout, int c, int r, bool bf16, cudaStream_t s, int split = 1, int group = 0);
// plain cast [n] fp32 -> 16-bit
cudaError_t pack_cast(const float* src, void* dst, long long n, bool bf16, cudaStream_t s);
// Snake: a = exp(alpha), ib = 1/(exp(beta)+1e-9)
cudaError_t pack_snake(const float* alpha, const float* beta, float* a, float* ib, int n, cudaStream_t s);
// RVQ: folded[code][d] = sum_j (embedding_sum[code][j] / max(cluster_usage[code],1e-5)) * Wout[d][j]   (fp32)
cudaError_t fold_codebook(const float* embedding_sum, const float* cluster_usage, const float* wout, float* folded, int ncodes, int dq, int d,
                          cudaStream_t s);

// ---- hot-path SIMT kernels
// codes int64 addressed as codes[b*sb + k*sk + t*st]; tables [Q][ncodes][D] fp32; out_act [B][T][D] 16-bit (optional);
// out_f32 [B][D][T] (optional)
// int16 code frames (packed codes on the wire) -> the int64 layout the gather kernel reads
cudaError_t widen_codes_i16(const short* in, long long* out, long long n, cudaStream_t s);
cudaError_t rvq_decode(const long long* codes, long long sb, long long sk, long long st, const float* tables, void* out_act, float* out_f32, int B,
                       int Q, int T, int ncodes, int D, bool bf16, cudaStream_t s, int out_bs_rows = 0, const int* src_index = nullptr,
                       int* bad_code = nullptr,   // bad_code (device int, optional): set to 1 when a code is outside [0, ncodes)
                       int split = 0);            // 1: out_act rows are [hi | lo] (2 * D wide): the split-precision A operand of pre_conv
// x fp32 [M][H] -> 16-bit [M][H]: w * x * rsqrt(mean(x^2) + eps)
// split = 1: out rows are [hi | lo] (2 * H wide)
cudaError_t rmsnorm(const float* x, const float* w, void* out, int M, int H, float eps, bool bf16, cudaStream_t s, int split = 0);
// fp32 [B][rows][C] (in_bs elements between items) -> 16-bit [B][.][2C] rows [hi | lo] (out_bs elements between items): turns an fp32
// main-path tensor into the split-precision A operand of the next GEMM
// with_lo = 0: plain 16-bit rows (C wide): the fp32 tensor rounded once, for a consumer with plain operands
cudaError_t split_f32(const float* in, long long in_bs, void* out, long long out_bs, int B, int rows, int C, bool bf16, cudaStream_t s, int with_lo = 1);
// rope table [maxT][hd/2] cos, sin (fp32)
cudaError_t rope_table(float* cos_t, float* sin_t, int maxT, int hd, float theta, cudaStream_t s);
// qkv 16-bit [B*T][3*nh*hd] (q | k | v), RoPE applied on the fly with position = t; causal, kv > q - window.
cudaError_t sliding_attention(const void* qkv, const float* cos_t, const float* sin_t, void* out, int B, int T, int nh, int hd, int window,
                              bool bf16, cudaStream_t s);
// ConvNeXt front half: depthwise causal conv k=7 + bias, LayerNorm(eps) over C.  x,out 16-bit [B][L][C]; dw weight [7][C] fp32.
cudaError_t dwconv_layernorm(const void* x, const float* dww, const float* dwb, const float* lnw, const float* lnb, void* out, int B, int L, int C,
                             float eps, bool bf16, cudaStream_t s, int in_bs_rows = 0, int hist = 0, bool x_f32 = false);  // x_f32: x is fp32
// per-device opt-in to > 48 KB dynamic shared memory for the attention kernels (called once per handle by engine_finalize)
cudaError_t attention_setup();
cudaError_t stream_attention_setup();
// ---- streaming (stateful per-packet decode; no counterpart in the reference, oracle = Decoder.forward on the whole sequence)
// qkv 16-bit [S*P][3*nh*hd] of the P new tokens of S streams; K (RoPE applied) and V are appended to the per-slot ring caches
// kc / vc [slots][ring][nh*hd] at (pos & (ring-1)); query i of stream s sits at absolute position slot_pos[slot_ids[s]] + i and sees
// keys kv <= q && kv > q - window.  ring >= window - 1 + P.
cudaError_t stream_attention(const void* qkv, void* kc, void* vc, const int* slot_ids, const int* slot_pos, const float* cos_t, const float* sin_t,
                             int rope_max, float theta, void* out, int S, int P, int nh, int hd, int window, int ring, bool bf16, cudaStream_t s);
// History rows of the streaming buffers <-> per-slot state.  descs (device): one StreamBufDesc per buffer.
struct StreamBufDesc {
  unsigned char* ptr;        // buffer base: [S][H + n][row_bytes]
  long long batch_stride;    // bytes between streams
  long long state_off;       // byte offset of this buffer's history inside a slot's state block
  int H, n, row_bytes, pad;
};
// save = 0: state -> rows [0, H) of every buffer;  save = 1: rows [n, n + H) -> state
cudaError_t stream_state_copy(const StreamBufDesc* descs, int ndesc, unsigned char* state, long long state_bytes, const int* slot_ids, int S, int save,
                              cudaStream_t s);
cudaError_t stream_advance_pos(int* slot_pos, const int* slot_ids, int S, int P, cudaStream_t s);
// out[index[b]][0:len] = in[b][0:len]  (fp32 rows)
cudaError_t scatter_rows_f32(const float* in, long long in_stride, float* out, long long out_stride, const int* index, int B, long long len, cudaStream_t s);
// probes: activation [B][L][C] (16-bit or fp32) -> fp32 [B][C][L]
cudaError_t to_channels_first_f32(const void* x, bool is_f32, float* out, int B, int L, int C, bool bf16, cudaStream_t s, int ld = 0);  // ld: row pitch (0 = C)
// dw weight [C][1][7] -> [7][C]
cudaError_t pack_dw_weight(const float* src, float* dst, int C, int k, cudaStream_t s);

}  // namespace q3
