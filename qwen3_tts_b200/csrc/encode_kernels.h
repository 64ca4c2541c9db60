// SIMT kernels of the 12 Hz encode path (encode_kernels.cu); the GEMMs between them are conv_gemm.cu's tensor-core kernels.
#pragma once
#include <cuda_runtime.h>

namespace q3 {

// first conv 1 -> C0 (k = 7, causal): wav [B][wav_bs] fp32 -> out [B][T][C0] fp32
// elu_split (optional): also ELU(out) as fp16 [B][T][2 C0] rows [hi | lo], the first ResnetBlock's conv operand
cudaError_t enc_conv0(const float* wav, long long wav_bs, const float* w, const float* bias, float* out, void* elu_split, int B, int T, int C0, cudaStream_t s);
// act(in) fp32 [B][L][C] -> fp16 [B][lead + ceil(L / r)][2 r C] super-rows [hi | lo]; act 0 none / 1 ELU / 2 GELU(erf); rows outside
// [0, L) are zero, or the nearest valid row when replicate != 0; cg (0 = r * C): K group size, every group its own [hi | lo] pair
cudaError_t enc_act_split(const float* in, void* out, int B, int L, int C, int act, int r, int lead, int replicate, int cg, cudaStream_t s);
// LayerNorm with bias over H, fp32 [M][H] -> fp16 [M][2H] rows [hi | lo]
cudaError_t enc_layernorm_split(const float* x, const float* w, const float* bias, void* out, long long M, int H, float eps, cudaStream_t s);
// cos / sin [maxT][hd / 2] from float32 inverse frequencies (device array of hd / 2)
cudaError_t enc_rope_table(float* cos_t, float* sin_t, const float* inv_freq_dev, int maxT, int hd, cudaStream_t s);
// causal sliding-window attention in fp32 with RoPE: qkv fp32 [B*T][3 nh hd] -> out fp16 [B*T][2 nh hd] rows [hi | lo]
cudaError_t enc_attention(const float* qkv, const float* cos_t, const float* sin_t, void* out, int B, int T, int nh, int hd, int window, cudaStream_t s);
// RVQ search, second half of a layer: S fp32 [M][NC] = r . e (from a GEMM) -> code = argmin (n2 - 2 S), lowest index on ties, stored at
// codes[(m / F) * code_stride_b + (m % F) * code_stride_m + code_off]; r[m][0:dq] (row pitch ldr) -= table[code]; a_next (optional):
// the new residual as fp16 [M][2 dq] rows [hi | lo]
cudaError_t enc_argmin_update(const float* S, const float* n2, const float* table, float* r, int ldr, void* a_next, long long* codes,
                              long long code_stride_m, long long code_stride_b, long long code_off, int M, int F, int NC, int dq, cudaStream_t s);
// fp32 [M][ld] columns [c0, c0 + C) -> fp16 [M][2C] rows [hi | lo]
cudaError_t enc_split_cols(const float* in, int ld, int c0, void* out, long long M, int C, cudaStream_t s);
// the same attention with both contractions on mma.sync tensor cores, every operand an fp16 hi + lo pair (three MMAs per product)
// qkvs: fp16 [B*T][6 nh hd] = [q_hi | q_lo | k_hi | k_lo | v_hi | v_lo], RoPE applied (enc_rope_split)
cudaError_t enc_attention_mma(const void* qkvs, void* out, int B, int T, int nh, int hd, int window, cudaStream_t s);
// qkv fp32 [M][3 nh hd] -> the fp16 operand rows above (M = B * T rows, position = row % T)
cudaError_t enc_rope_split(const float* qkv, const float* cos_t, const float* sin_t, void* out, long long M, int T, int nh, cudaStream_t s);
cudaError_t enc_attention_setup();   // per-device opt-in to its dynamic shared memory (encoder_front_finalize)
// fp32 [B][L][C] -> fp32 [B][C][L]
cudaError_t enc_to_channels_first(const float* in, float* out, int B, int L, int C, cudaStream_t s);

}  // namespace q3
