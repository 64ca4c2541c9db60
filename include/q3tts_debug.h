/* libq3tts_decode.so - debug / design-evidence exports.  NOT part of the drop-in boundary (include/q3tts_decode.h is):
 * device-side self-tests of the tcgen05 kernels against the SIMT cross-check, issue-rate and TMA microbenchmarks, and the
 * fused-unit timeline.  Used by tests/ (self-tests) and tools/ only; the Python package binds them through
 * qwen3_tts_b200/_capi.py DEBUG_EXPORTS.  The Q3TTS_* environment switches and GemmParams::dbg experiment bits the tools mention
 * exist only in the -DQ3TTS_DEBUG build (python -m qwen3_tts_b200.build --debug -> lib/libq3tts_decode_dbg.so). */
#ifndef Q3TTS_DEBUG_H_
#define Q3TTS_DEBUG_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Standalone GEMM self-test of the tcgen05 implicit-GEMM kernel against the SIMT cross-check
 * (device-side, no host math): returns max |tc - simt| over a [rows,n] problem via *max_abs_diff. */
int q3tts_selftest_gemm(int device, int rows, int cin, int n, int ntaps, int dilation, int bf16, float* max_abs_diff, float* max_abs_ref);

/* Self-test of the fused residual-unit kernel (conv7 -> Snake -> conv1 + residual in one launch, intermediate kept in
 * tensor memory) against two launches of the SIMT cross-check: out4 = {max|y-y_ref|, max|y_ref|, max|s-s_ref|, max|s_ref|}.
 * Returns Q3TTS_BAD_CONFIG when the channel count cannot be fused. */
int q3tts_selftest_unit(int device, int rows, int channels, int dilation, int bf16, float* out4);

/* Process-wide switch for programmatic dependent launch between the library's kernels (default on): same-box A/B only. */
int q3tts_debug_set_pdl(int on);

/* SnakeBeta exactly as the epilogues evaluate it (cosine form, MUFU.COS): out[i] = (x + h) - h cos(a2 x[i]) with a2 = 2 exp(alpha),
 * h = 0.5 / (exp(beta) + 1e-9), host arrays of n floats.  Lets a test bound the error of __cosf at large arguments. */
int q3tts_selftest_snake(int device, const float* x_host, int n, float a2, float h, float* out_host);

/* Debug timeline of the fused unit kernel (env Q3TTS_DBG bit 5): clock64 stamps [16 tiles][16 events] of CTA 0. */
int q3tts_unit_trace(long long* out256);

/* tcgen05.mma issue-rate microbenchmark (design evidence): SM cycles per MMA (M = 128, K = 16, fp16) for an N, operand form
 * (0 = A,B in shared memory, 1 = A in tensor memory), swizzle span (64 / 128 bytes), A row shift, accumulator count, and a tcgen05.commit / mbarrier wait every k groups of 8 MMAs. */
/* CTA-pair (tcgen05 cta_group::2) building blocks: D[256,n] = A[256,k] . W[n,k]^T on one 2-CTA cluster; ts != 0 also runs
 * the tensor-memory A operand form (k == n <= 128).  iters > 0: throughput run on every SM pair instead of a comparison.
 * out4 = {max |D - ref|, max |D2 - ref2|, cycles per M = 256 MMA, 0}. */
int q3tts_selftest_pair(int device, int n, int k, int iters, int ts, float* out4);
int q3tts_bench_mma(int device, int n, int span, int a_row_shift, int form, int nacc, int commit_every, int wait_every, float* clk_per_mma);

/* TMA box-shape microbenchmark (design evidence, DESIGN.md "Epilogue boxes"): aggregate GB/s of `warps` warps per SM moving
 * [box_rows x box_cols] 16-bit boxes between shared memory and a channels-last [rows][C] tensor. */
int q3tts_bench_tma(int device, int C, int box_cols, int box_rows, int is_load, int warps, int iters, float* gbps);

#ifdef __cplusplus
}
#endif
#endif /* Q3TTS_DEBUG_H_ */
