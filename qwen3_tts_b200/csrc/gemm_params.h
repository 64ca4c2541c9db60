// Parameters of the implicit-GEMM causal conv / transposed conv / linear kernel.
//
//   D[b, t, n] = sum_{tap} sum_{ci} A[b, t + tap_off[tap], ci] * W[tap, n, ci]         (fp32 accumulate)
//
// A is a channels-last activation [batches][rows][cin]; rows outside [0, rows) read as zero
// (TMA out-of-bounds fill), which IS the reference's causal left padding
// (modeling_qwen3_tts_tokenizer_v2.py:189-192).  A k-tap dilated causal conv uses
// tap_off = -(k-1-j)*dil; a stride-s, kernel-2s transposed conv with the reference's symmetric
// trim (modeling :200-207) is the same GEMM with two taps {0, +1} and N = s*Cout, because
//   y[q*s + p, co] = x[q] . W[:, co, p + s] + x[q+1] . W[:, co, p]
// so one accumulator row q holds s consecutive output rows (out_row_stride = s*Cout).
//
// Epilogue, applied per element in this order:
//   v = acc + bias[n % ncol_mod]; act (GELU erf | GLU silu(g)*u on interleaved column pairs);
//   v *= scale[n % ncol_mod]; v += res[b, t, n];
//   out_raw / out_f32 <- v;  out_snake <- v + snake_ib[c] * sin^2(snake_a[c] * v)   (SnakeBeta of the NEXT layer,
//   modeling :601-615; evaluated in the cosine form of epilogue.cuh snake() from the tables snake_a = 2 exp(alpha),
//   snake_ib = (1/2) / (exp(beta)+1e-9) precomputed at load).
#pragma once
#include <stdint.h>

#include "debug_env.h"

namespace q3 {

enum { ACT_NONE = 0, ACT_GELU = 1, ACT_GLU = 2, ACT_TAIL = 3, ACT_ELU_SPLIT = 4, ACT_GELU_SPLIT = 5 };
__host__ __device__ inline bool act_is_split(int act) { return act == ACT_ELU_SPLIT || act == ACT_GELU_SPLIT; }
// ACT_ELU_SPLIT / ACT_GELU_SPLIT (encode path): x = acc + bias (+ fp32 residual) and f(x), f = ELU or erf-GELU, leaves as an fp16 pair
// hi = fp16(f(x)), lo = fp16(f(x) - hi) in the operand layout of the next split-precision GEMM: out_raw is that tensor, rows of
// 2 * n_total values (pair_row_stride), column n + (n / pair_group) * pair_group for hi and pair_group further for lo (pair_group = the
// consumer's K group size, 0 = n_total: [hi(all) | lo(all)]).  out_snake must be non-null (it only selects the two-output paths).  Without fp32 tensors it is a 16-bit epilogue; with an fp32 residual and / or out_f32 (which
// receives x itself, before the ELU) it is the fp32 epilogue with one more 16-column store per half (epi_pair16).
// ACT_TAIL: accumulator column 0 + bias[0], clamp(-1, 1), stored as fp32 to out_f32[b*out_batch_stride + row - tail_skip]
// for row >= tail_skip (the 96 -> 1 output conv + clamp, reference modeling :862, :883; skip = chunk left context).

struct GemmParams {
  int batches;
  int rows;        // A rows per batch item (streaming: history rows + new rows)
  int a_row0;      // A row of accumulator row 0 (streaming: the history rows in front of the new ones; 0 otherwise)
  int rows_valid;  // accumulator rows per batch item that are stored
  int cin;
  int ntaps;
  int tap_off[8];
  int n_total;
  int bn;  // N tile, multiple of 16, <= 256, divides n_total
  int kb;  // K block in elements: 64 (SWIZZLE_128B) or 32 (SWIZZLE_64B)
  int ncol_mod;
  int act;
  int res_f32;
  int bf16;
  int tail_skip;
  int epi_warps;  // 4 or 8 epilogue warps (8: two per TMEM lane quadrant)
  int res_depth;  // residual prefetch ring depth per warp (2, 4 or 8 slots)
  // Packed tiles (streaming steps: many batch items with few rows each).  A 128-row tile is four 32-row chunks; chunk gc maps to
  //   mode 1 (rows_valid <= 16, padded to a power of two 1 << pack_rshift): 32 >> pack_rshift whole batch items, rows [0, 1 << pack_rshift)
  //   mode 2: batch item gc / pack_cpb, rows [(gc % pack_cpb) * 32, +32)
  // so a tile is filled with real rows instead of one batch item's few rows plus padding.  0 = one batch item per tile.
  int pack_mode, pack_rshift, pack_cpb;
  // Split-precision operands (more mantissa than one 16-bit value where it is cheap: DESIGN.md "Precision").  cin is the K the
  // MMAs walk (k_mult x the layer's real input channels); the A tensor in memory is a_cols wide and K blocks wrap around it:
  //   k_mult 2, a_cols = C  :  A = a            W' = [w_hi | w_lo]           (weights carried as hi + lo)
  //   k_mult 3, a_cols = 2C :  A = [a_hi | a_lo] W' = [w_hi | w_hi | w_lo]   (a_hi w_hi + a_lo w_hi + a_hi w_lo: ~21 bits each side)
  // a_cols = 0: plain (a_cols = cin, k_mult = 1).
  int a_cols, k_mult;
  // kgroup > 0 (k_mult 3 only): A rows are GROUPS of kgroup input values, each its own pair [hi(g) | lo(g)], and weight rows the matching
  // [hi(g) | hi(g) | lo(g)] (kernels.cu put_w): K' block k walks group k / (3 kgroup), A column g * 2 kgroup + (k % (3 kgroup)) % (2 kgroup).
  // a_cols is then the width of the groups this launch walks (2 kgroup * number of groups).  0: one group = the whole row.
  int kgroup;
  int skip_raw;  // fused unit kernel: y is not stored (last unit of a block: only snake_next(y) is read downstream)
  int dbg;  // debug experiments (env Q3TTS_DBG): bit 0 = epilogue only drains TMEM, bit 1 = MMA thread issues no MMAs
  const float* bias;
  const float* scale;
  const void* res;
  void* out_raw;
  float* out_f32;
  void* out_snake;
  const float* snake_a;
  const float* snake_ib;
  long long out_batch_stride;  // elements
  long long a_batch_stride;    // elements between batch items of A (0: rows * a_row_stride)
  long long a_row_stride;      // elements between rows of A (0: dense, = a_cols or cin); > a_cols when the launch walks one K group of wider rows
  long long w_row_stride;      // elements between rows of W (0: dense, = cin)
  long long res_batch_stride;  // residual tensor (0: same as out_batch_stride)
  long long snake_batch_stride;  // Snake'd output tensor (0: same as out_batch_stride)
  long long pair_row_stride;     // ACT_ELU_SPLIT: row pitch / batch pitch of the 16-bit pair tensor (0: out_row_stride / out_batch_stride);
  long long pair_batch_stride;   //   they differ from the fp32 output's when both are written
  int pair_group;                // split acts: K group size of the consumer (0 = n_total)
  long long out_row_stride;    // elements
  // raw operand pointers, used only by the SIMT cross-check kernel
  const void* a_ptr;
  const void* w_ptr;
};

// number of 128-row tiles over all batch items
__host__ __device__ inline int gemm_m_tiles_total(const GemmParams& p) {
  if (p.pack_mode == 1) {
    const int per_chunk = 32 >> p.pack_rshift;
    return ((p.batches + per_chunk - 1) / per_chunk + 3) / 4;
  }
  if (p.pack_mode == 2) return (p.batches * p.pack_cpb + 3) / 4;
  return p.batches * ((p.rows_valid + 127) / 128);
}
// 32-row chunk q of packed tile `mtile` -> first accumulator row and first batch item
__host__ __device__ inline void gemm_chunk_coords(const GemmParams& p, int mtile, int q, int& row0, int& b) {
  const int gc = mtile * 4 + q;
  if (p.pack_mode == 1) { row0 = 0; b = gc << (5 - p.pack_rshift); }
  else { b = gc / p.pack_cpb; row0 = (gc - b * p.pack_cpb) * 32; }
}

// ACT_ELU_SPLIT inside the fp32 epilogue flow (fp32 residual in and / or fp32 out next to the 16-bit pair)
// hi column of accumulator column n in the pair tensor (lo sits pair_group further)
__host__ __device__ inline int pair_hi_col(const GemmParams& p, int n) { return p.pair_group ? n + (n / p.pair_group) * p.pair_group : n; }
__host__ __device__ inline int pair_lo_off(const GemmParams& p) { return p.pair_group ? p.pair_group : p.n_total; }
__host__ __device__ inline bool epi_pair16(const GemmParams& p) { return act_is_split(p.act) && (p.out_f32 != nullptr || (p.res != nullptr && p.res_f32)); }

// A-tensor column of K' position k (elements): the split-precision wrap, per K group when kgroup is set
__host__ __device__ inline int gemm_a_col(const GemmParams& p, int k) {
  if (!p.a_cols) return k;
  if (!p.kgroup) return k % p.a_cols;
  const int g = k / (3 * p.kgroup);
  return g * 2 * p.kgroup + (k - g * 3 * p.kgroup) % (2 * p.kgroup);
}

}  // namespace q3
