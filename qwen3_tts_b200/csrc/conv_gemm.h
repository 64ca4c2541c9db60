#pragma once
#include <cuda.h>
#include <cuda_runtime.h>

#include "gemm_params.h"

namespace q3 {

enum { ENGINE_TC = 0, ENGINE_SIMT = 1 };

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

struct GemmOp {
  CUtensorMap tmA;
  CUtensorMap tmB;
  CUtensorMap tmRes;   // residual tile loads   (box 32 rows x 64 bytes, SWIZZLE_64B)
  CUtensorMap tmRaw;   // 16-bit raw output stores
  CUtensorMap tmSnake; // 16-bit Snake'd output stores
  CUtensorMap tmF32;   // fp32 output stores
  GemmParams p;
  int ws = 0;       // 1: weight-stationary slab kernel
  int variant = 0;  // 0 general (one TMA box pair per (K block, tap)), 1 weight-stationary, 2 slab-stream
  int nslabs = 0;   // slab ring depth
  int nstages = 0;  // slab-stream: weight tile ring depth
  int srp = 0;      // slab rows (128 + halo, rounded up to 16)
  int slab_lo = 0;  // slab rows before the tile's first row (-min tap_off)
  int pair = 0;     // slab-stream variant on CTA pairs (2-CTA clusters, 256-row tiles, tcgen05 cta_group::2)
  int retile = 0;   // 1: gemm_make_epilogue_maps may shrink the N tile until the grid covers the SMs (small-M streaming steps)
};


// Fused vocoder residual unit (reference modeling :618-634):  y = x + W1 . snake2(W7 (*) a1 + b7) + b1,
// a1 = snake1(x) written by the producing kernel's epilogue.  One launch replaces the conv7 and conv1 launches:
// the conv7 accumulator is Snake'd and packed back into tensor memory, where the 1x1 conv reads it as its A operand
// (tcgen05.mma with A in TMEM), so the C x L intermediate never touches shared or global memory.
struct UnitOp {
  CUtensorMap tmA;     // a1 slab loads (box kb x slab_box_rows)
  CUtensorMap tmW7;    // [7][C][C] K-major, box kb x C
  CUtensorMap tmW1;    // [1][C][C]
  CUtensorMap tmRes, tmRaw, tmSnake;
  GemmParams p;        // describes the 1x1 conv epilogue (bias = b1, res = x, out_raw = y, out_snake = next a1)
  const float* bias7 = nullptr;
  const float* snake2_a = nullptr;
  const float* snake2_ib = nullptr;
  int C = 0, dil = 1, ms = 1;         // ms: 128-row sub-tiles per CTA tile (2 for C <= 96)
  int srows = 0, nbox = 1;            // slab rows per K block (128*ms + 6*dil rounded up), TMA boxes per slab
  int nslabs = 0, nstages = 0;        // ring depths
  int ks = 1, nbuf = 1;               // K blocks per slab / weight stage; TMEM stages
  // streaming (history-prefixed a1 buffer): rows of a1 per batch item, row of output row 0, strides (0 = dense)
  int a_rows = 0, a_row0 = 0;
  long long a_batch_stride = 0, snake_batch_stride = 0;
  int variant = 0;                    // compiled instance (unit_kernel.cu: 0 = C 96, 1 / 2 = C 192 with 4 / 3 weight stages, 3..5 = CTA-pair forms)
  int pair = 0;                       // 1: CTA-pair form (2-CTA clusters, 256-row tiles, tcgen05 cta_group::2)
  // xin = 1: the activation operand is the raw residual stream x (= p.res, read from a DIFFERENT buffer than p.out_raw: tiles read each
  // other's halo rows); the kernel applies the input Snake (snake1_*) to each slab in shared memory and stores one tensor per launch:
  // y, or snake_next(y) when p.skip_raw is set.  Two HBM passes per unit instead of four.
  int xin = 0;
  const float* snake1_a = nullptr;
  const float* snake1_ib = nullptr;
  int enabled = 0;
};
// Shape test only (no device work): true when unit_make can fuse C channels.
bool unit_fusable(int C);
// true when the fused unit for C channels has the raw-input (xin) form
bool unit_xin_available(int C);
// Fills the maps and ring sizes; returns 0 when the shape can be fused (enabled = 1), 1 when it cannot, < 0 on error.
int unit_make(UnitOp* op, const void* a1, const void* w7, const void* w1, int batches, int rows);
cudaError_t launch_unit(const UnitOp& op, cudaStream_t stream);
int unit_setup();  // per-device shared-memory opt-in of every unit_kernel instance (engine_finalize)
int unit_trace_read(long long* out256);

// Fill the epilogue tensor maps from p (res / out_raw / out_snake / out_f32, strides, widths).
// `allow` = bit mask of kernel variants that may be chosen (bit 0 general, bit 1 weight-stationary, bit 2 slab-stream); also (re)builds tmA.
int gemm_make_epilogue_maps(GemmOp* op, int allow = 7);

int gemm_setup(int device);
int gemm_num_sms();
int gemm_smem_optin();
int tmap_init();
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box0, uint32_t box1, bool bf16);
// general 3-D map: element type 0 = fp16, 1 = bf16, 2 = fp32; strides in elements
int make_tmap_3d_strided(CUtensorMap* out, const void* base, int etype, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1, uint64_t stride2,
                         uint32_t box0, uint32_t box1, uint32_t box2 = 1);
// Output conv C -> 1 (k = 7) + clamp with the taps as the N dimension (tail_kernel.cu)
struct TailParams {
  int batches, rows_valid, a_row0, C, tail_skip, bf16;
  float bias;
  float* out;                  // [batches][out_batch_stride] fp32, row t - tail_skip
  short* out_i16;              // when set: PCM16 instead, lrintf(clamp(v) * 32767) (what libsndfile writes for a float waveform)
  long long out_batch_stride;
  // Voice-clone prefix cut folded into the decode (model wrapper :612-631): sample g = out_pos0 + (t - tail_skip) of item b is stored at
  // out[b][g - item_skip[b]] and only when g >= item_skip[b], so the prefix audio is never written anywhere.  item_skip = null: no cut.
  long long out_pos0;
  const long long* item_skip;
};
struct TailOp {
  CUtensorMap tmA, tmW;
  TailParams p;
  int nstages = 0;
  int enabled = 0;
};
bool tail_supported(int C);
int tail_make(TailOp* op, const void* a, const void* w16, int batches, int a_rows, long long a_batch_stride, int C, bool bf16);
cudaError_t launch_tail(const TailOp& op, cudaStream_t stream);
int tail_setup();  // per-device shared-memory opt-in
int gemm_stage_count(const GemmParams& p, bool pair = false);
cudaError_t launch_conv_gemm(const GemmOp& op, int engine, cudaStream_t stream);

}  // namespace q3
