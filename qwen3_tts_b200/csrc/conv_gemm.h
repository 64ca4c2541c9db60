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
  int ws = 0;      // 1: weight-stationary slab kernel
  int nslabs = 0;  // slab ring depth
  int srp = 0;     // slab rows (128 + halo, rounded up to 16)
};

// Fill the epilogue tensor maps from p (res / out_raw / out_snake / out_f32, strides, widths).
int gemm_make_epilogue_maps(GemmOp* op);

int gemm_setup(int device);
int gemm_num_sms();
int tmap_init();
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box0, uint32_t box1, bool bf16);
// general 3-D map: element type 0 = fp16, 1 = bf16, 2 = fp32; strides in elements
int make_tmap_3d_strided(CUtensorMap* out, const void* base, int etype, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1, uint64_t stride2,
                         uint32_t box0, uint32_t box1);
int gemm_stage_count(const GemmParams& p);
cudaError_t launch_conv_gemm(const GemmOp& op, int engine, cudaStream_t stream);

}  // namespace q3
