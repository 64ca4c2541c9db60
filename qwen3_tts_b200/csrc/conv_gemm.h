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
  GemmParams p;
};

int gemm_setup(int device);
int gemm_num_sms();
int tmap_init();
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box0, uint32_t box1, bool bf16);
int gemm_stage_count(const GemmParams& p);
cudaError_t launch_conv_gemm(const GemmOp& op, int engine, cudaStream_t stream);

}  // namespace q3
