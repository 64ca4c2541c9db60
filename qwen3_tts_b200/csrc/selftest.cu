// Device-side self-test of the tcgen05 implicit-GEMM kernel against the SIMT cross-check kernel.
#include <cmath>
#include <cstring>
#include <vector>

#include "../../include/q3tts_decode.h"
#include "conv_gemm.h"
#include "ptx.cuh"

using namespace q3;

namespace {

template <typename T>
__global__ void k_fill(T* p, long long n, uint32_t seed, float scale) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    uint32_t x = (uint32_t)i * 2654435761u + seed;
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    const float u = (float)(x & 0xFFFFFF) / 16777216.0f - 0.5f;
    p[i] = Elem<T>::from_f(u * scale);
  }
}

__global__ void k_maxdiff(const float* a, const float* b, long long n, float* out /*[2]: max|a-b|, max|b|*/) {
  float md = 0.f, mr = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float d = fabsf(a[i] - b[i]);
    if (!(d == d)) d = INFINITY;  // NaN must not hide behind fmaxf
    md = fmaxf(md, d);
    mr = fmaxf(mr, fabsf(b[i]));
  }
  for (int o = 16; o > 0; o >>= 1) {
    md = fmaxf(md, __shfl_xor_sync(0xffffffffu, md, o));
    mr = fmaxf(mr, __shfl_xor_sync(0xffffffffu, mr, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(reinterpret_cast<int*>(out), __float_as_int(md));  // non-negative floats order like ints
    atomicMax(reinterpret_cast<int*>(out) + 1, __float_as_int(mr));
  }
}

}  // namespace

extern "C" int q3tts_selftest_gemm(int device, int rows, int cin, int n, int ntaps, int dilation, int bf16, float* max_abs_diff,
                                   float* max_abs_ref) {
  if (rows < 1 || cin < 32 || n < 16 || n % 16 || ntaps < 1 || ntaps > 8 || !max_abs_diff) return Q3TTS_BAD_SHAPE;
  if (cudaSetDevice(device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  if (gemm_setup(device)) return Q3TTS_CUDA_ERROR;
  const int batches = 3;
  const int kb = cin % 64 == 0 ? 64 : 32;
  if (cin % kb) return Q3TTS_BAD_SHAPE;
  int bn = 0;
  for (int c = 256; c >= 16; c -= 16)
    if (n % c == 0) { bn = c; break; }
  const size_t na = (size_t)batches * rows * cin, nw = (size_t)ntaps * n * cin, no = (size_t)batches * rows * n;
  void *A = nullptr, *W = nullptr;
  float *o1 = nullptr, *o2 = nullptr, *res = nullptr;
  int rc = Q3TTS_OK;
  if (cudaMalloc(&A, na * 2) || cudaMalloc(&W, nw * 2) || cudaMalloc((void**)&o1, no * 4) || cudaMalloc((void**)&o2, no * 4) ||
      cudaMalloc((void**)&res, 8)) {
    rc = Q3TTS_OOM;
  } else {
    if (bf16) {
      k_fill<__nv_bfloat16><<<592, 256>>>((__nv_bfloat16*)A, (long long)na, 1u, 2.0f);
      k_fill<__nv_bfloat16><<<592, 256>>>((__nv_bfloat16*)W, (long long)nw, 7u, 0.25f);
    } else {
      k_fill<__half><<<592, 256>>>((__half*)A, (long long)na, 1u, 2.0f);
      k_fill<__half><<<592, 256>>>((__half*)W, (long long)nw, 7u, 0.25f);
    }
    cudaMemset(res, 0, 8);
    GemmOp op;
    GemmParams& p = op.p;
    memset(&p, 0, sizeof(p));
    p.batches = batches; p.rows = rows; p.rows_valid = rows; p.cin = cin; p.ntaps = ntaps;
    for (int j = 0; j < ntaps; ++j) p.tap_off[j] = -(ntaps - 1 - j) * dilation;
    p.n_total = n; p.bn = bn; p.kb = kb; p.ncol_mod = n; p.bf16 = bf16;
    p.out_row_stride = n; p.out_batch_stride = (long long)rows * n;
    p.a_ptr = A; p.w_ptr = W;
    if (make_tmap_3d(&op.tmA, A, cin, rows, batches, kb, 128, bf16 != 0) || make_tmap_3d(&op.tmB, W, cin, n, ntaps, kb, bn, bf16 != 0)) {
      rc = Q3TTS_CUDA_ERROR;
    } else {
      p.out_f32 = o2;
      cudaError_t e2 = launch_conv_gemm(op, ENGINE_SIMT, 0);
      p.out_f32 = o1;
      gemm_make_epilogue_maps(&op);  // picks the weight-stationary slab kernel when the shape allows it
      cudaError_t e1 = launch_conv_gemm(op, ENGINE_TC, 0);
      k_maxdiff<<<296, 256>>>(o1, o2, (long long)no, res);
      if (op.ws && e1 == cudaSuccess) {  // also exercise the general (per-tap TMA box) kernel on the same problem
        op.ws = 0;
        make_tmap_3d(&op.tmA, A, cin, rows, batches, kb, 128, bf16 != 0);
        cudaMemsetAsync(o1, 0xff, no * 4, 0);
        e1 = launch_conv_gemm(op, ENGINE_TC, 0);
        k_maxdiff<<<296, 256>>>(o1, o2, (long long)no, res);
      }
      float hres[2] = {0.f, 0.f};
      cudaError_t e3 = cudaMemcpy(hres, res, 8, cudaMemcpyDeviceToHost);
      if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) rc = Q3TTS_CUDA_ERROR;
      *max_abs_diff = hres[0];
      if (max_abs_ref) *max_abs_ref = hres[1];
    }
  }
  cudaFree(A); cudaFree(W); cudaFree(o1); cudaFree(o2); cudaFree(res);
  return rc;
}
