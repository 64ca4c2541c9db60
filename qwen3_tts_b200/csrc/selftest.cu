// Device-side self-test of the tcgen05 implicit-GEMM kernel against the SIMT cross-check kernel.
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/q3tts_decode.h"
#include "../../include/q3tts_debug.h"
#include "conv_gemm.h"
#include "epilogue.cuh"
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
      gemm_make_epilogue_maps(&op);  // picks weight-stationary / slab-stream / general as the shape allows
      cudaError_t e1 = launch_conv_gemm(op, ENGINE_TC, 0);
      k_maxdiff<<<296, 256>>>(o1, o2, (long long)no, res);
      // also exercise every lower-priority variant on the same problem: ws -> slab-stream -> general
      int allow = op.variant == 1 ? 5 : (op.variant == 2 ? 1 : 0);
      while (allow && e1 == cudaSuccess) {
        gemm_make_epilogue_maps(&op, allow);
        cudaMemsetAsync(o1, 0xff, no * 4, 0);
        e1 = launch_conv_gemm(op, ENGINE_TC, 0);
        k_maxdiff<<<296, 256>>>(o1, o2, (long long)no, res);
        allow = op.variant == 2 ? 1 : 0;
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

namespace {

template <typename T>
__global__ void k_maxdiff16(const T* a, const T* b, long long n, float* out) {
  float md = 0.f, mr = 0.f;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float x = Elem<T>::to_f(a[i]), y = Elem<T>::to_f(b[i]);
    float d = fabsf(x - y);
    if (!(d == d)) d = INFINITY;
    md = fmaxf(md, d);
    mr = fmaxf(mr, fabsf(y));
  }
  for (int o = 16; o > 0; o >>= 1) {
    md = fmaxf(md, __shfl_xor_sync(0xffffffffu, md, o));
    mr = fmaxf(mr, __shfl_xor_sync(0xffffffffu, mr, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(reinterpret_cast<int*>(out), __float_as_int(md));
    atomicMax(reinterpret_cast<int*>(out) + 1, __float_as_int(mr));
  }
}

__global__ void k_fillf(float* p, int n, uint32_t seed, float lo, float hi) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t x = (uint32_t)i * 2654435761u + seed;
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
  p[i] = lo + (hi - lo) * ((float)(x & 0xFFFFFF) / 16777216.0f);
}

template <typename T>
int run_unit_selftest(int rows, int C, int dil, float* out4) {
  const bool bf16 = Elem<T>::kBf16;
  const int batches = 3;
  const size_t nact = (size_t)batches * rows * C;
  T *a1 = nullptr, *x = nullptr, *w7 = nullptr, *w1 = nullptr, *a2 = nullptr, *y0 = nullptr, *s0 = nullptr, *y1 = nullptr, *s1 = nullptr;
  float *fv = nullptr, *res = nullptr;
  int rc = Q3TTS_OK;
  if (cudaMalloc((void**)&a1, nact * 2) || cudaMalloc((void**)&x, nact * 2) || cudaMalloc((void**)&a2, nact * 2) || cudaMalloc((void**)&y0, nact * 2) ||
      cudaMalloc((void**)&s0, nact * 2) || cudaMalloc((void**)&y1, nact * 2) || cudaMalloc((void**)&s1, nact * 2) ||
      cudaMalloc((void**)&w7, (size_t)7 * C * C * 2) || cudaMalloc((void**)&w1, (size_t)C * C * 2) || cudaMalloc((void**)&fv, (size_t)6 * C * 4) ||
      cudaMalloc((void**)&res, 16)) {
    rc = Q3TTS_OOM;
  } else {
    k_fill<T><<<592, 256>>>(a1, (long long)nact, 1u, 2.0f);
    k_fill<T><<<592, 256>>>(x, (long long)nact, 3u, 2.0f);
    k_fill<T><<<592, 256>>>(w7, (long long)7 * C * C, 7u, 1.5f / sqrtf(7.0f * C));
    k_fill<T><<<592, 256>>>(w1, (long long)C * C, 9u, 3.0f / sqrtf((float)C));
    float *b7 = fv, *b1 = fv + C, *s2a = fv + 2 * C, *s2ib = fv + 3 * C, *sna = fv + 4 * C, *snib = fv + 5 * C;
    k_fillf<<<(C + 127) / 128, 128>>>(b7, C, 11u, -0.1f, 0.1f);
    k_fillf<<<(C + 127) / 128, 128>>>(b1, C, 12u, -0.1f, 0.1f);
    k_fillf<<<(C + 127) / 128, 128>>>(s2a, C, 13u, 0.6f, 1.6f);
    k_fillf<<<(C + 127) / 128, 128>>>(s2ib, C, 14u, 0.5f, 1.5f);
    k_fillf<<<(C + 127) / 128, 128>>>(sna, C, 15u, 0.6f, 1.6f);
    k_fillf<<<(C + 127) / 128, 128>>>(snib, C, 16u, 0.5f, 1.5f);
    cudaMemset(res, 0, 16);
    cudaMemset(y1, 0xff, nact * 2);
    cudaMemset(s1, 0xff, nact * 2);
    // reference: the two SIMT cross-check launches
    GemmOp g;
    GemmParams& p = g.p;
    memset(&p, 0, sizeof(p));
    p.batches = batches; p.rows = rows; p.rows_valid = rows; p.cin = C; p.ntaps = 7;
    for (int j = 0; j < 7; ++j) p.tap_off[j] = -(6 - j) * dil;
    p.n_total = C; p.bn = C; p.kb = C % 64 == 0 ? 64 : 32; p.ncol_mod = C; p.bf16 = bf16;
    p.out_row_stride = C; p.out_batch_stride = (long long)rows * C;
    p.a_ptr = a1; p.w_ptr = w7; p.bias = b7; p.out_snake = a2; p.snake_a = s2a; p.snake_ib = s2ib;
    cudaError_t e1 = launch_conv_gemm(g, ENGINE_SIMT, 0);
    p.ntaps = 1; p.tap_off[0] = 0; p.a_ptr = a2; p.w_ptr = w1; p.bias = b1; p.res = x; p.out_raw = y0; p.out_snake = s0; p.snake_a = sna; p.snake_ib = snib;
    cudaError_t e2 = launch_conv_gemm(g, ENGINE_SIMT, 0);
    // fused unit kernel
    UnitOp op;
    memset(&op.p, 0, sizeof(op.p));
    op.C = C; op.dil = dil; op.bias7 = b7; op.snake2_a = s2a; op.snake2_ib = s2ib;
    { const char* d = Q3_ENV("Q3TTS_DBG"); op.p.dbg = d ? atoi(d) : 0; }
    op.p.bf16 = bf16; op.p.bias = b1; op.p.res = x; op.p.out_raw = y1; op.p.out_snake = s1; op.p.snake_a = sna; op.p.snake_ib = snib;
    const int mk = unit_make(&op, a1, w7, w1, batches, rows);
    cudaError_t e3 = mk == 0 ? launch_unit(op, 0) : cudaErrorInvalidConfiguration;
    k_maxdiff16<T><<<296, 256>>>(y1, y0, (long long)nact, res);
    k_maxdiff16<T><<<296, 256>>>(s1, s0, (long long)nact, res + 2);
    cudaError_t e4 = cudaMemcpy(out4, res, 16, cudaMemcpyDeviceToHost);
    if (mk != 0) rc = Q3TTS_BAD_CONFIG;
    else if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess || e4 != cudaSuccess) rc = Q3TTS_CUDA_ERROR;
  }
  cudaFree(a1); cudaFree(x); cudaFree(a2); cudaFree(y0); cudaFree(s0); cudaFree(y1); cudaFree(s1); cudaFree(w7); cudaFree(w1); cudaFree(fv); cudaFree(res);
  return rc;
}

}  // namespace

// Fused residual-unit kernel vs the two SIMT cross-check launches (conv7 + Snake, conv1 + residual + Snake).
// out4 = {max|y - y_ref|, max|y_ref|, max|s - s_ref|, max|s_ref|} over 16-bit outputs.
extern "C" int q3tts_selftest_unit(int device, int rows, int channels, int dilation, int bf16, float* out4) {
  if (rows < 1 || channels < 32 || dilation < 1 || !out4) return Q3TTS_BAD_SHAPE;
  if (cudaSetDevice(device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  if (gemm_setup(device) || unit_setup()) return Q3TTS_CUDA_ERROR;
  return bf16 ? run_unit_selftest<__nv_bfloat16>(rows, channels, dilation, out4) : run_unit_selftest<__half>(rows, channels, dilation, out4);
}

// ---------------------------------------------------------------------------------------------------------------------
// TMA box-shape microbenchmark: every warp of every CTA moves [box_rows x box_cols] 16-bit boxes between its shared-memory
// slot and a channels-last tensor [rows][C] (stores, or loads with an mbarrier), 2 slots in flight per warp.
// Reports aggregate GB/s.  Used to pick the epilogue box shape (DESIGN.md "Epilogue boxes").
namespace {

__global__ void __launch_bounds__(512, 1) k_tma_bench(const __grid_constant__ CUtensorMap tm, int C, int box_cols, int box_rows, int rows_total,
                                                       int is_load, int iters) {
  extern __shared__ __align__(1024) uint8_t sm_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(sm_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bars[16][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const uint32_t box_bytes = (uint32_t)box_cols * box_rows * 2;
  const uint32_t slot_bytes = (box_bytes + 1023) & ~1023u;
  uint8_t* my = sm + (size_t)warp * 2 * slot_bytes;
  if (lane == 0) { mbar_init(&bars[warp][0], 1); mbar_init(&bars[warp][1], 1); fence_barrier_init(); }
  for (uint32_t i = lane; i < 2 * slot_bytes / 4; i += 32) reinterpret_cast<uint32_t*>(my)[i] = i;
  fence_proxy_async_smem();
  __syncthreads();
  const int col_boxes = C / box_cols;
  const int row_boxes = rows_total / box_rows;
  long long g = ((long long)blockIdx.x * nwarps + warp);
  const long long stride = (long long)gridDim.x * nwarps;
  if (lane == 0) {
    for (int i = 0; i < iters; ++i, g += stride) {
      const int cb = (int)(g % col_boxes), rb = (int)((g / col_boxes) % row_boxes);
      const int slot = i & 1;
      if (is_load) {
        if (i >= 2) mbar_wait(&bars[warp][slot], ((i >> 1) - 1) & 1);
        mbar_expect_tx(&bars[warp][slot], box_bytes);
        tma_load_3d(my + slot * slot_bytes, &tm, &bars[warp][slot], cb * box_cols, rb * box_rows, 0);
      } else {
        tma_store_wait_read<1>();
        tma_store_3d(&tm, my + slot * slot_bytes, cb * box_cols, rb * box_rows, 0);
        tma_store_commit();
      }
    }
    if (is_load) {
      for (int i = iters > 2 ? iters - 2 : 0; i < iters; ++i) mbar_wait(&bars[warp][i & 1], (i >> 1) & 1);
    } else {
      tma_store_wait<0>();
    }
  }
}

}  // namespace

extern "C" int q3tts_bench_tma(int device, int C, int box_cols, int box_rows, int is_load, int warps, int iters, float* gbps) {
  if (!gbps || C < 16 || box_cols < 8 || C % box_cols || box_rows < 1 || box_rows > 256 || warps < 1 || warps > 16) return Q3TTS_BAD_SHAPE;
  if (cudaSetDevice(device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  if (gemm_setup(device)) return Q3TTS_CUDA_ERROR;
  const int rows_total = 1 << 22;
  void* buf = nullptr;
  if (cudaMalloc(&buf, (size_t)rows_total * C * 2) != cudaSuccess) return Q3TTS_OOM;
  cudaMemset(buf, 0, (size_t)rows_total * C * 2);
  CUtensorMap tm;
  int rc = Q3TTS_OK;
  if (make_tmap_3d_strided(&tm, buf, 0, (uint64_t)C, (uint64_t)rows_total, 1, (uint64_t)C, (uint64_t)rows_total * C, (uint32_t)box_cols, (uint32_t)box_rows)) {
    rc = Q3TTS_CUDA_ERROR;
  } else {
    const uint32_t slot = ((uint32_t)box_cols * box_rows * 2 + 1023) & ~1023u;
    const size_t smem = (size_t)warps * 2 * slot + 1024;
    if (smem + 1024 > (size_t)gemm_smem_optin() ||
        cudaFuncSetAttribute(k_tma_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) rc = Q3TTS_BAD_SHAPE;
    else {
      cudaEvent_t e0, e1;
      cudaEventCreate(&e0); cudaEventCreate(&e1);
      k_tma_bench<<<gemm_num_sms(), warps * 32, smem>>>(tm, C, box_cols, box_rows, rows_total, is_load, 64);
      cudaEventRecord(e0);
      k_tma_bench<<<gemm_num_sms(), warps * 32, smem>>>(tm, C, box_cols, box_rows, rows_total, is_load, iters);
      cudaEventRecord(e1);
      if (cudaGetLastError() != cudaSuccess || cudaEventSynchronize(e1) != cudaSuccess) rc = Q3TTS_CUDA_ERROR;
      float ms = 0.f;
      cudaEventElapsedTime(&ms, e0, e1);
      *gbps = (float)((double)gemm_num_sms() * warps * iters * box_cols * box_rows * 2.0 / (ms * 1e6));
      cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
  }
  cudaFree(buf);
  return rc;
}

// SnakeBeta as every epilogue evaluates it (epilogue.cuh snake(): cosine form on MUFU.COS) on n caller-supplied points:
// out[i] = snake(x[i], a2, h).  The test compares with x + (1/(exp(beta)+1e-9)) sin^2(x exp(alpha)) in float64 for |x a2| up to 1e3.
namespace {
__global__ void k_snake_probe(const float* __restrict__ x, float a2, float h, float* __restrict__ out, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) out[i] = q3::snake(x[i], a2, h);
}
}  // namespace
extern "C" int q3tts_selftest_snake(int device, const float* x_host, int n, float a2, float h, float* out_host) {
  if (!x_host || !out_host || n < 1) return Q3TTS_BAD_SHAPE;
  if (cudaSetDevice(device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  float *dx = nullptr, *dy = nullptr;
  if (cudaMalloc((void**)&dx, (size_t)n * 4) != cudaSuccess || cudaMalloc((void**)&dy, (size_t)n * 4) != cudaSuccess) { cudaFree(dx); return Q3TTS_OOM; }
  cudaMemcpy(dx, x_host, (size_t)n * 4, cudaMemcpyHostToDevice);
  k_snake_probe<<<(n + 255) / 256, 256>>>(dx, a2, h, dy, n);
  const cudaError_t e = cudaMemcpy(out_host, dy, (size_t)n * 4, cudaMemcpyDeviceToHost);
  cudaFree(dx); cudaFree(dy);
  return e == cudaSuccess ? Q3TTS_OK : Q3TTS_CUDA_ERROR;
}

#include "launch.cuh"
// Same-box A/B of programmatic dependent launch: 0 = every kernel serialises on the previous one (plain stream order), 1 = default.
extern "C" int q3tts_debug_set_pdl(int on) { q3::pdl_set(on != 0); return Q3TTS_OK; }

extern "C" int q3tts_unit_trace(long long* out256) { return unit_trace_read(out256); }

// ---------------------------------------------------------------------------------------------------------------------
// tcgen05.mma issue-rate microbenchmark: one CTA per SM, one thread issues `iters` x 8 MMAs (M = 128, K = 16, fp16) on
// whatever bytes are in shared memory (results are never read) and reports cycles per MMA.
//   form 0: A and B from shared memory (SWIZZLE_128B or 64B K-major descriptors), a_row_shift = rows the A start is shifted
//   form 1: A from tensor memory, B from shared memory
namespace {

__global__ void __launch_bounds__(128, 1) k_mma_bench(int n, int span, int a_row_shift, int form, int iters, int nacc, int commit_every, int wait_every,
                                                       long long* out) {
  extern __shared__ __align__(1024) uint8_t sm_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(sm_raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar, bar2, bar3;
  __shared__ uint32_t tbase;
  for (uint32_t i = threadIdx.x; i < 96 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); mbar_init(&bar2, 1); mbar_init(&bar3, 1); fence_barrier_init(); }
  if (threadIdx.x < 32) { tmem_alloc(&tbase, 512); tmem_relinquish(); }
  fence_proxy_async_smem();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  // `wait_every` doubles as the number of issuing warps when commit_every < 0 (multi-issuer experiment)
  const int issuers = commit_every < 0 ? wait_every : 1;
  __shared__ __align__(8) uint64_t ibar[4];
  if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(&ibar[i], 1); fence_barrier_init(); }
  __syncthreads();
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0 && w < issuers) {
    const uint32_t idesc = umma_idesc(128, n, false);
    const uint32_t hi = umma_desc_hi(span);
    const uint32_t a0 = umma_desc_lo(smem_u32(sm) + (uint32_t)a_row_shift * span + (uint32_t)w * 8192);
    const uint32_t b0 = umma_desc_lo(smem_u32(sm) + 48 * 1024 + (uint32_t)w * 8192);
    const uint32_t d0 = tbase + (uint32_t)(w * n);
    const long long t0 = clock64();
    if (commit_every < 0) {
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int j = 0; j < 8; ++j) umma_f16_lohi(d0, a0 + 2 * (j & 3), b0 + 2 * (j & 3), hi, idesc, 1);
      }
    } else {
      // per group of 8 MMAs: bit 0 of commit_every = tcgen05.commit, bit 1 = mbarrier wait that passes at once, bit 2 = tcgen05.fence::after_thread_sync
      for (int i = 0; i < iters; ++i) {
        if (form == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) umma_f16_lohi(d0, a0 + 2 * (j & 3), b0 + 2 * (j & 3), hi, idesc, 1);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) umma_f16_ts(d0, tbase + 480 + 8 * (j & 3), b0 + 2 * (j & 3), hi, idesc, 1);
        }
        if (commit_every & 1) umma_commit(&bar2);
        if (commit_every & 2) mbar_wait(&bar3, 1);  // bar3 never completes a phase: parity 1 passes at once
        if (commit_every & 4) tc_fence_after();
      }
    }
    umma_commit(&ibar[w]);
    mbar_wait(&ibar[w], 0);
    const long long t1 = clock64();
    if (blockIdx.x == 0 && w == 0) out[0] = t1 - t0;
  }
  tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { tc_fence_after(); tmem_dealloc(tbase, 512); }
}

}  // namespace

extern "C" int q3tts_bench_mma(int device, int n, int span, int a_row_shift, int form, int nacc, int commit_every, int wait_every, float* clk_per_mma) {
  if (!clk_per_mma || n < 16 || n > 256 || n % 16 || (span != 64 && span != 128) || nacc < 1 || nacc * n > 480 || (commit_every < 0 && (wait_every < 1 || wait_every > 4 || wait_every * n > 480)))
    return Q3TTS_BAD_SHAPE;
  if (cudaSetDevice(device) != cudaSuccess) return Q3TTS_CUDA_ERROR;
  if (gemm_setup(device)) return Q3TTS_CUDA_ERROR;
  long long* out = nullptr;
  if (cudaMalloc((void**)&out, 8) != cudaSuccess) return Q3TTS_OOM;
  const int iters = 2000;
  const size_t smem = 97 * 1024 + 1024;
  cudaFuncSetAttribute(k_mma_bench, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_mma_bench<<<gemm_num_sms(), 128, smem>>>(n, span, a_row_shift, form, 50, nacc, commit_every, wait_every, out);
  k_mma_bench<<<gemm_num_sms(), 128, smem>>>(n, span, a_row_shift, form, iters, nacc, commit_every, wait_every, out);
  long long h = 0;
  int rc = Q3TTS_OK;
  if (cudaGetLastError() != cudaSuccess || cudaMemcpy(&h, out, 8, cudaMemcpyDeviceToHost) != cudaSuccess) rc = Q3TTS_CUDA_ERROR;
  *clk_per_mma = (float)((double)h / (8.0 * iters));
  cudaFree(out);
  return rc;
}
