// ConvNeXt depthwise conv + LayerNorm, layout probes.
#include "kernels.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace q3 {

#define Q3_DISPATCH(bf16, ...)                                   \
  do {                                                            \
    if (bf16) { using T = __nv_bfloat16; __VA_ARGS__; }           \
    else { using T = __half; __VA_ARGS__; }                       \
  } while (0)

__device__ __forceinline__ float warp_sum2(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ============================================================================================ dwconv + LN
// One warp per output row; lane owns channels {4*lane + 128*i .. +3}.  Depthwise causal conv k=7
// (left zero padding, reference modeling :213-219, :189-192) then LayerNorm over C (modeling :220, :231).
template <typename T, int NI, typename XT>
__global__ void __launch_bounds__(256)
k_dwconv_ln(const XT* __restrict__ x, const float* __restrict__ dww, const float* __restrict__ dwb, const float* __restrict__ lnw,
            const float* __restrict__ lnb, T* __restrict__ out, int B, int L, int C, float eps, int in_bs_rows, int hist) {
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= (long long)B * L) return;
  const int l = (int)(row % L);
  const long long in_row = (row / L) * in_bs_rows + hist + l;  // row of x that holds sample l of this item (history rows in front)
  float acc[NI][4];
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const float4 bb = __ldg(reinterpret_cast<const float4*>(dwb + 4 * lane + 128 * i));
    acc[i][0] = bb.x; acc[i][1] = bb.y; acc[i][2] = bb.z; acc[i][3] = bb.w;
  }
#pragma unroll
  for (int j = 0; j < 7; ++j) {
    const int lt = hist + l - 6 + j;
    if (lt < 0) continue;
    const XT* xr = x + (in_row - 6 + j) * C;
#pragma unroll
    for (int i = 0; i < NI; ++i) {
      const int c = 4 * lane + 128 * i;
      float2 a, bq;
      if (sizeof(XT) == 4) {  // fp32 main-path tensor (split-precision plans keep the ConvNeXt input in fp32)
        const float4 f = *reinterpret_cast<const float4*>(xr + c);
        a = make_float2(f.x, f.y); bq = make_float2(f.z, f.w);
      } else {
        const uint2 u = *reinterpret_cast<const uint2*>(xr + c);
        a = Elem<T>::unpack2(u.x); bq = Elem<T>::unpack2(u.y);
      }
      const float4 w = __ldg(reinterpret_cast<const float4*>(dww + (long long)j * C + c));
      acc[i][0] = fmaf(w.x, a.x, acc[i][0]);
      acc[i][1] = fmaf(w.y, a.y, acc[i][1]);
      acc[i][2] = fmaf(w.z, bq.x, acc[i][2]);
      acc[i][3] = fmaf(w.w, bq.y, acc[i][3]);
    }
  }
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NI; ++i) s += acc[i][0] + acc[i][1] + acc[i][2] + acc[i][3];
  const float mean = warp_sum2(s) / (float)C;
  float v = 0.f;
#pragma unroll
  for (int i = 0; i < NI; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) { const float d = acc[i][e] - mean; v = fmaf(d, d, v); }
  const float rstd = rsqrtf(warp_sum2(v) / (float)C + eps);
#pragma unroll
  for (int i = 0; i < NI; ++i) {
    const int c = 4 * lane + 128 * i;
    const float4 g = __ldg(reinterpret_cast<const float4*>(lnw + c));
    const float4 bt = __ldg(reinterpret_cast<const float4*>(lnb + c));
    uint2 o;
    o.x = Elem<T>::pack2((acc[i][0] - mean) * rstd * g.x + bt.x, (acc[i][1] - mean) * rstd * g.y + bt.y);
    o.y = Elem<T>::pack2((acc[i][2] - mean) * rstd * g.z + bt.z, (acc[i][3] - mean) * rstd * g.w + bt.w);
    *reinterpret_cast<uint2*>(out + row * C + c) = o;
  }
}
cudaError_t dwconv_layernorm(const void* x, const float* dww, const float* dwb, const float* lnw, const float* lnb, void* out, int B, int L, int C,
                             float eps, bool bf16, cudaStream_t s, int in_bs_rows, int hist, bool x_f32) {
  if (in_bs_rows <= 0) in_bs_rows = L;
  if (C != 1024) return cudaErrorInvalidValue;  // NI = C / 128 is a compile-time unroll
  const long long rows = (long long)B * L;
  if (rows <= 0) return cudaSuccess;
  const int blocks = (int)((rows + 7) / 8);
  cudaError_t err;
  if (x_f32) Q3_DISPATCH(bf16, (err = launch_k(k_dwconv_ln<T, 8, float>, dim3(blocks), dim3(256), 0, s, 1, (const float*)x, dww, dwb, lnw, lnb, (T*)out, B, L, C, eps, in_bs_rows, hist)));
  else Q3_DISPATCH(bf16, (err = launch_k(k_dwconv_ln<T, 8, T>, dim3(blocks), dim3(256), 0, s, 1, (const T*)x, dww, dwb, lnw, lnb, (T*)out, B, L, C, eps, in_bs_rows, hist)));
  return err;
}

// ============================================================================================ fp32 -> [hi | lo]
template <typename T>
__global__ void __launch_bounds__(256)
k_split_f32(const float* __restrict__ in, long long in_bs, T* __restrict__ out, long long out_bs, int rows, int C, int with_lo) {
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int b = blockIdx.y;
  const int c4 = C >> 2;
  const long long n = (long long)rows * c4;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / c4;
    const int c = (int)(i - r * c4) * 4;
    const float4 v = *reinterpret_cast<const float4*>(in + b * in_bs + r * C + c);
    uint2 h, l;
    h.x = Elem<T>::pack2(v.x, v.y); h.y = Elem<T>::pack2(v.z, v.w);
    const float2 h0 = Elem<T>::unpack2(h.x), h1 = Elem<T>::unpack2(h.y);
    l.x = Elem<T>::pack2(v.x - h0.x, v.y - h0.y); l.y = Elem<T>::pack2(v.z - h1.x, v.w - h1.y);
    T* o = out + b * out_bs + r * (with_lo ? 2ll * C : (long long)C);
    *reinterpret_cast<uint2*>(o + c) = h;
    if (with_lo) *reinterpret_cast<uint2*>(o + C + c) = l;
  }
}
cudaError_t split_f32(const float* in, long long in_bs, void* out, long long out_bs, int B, int rows, int C, bool bf16, cudaStream_t s, int with_lo) {
  if (B <= 0 || rows <= 0) return cudaSuccess;
  if (C % 4) return cudaErrorInvalidValue;
  long long bx = ((long long)rows * (C >> 2) + 255) / 256;
  if (bx > 148 * 8) bx = 148 * 8;
  dim3 grid((unsigned)bx, (unsigned)B);
  cudaError_t err;
  Q3_DISPATCH(bf16, (err = launch_k(k_split_f32<T>, grid, dim3(256), 0, s, 1, in, in_bs, (T*)out, out_bs, rows, C, with_lo)));
  return err;
}

// ============================================================================================ probes
template <typename T>
__global__ void k_to_cf(const void* __restrict__ x, bool is_f32, float* __restrict__ out, int L, int C, int ld) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z;
  const int l0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int l = l0 + r, c = c0 + threadIdx.x;
    float v = 0.f;
    if (l < L && c < C) {
      const long long idx = ((long long)b * L + l) * ld + c;
      v = is_f32 ? reinterpret_cast<const float*>(x)[idx] : Elem<T>::to_f(reinterpret_cast<const T*>(x)[idx]);
    }
    tile[r][threadIdx.x] = v;
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int c = c0 + r, l = l0 + threadIdx.x;
    if (l < L && c < C) out[((long long)b * C + c) * L + l] = tile[threadIdx.x][r];
  }
}
cudaError_t to_channels_first_f32(const void* x, bool is_f32, float* out, int B, int L, int C, bool bf16, cudaStream_t s, int ld) {
  if (ld <= 0) ld = C;
  if (B <= 0 || L <= 0) return cudaSuccess;
  dim3 grid((L + 31) / 32, (C + 31) / 32, B), block(32, 8);
  Q3_DISPATCH(bf16, (k_to_cf<T><<<grid, block, 0, s>>>(x, is_f32, out, L, C, ld)));
  return cudaGetLastError();
}

}  // namespace q3
