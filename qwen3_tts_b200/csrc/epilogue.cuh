// Fused epilogue shared by the tcgen05 implicit-GEMM kernel and the SIMT cross-check kernel.
// One call handles 16 consecutive accumulator columns of one output row.
#pragma once
#include "gemm_params.h"
#include "ptx.cuh"

namespace q3 {

__device__ __forceinline__ float gelu_erf(float v) { return 0.5f * v * (1.0f + erff(v * 0.70710678118654752440f)); }
__device__ __forceinline__ float silu(float v) { return v / (1.0f + __expf(-v)); }
// ELU (alpha 1) of the encode path; exp(v) - 1 with the hardware exponential (absolute error ~2e-7, the level of the fp16 pair it is stored as)
__device__ __forceinline__ float elu_fast(float v) { return v > 0.f ? v : __expf(v) - 1.0f; }
// SnakeBeta (reference modeling :601-615): x + 1/(exp(beta)+1e-9) * sin(x*exp(alpha))^2, evaluated as
//   (x + h) - h * cos(a2 * x),  a2 = 2 exp(alpha), h = (1/2) / (exp(beta)+1e-9)   (tables from pack_snake)
// which is one multiply less per element and has a packed two-lane form (snake2) of FMUL2 / FADD2 / FFMA2.
// Every kernel goes through these two functions, lane for lane the same operations, so bulk, streaming and the
// SIMT cross-check stay bit-identical.
__device__ __forceinline__ float snake(float v, float a2, float h) {
  return fmaf(-h, __cosf(v * a2), v + h);
}
__device__ __forceinline__ float2 snake2(float2 v, float2 a2, float2 h) {
  const float2 t = mul2(v, a2);
  const float2 c = make_float2(__cosf(t.x), __cosf(t.y));
  return fma2(make_float2(-h.x, -h.y), c, add2(v, h));
}

__device__ __forceinline__ void load16f(const float* __restrict__ p, float (&o)[16]) {
  const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4 t = __ldg(q + i);
    o[4 * i + 0] = t.x; o[4 * i + 1] = t.y; o[4 * i + 2] = t.z; o[4 * i + 3] = t.w;
  }
}

template <typename T>
__device__ __forceinline__ void store16(T* __restrict__ dst, const float (&v)[16]) {
  uint4 a, b;
  a.x = Elem<T>::pack2(v[0], v[1]);   a.y = Elem<T>::pack2(v[2], v[3]);
  a.z = Elem<T>::pack2(v[4], v[5]);   a.w = Elem<T>::pack2(v[6], v[7]);
  b.x = Elem<T>::pack2(v[8], v[9]);   b.y = Elem<T>::pack2(v[10], v[11]);
  b.z = Elem<T>::pack2(v[12], v[13]); b.w = Elem<T>::pack2(v[14], v[15]);
  uint4* d = reinterpret_cast<uint4*>(dst);
  d[0] = a; d[1] = b;
}

// v: 16 fp32 accumulators for columns [n, n+16) of output row (b, row).  n % 16 == 0.
template <typename T>
__device__ __forceinline__ void epilogue16(const GemmParams& p, float (&v)[16], int b, int row, int n) {
  const int c0 = n % p.ncol_mod;  // ncol_mod % 16 == 0, so the 16 columns never wrap
  if (p.bias) {
    float t[16]; load16f(p.bias + c0, t);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] += t[i];
  }
  if (p.act == ACT_GLU) {
    // columns are interleaved (gate, up) pairs -> 8 outputs at column n/2 (reference modeling :367-369)
    uint4 o;
    o.x = Elem<T>::pack2(silu(v[0]) * v[1], silu(v[2]) * v[3]);
    o.y = Elem<T>::pack2(silu(v[4]) * v[5], silu(v[6]) * v[7]);
    o.z = Elem<T>::pack2(silu(v[8]) * v[9], silu(v[10]) * v[11]);
    o.w = Elem<T>::pack2(silu(v[12]) * v[13], silu(v[14]) * v[15]);
    T* dst = reinterpret_cast<T*>(p.out_raw) + (long long)b * p.out_batch_stride + (long long)row * p.out_row_stride + (n >> 1);
    *reinterpret_cast<uint4*>(dst) = o;
    return;
  }
  if (p.act == ACT_GELU) {
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);
  }
  if (p.scale) {
    float t[16]; load16f(p.scale + c0, t);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] *= t[i];
  }
  const long long off = (long long)b * p.out_batch_stride + (long long)row * p.out_row_stride + n;
  if (p.res) {
    if (p.res_f32) {
      const float4* r = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.res) + off);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float4 t = r[i];
        v[4 * i + 0] += t.x; v[4 * i + 1] += t.y; v[4 * i + 2] += t.z; v[4 * i + 3] += t.w;
      }
    } else {
      const uint4* r = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(p.res) + off);
      uint4 ra = r[0], rb = r[1];
      const uint32_t u[8] = {ra.x, ra.y, ra.z, ra.w, rb.x, rb.y, rb.z, rb.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float2 f = Elem<T>::unpack2(u[i]);
        v[2 * i] += f.x; v[2 * i + 1] += f.y;
      }
    }
  }
  if (p.out_f32) {
    float4* d = reinterpret_cast<float4*>(p.out_f32 + off);
#pragma unroll
    for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
  }
  if (act_is_split(p.act)) {  // f(x) as an fp16 pair (gemm_params.h); the pair tensor has its own pitches and column map
    const long long o = (long long)b * (p.pair_batch_stride ? p.pair_batch_stride : p.out_batch_stride) +
                        (long long)row * (p.pair_row_stride ? p.pair_row_stride : p.out_row_stride) + pair_hi_col(p, n);
    float lo[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (p.act == ACT_ELU_SPLIT) v[i] = elu_fast(v[i]); else v[i] = gelu_erf(v[i]);
      lo[i] = v[i] - Elem<T>::to_f(Elem<T>::from_f(v[i]));
    }
    store16<T>(reinterpret_cast<T*>(p.out_raw) + o, v);
    store16<T>(reinterpret_cast<T*>(p.out_raw) + o + pair_lo_off(p), lo);
    return;
  }
  if (p.out_raw) store16<T>(reinterpret_cast<T*>(p.out_raw) + off, v);
  if (p.out_snake) {
    float a[16], ib[16];
    load16f(p.snake_a + c0, a);
    load16f(p.snake_ib + c0, ib);
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = snake(v[i], a[i], ib[i]);
    store16<T>(reinterpret_cast<T*>(p.out_snake) + off, v);
  }
}

}  // namespace q3
