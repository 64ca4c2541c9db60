// Bandwidth-bound stages of the decode path: RVQ gather-sum, RMSNorm, RoPE + sliding-window
// attention, depthwise conv + LayerNorm, tail conv + clamp, layout probes.
#include "kernels.h"
#include "launch.cuh"
#include "ptx.cuh"
#include "attention_mma.cuh"

namespace q3 {

#define Q3_DISPATCH(bf16, ...)                                   \
  do {                                                            \
    if (bf16) { using T = __nv_bfloat16; __VA_ARGS__; }           \
    else { using T = __half; __VA_ARGS__; }                       \
  } while (0)

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ============================================================================================ RVQ
// One block per frame; thread i owns 4 consecutive channels (one float4 of each table row: a warp reads 512 contiguous bytes of a
// row).  The frame's Q code indices are staged in shared memory by the first Q threads (one coalesced read of the codes), then
// every thread issues its row loads in groups of 8 BEFORE summing any of them, so 8 independent 16-byte loads are in flight per
// thread instead of one load waiting behind each code fetch (ncu r2: the serial form was latency-bound at 2.2 TB/s).
// Sum order: book 0 (rvq_first) plus the left-to-right sum of books 1..Q-1 (rvq_rest), as the reference sums them
// (modeling :720-726, :817-819) - unchanged, so the result is bit-identical to the serial form.
template <typename T>
__global__ void __launch_bounds__(1024)
k_rvq(const long long* __restrict__ codes, long long sb, long long sk, long long st, const float* __restrict__ tables,
                      T* __restrict__ out_act, float* __restrict__ out_f32, int Q, int Tn, int ncodes, int D, int out_bs_rows,
                      const int* __restrict__ src_index, int* __restrict__ bad_code, int split) {
  __shared__ int s_code[64];  // Q <= 64 (q3tts_create)
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int t = blockIdx.x, b = blockIdx.y;
  const int bs = src_index ? src_index[b] : b;  // row of the caller's codes tensor this batch item reads
  if ((int)threadIdx.x < Q) {
    long long c = codes[bs * sb + threadIdx.x * sk + t * st];
    if (c < 0 || c >= ncodes) {
      // never fault on a bad code (the reference's F.embedding would assert, modeling :675-678): decode a clamped index and raise the
      // handle's device-side flag, which the checked entries (q3tts_set_check_codes, q3tts_decode_host) turn into Q3TTS_CODE_OUT_OF_RANGE
      if (bad_code) atomicMax(bad_code, 1);
      c = c < 0 ? 0 : ncodes - 1;
    }
    s_code[threadIdx.x] = (int)c;
  }
  __syncthreads();
  const int d = threadIdx.x * 4;
  if (d >= D) return;
  float4 first = make_float4(0.f, 0.f, 0.f, 0.f), rest = first;
  for (int k0 = 0; k0 < Q; k0 += 8) {
    float4 e[8];
#pragma unroll
    for (int j = 0; j < 8; ++j)
      if (k0 + j < Q) e[j] = __ldg(reinterpret_cast<const float4*>(tables + ((long long)(k0 + j) * ncodes + s_code[k0 + j]) * D + d));
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (k0 + j >= Q) break;
      if (k0 + j == 0) first = e[j];
      else { rest.x += e[j].x; rest.y += e[j].y; rest.z += e[j].z; rest.w += e[j].w; }
    }
  }
  const float4 v = make_float4(first.x + rest.x, first.y + rest.y, first.z + rest.z, first.w + rest.w);
  if (out_act) {
    uint2 o;
    o.x = Elem<T>::pack2(v.x, v.y);
    o.y = Elem<T>::pack2(v.z, v.w);
    T* row = out_act + ((long long)b * out_bs_rows + t) * (split ? 2 * D : D);
    *reinterpret_cast<uint2*>(row + d) = o;
    if (split) {  // rows are [hi | lo]: what the first 16-bit value lost goes into a second one (split-precision A operand)
      const float2 h0 = Elem<T>::unpack2(o.x), h1 = Elem<T>::unpack2(o.y);
      uint2 l;
      l.x = Elem<T>::pack2(v.x - h0.x, v.y - h0.y);
      l.y = Elem<T>::pack2(v.z - h1.x, v.w - h1.y);
      *reinterpret_cast<uint2*>(row + D + d) = l;
    }
  }
  if (out_f32) {
    float* o = out_f32 + ((long long)b * D + d) * Tn + t;
    o[0] = v.x; o[Tn] = v.y; o[2ll * Tn] = v.z; o[3ll * Tn] = v.w;
  }
}
cudaError_t rvq_decode(const long long* codes, long long sb, long long sk, long long st, const float* tables, void* out_act, float* out_f32, int B,
                       int Q, int Tn, int ncodes, int D, bool bf16, cudaStream_t s, int out_bs_rows, const int* src_index, int* bad_code, int split) {
  if (B <= 0 || Tn <= 0) return cudaSuccess;
  if (out_bs_rows <= 0) out_bs_rows = Tn;
  dim3 grid(Tn, B);
  int threads = ((D / 4 + 31) / 32) * 32;
  if (threads < ((Q + 31) / 32) * 32) threads = ((Q + 31) / 32) * 32;  // the first Q threads stage the frame's codes
  cudaError_t err;
  Q3_DISPATCH(bf16, (err = launch_k(k_rvq<T>, grid, dim3(threads), 0, s, 1, codes, sb, sk, st, tables, (T*)out_act, out_f32, Q, Tn, ncodes, D, out_bs_rows, src_index,
                                    bad_code, split)));
  return err;
}

__global__ void k_widen_i16(const short* __restrict__ in, long long* __restrict__ out, long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) out[i] = in[i];
}
cudaError_t widen_codes_i16(const short* in, long long* out, long long n, cudaStream_t s) {
  if (n <= 0) return cudaSuccess;
  long long blocks = (n + 255) / 256;
  if (blocks > 1184) blocks = 1184;
  k_widen_i16<<<(int)blocks, 256, 0, s>>>(in, out, n);
  return cudaGetLastError();
}

// ============================================================================================ RMSNorm
// One warp per row (reference modeling :382-387, computed in fp32).
template <typename T>
__global__ void k_rmsnorm(const float* __restrict__ x, const float* __restrict__ w, T* __restrict__ out, int M, int H, float eps, int split) {
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= M) return;
  const float4* xr = reinterpret_cast<const float4*>(x + (long long)row * H);
  const int n4 = H >> 2;
  float ss = 0.f;
  for (int i = lane; i < n4; i += 32) {
    const float4 v = xr[i];
    ss += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  ss = warp_sum(ss);
  const float r = rsqrtf(ss / (float)H + eps);
  const float4* wr = reinterpret_cast<const float4*>(w);
  for (int i = lane; i < n4; i += 32) {
    const float4 v = xr[i];
    const float4 g = __ldg(wr + i);
    const float y0 = g.x * (v.x * r), y1 = g.y * (v.y * r), y2 = g.z * (v.z * r), y3 = g.w * (v.w * r);
    uint2 o;
    o.x = Elem<T>::pack2(y0, y1);
    o.y = Elem<T>::pack2(y2, y3);
    T* orow = out + (long long)row * (split ? 2 * H : H);
    *reinterpret_cast<uint2*>(orow + i * 4) = o;
    if (split) {
      const float2 h0 = Elem<T>::unpack2(o.x), h1 = Elem<T>::unpack2(o.y);
      uint2 l;
      l.x = Elem<T>::pack2(y0 - h0.x, y1 - h0.y);
      l.y = Elem<T>::pack2(y2 - h1.x, y3 - h1.y);
      *reinterpret_cast<uint2*>(orow + H + i * 4) = l;
    }
  }
}
cudaError_t rmsnorm(const float* x, const float* w, void* out, int M, int H, float eps, bool bf16, cudaStream_t s, int split) {
  if (M <= 0) return cudaSuccess;
  cudaError_t err;
  Q3_DISPATCH(bf16, (err = launch_k(k_rmsnorm<T>, dim3((M + 7) / 8), dim3(256), 0, s, 1, x, w, (T*)out, M, H, eps, split)));
  return err;
}

// ============================================================================================ RoPE table
// cos/sin of pos * theta^(-2i/hd), fp32, as DecoderRotatoryEmbedding computes them (modeling :268-276).
__global__ void k_rope_table(float* cos_t, float* sin_t, int maxT, int half, float theta, int hd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= maxT * half) return;
  const int pos = i / half, j = i % half;
  const float inv = 1.0f / powf(theta, (float)(2 * j) / (float)hd);
  const float f = (float)pos * inv;
  cos_t[i] = cosf(f);
  sin_t[i] = sinf(f);
}
cudaError_t rope_table(float* cos_t, float* sin_t, int maxT, int hd, float theta, cudaStream_t s) {
  const int n = maxT * (hd / 2);
  k_rope_table<<<(n + 255) / 256, 256, 0, s>>>(cos_t, sin_t, maxT, hd / 2, theta, hd);
  return cudaGetLastError();
}

// ============================================================================================ attention
// RoPE + sliding-window causal attention (modeling :281-353, mask :546-550).  Block = (64 queries, head, item), 4 warps of
// 16 queries each on mma.sync tensor-core tiles (attention_mma.cuh).  K (RoPE applied) and V of the <= 71 + 64 visible keys
// are staged once per block at row (t - kbase), kbase = 16-aligned; rows outside [k_lo, k_hi) are zero.
static constexpr int ATT_QT = 64;
static constexpr int ATT_ROWS = 160;  // ((64 + 71 - 1 + 15) | 15) + 1 key rows at most

template <typename T>
__global__ void __launch_bounds__(128)
k_attention(const T* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t, T* __restrict__ out, int Tn, int nh,
            int window) {
  extern __shared__ __align__(16) uint8_t att_smem[];
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int q0 = blockIdx.x * ATT_QT, h = blockIdx.y, b = blockIdx.z;
  const int k_lo = max(0, q0 - window + 1);
  const int k_hi = min(Tn, q0 + ATT_QT);  // exclusive
  const int kbase = k_lo & ~15;
  const int nrows = ((k_hi - 1 - kbase) | 15) + 1;
  T* sK = reinterpret_cast<T*>(att_smem);
  T* sV = sK + ATT_ROWS * AM_PITCH;
  T* sQ = sV + ATT_ROWS * AM_PITCH;
  const int ld = 3 * nh * AM_HD;
  const T* base = qkv + (long long)b * Tn * ld + h * AM_HD;
  for (int i = threadIdx.x; i < nrows * 4; i += blockDim.x) {
    const int r = i >> 2, part = i & 3, t = kbase + r;
    if (t >= k_lo && t < k_hi) {
      float cs[8], sn[8];
      {
        const float4* c4 = reinterpret_cast<const float4*>(cos_t + t * 32 + part * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sin_t + t * 32 + part * 8);
        const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
        cs[0] = ca.x; cs[1] = ca.y; cs[2] = ca.z; cs[3] = ca.w; cs[4] = cb.x; cs[5] = cb.y; cs[6] = cb.z; cs[7] = cb.w;
        sn[0] = sa.x; sn[1] = sa.y; sn[2] = sa.z; sn[3] = sa.w; sn[4] = sb.x; sn[5] = sb.y; sn[6] = sb.z; sn[7] = sb.w;
      }
      rope_row8<T>(base + (long long)t * ld + nh * AM_HD, cs, sn, sK + r * AM_PITCH, part);
      const uint4* vr = reinterpret_cast<const uint4*>(base + (long long)t * ld + 2 * nh * AM_HD);
      reinterpret_cast<uint4*>(sV + r * AM_PITCH)[part] = vr[part];
      reinterpret_cast<uint4*>(sV + r * AM_PITCH)[part + 4] = vr[part + 4];
    } else {
      zero_row8(sK + r * AM_PITCH, part);
      zero_row8(sV + r * AM_PITCH, part);
    }
  }
  for (int i = threadIdx.x; i < ATT_QT * 4; i += blockDim.x) {
    const int r = i >> 2, part = i & 3, t = q0 + r;
    if (t < Tn) {
      float cs[8], sn[8];
      {
        const float4* c4 = reinterpret_cast<const float4*>(cos_t + t * 32 + part * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sin_t + t * 32 + part * 8);
        const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
        cs[0] = ca.x; cs[1] = ca.y; cs[2] = ca.z; cs[3] = ca.w; cs[4] = cb.x; cs[5] = cb.y; cs[6] = cb.z; cs[7] = cb.w;
        sn[0] = sa.x; sn[1] = sa.y; sn[2] = sa.z; sn[3] = sa.w; sn[4] = sb.x; sn[5] = sb.y; sn[6] = sb.z; sn[7] = sb.w;
      }
      rope_row8<T>(base + (long long)t * ld, cs, sn, sQ + r * AM_PITCH, part);
    } else {
      zero_row8(sQ + r * AM_PITCH, part);
    }
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5;
  const int tq0 = q0 + warp * 16;
  const int nq = min(16, Tn - tq0);
  if (nq > 0)
    attn_warp16<T>(sQ + warp * 16 * AM_PITCH, sK, sV, tq0, nq, kbase, window, out + ((long long)b * Tn + tq0) * (nh * AM_HD) + h * AM_HD,
                   (long long)nh * AM_HD);
}
static constexpr size_t ATT_SMEM = (size_t)(2 * ATT_ROWS + ATT_QT) * AM_PITCH * 2;
// > 48 KB of dynamic shared memory: opt in per device (called from engine_finalize for the handle's device)
cudaError_t attention_setup() {
  cudaError_t e = cudaFuncSetAttribute(k_attention<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATT_SMEM);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_attention<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ATT_SMEM);
  return e;
}
cudaError_t sliding_attention(const void* qkv, const float* cos_t, const float* sin_t, void* out, int B, int Tn, int nh, int hd, int window,
                              bool bf16, cudaStream_t s) {
  if (hd != AM_HD || window < 1 || window > 72) return cudaErrorInvalidValue;  // AM_MAXCH chunks cover a window of at most 72
  if (B <= 0 || Tn <= 0) return cudaSuccess;
  const size_t smem = ATT_SMEM;
  dim3 grid((Tn + ATT_QT - 1) / ATT_QT, nh, B);
  cudaError_t err;
  Q3_DISPATCH(bf16, (err = launch_k(k_attention<T>, grid, dim3(128), smem, s, 1, (const T*)qkv, cos_t, sin_t, (T*)out, Tn, nh, window)));
  return err;
}

}  // namespace q3
