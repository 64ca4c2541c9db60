// Streaming (stateful per-packet) decode: KV-cache attention, history-row state copies, position bookkeeping.
// The reference has no stateful decode (its streaming-shaped call re-decodes 25 frames of left context per chunk,
// modeling :885-895); these kernels make every stage causal-with-state so that the concatenated stream output equals
// Decoder.forward(whole sequence) (DESIGN.md "Streaming").
#include "kernels.h"
#include "ptx.cuh"

namespace q3 {

#define Q3_DISPATCH(bf16, ...)                                   \
  do {                                                            \
    if (bf16) { using T = __nv_bfloat16; __VA_ARGS__; }           \
    else { using T = __half; __VA_ARGS__; }                       \
  } while (0)

__device__ __forceinline__ float wsum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float wmax(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// cos/sin of pos * theta^(-2d/64): table below rope_max (bit-identical to the bulk path), direct evaluation above
__device__ __forceinline__ void rope_cs(const float* __restrict__ cos_t, const float* __restrict__ sin_t, int rope_max, float theta, int pos, int d,
                                        float& c, float& s) {
  if (pos < rope_max) {
    c = cos_t[pos * 32 + d];
    s = sin_t[pos * 32 + d];
  } else {
    const float f = (float)pos * (1.0f / powf(theta, (float)(2 * d) / 64.0f));
    c = cosf(f);
    s = sinf(f);
  }
}

static constexpr int SA_HD = 64;
static constexpr int SA_KSTR = SA_HD + 2;  // odd word pitch: conflict-free lane-per-key reads
static constexpr int SA_WARPS = 4;

// Block = (head, stream).  Phase 1 appends the P new keys (RoPE applied) and values to the slot's ring cache, phase 2 stages the
// visible keys / values in shared memory, phase 3: one warp per query, one lane per key, fp32 softmax (reference modeling :134-141).
template <typename T>
__global__ void __launch_bounds__(SA_WARPS * 32)
k_stream_attention(const T* __restrict__ qkv, T* __restrict__ kc, T* __restrict__ vc, const int* __restrict__ slot_ids, const int* __restrict__ slot_pos,
                   const float* __restrict__ cos_t, const float* __restrict__ sin_t, int rope_max, float theta, T* __restrict__ out, int P, int nh,
                   int window, int ring) {
  extern __shared__ __align__(16) uint8_t sa_smem[];
  const int h = blockIdx.x, s = blockIdx.y;
  const int slot = slot_ids[s];
  const int pos0 = slot_pos[slot];
  const int H = nh * SA_HD, ld = 3 * H;
  const int maxk = window - 1 + P;
  T* sK = reinterpret_cast<T*>(sa_smem);
  T* sV = sK + (((size_t)maxk * SA_KSTR + 7) & ~(size_t)7);
  float* sQ = reinterpret_cast<float*>(sV + (size_t)maxk * SA_HD);  // [warps][64]
  const int pstride = ((maxk + 31) / 32) * 32;
  float* sP = sQ + SA_WARPS * SA_HD;                                 // [warps][pstride]
  const T* base = qkv + (long long)s * P * ld;
  T* kslot = kc + (long long)slot * ring * H + h * SA_HD;
  T* vslot = vc + (long long)slot * ring * H + h * SA_HD;
  // ---- phase 1: append
  for (int i = threadIdx.x; i < P * 32; i += blockDim.x) {
    const int t = i >> 5, d = i & 31;
    const int pos = pos0 + t;
    const T* kr = base + (long long)t * ld + H + h * SA_HD;
    const float x1 = Elem<T>::to_f(kr[d]), x2 = Elem<T>::to_f(kr[d + 32]);
    float c, sn;
    rope_cs(cos_t, sin_t, rope_max, theta, pos, d, c, sn);
    T* dst = kslot + (long long)(pos & (ring - 1)) * H;
    dst[d] = Elem<T>::from_f(x1 * c - x2 * sn);
    dst[d + 32] = Elem<T>::from_f(x2 * c + x1 * sn);
  }
  for (int i = threadIdx.x; i < P * 8; i += blockDim.x) {
    const int t = i >> 3, c8 = i & 7;
    const T* vr = base + (long long)t * ld + 2 * H + h * SA_HD;
    reinterpret_cast<uint4*>(vslot + (long long)((pos0 + t) & (ring - 1)) * H)[c8] = reinterpret_cast<const uint4*>(vr)[c8];
  }
  __syncthreads();
  // ---- phase 2: stage the keys every new query can see
  const int k_lo = max(0, pos0 - window + 1);
  const int nk = pos0 + P - k_lo;
  for (int i = threadIdx.x; i < nk * 32; i += blockDim.x) {
    const int kk = i >> 5, d2 = i & 31;
    const uint32_t w = reinterpret_cast<const uint32_t*>(kslot + (long long)((k_lo + kk) & (ring - 1)) * H)[d2];
    reinterpret_cast<uint32_t*>(sK + (size_t)kk * SA_KSTR)[d2] = w;
  }
  for (int i = threadIdx.x; i < nk * 8; i += blockDim.x) {
    const int kk = i >> 3, c8 = i & 7;
    reinterpret_cast<uint4*>(sV + (size_t)kk * SA_HD)[c8] = reinterpret_cast<const uint4*>(vslot + (long long)((k_lo + kk) & (ring - 1)) * H)[c8];
  }
  __syncthreads();
  // ---- phase 3
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* q = sQ + warp * SA_HD;
  float* pr = sP + warp * pstride;
  for (int qi = warp; qi < P; qi += SA_WARPS) {
    const int tq = pos0 + qi;
    {
      const T* qr = base + (long long)qi * ld + h * SA_HD;
      const float x1 = Elem<T>::to_f(qr[lane]), x2 = Elem<T>::to_f(qr[lane + 32]);
      float c, sn;
      rope_cs(cos_t, sin_t, rope_max, theta, tq, lane, c, sn);
      q[lane] = x1 * c - x2 * sn;
      q[lane + 32] = x2 * c + x1 * sn;
    }
    __syncwarp();
    const int lo = max(0, tq - window + 1);
    const int nvis = tq - lo + 1;
    float mx = -INFINITY;
    for (int j0 = 0; j0 < nvis; j0 += 32) {
      const int j = j0 + lane;
      float sc = -INFINITY;
      if (j < nvis) {
        const uint32_t* kr = reinterpret_cast<const uint32_t*>(sK + (size_t)(lo - k_lo + j) * SA_KSTR);
        float acc = 0.f;
#pragma unroll
        for (int d2 = 0; d2 < 32; ++d2) {
          const float2 kv = Elem<T>::unpack2(kr[d2]);
          acc = fmaf(q[2 * d2], kv.x, acc);
          acc = fmaf(q[2 * d2 + 1], kv.y, acc);
        }
        sc = acc * 0.125f;
      }
      pr[j] = sc;
      mx = fmaxf(mx, sc);
    }
    mx = wmax(mx);
    float sum = 0.f;
    for (int j0 = 0; j0 < nvis; j0 += 32) {
      const int j = j0 + lane;
      if (j < nvis) {
        const float e = __expf(pr[j] - mx);
        pr[j] = e;
        sum += e;
      }
    }
    sum = wsum(sum);
    __syncwarp();
    const float inv = 1.0f / sum;
    float o0 = 0.f, o1 = 0.f;
    const uint32_t* vb = reinterpret_cast<const uint32_t*>(sV + (size_t)(lo - k_lo) * SA_HD) + lane;
    for (int j = 0; j < nvis; ++j) {
      const float pj = pr[j];
      const float2 vv = Elem<T>::unpack2(vb[j * (SA_HD / 2)]);
      o0 = fmaf(pj, vv.x, o0);
      o1 = fmaf(pj, vv.y, o1);
    }
    reinterpret_cast<uint32_t*>(out + ((long long)s * P + qi) * H + h * SA_HD)[lane] = Elem<T>::pack2(o0 * inv, o1 * inv);
    __syncwarp();
  }
}

cudaError_t stream_attention(const void* qkv, void* kc, void* vc, const int* slot_ids, const int* slot_pos, const float* cos_t, const float* sin_t,
                             int rope_max, float theta, void* out, int S, int P, int nh, int hd, int window, int ring, bool bf16, cudaStream_t s) {
  if (hd != SA_HD || (ring & (ring - 1)) || ring < window - 1 + P) return cudaErrorInvalidValue;
  if (S <= 0 || P <= 0) return cudaSuccess;
  const int maxk = window - 1 + P;
  const int pstride = ((maxk + 31) / 32) * 32;
  const size_t smem = ((((size_t)maxk * SA_KSTR + 7) & ~(size_t)7) + (size_t)maxk * SA_HD) * 2 + SA_WARPS * SA_HD * 4 + (size_t)SA_WARPS * pstride * 4;
  if (smem > 48 * 1024) return cudaErrorInvalidValue;
  dim3 grid(nh, S);
  Q3_DISPATCH(bf16, (k_stream_attention<T><<<grid, SA_WARPS * 32, smem, s>>>((const T*)qkv, (T*)kc, (T*)vc, slot_ids, slot_pos, cos_t, sin_t, rope_max, theta,
                                                                               (T*)out, P, nh, window, ring)));
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------------------------- state copies
// grid (stream, buffer); 16-byte chunks.  save = 0: state -> rows [0, H);  save = 1: rows [n, n + H) -> state.
__global__ void __launch_bounds__(256)
k_state_copy(const StreamBufDesc* __restrict__ descs, unsigned char* __restrict__ state, long long state_bytes, const int* __restrict__ slot_ids, int save) {
  const StreamBufDesc d = descs[blockIdx.y];
  const int s = blockIdx.x;
  const long long bytes = (long long)d.H * d.row_bytes;
  unsigned char* st = state + (long long)slot_ids[s] * state_bytes + d.state_off;
  unsigned char* buf = d.ptr + (long long)s * d.batch_stride + (save ? (long long)d.n * d.row_bytes : 0);
  const uint4* src = reinterpret_cast<const uint4*>(save ? buf : st);
  uint4* dst = reinterpret_cast<uint4*>(save ? st : buf);
  for (long long i = threadIdx.x; i < bytes / 16; i += blockDim.x) dst[i] = src[i];
}
cudaError_t stream_state_copy(const StreamBufDesc* descs, int ndesc, unsigned char* state, long long state_bytes, const int* slot_ids, int S, int save,
                              cudaStream_t s) {
  if (S <= 0 || ndesc <= 0) return cudaSuccess;
  dim3 grid(S, ndesc);
  k_state_copy<<<grid, 256, 0, s>>>(descs, state, state_bytes, slot_ids, save);
  return cudaGetLastError();
}

__global__ void k_advance_pos(int* slot_pos, const int* slot_ids, int S, int P) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < S) slot_pos[slot_ids[i]] += P;
}
cudaError_t stream_advance_pos(int* slot_pos, const int* slot_ids, int S, int P, cudaStream_t s) {
  if (S <= 0) return cudaSuccess;
  k_advance_pos<<<(S + 127) / 128, 128, 0, s>>>(slot_pos, slot_ids, S, P);
  return cudaGetLastError();
}

__global__ void __launch_bounds__(256)
k_scatter_rows(const float* __restrict__ in, long long in_stride, float* __restrict__ out, long long out_stride, const int* __restrict__ index, long long len) {
  const int b = blockIdx.y;
  const float* src = in + (long long)b * in_stride;
  float* dst = out + (long long)index[b] * out_stride;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (long long)gridDim.x * blockDim.x) dst[i] = src[i];
}
cudaError_t scatter_rows_f32(const float* in, long long in_stride, float* out, long long out_stride, const int* index, int B, long long len, cudaStream_t s) {
  if (B <= 0 || len <= 0) return cudaSuccess;
  int bx = (int)((len + 255) / 256);
  if (bx > 64) bx = 64;
  dim3 grid(bx, B);
  k_scatter_rows<<<grid, 256, 0, s>>>(in, in_stride, out, out_stride, index, len);
  return cudaGetLastError();
}

}  // namespace q3
