// Streaming (stateful per-packet) decode: KV-cache attention, history-row state copies, position bookkeeping.
// The reference has no stateful decode (its streaming-shaped call re-decodes 25 frames of left context per chunk,
// modeling :885-895); these kernels make every stage causal-with-state so that the concatenated stream output equals
// Decoder.forward(whole sequence) (DESIGN.md "Streaming").
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

static constexpr int SA_WARPS = 4;  // 16 queries per warp: P <= 64 new frames per step

// Block = (head, stream).  Phase 1 appends the P new keys (RoPE applied) and values to the slot's ring cache, phase 2 stages the
// visible keys / values in shared memory at row (t - kbase), kbase 16-aligned, phase 3 is the same warp-level tensor-core
// routine as the bulk kernel (attention_mma.cuh) - same absolute key tiles, hence bit-identical results.
template <typename T>
__global__ void __launch_bounds__(SA_WARPS * 32)
k_stream_attention(const T* __restrict__ qkv, T* __restrict__ kc, T* __restrict__ vc, const int* __restrict__ slot_ids, const int* __restrict__ slot_pos,
                   const float* __restrict__ cos_t, const float* __restrict__ sin_t, int rope_max, float theta, T* __restrict__ out, int P, int nh,
                   int window, int ring, int krows) {
  extern __shared__ __align__(16) uint8_t sa_smem[];
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int h = blockIdx.x, s = blockIdx.y;
  const int slot = slot_ids[s];
  const int pos0 = slot_pos[slot];
  const int H = nh * AM_HD, ld = 3 * H;
  T* sK = reinterpret_cast<T*>(sa_smem);
  T* sV = sK + krows * AM_PITCH;
  T* sQ = sV + krows * AM_PITCH;
  const T* base = qkv + (long long)s * P * ld + h * AM_HD;
  T* kslot = kc + (long long)slot * ring * H + h * AM_HD;
  T* vslot = vc + (long long)slot * ring * H + h * AM_HD;
  const int k_lo = max(0, pos0 - window + 1);
  const int kbase = k_lo & ~15;
  const int nrows = ((pos0 + P - 1 - kbase) | 15) + 1;
  const int qrows = (P + 15) & ~15;
  // ---- phase 1: the packet's keys (rotated) and values go to the ring and, in the same pass, to their shared-memory rows;
  //      the packet's queries are rotated into sQ
  for (int i = threadIdx.x; i < qrows * 4; i += blockDim.x) {
    const int t = i >> 2, part = i & 3;
    if (t < P) {
      const int pos = pos0 + t;
      float cs[8], sn[8];
      if (pos < rope_max) {  // table rows as two float4 each (the same values the bulk kernel reads)
        const float4* c4 = reinterpret_cast<const float4*>(cos_t + pos * 32 + part * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sin_t + pos * 32 + part * 8);
        const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
        cs[0] = ca.x; cs[1] = ca.y; cs[2] = ca.z; cs[3] = ca.w; cs[4] = cb.x; cs[5] = cb.y; cs[6] = cb.z; cs[7] = cb.w;
        sn[0] = sa.x; sn[1] = sa.y; sn[2] = sa.z; sn[3] = sa.w; sn[4] = sb.x; sn[5] = sb.y; sn[6] = sb.z; sn[7] = sb.w;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) rope_cs(cos_t, sin_t, rope_max, theta, pos, part * 8 + j, cs[j], sn[j]);
      }
      T* krow = sK + (pos - kbase) * AM_PITCH;
      rope_row8<T>(base + (long long)t * ld + H, cs, sn, krow, part);
      T* kdst = kslot + (long long)(pos & (ring - 1)) * H;
      reinterpret_cast<uint4*>(kdst)[part] = reinterpret_cast<const uint4*>(krow)[part];
      reinterpret_cast<uint4*>(kdst)[part + 4] = reinterpret_cast<const uint4*>(krow)[part + 4];
      const uint4* vr = reinterpret_cast<const uint4*>(base + (long long)t * ld + 2 * H);
      T* vdst = vslot + (long long)(pos & (ring - 1)) * H;
      const uint4 v0 = vr[part], v1 = vr[part + 4];
      reinterpret_cast<uint4*>(vdst)[part] = v0; reinterpret_cast<uint4*>(vdst)[part + 4] = v1;
      reinterpret_cast<uint4*>(sV + (pos - kbase) * AM_PITCH)[part] = v0; reinterpret_cast<uint4*>(sV + (pos - kbase) * AM_PITCH)[part + 4] = v1;
      rope_row8<T>(base + (long long)t * ld, cs, sn, sQ + t * AM_PITCH, part);
    } else {
      zero_row8(sQ + t * AM_PITCH, part);
    }
  }
  // ---- phase 2: older visible keys / values from the ring; rows outside [k_lo, pos0 + P) are zero
  for (int i = threadIdx.x; i < nrows * 4; i += blockDim.x) {
    const int r = i >> 2, part = i & 3, t = kbase + r;
    if (t >= pos0 && t < pos0 + P) continue;  // written by phase 1
    if (t >= k_lo && t < pos0) {
      const uint4* kr = reinterpret_cast<const uint4*>(kslot + (long long)(t & (ring - 1)) * H);
      const uint4* vr = reinterpret_cast<const uint4*>(vslot + (long long)(t & (ring - 1)) * H);
      reinterpret_cast<uint4*>(sK + r * AM_PITCH)[part] = kr[part]; reinterpret_cast<uint4*>(sK + r * AM_PITCH)[part + 4] = kr[part + 4];
      reinterpret_cast<uint4*>(sV + r * AM_PITCH)[part] = vr[part]; reinterpret_cast<uint4*>(sV + r * AM_PITCH)[part + 4] = vr[part + 4];
    } else {
      zero_row8(sK + r * AM_PITCH, part);
      zero_row8(sV + r * AM_PITCH, part);
    }
  }
  __syncthreads();
  // ---- phase 3
  const int warp = threadIdx.x >> 5;
  const int nq = min(16, P - warp * 16);
  if (nq > 0)
    attn_warp16<T>(sQ + warp * 16 * AM_PITCH, sK, sV, pos0 + warp * 16, nq, kbase, window,
                   out + ((long long)s * P + warp * 16) * H + h * AM_HD, (long long)H);
}

// shared memory of the largest step (P = 64 new frames): opted in per device by stream_attention_setup (engine_finalize)
static constexpr int SA_KROWS_MAX = ((72 - 1 + SA_WARPS * 16 - 1 + 15) | 15) + 1;
static constexpr size_t SA_SMEM_MAX = (size_t)(2 * SA_KROWS_MAX + SA_WARPS * 16) * AM_PITCH * 2;
cudaError_t stream_attention_setup() {
  cudaError_t e = cudaFuncSetAttribute(k_stream_attention<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SA_SMEM_MAX);
  if (e == cudaSuccess) e = cudaFuncSetAttribute(k_stream_attention<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SA_SMEM_MAX);
  return e;
}
cudaError_t stream_attention(const void* qkv, void* kc, void* vc, const int* slot_ids, const int* slot_pos, const float* cos_t, const float* sin_t,
                             int rope_max, float theta, void* out, int S, int P, int nh, int hd, int window, int ring, bool bf16, cudaStream_t s) {
  if (hd != AM_HD || (ring & (ring - 1)) || ring < window - 1 + P || window < 1 || window > 72 || P > SA_WARPS * 16) return cudaErrorInvalidValue;
  if (S <= 0 || P <= 0) return cudaSuccess;
  const int krows = ((window - 1 + P - 1 + 15) | 15) + 1;  // key rows at most: alignment slack + (window - 1) older keys + P new ones
  const size_t smem = (size_t)(2 * krows + ((P + 15) & ~15)) * AM_PITCH * 2;
  if (smem > SA_SMEM_MAX) return cudaErrorInvalidValue;
  dim3 grid(nh, S);
  cudaError_t err;
  Q3_DISPATCH(bf16, (err = launch_k(k_stream_attention<T>, grid, dim3(SA_WARPS * 32), smem, s, 1, (const T*)qkv, (T*)kc, (T*)vc, slot_ids, slot_pos, cos_t, sin_t, rope_max,
                                    theta, (T*)out, P, nh, window, ring, krows)));
  return err;
}

// ---------------------------------------------------------------------------------------------------------------- state copies
// grid (stream, buffer); 16-byte chunks.  save = 0: state -> rows [0, H);  save = 1: rows [n, n + H) -> state.
__global__ void __launch_bounds__(256)
k_state_copy(const StreamBufDesc* __restrict__ descs, unsigned char* __restrict__ state, long long state_bytes, const int* __restrict__ slot_ids, int save) {
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
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
  return launch_k(k_state_copy, grid, dim3(256), 0, s, 1, descs, state, state_bytes, slot_ids, save);
}

__global__ void k_advance_pos(int* slot_pos, const int* slot_ids, int S, int P) {
  pdl_trigger(); pdl_wait();  // programmatic dependent launch (launch.cuh): no prologue worth overlapping, only the launch latency
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < S) slot_pos[slot_ids[i]] += P;
}
cudaError_t stream_advance_pos(int* slot_pos, const int* slot_ids, int S, int P, cudaStream_t s) {
  if (S <= 0) return cudaSuccess;
  return launch_k(k_advance_pos, dim3((S + 127) / 128), dim3(128), 0, s, 1, slot_pos, slot_ids, S, P);
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
