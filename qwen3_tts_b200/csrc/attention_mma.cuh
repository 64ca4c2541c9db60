// Sliding-window causal attention on the warp-level tensor-core path (mma.sync m16n8k16, fp32 accumulate), shared by the
// bulk kernel (kernels_hot.cu) and the streaming kernel (kernels_stream.cu).  Reference: modeling :134-141 (eager attention:
// softmax in fp32 over scores / sqrt(64)), :546-550 (mask kv <= q and kv > q - window).
//
// The work per (item, head) is tiny (125 x <= 72 x 64, twice), so tcgen05 with its 128-row tiles and TMEM round trips does
// not pay here; one warp owns 16 queries and keeps scores, probabilities and the output in registers.
//
// Canonical arithmetic (what makes bulk and streaming bit-identical): keys sit in shared memory at row (t - kbase) with
// kbase a multiple of 16, so every 8-key score tile and every 16-key P.V step covers the same ABSOLUTE key positions in
// both kernels; tiles are visited in ascending order; masked keys get probability exactly 0, so tiles that only one of
// the two kernels visits add exact zeros; an output element depends only on its own query row and key / value rows.
#pragma once
#include "ptx.cuh"

namespace q3 {

static constexpr int AM_HD = 64;
static constexpr int AM_PITCH = AM_HD + 8;  // elements; 144-byte rows: 16-byte aligned for ldmatrix, conflict-free over 8 rows
static constexpr int AM_MAXCH = 7;          // 16-key chunks a warp can see: (window - 1) + 16 queries + alignment slack <= 112 keys

template <typename T> struct MmaOp;
template <> struct MmaOp<__half> {
  __device__ static __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
};
template <> struct MmaOp<__nv_bfloat16> {
  __device__ static __forceinline__ void mma(float (&d)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
};
__device__ __forceinline__ void ldsm_x4(uint32_t saddr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t saddr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(saddr));
}

// RoPE, half-rotation form (modeling :75-110): (x1, x2) = channels (d, d + 32).  Explicit fma so that every call site rounds alike.
__device__ __forceinline__ void rope_pair(float x1, float x2, float c, float s, float& o1, float& o2) {
  o1 = fmaf(x1, c, -(x2 * s));
  o2 = fmaf(x2, c, x1 * s);
}
// One thread rotates 8 + 8 channels (d0..d0+7 and d0+32..d0+39, d0 = 8 * part, part in [0, 4)) of a 64-wide row into a
// shared-memory row.  cs / sn: this position's cos / sin for channels d0..d0+7.
template <typename T>
__device__ __forceinline__ void rope_row8(const T* __restrict__ src, const float (&cs)[8], const float (&sn)[8], T* dst, int part) {
  const uint4 lo = reinterpret_cast<const uint4*>(src)[part], hi = reinterpret_cast<const uint4*>(src)[part + 4];
  const uint32_t l[4] = {lo.x, lo.y, lo.z, lo.w}, h[4] = {hi.x, hi.y, hi.z, hi.w};
  uint32_t ol[4], oh[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 x1 = Elem<T>::unpack2(l[i]), x2 = Elem<T>::unpack2(h[i]);
    float a0, b0, a1, b1;
    rope_pair(x1.x, x2.x, cs[2 * i], sn[2 * i], a0, b0);
    rope_pair(x1.y, x2.y, cs[2 * i + 1], sn[2 * i + 1], a1, b1);
    ol[i] = Elem<T>::pack2(a0, a1);
    oh[i] = Elem<T>::pack2(b0, b1);
  }
  reinterpret_cast<uint4*>(dst)[part] = make_uint4(ol[0], ol[1], ol[2], ol[3]);
  reinterpret_cast<uint4*>(dst)[part + 4] = make_uint4(oh[0], oh[1], oh[2], oh[3]);
}
__device__ __forceinline__ void zero_row8(void* dst, int part) {
  reinterpret_cast<uint4*>(dst)[part] = make_uint4(0, 0, 0, 0);
  reinterpret_cast<uint4*>(dst)[part + 4] = make_uint4(0, 0, 0, 0);
}

// One warp: queries at absolute times tq0 .. tq0 + nq - 1 (nq <= 16) in sQ rows 0..15 (RoPE applied, rows >= nq zero), keys /
// values of absolute time kbase + i in row i of sK (RoPE applied) / sV; rows the warp may touch but that hold no visible key
// are finite (zero-filled or older keys).  Writes out[row * out_ld + 0..63] for the valid rows.
template <typename T>
__device__ __forceinline__ void attn_warp16(const T* sQ, const T* sK, const T* sV, int tq0, int nq, int kbase, int window, T* out, long long out_ld) {
  const int lane = threadIdx.x & 31, g = lane >> 2, c = lane & 3;
  // ---- Q fragments: 4 k-steps of 16 channels
  uint32_t qf[4][4];
  {
    const uint32_t qa = smem_u32(sQ + ((lane & 7) + ((lane >> 3) & 1) * 8) * AM_PITCH + (lane >> 4) * 8);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) ldsm_x4(qa + ks * 32, qf[ks]);
  }
  const int k_first = max(0, tq0 - window + 1);
  const int ch0 = (k_first - kbase) >> 4;                    // first 16-key chunk with a visible key
  const int nch = ((tq0 + nq - 1 - kbase) >> 4) - ch0 + 1;   // chunks up to the last query's own key
  const int row_lo = tq0 + min(g, nq - 1), row_hi = tq0 + min(g + 8, nq - 1);  // padded rows repeat the last query (never stored)
  // ---- scores: S[ch][half][4], half = 8-key tile inside the chunk
  float sc[AM_MAXCH][2][4];
  float mx_lo = -INFINITY, mx_hi = -INFINITY;
#pragma unroll
  for (int ch = 0; ch < AM_MAXCH; ++ch) {
    if (ch < nch) {
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int krow = (ch0 + ch) * 16 + hf * 8;
        const uint32_t ka = smem_u32(sK + (krow + (lane & 7)) * AM_PITCH + (lane >> 3) * 8);
        uint32_t kf0[4], kf1[4];
        ldsm_x4(ka, kf0);        // channels 0..31: (b0, b1) of k-steps 0, 1
        ldsm_x4(ka + 64, kf1);   // channels 32..63: k-steps 2, 3
        float (&d)[4] = sc[ch][hf];
        d[0] = d[1] = d[2] = d[3] = 0.f;
        MmaOp<T>::mma(d, qf[0], kf0[0], kf0[1]);
        MmaOp<T>::mma(d, qf[1], kf0[2], kf0[3]);
        MmaOp<T>::mma(d, qf[2], kf1[0], kf1[1]);
        MmaOp<T>::mma(d, qf[3], kf1[2], kf1[3]);
        const int key = kbase + krow + 2 * c;  // absolute key of d[0] / d[2]; d[1] / d[3] are key + 1
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int kk = key + (e & 1), tq = e < 2 ? row_lo : row_hi;
          const float v = (kk <= tq && kk > tq - window) ? d[e] * 0.125f : -INFINITY;
          d[e] = v;
          if (e < 2) mx_lo = fmaxf(mx_lo, v); else mx_hi = fmaxf(mx_hi, v);
        }
      }
    }
  }
  mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 1)); mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 2));
  mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 1)); mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 2));
  // ---- probabilities (fp32 exp, sums over the unrounded values), P.V with P rounded to the operand type
  float sum_lo = 0.f, sum_hi = 0.f;
  float o[8][4];
#pragma unroll
  for (int n = 0; n < 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
#pragma unroll
  for (int ch = 0; ch < AM_MAXCH; ++ch) {
    if (ch < nch) {
      uint32_t pf[4];
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const float p0 = __expf(sc[ch][hf][0] - mx_lo), p1 = __expf(sc[ch][hf][1] - mx_lo);
        const float p2 = __expf(sc[ch][hf][2] - mx_hi), p3 = __expf(sc[ch][hf][3] - mx_hi);
        sum_lo += p0 + p1;
        sum_hi += p2 + p3;
        pf[2 * hf] = Elem<T>::pack2(p0, p1);      // a0 / a2: row g
        pf[2 * hf + 1] = Elem<T>::pack2(p2, p3);  // a1 / a3: row g + 8
      }
      const int krow = (ch0 + ch) * 16;
      const uint32_t va = smem_u32(sV + (krow + (lane & 7) + ((lane >> 3) & 1) * 8) * AM_PITCH + (lane >> 4) * 8);
#pragma unroll
      for (int n2 = 0; n2 < 4; ++n2) {  // two 8-channel output tiles per ldmatrix
        uint32_t vf[4];
        ldsm_x4_t(va + n2 * 32, vf);
        MmaOp<T>::mma(o[2 * n2], pf, vf[0], vf[1]);
        MmaOp<T>::mma(o[2 * n2 + 1], pf, vf[2], vf[3]);
      }
    }
  }
  sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 1); sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 2);
  sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 1); sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 2);
  const float inv_lo = 1.0f / sum_lo, inv_hi = 1.0f / sum_hi;
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    if (g < nq) *reinterpret_cast<uint32_t*>(out + (long long)g * out_ld + n * 8 + 2 * c) = Elem<T>::pack2(o[n][0] * inv_lo, o[n][1] * inv_lo);
    if (g + 8 < nq) *reinterpret_cast<uint32_t*>(out + (long long)(g + 8) * out_ld + n * 8 + 2 * c) = Elem<T>::pack2(o[n][2] * inv_hi, o[n][3] * inv_hi);
  }
}

}  // namespace q3
