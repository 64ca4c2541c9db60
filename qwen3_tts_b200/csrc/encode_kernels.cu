// SIMT kernels of the 12 Hz ENCODE path: everything between the tensor-core GEMMs of encode_plan.cu.
//
// Reference: Qwen3TTSTokenizerV2Model.encode (modeling :960-990) -> MimiModel.encode of the `transformers` package (not vendored;
// transformers/models/mimi/modeling_mimi.py): MimiEncoder (:454-496), MimiTransformerLayer (:926-993), MimiAttention (:645-738),
// MimiRotaryEmbedding (:515-576), the replicate-padded down-sampling MimiConv1d (:1421-1431).
//
// The encode path ends in an argmin over codebooks, so its arithmetic has to stay at fp32 level: activations live in HBM as fp32
// [B][L][C] (channels last) and every GEMM reads them as fp16 [hi | lo] pairs (gemm_params.h, k_mult 3: ~22 bits per operand).
// These kernels apply the element-wise function that sits in front of a GEMM (ELU, GELU, LayerNorm, attention) and write that
// operand directly, so each activation makes one fp32 round trip per layer.
#include <cuda_fp16.h>

#include "attention_mma.cuh"
#include "encode_kernels.h"
#include "launch.cuh"

namespace q3 {

static inline int enc_blocks(long long n, int threads, int cap = 148 * 16) {
  long long b = (n + threads - 1) / threads;
  if (b > cap) b = cap;
  return b < 1 ? 1 : (int)b;
}

__device__ __forceinline__ void split_hi_lo(float v, __half& hi, __half& lo) {
  hi = __float2half_rn(v);
  lo = __float2half_rn(v - __half2float(hi));
}
// four values at once with the packed converts: hi[0..3] and lo[0..3] as two 8-byte words each
__device__ __forceinline__ void split4(const float4 v, uint2& hi, uint2& lo) {
  const __half2 h01 = __floats2half2_rn(v.x, v.y), h23 = __floats2half2_rn(v.z, v.w);
  const float2 b01 = __half22float2(h01), b23 = __half22float2(h23);
  const __half2 l01 = __floats2half2_rn(v.x - b01.x, v.y - b01.y), l23 = __floats2half2_rn(v.z - b23.x, v.w - b23.y);
  hi.x = *reinterpret_cast<const unsigned*>(&h01); hi.y = *reinterpret_cast<const unsigned*>(&h23);
  lo.x = *reinterpret_cast<const unsigned*>(&l01); lo.y = *reinterpret_cast<const unsigned*>(&l23);
}

template <int ACT>
__device__ __forceinline__ float enc_act(float v) {
  // ELU (alpha 1).  exp(v) - 1 with the hardware exponential: absolute error <= ~2.4e-7 (2 ulp of a value below 1), the level of the
  // fp16 hi + lo pair the result is stored as; expm1f costs ~10x the instructions and made these passes instruction-bound (2.9 TB/s)
  if (ACT == 1) return v > 0.f ? v : __expf(v) - 1.0f;
  if (ACT == 2) return 0.5f * v * (1.0f + erff(v * 0.70710678118654752440f));
  return v;
}
// ------------------------------------------------------------------------------------------------ first conv: 1 -> C0 channels, k = 7
// wav [B][wav_bs] fp32 (samples beyond T are not read), w [C0][1][7], out [B][T][C0] fp32.  Causal: x[t - 6 + j] (MimiConv1d :331-352).
__global__ void __launch_bounds__(256) k_enc_conv0(const float* __restrict__ wav, long long wav_bs, const float* __restrict__ w,
                                                    const float* __restrict__ bias, float* __restrict__ out, __half* __restrict__ elu_split, int T,
                                                    int C0) {
  // grid (sample tiles, items); a block = (256 / (C0 / 4)) * 8 consecutive samples of one item; thread = 4 channels, 8 samples apart
  extern __shared__ float sw[];  // [7][C0] weights, [C0] bias, then the block's input samples (+ 6 of history)
  const int c4n = C0 >> 2, spb = 256 / c4n, tile = spb * 8;
  float* sx = sw + 8 * C0;
  for (int i = threadIdx.x; i < 7 * C0; i += blockDim.x) { const int co = i % C0, j = i / C0; sw[i] = w[co * 7 + j]; }
  for (int i = threadIdx.x; i < C0; i += blockDim.x) sw[7 * C0 + i] = bias[i];
  const long long b = blockIdx.y;
  const int t0 = blockIdx.x * tile;
  const float* x = wav + b * wav_bs;
  for (int i = threadIdx.x; i < tile + 6; i += blockDim.x) {
    const int ts = t0 - 6 + i;
    sx[i] = (ts >= 0 && ts < T) ? __ldg(x + ts) : 0.f;     // causal left padding (MimiConv1d :331-352)
  }
  __syncthreads();
  const int c4 = threadIdx.x % c4n, s0 = threadIdx.x / c4n;
  float4 wv[7];
#pragma unroll
  for (int j = 0; j < 7; ++j) wv[j] = *reinterpret_cast<const float4*>(sw + j * C0 + 4 * c4);
  const float4 bv = *reinterpret_cast<const float4*>(sw + 7 * C0 + 4 * c4);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int ls = s0 + k * spb, t = t0 + ls;
    if (t >= T) break;
    float4 acc = bv;
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      const float xv = sx[ls + j];
      acc.x = fmaf(xv, wv[j].x, acc.x); acc.y = fmaf(xv, wv[j].y, acc.y); acc.z = fmaf(xv, wv[j].z, acc.z); acc.w = fmaf(xv, wv[j].w, acc.w);
    }
    const long long bt = b * T + t;
    *reinterpret_cast<float4*>(out + bt * C0 + 4 * c4) = acc;
    if (elu_split) {  // the first ResnetBlock's conv operand: ELU(x) as [hi | lo] (saves a pass over the largest tensor of the path)
      uint2 h, l;
      split4(make_float4(enc_act<1>(acc.x), enc_act<1>(acc.y), enc_act<1>(acc.z), enc_act<1>(acc.w)), h, l);
      __half* o = elu_split + bt * 2 * C0 + 4 * c4;
      *reinterpret_cast<uint2*>(o) = h;
      *reinterpret_cast<uint2*>(o + C0) = l;
    }
  }
}
cudaError_t enc_conv0(const float* wav, long long wav_bs, const float* w, const float* bias, float* out, void* elu_split, int B, int T, int C0,
                      cudaStream_t s) {
  if (C0 % 4 || 256 % (C0 / 4) || B > 65535) return cudaErrorInvalidValue;
  const int tile = 256 / (C0 / 4) * 8;
  dim3 grid((T + tile - 1) / tile, B);
  k_enc_conv0<<<grid, 256, (size_t)(8 * C0 + tile + 6) * sizeof(float), s>>>(wav, wav_bs, w, bias, out, (__half*)elu_split, T, C0);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ act + split (+ stride regrouping)
// in fp32 [B][L][C] -> out fp16 [B][lead + ceil(L / r)][2 * r * C]: super-row s holds input rows (s - lead) * r + p, p < r, as
// [hi(p, c) | lo(p, c)].  A stride-r, kernel-2r causal conv is then a two-tap GEMM over super-rows (encode_plan.cu).  Rows outside
// [0, L) are zero (the conv's constant padding, applied AFTER the activation like the reference does) or, with replicate = 1, the
// nearest valid row (the down-sampling conv's pad_mode "replicate").  act: 0 none, 1 ELU (alpha 1), 2 GELU (erf).
// cg (K group, a divisor of r * C): the super-row is written group by group, [hi(g) | lo(g)] for every cg input values, matching the
// grouped weight rows (kernels.cu put_w) - one GEMM launch then walks one group.  cg = r * C: a single [hi | lo] pair per row.
template <int ACT>
__global__ void __launch_bounds__(256) k_act_split(const float* __restrict__ in, __half* __restrict__ out, int B, int L, int C, int r, int lead,
                                                    int replicate, int cg, int rows_per_block) {
  // block (tx, ty): tx walks the float4 groups of a super-row, ty the block's rows_per_block consecutive super-rows.  Every division
  // sits outside the row loop (this pass was bound by its index arithmetic): a thread's column decomposition is fixed, its
  // (item, row) pair is advanced incrementally.  32-bit arithmetic: B * ns < 2^31 is checked by the launcher.
  const int ns = lead + (L + r - 1) / r;
  const int rc = r * C, w4 = rc >> 2;
  const int nrows = B * ns;
  const int row_end = min(nrows, ((int)blockIdx.x + 1) * rows_per_block);
  const bool one_group = cg >= rc;
  for (int j4 = threadIdx.x; j4 < w4; j4 += blockDim.x) {
    const int j = j4 << 2;
    const int p = r == 1 ? 0 : j / C;
    const int c = j - p * C;
    const int g = one_group ? 0 : j / cg;
    const int ocol = g * 2 * cg + (j - g * cg);
    int srow = blockIdx.x * rows_per_block + threadIdx.y;
    int b = srow / ns, sr = srow - b * ns;
    for (; srow < row_end; srow += blockDim.y) {
      int row = (sr - lead) * r + p;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      bool valid = row >= 0 && row < L;
      if (!valid && replicate) { row = row < 0 ? 0 : L - 1; valid = true; }
      if (valid) {
        v = __ldcs(reinterpret_cast<const float4*>(in + (((long long)b * L + row) * C + c)));
        v.x = enc_act<ACT>(v.x); v.y = enc_act<ACT>(v.y); v.z = enc_act<ACT>(v.z); v.w = enc_act<ACT>(v.w);
      }
      uint2 h, l;
      split4(v, h, l);
      __half* o = out + (long long)srow * (2 * rc) + ocol;
      *reinterpret_cast<uint2*>(o) = h;
      *reinterpret_cast<uint2*>(o + cg) = l;
      sr += blockDim.y;
      while (sr >= ns) { sr -= ns; ++b; }
    }
  }
}
cudaError_t enc_act_split(const float* in, void* out, int B, int L, int C, int act, int r, int lead, int replicate, int cg, cudaStream_t s) {
  if (cg <= 0) cg = r * C;
  if ((r * C) % cg || cg % 4 || C % 4) return cudaErrorInvalidValue;
  const long long nrows = (long long)B * (lead + (L + r - 1) / r);
  const int w4 = (r * C) / 4;
  int tx = 1;
  while (tx < w4 && tx < 256) tx <<= 1;        // power of two >= min(w4, 256): idle lanes only when w4 is not one (r = 5, 6)
  const int ty = 256 / tx;
  const int rpb = ty * 8;                      // eight rows per thread
  const long long blocks = (nrows + rpb - 1) / rpb;
  if (blocks > 0x7fffffffll || nrows > 0x7fffffffll) return cudaErrorInvalidValue;
  const dim3 block(tx, ty);
  const unsigned g = (unsigned)(blocks < 1 ? 1 : blocks);
  if (act == 1) k_act_split<1><<<g, block, 0, s>>>(in, (__half*)out, B, L, C, r, lead, replicate, cg, rpb);
  else if (act == 2) k_act_split<2><<<g, block, 0, s>>>(in, (__half*)out, B, L, C, r, lead, replicate, cg, rpb);
  else k_act_split<0><<<g, block, 0, s>>>(in, (__half*)out, B, L, C, r, lead, replicate, cg, rpb);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ LayerNorm (with bias) -> [hi | lo]
// x fp32 [M][H] -> out fp16 [M][2H]; one warp per row, the row stays in registers between the two passes (H <= 1024, H % 128 == 0).
__global__ void __launch_bounds__(256) k_layernorm_split(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                                          __half* __restrict__ out, long long M, int H, float eps) {
  const int lane = threadIdx.x & 31;
  const long long row = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= M) return;
  const int nv = H >> 7;  // float4 per lane
  float4 v[8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (i >= nv) break;
    v[i] = *reinterpret_cast<const float4*>(x + row * H + (i * 32 + lane) * 4);
    sum += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  const float mean = sum / (float)H;
  float sq = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (i >= nv) break;
    v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
    sq = fmaf(v[i].x, v[i].x, sq); sq = fmaf(v[i].y, v[i].y, sq); sq = fmaf(v[i].z, v[i].z, sq); sq = fmaf(v[i].w, v[i].w, sq);
  }
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
  const float rstd = 1.0f / sqrtf(sq / (float)H + eps);   // biased variance, eps inside the root (torch.nn.LayerNorm)
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (i >= nv) break;
    const int c = (i * 32 + lane) * 4;
    const float4 wv = *reinterpret_cast<const float4*>(w + c);
    const float4 bv = *reinterpret_cast<const float4*>(bias + c);
    const float y[4] = {fmaf(v[i].x * rstd, wv.x, bv.x), fmaf(v[i].y * rstd, wv.y, bv.y), fmaf(v[i].z * rstd, wv.z, bv.z), fmaf(v[i].w * rstd, wv.w, bv.w)};
    __half h[4], l[4];
    for (int k = 0; k < 4; ++k) split_hi_lo(y[k], h[k], l[k]);
    *reinterpret_cast<uint2*>(out + row * 2 * H + c) = *reinterpret_cast<const uint2*>(h);
    *reinterpret_cast<uint2*>(out + row * 2 * H + H + c) = *reinterpret_cast<const uint2*>(l);
  }
}
cudaError_t enc_layernorm_split(const float* x, const float* w, const float* bias, void* out, long long M, int H, float eps, cudaStream_t s) {
  if (H % 128 || H > 1024) return cudaErrorInvalidValue;
  k_layernorm_split<<<(unsigned)((M + 7) / 8), 256, 0, s>>>(x, w, bias, (__half*)out, M, H, eps);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ RoPE table
// cos / sin [maxT][hd / 2] with the reference's float32 arithmetic (MimiRotaryEmbedding :515-576): inv_freq = 1 / theta^(2i / hd) in
// float32, angle = float32(pos) * inv_freq in float32, cos / sin of that float32 angle.
__global__ void k_enc_rope_table(float* __restrict__ cos_t, float* __restrict__ sin_t, const float* __restrict__ inv_freq, int maxT, int half) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= maxT * half) return;
  const float ang = __fmul_rn((float)(i / half), inv_freq[i % half]);
  cos_t[i] = (float)cos((double)ang);
  sin_t[i] = (float)sin((double)ang);
}
cudaError_t enc_rope_table(float* cos_t, float* sin_t, const float* inv_freq_dev, int maxT, int hd, cudaStream_t s) {
  const int n = maxT * (hd / 2);
  k_enc_rope_table<<<(n + 255) / 256, 256, 0, s>>>(cos_t, sin_t, inv_freq_dev, maxT, hd / 2);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ attention, fp32
// qkv fp32 [B*T][3*H] (q | k | v, H = nh * 64) -> out fp16 [B*T][2H] rows [hi | lo] (the A operand of o_proj).
// Block = (32-query tile, head, item), 8 warps x 4 queries.  Keys are walked in chunks of 64 staged in shared memory (RoPE applied on
// the way in), with a running max / sum per query (the usual online softmax); query i sees keys j with 0 <= i - j < window.
static constexpr int EA_Q = 32, EA_K = 64, EA_HD = 64, EA_PAD = EA_HD + 4;  // row pitch 68 floats: float4 reads of 8 consecutive rows hit 8 different bank groups
__global__ void __launch_bounds__(256) k_enc_attention(const float* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                                        __half* __restrict__ out, int T, int nh, int window) {
  __shared__ __align__(16) float sq[EA_Q][EA_HD];
  __shared__ __align__(16) float sk[EA_K][EA_PAD];
  __shared__ __align__(16) float sv[EA_K][EA_HD];
  const int q0 = blockIdx.x * EA_Q, head = blockIdx.y, b = blockIdx.z;
  const int H = nh * EA_HD;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const float* base = qkv + (long long)b * T * 3 * H + head * EA_HD;
  // queries of the tile, rotated: out[d] = x[d] cos - x[d + 32] sin, out[d + 32] = x[d + 32] cos + x[d] sin   (rotate_half, :579-584)
  for (int i = tid; i < EA_Q * 32; i += 256) {
    const int qi = i >> 5, d = i & 31, t = q0 + qi;
    float a = 0.f, c = 0.f;
    if (t < T) {
      const float x1 = base[(long long)t * 3 * H + d], x2 = base[(long long)t * 3 * H + d + 32];
      const float cs = cos_t[t * 32 + d], sn = sin_t[t * 32 + d];
      a = x1 * cs - x2 * sn; c = x2 * cs + x1 * sn;
    }
    sq[qi][d] = a; sq[qi][d + 32] = c;
  }
  float m[4], l[4], acc0[4], acc1[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { m[i] = -3.0e38f; l[i] = 0.f; acc0[i] = 0.f; acc1[i] = 0.f; }
  const int qlast = min(q0 + EA_Q - 1, T - 1);
  int kbeg = q0 - window + 1;
  if (kbeg < 0) kbeg = 0;
  for (int k0 = kbeg; k0 <= qlast; k0 += EA_K) {
    __syncthreads();  // previous chunk fully consumed (and sq written, first time round)
    for (int i = tid; i < EA_K * 32; i += 256) {
      const int ki = i >> 5, d = i & 31, t = k0 + ki;
      float a = 0.f, c = 0.f, v1 = 0.f, v2 = 0.f;
      if (t <= qlast) {
        const float* row = base + (long long)t * 3 * H;
        const float x1 = row[H + d], x2 = row[H + d + 32];
        const float cs = cos_t[t * 32 + d], sn = sin_t[t * 32 + d];
        a = x1 * cs - x2 * sn; c = x2 * cs + x1 * sn;
        v1 = row[2 * H + d]; v2 = row[2 * H + d + 32];
      }
      sk[ki][d] = a; sk[ki][d + 32] = c;
      sv[ki][d] = v1; sv[ki][d + 32] = v2;
    }
    __syncthreads();
    float pr0[4], pr1[4];  // this lane's two probabilities per query (keys k0 + lane, k0 + lane + 32); broadcast by shuffle below
    float sc0[4] = {0.f, 0.f, 0.f, 0.f}, sc1[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int d = 0; d < EA_HD; d += 4) {      // one float4 of each of this lane's two keys against the warp's four queries
      const float4 ka = *reinterpret_cast<const float4*>(&sk[lane][d]);
      const float4 kb = *reinterpret_cast<const float4*>(&sk[lane + 32][d]);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 qv = *reinterpret_cast<const float4*>(&sq[warp * 4 + i][d]);
        sc0[i] = fmaf(qv.x, ka.x, sc0[i]); sc0[i] = fmaf(qv.y, ka.y, sc0[i]); sc0[i] = fmaf(qv.z, ka.z, sc0[i]); sc0[i] = fmaf(qv.w, ka.w, sc0[i]);
        sc1[i] = fmaf(qv.x, kb.x, sc1[i]); sc1[i] = fmaf(qv.y, kb.y, sc1[i]); sc1[i] = fmaf(qv.z, kb.z, sc1[i]); sc1[i] = fmaf(qv.w, kb.w, sc1[i]);
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int qi = warp * 4 + i, t = q0 + qi;
      float s0 = sc0[i], s1 = sc1[i];
      const int j0 = k0 + lane, j1 = k0 + lane + 32;
      const bool ok0 = t < T && j0 <= t && t - j0 < window, ok1 = t < T && j1 <= t && t - j1 < window;
      s0 = ok0 ? s0 * 0.125f : -3.0e38f;   // scaling = 1 / sqrt(head_dim) (:665)
      s1 = ok1 ? s1 * 0.125f : -3.0e38f;
      float mx = fmaxf(s0, s1);
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      const float mnew = fmaxf(m[i], mx);
      const float p0 = ok0 ? expf(s0 - mnew) : 0.f, p1 = ok1 ? expf(s1 - mnew) : 0.f;
      float ps = p0 + p1;
      for (int o = 16; o > 0; o >>= 1) ps += __shfl_xor_sync(0xffffffffu, ps, o);
      const float corr = expf(m[i] - mnew);  // first chunk: exp(-3e38 - mnew) = 0 and l, acc are 0 anyway
      l[i] = l[i] * corr + ps;
      acc0[i] *= corr; acc1[i] *= corr;
      m[i] = mnew;
      pr0[i] = p0; pr1[i] = p1;
    }
#pragma unroll 4
    for (int k = 0; k < 32; ++k) {            // two value rows per step, shared by the warp's four queries
      const float va0 = sv[k][lane], va1 = sv[k][lane + 32], vb0 = sv[k + 32][lane], vb1 = sv[k + 32][lane + 32];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float p = __shfl_sync(0xffffffffu, pr0[i], k), pb = __shfl_sync(0xffffffffu, pr1[i], k);
        acc0[i] = fmaf(p, va0, acc0[i]); acc1[i] = fmaf(p, va1, acc1[i]);
        acc0[i] = fmaf(pb, vb0, acc0[i]); acc1[i] = fmaf(pb, vb1, acc1[i]);
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = q0 + warp * 4 + i;
    if (t >= T) continue;
    const float inv = 1.0f / l[i];
    __half h0, l0, h1, l1;
    split_hi_lo(acc0[i] * inv, h0, l0);
    split_hi_lo(acc1[i] * inv, h1, l1);
    __half* o = out + ((long long)b * T + t) * 2 * H + head * EA_HD;
    o[lane] = h0; o[lane + 32] = h1;
    o[H + lane] = l0; o[H + lane + 32] = l1;
  }
}
cudaError_t enc_attention(const float* qkv, const float* cos_t, const float* sin_t, void* out, int B, int T, int nh, int hd, int window, cudaStream_t s) {
  if (hd != EA_HD) return cudaErrorInvalidValue;
  dim3 grid((T + EA_Q - 1) / EA_Q, nh, B);
  k_enc_attention<<<grid, 256, 0, s>>>(qkv, cos_t, sin_t, (__half*)out, T, nh, window);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ attention on mma.sync, fp32 level
// The same attention with the two contractions on the warp-level tensor cores (mma.sync m16n8k16, fp32 accumulate) and every operand an
// fp16 hi + lo pair, three MMAs per product (q_hi k_hi + q_lo k_hi + q_hi k_lo; p_hi v_hi + p_lo v_hi + p_hi v_lo): ~22 bits per
// operand, like the path's GEMMs.  Block = 64 queries of one (item, head), four warps of 16 queries; keys / values are staged 64 at a
// time as four fp16 arrays (K hi / lo with RoPE applied, V hi / lo); a warp walks the 16-key sub-chunks its queries can see with a
// running max / sum per query row (fragment layouts as in attention_mma.cuh).
// Pre-pass: qkv fp32 [M][3H] -> fp16 [M][6H] = [q_hi | q_lo | k_hi | k_lo | v_hi | v_lo] with RoPE applied to q and k, so that the
// attention blocks stage plain 16-byte copies (every key / value row is staged by ~5 query blocks; rotating and splitting it there
// made the staging the larger half of the kernel).
__global__ void __launch_bounds__(256) k_enc_rope_split(const float* __restrict__ qkv, const float* __restrict__ cos_t, const float* __restrict__ sin_t,
                                                         __half* __restrict__ out, long long M, int T, int nh) {
  const int H = nh * EA_HD;
  const long long n = M * nh * 8;                        // (row, head, group of 4 channels d..d+3 with their partners d+32..)
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int d = (int)(i & 7) << 2;
    const long long rh = i >> 3;
    const int head = (int)(rh % nh);
    const long long row = rh / nh;
    const int t = (int)(row % T);
    const float* src = qkv + row * 3 * H + head * EA_HD + d;
    const float4 cs = *reinterpret_cast<const float4*>(cos_t + t * 32 + d), sn = *reinterpret_cast<const float4*>(sin_t + t * 32 + d);
    __half* dst = out + row * 6 * H + head * EA_HD + d;
#pragma unroll
    for (int w = 0; w < 3; ++w) {                        // q, k, v
      const float4 x1 = *reinterpret_cast<const float4*>(src + w * H), x2 = *reinterpret_cast<const float4*>(src + w * H + 32);
      float4 a = x1, bb = x2;
      if (w < 2) {   // rotate_half (:579-584): out[d] = x[d] cos - x[d + 32] sin, out[d + 32] = x[d + 32] cos + x[d] sin
        a = make_float4(x1.x * cs.x - x2.x * sn.x, x1.y * cs.y - x2.y * sn.y, x1.z * cs.z - x2.z * sn.z, x1.w * cs.w - x2.w * sn.w);
        bb = make_float4(x2.x * cs.x + x1.x * sn.x, x2.y * cs.y + x1.y * sn.y, x2.z * cs.z + x1.z * sn.z, x2.w * cs.w + x1.w * sn.w);
      }
      uint2 h, l;
      split4(a, h, l);
      *reinterpret_cast<uint2*>(dst + (2 * w) * H) = h;
      *reinterpret_cast<uint2*>(dst + (2 * w + 1) * H) = l;
      split4(bb, h, l);
      *reinterpret_cast<uint2*>(dst + (2 * w) * H + 32) = h;
      *reinterpret_cast<uint2*>(dst + (2 * w + 1) * H + 32) = l;
    }
  }
}
cudaError_t enc_rope_split(const float* qkv, const float* cos_t, const float* sin_t, void* out, long long M, int T, int nh, cudaStream_t s) {
  k_enc_rope_split<<<enc_blocks(M * nh * 8, 256, 148 * 32), 256, 0, s>>>(qkv, cos_t, sin_t, (__half*)out, M, T, nh);
  return cudaGetLastError();
}

static constexpr int EM_Q = 64, EM_K = 64, EM_THREADS = 128;
static constexpr int EM_SMEM = (2 * EM_Q + 4 * EM_K) * AM_PITCH * 2;   // bytes: Q hi / lo, K hi / lo, V hi / lo

__global__ void __launch_bounds__(EM_THREADS) k_enc_attention_mma(const __half* __restrict__ qkvs, __half* __restrict__ out, int T, int nh, int window) {
  extern __shared__ __align__(16) unsigned char em_smem[];
  __half* sQh = reinterpret_cast<__half*>(em_smem);
  __half* sQl = sQh + EM_Q * AM_PITCH;
  __half* sKh = sQl + EM_Q * AM_PITCH;
  __half* sKl = sKh + EM_K * AM_PITCH;
  __half* sVh = sKl + EM_K * AM_PITCH;
  __half* sVl = sVh + EM_K * AM_PITCH;
  const int q0 = blockIdx.x * EM_Q, head = blockIdx.y, b = blockIdx.z;
  const int H = nh * EA_HD;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, c = lane & 3;
  const __half* base = qkvs + (long long)b * T * 6 * H + head * EA_HD;   // rows of [q_hi | q_lo | k_hi | k_lo | v_hi | v_lo] (k_enc_rope_split)
  // ---- queries: 64 rows x (hi, lo) x 8 chunks of 16 bytes
  for (int i = tid; i < EM_Q * 16; i += EM_THREADS) {
    const int qi = i >> 4, part = (i >> 3) & 1, ch = i & 7, t = q0 + qi;
    uint4 v = make_uint4(0, 0, 0, 0);
    if (t < T) v = *reinterpret_cast<const uint4*>(base + (long long)t * 6 * H + part * H + ch * 8);
    *reinterpret_cast<uint4*>((part ? sQl : sQh) + qi * AM_PITCH + ch * 8) = v;
  }
  __syncthreads();
  const int tq0 = q0 + warp * 16;                       // this warp's first query
  uint32_t qh[4][4], ql[4][4];
  {
    const int off = (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * AM_PITCH + (lane >> 4) * 8;
    const uint32_t ah = smem_u32(sQh + off), al = smem_u32(sQl + off);
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) { ldsm_x4(ah + ks * 32, qh[ks]); ldsm_x4(al + ks * 32, ql[ks]); }
  }
  const int row_lo = tq0 + g, row_hi = tq0 + g + 8;     // the two query rows this lane's accumulators belong to
  float m_lo = -1.0e30f, m_hi = -1.0e30f, l_lo = 0.f, l_hi = 0.f;
  float o[8][4];
#pragma unroll
  for (int n = 0; n < 8; ++n) o[n][0] = o[n][1] = o[n][2] = o[n][3] = 0.f;
  const int qlast = min(q0 + EM_Q - 1, T - 1);
  int kbeg = q0 - window + 1;
  kbeg = kbeg < 0 ? 0 : (kbeg & ~15);
  for (int k0 = kbeg; k0 <= qlast; k0 += EM_K) {
    __syncthreads();                                    // the previous chunk has been consumed by every warp
    for (int i = tid; i < EM_K * 32; i += EM_THREADS) {   // 64 rows x (K hi, K lo, V hi, V lo) x 8 chunks of 16 bytes
      const int ki = i >> 5, arr = (i >> 3) & 3, ch = i & 7, t = k0 + ki;
      uint4 v = make_uint4(0, 0, 0, 0);
      if (t <= qlast) v = *reinterpret_cast<const uint4*>(base + (long long)t * 6 * H + (2 + arr) * H + ch * 8);
      __half* dst = arr == 0 ? sKh : (arr == 1 ? sKl : (arr == 2 ? sVh : sVl));
      *reinterpret_cast<uint4*>(dst + ki * AM_PITCH + ch * 8) = v;
    }
    __syncthreads();
    if (tq0 >= T) continue;                             // a warp without queries only helps with the staging
#pragma unroll 1
    for (int sc = 0; sc < EM_K / 16; ++sc) {
      const int ks0 = k0 + sc * 16;                     // absolute position of the sub-chunk's first key
      if (ks0 > tq0 + 15 || ks0 + 15 < tq0 - window + 1) continue;   // nothing the warp's queries can see (warp-uniform)
      // ---- scores of 16 queries x 16 keys: two 8-key tiles, 12 MMAs each
      float s[2][4];
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        const int koff = (sc * 16 + hf * 8 + (lane & 7)) * AM_PITCH + (lane >> 3) * 8;
        uint32_t kh0[4], kh1[4], kl0[4], kl1[4];
        ldsm_x4(smem_u32(sKh + koff), kh0); ldsm_x4(smem_u32(sKh + koff) + 64, kh1);
        ldsm_x4(smem_u32(sKl + koff), kl0); ldsm_x4(smem_u32(sKl + koff) + 64, kl1);
        float (&d)[4] = s[hf];
        d[0] = d[1] = d[2] = d[3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t bh0 = ks < 2 ? kh0[2 * ks] : kh1[2 * ks - 4], bh1 = ks < 2 ? kh0[2 * ks + 1] : kh1[2 * ks - 3];
          const uint32_t bl0 = ks < 2 ? kl0[2 * ks] : kl1[2 * ks - 4], bl1 = ks < 2 ? kl0[2 * ks + 1] : kl1[2 * ks - 3];
          MmaOp<__half>::mma(d, ql[ks], bh0, bh1);      // small terms first
          MmaOp<__half>::mma(d, qh[ks], bl0, bl1);
          MmaOp<__half>::mma(d, qh[ks], bh0, bh1);
        }
      }
      // ---- mask, running max
      bool ok[2][4];
      float cm_lo = -1.0e30f, cm_hi = -1.0e30f;
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int key = ks0 + hf * 8 + 2 * c + (e & 1), tq = e < 2 ? row_lo : row_hi;
          ok[hf][e] = tq < T && key <= tq && tq - key < window;
          s[hf][e] = ok[hf][e] ? s[hf][e] * 0.125f : -1.0e30f;   // scaling = 1 / sqrt(head_dim) (:665)
          if (e < 2) cm_lo = fmaxf(cm_lo, s[hf][e]); else cm_hi = fmaxf(cm_hi, s[hf][e]);
        }
      }
      cm_lo = fmaxf(cm_lo, __shfl_xor_sync(0xffffffffu, cm_lo, 1)); cm_lo = fmaxf(cm_lo, __shfl_xor_sync(0xffffffffu, cm_lo, 2));
      cm_hi = fmaxf(cm_hi, __shfl_xor_sync(0xffffffffu, cm_hi, 1)); cm_hi = fmaxf(cm_hi, __shfl_xor_sync(0xffffffffu, cm_hi, 2));
      const float mn_lo = fmaxf(m_lo, cm_lo), mn_hi = fmaxf(m_hi, cm_hi);
      const float cr_lo = expf(m_lo - mn_lo), cr_hi = expf(m_hi - mn_hi);   // 1 while nothing has been seen yet (both -1e30)
      m_lo = mn_lo; m_hi = mn_hi;
      l_lo *= cr_lo; l_hi *= cr_hi;
#pragma unroll
      for (int n = 0; n < 8; ++n) { o[n][0] *= cr_lo; o[n][1] *= cr_lo; o[n][2] *= cr_hi; o[n][3] *= cr_hi; }
      // ---- probabilities as fp16 pairs (A fragments of P . V)
      uint32_t ph[4], pl[4];
#pragma unroll
      for (int hf = 0; hf < 2; ++hf) {
        float p[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) p[e] = ok[hf][e] ? expf(s[hf][e] - (e < 2 ? mn_lo : mn_hi)) : 0.f;
        l_lo += p[0] + p[1];
        l_hi += p[2] + p[3];
        const __half2 h01 = __floats2half2_rn(p[0], p[1]), h23 = __floats2half2_rn(p[2], p[3]);
        const float2 b01 = __half22float2(h01), b23 = __half22float2(h23);
        const __half2 l01 = __floats2half2_rn(p[0] - b01.x, p[1] - b01.y), l23 = __floats2half2_rn(p[2] - b23.x, p[3] - b23.y);
        ph[2 * hf] = *reinterpret_cast<const uint32_t*>(&h01); ph[2 * hf + 1] = *reinterpret_cast<const uint32_t*>(&h23);
        pl[2 * hf] = *reinterpret_cast<const uint32_t*>(&l01); pl[2 * hf + 1] = *reinterpret_cast<const uint32_t*>(&l23);
      }
      // ---- P . V: 8 output tiles of 8 channels, three MMAs each
      const int voff = (sc * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * AM_PITCH + (lane >> 4) * 8;
#pragma unroll
      for (int n2 = 0; n2 < 4; ++n2) {
        uint32_t vh[4], vl[4];
        ldsm_x4_t(smem_u32(sVh + voff) + n2 * 32, vh);
        ldsm_x4_t(smem_u32(sVl + voff) + n2 * 32, vl);
        MmaOp<__half>::mma(o[2 * n2], pl, vh[0], vh[1]);
        MmaOp<__half>::mma(o[2 * n2], ph, vl[0], vl[1]);
        MmaOp<__half>::mma(o[2 * n2], ph, vh[0], vh[1]);
        MmaOp<__half>::mma(o[2 * n2 + 1], pl, vh[2], vh[3]);
        MmaOp<__half>::mma(o[2 * n2 + 1], ph, vl[2], vl[3]);
        MmaOp<__half>::mma(o[2 * n2 + 1], ph, vh[2], vh[3]);
      }
    }
  }
  if (tq0 >= T) return;
  l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 1); l_lo += __shfl_xor_sync(0xffffffffu, l_lo, 2);
  l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 1); l_hi += __shfl_xor_sync(0xffffffffu, l_hi, 2);
  const float inv_lo = 1.0f / l_lo, inv_hi = 1.0f / l_hi;
#pragma unroll
  for (int n = 0; n < 8; ++n) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int t = r ? row_hi : row_lo;
      if (t >= T) continue;
      const float v0 = o[n][2 * r] * (r ? inv_hi : inv_lo), v1 = o[n][2 * r + 1] * (r ? inv_hi : inv_lo);
      const __half2 h = __floats2half2_rn(v0, v1);
      const float2 bk = __half22float2(h);
      const __half2 l = __floats2half2_rn(v0 - bk.x, v1 - bk.y);
      __half* op = out + ((long long)b * T + t) * 2 * H + head * EA_HD + n * 8 + 2 * c;
      *reinterpret_cast<__half2*>(op) = h;
      *reinterpret_cast<__half2*>(op + H) = l;
    }
  }
}
cudaError_t enc_attention_setup() {
  return cudaFuncSetAttribute(k_enc_attention_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, EM_SMEM);
}
cudaError_t enc_attention_mma(const void* qkvs, void* out, int B, int T, int nh, int hd, int window, cudaStream_t s) {
  if (hd != EA_HD) return cudaErrorInvalidValue;
  dim3 grid((T + EM_Q - 1) / EM_Q, nh, B);
  k_enc_attention_mma<<<grid, EM_THREADS, EM_SMEM, s>>>((const __half*)qkvs, (__half*)out, T, nh, window);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ RVQ search, step 2 of a layer
// The layer's scores come from a tensor-core GEMM (encode_plan.cu): S[m][c] = r_m . e_c, fp32 [M][NC].  One warp per frame m:
// code = argmin_c (|e_c|^2 - 2 S[m][c]) with the LOWEST index winning ties (torch.argmin over torch.cdist, modeling_mimi.py:1205-1219;
// the |r|^2 term is the same for every code), r_m -= e_code in fp32, and the next layer's GEMM operand [hi | lo] of the new residual.
__global__ void __launch_bounds__(256) k_enc_argmin_update(const float* __restrict__ S, const float* __restrict__ n2, const float* __restrict__ table,
                                                            float* __restrict__ r, int ldr, __half* __restrict__ a_next, long long* __restrict__ codes,
                                                            long long code_stride_m, long long code_stride_b, long long code_off, int M, int F, int NC,
                                                            int dq) {
  const int lane = threadIdx.x & 31;
  const long long m = blockIdx.x * (long long)(blockDim.x >> 5) + (threadIdx.x >> 5);
  if (m >= M) return;
  const float* s = S + m * NC;
  float best = 3.4e38f;
  int bidx = 0x7fffffff;
  for (int c = lane * 4; c < NC; c += 128) {   // indices ascend per lane: strict < keeps the lowest
    const float4 sv = __ldcs(reinterpret_cast<const float4*>(s + c));
    const float4 nv = *reinterpret_cast<const float4*>(n2 + c);
    const float v0 = fmaf(-2.f, sv.x, nv.x), v1 = fmaf(-2.f, sv.y, nv.y), v2 = fmaf(-2.f, sv.z, nv.z), v3 = fmaf(-2.f, sv.w, nv.w);
    if (v0 < best) { best = v0; bidx = c; }
    if (v1 < best) { best = v1; bidx = c + 1; }
    if (v2 < best) { best = v2; bidx = c + 2; }
    if (v3 < best) { best = v3; bidx = c + 3; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float b2 = __shfl_xor_sync(0xffffffffu, best, o);
    const int i2 = __shfl_xor_sync(0xffffffffu, bidx, o);
    if (b2 < best || (b2 == best && i2 < bidx)) { best = b2; bidx = i2; }
  }
  if (lane == 0) codes[(m / F) * code_stride_b + (m % F) * code_stride_m + code_off] = bidx;
  float* rr = r + m * ldr;
  const float* e = table + (long long)bidx * dq;
  for (int j = lane * 4; j < dq; j += 128) {
    float4 rv = *reinterpret_cast<const float4*>(rr + j);
    const float4 ev = *reinterpret_cast<const float4*>(e + j);
    rv.x -= ev.x; rv.y -= ev.y; rv.z -= ev.z; rv.w -= ev.w;
    *reinterpret_cast<float4*>(rr + j) = rv;
    if (a_next) {
      __half h[4], l[4];
      split_hi_lo(rv.x, h[0], l[0]); split_hi_lo(rv.y, h[1], l[1]); split_hi_lo(rv.z, h[2], l[2]); split_hi_lo(rv.w, h[3], l[3]);
      *reinterpret_cast<uint2*>(a_next + m * 2 * dq + j) = *reinterpret_cast<const uint2*>(h);
      *reinterpret_cast<uint2*>(a_next + m * 2 * dq + dq + j) = *reinterpret_cast<const uint2*>(l);
    }
  }
}
cudaError_t enc_argmin_update(const float* S, const float* n2, const float* table, float* r, int ldr, void* a_next, long long* codes,
                              long long code_stride_m, long long code_stride_b, long long code_off, int M, int F, int NC, int dq, cudaStream_t s) {
  if (NC % 128 || dq % 4 || ldr % 4) return cudaErrorInvalidValue;
  k_enc_argmin_update<<<(unsigned)((M + 7) / 8), 256, 0, s>>>(S, n2, table, r, ldr, (__half*)a_next, codes, code_stride_m, code_stride_b, code_off, M, F, NC, dq);
  return cudaGetLastError();
}
// fp32 [M][ld] columns [c0, c0 + C) -> fp16 [M][2C] rows [hi | lo] (first operand of a quantiser group: its projected latents)
__global__ void k_enc_split_cols(const float* __restrict__ in, int ld, int c0, __half* __restrict__ out, long long M, int C) {
  const int c4n = C >> 2;
  const long long n = M * c4n;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long m = i / c4n;
    const int c = (int)(i - m * c4n) << 2;
    const float4 v = *reinterpret_cast<const float4*>(in + m * ld + c0 + c);
    __half h[4], l[4];
    split_hi_lo(v.x, h[0], l[0]); split_hi_lo(v.y, h[1], l[1]); split_hi_lo(v.z, h[2], l[2]); split_hi_lo(v.w, h[3], l[3]);
    *reinterpret_cast<uint2*>(out + m * 2 * C + c) = *reinterpret_cast<const uint2*>(h);
    *reinterpret_cast<uint2*>(out + m * 2 * C + C + c) = *reinterpret_cast<const uint2*>(l);
  }
}
cudaError_t enc_split_cols(const float* in, int ld, int c0, void* out, long long M, int C, cudaStream_t s) {
  k_enc_split_cols<<<enc_blocks(M * (C / 4), 256), 256, 0, s>>>(in, ld, c0, (__half*)out, M, C);
  return cudaGetLastError();
}

// ------------------------------------------------------------------------------------------------ debug / parity tap
// fp32 [B][L][C] -> fp32 [B][C][L] (the reference's channels-first layout), for the latents the parity tests compare
__global__ void k_enc_to_cf(const float* __restrict__ in, float* __restrict__ out, int B, int L, int C) {
  const long long n = (long long)B * L * C;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const int t = (int)(i % L);
    const long long bc = i / L;
    const int c = (int)(bc % C);
    const long long b = bc / C;
    out[i] = in[(b * L + t) * C + c];
  }
}
cudaError_t enc_to_channels_first(const float* in, float* out, int B, int L, int C, cudaStream_t s) {
  k_enc_to_cf<<<enc_blocks((long long)B * L * C, 256), 256, 0, s>>>(in, out, B, L, C);
  return cudaGetLastError();
}

}  // namespace q3
