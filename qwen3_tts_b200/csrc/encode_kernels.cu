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
