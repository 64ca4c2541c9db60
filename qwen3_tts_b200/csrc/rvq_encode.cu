// Encode side, first kernel: residual-vector-quantiser nearest-neighbour search.
//
// Reference: Qwen3TTSTokenizerV2Model.encode (modeling :960-990) calls MimiModel.encode of the `transformers` package (pinned
// 4.57.3, not vendored), whose last stage is MimiSplitResidualVectorQuantizer.encode (transformers/models/mimi/modeling_mimi.py):
//   semantic RVQ (1 codebook) and acoustic RVQ (the next K-1 codebooks) each take the SAME latents x [B, 512, T]:
//     r = input_proj(x)                       Conv1d(512 -> 256, k = 1, no bias)
//     for each layer:  code = argmin_c || r - e_c ||   (torch.cdist + argmin; e = embed_sum / clamp(cluster_usage, 1e-5))
//                      r = r - e_code
// Only the first encoder_valid_num_quantizers = 16 codebooks are kept by the reference (modeling :982), so only those are searched.
//
// Kernel: one block = FR frames; residuals live in shared memory; thread t scores codes t, t + 256, ... against all FR frames at once
// (each table row is read once per block, not once per frame), score = ||e||^2 - 2 r.e in fp32 (the ||r||^2 term is the same for
// every code), block-wide argmin with the LOWEST index winning ties (torch.argmin returns the first minimum).  Integer output:
// the bar is identical indices; frames whose two best scores are closer than fp32 rounding are the only admissible differences
// (tests/test_gpu_rvq_encode.py counts them against the float64 oracle).
#include <string>

#include "engine.h"

namespace q3 {

static constexpr int ENC_FR = 8;      // frames per block
static constexpr int ENC_THREADS = 256;

// tables [K][ncodes][dq] fp32, norms [K][ncodes] = ||e||^2, win [2][dq][D] (semantic, acoustic input_proj)
__global__ void __launch_bounds__(ENC_THREADS)
k_rvq_encode(const float* __restrict__ x, const float* __restrict__ tables, const float* __restrict__ norms, const float* __restrict__ win,
             long long* __restrict__ codes, int B, int T, int D, int dq, int ncodes, int K, int n_sem, int x_cl, int codes_tq) {
  extern __shared__ float sm[];
  float* r = sm;                         // [ENC_FR][dq] residuals
  float* xs = r + ENC_FR * dq;           // [ENC_FR][D] latents of the block's frames
  __shared__ float s_best[ENC_FR][ENC_THREADS / 32];
  __shared__ int s_idx[ENC_FR][ENC_THREADS / 32];
  __shared__ int s_code[ENC_FR];
  const long long f0 = (long long)blockIdx.x * ENC_FR;   // first flat frame (b * T + t)
  const long long nfr = (long long)B * T;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  for (int i = tid; i < ENC_FR * D; i += ENC_THREADS) {
    const int f = i / D, c = i - f * D;
    const long long g = f0 + f;
    float v = 0.f;
    if (g < nfr) { const long long b = g / T, t = g - b * T; v = x_cl ? x[g * D + c] : x[(b * D + c) * T + t]; }
    xs[f * D + c] = v;
  }
  __syncthreads();
  for (int k = 0; k < K; ++k) {
    if (k == 0 || k == n_sem) {
      // a new quantiser group starts from its own projection of the latents (semantic group, then acoustic group)
      const float* w = win + (size_t)(k == 0 ? 0 : 1) * dq * D;
      for (int i = tid; i < ENC_FR * dq; i += ENC_THREADS) {
        const int f = i / dq, j = i - f * dq;
        const float* wr = w + (size_t)j * D;
        float acc = 0.f;
        for (int c = 0; c < D; ++c) acc = fmaf(wr[c], xs[f * D + c], acc);
        r[f * dq + j] = acc;
      }
      __syncthreads();
    }
    const float* tab = tables + (size_t)k * ncodes * dq;
    const float* nrm = norms + (size_t)k * ncodes;
    float best[ENC_FR];
    int bidx[ENC_FR];
#pragma unroll
    for (int f = 0; f < ENC_FR; ++f) { best[f] = 3.4e38f; bidx[f] = 0x7fffffff; }
    for (int c = tid; c < ncodes; c += ENC_THREADS) {
      const float4* e4 = reinterpret_cast<const float4*>(tab + (size_t)c * dq);
      float dot[ENC_FR];
#pragma unroll
      for (int f = 0; f < ENC_FR; ++f) dot[f] = 0.f;
      for (int j = 0; j < dq / 4; ++j) {
        const float4 e = __ldg(e4 + j);
#pragma unroll
        for (int f = 0; f < ENC_FR; ++f) {
          const float4 rv = *reinterpret_cast<const float4*>(r + f * dq + 4 * j);
          dot[f] = fmaf(e.x, rv.x, dot[f]); dot[f] = fmaf(e.y, rv.y, dot[f]);
          dot[f] = fmaf(e.z, rv.z, dot[f]); dot[f] = fmaf(e.w, rv.w, dot[f]);
        }
      }
      const float n2 = nrm[c];
#pragma unroll
      for (int f = 0; f < ENC_FR; ++f) {
        const float s = fmaf(-2.f, dot[f], n2);
        if (s < best[f]) { best[f] = s; bidx[f] = c; }   // codes ascend per thread: strict < keeps the lowest index
      }
    }
    // block argmin per frame: warp shuffle, then across the 8 warps; ties -> lowest index
#pragma unroll
    for (int f = 0; f < ENC_FR; ++f) {
      float s = best[f]; int ix = bidx[f];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
        const int i2 = __shfl_xor_sync(0xffffffffu, ix, o);
        if (s2 < s || (s2 == s && i2 < ix)) { s = s2; ix = i2; }
      }
      if (lane == 0) { s_best[f][warp] = s; s_idx[f][warp] = ix; }
    }
    __syncthreads();
    if (tid < ENC_FR) {
      float s = s_best[tid][0]; int ix = s_idx[tid][0];
      for (int w = 1; w < ENC_THREADS / 32; ++w)
        if (s_best[tid][w] < s || (s_best[tid][w] == s && s_idx[tid][w] < ix)) { s = s_best[tid][w]; ix = s_idx[tid][w]; }
      s_code[tid] = ix;
      const long long g = f0 + tid;
      if (g < nfr) { const long long b = g / T, t = g - b * T; codes[codes_tq ? g * K + k : (b * K + k) * T + t] = ix; }
    }
    __syncthreads();
    for (int i = tid; i < ENC_FR * dq; i += ENC_THREADS) {   // residual = residual - quantized
      const int f = i / dq, j = i - f * dq;
      r[f * dq + j] -= tab[(size_t)s_code[f] * dq + j];
    }
    __syncthreads();
  }
}

// embed = embed_sum / clamp(cluster_usage, 1e-5) and its squared norm; one block per code
__global__ void k_enc_table(const float* __restrict__ es, const float* __restrict__ usage, float* __restrict__ tab, float* __restrict__ nrm, int dq) {
  __shared__ float red[32];
  const int c = blockIdx.x;
  const float u = fmaxf(usage[c], 1e-5f);
  float acc = 0.f;
  for (int j = threadIdx.x; j < dq; j += blockDim.x) {
    const float v = es[(size_t)c * dq + j] / u;
    tab[(size_t)c * dq + j] = v;
    acc = fmaf(v, v, acc);
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) s += red[w];
    nrm[c] = s;
  }
}

// Packs the encoder quantiser from the raw state_dict tensors (reference keys under `encoder.quantizer.`); called by engine_finalize
// when those keys were loaded.  Returns 0 when there is nothing to pack.
int encode_finalize(Engine* e) {
  const q3tts_config& c = e->cfg;
  const int K = c.num_quantizers, dq = c.codebook_dim / 2, D = c.codebook_dim, nc = c.codebook_size;
  const std::string qp = "encoder.quantizer.";
  const std::string grp[2] = {qp + "semantic_residual_vector_quantizer.", qp + "acoustic_residual_vector_quantizer."};
  if (e->raw.find(grp[0] + "input_proj.weight") == e->raw.end()) return Q3TTS_OK;
  if (e->enc_tables) { cudaFree(e->enc_tables); cudaFree(e->enc_norms); cudaFree(e->enc_win); e->enc_tables = e->enc_norms = e->enc_win = nullptr; }
  if (cudaMalloc((void**)&e->enc_tables, (size_t)K * nc * dq * 4) != cudaSuccess || cudaMalloc((void**)&e->enc_norms, (size_t)K * nc * 4) != cudaSuccess ||
      cudaMalloc((void**)&e->enc_win, (size_t)2 * dq * D * 4) != cudaSuccess) {
    cudaGetLastError();
    return e->fail(Q3TTS_OOM, "encoder quantiser allocation failed");
  }
  for (int g = 0; g < 2; ++g) {
    auto it = e->raw.find(grp[g] + "input_proj.weight");
    if (it == e->raw.end() || it->second.numel != (long long)dq * D) return e->fail(Q3TTS_BAD_KEY, "missing or mis-shaped " + grp[g] + "input_proj.weight");
    cudaMemcpy(e->enc_win + (size_t)g * dq * D, it->second.d, (size_t)dq * D * 4, cudaMemcpyDeviceToDevice);
  }
  const int n_sem = 1;
  for (int k = 0; k < K; ++k) {
    const std::string p = grp[k < n_sem ? 0 : 1] + "layers." + std::to_string(k < n_sem ? k : k - n_sem) + ".codebook.";
    auto es = e->raw.find(p + "embed_sum"), cu = e->raw.find(p + "cluster_usage");
    if (es == e->raw.end() || cu == e->raw.end() || es->second.numel != (long long)nc * dq || cu->second.numel != nc)
      return e->fail(Q3TTS_BAD_KEY, "missing or mis-shaped " + p + "{embed_sum, cluster_usage}");
    k_enc_table<<<nc, 64>>>(es->second.d, cu->second.d, e->enc_tables + (size_t)k * nc * dq, e->enc_norms + (size_t)k * nc, dq);
  }
  if (cudaDeviceSynchronize() != cudaSuccess) return e->fail(Q3TTS_CUDA_ERROR, "encoder quantiser packing failed");
  e->enc_ready = true;
  return Q3TTS_OK;
}

cudaError_t rvq_encode_launch(Engine* e, const float* latents, int B, int T, long long* codes, int x_cl, int codes_tq, cudaStream_t s) {
  const q3tts_config& c = e->cfg;
  const int dq = c.codebook_dim / 2;
  const long long nfr = (long long)B * T;
  const int blocks = (int)((nfr + ENC_FR - 1) / ENC_FR);
  const size_t smem = (size_t)ENC_FR * (dq + c.codebook_dim) * sizeof(float);
  k_rvq_encode<<<blocks, ENC_THREADS, smem, s>>>(latents, e->enc_tables, e->enc_norms, e->enc_win, codes, B, T, c.codebook_dim, dq, c.codebook_size,
                                                 c.num_quantizers, 1, x_cl, codes_tq);
  return cudaGetLastError();
}

}  // namespace q3

extern "C" int q3tts_rvq_encode(q3tts_handle* h, const float* latents, int B, int T, int64_t* codes, void* stream) {
  if (!h) return Q3TTS_BAD_CONFIG;
  q3::Engine& e = h->e;
  if (!e.finalized || !e.enc_ready) return e.fail(Q3TTS_NOT_READY, "rvq_encode: the encoder quantiser weights (encoder.quantizer.*) are not loaded");
  if (!latents || !codes || B < 1 || T < 1) return e.fail(Q3TTS_BAD_SHAPE, "rvq_encode: latents must be [B>=1, codebook_dim, T>=1]");
  const q3tts_config& c = e.cfg;
  const int dq = c.codebook_dim / 2;
  if (dq % 4) return e.fail(Q3TTS_BAD_CONFIG, "rvq_encode: codebook_dim / 2 must be a multiple of 4");
  q3::DeviceGuard g(e.device);
  const cudaError_t err = q3::rvq_encode_launch(&e, latents, B, T, (long long*)codes, 0, 0, (cudaStream_t)stream);
  ++e.launches;
  if (err != cudaSuccess) return e.fail(Q3TTS_CUDA_ERROR, std::string("rvq_encode: ") + cudaGetErrorString(err));
  return Q3TTS_OK;
}
