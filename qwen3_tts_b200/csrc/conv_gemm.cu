// Implicit-GEMM causal conv / transposed conv / linear on tcgen05 tensor cores (sm_100a).
//
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread
// tcgen05.mma issuer, warps 2..5 = epilogue (TMEM -> registers -> fused epilogue -> global).
// One CTA per SM; accumulators are double-buffered in TMEM (2 x 256 fp32 columns) so the
// epilogue of tile i overlaps the MMAs of tile i+1.  Operand tiles are K-major, written by TMA
// with 128B (KB=64) or 64B (KB=32, the Cin=96 layers) swizzle; conv taps are time-shifted TMA
// boxes of the same channels-last activation, with out-of-range rows zero-filled by TMA.
#include "conv_gemm.h"
#include "epilogue.cuh"

namespace q3 {

static constexpr int BM = 128;
static constexpr int kGemmThreads = 192;
static constexpr int kMaxStages = 8;

struct __align__(8) GemmBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint32_t tmem_base;
  uint32_t pad;
};

template <typename T>
__global__ void __launch_bounds__(kGemmThreads, 1)
conv_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const GemmParams p, const int nstages) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int KB = p.kb;
  const uint32_t a_bytes = BM * KB * 2;
  const uint32_t b_bytes = p.bn * KB * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  GemmBarriers* bars = reinterpret_cast<GemmBarriers*>(smem + (size_t)nstages * stage_bytes);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  const int m_tiles = (p.rows_valid + BM - 1) / BM;
  const int n_tiles = p.n_total / p.bn;
  const int total_tiles = p.batches * m_tiles * n_tiles;
  const int kblocks = p.cin / KB;
  const int ksteps = kblocks * p.ntaps;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    for (int s = 0; s < nstages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->tmem_full[s], 1); mbar_init(&bars->tmem_empty[s], 4); }
    fence_barrier_init();
  }
  if (warp == 1) {
    tmem_alloc(&bars->tmem_base, 512);
    tmem_relinquish();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int nt = tile % n_tiles;
        const int mt = (tile / n_tiles) % m_tiles;
        const int b = tile / (n_tiles * m_tiles);
        const int m0 = mt * BM, n0 = nt * p.bn;
        for (int kb = 0; kb < kblocks; ++kb) {
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait(&bars->empty[s], ph ^ 1);
            uint8_t* sa = smem + (size_t)s * stage_bytes;
            mbar_expect_tx(&bars->full[s], stage_bytes);
            tma_load_3d(sa, &tmA, &bars->full[s], kb * KB, m0 + p.tap_off[tap], b);
            tma_load_3d(sa + a_bytes, &tmB, &bars->full[s], kb * KB, n0, tap);
            if (++s == nstages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread)
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(BM, p.bn, p.bf16 != 0);
      const uint32_t span = KB * 2;
      int s = 0; uint32_t ph = 0; int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        const uint32_t tph = (it >> 1) & 1;
        mbar_wait(&bars->tmem_empty[buf], tph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        for (int ks = 0; ks < ksteps; ++ks) {
          mbar_wait(&bars->full[s], ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + (size_t)s * stage_bytes);
          const uint64_t ad = umma_smem_desc(sa, span);
          const uint64_t bd = umma_smem_desc(sa + a_bytes, span);
          for (int k = 0; k < KB / 16; ++k) {
            // advance 16 elements (32 bytes) along K inside the swizzle span: +2 in the >>4 address field
            umma_f16(d_tmem, ad + 2 * k, bd + 2 * k, idesc, (ks | k) != 0);
          }
          umma_commit(&bars->empty[s]);
          if (++s == nstages) { s = 0; ph ^= 1; }
        }
        umma_commit(&bars->tmem_full[buf]);
      }
    }
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int quad = warp & 3;  // TMEM lane quadrant this warp may access
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      const int nt = tile % n_tiles;
      const int mt = (tile / n_tiles) % m_tiles;
      const int b = tile / (n_tiles * m_tiles);
      const int buf = it & 1;
      const uint32_t tph = (it >> 1) & 1;
      const int row = mt * BM + quad * 32 + lane;
      const bool row_ok = row < p.rows_valid;
      mbar_wait(&bars->tmem_full[buf], tph);
      tc_fence_after();
      const uint32_t t0 = tmem_base + ((uint32_t)(quad * 32) << 16) + buf * 256;
      for (int c = 0; c < p.bn; c += 16) {
        uint32_t r[16];
        tmem_ld16(t0 + c, r);
        tmem_ld_wait();
        if (row_ok) {
          float v[16];
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
          epilogue16<T>(p, v, b, row, nt * p.bn + c);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->tmem_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// SIMT cross-check: same math, same epilogue, no tensor cores / TMA.  Debug engine only
// (Q3TTS_ENGINE=simt); one thread per (row, 16-column group).
template <typename T>
__global__ void conv_gemm_simt_kernel(const GemmParams p) {
  const int ngroups = p.n_total / 16;
  const long long total = (long long)p.batches * p.rows_valid * ngroups;
  for (long long idx = blockIdx.x * (long long)blockDim.x + threadIdx.x; idx < total; idx += (long long)gridDim.x * blockDim.x) {
    const int g = (int)(idx % ngroups);
    const int row = (int)((idx / ngroups) % p.rows_valid);
    const int b = (int)(idx / ((long long)ngroups * p.rows_valid));
    const int n = g * 16;
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    const T* A = reinterpret_cast<const T*>(p.a_ptr);
    const T* W = reinterpret_cast<const T*>(p.w_ptr);
    for (int tap = 0; tap < p.ntaps; ++tap) {
      const int r = row + p.tap_off[tap];
      if (r < 0 || r >= p.rows) continue;
      const T* a = A + ((long long)b * p.rows + r) * p.cin;
      const T* w = W + ((long long)tap * p.n_total + n) * p.cin;
      for (int k = 0; k < p.cin; ++k) {
        const float av = Elem<T>::to_f(a[k]);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(av, Elem<T>::to_f(w[(long long)i * p.cin + k]), acc[i]);
      }
    }
    epilogue16<T>(p, acc, b, row, n);
  }
}

// ------------------------------------------------------------------------------------------- host
static PFN_encodeTiled g_encode = nullptr;

int tmap_init() {
  if (g_encode) return 0;
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &qres);
  if (e != cudaSuccess || qres != cudaDriverEntryPointSuccess || !fn) return -1;
  g_encode = reinterpret_cast<PFN_encodeTiled>(fn);
  return 0;
}

// 3-D tensor map over a 16-bit tensor [d2][d1][d0] (d0 contiguous), box [1][box1][box0].
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box0, uint32_t box1, bool bf16) {
  if (tmap_init()) return -1;
  cuuint64_t dims[3] = {d0, d1, d2};
  cuuint64_t strides[2] = {d0 * 2, d0 * d1 * 2};
  cuuint32_t box[3] = {box0, box1, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUtensorMapSwizzle sw = box0 * 2 == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (box0 * 2 == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUresult r = g_encode(out, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), dims,
                        strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

static int g_num_sms = 0;
static int g_smem_optin = 0;

int gemm_stage_count(const GemmParams& p) {
  const int stage = (BM + p.bn) * p.kb * 2;
  int n = (g_smem_optin - 1024 - (int)sizeof(GemmBarriers) - 1024) / stage;
  if (n > kMaxStages) n = kMaxStages;
  return n;
}

int gemm_setup(int device) {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1;
  g_num_sms = prop.multiProcessorCount;
  g_smem_optin = (int)prop.sharedMemPerBlockOptin;
  cudaError_t e1 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e2 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e1 != cudaSuccess || e2 != cudaSuccess) return -2;
  return tmap_init();
}

int gemm_num_sms() { return g_num_sms; }

cudaError_t launch_conv_gemm(const GemmOp& op, int engine, cudaStream_t stream) {
  const GemmParams& p = op.p;
  if (engine == ENGINE_SIMT) {
    const long long total = (long long)p.batches * p.rows_valid * (p.n_total / 16);
    int blocks = (int)((total + 127) / 128);
    if (blocks > g_num_sms * 16) blocks = g_num_sms * 16;
    if (blocks < 1) blocks = 1;
    if (p.bf16) conv_gemm_simt_kernel<__nv_bfloat16><<<blocks, 128, 0, stream>>>(p);
    else conv_gemm_simt_kernel<__half><<<blocks, 128, 0, stream>>>(p);
    return cudaGetLastError();
  }
  const int m_tiles = (p.rows_valid + BM - 1) / BM;
  const int total_tiles = p.batches * m_tiles * (p.n_total / p.bn);
  if (total_tiles <= 0) return cudaSuccess;
  const int grid = total_tiles < g_num_sms ? total_tiles : g_num_sms;
  const int nstages = gemm_stage_count(p);
  if (nstages < 2) return cudaErrorInvalidConfiguration;
  const size_t smem = (size_t)nstages * (BM + p.bn) * p.kb * 2 + sizeof(GemmBarriers) + 1024;
  if (p.bf16) conv_gemm_tc_kernel<__nv_bfloat16><<<grid, kGemmThreads, smem, stream>>>(op.tmA, op.tmB, p, nstages);
  else conv_gemm_tc_kernel<__half><<<grid, kGemmThreads, smem, stream>>>(op.tmA, op.tmB, p, nstages);
  return cudaGetLastError();
}

}  // namespace q3
