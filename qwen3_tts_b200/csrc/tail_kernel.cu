// Output conv of the vocoder: decoder[6] = CausalConv(C -> 1, k = 7) + clamp(-1, 1)   (reference modeling :862, :883).
//
// As an implicit GEMM with N = 1 (padded to the smallest UMMA N, 16) the conv costs 7 taps x C/16 MMAs of ~56 issue cycles per
// 128 rows: MMA-issue bound at 3.6x the HBM time of reading the [rows][C] input once.  Here the taps ARE the N dimension:
//     P[t', j] = sum_c a[t', c] * w[j, c]           one M=128, N=16 (7 used), K=C product per tile: C/16 MMAs instead of 7*C/16
//     out[t]   = clamp(bias + sum_j P[t - 6 + j, j])  a 7-term shifted sum, done by the epilogue warps through shared memory
// A tile reads 128 input rows [t0 - 6, t0 + 122) and emits the 122 outputs [t0, t0 + 122); rows before the start of a batch
// item are zero-filled by TMA (the causal padding) or are the stream's history rows (a_row0 = 6).
// P[t', j] depends only on row t' and the sum runs j = 0..6 in order, so the result does not depend on where tiles start:
// streaming steps and bulk decodes stay bit-identical.
//
// Roles: warp 0 TMA producer (ring of input tiles), warp 1 MMA issuer (ring of 16-column accumulators in tensor memory),
// warps 2..5 epilogue.  HBM-bound by construction: 24.6 KB in, 0.5 KB out per tile.
#include "conv_gemm.h"
#include "launch.cuh"
#include "ptx.cuh"

namespace q3 {

static constexpr int T_ROWS = 128;          // input rows per tile
static constexpr int T_OUT = T_ROWS - 6;    // outputs per tile
static constexpr int T_NACC = 8;            // accumulator ring (16 columns each)
static constexpr int T_MAXSTAGES = 8;
static constexpr int T_PPITCH = T_ROWS + 4;

struct __align__(8) TailBarriers {
  uint64_t full[T_MAXSTAGES], empty[T_MAXSTAGES];
  uint64_t acc_full[T_NACC], acc_empty[T_NACC];
  uint64_t w_full;
  uint32_t tmem_base;
  uint32_t pad;
};

template <typename T>
__global__ void __launch_bounds__(192, 1)
tail_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW, const TailParams p, const int nstages) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset on the shared array: keeps LDS / STS (a uintptr_t round trip makes every access generic)
  const int KS = p.C / 32;                     // K blocks of 32 channels (64-byte rows, SWIZZLE_64B)
  const uint32_t a_sub = T_ROWS * 64, stage_bytes = (uint32_t)KS * a_sub;
  uint8_t* sW = smem;                          // [KS][16 rows][64 B], 1 KB per K block
  uint8_t* sA = smem + 8 * 1024;               // [nstages][KS][128 rows][64 B]
  float* sP = reinterpret_cast<float*>(sA + (size_t)nstages * stage_bytes);  // [2][7][T_PPITCH]
  TailBarriers* bars = reinterpret_cast<TailBarriers*>(sP + 2 * 7 * T_PPITCH);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tiles_per_item = (p.rows_valid + T_OUT - 1) / T_OUT;
  const int total_tiles = p.batches * tiles_per_item;
  const FastDiv fdt((uint32_t)tiles_per_item);
  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW);
    for (int s = 0; s < nstages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    for (int a = 0; a < T_NACC; ++a) { mbar_init(&bars->acc_full[a], 1); mbar_init(&bars->acc_empty[a], 4); }
    mbar_init(&bars->w_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(&bars->tmem_base, 128); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  pdl_trigger();  // programmatic dependent launch (launch.cuh): after the TMEM allocation
  pdl_wait();
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(&bars->w_full, (uint32_t)KS * 1024);
      for (int kk = 0; kk < KS; ++kk) tma_load_3d(sW + kk * 1024, &tmW, &bars->w_full, kk * 32, 0, 0);
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        int b, t0; fdt.divmod(tile, b, t0); t0 *= T_OUT;
        mbar_wait_backoff(&bars->empty[s], ph ^ 1);
        mbar_expect_tx(&bars->full[s], stage_bytes);
        for (int kk = 0; kk < KS; ++kk)
          tma_load_3d(sA + (size_t)s * stage_bytes + (size_t)kk * a_sub, &tmA, &bars->full[s], kk * 32, p.a_row0 + t0 - 6, b);
        if (++s == nstages) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(T_ROWS, 16, p.bf16 != 0);
      const uint32_t hi = umma_desc_hi(64);
      const uint32_t w0 = umma_desc_lo(smem_u32(sW)), a0 = umma_desc_lo(smem_u32(sA));
      mbar_wait(&bars->w_full, 0);
      int s = 0; uint32_t ph = 0; int it = 0;
      for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
        const int a = it & (T_NACC - 1);
        mbar_wait_backoff(&bars->acc_empty[a], ((uint32_t)(it / T_NACC) & 1u) ^ 1u);
        mbar_wait_backoff(&bars->full[s], ph);
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)(a * 16);
        const uint32_t as = a0 + (uint32_t)s * (stage_bytes >> 4);
        for (int kk = 0; kk < KS; ++kk) {
          umma_f16_lohi(d, as + (uint32_t)kk * (a_sub >> 4), w0 + (uint32_t)kk * (1024 >> 4), hi, idesc, kk != 0);
          umma_f16_lohi(d, as + (uint32_t)kk * (a_sub >> 4) + 2, w0 + (uint32_t)kk * (1024 >> 4) + 2, hi, idesc, 1);
        }
        umma_commit(&bars->empty[s]);
        umma_commit(&bars->acc_full[a]);
        if (++s == nstages) { s = 0; ph ^= 1; }
      }
    }
  } else {
    const int quad = warp & 3;  // TMEM lane quadrant of this warp
    const int i = quad * 32 + lane;  // tile-local input row = output index
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    int it = 0;
    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x, ++it) {
      int b, t0; fdt.divmod(tile, b, t0); t0 *= T_OUT;
      const int a = it & (T_NACC - 1);
      const long long cut = p.item_skip ? __ldg(p.item_skip + b) : 0;
      mbar_wait(&bars->acc_full[a], (uint32_t)(it / T_NACC) & 1u);
      tc_fence_after();
      uint32_t r[16];
      __syncwarp();
      tmem_ld16(tmem_base + lane_base + (uint32_t)(a * 16), r);
      tmem_ld_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc_empty[a]);
      float* P = sP + (it & 1) * 7 * T_PPITCH;
#pragma unroll
      for (int j = 0; j < 7; ++j) P[j * T_PPITCH + i] = __uint_as_float(r[j]);
      asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps; the other buffer is rewritten only after the next barrier
      if (i < T_OUT) {
        float v = p.bias;
#pragma unroll
        for (int j = 0; j < 7; ++j) v += P[j * T_PPITCH + i + j];
        const int t = t0 + i;
        const long long g = p.out_pos0 + (t - p.tail_skip);  // position in the item's whole (chunk-concatenated) waveform
        if (t < p.rows_valid && t >= p.tail_skip && g >= cut) {
          const float c = fminf(1.0f, fmaxf(-1.0f, v));
          const long long o = (long long)b * p.out_batch_stride + (g - cut);
          if (p.out_i16) p.out_i16[o] = (short)__float2int_rn(c * 32767.0f);
          else p.out[o] = c;
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 128); }
}

static size_t tail_smem(int C, int nstages) {
  return 8 * 1024 + (size_t)nstages * (C / 32) * T_ROWS * 64 + (size_t)2 * 7 * T_PPITCH * sizeof(float) + sizeof(TailBarriers) + 1024;
}

bool tail_supported(int C) { return C % 32 == 0 && C >= 32 && C <= 256 && !Q3_ENV("Q3TTS_OLD_TAIL"); }

// a: [batches][a_rows][C] 16-bit (a_batch_stride elements between items, 0 = dense); w16: [16][C] 16-bit, rows 0..6 = taps, rest zero
int tail_make(TailOp* op, const void* a, const void* w16, int batches, int a_rows, long long a_batch_stride, int C, bool bf16) {
  op->enabled = 0;
  if (!tail_supported(C)) return 1;
  int ns = T_MAXSTAGES;
  while (ns > 2 && tail_smem(C, ns) > (size_t)gemm_smem_optin()) --ns;
  if (tail_smem(C, ns) > (size_t)gemm_smem_optin()) return 1;
  op->nstages = ns;
  int rc = make_tmap_3d_strided(&op->tmA, a, bf16 ? 1 : 0, (uint64_t)C, (uint64_t)a_rows, (uint64_t)batches, (uint64_t)C,
                                a_batch_stride ? (uint64_t)a_batch_stride : (uint64_t)a_rows * C, 32, T_ROWS);
  rc |= make_tmap_3d(&op->tmW, w16, (uint64_t)C, 16, 1, 32, 16, bf16);
  if (rc) return -1;
  op->p.C = C; op->p.bf16 = bf16 ? 1 : 0; op->p.batches = batches; op->p.out = nullptr; op->p.out_i16 = nullptr; op->p.out_pos0 = 0; op->p.item_skip = nullptr;
  op->enabled = 1;
  return 0;
}

int tail_setup() {  // per device (see unit_setup)
  if (cudaFuncSetAttribute(tail_conv_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_optin()) != cudaSuccess ||
      cudaFuncSetAttribute(tail_conv_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_optin()) != cudaSuccess)
    return -2;
  return 0;
}

cudaError_t launch_tail(const TailOp& op, cudaStream_t stream) {
  if (!op.enabled) return cudaErrorInvalidConfiguration;
  const TailParams& p = op.p;
  const int tiles = p.batches * ((p.rows_valid + T_OUT - 1) / T_OUT);
  if (tiles <= 0) return cudaSuccess;
  const int grid = tiles < gemm_num_sms() ? tiles : gemm_num_sms();
  const size_t smem = tail_smem(p.C, op.nstages);
  if (p.bf16) return launch_k(tail_conv_kernel<__nv_bfloat16>, dim3(grid), dim3(192), smem, stream, 1, op.tmA, op.tmW, p, op.nstages);
  return launch_k(tail_conv_kernel<__half>, dim3(grid), dim3(192), smem, stream, 1, op.tmA, op.tmW, p, op.nstages);
}

}  // namespace q3
