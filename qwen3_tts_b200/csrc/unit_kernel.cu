// Fused vocoder residual unit on tcgen05 tensor cores (sm_100a):
//     y = x + W1 . snake2(W7 (*) a1 + b7) + b1        (reference modeling :618-634, dilation 1 / 3 / 9)
// a1 = snake1(x) and x arrive as two 16-bit channels-last tensors; y and the next layer's Snake'd copy leave the same way.
//
// Warp roles (one CTA per SM, persistent over tiles of 128*MS rows of one batch item):
//   warp 0      TMA producer: per K block one a1 slab (tile rows + 6*dil halo rows), then the 7 tap weight tiles;
//               after the conv7 operands the K blocks of W1.  Two rings: slabs and weight tiles.
//   warp 1      TMEM allocator + single-thread MMA issuer:
//                 MMA1  acc1[ms] += slab(rows shifted by the tap) . W7[tap]      (A and B in shared memory)
//                 MMA2  acc2[ms]  = A2[ms] . W1                                  (A in TENSOR MEMORY, B in shared memory)
//   warps 2..9  epilogue, two phases per tile:
//                 phase 1  acc1 -> + b7 -> SnakeBeta -> 16-bit pairs -> tcgen05.st into A2 (never leaves the SM)
//                 phase 2  acc2 -> + b1 + residual x (TMA ring) -> y (TMA store) and snake_next(y) (TMA store)
// TMEM columns: MS * (C fp32 acc1 + C fp32 acc2 + C/2 packed A2) = 2.5 * MS * C <= 512  (C = 96: MS = 2; C = 192: MS = 1).
// Accumulators are single-buffered, so the tensor pipe idles during phase 1 of every tile; phase 2 overlaps the next
// tile's MMA1.  The layer is bound by HBM / epilogue ALU at C = 96 and is ~88 % tensor-busy at C = 192.
#include "conv_gemm.h"
#include "tc_epilogue.cuh"

#include <cstdlib>
#include <cstring>

namespace q3 {

static constexpr int U_BM = 128;
static constexpr int kUnitThreads = 64 + 32 * 8;
static constexpr int kUMaxSlabs = 6;
static constexpr int kUMaxStages = 12;

struct __align__(8) UnitBarriers {
  uint64_t sfull[kUMaxSlabs], sempty[kUMaxSlabs];
  uint64_t wfull[kUMaxStages], wempty[kUMaxStages];
  uint64_t acc1_full, a2_full, acc2_full, acc2_empty;
  uint64_t res_bar[kMaxEpiWarps][kMaxResDepth];
  uint32_t tmem_base;
  uint32_t pad;
};

struct UnitArgs {
  const float* bias7;
  const float* s2a;
  const float* s2ib;
  int C, dil, ms, srows, nbox, nslabs, nstages;
};

template <typename T>
__global__ void __launch_bounds__(kUnitThreads, 1)
unit_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW7, const __grid_constant__ CUtensorMap tmW1,
            const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmRaw, const __grid_constant__ CUtensorMap tmSnake,
            const GemmParams p, const UnitArgs u) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int KB = p.kb, C = u.C, MS = u.ms;
  const int nkb = C / KB;
  const int halo = 6 * u.dil;
  const int TM = U_BM * MS;
  const uint32_t span = KB * 2;
  const uint32_t ssub = u.srows * span;  // one slab sub-block (all rows of one K block)
  const uint32_t wsub = C * span;        // one weight tile (C output channels x KB)
  uint8_t* slabs = smem;
  uint8_t* wst = slabs + (size_t)u.nslabs * ssub;
  uint8_t* epi = wst + (size_t)u.nstages * wsub;
  UnitBarriers* bars = reinterpret_cast<UnitBarriers*>(epi + (size_t)8 * epi_slots_per_warp(p) * kEpiSlot);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  TileSched ts;
  ts.m_tiles = (p.rows_valid + TM - 1) / TM;
  ts.n_tiles = 1;
  ts.total_tiles = p.batches * ts.m_tiles;

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW7); tma_prefetch_desc(&tmW1);
    for (int s = 0; s < u.nslabs; ++s) { mbar_init(&bars->sfull[s], 1); mbar_init(&bars->sempty[s], 1); }
    for (int s = 0; s < u.nstages; ++s) { mbar_init(&bars->wfull[s], 1); mbar_init(&bars->wempty[s], 1); }
    mbar_init(&bars->acc1_full, 1); mbar_init(&bars->a2_full, 8); mbar_init(&bars->acc2_full, 1); mbar_init(&bars->acc2_empty, 8);
    for (int w = 0; w < kMaxEpiWarps; ++w)
      for (int s = 0; s < kMaxResDepth; ++s) mbar_init(&bars->res_bar[w][s], 1);
    fence_barrier_init();
  }
  if (warp == 1) { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = bars->tmem_base;
  // TMEM column map
  const uint32_t c_acc1 = 0, c_acc2 = (uint32_t)(MS * C), c_a2 = (uint32_t)(2 * MS * C);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (elect_one()) {
      int ss = 0, ws = 0; uint32_t sph = 0, wph = 0;
      const int boxrows = u.srows / u.nbox;
      for (int tile = blockIdx.x; tile < ts.total_tiles; tile += gridDim.x) {
        int nt, mt, b;
        ts.decode(tile, nt, mt, b);
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->sempty[ss], sph ^ 1);
          mbar_expect_tx(&bars->sfull[ss], ssub);
          for (int bx = 0; bx < u.nbox; ++bx)
            tma_load_3d(slabs + (size_t)ss * ssub + (size_t)bx * boxrows * span, &tmA, &bars->sfull[ss], kb * KB, mt * TM - halo + bx * boxrows, b);
          if (++ss == u.nslabs) { ss = 0; sph ^= 1; }
          for (int tap = 0; tap < 7; ++tap) {
            mbar_wait_backoff(&bars->wempty[ws], wph ^ 1);
            mbar_expect_tx(&bars->wfull[ws], wsub);
            tma_load_3d(wst + (size_t)ws * wsub, &tmW7, &bars->wfull[ws], kb * KB, 0, tap);
            if (++ws == u.nstages) { ws = 0; wph ^= 1; }
          }
        }
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->wempty[ws], wph ^ 1);
          mbar_expect_tx(&bars->wfull[ws], wsub);
          tma_load_3d(wst + (size_t)ws * wsub, &tmW1, &bars->wfull[ws], kb * KB, 0, 0);
          if (++ws == u.nstages) { ws = 0; wph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread)
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(U_BM, C, p.bf16 != 0);
      const uint32_t hi = umma_desc_hi(span);
      const uint32_t s_lo0 = umma_desc_lo(smem_u32(slabs)), s_off = ssub >> 4;
      const uint32_t w_lo0 = umma_desc_lo(smem_u32(wst)), w_off = wsub >> 4;
      const int ksteps = KB / 16;  // MMAs (K = 16) per stage
      int ss = 0, ws = 0; uint32_t sph = 0, wph = 0; int it = 0;
      for (int tile = blockIdx.x; tile < ts.total_tiles; tile += gridDim.x, ++it) {
        // ---- MMA1: dilated conv7 into acc1 (acc1 is free: a2_full of the previous tile was waited on before its MMA2)
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->sfull[ss], sph);
          const uint32_t slab = s_lo0 + (uint32_t)ss * s_off;
          for (int tap = 0; tap < 7; ++tap) {
            mbar_wait_backoff(&bars->wfull[ws], wph);
            tc_fence_after();
            const uint32_t b = w_lo0 + (uint32_t)ws * w_off;
            for (int ms = 0; ms < MS; ++ms) {
              // rows of this sub-tile for this tap start (ms*128 + halo - (6 - tap)*dil) rows into the slab
              const uint32_t a = slab + (((uint32_t)(ms * U_BM + halo - (6 - tap) * u.dil) * span) >> 4);
              const uint32_t d = tmem_base + c_acc1 + (uint32_t)(ms * C);
              for (int k = 0; k < ksteps; ++k) umma_f16_lohi(d, a + 2 * k, b + 2 * k, hi, idesc, (kb | tap | k) != 0);
            }
            umma_commit(&bars->wempty[ws]);
            if (++ws == u.nstages) { ws = 0; wph ^= 1; }
          }
          umma_commit(&bars->sempty[ss]);
          if (++ss == u.nslabs) { ss = 0; sph ^= 1; }
        }
        umma_commit(&bars->acc1_full);
        // ---- MMA2: 1x1 conv, A = Snake'd conv7 output packed in tensor memory by epilogue phase 1
        mbar_wait_backoff(&bars->a2_full, it & 1);
        mbar_wait_backoff(&bars->acc2_empty, (it & 1) ^ 1);
        tc_fence_after();
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->wfull[ws], wph);
          tc_fence_after();
          const uint32_t b = w_lo0 + (uint32_t)ws * w_off;
          for (int ms = 0; ms < MS; ++ms) {
            const uint32_t d = tmem_base + c_acc2 + (uint32_t)(ms * C);
            const uint32_t a = tmem_base + c_a2 + (uint32_t)(ms * (C / 2)) + (uint32_t)(kb * (KB / 2));
            for (int k = 0; k < ksteps; ++k) umma_f16_ts(d, a + 8 * k, b + 2 * k, hi, idesc, (kb | k) != 0);
          }
          umma_commit(&bars->wempty[ws]);
          if (++ws == u.nstages) { ws = 0; wph ^= 1; }
        }
        umma_commit(&bars->acc2_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int quad = warp & 3, ew = warp - 2, half = ew >> 2;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    // phase-2 machinery (residual prefetch ring, TMA stores) is the shared asynchronous epilogue
    EpiWarp<T> ep(p, &tmRes, &tmRaw, &tmSnake, &tmRaw, epi + (size_t)ew * epi_slots_per_warp(p) * kEpiSlot, bars->res_bar[ew], ts, lane, quad,
                  MS == 2 ? 0 : half, MS == 2 ? 1 : 2);
    if (MS == 2) { ep.mt_mul = 2; ep.mt_add = half; }
    ep.start();
    const int units = C >> 4;
    const int ms1 = MS == 2 ? half : 0;                        // sub-tile this warp serves
    const int u_lo = MS == 2 ? 0 : half * (units / 2);         // 16-column units of phase 1
    const int u_hi = MS == 2 ? units : u_lo + units / 2;
    int it = 0;
    for (int tile = blockIdx.x; tile < ts.total_tiles; tile += gridDim.x, ++it) {
      int nt, mt, b;
      ts.decode(tile, nt, mt, b);
      // ---- phase 1
      mbar_wait(&bars->acc1_full, it & 1);
      tc_fence_after();
      {
        const uint32_t src = tmem_base + lane_base + c_acc1 + (uint32_t)(ms1 * C);
        const uint32_t dst = tmem_base + lane_base + c_a2 + (uint32_t)(ms1 * (C / 2));
        for (int uu = u_lo; uu < u_hi; uu += 2) {
          uint32_t ra[16], rb[16];
          tmem_ld16(src + uu * 16, ra);
          tmem_ld16(src + uu * 16 + 16, rb);
          float ba[16], bb[16], sa[16], sb[16], ia[16], ib[16];
          load16f(u.bias7 + uu * 16, ba); load16f(u.bias7 + uu * 16 + 16, bb);
          load16f(u.s2a + uu * 16, sa); load16f(u.s2a + uu * 16 + 16, sb);
          load16f(u.s2ib + uu * 16, ia); load16f(u.s2ib + uu * 16 + 16, ib);
          tmem_ld_wait();
          uint32_t pa[8], pb[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float a0 = snake(__uint_as_float(ra[2 * i]) + ba[2 * i], sa[2 * i], ia[2 * i]);
            const float a1 = snake(__uint_as_float(ra[2 * i + 1]) + ba[2 * i + 1], sa[2 * i + 1], ia[2 * i + 1]);
            const float b0 = snake(__uint_as_float(rb[2 * i]) + bb[2 * i], sb[2 * i], ib[2 * i]);
            const float b1 = snake(__uint_as_float(rb[2 * i + 1]) + bb[2 * i + 1], sb[2 * i + 1], ib[2 * i + 1]);
            pa[i] = Elem<T>::pack2(a0, a1);
            pb[i] = Elem<T>::pack2(b0, b1);
          }
          tmem_st8(dst + uu * 8, pa);
          tmem_st8(dst + uu * 8 + 8, pb);
        }
        tmem_st_wait();
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&bars->a2_full);
      }
      // ---- phase 2
      mbar_wait(&bars->acc2_full, it & 1);
      tc_fence_after();
      ep.tile(tmem_base + lane_base + c_acc2 + (uint32_t)(ms1 * C), it, 0, mt * MS + (MS == 2 ? half : 0), b);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&bars->acc2_empty);
    }
    ep.finish();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------- host
static size_t unit_smem_bytes(const UnitOp& op) {
  const size_t span = (size_t)op.p.kb * 2;
  return (size_t)op.nslabs * op.srows * span + (size_t)op.nstages * op.C * span + (size_t)8 * epi_slots_per_warp(op.p) * kEpiSlot +
         sizeof(UnitBarriers) + 1024;
}

bool unit_fusable(int C) {
  if (getenv("Q3TTS_NO_FUSE")) return false;
  if (C % 32 || C > 256 || C < 32) return false;
  const int ms = C <= 96 ? 2 : 1;
  return 5 * ms * C / 2 <= 512;
}

int unit_make(UnitOp* op, const void* a1, const void* w7, const void* w1, int batches, int rows) {
  GemmParams& p = op->p;
  op->enabled = 0;
  const int C = op->C;
  if (getenv("Q3TTS_NO_FUSE")) return 1;
  if (C % 32 || C > 256 || C < 32) return 1;
  op->ms = C <= 96 ? 2 : 1;
  if (5 * op->ms * C / 2 > 512) return 1;  // TMEM columns: acc1 + acc2 + packed A2
  p.kb = (C % 64 == 0) ? 64 : 32;
  p.epi_warps = 8;
  p.res_depth = 2;
  p.batches = batches; p.rows = rows; p.rows_valid = rows; p.cin = C; p.ntaps = 1; p.tap_off[0] = 0;
  p.n_total = C; p.bn = C; p.ncol_mod = C; p.act = ACT_NONE; p.res_f32 = 0; p.out_f32 = nullptr; p.scale = nullptr;
  p.out_row_stride = C; p.out_batch_stride = (long long)rows * C;
  p.a_ptr = a1;
  if (!p.res || !p.out_raw || !p.out_snake) return 1;
  const int span = p.kb * 2;
  const int need = U_BM * op->ms + 6 * op->dil;
  op->srows = (need + 15) / 16 * 16;
  op->nbox = op->srows > 256 ? 2 : 1;
  if (op->srows / op->nbox > 256 || (op->srows / op->nbox) % 8) return 1;
  const long long fixed = 1024 + (long long)8 * epi_slots_per_warp(p) * kEpiSlot + (long long)sizeof(UnitBarriers) + 256;
  const long long avail = (long long)gemm_smem_optin() - fixed;
  const long long slab = (long long)op->srows * span, wst = (long long)C * span;
  const int want_st = wst >= 16384 ? 3 : 6;
  op->nslabs = 0;
  for (int ns = 4; ns >= 2; --ns) {
    long long st = (avail - ns * slab) / wst;
    if (st >= want_st) { op->nslabs = ns; op->nstages = (int)(st > kUMaxStages ? kUMaxStages : st); break; }
  }
  if (!op->nslabs) return 1;
  const int et = p.bf16 ? 1 : 0;
  int rc = 0;
  rc |= make_tmap_3d(&op->tmA, a1, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, (uint32_t)p.kb, (uint32_t)(op->srows / op->nbox), p.bf16 != 0);
  rc |= make_tmap_3d(&op->tmW7, w7, (uint64_t)C, (uint64_t)C, 7, (uint32_t)p.kb, (uint32_t)C, p.bf16 != 0);
  rc |= make_tmap_3d(&op->tmW1, w1, (uint64_t)C, (uint64_t)C, 1, (uint32_t)p.kb, (uint32_t)C, p.bf16 != 0);
  const uint64_t rs = (uint64_t)p.out_row_stride, bs = (uint64_t)p.out_batch_stride;
  rc |= make_tmap_3d_strided(&op->tmRes, p.res, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs, bs, 32, 32);
  rc |= make_tmap_3d_strided(&op->tmRaw, p.out_raw, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs, bs, 32, 32);
  rc |= make_tmap_3d_strided(&op->tmSnake, p.out_snake, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs, bs, 32, 32);
  if (rc) return -1;
  static bool attr_done = false;
  if (!attr_done) {
    if (cudaFuncSetAttribute(unit_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_optin()) != cudaSuccess ||
        cudaFuncSetAttribute(unit_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, gemm_smem_optin()) != cudaSuccess)
      return -2;
    attr_done = true;
  }
  op->enabled = 1;
  return 0;
}

cudaError_t launch_unit(const UnitOp& op, cudaStream_t stream) {
  if (!op.enabled) return cudaErrorInvalidConfiguration;
  const GemmParams& p = op.p;
  const int TM = U_BM * op.ms;
  const int tiles = p.batches * ((p.rows_valid + TM - 1) / TM);
  if (tiles <= 0) return cudaSuccess;
  const int grid = tiles < gemm_num_sms() ? tiles : gemm_num_sms();
  UnitArgs u{op.bias7, op.snake2_a, op.snake2_ib, op.C, op.dil, op.ms, op.srows, op.nbox, op.nslabs, op.nstages};
  const size_t smem = unit_smem_bytes(op);
  if (p.bf16)
    unit_kernel<__nv_bfloat16><<<grid, kUnitThreads, smem, stream>>>(op.tmA, op.tmW7, op.tmW1, op.tmRes, op.tmRaw, op.tmSnake, p, u);
  else
    unit_kernel<__half><<<grid, kUnitThreads, smem, stream>>>(op.tmA, op.tmW7, op.tmW1, op.tmRes, op.tmRaw, op.tmSnake, p, u);
  return cudaGetLastError();
}

}  // namespace q3
