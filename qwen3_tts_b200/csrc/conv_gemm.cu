// Implicit-GEMM causal conv / transposed conv / linear on tcgen05 tensor cores (sm_100a).
//
// Two kernels share one warp-role layout (warp 0 = TMA producer, warp 1 = TMEM allocator +
// single-thread tcgen05.mma issuer, warps 2..5 = asynchronous epilogue, tc_epilogue.cuh), one CTA
// per SM, persistent over tiles, accumulators double-buffered in TMEM (2 x 256 fp32 columns):
//
//  conv_gemm_tc_kernel  general: pipeline stage = one (K block, tap) pair: a time-shifted 128-row
//                       activation box plus the tap's weight tile.  Used when weights are too big
//                       to stay resident (C >= 192).
//  ws_conv_kernel       weight-stationary: all taps' weights are loaded into shared memory once per
//                       CTA; per tile ONE activation slab (128 + halo rows) is loaded and every tap
//                       reads it through a row-offset UMMA descriptor, so L2->smem traffic per tile
//                       is the slab alone.  Used for the C = 96 layers and the 96 -> 1 output conv.
//
// Operand tiles are K-major, written by TMA with 128B (KB = 64) or 64B (KB = 32) swizzle; rows
// outside [0, rows) are zero-filled by TMA (= the reference's causal left padding).
#include "conv_gemm.h"
#include "launch.cuh"
#include "pair_ptx.cuh"
#include "tc_epilogue.cuh"

#include <cstdlib>

namespace q3 {

static constexpr int BM = 128;
static constexpr int kMaxGemmThreads = 64 + 32 * kMaxEpiWarps;
static constexpr int kMaxStages = 8;
static constexpr int kMaxSlabs = 4;

struct __align__(8) GemmBarriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t res_bar[kMaxEpiWarps][kMaxResDepth];
  uint64_t w_full;
  uint64_t sfull[kMaxSlabs];   // slab-stream kernel: activation slab ring
  uint64_t sempty[kMaxSlabs];
  uint32_t tmem_base;
  uint32_t pad;
  uint32_t a_tab[64];  // ws kernel: per (K block, tap) A descriptor offset (>> 4) inside a slab
  uint32_t b_tab[64];  // ws kernel: per (K block, tap) B descriptor low word
};

// pair: CTA-pair form (2-CTA cluster, tcgen05 cta_group::2): the leader's tmem_empty collects both CTAs' epilogue warps
__device__ __forceinline__ void init_common(GemmBarriers* bars, int nring, int warp, int epi_warps, bool pair = false) {
  if (threadIdx.x == 0) {
    for (int s = 0; s < nring; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&bars->tmem_full[s], 1); mbar_init(&bars->tmem_empty[s], pair ? 2 * epi_warps : epi_warps); }
    for (int w = 0; w < kMaxEpiWarps; ++w)
      for (int s = 0; s < kMaxResDepth; ++s) mbar_init(&bars->res_bar[w][s], 1);
    mbar_init(&bars->w_full, 1);
    for (int s = 0; s < kMaxSlabs; ++s) { mbar_init(&bars->sfull[s], 1); mbar_init(&bars->sempty[s], 1); }
    fence_barrier_init();
  }
  if (warp == 1) {
    if (pair) { tmem_alloc2(&bars->tmem_base, 512); tmem_relinquish2(); }
    else { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (pair) cluster_sync_all();  // the peer's barriers exist before anything arrives on them
  tc_fence_after();
  // Programmatic dependent launch (launch.cuh).  Trigger only now, AFTER this CTA holds its tensor memory: a kernel two launches ahead
  // could otherwise allocate TMEM on this SM first and then wait for us forever.  Everything above ran in the shadow of the previous
  // kernel; after the wait its outputs may be read and its inputs overwritten.
  pdl_trigger();
  pdl_wait();
}

template <typename T>
__device__ __forceinline__ void epilogue_role(const GemmParams& p, const CUtensorMap* tmRes, const CUtensorMap* tmRaw, const CUtensorMap* tmSnake,
                                              const CUtensorMap* tmF32, uint8_t* epi, GemmBarriers* bars, const TileSched& ts, uint32_t tmem_base,
                                              int warp, int lane, bool pair = false, int rank = 0) {
  const int quad = warp & 3;  // TMEM lane quadrant this warp may access
  const int ew = warp - 2;
  EpiWarp<T> ep(p, tmRes, tmRaw, tmSnake, tmF32, epi + (size_t)ew * epi_slots_per_warp(p) * kEpiSlot, bars->res_bar[ew], ts, lane, quad, ew >> 2,
                p.epi_warps >> 2);
  if (pair) { ep.mt_mul = 2; ep.mt_add = rank; ep.tile0 = (int)(blockIdx.x >> 1); ep.tile_stride = (int)(gridDim.x >> 1); }
  ep.start();
  int it = 0;
  for (int tile = ep.tile0; tile < ts.total_tiles; tile += ep.tile_stride, ++it) {
    int nt, mt, b;
    ts.decode(tile, nt, mt, b);
    const int buf = it & 1;
    mbar_wait(&bars->tmem_full[buf], (it >> 1) & 1);
    tc_fence_after();
    ep.tile(tmem_base + ((uint32_t)(quad * 32) << 16) + buf * 256, it, nt, mt, b);
    tc_fence_before();
    __syncwarp();
    if (lane == 0) { if (pair) mbar_arrive_leader(&bars->tmem_empty[buf]); else mbar_arrive(&bars->tmem_empty[buf]); }
  }
  ep.finish();
}

// PAIR: CTA-pair form (see slab_conv_kernel): 256-row tiles, each CTA loads its 128 rows of A and half of the weight tile, the
// even CTA issues M = 256 tcgen05.mma.cta_group::2.  Not combined with packed tiles.
template <typename T, bool PAIR>
__global__ void __launch_bounds__(kMaxGemmThreads, 1)
conv_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmRaw,
                    const __grid_constant__ CUtensorMap tmSnake, const __grid_constant__ CUtensorMap tmF32,
                    const GemmParams p, const int nstages) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset on the shared array: keeps LDS / STS (a uintptr_t round trip makes every access generic)
  const int KB = p.kb;
  constexpr uint32_t NCTA = PAIR ? 2u : 1u;
  constexpr int TM = PAIR ? 2 * BM : BM;
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int wid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int nworkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  const int bn_cta = PAIR ? p.bn / 2 : p.bn;
  const uint32_t a_bytes = BM * KB * 2;
  const uint32_t b_bytes = bn_cta * KB * 2;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  uint8_t* epi = smem + (size_t)nstages * stage_bytes;
  GemmBarriers* bars = reinterpret_cast<GemmBarriers*>(epi + (size_t)p.epi_warps * epi_slots_per_warp(p) * kEpiSlot);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  TileSched ts;
  ts.m_tiles = p.pack_mode ? 1 : (p.rows_valid + TM - 1) / TM;  // packed: decode() returns the tile index as "batch"
  ts.n_tiles = p.n_total / p.bn;
  ts.total_tiles = (PAIR ? p.batches * ts.m_tiles : gemm_m_tiles_total(p)) * ts.n_tiles;
  ts.init();
  const int kblocks = p.cin / KB;
  const int ksteps = kblocks * p.ntaps;
  if (threadIdx.x == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  init_common(bars, nstages, warp, p.epi_warps, PAIR);
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer
    if (elect_one()) {
      auto load = [&](uint8_t* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_pair(smem_u32(dst), map, leader_addr(smem_u32(bar)), c0, c1, c2);
        else tma_load_3d(dst, map, bar, c0, c1, c2);
      };
      int s = 0; uint32_t ph = 0;
      for (int tile = wid; tile < ts.total_tiles; tile += nworkers) {
        int nt, mt, b;
        ts.decode(tile, nt, mt, b);
        const int m0 = mt * TM + rank * BM, n0 = nt * p.bn + rank * bn_cta;
        int cr[4] = {0, 0, 0, 0}, cb[4] = {0, 0, 0, 0};
        if (!PAIR && p.pack_mode)
          for (int q = 0; q < 4; ++q) gemm_chunk_coords(p, b, q, cr[q], cb[q]);
        for (int kb = 0; kb < kblocks; ++kb) {
          const int acol = gemm_a_col(p, kb * KB);  // split-precision operands: K blocks wrap around the A tensor (per K group)
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait_backoff(&bars->empty[s], ph ^ 1);
            uint8_t* sa = smem + (size_t)s * stage_bytes;
            if (rank == 0) mbar_expect_tx(&bars->full[s], stage_bytes * NCTA);
            if (!PAIR && p.pack_mode) {  // four 32-row chunk boxes fill the 128-row A tile
              for (int q = 0; q < 4; ++q)
                tma_load_3d(sa + q * (a_bytes >> 2), &tmA, &bars->full[s], acol, cr[q] + p.a_row0 + p.tap_off[tap], cb[q]);
            } else {
              load(sa, &tmA, &bars->full[s], acol, m0 + p.a_row0 + p.tap_off[tap], b);
            }
            load(sa + a_bytes, &tmB, &bars->full[s], kb * KB, n0, tap);
            if (++s == nstages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (one thread; pair: of the even CTA only)
    if (rank == 0 && elect_one()) {
      const uint32_t idesc = umma_idesc(TM, p.bn, p.bf16 != 0);
      const uint32_t hi = umma_desc_hi(KB * 2);
      const uint32_t a_lo0 = umma_desc_lo(smem_u32(smem)), b_off = a_bytes >> 4, st_off = stage_bytes >> 4;
      const bool k4 = KB == 64;
      auto mma = [&](uint32_t d, uint32_t a, uint32_t b, uint32_t acc) {
        if (PAIR) umma2_f16_lohi(d, a, b, hi, idesc, acc); else umma_f16_lohi(d, a, b, hi, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) { if (PAIR) umma2_commit(bar); else umma_commit(bar); };
      int s = 0; uint32_t ph = 0; int it = 0;
      for (int tile = wid; tile < ts.total_tiles; tile += nworkers, ++it) {
        const int buf = it & 1;
        if (PAIR) mbar_wait_cluster(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        else mbar_wait_backoff(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        for (int ks = 0; ks < ksteps; ++ks) {
          mbar_wait_backoff(&bars->full[s], ph);
          tc_fence_after();
          const uint32_t a = a_lo0 + (uint32_t)s * st_off, b = a + b_off;
          if (!Q3_DBG(p, 2)) {
            // +2 in the >>4 address field advances 16 elements (32 bytes) along K inside the swizzle span
            mma(d_tmem, a, b, ks != 0);
            mma(d_tmem, a + 2, b + 2, 1);
            if (k4) {
              mma(d_tmem, a + 4, b + 4, 1);
              mma(d_tmem, a + 6, b + 6, 1);
            }
          }
          commit(&bars->empty[s]);
          if (++s == nstages) { s = 0; ph ^= 1; }
        }
        commit(&bars->tmem_full[buf]);
      }
    }
  } else {
    epilogue_role<T>(p, &tmRes, &tmRaw, &tmSnake, &tmF32, epi, bars, ts, tmem_base, warp, lane, PAIR, rank);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc2(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
  }
}

// Weight-stationary slab kernel (single N tile: bn == n_total; all tap offsets <= 0).
//   smem: [ntaps*nkb weight sub-tiles of bn x KB][nslabs x nkb slab sub-blocks of srp x KB][epilogue slots][barriers]
// The slab row for output row m and tap j is m + tap_off[j] + halo; the UMMA descriptor of a tap simply
// starts (tap_off[j] + halo) rows into the slab sub-block -- the swizzle is a function of the shared-memory
// address bits, so a row-shifted start address reads the rows TMA wrote.
template <typename T>
__global__ void __launch_bounds__(kMaxGemmThreads, 1)
ws_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmRaw,
               const __grid_constant__ CUtensorMap tmSnake, const __grid_constant__ CUtensorMap tmF32,
               const GemmParams p, const int nslabs, const int srp) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset on the shared array: keeps LDS / STS (a uintptr_t round trip makes every access generic)
  const int KB = p.kb;
  const int nkb = p.cin / KB;
  const int halo = -p.tap_off[0];
  const uint32_t wsub = p.bn * KB * 2;
  const uint32_t ssub = srp * KB * 2;
  const uint32_t w_bytes = p.ntaps * nkb * wsub;
  const uint32_t slab_bytes = nkb * ssub;
  uint8_t* wsm = smem;
  uint8_t* slabs = smem + w_bytes;
  uint8_t* epi = slabs + (size_t)nslabs * slab_bytes;
  GemmBarriers* bars = reinterpret_cast<GemmBarriers*>(epi + (size_t)p.epi_warps * epi_slots_per_warp(p) * kEpiSlot);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  TileSched ts;
  ts.m_tiles = (p.rows_valid + BM - 1) / BM;
  ts.n_tiles = 1;
  ts.total_tiles = p.batches * ts.m_tiles;
  ts.init();
  if (threadIdx.x == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  init_common(bars, nslabs, warp, p.epi_warps);
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    if (elect_one()) {
      mbar_expect_tx(&bars->w_full, w_bytes);
      for (int tap = 0; tap < p.ntaps; ++tap)
        for (int kb = 0; kb < nkb; ++kb) tma_load_3d(wsm + (size_t)(tap * nkb + kb) * wsub, &tmB, &bars->w_full, kb * KB, 0, tap);
      int s = 0; uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < ts.total_tiles; tile += gridDim.x) {
        int nt, mt, b;
        ts.decode(tile, nt, mt, b);
        mbar_wait_backoff(&bars->empty[s], ph ^ 1);
        mbar_expect_tx(&bars->full[s], slab_bytes);
        for (int kb = 0; kb < nkb; ++kb)
          tma_load_3d(slabs + (size_t)s * slab_bytes + (size_t)kb * ssub, &tmA, &bars->full[s], gemm_a_col(p, kb * KB),
                      mt * BM + p.a_row0 - halo, b);
        if (++s == nslabs) { s = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    if (elect_one()) {
      const uint32_t idesc = umma_idesc(BM, p.bn, p.bf16 != 0);
      const uint32_t span = KB * 2;
      const uint32_t hi = umma_desc_hi(span);
      const int nsteps = nkb * p.ntaps;
      // descriptor tables: everything that does not depend on the slab slot is computed once
      for (int kb = 0; kb < nkb; ++kb)
        for (int tap = 0; tap < p.ntaps; ++tap) {
          bars->a_tab[kb * p.ntaps + tap] = (kb * ssub + (uint32_t)(p.tap_off[tap] + halo) * span) >> 4;
          bars->b_tab[kb * p.ntaps + tap] = umma_desc_lo(smem_u32(wsm) + (uint32_t)(tap * nkb + kb) * wsub);
        }
      const bool k4 = KB == 64;
      mbar_wait_backoff(&bars->w_full, 0);
      int s = 0; uint32_t ph = 0; int it = 0;
      for (int tile = blockIdx.x; tile < ts.total_tiles; tile += gridDim.x, ++it) {
        const int buf = it & 1;
        mbar_wait_backoff(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        mbar_wait_backoff(&bars->full[s], ph);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        const uint32_t slab_lo = umma_desc_lo(smem_u32(slabs + (size_t)s * slab_bytes));
        if (!Q3_DBG(p, 2)) {
#pragma unroll 4
          for (int i = 0; i < nsteps; ++i) {
            const uint32_t a = slab_lo + bars->a_tab[i], b = bars->b_tab[i];
            umma_f16_lohi(d_tmem, a, b, hi, idesc, i != 0);
            umma_f16_lohi(d_tmem, a + 2, b + 2, hi, idesc, 1);
            if (k4) {
              umma_f16_lohi(d_tmem, a + 4, b + 4, hi, idesc, 1);
              umma_f16_lohi(d_tmem, a + 6, b + 6, hi, idesc, 1);
            }
          }
        }
        umma_commit(&bars->empty[s]);
        umma_commit(&bars->tmem_full[buf]);
        if (++s == nslabs) { s = 0; ph ^= 1; }
      }
    }
  } else {
    epilogue_role<T>(p, &tmRes, &tmRaw, &tmSnake, &tmF32, epi, bars, ts, tmem_base, warp, lane);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}


// Slab-stream kernel (multi-tap convs whose weights do not fit in shared memory: C >= 192).
//   smem: [nslabs x (srp x KB) activation slabs][nstages x (bn x KB) weight tiles][epilogue slots][barriers]
// Per (tile, K block) ONE activation slab (128 + halo rows, time-shifted taps read it through row-offset UMMA
// descriptors exactly like ws_conv_kernel) and ntaps weight tiles are loaded, so the L2 -> shared-memory
// traffic per MMA drops from (128 + bn) to (srp / ntaps + bn) rows of K: the general kernel re-loads the
// activation box once per tap and is bound by that traffic, not by the tensor pipe (profiles/r1_*).
// slab_lo = rows of halo before the tile (-min tap_off), so it also serves the {0, +1} taps of a transposed conv.
// PAIR: CTA-pair form.  A 2-CTA cluster works on 256-row tiles: each CTA loads its own 128-row slabs and HALF of every weight
// tile (bn/2 output channels); the even CTA issues one M = 256 tcgen05.mma.cta_group::2 stream for both.  The single-CTA form is
// bound by shared-memory fill (one weight tile per 128 rows: 70 B/clk wanted at C = 384, ~43 B/clk available); halving the
// weight bytes per SM moves these layers back to the tensor pipe.
template <typename T, bool PAIR>
__global__ void __launch_bounds__(kMaxGemmThreads, 1)
slab_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmRaw,
                 const __grid_constant__ CUtensorMap tmSnake, const __grid_constant__ CUtensorMap tmF32,
                 const GemmParams p, const int nslabs, const int nstages, const int srp, const int slab_lo) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset on the shared array: keeps LDS / STS (a uintptr_t round trip makes every access generic)
  const int KB = p.kb;
  const int nkb = p.cin / KB;
  const int bn_cta = PAIR ? p.bn / 2 : p.bn;   // weight rows this CTA holds
  const uint32_t wsub = bn_cta * KB * 2;
  const uint32_t ssub = srp * KB * 2;
  constexpr uint32_t NCTA = PAIR ? 2u : 1u;
  constexpr int TM = PAIR ? 2 * BM : BM;
  const int rank = PAIR ? (int)cluster_ctarank() : 0;
  const int wid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int nworkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  uint8_t* slabs = smem;
  uint8_t* wst = smem + (size_t)nslabs * ssub;
  uint8_t* epi = wst + (size_t)nstages * wsub;
  GemmBarriers* bars = reinterpret_cast<GemmBarriers*>(epi + (size_t)p.epi_warps * epi_slots_per_warp(p) * kEpiSlot);
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  TileSched ts;
  ts.m_tiles = (p.rows_valid + TM - 1) / TM;
  ts.n_tiles = p.n_total / p.bn;
  ts.total_tiles = p.batches * ts.m_tiles * ts.n_tiles;
  ts.init();
  if (threadIdx.x == 0) { tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmB); }
  init_common(bars, nstages, warp, p.epi_warps, PAIR);
  const uint32_t tmem_base = bars->tmem_base;

  if (warp == 0) {
    if (elect_one()) {
      // pair: loads land in this CTA's shared memory and are counted on the leader's full barriers (it issues the MMAs)
      auto load = [&](uint8_t* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_pair(smem_u32(dst), map, leader_addr(smem_u32(bar)), c0, c1, c2);
        else tma_load_3d(dst, map, bar, c0, c1, c2);
      };
      int s = 0, ss = 0; uint32_t ph = 0, sph = 0;
      for (int tile = wid; tile < ts.total_tiles; tile += nworkers) {
        int nt, mt, b;
        ts.decode(tile, nt, mt, b);
        const int n0 = nt * p.bn + rank * bn_cta;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->sempty[ss], sph ^ 1);
          if (rank == 0) mbar_expect_tx(&bars->sfull[ss], ssub * NCTA);
          load(slabs + (size_t)ss * ssub, &tmA, &bars->sfull[ss], gemm_a_col(p, kb * KB), mt * TM + rank * BM + p.a_row0 - slab_lo, b);
          if (++ss == nslabs) { ss = 0; sph ^= 1; }
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait_backoff(&bars->empty[s], ph ^ 1);
            if (rank == 0) mbar_expect_tx(&bars->full[s], wsub * NCTA);
            load(wst + (size_t)s * wsub, &tmB, &bars->full[s], kb * KB, n0, tap);
            if (++s == nstages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    if (rank == 0 && elect_one()) {
      const uint32_t idesc = umma_idesc(TM, p.bn, p.bf16 != 0);
      const uint32_t span = KB * 2;
      const uint32_t hi = umma_desc_hi(span);
      for (int tap = 0; tap < p.ntaps; ++tap) bars->a_tab[tap] = ((uint32_t)(p.tap_off[tap] + slab_lo) * span) >> 4;
      const uint32_t w_lo0 = umma_desc_lo(smem_u32(wst)), w_off = wsub >> 4;
      const uint32_t s_lo0 = umma_desc_lo(smem_u32(slabs)), s_off = ssub >> 4;
      const bool k4 = KB == 64;
      auto mma = [&](uint32_t d, uint32_t a, uint32_t b, uint32_t acc) {
        if (PAIR) umma2_f16_lohi(d, a, b, hi, idesc, acc); else umma_f16_lohi(d, a, b, hi, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) { if (PAIR) umma2_commit(bar); else umma_commit(bar); };
      int s = 0, ss = 0; uint32_t ph = 0, sph = 0; int it = 0;
      for (int tile = wid; tile < ts.total_tiles; tile += nworkers, ++it) {
        const int buf = it & 1;
        if (PAIR) mbar_wait_cluster(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        else mbar_wait_backoff(&bars->tmem_empty[buf], ((it >> 1) & 1) ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * 256;
        for (int kb = 0; kb < nkb; ++kb) {
          mbar_wait_backoff(&bars->sfull[ss], sph);
          const uint32_t slab_lo_desc = s_lo0 + (uint32_t)ss * s_off;
          for (int tap = 0; tap < p.ntaps; ++tap) {
            mbar_wait_backoff(&bars->full[s], ph);
            tc_fence_after();
            const uint32_t a = slab_lo_desc + bars->a_tab[tap], b = w_lo0 + (uint32_t)s * w_off;
            if (!Q3_DBG(p, 2)) {
              mma(d_tmem, a, b, (kb | tap) != 0);
              mma(d_tmem, a + 2, b + 2, 1);
              if (k4) {
                mma(d_tmem, a + 4, b + 4, 1);
                mma(d_tmem, a + 6, b + 6, 1);
              }
            }
            commit(&bars->empty[s]);
            if (++s == nstages) { s = 0; ph ^= 1; }
          }
          commit(&bars->sempty[ss]);
          if (++ss == nslabs) { ss = 0; sph ^= 1; }
        }
        commit(&bars->tmem_full[buf]);
      }
    }
  } else if (warp >= 2) {
    epilogue_role<T>(p, &tmRes, &tmRaw, &tmSnake, &tmF32, epi, bars, ts, tmem_base, warp, lane, PAIR, rank);
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the leader's MMAs read the peer's shared and tensor memory until its last commit
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc2(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
  }
}

// SIMT cross-check: same math, direct global I/O, no tensor cores / TMA.  Debug engine only
// (Q3TTS_ENGINE_SIMT); one thread per (row, 16-column group).
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
      const int r = row + p.a_row0 + p.tap_off[tap];
      if (r < 0 || r >= p.rows) continue;
      const int aw = p.a_cols ? p.a_cols : p.cin;
      const long long ars = p.a_row_stride ? p.a_row_stride : aw, wrs = p.w_row_stride ? p.w_row_stride : p.cin;
      const T* a = A + (long long)b * (p.a_batch_stride ? p.a_batch_stride : (long long)p.rows * ars) + (long long)r * ars;
      const T* w = W + ((long long)tap * p.n_total + n) * wrs;
      for (int k = 0; k < p.cin; ++k) {
        const float av = Elem<T>::to_f(a[gemm_a_col(p, k)]);
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = fmaf(av, Elem<T>::to_f(w[(long long)i * wrs + k]), acc[i]);
      }
    }
    if (p.act == ACT_TAIL) {
      if (row >= p.tail_skip)
        p.out_f32[(long long)b * p.out_batch_stride + (row - p.tail_skip)] = fminf(1.0f, fmaxf(-1.0f, acc[0] + p.bias[0]));
      continue;
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

int make_tmap_3d_strided(CUtensorMap* out, const void* base, int etype, uint64_t d0, uint64_t d1, uint64_t d2, uint64_t stride1,
                         uint64_t stride2, uint32_t box0, uint32_t box1, uint32_t box2) {
  if (tmap_init()) return -1;
  const uint64_t esz = etype == 2 ? 4 : 2;
  cuuint64_t dims[3] = {d0, d1 ? d1 : 1, d2 ? d2 : 1};
  cuuint64_t strides[2] = {stride1 * esz, stride2 * esz};
  if (strides[1] == 0) strides[1] = strides[0] * dims[1];
  cuuint32_t box[3] = {box0, box1, box2 ? box2 : 1};
  cuuint32_t estr[3] = {1, 1, 1};
  const uint32_t inner = box0 * (uint32_t)esz;
  CUtensorMapSwizzle sw = inner == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                          : (inner == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : (inner == 32 ? CU_TENSOR_MAP_SWIZZLE_32B : CU_TENSOR_MAP_SWIZZLE_NONE));
  CUtensorMapDataType dt = etype == 2 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : (etype == 1 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
  CUresult r = g_encode(out, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS ? 0 : -2;
}

// 3-D tensor map over a dense 16-bit tensor [d2][d1][d0] (d0 contiguous), box [1][box1][box0].
int make_tmap_3d(CUtensorMap* out, const void* base, uint64_t d0, uint64_t d1, uint64_t d2, uint32_t box0, uint32_t box1, bool bf16) {
  return make_tmap_3d_strided(out, base, bf16 ? 1 : 0, d0, d1, d2, d0, d0 * d1, box0, box1, 1);
}

// weight map [taps][n_total][cin] with an optional row pitch (K groups of wider packed rows)
static int make_tmap_w(CUtensorMap* out, const GemmParams& p, uint32_t box_n) {
  const uint64_t wrs = p.w_row_stride ? (uint64_t)p.w_row_stride : (uint64_t)p.cin;
  return make_tmap_3d_strided(out, p.w_ptr, p.bf16 ? 1 : 0, (uint64_t)p.cin, (uint64_t)p.n_total, (uint64_t)p.ntaps, wrs, wrs * (uint64_t)p.n_total, (uint32_t)p.kb, box_n, 1);
}

static int g_num_sms = 0;
static int g_smem_optin = 0;

static size_t epi_bytes(const GemmParams& p) { return (size_t)p.epi_warps * epi_slots_per_warp(p) * kEpiSlot; }

int gemm_stage_count(const GemmParams& p, bool pair) {
  const int stage = (BM + (pair ? p.bn / 2 : p.bn)) * p.kb * 2;
  int n = (int)((g_smem_optin - 1024 - (long long)epi_bytes(p) - (long long)sizeof(GemmBarriers) - 256) / stage);
  if (n > kMaxStages) n = kMaxStages;
  return n;
}

// Kernel variant: weight-stationary (single N tile, causal taps, weights + >= 2 slabs + epilogue slots fit in shared
// memory), else slab-stream (any multi-tap conv), else the general per-stage kernel (1-tap GEMMs).
static void plan_variant(GemmOp* op, int allow) {
  const GemmParams& p = op->p;
  op->ws = 0; op->nslabs = 0; op->srp = 0; op->variant = 0; op->nstages = 0; op->slab_lo = 0; op->pair = 0;
  if (p.pack_mode) return;  // packed tiles exist in the general kernel only
  int lo = 0, hi = 0;
  for (int j = 0; j < p.ntaps; ++j) {
    if (-p.tap_off[j] > lo) lo = -p.tap_off[j];
    if (p.tap_off[j] > hi) hi = p.tap_off[j];
  }
  const long long fixed = 1024 + (long long)epi_bytes(p) + (long long)sizeof(GemmBarriers) + 256;
  if ((allow & 2) && !Q3_ENV("Q3TTS_NO_WS") && p.n_total == p.bn && p.ntaps >= 1 && hi == 0 && lo == -p.tap_off[0]) {
    const int srp = (BM + lo + 15) / 16 * 16;
    const long long w_bytes = (long long)p.ntaps * p.cin * p.bn * 2;
    const long long slab = (long long)srp * p.cin * 2;
    const long long avail = (long long)g_smem_optin - fixed - w_bytes;
    if (srp <= 256 && avail >= 2 * slab && p.ntaps * (p.cin / p.kb) <= 64) {
      int ns = (int)(avail / slab);
      if (ns > kMaxSlabs) ns = kMaxSlabs;
      op->ws = 1; op->variant = 1; op->nslabs = ns; op->srp = srp; op->slab_lo = lo;
      return;
    }
  }
  if ((allow & 4) && !Q3_ENV("Q3TTS_NO_SLAB") && p.ntaps >= 2) {
    const int srp = (BM + lo + hi + 15) / 16 * 16;
    // CTA pairs (256-row tiles, half a weight tile per SM) when a batch item's rows fill 256-row tiles well
    const int mt2 = (p.rows_valid + 2 * BM - 1) / (2 * BM);
    // ... and the tile has enough K work for the weight traffic to matter: short-K layers (the last two transposed convs) are
    // bound by their Snake epilogue and HBM, and the cross-CTA handshakes only cost them time (measured: 1.88 -> 2.03 ms, 2.97 -> 3.32 ms).
    const bool pair = !Q3_ENV("Q3TTS_NO_PAIR") && p.bn % 32 == 0 && p.rows_valid * 10 >= mt2 * 2 * BM * 9 && p.ntaps * p.cin >= 1536;
    const long long slab = (long long)srp * p.kb * 2, wst = (long long)(pair ? p.bn / 2 : p.bn) * p.kb * 2;
    const long long avail = (long long)g_smem_optin - fixed;
    op->pair = pair ? 1 : 0;
    if (srp <= 256) {
      int ns = 3;
      int nst = (int)((avail - ns * slab) / wst);
      if (nst < 4) { ns = 2; nst = (int)((avail - ns * slab) / wst); }
      if (nst > kMaxStages) nst = kMaxStages;
      if (nst >= 2) { op->variant = 2; op->nslabs = ns; op->nstages = nst; op->srp = srp; op->slab_lo = lo; }
    }
  }
}

// Fill the epilogue tensor maps from p (res / out_raw / out_snake / out_f32, strides, widths), choose the kernel
// variant and (for the weight-stationary variant) rebuild the activation map with a slab-sized box.
int gemm_make_epilogue_maps(GemmOp* op, int allow) {
  {  // packed tiles when a batch item's rows would fill less than 60 % of its 128-row tiles
    GemmParams& q = op->p;
    q.pack_mode = 0; q.pack_rshift = 0; q.pack_cpb = 1;
    const int mt = (q.rows_valid + BM - 1) / BM;
    if (q.act != ACT_TAIL && q.batches > 1 && q.rows_valid > 0 && !Q3_ENV("Q3TTS_NO_PACK") && q.rows_valid * 10 < mt * BM * 6) {
      if (q.rows_valid <= 16) {
        q.pack_mode = 1;
        while ((1 << q.pack_rshift) < q.rows_valid) ++q.pack_rshift;
      } else {
        q.pack_mode = 2;
        q.pack_cpb = (q.rows_valid + 31) / 32;
      }
    }
    // Small-M steps: a handful of 128 x bn tiles leaves most SMs idle and each CTA walks all of K alone, so shrink the N tile
    // until the grid covers the SMs (the A panel is re-read from L2 by more CTAs, which is cheap at these sizes).
    if (op->retile && q.act != ACT_TAIL && !Q3_ENV("Q3TTS_NO_RETILE")) {
      const int step = q.act == ACT_GLU ? 64 : ((q.out_raw || q.out_snake || (q.res && !q.res_f32)) ? 32 : 16);
      const int mtt = gemm_m_tiles_total(q);
      // cost model (operand-load bound): waves over the SMs x bytes a CTA loads per K block
      int best = q.bn;
      long long best_cost = -1;
      for (int bn = q.bn; bn >= step; bn -= step) {
        if (q.n_total % bn) continue;
        const long long tiles = (long long)mtt * (q.n_total / bn);
        const long long cost = ((tiles + g_num_sms - 1) / g_num_sms) * (BM + bn);
        if (best_cost < 0 || cost < best_cost) { best = bn; best_cost = cost; }
      }
      if (best != q.bn) {
        q.bn = best;
        if (make_tmap_w(&op->tmB, q, (uint32_t)q.bn)) return -3;
      }
    }
  }
  {
    GemmParams& q = op->p;
    const char* ew = Q3_ENV("Q3TTS_EPI_WARPS");
    // 12 epilogue warps (three per TMEM lane quadrant) for short-K layers whose Snake epilogue outweighs their MMAs: the transposed conv
    // 192 -> 96 2.96 -> 2.56 ms, the 1x1 convs of the C = 384 units 3.63 -> 2.85 ms (with one output slot per tensor and warp, see
    // epi_out_bufs); measured slower on the K >= 768 transposed convs, neutral on the MMA-bound conv7 layers.
    // (1x1 conv C = 768: 1.62 -> 1.41 ms with the N tile narrowed 256 -> 192 so that three operand stages still fit)
    const bool short_k = q.ntaps * q.cin <= 512 || (q.ntaps == 1 && q.cin <= 768);
    // (ConvNeXt pw1: the erf GELU epilogue outweighs K = 1024 of MMAs per tile too: 0.70 -> 0.56 ms)
    q.epi_warps = ew ? atoi(ew) : ((short_k && q.out_snake) || q.act == ACT_GELU ? 12 : 8);
    if (q.epi_warps != 4 && q.epi_warps != 8 && q.epi_warps != 12) q.epi_warps = 8;
    // 12 warps: 16-bit epilogues, and the encode path's pair epilogue with an fp32 residual but no fp32 output (its items are short
    // dependent chains: ncu had the k1 conv + residual at 37 % issue-slot use with 8 warps)
    const bool pair_res_only = epi_pair16(q) && !q.out_f32;   // (with an fp32 output as well, 12 warps measured neutral)
    if (q.epi_warps == 12 && (q.act == ACT_TAIL || q.act == ACT_GLU || q.out_f32 || (q.res_f32 && !pair_res_only) || q.pack_mode)) q.epi_warps = 8;
    q.res_depth = q.res ? (q.epi_warps == 12 ? 2 : 4) : 0;
    if (q.epi_warps == 12 && q.res && gemm_stage_count(q) < 3) {
      // the extra slots would starve the operand pipeline: try a narrower N tile (three items per warp trio), else stay at 8 warps
      const int bn0 = q.bn;
      bool ok = false;
      for (int bn = bn0 - 32; bn >= 96 && !ok; bn -= 32) {
        if (q.n_total % bn || bn % 96) continue;
        q.bn = bn;
        ok = gemm_stage_count(q) >= 3;
      }
      if (!ok) { q.bn = bn0; q.epi_warps = 8; q.res_depth = 4; }
    }
    // keep at least three pipeline stages for the general kernel: shrink the residual ring before the stage count
    if (q.res && gemm_stage_count(q) < 3) q.res_depth = 2;
  }
  const GemmParams& p = op->p;
  const int et = p.bf16 ? 1 : 0;
  const uint64_t width = p.act == ACT_GLU ? p.n_total / 2 : p.n_total;
  const uint64_t rows = p.rows_valid, nb = p.batches;
  const uint64_t rs = (uint64_t)p.out_row_stride, bs = (uint64_t)p.out_batch_stride;
  // epilogue boxes: 32 accumulator rows = one batch item's rows, or (packed mode 1) all rows of 32 >> rshift batch items
  const uint32_t br = p.pack_mode == 1 ? (1u << p.pack_rshift) : 32u, bb = p.pack_mode == 1 ? (32u >> p.pack_rshift) : 1u;
  int rc = 0;
  bool has_res = false, has_raw = false, has_snake = false, has_f32 = false;
  if (p.act != ACT_TAIL) {
    if (p.res) { has_res = true; rc |= make_tmap_3d_strided(&op->tmRes, p.res, p.res_f32 ? 2 : et, width, rows, nb, rs, p.res_batch_stride ? (uint64_t)p.res_batch_stride : bs, p.res_f32 ? 16 : 32, br, bb); }
    if (act_is_split(p.act)) {
      // the fp16 pair tensor (2 * n_total wide, its own pitches): both halves are stored through the same map; 16-column boxes inside
      // the fp32 epilogue flow, 32 in the 16-bit one
      const uint64_t prs = p.pair_row_stride ? (uint64_t)p.pair_row_stride : rs, pbs = p.pair_batch_stride ? (uint64_t)p.pair_batch_stride : bs;
      const uint32_t pbox = epi_pair16(p) ? 16 : 32;
      has_raw = has_snake = true;
      rc |= make_tmap_3d_strided(&op->tmRaw, p.out_raw, et, 2 * width, rows, nb, prs, pbs, pbox, br, bb);
      op->tmSnake = op->tmRaw;
    } else {
      if (p.out_raw) { has_raw = true; rc |= make_tmap_3d_strided(&op->tmRaw, p.out_raw, et, width, rows, nb, rs, bs, 32, br, bb); }
      if (p.out_snake) { has_snake = true; rc |= make_tmap_3d_strided(&op->tmSnake, p.out_snake, et, width, rows, nb, rs, p.snake_batch_stride ? (uint64_t)p.snake_batch_stride : bs, 32, br, bb); }
    }
    if (p.out_f32) { has_f32 = true; rc |= make_tmap_3d_strided(&op->tmF32, p.out_f32, 2, width, rows, nb, rs, bs, 16, br, bb); }
  }
  plan_variant(op, allow);
  if (op->variant == 1) op->pair = 0;
  if (op->variant == 0) {
    // general kernel on CTA pairs: long-K GEMMs whose rows fill 256-row tiles (transformer and 1x1 conv layers in bulk decode)
    const int mt2 = (p.rows_valid + 2 * BM - 1) / (2 * BM);
    op->pair = (!Q3_ENV("Q3TTS_NO_PAIR") && !Q3_ENV("Q3TTS_NO_PAIR_GEMM") && !p.pack_mode && p.act != ACT_TAIL && p.bn % 32 == 0 &&
                p.rows_valid * 10 >= mt2 * 2 * BM * 9 && p.ntaps * p.cin >= 3072) ? 1 : 0;  // measured: -32 % at K = 3072, neutral or worse at K <= 1024
  }
  // the weight map follows the plan: a CTA of a pair loads half of a weight tile
  if (make_tmap_w(&op->tmB, p, (uint32_t)(op->pair ? p.bn / 2 : p.bn))) rc |= 1;
  const uint64_t aw = p.a_cols ? (uint64_t)p.a_cols : (uint64_t)p.cin;  // width of the A tensor in memory
  const uint64_t ars = p.a_row_stride ? (uint64_t)p.a_row_stride : aw;  // row pitch (wider than aw when this launch walks one K group)
  rc |= make_tmap_3d_strided(&op->tmA, p.a_ptr, p.bf16 ? 1 : 0, aw, (uint64_t)p.rows, (uint64_t)p.batches, ars,
                             p.a_batch_stride ? (uint64_t)p.a_batch_stride : (uint64_t)p.rows * ars, (uint32_t)p.kb,
                             p.pack_mode ? br : (uint32_t)(op->variant ? op->srp : BM), bb);
  // unused maps hold a valid placeholder (the kernels take all of them as __grid_constant__ arguments)
  if (!has_res) op->tmRes = op->tmA;
  if (!has_raw) op->tmRaw = op->tmA;
  if (!has_snake) op->tmSnake = op->tmA;
  if (!has_f32) op->tmF32 = op->tmA;
  return rc;
}

int gemm_setup(int device) {
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return -1;
  g_num_sms = prop.multiProcessorCount;
  g_smem_optin = (int)prop.sharedMemPerBlockOptin;
  cudaError_t e1 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e2 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e1 == cudaSuccess) e1 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e2 == cudaSuccess) e2 = cudaFuncSetAttribute(conv_gemm_tc_kernel<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e3 = cudaFuncSetAttribute(ws_conv_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e4 = cudaFuncSetAttribute(ws_conv_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e5 = cudaFuncSetAttribute(slab_conv_kernel<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  cudaError_t e6 = cudaFuncSetAttribute(slab_conv_kernel<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e5 == cudaSuccess) e5 = cudaFuncSetAttribute(slab_conv_kernel<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e6 == cudaSuccess) e6 = cudaFuncSetAttribute(slab_conv_kernel<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, g_smem_optin);
  if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess || e4 != cudaSuccess || e5 != cudaSuccess || e6 != cudaSuccess) return -2;
  return tmap_init();
}

static bool g_pdl = true;
static thread_local bool t_pdl_scope = false;
bool pdl_enabled() { return g_pdl && t_pdl_scope; }
void pdl_set(bool on) { g_pdl = on; }
PdlScope::PdlScope(bool on) : prev(t_pdl_scope) { t_pdl_scope = on; }
PdlScope::~PdlScope() { t_pdl_scope = prev; }

int gemm_num_sms() { return g_num_sms; }
int gemm_smem_optin() { return g_smem_optin; }

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
  // epilogue slot geometry: 16-bit tensors move in 32-column boxes, GLU in 64 accumulator columns
  const bool t16 = (p.res && !p.res_f32) || p.out_raw || p.out_snake;
  if (p.res && !p.res_f32 && p.out_f32 && p.act != ACT_TAIL) return cudaErrorInvalidConfiguration;  // 16-bit residual with fp32 output: unused combination
  if (t16 && p.ncol_mod % 32) return cudaErrorInvalidConfiguration;
  if ((t16 && p.bn % 32) || (p.act == ACT_GLU && p.bn % 64) || p.bn % 16 || p.bn > 256) return cudaErrorInvalidConfiguration;
  const int total_tiles = gemm_m_tiles_total(p) * (p.n_total / p.bn);
  if (total_tiles <= 0) return cudaSuccess;
  const int grid = total_tiles < g_num_sms ? total_tiles : g_num_sms;
  const int threads = 64 + 32 * p.epi_warps;
  if (op.ws) {
    const size_t smem = (size_t)p.ntaps * p.cin * p.bn * 2 + (size_t)op.nslabs * op.srp * p.cin * 2 + epi_bytes(p) + sizeof(GemmBarriers) + 1024;
    if (p.bf16) return launch_k(ws_conv_kernel<__nv_bfloat16>, dim3(grid), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.srp);
    return launch_k(ws_conv_kernel<__half>, dim3(grid), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.srp);
  }
  if (op.variant == 2) {
    const size_t smem = (size_t)op.nslabs * op.srp * p.kb * 2 + (size_t)op.nstages * (op.pair ? p.bn / 2 : p.bn) * p.kb * 2 + epi_bytes(p) + sizeof(GemmBarriers) + 1024;
    int g2 = grid, cl = 1;
    if (op.pair) {
      const int tiles2 = p.batches * ((p.rows_valid + 2 * BM - 1) / (2 * BM)) * (p.n_total / p.bn);
      const int pairs = g_num_sms / 2;
      g2 = 2 * (tiles2 < pairs ? tiles2 : pairs); cl = 2;
    }
    if (op.pair) {
      if (p.bf16) return launch_k(slab_conv_kernel<__nv_bfloat16, true>, dim3(g2), dim3(threads), smem, stream, cl, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.nstages, op.srp, op.slab_lo);
      return launch_k(slab_conv_kernel<__half, true>, dim3(g2), dim3(threads), smem, stream, cl, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.nstages, op.srp, op.slab_lo);
    }
    if (p.bf16) return launch_k(slab_conv_kernel<__nv_bfloat16, false>, dim3(g2), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.nstages, op.srp, op.slab_lo);
    return launch_k(slab_conv_kernel<__half, false>, dim3(g2), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, op.nslabs, op.nstages, op.srp, op.slab_lo);
  }
  const int nstages = gemm_stage_count(p, op.pair != 0);
  if (nstages < 2) return cudaErrorInvalidConfiguration;
  const size_t smem = (size_t)nstages * (BM + (op.pair ? p.bn / 2 : p.bn)) * p.kb * 2 + epi_bytes(p) + sizeof(GemmBarriers) + 1024;
  if (op.pair) {
    const int tiles2 = p.batches * ((p.rows_valid + 2 * BM - 1) / (2 * BM)) * (p.n_total / p.bn);
    const int pairs = g_num_sms / 2;
    const int g2 = 2 * (tiles2 < pairs ? tiles2 : pairs);
    if (p.bf16) return launch_k(conv_gemm_tc_kernel<__nv_bfloat16, true>, dim3(g2), dim3(threads), smem, stream, 2, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, nstages);
    return launch_k(conv_gemm_tc_kernel<__half, true>, dim3(g2), dim3(threads), smem, stream, 2, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, nstages);
  }
  if (p.bf16) return launch_k(conv_gemm_tc_kernel<__nv_bfloat16, false>, dim3(grid), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, nstages);
  return launch_k(conv_gemm_tc_kernel<__half, false>, dim3(grid), dim3(threads), smem, stream, 1, op.tmA, op.tmB, op.tmRes, op.tmRaw, op.tmSnake, op.tmF32, p, nstages);
}

}  // namespace q3
