// Fused vocoder residual unit on tcgen05 tensor cores (sm_100a):
//     y = x + W1 . snake2(W7 (*) a1 + b7) + b1        (reference modeling :618-634, dilation 1 / 3 / 9)
// a1 = snake1(x) and x arrive as two 16-bit channels-last tensors; y and the next layer's Snake'd copy leave the same way.
//
// Warp roles (one CTA per SM, persistent over tiles of 128*MS rows of one batch item):
//   warp 0        TMA producer: activation slabs (tile rows + 6*dil halo rows) and weight stages (one tap of W7, or W1),
//                 each covering KS K blocks; two rings.
//   warp 1        TMEM allocator + single-thread MMA issuer:
//                   MMA1  acc1[ms] += slab(rows shifted by the tap) . W7[tap]    (A and B in shared memory)
//                   MMA2  acc2[ms]  = A2[ms] . W1                                (A in TENSOR MEMORY, B in shared memory)
//   warps 2..9    phase 1: acc1 -> + b7 -> SnakeBeta -> 16-bit pairs -> tcgen05.st into A2 (never leaves the SM)
//   warps 10..17  phase 2: acc2 -> + b1 + residual x (TMA ring) -> y (TMA store) and snake_next(y) (TMA store)
// TMEM: one stage = MS * (C fp32 acc1 + C fp32 acc2 + C/2 packed A2) = 2.5 * MS * C columns.
//   C = 96:  MS = 1, TWO stages (2 x 256 columns): MMA1 of tile i+1 runs while phase 1 of tile i converts, so the tensor
//            pipe never waits for the epilogue; the two warp groups of each phase alternate tiles (group = stage).
//   C = 192: MS = 1, one stage (480 columns): the tensor pipe idles during phase 1; the two warp groups split the columns.
// Measured (tools/mma_bench.py): a tcgen05.mma with M = 128, K = 16 occupies the pipe for >= 117.5 cycles whatever N is,
// so a C = 96 layer cannot exceed 41 % of the tensor peak in this operand arrangement; the kernel is built to sit on
// that floor (DESIGN.md "Fused residual unit").
#include "conv_gemm.h"
#include "launch.cuh"
#include "pair_ptx.cuh"
#include "tc_epilogue.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>

namespace q3 {

static constexpr int U_BM = 128;
static constexpr int kP1Warps = 8, kP2Warps = 8;
static constexpr int kUnitThreads = 64 + 32 * (kP1Warps + kP2Warps);
static constexpr int kUMaxSlabs = 6;
static constexpr int kUMaxStages = 12;
static constexpr int kURes = 2;                      // residual ring depth per phase-2 warp
static constexpr int kUOutBufs = 1;                  // output slot pairs per phase-2 warp: the TMA stores of item i drain while item i+1 is computed
static constexpr int kUSlotsPerWarp = kURes + 2 * kUOutBufs;  // residual ring + (raw, Snake'd) output slots

struct __align__(8) UnitBarriers {
  uint64_t sfull[kUMaxSlabs], sempty[kUMaxSlabs];
  uint64_t sland[kUMaxSlabs];  // XIN: a raw slab has landed in THIS CTA (its phase-1 warps then apply the input Snake in place)
  uint64_t wfull[kUMaxStages], wempty[kUMaxStages];
  uint64_t acc1_full[2], a2_full[2], acc2_full[2], acc2_empty[2];
  uint64_t res_bar[12][kURes];  // up to 12 phase-2 warps (P2X form)
  uint32_t tmem_base;
  uint32_t pad;
};

// Debug timeline (Q3TTS_DBG bit 5): clock64 stamps of CTA 0, first 16 tiles: [tile][event]
__device__ long long g_unit_trace[16 * 16];
#define UTRACE(it, ev) do { if (Q3_DBG(p, 32) && blockIdx.x == 0 && (it) < 16) g_unit_trace[(it) * 16 + (ev)] = clock64(); } while (0)

struct UnitArgs {
  const float* bias7;
  const float* s2a;
  const float* s2ib;
  const float* s1a;   // XIN: tables of the unit's input Snake (act1)
  const float* s1ib;
  int C, dil, ms, srows, nbox, nslabs, nstages;
  int ks;    // K blocks per slab / weight stage
  int nbuf;  // TMEM stages (1 or 2)
  int a_row0;  // row of the a1 tensor that belongs to output row 0 (streaming history rows in front)
};

__device__ __forceinline__ float4 lds_f4(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_u4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tma_load_3d_s(uint32_t smem_dst, const void* map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d_s(const void* map, uint32_t smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::"l"(reinterpret_cast<uint64_t>(map)),
               "r"(smem_src), "r"(c0), "r"(c1), "r"(c2)
               : "memory");
}

// Compile-time shape of one fused-unit instance.  Everything the MMA-issuing thread touches is a constant so that its code is
// straight-line: tools/mma_bench.py shows the tensor pipe retires an N = 96 MMA every 56 cycles and an N = 192 MMA every 96,
// but only if the single issuing thread spends no more than that per MMA (a passing mbarrier wait alone costs ~180 cycles).
// PAIR_ = 1: the CTA-pair form.  A 2-CTA cluster works on 256-row tiles with one M = 256 tcgen05.mma.cta_group::2 stream issued by
// the even CTA; each CTA keeps its own 128 rows (slab, tensor memory, phase-1/2 warps) and loads only HALF of every weight stage
// (its C/2 output channels).  Weight re-streaming is what bounds the single-CTA form (C = 192: 590 KB of weights per 128-row tile
// against ~43 B/clk of L2->SM fill), so halving it per SM is the point.
// XIN_ = 1: the activation operand is the RAW residual stream x; the phase-1 warps apply the input Snake to each slab in shared
// memory before its MMAs, and phase 2 stores ONE tensor (y, or snake_next(y) for the last unit of a block).  HBM traffic of a unit
// drops from four passes (a1 + x in, y + snake(y) out) to two.
// P2X_ = 1 (two TMEM stages only): 4 phase-1 warps serve every tile and 12 phase-2 warps own ONE 32-column item each (3 per TMEM lane
// quadrant), so a tile's phase 2 is one item deep instead of three.  The timeline shows the unit waits on the
// serial per-item chain of its phase-2 warps while the phase-1 warps idle 75 % of a tile period.
// P1A_ = 1 (one TMEM stage only, C = 192): the tensor pipe idles while phase 1 converts acc1, so ALL 16 epilogue warps take a 48-column
// slice of it (the phase-2 warps finish the previous tile's phase 2 well inside the conv7 MMAs and are free by then).
template <int C_, int KB_, int KS_, int NST_, int NSL_, int NBUF_, int PAIR_ = 0, int XIN_ = 0, int P2X_ = 0, int P1A_ = 0>
struct UnitCfg {
  static constexpr int C = C_, KB = KB_, KS = KS_, NST = NST_, NSL = NSL_, NBUF = NBUF_, PAIR = PAIR_, XIN = XIN_, P2X = P2X_, P1A = P1A_;
  static_assert(!P1A || (NBUF == 1 && C % 64 == 0 && !XIN && !P2X), "P1A: single TMEM stage, four 16-column-aligned slices");
  static constexpr int P1W = P2X ? 4 : kP1Warps, P2W = P2X ? 12 : kP2Warps;  // same thread count either way
  static constexpr int URES = P2X ? 1 : kURes;                               // residual ring depth per phase-2 warp
  static constexpr int USLOTS = URES + 2 * kUOutBufs;
  static_assert(!P2X || (NBUF == 2 && C == 96 && !XIN), "P2X: C = 96 with two TMEM stages");
  static constexpr int NKB = C / KB, NKG = NKB / KS, KSTEPS = KB / 16;
  static constexpr int SPAN = KB * 2;
  static constexpr int M1_STAGES = NKG * 7, M2_STAGES = NKG, TILE_STAGES = M1_STAGES + M2_STAGES;
  static constexpr uint32_t WSUB = (PAIR ? C / 2 : C) * SPAN, STAGE_BYTES = KS * WSUB;  // bytes this CTA holds
  static_assert(TILE_STAGES % NST == 0, "weight ring slots must be static per tile");
  static constexpr bool PAR_FLIPS = (TILE_STAGES / NST) % 2 != 0;  // ring parity alternates from tile to tile (else it is static too)
  static_assert(NKG % NSL == 0 || NSL % NKG == 0, "slab ring");
  static_assert(NBUF == 1 || 5 * C / 2 <= 256, "two TMEM stages need 2.5*C <= 256 columns");
};

template <typename T, typename Cfg>
__global__ void __launch_bounds__(kUnitThreads, 1)
unit_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW7, const __grid_constant__ CUtensorMap tmW1,
            const __grid_constant__ CUtensorMap tmRes, const __grid_constant__ CUtensorMap tmRaw, const __grid_constant__ CUtensorMap tmSnake,
            const GemmParams p, const UnitArgs u) {
  constexpr int C = Cfg::C, KB = Cfg::KB, KS = Cfg::KS, NST = Cfg::NST, NSL = Cfg::NSL, NBUF = Cfg::NBUF, NKG = Cfg::NKG, KSTEPS = Cfg::KSTEPS;
  constexpr bool PAIR = Cfg::PAIR != 0;
  constexpr bool XIN = Cfg::XIN != 0;
  constexpr bool P2X = Cfg::P2X != 0;
  constexpr bool P1A = Cfg::P1A != 0;
  constexpr int P1W = Cfg::P1W, P2W = Cfg::P2W, URES = Cfg::URES, USLOTS = Cfg::USLOTS;
  static_assert(!XIN || (NBUF == 2 && NKG == 1), "XIN: one slab per tile, two phase-1 groups");
  const int rank = PAIR ? (int)cluster_ctarank() : 0;               // 0 = leader (issues the MMAs of both CTAs)
  const int wid = PAIR ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;  // worker (CTA or CTA pair) index
  const int nworkers = PAIR ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr uint32_t NCTA = PAIR ? 2u : 1u;
  constexpr uint32_t span = Cfg::SPAN, wsub = Cfg::WSUB, stage_bytes = Cfg::STAGE_BYTES;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // offset on the shared array: keeps LDS / STS (a uintptr_t round trip makes every access generic)
  const int halo = 6 * u.dil;
  constexpr int TM = PAIR ? 2 * U_BM : U_BM;  // rows of a scheduler tile
  const int rrow = rank * U_BM;               // this CTA's rows inside a tile
  const uint32_t ssub = u.srows * span;  // one K block of a slab
  const uint32_t slab_bytes = KS * ssub;
  uint8_t* wst = smem;                                   // weight ring first: its slot addresses are compile-time offsets
  uint8_t* slabs = wst + (size_t)NST * stage_bytes;
  uint8_t* epi = slabs + (size_t)NSL * slab_bytes;
  float* prm = reinterpret_cast<float*>(epi + (size_t)P2W * USLOTS * kEpiSlot);  // [6][C]: b7, s2a, s2ib, b1, sna, snib
  UnitBarriers* bars = reinterpret_cast<UnitBarriers*>(prm + 8 * C);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m_tiles = (p.rows_valid + TM - 1) / TM;
  const FastDiv fdm((uint32_t)m_tiles);  // tile -> (batch item, M tile) by multiply-shift (ptx.cuh): the decode runs per tile and per residual prefetch
  const int total_tiles = p.batches * m_tiles;
  const int my_tiles = total_tiles > wid ? (total_tiles - wid + nworkers - 1) / nworkers : 0;
  constexpr int grp_warps = NBUF == 2 ? 4 : 8;  // epilogue warps that serve one TMEM stage

  if (threadIdx.x == 0) {
    tma_prefetch_desc(&tmA); tma_prefetch_desc(&tmW7); tma_prefetch_desc(&tmW1);
    tma_prefetch_desc(&tmRes); tma_prefetch_desc(&tmRaw); tma_prefetch_desc(&tmSnake);
    // XIN: sfull counts the phase-1 warps (of both CTAs) that transformed the slab; the TMA bytes are counted on each CTA's own sland
    for (int s = 0; s < NSL; ++s) { mbar_init(&bars->sfull[s], XIN ? 4 * NCTA : 1); mbar_init(&bars->sempty[s], 1); mbar_init(&bars->sland[s], 1); }
    for (int s = 0; s < NST; ++s) { mbar_init(&bars->wfull[s], 1); mbar_init(&bars->wempty[s], 1); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&bars->acc1_full[s], 1); mbar_init(&bars->a2_full[s], (P1A ? 16 : grp_warps) * NCTA);  // pair: both CTAs' warps arrive on the leader's copy
      mbar_init(&bars->acc2_full[s], 1); mbar_init(&bars->acc2_empty[s], (P2X ? P2W : grp_warps) * NCTA);
    }
    for (int w = 0; w < P2W; ++w)
      for (int s = 0; s < URES; ++s) mbar_init(&bars->res_bar[w][s], 1);
    fence_barrier_init();
  }
  for (int i = threadIdx.x; i < C; i += blockDim.x) {
    prm[i] = u.bias7[i]; prm[C + i] = u.s2a[i]; prm[2 * C + i] = u.s2ib[i];
    prm[3 * C + i] = p.bias ? p.bias[i] : 0.f;
    prm[4 * C + i] = p.snake_a ? p.snake_a[i] : 0.f; prm[5 * C + i] = p.snake_ib ? p.snake_ib[i] : 0.f;
    if (XIN) { prm[6 * C + i] = u.s1a[i]; prm[7 * C + i] = u.s1ib[i]; }
  }
  if (warp == 1) {
    if (PAIR) { tmem_alloc2(&bars->tmem_base, 512); tmem_relinquish2(); }
    else { tmem_alloc(&bars->tmem_base, 512); tmem_relinquish(); }
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the peer's barriers are initialised before anything arrives on them
  tc_fence_after();
  pdl_trigger();  // programmatic dependent launch (launch.cuh): after the TMEM allocation, see conv_gemm.cu init_common
  pdl_wait();     // the parameter tables above are weights; activations are only touched below
  const uint32_t tmem_base = bars->tmem_base;
  // TMEM column map inside one stage; stage s starts at column s * 256
  constexpr uint32_t c_acc1 = 0, c_acc2 = C, c_a2 = 2 * C;
  // P1A: phase 1 of one 48-column slice (slice cs of four) of this warp's 32 rows: acc1 -> + b7 -> SnakeBeta -> packed pairs -> A2
  auto p1_slice = [&](int cs, int jt) {
    const bool tr = (warp & 3) == 2 && lane == 0 && cs < 2;
    constexpr int SL = C / 4;
    const uint32_t lane_b = (uint32_t)((warp & 3) * 32) << 16;
    const uint32_t src = tmem_base + lane_b + c_acc1 + (uint32_t)(cs * SL);
    const uint32_t dst = tmem_base + lane_b + c_a2 + (uint32_t)(cs * SL / 2);
    const uint32_t prm_b = smem_u32(prm) + (uint32_t)(cs * SL) * 4u;
    // all loads of the slice first, one wait: the slice is latency-bound (three serial ld -> math -> st chains took as long as six)
    constexpr int NU = SL / 16;
    uint32_t ra[NU][16];
    __syncwarp();  // .sync.aligned
#pragma unroll
    for (int un = 0; un < NU; ++un) tmem_ld16(src + 16 * un, ra[un]);
    tmem_ld_wait();
    if (tr) UTRACE(jt, 13);
#pragma unroll
    for (int un = 0; un < NU; ++un) {
      const int g = 16 * un;
      uint32_t pk[8];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint32_t pa = prm_b + (uint32_t)(g + 4 * q) * 4u;
        const float4 b = lds_f4(pa), a = lds_f4(pa + (uint32_t)C * 4u), ib = lds_f4(pa + (uint32_t)C * 8u);
        const float2 v01 = snake2(add2(make_float2(__uint_as_float(ra[un][4 * q]), __uint_as_float(ra[un][4 * q + 1])), make_float2(b.x, b.y)),
                                  make_float2(a.x, a.y), make_float2(ib.x, ib.y));
        const float2 v23 = snake2(add2(make_float2(__uint_as_float(ra[un][4 * q + 2]), __uint_as_float(ra[un][4 * q + 3])), make_float2(b.z, b.w)),
                                  make_float2(a.z, a.w), make_float2(ib.z, ib.w));
        pk[2 * q] = Elem<T>::pack2(v01.x, v01.y);
        pk[2 * q + 1] = Elem<T>::pack2(v23.x, v23.y);
      }
      tmem_st8(dst + (g >> 1), pk);
    }
    if (tr) UTRACE(jt, 14);
    tmem_st_wait();
    if (tr) UTRACE(jt, 15);
    tc_fence_before();
    __syncwarp();
    if (lane == 0) { if (PAIR) mbar_arrive_leader(&bars->a2_full[0]); else mbar_arrive(&bars->a2_full[0]); }
  };

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (same order as the MMA issuer consumes)
    if (elect_one()) {
      int ss = 0, ws = 0; uint32_t sph = 0, wph = 0;
      const int boxrows = u.srows / u.nbox;
      // pair: every load lands in this CTA's shared memory but is counted on the LEADER's full barrier (it issues the MMAs);
      // the leader expects both CTAs' bytes.  The empty barriers are per CTA (multicast commit by the leader's MMA thread).
      auto load = [&](uint8_t* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
        if (PAIR) tma_load_3d_pair(smem_u32(dst), map, leader_addr(smem_u32(bar)), c0, c1, c2);
        else tma_load_3d(dst, map, bar, c0, c1, c2);
      };
      const int wrow = PAIR ? rank * (C / 2) : 0;  // this CTA's output-channel rows of a weight stage
      auto load_m1 = [&](int j) {  // slabs and W7 stages of this CTA's j-th tile
        const int tile = wid + j * nworkers;
        int b, mt; fdm.divmod(tile, b, mt);
        for (int kg = 0; kg < NKG; ++kg) {
          if (!XIN) {
            mbar_wait(&bars->sempty[ss], sph ^ 1);
            if (rank == 0) mbar_expect_tx(&bars->sfull[ss], slab_bytes * NCTA);
            for (int kk = 0; kk < KS; ++kk)
              for (int bx = 0; bx < u.nbox; ++bx)
                load(slabs + (size_t)ss * slab_bytes + (size_t)kk * ssub + (size_t)bx * boxrows * span, &tmA, &bars->sfull[ss],
                     (kg * KS + kk) * KB, mt * TM + rrow - halo + u.a_row0 + bx * boxrows, b);
            if (++ss == NSL) { ss = 0; sph ^= 1; }
          }
          for (int tap = 0; tap < 7; ++tap) {
            mbar_wait(&bars->wempty[ws], wph ^ 1);
            if (rank == 0) mbar_expect_tx(&bars->wfull[ws], stage_bytes * NCTA);
            for (int kk = 0; kk < KS; ++kk)
              load(wst + (size_t)ws * stage_bytes + (size_t)kk * wsub, &tmW7, &bars->wfull[ws], (kg * KS + kk) * KB, wrow, tap);
            if (++ws == NST) { ws = 0; wph ^= 1; }
          }
        }
      };
      // XIN: raw slabs run two tiles ahead of the weight ring (a slab has to land AND be transformed by the phase-1 warps before its
      // tile's MMAs), so they are issued on their own, not in MMA order; each lands on this CTA's own barrier.
      auto load_slab_x = [&](int j) {
        const int tile = wid + j * nworkers;
        int b, mt; fdm.divmod(tile, b, mt);
        mbar_wait(&bars->sempty[ss], sph ^ 1);
        mbar_expect_tx(&bars->sland[ss], slab_bytes);
        for (int kk = 0; kk < KS; ++kk)
          for (int bx = 0; bx < u.nbox; ++bx)
            tma_load_3d(slabs + (size_t)ss * slab_bytes + (size_t)kk * ssub + (size_t)bx * boxrows * span, &tmA, &bars->sland[ss],
                        kk * KB, mt * TM + rrow - halo + u.a_row0 + bx * boxrows, b);
        if (++ss == NSL) { ss = 0; sph ^= 1; }
      };
      auto load_m2 = [&]() {
        for (int kg = 0; kg < NKG; ++kg) {
          mbar_wait(&bars->wempty[ws], wph ^ 1);
          if (rank == 0) mbar_expect_tx(&bars->wfull[ws], stage_bytes * NCTA);
          for (int kk = 0; kk < KS; ++kk)
            load(wst + (size_t)ws * stage_bytes + (size_t)kk * wsub, &tmW1, &bars->wfull[ws], (kg * KS + kk) * KB, wrow, 0);
          if (++ws == NST) { ws = 0; wph ^= 1; }
        }
      };
      if (XIN) { if (my_tiles > 0) load_slab_x(0); if (my_tiles > 1) load_slab_x(1); }
      if (my_tiles > 0) load_m1(0);
      for (int j = 0; j < my_tiles; ++j) {
        if (XIN && j + 2 < my_tiles) load_slab_x(j + 2);
        if (NBUF == 2 && j + 1 < my_tiles) load_m1(j + 1);
        load_m2();
        if (NBUF == 1 && j + 1 < my_tiles) load_m1(j + 1);
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (one thread, straight-line per tile)
    if (rank == 0 && elect_one()) {
      const uint32_t idesc = umma_idesc(PAIR ? 2 * U_BM : U_BM, C, p.bf16 != 0);
      auto mma_ss = [&](uint32_t d, uint32_t a_lo, uint32_t b_lo, uint32_t hi_, uint32_t acc) {
        if (PAIR) umma2_f16_lohi(d, a_lo, b_lo, hi_, idesc, acc); else umma_f16_lohi(d, a_lo, b_lo, hi_, idesc, acc);
      };
      auto mma_ts = [&](uint32_t d, uint32_t a_t, uint32_t b_lo, uint32_t hi_, uint32_t acc) {
        if (PAIR) umma2_f16_ts(d, a_t, b_lo, hi_, idesc, acc); else umma_f16_ts(d, a_t, b_lo, hi_, idesc, acc);
      };
      auto commit = [&](uint64_t* bar) { if (PAIR) umma2_commit(bar); else umma_commit(bar); };  // pair: arrives in both CTAs
      const uint32_t hi = umma_desc_hi(span);
      const uint32_t w_lo0 = umma_desc_lo(smem_u32(wst));
      const uint32_t s_lo0 = umma_desc_lo(smem_u32(slabs)), s_off = slab_bytes >> 4, sk_off = ssub >> 4;
      const uint32_t tap_step = ((uint32_t)u.dil * span) >> 4;  // descriptor offset of one tap: dil rows
      // `g` = index of a weight stage inside the (prologue + steady-state) sequence; slot and parity follow from it at compile time
      // MMA1 of this CTA's j-th tile: dilated conv7 into acc1 of TMEM stage j % NBUF (free: a2_full of tile j - NBUF was waited
      // on before its MMA2, which was issued earlier).  G0 = first weight-stage index of this call.
      auto mma1 = [&](int j, auto g0_tag, uint32_t flip) {
        constexpr int G0 = decltype(g0_tag)::value;
        const uint32_t d = tmem_base + (uint32_t)((j % NBUF) * 256) + c_acc1;
        UTRACE(j, 0);
#pragma unroll
        for (int kg = 0; kg < NKG; ++kg) {
          int sl; uint32_t spar;
          if (NKG % NSL == 0) { sl = kg % NSL; spar = (uint32_t)(j * (NKG / NSL) + kg / NSL) & 1u; }
          else { constexpr int R = NSL / NKG; sl = (j % R) * NKG + kg; spar = (uint32_t)(j / R) & 1u; }
          if (XIN && PAIR) mbar_wait_cluster(&bars->sfull[sl], spar); else mbar_wait(&bars->sfull[sl], spar);
          const uint32_t slab = s_lo0 + (uint32_t)sl * s_off;
          bool ready = false;  // result of the early poll of this tap's weight stage (issued before the previous tap's MMAs)
#pragma unroll
          for (int tap = 0; tap < 7; ++tap) {
            const int g = G0 + kg * 7 + tap;          // compile-time after unrolling
            const int ws = g % NST;
            const uint32_t par = ((uint32_t)(g / NST) & 1u) ^ flip;
            if (!ready) mbar_wait(&bars->wfull[ws], par);
            // poll the next tap's stage now: the ~180-cycle latency of a passing mbarrier test then overlaps the MMA issue below
            if (tap + 1 < 7 && !Q3_DBG(p, 256)) ready = mbar_try_wait(&bars->wfull[(g + 1) % NST], ((uint32_t)((g + 1) / NST) & 1u) ^ flip);
            const uint32_t wb = w_lo0 + (uint32_t)ws * (stage_bytes >> 4);
            const uint32_t at = slab + (uint32_t)tap * tap_step;
#pragma unroll
            for (int kk = 0; kk < KS; ++kk) {
#pragma unroll
              for (int k = 0; k < KSTEPS; ++k)
                mma_ss(d, at + (uint32_t)kk * sk_off + 2 * k, wb + (uint32_t)kk * (wsub >> 4) + 2 * k, hi, (kg | tap | kk | k) != 0);
            }
            commit(&bars->wempty[ws]);
          }
          commit(&bars->sempty[sl]);
        }
        commit(&bars->acc1_full[j % NBUF]);
        UTRACE(j, 2);
      };
      // MMA2: 1x1 conv, A = Snake'd conv7 output packed in tensor memory by the phase-1 warps
      auto mma2 = [&](int j, auto g0_tag, uint32_t flip) {
        constexpr int G0 = decltype(g0_tag)::value;
        const int sb = j % NBUF;
        const uint32_t use = (uint32_t)(j / NBUF);  // how many times this TMEM stage has been used before
        const uint32_t st = tmem_base + (uint32_t)(sb * 256);
        if (PAIR) mbar_wait_cluster(&bars->a2_full[sb], use & 1); else mbar_wait_backoff(&bars->a2_full[sb], use & 1);
        UTRACE(j, 3);
        if (PAIR) mbar_wait_cluster(&bars->acc2_empty[sb], (use & 1) ^ 1); else mbar_wait_backoff(&bars->acc2_empty[sb], (use & 1) ^ 1);
        tc_fence_after();
        UTRACE(j, 4);
#pragma unroll
        for (int kg = 0; kg < NKG; ++kg) {
          const int g = G0 + kg;
          const int ws = g % NST;
          const uint32_t par = ((uint32_t)(g / NST) & 1u) ^ flip;
          mbar_wait(&bars->wfull[ws], par);
          const uint32_t wb = w_lo0 + (uint32_t)ws * (stage_bytes >> 4);
#pragma unroll
          for (int kk = 0; kk < KS; ++kk) {
#pragma unroll
            for (int k = 0; k < KSTEPS; ++k)
              mma_ts(st + c_acc2, st + c_a2 + (uint32_t)((kg * KS + kk) * (KB / 2)) + 8 * k, wb + (uint32_t)kk * (wsub >> 4) + 2 * k, hi, (kg | kk | k) != 0);
          }
          commit(&bars->wempty[ws]);
        }
        commit(&bars->acc2_full[sb]);
        UTRACE(j, 5);
      };
      // Weight-stage indices: prologue MMA1(0) uses [0, M1); steady state per tile is TILE_STAGES stages, so the index modulo
      // (2 * NST) -- all the slot / parity arithmetic needs -- is the same in every iteration (TILE_STAGES % (2*NST) == 0).
      constexpr int M1S = Cfg::M1_STAGES;
      if (my_tiles > 0) mma1(0, std::integral_constant<int, 0>{}, 0u);
      for (int j = 0; j < my_tiles; ++j) {
        const uint32_t flip = Cfg::PAR_FLIPS ? (uint32_t)(j & 1) : 0u;  // global stage index = j * TILE_STAGES + (index inside the iteration)
        if (NBUF == 2) {
          if (j + 1 < my_tiles) { mma1(j + 1, std::integral_constant<int, M1S>{}, flip); mma2(j, std::integral_constant<int, 2 * M1S>{}, flip); }
          else mma2(j, std::integral_constant<int, M1S>{}, flip);
        } else {
          mma2(j, std::integral_constant<int, M1S>{}, flip);
          if (j + 1 < my_tiles) mma1(j + 1, std::integral_constant<int, M1S + Cfg::M2_STAGES>{}, flip);
        }
      }
    }
  } else if (warp < 2 + P1W) {
    // ------------------------------------------------------------------ phase-1 warps: acc1 -> + b7 -> SnakeBeta -> 16-bit pairs -> A2 (TMEM)
    const int quad = warp & 3, grp = (warp - 2) >> 2;
    constexpr int ncols = NBUF == 2 ? C : C / 2;  // columns this warp serves (a multiple of 32)
    const uint32_t tcol = NBUF == 2 ? 0u : (uint32_t)(grp * ncols);  // one TMEM stage: the two warp groups split the columns
    const int pcol0 = NBUF == 2 ? 0 : grp * ncols;
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t prm_s = smem_u32(prm);
    // XIN: input Snake of tile j's raw slab, in place, by the four warps of the group that owns the tile.  SWIZZLE_64B slab rows:
    // the 16-byte chunk at physical position pch of row r holds channels 8 * (pch ^ ((r >> 1) & 3)) .. + 7 of its K block, and a
    // thread walks rows r0, r0 + 32, ... so its channels (and their Snake tables) are fixed.  Zero rows (TMA fill outside the item)
    // stay zero: snake(0) = 0.
    auto transform = [&](int j) {
      static_assert(!XIN || KB == 32, "XIN transform is written for 64-byte slab rows");
      const int sl = j % NSL;
      if (quad == 2 && lane == 0) UTRACE(j, 12);
      mbar_wait(&bars->sland[sl], (uint32_t)(j / NSL) & 1u);
      if (quad == 2 && lane == 0) UTRACE(j, 10);
      const int tid = ((warp - 2) & 3) * 32 + lane;
      const int pch = tid & 3, r0 = tid >> 2;
      const int lc = pch ^ ((r0 >> 1) & 3);
      const uint32_t sbase = smem_u32(slabs) + (uint32_t)sl * slab_bytes + (uint32_t)pch * 16u;
#pragma unroll
      for (int kk = 0; kk < KS; ++kk) {
        const uint32_t pa = prm_s + (uint32_t)(6 * C + kk * KB + lc * 8) * 4u;
        const float4 a0 = lds_f4(pa), a1 = lds_f4(pa + 16u), h0 = lds_f4(pa + (uint32_t)C * 4u), h1 = lds_f4(pa + (uint32_t)C * 4u + 16u);
        auto chunk = [&](uint4 v) {
          const float2 x0 = Elem<T>::unpack2(v.x), x1 = Elem<T>::unpack2(v.y), x2 = Elem<T>::unpack2(v.z), x3 = Elem<T>::unpack2(v.w);
          const float2 y0 = snake2(x0, make_float2(a0.x, a0.y), make_float2(h0.x, h0.y));
          const float2 y1 = snake2(x1, make_float2(a0.z, a0.w), make_float2(h0.z, h0.w));
          const float2 y2 = snake2(x2, make_float2(a1.x, a1.y), make_float2(h1.x, h1.y));
          const float2 y3 = snake2(x3, make_float2(a1.z, a1.w), make_float2(h1.z, h1.w));
          return make_uint4(Elem<T>::pack2(y0.x, y0.y), Elem<T>::pack2(y1.x, y1.y), Elem<T>::pack2(y2.x, y2.y), Elem<T>::pack2(y3.x, y3.y));
        };
        const uint32_t kbase = sbase + (uint32_t)kk * ssub;
        int r = r0;
        for (; r + 96 < u.srows; r += 128) {  // four rows per pass: independent chains hide the ld.shared / MUFU latencies
          const uint32_t ad = kbase + (uint32_t)r * 64u;
          const uint4 v0 = lds_u4(ad), v1 = lds_u4(ad + 2048u), v2 = lds_u4(ad + 4096u), v3 = lds_u4(ad + 6144u);
          const uint4 w0 = chunk(v0), w1 = chunk(v1), w2 = chunk(v2), w3 = chunk(v3);
          sts_u4(ad, w0.x, w0.y, w0.z, w0.w); sts_u4(ad + 2048u, w1.x, w1.y, w1.z, w1.w);
          sts_u4(ad + 4096u, w2.x, w2.y, w2.z, w2.w); sts_u4(ad + 6144u, w3.x, w3.y, w3.z, w3.w);
        }
        for (; r < u.srows; r += 32) {
          const uint32_t ad = kbase + (uint32_t)r * 64u;
          const uint4 w = chunk(lds_u4(ad));
          sts_u4(ad, w.x, w.y, w.z, w.w);
        }
      }
      fence_proxy_async_smem();  // the tensor core reads the slab through the asynchronous proxy
      __syncwarp();
      if (lane == 0) { if (PAIR) mbar_arrive_leader(&bars->sfull[sl]); else mbar_arrive(&bars->sfull[sl]); }
      if (quad == 2 && lane == 0) UTRACE(j, 11);
    };
    if (XIN && grp < my_tiles) transform(grp);
    for (int j = (NBUF == 2 && !P2X) ? grp : 0; j < my_tiles; j += (P2X ? 1 : NBUF)) {
      const int sb = NBUF == 2 ? (P2X ? (j & 1) : grp) : 0;  // TMEM stage of this tile
      const uint32_t st = tmem_base + lane_base + (uint32_t)(sb * 256);
      const uint32_t src = st + c_acc1 + tcol;
      const uint32_t dst = st + c_a2 + (tcol >> 1);
      // the group's NEXT tile first: its slab lands while this tile's conv7 MMAs run, which is when these warps are idle
      if (XIN && j + NBUF < my_tiles) transform(j + NBUF);
      mbar_wait(&bars->acc1_full[sb], (uint32_t)(j / NBUF) & 1);
      tc_fence_after();
      if (quad == 2 && lane == 0) UTRACE(j, 6);
      if (P1A) {
        p1_slice(grp, j);  // slices 0, 1 (the phase-2 warps take 2, 3)
        if (quad == 2 && lane == 0) UTRACE(j, 7);
        continue;
      }
      for (int g = 0; g < ncols; g += 32) {
        uint32_t ra[16], rb[16];
        __syncwarp();  // .sync.aligned: reconverge after the barrier spin loop
        tmem_ld16(src + g, ra);
        tmem_ld16(src + g + 16, rb);
        tmem_ld_wait();
        uint32_t pk[16];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const uint32_t pa = prm_s + (uint32_t)(pcol0 + g + 4 * q) * 4u;
          const float4 b = lds_f4(pa), a = lds_f4(pa + (uint32_t)C * 4u), ib = lds_f4(pa + (uint32_t)C * 8u);
          const uint32_t* r = q < 4 ? &ra[4 * q] : &rb[4 * (q - 4)];
          const float2 v01 = snake2(add2(make_float2(__uint_as_float(r[0]), __uint_as_float(r[1])), make_float2(b.x, b.y)),
                                    make_float2(a.x, a.y), make_float2(ib.x, ib.y));
          const float2 v23 = snake2(add2(make_float2(__uint_as_float(r[2]), __uint_as_float(r[3])), make_float2(b.z, b.w)),
                                    make_float2(a.z, a.w), make_float2(ib.z, ib.w));
          pk[2 * q] = Elem<T>::pack2(v01.x, v01.y);
          pk[2 * q + 1] = Elem<T>::pack2(v23.x, v23.y);
        }
        const uint32_t (&p0)[8] = *reinterpret_cast<const uint32_t (*)[8]>(&pk[0]);
        const uint32_t (&p1)[8] = *reinterpret_cast<const uint32_t (*)[8]>(&pk[8]);
        tmem_st8(dst + (g >> 1), p0);
        tmem_st8(dst + (g >> 1) + 8, p1);
      }
      tmem_st_wait();
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (PAIR) mbar_arrive_leader(&bars->a2_full[sb]); else mbar_arrive(&bars->a2_full[sb]); }
      if (quad == 2 && lane == 0) UTRACE(j, 7);
    }
  } else {
    // ------------------------------------------------------------------ phase-2 warps: acc2 + b1 + x -> y, snake_next(y) (TMA in / out)
    const int pw = warp - 2 - P1W;
    const int quad = warp & 3, grp = pw >> 2;  // P2X: grp = the item (32 columns) this warp owns in every tile
    constexpr int ncols = NBUF == 2 ? C : C / 2;
    constexpr int nitems = P2X ? 1 : ncols >> 5;
    const int item0 = P2X ? grp : 0;
    const uint32_t tcol = NBUF == 2 ? 0u : (uint32_t)(grp * ncols);
    const int pcol0 = NBUF == 2 ? 0 : grp * ncols;
    const int rsub = rrow + quad * 32;  // row offset of this warp inside a tile
    const uint32_t lane_base = (uint32_t)(quad * 32) << 16;
    const uint32_t prm_s = smem_u32(prm);
    const uint32_t eb = smem_u32(epi) + (uint32_t)pw * USLOTS * kEpiSlot;  // [URES residual][raw][snake] slots of 2 KB
    const uint32_t rbar = smem_u32(&bars->res_bar[pw][0]);
    const int j0 = (NBUF == 2 && !P2X) ? grp : 0;
    constexpr int jstep = P2X ? 1 : NBUF;
    const bool wr_raw = p.skip_raw == 0;  // the last unit of a block stores only snake_next(y)
    const bool wr_snk = XIN ? p.skip_raw != 0 : true;  // XIN: exactly one tensor leaves (the next unit applies its own input Snake)
    uint32_t off[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) off[c] = (uint32_t)(lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4));
    // residual prefetch cursor over this warp's (tile, item) sequence
    int pf_j = j0, pf_item = 0;
    uint32_t issued = 0, cons = 0, ob = 0;
    auto prefetch = [&]() {
      if (pf_j >= my_tiles || Q3_DBG(p, 128)) return;
      if (lane == 0) {
        const int tile = wid + pf_j * nworkers;
        int b, mt; fdm.divmod(tile, b, mt);
        const uint32_t slot = issued & (URES - 1);
        const uint32_t bar = rbar + slot * 8u;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"((uint32_t)kEpiSlot) : "memory");
        tma_load_3d_s(eb + slot * kEpiSlot, &tmRes, bar, pcol0 + 32 * (item0 + pf_item), mt * TM + rsub, b);
      }
      ++issued;
      if (++pf_item == nitems) { pf_item = 0; pf_j += jstep; }
    };
    for (int i = 0; i < URES; ++i) prefetch();
    for (int j = j0; j < my_tiles; j += jstep) {
      const int sb = NBUF == 2 ? (P2X ? (j & 1) : grp) : 0;
      const uint32_t acc = tmem_base + lane_base + (uint32_t)(sb * 256) + c_acc2 + tcol + (uint32_t)(32 * item0);
      const int tile = wid + j * nworkers;
      int b, mt; fdm.divmod(tile, b, mt);
      const int row0 = mt * TM + rsub;
      if (P1A) {  // this warp's slice of phase 1 first: the tensor pipe waits for it
        mbar_wait(&bars->acc1_full[0], (uint32_t)j & 1);
        tc_fence_after();
        p1_slice(2 + grp, j);
      }
      mbar_wait(&bars->acc2_full[sb], (uint32_t)(j / NBUF) & 1);
      tc_fence_after();
      if (quad == 2 && lane == 0) UTRACE(j, 8);
      for (int item = 0; item < nitems; ++item) {
        const uint32_t slot = cons & (URES - 1);
        if (!Q3_DBG(p, 128)) mbar_wait(&bars->res_bar[pw][slot], (cons / URES) & 1);  // debug bit 7: phase 2 without residual loads
        const uint32_t rs = eb + slot * kEpiSlot;
        uint4 res[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) res[c] = lds_u4(rs + off[c]);
        ++cons;
        const uint32_t raw_s = eb + (URES + 2 * (ob & (kUOutBufs - 1))) * kEpiSlot, snk_s = raw_s + kEpiSlot;
        // asynchronous TMEM loads and their wait stay adjacent: nothing may touch the destination registers in between
        uint32_t ra[16], rb[16];
        __syncwarp();  // .sync.aligned: reconverge after the lane-0 branch in prefetch()
        tmem_ld16(acc + 32 * item, ra);
        tmem_ld16(acc + 32 * item + 16, rb);
        tmem_ld_wait();
        uint32_t oraw[16], osnk[16];
#pragma unroll
        for (int c = 0; c < 4; ++c) {  // chunk c = 8 columns
          const uint32_t* r = c < 2 ? &ra[8 * c] : &rb[8 * (c - 2)];
          const uint32_t rw[4] = {res[c].x, res[c].y, res[c].z, res[c].w};
#pragma unroll
          for (int q = 0; q < 2; ++q) {
            const uint32_t pa = prm_s + (uint32_t)(3 * C + pcol0 + 32 * (item0 + item) + 8 * c + 4 * q) * 4u;
            const float4 bb = lds_f4(pa), a = lds_f4(pa + (uint32_t)C * 4u), ib = lds_f4(pa + (uint32_t)C * 8u);
            const float2 x0 = Elem<T>::unpack2(rw[2 * q]), x1 = Elem<T>::unpack2(rw[2 * q + 1]);
            const float2 y01 = add2(add2(make_float2(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1])), make_float2(bb.x, bb.y)), x0);
            const float2 y23 = add2(add2(make_float2(__uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3])), make_float2(bb.z, bb.w)), x1);
            oraw[4 * c + 2 * q] = Elem<T>::pack2(y01.x, y01.y);
            oraw[4 * c + 2 * q + 1] = Elem<T>::pack2(y23.x, y23.y);
            if (wr_snk) {
              const float2 s01 = snake2(y01, make_float2(a.x, a.y), make_float2(ib.x, ib.y));
              const float2 s23 = snake2(y23, make_float2(a.z, a.w), make_float2(ib.z, ib.w));
              osnk[4 * c + 2 * q] = Elem<T>::pack2(s01.x, s01.y);
              osnk[4 * c + 2 * q + 1] = Elem<T>::pack2(s23.x, s23.y);
            }
          }
        }
        const bool do_store = !Q3_DBG(p, 64);  // debug: phase 2 without its output stores
        if (do_store) {
          if (lane == 0) tma_store_wait_read<kUOutBufs - 1>();  // the stores that last used this slot pair have read it
          __syncwarp();
        }
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          if (wr_raw) sts_u4(raw_s + off[c], oraw[4 * c], oraw[4 * c + 1], oraw[4 * c + 2], oraw[4 * c + 3]);
          if (wr_snk) sts_u4(snk_s + off[c], osnk[4 * c], osnk[4 * c + 1], osnk[4 * c + 2], osnk[4 * c + 3]);
        }
        if (do_store) {
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (wr_raw) tma_store_3d_s(&tmRaw, raw_s, pcol0 + 32 * (item0 + item), row0, b);
            if (wr_snk) tma_store_3d_s(&tmSnake, snk_s, pcol0 + 32 * (item0 + item), row0, b);
            tma_store_commit();
          }
        }
        // refill the residual slot only now: the st.shared above consumed the residual registers (in-order issue), so every
        // ld.shared of the slot has returned and the __syncwarp() above covers all lanes
        prefetch();
        ++ob;
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (PAIR) mbar_arrive_leader(&bars->acc2_empty[sb]); else mbar_arrive(&bars->acc2_empty[sb]); }
      if (quad == 2 && lane == 0) UTRACE(j, 9);
    }
    if (lane == 0) tma_store_wait<0>();
  }
  tc_fence_before();
  __syncthreads();
  if (PAIR) cluster_sync_all();  // the leader's MMAs read the peer's shared and tensor memory until its last commit
  if (warp == 1) {
    tc_fence_after();
    if (PAIR) tmem_dealloc2(tmem_base, 512); else tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------- host
using Cfg96 = UnitCfg<96, 32, 3, 4, 2, 2>;       // one weight stage = one tap over all of K (18 KB), whole-tile slabs, two TMEM stages
using Cfg192a = UnitCfg<192, 64, 1, 4, 3, 1>;    // stage = (K block, tap) 24 KB; 4 stages when the slabs are small (dilation 1, 3)
using Cfg192b = UnitCfg<192, 64, 1, 3, 3, 1>;    // 3 stages for the dilation-9 slabs
// CTA-pair forms (variants 3..5): half-size weight stages (9 KB / 12 KB), so the ring can be deeper
using Cfg96p = UnitCfg<96, 32, 3, 8, 2, 2, 1>;
using Cfg192pa = UnitCfg<192, 64, 1, 6, 3, 1, 1>;
using Cfg192pb = UnitCfg<192, 64, 1, 4, 3, 1, 1>;
// raw-input form (variant 6): CTA pairs, three slabs in flight (a slab must land AND be transformed before its tile's MMAs), 4 weight stages
using Cfg96x = UnitCfg<96, 32, 3, 4, 3, 2, 1, 1>;
// phase 2 by item (variant 7): CTA pairs, 4 phase-1 + 12 phase-2 warps
using Cfg96q = UnitCfg<96, 32, 3, 8, 2, 2, 1, 0, 1>;
// phase 1 on all 16 epilogue warps (variants 8, 9): CTA pairs, C = 192
using Cfg192qa = UnitCfg<192, 64, 1, 6, 3, 1, 1, 0, 0, 1>;
using Cfg192qb = UnitCfg<192, 64, 1, 4, 3, 1, 1, 0, 0, 1>;

static size_t unit_smem_bytes(const UnitOp& op) {
  const size_t span = (size_t)op.p.kb * 2;
  return (size_t)op.nslabs * op.ks * op.srows * span + (size_t)op.nstages * op.ks * (op.pair ? op.C / 2 : op.C) * span + (size_t)(op.variant == 7 ? Cfg96q::P2W * Cfg96q::USLOTS : kP2Warps * kUSlotsPerWarp) * kEpiSlot +
         (size_t)8 * op.C * sizeof(float) + sizeof(UnitBarriers) + 1024;
}

// Opt-in experiment (Q3TTS_XIN=1), off by default: measured SLOWER than the two-tensor form on B200 (C = 96 units 8.9 -> 10.4-11.4 ms
// per 64 x 10 s step, transposed conv 192 -> 96 2.8 -> 2.1 ms).  The unit is bound by its phase-2 warps, not by HBM, and
// the in-kernel Snake adds CUDA-core work to the same schedulers.  The streaming plans keep the two-tensor form, so streaming output is
// not bit-identical to bulk output while the switch is on.
bool unit_xin_available(int C) {
  static const bool on = Q3_ENV("Q3TTS_XIN") != nullptr && Q3_ENV("Q3TTS_XIN")[0] == '1';
  return on && C == 96 && !Q3_ENV("Q3TTS_NO_FUSE") && !Q3_ENV("Q3TTS_NO_PAIR");
}

bool unit_fusable(int C) {
  if (Q3_ENV("Q3TTS_NO_FUSE")) return false;
  return C == 96 || C == 192;  // instantiated shapes (the vocoder's two HBM-bound levels)
}

int unit_make(UnitOp* op, const void* a1, const void* w7, const void* w1, int batches, int rows) {
  GemmParams& p = op->p;
  op->enabled = 0;
  const int C = op->C;
  if (!unit_fusable(C)) return 1;
  op->ms = 1;
  p.kb = C == 96 ? 32 : 64;
  p.epi_warps = 8;
  p.res_depth = 2;
  p.batches = batches; p.rows = rows; p.rows_valid = rows; p.cin = C; p.ntaps = 1; p.tap_off[0] = 0;
  p.n_total = C; p.bn = C; p.ncol_mod = C; p.act = ACT_NONE; p.res_f32 = 0; p.out_f32 = nullptr; p.scale = nullptr;
  p.out_row_stride = C; p.out_batch_stride = (long long)rows * C;
  p.a_ptr = a1;
  if (op->xin) {
    if (!unit_xin_available(C) || !p.res || p.res != a1 || !p.out_raw || p.out_raw == p.res || !op->snake1_a || !op->snake1_ib) return 1;
    if (p.out_snake && (!p.snake_a || !p.snake_ib)) return 1;
  } else if (!p.res || !p.out_raw || !p.out_snake || !p.snake_a || !p.snake_ib) return 1;
  const int span = p.kb * 2;
  const int need = U_BM + 6 * op->dil;
  op->srows = (need + 7) / 8 * 8;
  op->nbox = 1;
  if (op->srows > 256) { if (Q3_ENV("Q3TTS_VERBOSE")) fprintf(stderr, "unit_make: srows %d\n", op->srows); return 1; }
  op->pair = Q3_ENV("Q3TTS_NO_PAIR") ? 0 : 1;
  if (op->xin) { op->pair = 1; op->ks = Cfg96x::KS; op->nstages = Cfg96x::NST; op->nslabs = Cfg96x::NSL; op->nbuf = Cfg96x::NBUF; op->variant = 6; }
  else if (C == 96) {
    op->ks = Cfg96::KS; op->nstages = op->pair ? Cfg96p::NST : Cfg96::NST; op->nslabs = Cfg96::NSL; op->nbuf = Cfg96::NBUF; op->variant = op->pair ? 3 : 0;
    if (op->pair && !Q3_ENV("Q3TTS_NO_P2X")) op->variant = 7;
  }
  else if (op->pair) {
    op->ks = 1; op->nslabs = 3; op->nbuf = 1;
    op->nstages = Cfg192pa::NST; op->variant = 4;
    if (unit_smem_bytes(*op) > (size_t)gemm_smem_optin()) { op->nstages = Cfg192pb::NST; op->variant = 5; }
    if (!Q3_ENV("Q3TTS_NO_P1A")) op->variant += 4;  // 8 / 9: same rings, phase 1 on all 16 epilogue warps
  } else {
    op->ks = 1; op->nslabs = 3; op->nbuf = 1;
    op->nstages = 4; op->variant = 1;
    if (unit_smem_bytes(*op) > (size_t)gemm_smem_optin()) { op->nstages = 3; op->variant = 2; }
  }
  if (unit_smem_bytes(*op) > (size_t)gemm_smem_optin()) {
    if (Q3_ENV("Q3TTS_VERBOSE")) fprintf(stderr, "unit_make: smem %zu > %d (C %d dil %d srows %d)\n", unit_smem_bytes(*op), gemm_smem_optin(), C, op->dil, op->srows);
    return 1;
  }
  const int et = p.bf16 ? 1 : 0;
  int rc = 0;
  const uint64_t a_rows = op->a_rows ? (uint64_t)op->a_rows : (uint64_t)rows;
  rc |= make_tmap_3d_strided(&op->tmA, a1, et, (uint64_t)C, a_rows, (uint64_t)batches, (uint64_t)C,
                             op->a_batch_stride ? (uint64_t)op->a_batch_stride : a_rows * C, (uint32_t)p.kb, (uint32_t)(op->srows / op->nbox));
  const uint32_t wbox = (uint32_t)(op->pair ? C / 2 : C);  // pair: each CTA loads its half of the output channels
  rc |= make_tmap_3d(&op->tmW7, w7, (uint64_t)C, (uint64_t)C, 7, (uint32_t)p.kb, wbox, p.bf16 != 0);
  rc |= make_tmap_3d(&op->tmW1, w1, (uint64_t)C, (uint64_t)C, 1, (uint32_t)p.kb, wbox, p.bf16 != 0);
  const uint64_t rs = (uint64_t)p.out_row_stride, bs = (uint64_t)p.out_batch_stride;
  rc |= make_tmap_3d_strided(&op->tmRes, p.res, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs, bs, 32, 32);
  rc |= make_tmap_3d_strided(&op->tmRaw, p.out_raw, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs, bs, 32, 32);
  if (p.out_snake)
    rc |= make_tmap_3d_strided(&op->tmSnake, p.out_snake, et, (uint64_t)C, (uint64_t)rows, (uint64_t)batches, rs,
                               op->snake_batch_stride ? (uint64_t)op->snake_batch_stride : bs, 32, 32);
  else op->tmSnake = op->tmRaw;  // xin, not the last unit of its block: never stored through
  if (rc) { if (Q3_ENV("Q3TTS_VERBOSE")) fprintf(stderr, "unit_make: tensor map failed (C %d dil %d srows %d)\n", C, op->dil, op->srows); return -1; }
  op->enabled = 1;
  return 0;
}


// Opt-in to the full dynamic shared memory for every instance.  Function attributes belong to the device the call is made on, so this
// runs once per handle (engine_finalize, next to gemm_setup), not once per process.
int unit_setup() {
  const int mx = gemm_smem_optin();
  cudaError_t e = cudaSuccess;
  auto set = [&](const void* f) { if (e == cudaSuccess) e = cudaFuncSetAttribute(f, cudaFuncAttributeMaxDynamicSharedMemorySize, mx); };
  set((const void*)unit_kernel<__half, Cfg96>); set((const void*)unit_kernel<__nv_bfloat16, Cfg96>);
  set((const void*)unit_kernel<__half, Cfg192a>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192a>);
  set((const void*)unit_kernel<__half, Cfg192b>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192b>);
  set((const void*)unit_kernel<__half, Cfg96p>); set((const void*)unit_kernel<__nv_bfloat16, Cfg96p>);
  set((const void*)unit_kernel<__half, Cfg192pa>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192pa>);
  set((const void*)unit_kernel<__half, Cfg192pb>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192pb>);
  set((const void*)unit_kernel<__half, Cfg96x>); set((const void*)unit_kernel<__nv_bfloat16, Cfg96x>);
  set((const void*)unit_kernel<__half, Cfg96q>); set((const void*)unit_kernel<__nv_bfloat16, Cfg96q>);
  set((const void*)unit_kernel<__half, Cfg192qa>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192qa>);
  set((const void*)unit_kernel<__half, Cfg192qb>); set((const void*)unit_kernel<__nv_bfloat16, Cfg192qb>);
  return e == cudaSuccess ? 0 : -2;
}

// Copies the debug timeline of the last traced launch (Q3TTS_DBG bit 5) to the host.
int unit_trace_read(long long* out256) { return cudaMemcpyFromSymbol(out256, g_unit_trace, sizeof(long long) * 256) == cudaSuccess ? 0 : -1; }

template <typename T, typename Cfg>
static cudaError_t launch_one(const UnitOp& op, const UnitArgs& u, int grid, size_t smem, cudaStream_t stream) {
  return launch_k(unit_kernel<T, Cfg>, dim3((unsigned)grid), dim3(kUnitThreads), smem, stream, Cfg::PAIR ? 2 : 1, op.tmA, op.tmW7, op.tmW1, op.tmRes, op.tmRaw,
                  op.tmSnake, op.p, u);
}
template <typename Cfg>
static cudaError_t launch_cfg(const UnitOp& op, const UnitArgs& u, int grid, size_t smem, cudaStream_t stream) {
  return op.p.bf16 ? launch_one<__nv_bfloat16, Cfg>(op, u, grid, smem, stream) : launch_one<__half, Cfg>(op, u, grid, smem, stream);
}

cudaError_t launch_unit(const UnitOp& op, cudaStream_t stream) {
  if (!op.enabled) return cudaErrorInvalidConfiguration;
  const GemmParams& p = op.p;
  const int tm = op.pair ? 2 * U_BM : U_BM;
  const int tiles = p.batches * ((p.rows_valid + tm - 1) / tm);
  if (tiles <= 0) return cudaSuccess;
  int grid;
  if (op.pair) { const int pairs = gemm_num_sms() / 2; grid = 2 * (tiles < pairs ? tiles : pairs); }
  else grid = tiles < gemm_num_sms() ? tiles : gemm_num_sms();
  UnitArgs u{op.bias7, op.snake2_a, op.snake2_ib, op.snake1_a, op.snake1_ib, op.C, op.dil, op.ms, op.srows, op.nbox, op.nslabs, op.nstages, op.ks, op.nbuf, op.a_row0};
  const size_t smem = unit_smem_bytes(op);
  switch (op.variant) {
    case 0: return launch_cfg<Cfg96>(op, u, grid, smem, stream);
    case 1: return launch_cfg<Cfg192a>(op, u, grid, smem, stream);
    case 2: return launch_cfg<Cfg192b>(op, u, grid, smem, stream);
    case 3: return launch_cfg<Cfg96p>(op, u, grid, smem, stream);
    case 4: return launch_cfg<Cfg192pa>(op, u, grid, smem, stream);
    case 6: return launch_cfg<Cfg96x>(op, u, grid, smem, stream);
    case 7: return launch_cfg<Cfg96q>(op, u, grid, smem, stream);
    case 8: return launch_cfg<Cfg192qa>(op, u, grid, smem, stream);
    case 9: return launch_cfg<Cfg192qb>(op, u, grid, smem, stream);
    default: return launch_cfg<Cfg192pb>(op, u, grid, smem, stream);
  }
}

}  // namespace q3
