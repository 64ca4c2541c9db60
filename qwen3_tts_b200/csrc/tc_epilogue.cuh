// Asynchronous per-warp epilogue shared by the tcgen05 kernels.
//
// 4 or 8 epilogue warps.  A warp may only read its TMEM lane quadrant (warp % 4), i.e. 32 accumulator
// rows; with 8 warps the two warps of a quadrant split a tile's columns item by item.  A work item is
// 32 accumulator columns (two independent 16-column halves, for instruction-level parallelism), 64 for
// the GLU epilogue, 16 when an fp32 tensor is involved.  Every warp has a private ring of 2 KB
// shared-memory slots (32 rows x 64 bytes, SWIZZLE_64B) and private mbarriers: residual tiles arrive by
// TMA, prefetched `res_depth` slots ahead across tile boundaries; results are staged in slots and leave
// by TMA store.  Nothing blocks on a global load or store; rows/columns past the end of a tensor are
// clipped by the tensor maps.
#pragma once
#include "epilogue.cuh"

namespace q3 {

static constexpr int kEpiSlot = 2048;  // 32 rows x 64 bytes
static constexpr int kMaxEpiWarps = 12;
static constexpr int kMaxResDepth = 8;

__host__ __device__ inline int epi_res_depth(const GemmParams& p) { return p.res ? (p.res_depth > 0 ? p.res_depth : 4) : 0; }
// output slots per 16-bit output tensor and warp: 2 (the stores of item i drain while item i+1 is computed); 1 with 12 epilogue warps
// and a residual ring, so that the slots of 12 warps still leave room for three operand stages
__host__ __device__ inline int epi_out_bufs(const GemmParams& p) { return (p.epi_warps == 12 && p.res && p.act != ACT_GLU) ? 1 : 2; }
// slots per warp for a given epilogue configuration
__host__ __device__ inline int epi_slots_per_warp(const GemmParams& p) {
  if (p.act == ACT_TAIL) return 0;
  int n = epi_res_depth(p);
  if (p.act == ACT_GLU) return n + 2;
  if (epi_pair16(p)) return n + 2 + (p.out_f32 ? 2 : 0);   // pair ring and fp32 ring, two slots each (one store group per item)
  if (p.out_raw) n += epi_out_bufs(p);
  if (p.out_snake) n += epi_out_bufs(p);
  if (p.out_f32) n += 4;
  return n;
}

// 16-byte chunk `c` of row `lane` inside a 32-row x 64-byte SWIZZLE_64B slot
__device__ __forceinline__ uint32_t slot_off(int lane, int c) { return (uint32_t)(lane * 64 + ((c ^ ((lane >> 1) & 3)) << 4)); }

struct TileSched {
  int m_tiles, n_tiles, total_tiles;
  FastDiv fn, fm;  // division by n_tiles / m_tiles as multiply-shift (set by init())
  __device__ __forceinline__ void init() { fn = FastDiv((uint32_t)n_tiles); fm = FastDiv((uint32_t)m_tiles); }
  __device__ __forceinline__ void decode(int tile, int& nt, int& mt, int& b) const {  // 32-bit: total_tiles < 2^31
    int q = tile;
    nt = 0;
    if (n_tiles > 1) fn.divmod(tile, q, nt);
    fm.divmod(q, b, mt);
  }
};

template <typename T>
struct EpiWarp {
  const GemmParams& p;
  const CUtensorMap* tmRes; const CUtensorMap* tmRaw; const CUtensorMap* tmSnake; const CUtensorMap* tmF32;
  uint8_t* eb;        // this warp's slots
  uint64_t* rbar;     // this warp's residual barriers
  TileSched ts;
  int lane, quad, half, nsplit;
  bool has_res, rf32, glu, tail, of32, o16, pair16;
  int units, item_units, nitems, rd, rd_shift;
  bool wraps;  // a tile's columns can wrap around ncol_mod (transposed convs: n_total = stride * cout)
  int tile_c0 = 0;  // (first column of the current tile) % ncol_mod: one division per tile instead of one per item
  int s_raw, s_snake, s_f32;  // first slot of each output ring
  int obufs = 2;              // slots per 16-bit output ring (item_pair)
  uint32_t rb_issue = 0, rb_cons = 0, ob = 0, fb = 0;
  int pf_it = 0, pf_item = 0;  // residual prefetch cursor (tile iteration, item)
  bool pf_done = false;
  int mt_mul = 1, mt_add = 0;  // 128-row block of a scheduler tile: mt * mt_mul + mt_add (CTA pairs: 2 blocks per tile, one per CTA)
  int tile0 = (int)blockIdx.x, tile_stride = (int)gridDim.x;  // this worker's tiles: tile0 + it * tile_stride (CTA pairs: pair index / count)

  __device__ EpiWarp(const GemmParams& p_, const CUtensorMap* r, const CUtensorMap* a, const CUtensorMap* s, const CUtensorMap* f,
                     uint8_t* eb_, uint64_t* rbar_, TileSched ts_, int lane_, int quad_, int half_, int nsplit_)
      : p(p_), tmRes(r), tmRaw(a), tmSnake(s), tmF32(f), eb(eb_), rbar(rbar_), ts(ts_), lane(lane_), quad(quad_), half(half_), nsplit(nsplit_) {
    has_res = p.res != nullptr && !Q3_DBG(p, 8);
    rf32 = p.res_f32 != 0;
    glu = p.act == ACT_GLU;
    tail = p.act == ACT_TAIL;
    of32 = p.out_f32 != nullptr && !tail;
    o16 = !glu && (p.out_raw != nullptr || p.out_snake != nullptr);
    pair16 = epi_pair16(p);
    units = p.bn >> 4;
    item_units = glu ? 4 : ((of32 || (has_res && rf32) || tail) ? 1 : 2);
    nitems = units / item_units;
    rd = epi_res_depth(p);  // 2, 4 or 8
    rd_shift = rd == 8 ? 3 : (rd == 4 ? 2 : 1);
    wraps = p.ncol_mod != p.n_total;
    int nslot = rd;
    s_raw = nslot;
    obufs = epi_out_bufs(p);
    if (pair16) nslot += 2;   // one slot per item holds both halves of the pair (hi in its first KB, lo in its second)
    else if (glu || p.out_raw) nslot += glu ? 2 : obufs;
    s_snake = nslot;
    if (!glu && !pair16 && p.out_snake) nslot += obufs;
    s_f32 = nslot;
  }

  // first item this warp owns in tile iteration `it` (the two warps of a quadrant alternate so odd item counts balance)
  __device__ __forceinline__ int first_item(int it) const {  // nsplit is 1, 2 or 3
    return nsplit == 3 ? (half + 2 * (it % 3)) % 3 : (half - it) & (nsplit - 1);
  }

  __device__ __forceinline__ void prefetch_one() {  // warp-uniform control flow; lane 0 issues
    if (pf_done) return;
    while (pf_item >= nitems) {
      ++pf_it;
      pf_item = first_item(pf_it);
      if (tile0 + pf_it * tile_stride >= ts.total_tiles) { pf_done = true; return; }
    }
    const int tile = tile0 + pf_it * tile_stride;
    if (tile >= ts.total_tiles) { pf_done = true; return; }
    if (lane == 0) {
      int nt, mt, b;
      ts.decode(tile, nt, mt, b);
      const uint32_t slot = rb_issue & (rd - 1);
      mbar_expect_tx(&rbar[slot], kEpiSlot);
      int row0 = (mt * mt_mul + mt_add) * 128 + quad * 32;
      if (p.pack_mode) gemm_chunk_coords(p, b, quad, row0, b);
      tma_load_3d(eb + slot * kEpiSlot, tmRes, &rbar[slot], nt * p.bn + pf_item * item_units * 16, row0, b);
    }
    ++rb_issue;
    pf_item += nsplit;
  }

  __device__ __forceinline__ void start() {
    if (has_res && !Q3_DBG(p, 1)) {
      pf_it = 0;
      pf_item = first_item(0);
      for (int i = 0; i < rd; ++i) prefetch_one();
    }
  }

  __device__ __forceinline__ void pack_store(uint8_t* o8, int cb, const float (&v)[16]) {
    uint4 a, c;
    a.x = Elem<T>::pack2(v[0], v[1]);   a.y = Elem<T>::pack2(v[2], v[3]);
    a.z = Elem<T>::pack2(v[4], v[5]);   a.w = Elem<T>::pack2(v[6], v[7]);
    c.x = Elem<T>::pack2(v[8], v[9]);   c.y = Elem<T>::pack2(v[10], v[11]);
    c.z = Elem<T>::pack2(v[12], v[13]); c.w = Elem<T>::pack2(v[14], v[15]);
    *reinterpret_cast<uint4*>(o8 + slot_off(lane, cb)) = a;
    *reinterpret_cast<uint4*>(o8 + slot_off(lane, cb + 1)) = c;
  }
  __device__ __forceinline__ void load_res16(const uint8_t* rbuf, int cb, float (&rs)[16]) {
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const uint4 t = *reinterpret_cast<const uint4*>(rbuf + slot_off(lane, cb + c));
      const uint32_t w4[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 f = Elem<T>::unpack2(w4[i]);
        rs[8 * c + 2 * i] = f.x; rs[8 * c + 2 * i + 1] = f.y;
      }
    }
  }
  // bias -> (GELU) -> scale -> + residual
  __device__ __forceinline__ void math_unit(float (&v)[16], const float (&rs)[16], int c0) {
    if (p.bias) {
      float t[16]; load16f(p.bias + c0, t);
#pragma unroll
      for (int i = 0; i < 16; i += 2) { const float2 r = add2(make_float2(v[i], v[i + 1]), make_float2(t[i], t[i + 1])); v[i] = r.x; v[i + 1] = r.y; }
    }
    if (p.act == ACT_GELU) {
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);
    }
    if (p.scale) {
      float t[16]; load16f(p.scale + c0, t);
#pragma unroll
      for (int i = 0; i < 16; i += 2) { const float2 r = mul2(make_float2(v[i], v[i + 1]), make_float2(t[i], t[i + 1])); v[i] = r.x; v[i + 1] = r.y; }
    }
    if (has_res) {
#pragma unroll
      for (int i = 0; i < 16; i += 2) { const float2 r = add2(make_float2(v[i], v[i + 1]), make_float2(rs[i], rs[i + 1])); v[i] = r.x; v[i + 1] = r.y; }
    }
  }
  __device__ __forceinline__ void snake_unit(float (&v)[16], int c0) {
    if (act_is_split(p.act)) {  // second tensor = the low half of the fp16 pair (gemm_params.h)
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] -= Elem<T>::to_f(Elem<T>::from_f(v[i]));
      return;
    }
    float sa[16], sib[16];
    load16f(p.snake_a + c0, sa);
    load16f(p.snake_ib + c0, sib);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      const float2 r = snake2(make_float2(v[i], v[i + 1]), make_float2(sa[i], sa[i + 1]), make_float2(sib[i], sib[i + 1]));
      v[i] = r.x; v[i + 1] = r.y;
    }
  }
  __device__ __forceinline__ void res_consumed() {  // slot fully read by every lane: refill it rd slots ahead
    __syncwarp();
    prefetch_one();
    ++rb_cons;
  }

  // ---- 16-bit flow: 32 accumulator columns, residual / outputs all 16-bit
  __device__ __forceinline__ void item_pair(uint32_t t0, int nt, int row0, int b, int item) {
    const int u0 = item * 2;
    const int n0 = nt * p.bn + u0 * 16;
    int c0 = n0;  // ncol_mod % 32 == 0 for every 16-bit flow, so the 32 columns never wrap
    if (wraps) { c0 = tile_c0 + u0 * 16; while (c0 >= p.ncol_mod) c0 -= p.ncol_mod; }
    float rsa[16], rsb[16];
    if (has_res) {
      const uint32_t slot = rb_cons & (rd - 1);
      mbar_wait(&rbar[slot], (rb_cons >> rd_shift) & 1);
      const uint8_t* rbuf = eb + slot * kEpiSlot;
      load_res16(rbuf, 0, rsa);
      load_res16(rbuf, 2, rsb);
    }
    // The asynchronous TMEM loads and their wait stay adjacent: the compiler treats the destination registers as defined
    // by the load statement, so any code in between could copy them before the data has arrived (seen as a rare glitch).
    uint32_t ra[16], rb[16];
    __syncwarp();  // tcgen05.ld is .sync.aligned: reconverge after the lane-0 branches / barrier spin loops above
    tmem_ld16(t0 + u0 * 16, ra);
    tmem_ld16(t0 + u0 * 16 + 16, rb);
    tmem_ld_wait();
    float va[16], vb[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) { va[i] = __uint_as_float(ra[i]); vb[i] = __uint_as_float(rb[i]); }
    math_unit(va, rsa, c0);
    math_unit(vb, rsb, c0 + 16);
    if (p.act == ACT_ELU_SPLIT) {
#pragma unroll
      for (int i = 0; i < 16; ++i) { va[i] = elu_fast(va[i]); vb[i] = elu_fast(vb[i]); }
    } else if (p.act == ACT_GELU_SPLIT) {
#pragma unroll
      for (int i = 0; i < 16; ++i) { va[i] = gelu_erf(va[i]); vb[i] = gelu_erf(vb[i]); }
    }
    const uint32_t sl = ob & (uint32_t)(obufs - 1);
    const bool st = !Q3_DBG(p, 4);
    if (st) {
      if (lane == 0) { if (obufs == 1) tma_store_wait_read<0>(); else tma_store_wait_read<1>(); }
      __syncwarp();
    }
    if (p.out_raw) {
      pack_store(eb + (s_raw + sl) * kEpiSlot, 0, va);
      pack_store(eb + (s_raw + sl) * kEpiSlot, 2, vb);
    }
    if (p.out_snake) {
      snake_unit(va, c0);
      snake_unit(vb, c0 + 16);
      pack_store(eb + (s_snake + sl) * kEpiSlot, 0, va);
      pack_store(eb + (s_snake + sl) * kEpiSlot, 2, vb);
    }
    if (st) {
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        int craw = n0, csnake = n0;
        if (act_is_split(p.act)) { craw = pair_hi_col(p, n0); csnake = craw + pair_lo_off(p); }   // both halves go through the pair tensor's map
        if (p.out_raw) tma_store_3d(tmRaw, eb + (s_raw + sl) * kEpiSlot, craw, row0, b);
        if (p.out_snake) tma_store_3d(tmSnake, eb + (s_snake + sl) * kEpiSlot, csnake, row0, b);
        tma_store_commit();
      }
    }
    // Hand the residual slot back to the TMA engine only now: the st.shared above consumed values computed from the residual
    // registers, and a warp issues in order, so every ld.shared of the slot has returned its data.  (Refilling right after
    // the loads were ISSUED let the asynchronous-proxy write overtake a queued load: a rare one-chunk glitch.)
    if (has_res) res_consumed();
    ++ob;
  }

  // ---- fp32 flow: 16 accumulator columns; residual and/or output are fp32 (transformer residual stream)
  __device__ __forceinline__ void item_f32(uint32_t t0, int nt, int row0, int b, int item) {
    const int n0 = nt * p.bn + item * 16;
    int c0 = n0;
    if (wraps) { c0 = tile_c0 + item * 16; while (c0 >= p.ncol_mod) c0 -= p.ncol_mod; }  // tile_c0: one modulo per tile, not per item
    float rs[16];
    if (has_res) {
      const uint32_t slot = rb_cons & (rd - 1);
      mbar_wait(&rbar[slot], (rb_cons >> rd_shift) & 1);
      const uint8_t* rbuf = eb + slot * kEpiSlot;
      if (rf32) {
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const float4 t = *reinterpret_cast<const float4*>(rbuf + slot_off(lane, c));
          rs[4 * c] = t.x; rs[4 * c + 1] = t.y; rs[4 * c + 2] = t.z; rs[4 * c + 3] = t.w;
        }
      } else {
        load_res16(rbuf, 0, rs);  // a 16-bit residual box of 16 columns sits in chunks 0..1
      }
      if (Q3_DBG(p, 512) && rf32) {  // debug: residual straight from global memory (race localisation)
        const int row = row0 + lane;
        if (row < p.rows_valid) {
          const float4* g = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(p.res) + (long long)b * p.out_batch_stride +
                                                            (long long)row * p.out_row_stride + n0);
#pragma unroll
          for (int c = 0; c < 4; ++c) { const float4 t = g[c]; rs[4 * c] = t.x; rs[4 * c + 1] = t.y; rs[4 * c + 2] = t.z; rs[4 * c + 3] = t.w; }
        }
      }
    }
    uint32_t r[16];
    __syncwarp();                  // .sync.aligned instruction: the warp must be converged
    tmem_ld16(t0 + item * 16, r);  // load and wait adjacent (see item_pair)
    tmem_ld_wait();
    float v[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
    if (tail) {
      if (has_res) res_consumed();
      if (p.bias) v[0] += __ldg(p.bias);
      const int row = row0 + lane;
      if (row < p.rows_valid && row >= p.tail_skip)
        p.out_f32[(long long)b * p.out_batch_stride + (row - p.tail_skip)] = fminf(1.0f, fmaxf(-1.0f, v[0]));
      return;
    }
    math_unit(v, rs, c0);
    if (of32 && Q3_DBG(p, 1024)) {  // debug: output straight to global memory (race localisation)
      const int row = row0 + lane;
      if (row < p.rows_valid) {
        float4* g = reinterpret_cast<float4*>(p.out_f32 + (long long)b * p.out_batch_stride + (long long)row * p.out_row_stride + n0);
#pragma unroll
        for (int c = 0; c < 4; ++c) g[c] = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      }
      if (has_res) res_consumed();
      return;
    }
    if (pair16) {
      // fp32 flow of ACT_ELU_SPLIT: x goes out as fp32 (when asked for) and ELU(x) as the fp16 pair, all in ONE store group per item;
      // both rings have two slots, so "at most one group still reading" frees the slots written two items ago.
      // The pair boxes are 32 rows x 16 columns (32 bytes, SWIZZLE_32B: 16-byte chunk c of row r sits at r * 32 + ((c ^ (r >> 2 & 1)) << 4)).
      const uint32_t sl = fb & 1;
      uint8_t* f8 = eb + (s_f32 + sl) * kEpiSlot;
      uint8_t* h8 = eb + (s_raw + sl) * kEpiSlot;
      if (lane == 0) tma_store_wait_read<1>();
      __syncwarp();
      if (of32) {
#pragma unroll
        for (int c = 0; c < 4; ++c)
          *reinterpret_cast<float4*>(f8 + slot_off(lane, c)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      }
      float lo[16];
      if (p.act == ACT_ELU_SPLIT) {   // the branch stays outside the loops: a select inside them evaluates erff for every element
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = elu_fast(v[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = gelu_erf(v[i]);
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) lo[i] = v[i] - Elem<T>::to_f(Elem<T>::from_f(v[i]));
      const uint32_t sw = (uint32_t)((lane >> 2) & 1);
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint4 a, l4;
        a.x = Elem<T>::pack2(v[8 * c], v[8 * c + 1]);      a.y = Elem<T>::pack2(v[8 * c + 2], v[8 * c + 3]);
        a.z = Elem<T>::pack2(v[8 * c + 4], v[8 * c + 5]);  a.w = Elem<T>::pack2(v[8 * c + 6], v[8 * c + 7]);
        l4.x = Elem<T>::pack2(lo[8 * c], lo[8 * c + 1]);     l4.y = Elem<T>::pack2(lo[8 * c + 2], lo[8 * c + 3]);
        l4.z = Elem<T>::pack2(lo[8 * c + 4], lo[8 * c + 5]); l4.w = Elem<T>::pack2(lo[8 * c + 6], lo[8 * c + 7]);
        const uint32_t off = (uint32_t)(lane * 32 + (((uint32_t)c ^ sw) << 4));
        *reinterpret_cast<uint4*>(h8 + off) = a;
        *reinterpret_cast<uint4*>(h8 + 1024 + off) = l4;
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        if (of32) tma_store_3d(tmF32, f8, n0, row0, b);
        const int chi = pair_hi_col(p, n0);
        tma_store_3d(tmRaw, h8, chi, row0, b);
        tma_store_3d(tmSnake, h8 + 1024, chi + pair_lo_off(p), row0, b);
        tma_store_commit();
      }
      ++fb;
      if (has_res) res_consumed();
      return;
    }
    if (of32) {
      uint8_t* o8 = eb + (s_f32 + (fb & 3)) * kEpiSlot;
      if (lane == 0) tma_store_wait_read<3>();
      __syncwarp();
#pragma unroll
      for (int c = 0; c < 4; ++c)
        *reinterpret_cast<float4*>(o8 + slot_off(lane, c)) = make_float4(v[4 * c], v[4 * c + 1], v[4 * c + 2], v[4 * c + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_store_3d(tmF32, o8, n0, row0, b);
        tma_store_commit();
      }
      ++fb;
    }
    if (has_res) res_consumed();  // after the st.shared that consumed the residual (see item_pair)
  }

  // ---- GLU flow: 64 interleaved (gate, up) accumulator columns -> 32 outputs (reference modeling :367-369)
  __device__ __forceinline__ uint4 glu8(const uint32_t (&r)[16]) {
    uint4 o;
    o.x = Elem<T>::pack2(silu(__uint_as_float(r[0])) * __uint_as_float(r[1]), silu(__uint_as_float(r[2])) * __uint_as_float(r[3]));
    o.y = Elem<T>::pack2(silu(__uint_as_float(r[4])) * __uint_as_float(r[5]), silu(__uint_as_float(r[6])) * __uint_as_float(r[7]));
    o.z = Elem<T>::pack2(silu(__uint_as_float(r[8])) * __uint_as_float(r[9]), silu(__uint_as_float(r[10])) * __uint_as_float(r[11]));
    o.w = Elem<T>::pack2(silu(__uint_as_float(r[12])) * __uint_as_float(r[13]), silu(__uint_as_float(r[14])) * __uint_as_float(r[15]));
    return o;
  }
  __device__ __forceinline__ void item_glu(uint32_t t0, int nt, int row0, int b, int item) {
    const int u0 = item * 4;
    uint8_t* o8 = eb + (s_raw + (ob & 1)) * kEpiSlot;
    if (lane == 0) tma_store_wait_read<1>();
    __syncwarp();
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      uint32_t ra[16], rb[16];
      __syncwarp();
      tmem_ld16(t0 + (u0 + 2 * h) * 16, ra);
      tmem_ld16(t0 + (u0 + 2 * h) * 16 + 16, rb);
      tmem_ld_wait();
      *reinterpret_cast<uint4*>(o8 + slot_off(lane, 2 * h)) = glu8(ra);
      *reinterpret_cast<uint4*>(o8 + slot_off(lane, 2 * h + 1)) = glu8(rb);
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_3d(tmRaw, o8, (nt * p.bn + u0 * 16) >> 1, row0, b);
      tma_store_commit();
    }
    ++ob;
  }

  // one accumulator tile: t0 = TMEM address of (this warp's lane quadrant, first column of the tile)
  __device__ __forceinline__ void tile(uint32_t t0, int it, int nt, int mt, int b) {
    int row0 = (mt * mt_mul + mt_add) * 128 + quad * 32;
    if (wraps) tile_c0 = (nt * p.bn) % p.ncol_mod;
    if (p.pack_mode) gemm_chunk_coords(p, b, quad, row0, b);  // packed tiles: the scheduler's "batch" is the tile index
    if (Q3_DBG(p, 1)) {
      uint32_t r[16];
      for (int u = half; u < units; u += nsplit) { tmem_ld16(t0 + u * 16, r); tmem_ld_wait(); }
      return;
    }
    for (int item = first_item(it); item < nitems; item += nsplit) {
      if (glu) item_glu(t0, nt, row0, b, item);
      else if (item_units == 2) item_pair(t0, nt, row0, b, item);
      else item_f32(t0, nt, row0, b, item);
    }
  }

  __device__ __forceinline__ void finish() {
    if (lane == 0) tma_store_wait<0>();
  }
};

}  // namespace q3
