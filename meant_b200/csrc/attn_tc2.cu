// Persistent, warp-specialised fused rotary/xPos attention for sequences that fit one 128-row
// tile (Sq, Sk <= 128: the MEANT text path S = 128, TimeSformer time groups), forward and
// backward, on tcgen05 tensor cores with TMEM accumulators.  bf16 in, fp32 accumulate.
//
// One CTA per SM walks (sequence, head) work items blockIdx.x, blockIdx.x + gridDim.x, ...
// (consecutive CTAs take the heads of one sequence, so the packed `b s (h d)` rows they read
// are contiguous in DRAM).  384 threads, two roles:
//   loader warps 8-11  cp.async (LDGSTS, 16 B, zero-filled past S) Q/K/V[/dO] tiles straight from
//                      the packed projections into SWIZZLE_128B / SWIZZLE_64B operand sub-tiles
//                      (whole rows written contiguously, read by the tensor core as is), up to
//                      three items ahead, completion signalled by cp.async.mbarrier.arrive; the
//                      padding-mask bias / LSE row; later drain the staged results to global
//                      memory with coalesced 16-byte stores.
//   compute warps 0-7  warps w and w+4 share TMEM lanes (query rows) 32(w%4).. : thread (row t, half)
//                      rotates + xPos-scales row t of Q (half 0) or K (half 1) IN PLACE from
//                      per-thread register tables (fp16 pairs, loaded once), issues the MMAs,
//                      and the halves split the key columns in interleaved 32-column blocks
//                      (balanced under the causal mask) for the softmax (forward) or the P/dS
//                      recompute (backward) straight out of TMEM, with one shared-memory exchange
//                      of the row max / row sums; accumulators go TMEM -> (un-rotate dQ/dK in
//                      fp32 registers) -> bf16 -> staging.
// Hand-offs are mbarriers (tcgen05.commit for MMA completion, mbarrier.arrive for data), never
// __syncthreads, so loads of item j+1/j+2, the tensor core on item j and the stores of item
// j-1 overlap.  Both kernels are HBM-bound at S = 128 (see DESIGN.md for the byte counts).
#include "attn_common.cuh"
#include "tc_utils.cuh"
#include <cuda_fp16.h>
#include <math_constants.h>

namespace meant {
namespace t2 {

constexpr uint32_t SUB_A = 128 * 128;     // 128 rows x 64 columns, SWIZZLE_128B sub-tile
constexpr uint32_t SUB_B = 128 * 64;      // 128 rows x 32 columns, SWIZZLE_64B sub-tile
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr int BAR_C = 1, BAR_L = 2;      // named barriers of the compute (256 threads) / loader (128) groups
constexpr int NC = 256, NL = 128, NT = NC + NL;
// The thread that issues the MMAs: lane 0 of warp 4 (row group 0, column half 1), the compute warp that has no
// key block under the causal mask, so the serial issue code hides in its idle time.
constexpr int ISSUER = 128;

// Optional in-kernel phase timing (build with -DMEANT_PROF): CTA 0 accumulates clock64 deltas per phase
// for one compute thread (slots 0..15) and one loader thread (slots 16..31); read with meant_debug_prof().
#ifdef MEANT_PROF
__device__ unsigned long long g_prof[64];
struct Prof {
  long long last; bool on; int base;
  __device__ __forceinline__ void init(bool enable, int b) { on = enable; base = b; last = clock64(); }
  __device__ __forceinline__ void tick(int slot) {
    if (on) { const long long now = clock64(); atomicAdd(&g_prof[base + slot], (unsigned long long)(now - last)); last = now; }
  }
};
#else
struct Prof {
  __device__ __forceinline__ void init(bool, int) {}
  __device__ __forceinline__ void tick(int) {}
};
#endif

template <int DH>
struct Geo {
  static constexpr int CPR = DH / 8;                 // 16-byte chunks per row
  static constexpr bool TAIL = DH > 64;              // columns [64, 96) live in a SWIZZLE_64B sub-tile
  static constexpr uint32_t OPER = SUB_A + (TAIL ? SUB_B : 0u);   // one 128 x DH operand tile
  static constexpr uint32_t PBUF = 2 * SUB_A;        // one 128 x 128 P / dS tile (two SWIZZLE_128B sub-tiles)
  static constexpr uint32_t PITCH = DH * 2 + 16;     // staging row pitch (bank-conflict free)
  static_assert(DH == 64 || DH == 96, "operand tile = one 64-column sub-tile (+ one 32-column sub-tile)");
};

// Shared-memory operand layout (conventions verified by tc_probe.cu::tc_probe_sw_kernel): every tile is a
// set of 128-row sub-tiles whose rows are contiguous (128 B: SWIZZLE_128B, 64 B: SWIZZLE_64B) with the
// 16-byte chunks of a row XOR-permuted, so that a row-walking cp.async stream writes whole rows (6.6 TB/s
// through LDGSTS, tools/ldbench.cu) and the tensor core reads the tile without a transposing copy.
// byte offset of 16-byte chunk ch of row r in a 128 x DH operand tile
__device__ __forceinline__ uint32_t oper_off(int r, int ch) {
  return ch < 8 ? tc::sw128_off(r, ch) : SUB_A + tc::sw64_off(r, ch - 8);
}
// byte offset of 16-byte chunk c (8 columns) of row r in a 128 x 128 P / dS tile
__device__ __forceinline__ uint32_t pbuf_off(int r, int c) { return (c >> 3) * SUB_A + tc::sw128_off(r, c & 7); }

// K-major view (K = head dimension) of an operand tile, K-step ks of 16 columns
__device__ __forceinline__ uint64_t desc_oper_k(uint32_t base, int ks) {
  return ks < 4 ? tc::make_desc_sw(base + ks * 32, 16, 1024, 2) : tc::make_desc_sw(base + SUB_A + (ks - 4) * 32, 16, 512, 4);
}
// MN-major view (K = rows) of an operand tile, K-step ks of 16 rows: columns [0, 64) / [64, 96)
__device__ __forceinline__ uint64_t desc_oper_mn_a(uint32_t base, int ks) { return tc::make_desc_sw(base + ks * 2048, SUB_A, 1024, 2); }
__device__ __forceinline__ uint64_t desc_oper_mn_b(uint32_t base, int ks) { return tc::make_desc_sw(base + SUB_A + ks * 1024, SUB_B, 512, 4); }
// P / dS tile: K-major (K = key columns) and MN-major (K = query rows; the two sub-tiles are LBO apart)
__device__ __forceinline__ uint64_t desc_p_k(uint32_t base, int ks) { return tc::make_desc_sw(base + (ks >> 2) * SUB_A + (ks & 3) * 32, 16, 1024, 2); }
__device__ __forceinline__ uint64_t desc_p_mn(uint32_t base, int ks) { return tc::make_desc_sw(base + ks * 2048, SUB_A, 1024, 2); }

// D[128 x DH] (+)= A * B with B an MN-major operand tile: one MMA per sub-tile (N = 64, then N = 32)
template <int DH>
__device__ __forceinline__ void mma_oper_mn(uint32_t tmem_d, uint64_t adesc, uint32_t sb, int ks, int a_mn, bool acc) {
  tc::mma_ss(tmem_d, adesc, desc_oper_mn_a(sb, ks), tc::make_idesc(128, 64, a_mn, 1), acc);
  if (Geo<DH>::TAIL) tc::mma_ss(tmem_d + 64, adesc, desc_oper_mn_b(sb, ks), tc::make_idesc(128, 32, a_mn, 1), acc);
}

__device__ __forceinline__ const __nv_bfloat16* head_base(const HeadView& v, int n, int n2c, int h) {
  const int n1 = n / n2c, n2 = n - n1 * n2c;
  return static_cast<const __nv_bfloat16*>(v.p) + (int64_t)n1 * v.sn1 + (int64_t)n2 * v.sn2 + (int64_t)h * v.sh;
}

// Loader thread t of 128 moves chunks idx = t + 128 k (k < CPR) of every 128 x DH tile: lanes walk the 16-byte
// chunks of a row first, so a warp reads ~2.7 whole rows (full 64-byte segments, few 128-byte lines) and
// writes them to contiguous (swizzled) shared rows.  The (row, chunk, shared offset) split is the same for
// every tile and item, so it is computed once.
template <int DH>
struct ChunkMap {
  static constexpr int CPR = DH / 8;
  int rc[CPR];                                     // row | chunk << 8 | shared offset << 16
  __device__ __forceinline__ void init(int t) {
#pragma unroll
    for (int k = 0; k < CPR; ++k) {
      const int idx = t + k * 128, r = idx / CPR, ch = idx - r * CPR;
      rc[k] = r | (ch << 8) | (int)(oper_off(r, ch) << 16);
    }
  }
  __device__ __forceinline__ int row(int k) const { return rc[k] & 0xff; }
  __device__ __forceinline__ int gcol(int k) const { return ((rc[k] >> 8) & 0xff) * 8; }             // element column
  __device__ __forceinline__ uint32_t soff(int k) const { return (uint32_t)rc[k] >> 16; }            // swizzled operand tile
  __device__ __forceinline__ uint32_t stg(int k) const { return (rc[k] & 0xff) * Geo<DH>::PITCH + ((rc[k] >> 8) & 0xff) * 16; }
};

// rows [0,128) of head (n, h) -> swizzled operand tile at shared address sbase; rows >= S are zero filled.
template <int DH>
__device__ __forceinline__ void issue_tile(const ChunkMap<DH>& cm, uint32_t sbase, const HeadView& v, int n, int n2c, int h, int S) {
  const __nv_bfloat16* base = head_base(v, n, n2c, h);
  const int ss = (int)v.ss;
#pragma unroll
  for (int k = 0; k < DH / 8; ++k) {
    const bool ok = cm.row(k) < S;
    tc::cp_async16(sbase + cm.soff(k), base + (ok ? cm.row(k) * ss : 0) + cm.gcol(k), ok ? 16u : 0u);
  }
}

// staged rows (row-major, PITCH bytes apart) -> head (n, h) of a global view, coalesced.
template <int DH>
__device__ __forceinline__ void drain_rows(const ChunkMap<DH>& cm, const uint8_t* stg, const HeadView& v, int n, int n2c, int h, int S) {
  __nv_bfloat16* base = const_cast<__nv_bfloat16*>(head_base(v, n, n2c, h));
  const int ss = (int)v.ss;
#pragma unroll
  for (int k = 0; k < DH / 8; ++k) {
    if (cm.row(k) < S) {
      const uint4 val = *reinterpret_cast<const uint4*>(stg + cm.stg(k));
      *reinterpret_cast<uint4*>(base + cm.row(k) * ss + cm.gcol(k)) = val;
    }
  }
}

// Per-thread rotation table of one position: (C[2i], C[2i+1]) / (S[2i], S[2i+1]) pairs in fp16
// (|values| <= ~1.3; the 2^-12 relative error is an order below the bf16 rounding of the result).
template <int RC>
struct RotTab {
  __half2 c[RC * 4], s[RC * 4];
};

template <int RC>
__device__ __forceinline__ void load_tab(RotTab<RC>& T, const float* cs, const float* sn, int pos, int rot_dim) {
  const float2* c2 = reinterpret_cast<const float2*>(cs + (int64_t)pos * rot_dim);
  const float2* s2 = reinterpret_cast<const float2*>(sn + (int64_t)pos * rot_dim);
#pragma unroll
  for (int i = 0; i < RC * 4; ++i) {
    float2 a = make_float2(1.f, 1.f), b = make_float2(0.f, 0.f);
    if (2 * i < rot_dim) { a = __ldg(c2 + i); b = __ldg(s2 + i); }
    T.c[i] = __floats2half2_rn(a.x, a.y);
    T.s[i] = __floats2half2_rn(b.x, b.y);
  }
}

// Rotate (INV = false) or un-rotate (transpose rotation, INV = true) the first rc chunks of one
// row r of an operand tile in place (RC <= 8: the rotated chunks all live in the SWIZZLE_128B sub-tile).
template <int RC, bool INV>
__device__ __forceinline__ void rot_row(uint8_t* tile, int r, int rc, const RotTab<RC>& T) {
  static_assert(RC <= 8, "rotated chunks must stay in the first sub-tile");
#pragma unroll
  for (int ch = 0; ch < RC; ++ch) {
    if (ch < rc) {
      uint4* p = reinterpret_cast<uint4*>(tile + tc::sw128_off(r, ch));
      uint4 raw = *p;
      __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 x = __bfloat1622float2(h2[i]);
        const float2 c = __half22float2(T.c[ch * 4 + i]), s = __half22float2(T.s[ch * 4 + i]);
        float y0, y1;
        if (!INV) { y0 = x.x * c.x - x.y * s.x; y1 = x.y * c.y + x.x * s.y; }
        else      { y0 = x.x * c.x + x.y * s.y; y1 = x.y * c.y - x.x * s.x; }
        h2[i] = __floats2bfloat162_rn(y0, y1);
      }
      *p = raw;
    }
  }
}

// TMEM accumulator (this thread's lane, NCOL fp32 columns from taddr; NCOL = 32 or 48) -> bf16 at dst
template <int NCOL>
__device__ __forceinline__ void stage_acc_cols(uint32_t taddr, uint8_t* dst, float mul) {
  float o[32];
  tc::tmem_ld32(taddr, o);
  tc::tmem_ld_wait();
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint4 pk;
    __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
    for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(o[q * 8 + 2 * i] * mul, o[q * 8 + 2 * i + 1] * mul);
    *reinterpret_cast<uint4*>(dst + q * 16) = pk;
  }
  if (NCOL > 32) {
    tc::tmem_ld16(taddr + 32, o);
    tc::tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      uint4 pk;
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(o[q * 8 + 2 * i] * mul, o[q * 8 + 2 * i + 1] * mul);
      *reinterpret_cast<uint4*>(dst + 64 + q * 16) = pk;
    }
  }
}

// Full accumulator row (DH fp32 columns) -> un-rotate the first rc chunks in fp32 registers -> bf16 row at dst
template <int DH, int RC>
__device__ __forceinline__ void stage_acc_unrot(uint32_t taddr, uint8_t* dst, int rc, const RotTab<RC>& T) {
#pragma unroll
  for (int c0 = 0; c0 < DH; c0 += 32) {
    float o[32];
    tc::tmem_ld32(taddr + c0, o);
    tc::tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ch = c0 / 8 + q;
      uint4 pk;
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        float g0 = o[q * 8 + 2 * i], g1 = o[q * 8 + 2 * i + 1];
        if (ch < RC) {
          if (ch < rc) {
            const float2 c = __half22float2(T.c[(ch < RC ? ch : 0) * 4 + i]), s = __half22float2(T.s[(ch < RC ? ch : 0) * 4 + i]);
            const float y0 = g0 * c.x + g1 * s.y, y1 = g1 * c.y - g0 * s.x;
            g0 = y0;
            g1 = y1;
          }
        }
        ph[i] = __floats2bfloat162_rn(g0, g1);
      }
      *reinterpret_cast<uint4*>(dst + ch * 16) = pk;
    }
  }
}

// additive key bias in the exp2 domain: padding mask (meant/xPosAttention.py:54-56) and keys past Sk
__device__ __forceinline__ float key_bias2(float mval, int k, int Sk) {
  return k < Sk ? (1.0f - mval) * -1e9f * LOG2E : -CUDART_INF_F;
}

// ============================================================================== forward
template <int DH>
struct FwdL {
  using G = Geo<DH>;
  static constexpr int NST = 3;                                // operand stages in flight
  static constexpr uint32_t STAGE = 3 * G::OPER;               // Q | K | V   (P and the O staging alias Q | K)
  static constexpr uint32_t AUX = NST * STAGE;                 // [NST][128] fp32 additive key bias (exp2 domain)
  static constexpr uint32_t RED = AUX + NST * 512;             // [2][2][128] fp32 row max / row sum exchange
  static constexpr uint32_t BARS = RED + 2048;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;          // + alignment slack
  static_assert(G::PBUF <= 2 * G::OPER && 128 * G::PITCH <= 2 * G::OPER, "P / staging must fit in Q | K");
};

template <int DH, int RC>
__global__ void __launch_bounds__(NT, 1) attn2_fwd_kernel(AttnArgs a, int n_items) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms are 1024-byte aligned
  using L = FwdL<DH>;
  using G = Geo<DH>;
  constexpr int NST = L::NST;
  constexpr int HALF = DH / 2;                                 // accumulator columns per compute half
  // TMEM: one 128-column buffer per stage: S, then (once P is in shared memory) O over its first DH columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed = bars, *s_full = bars + NST, *o_full = bars + 2 * NST, *o_staged = bars + 3 * NST,
           *aux_ready = bars + 4 * NST;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&landed[i], NL);
      tc::mbar_init(&aux_ready[i], NL);
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&o_full[i], 1);
      tc::mbar_init(&o_staged[i], NC);
    }
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = a.qcos ? a.rot_dim / 8 : 0;
  const uint32_t s0 = tc::smem_u32(smem);
  const int n_s = (a.Sk + 15) & ~15;                            // PV K extent (keys, rounded to one MMA K step)

  if (tid >= NC) {
    // ------------------------------------------------------------------ loader group
    const int t = tid - NC;
    ChunkMap<DH> cm;
    cm.init(t);
    Prof pf;
    pf.init(blockIdx.x == 0 && t == 0, 16);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      pf.tick(0);
      if (j >= NST) {   // the stage still holds item j - NST: drain its staged output first
        const int pitem = item - NST * (int)gridDim.x, pn = pitem / a.H, ph = pitem - pn * a.H;
        tc::mbar_wait(&o_staged[st], ((j - NST) / NST) & 1);
        pf.tick(1);
        drain_rows<DH>(cm, smem + st * L::STAGE, a.o, pn, a.n2_count, ph, a.Sq);
        pf.tick(2);
        tc::bar_sync(BAR_L, NL);    // every staged row has been read before the stage is refilled
        pf.tick(3);
      }
      float mval = 1.f;
      if (a.key_mask && t < a.Sk) mval = __ldg(a.key_mask + (int64_t)n * a.Sk + t);
      const uint32_t sb = s0 + st * L::STAGE;
      issue_tile<DH>(cm, sb, a.q, n, a.n2_count, h, a.Sq);
      issue_tile<DH>(cm, sb + G::OPER, a.k, n, a.n2_count, h, a.Sk);
      issue_tile<DH>(cm, sb + 2 * G::OPER, a.v, n, a.n2_count, h, a.Sk);
      tc::cp_async_arrive(&landed[st]);                           // fires when this thread's copies have landed
      pf.tick(4);
      reinterpret_cast<float*>(smem + L::AUX + st * 512)[t] = key_bias2(mval, t, a.Sk);
      tc::mbar_arrive(&aux_ready[st]);
      pf.tick(5);
    }
    for (int j = max(0, n_my - NST); j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      tc::mbar_wait(&o_staged[st], (j / NST) & 1);
      drain_rows<DH>(cm, smem + st * L::STAGE, a.o, n, a.n2_count, h, a.Sq);
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;                    // TMEM lane quarter / column half
    const int t = rg * 32 + lane;                                 // query row (and the Q or K row this thread rotates)
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    float* red_max = reinterpret_cast<float*>(smem + L::RED);     // [2][128]
    float* red_sum = red_max + 256;                               // [2][128]
    int nblk = (a.Sk + 31) >> 5;                                  // 32-key blocks with any unmasked key for this warp's rows
    if (a.causal) nblk = min(nblk, rg + 1);
    RotTab<RC> tab;                                               // half 0: q tables, half 1: k tables
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_h = -1;
    if (rc && !a.pos_is_head) load_tab<RC>(tab, tcs, tsn, min(t, (half ? a.Sk : a.Sq) - 1), a.rot_dim);
    // rotate item j's Q / K rows in place, then start its score MMA
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    auto prepare = [&](int j) {
      const int st = j % NST;
      pf.tick(8);
      tc::mbar_wait(&landed[st], (j / NST) & 1);
      pf.tick(9);
      if (rc) {
        if (a.pos_is_head) {
          const int h = (blockIdx.x + j * gridDim.x) % a.H;
          if (h != cur_h) { load_tab<RC>(tab, tcs, tsn, h, a.rot_dim); cur_h = h; }
        }
        rot_row<RC, false>(smem + st * L::STAGE + half * G::OPER, t, rc, tab);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      pf.tick(10);
      tc::bar_sync(BAR_C, NC);
      pf.tick(11);
      if (tid == ISSUER) {
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + st * 128, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
        tc::mma_commit(&s_full[st]);
      }
    };
    if (n_my > 0) prepare(0);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      const uint32_t par = (j / NST) & 1;
      pf.tick(0);
      tc::mbar_wait(&aux_ready[st], par);
      tc::mbar_wait(&s_full[st], par);
      tc::fence_after_sync();
      pf.tick(1);
      const float4* b4 = reinterpret_cast<const float4*>(smem + L::AUX + st * 512);
      const uint32_t ts = lane_base + st * 128;
      // two passes over the scores in TMEM (re-reading them is cheaper than 64 live registers per thread)
      float m = -CUDART_INF_F;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const int lim = (a.causal && blk == rg) ? lane : 31;      // diagonal block: keys after the query are masked
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 b = b4[(blk * 32 + e) >> 2];
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float v = fmaf(sv[e + u], scale2, bb[u]);
              m = fmaxf(m, e + u > lim ? -CUDART_INF_F : v);
            }
          }
        }
      }
      red_max[half * 128 + t] = m;
      pf.tick(2);
      tc::bar_sync(BAR_C, NC);
      pf.tick(3);
      m = fmaxf(m, red_max[(half ^ 1) * 128 + t]);
      const float m_use = m == -CUDART_INF_F ? 0.f : m;
      float l0 = 0.f, l1 = 0.f;
      uint8_t* pbuf = smem + st * L::STAGE;                         // P over the consumed Q | K
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const int lim = (a.causal && blk == rg) ? lane : 31;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 ba = b4[(blk * 32 + q * 8) >> 2], bc = b4[((blk * 32 + q * 8) >> 2) + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u]) - m_use), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1]) - m_use);
              if (e > lim) p0 = 0.f;
              if (e + 1 > lim) p1 = 0.f;
              l0 += p0;
              l1 += p1;
              ph[u] = __floats2bfloat162_rn(p0, p1);
            }
            *reinterpret_cast<uint4*>(pbuf + pbuf_off(t, blk * 4 + q)) = pk;
          }
        } else if (blk * 32 < n_s) {
#pragma unroll
          for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(pbuf + pbuf_off(t, blk * 4 + q)) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      float l = l0 + l1;
      red_sum[half * 128 + t] = l;
      tc::fence_async_smem();
      tc::fence_before_sync();
      pf.tick(4);
      tc::bar_sync(BAR_C, NC);
      pf.tick(5);
      if (tid == ISSUER) {
        tc::fence_after_sync();
        const uint32_t sp = s0 + st * L::STAGE, sv = sp + 2 * G::OPER;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks * 16 < n_s) mma_oper_mn<DH>(tmem + st * 128, desc_p_k(sp, ks), sv, ks, 0, ks > 0);
        tc::mma_commit(&o_full[st]);
      }
      l += red_sum[(half ^ 1) * 128 + t];
      if (j + 1 < n_my) prepare(j + 1);                             // next item's rotation + score MMA hide the PV latency
      pf.tick(12);
      tc::mbar_wait(&o_full[st], par);
      tc::fence_after_sync();
      pf.tick(6);
      stage_acc_cols<HALF>(ts + half * HALF, smem + st * L::STAGE + t * G::PITCH + half * HALF * 2, l > 0.f ? 1.0f / l : 0.f);
      if (half == 0 && t < a.Sq && a.lse) a.lse[((int64_t)n * a.H + h) * a.Sq + t] = m_use * LN2 + logf(l);
      tc::fence_before_sync();
      tc::mbar_arrive(&o_staged[st]);
      pf.tick(7);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== backward
template <int DH>
struct BwdL {
  using G = Geo<DH>;
  static constexpr uint32_t Q = 0, K = G::OPER, V = 2 * G::OPER, DO = 3 * G::OPER, P = 4 * G::OPER, DS = P + G::PBUF;
  static constexpr uint32_t STG_Z = DS + G::PBUF;              // dK staging (dQ stages over P, dV over dS)
  static constexpr uint32_t AUX = STG_Z + 128 * G::PITCH;      // bias2[128] | lse*log2e[128] | delta exchange [2][128]
  static constexpr uint32_t BARS = AUX + 2048;
  static constexpr uint32_t TOTAL = BARS + 128 + 1024;          // + alignment slack
  static_assert(128 * G::PITCH <= G::PBUF, "gradient staging must fit in the P / dS buffers");
};

template <int DH, int RC>
__global__ void __launch_bounds__(NT, 1) attn2_bwd_kernel(AttnArgs a, int n_items) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms are 1024-byte aligned
  using L = BwdL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 352, T_DQ = 0;   // dQ reuses the consumed S columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed_qk = bars, *landed_dv = bars + 1, *sdp_full = bars + 2, *grads_full = bars + 3, *staged = bars + 4,
           *stg_free = bars + 5, *aux_ready = bars + 6;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    tc::mbar_init(landed_qk, NL);
    tc::mbar_init(landed_dv, NL);
    tc::mbar_init(aux_ready, NL);
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(grads_full, 1);
    tc::mbar_init(staged, NC);
    tc::mbar_init(stg_free, NL);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = a.qcos ? a.rot_dim / 8 : 0;
  const uint32_t s0 = tc::smem_u32(smem);
  const uint32_t sq = s0 + L::Q, sk = s0 + L::K, sv = s0 + L::V, sdo = s0 + L::DO, sp = s0 + L::P, sds = s0 + L::DS;
  uint8_t* stg_x = smem + L::P;       // dQ staging (free once the gradient MMAs are done)
  uint8_t* stg_y = smem + L::DS;      // dV staging
  uint8_t* stg_z = smem + L::STG_Z;   // dK staging
  float* aux_bias = reinterpret_cast<float*>(smem + L::AUX);
  float* aux_lse2 = aux_bias + 128;
  float* red_delta = aux_bias + 256;                                // [2][128]

  if (tid >= NC) {
    // ------------------------------------------------------------------ loader group
    const int t = tid - NC;
    ChunkMap<DH> cm;
    cm.init(t);
    Prof pf;
    pf.init(blockIdx.x == 0 && t == 0, 16);
    auto drain = [&](int j) {   // the three gradients of item j, staged by the compute group
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H;
      pf.tick(4);
      tc::mbar_wait(staged, j & 1);
      pf.tick(5);
      drain_rows<DH>(cm, stg_x, a.dq, n, a.n2_count, h, a.Sq);
      drain_rows<DH>(cm, stg_y, a.dv, n, a.n2_count, h, a.Sk);
      drain_rows<DH>(cm, stg_z, a.dk, n, a.n2_count, h, a.Sk);
      tc::mbar_arrive(stg_free);
      pf.tick(6);
    };
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H;
      pf.tick(0);
      if (j > 0) tc::mbar_wait(grads_full, (j - 1) & 1);           // operand tiles of item j-1 are consumed
      pf.tick(1);
      float mval = 1.f, lse = 0.f;
      if (a.key_mask && t < a.Sk) mval = __ldg(a.key_mask + (int64_t)n * a.Sk + t);
      if (t < a.Sq) lse = __ldg(a.lse + ((int64_t)n * a.H + h) * a.Sq + t);
      issue_tile<DH>(cm, sq, a.q, n, a.n2_count, h, a.Sq);
      issue_tile<DH>(cm, sk, a.k, n, a.n2_count, h, a.Sk);
      tc::cp_async_arrive(landed_qk);
      issue_tile<DH>(cm, sdo, a.o, n, a.n2_count, h, a.Sq);
      issue_tile<DH>(cm, sv, a.v, n, a.n2_count, h, a.Sk);
      tc::cp_async_arrive(landed_dv);
      pf.tick(2);
      aux_bias[t] = key_bias2(mval, t, a.Sk);
      aux_lse2[t] = lse * LOG2E;
      tc::mbar_arrive(aux_ready);
      pf.tick(3);
      if (j > 0) drain(j - 1);
    }
    if (n_my > 0) drain(n_my - 1);
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale = a.scale, scale2 = a.scale * LOG2E;
    const float4* b4 = reinterpret_cast<const float4*>(aux_bias);
    const bool qrow = t < a.Sq;
    int nblk = (a.Sk + 31) >> 5;
    if (a.causal) nblk = min(nblk, rg + 1);
    RotTab<RC> tab;                                               // half 0: q tables (Q, dQ), half 1: k tables (K, dK)
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_h = -1;
    if (rc && !a.pos_is_head) load_tab<RC>(tab, tcs, tsn, min(t, (half ? a.Sk : a.Sq) - 1), a.rot_dim);
    Prof pc;
    pc.init(blockIdx.x == 0 && tid == 96, 0);
    for (int j = 0; j < n_my; ++j) {
      pc.tick(0);
      // ---- rotate Q / K rows in place, start S = Q K^T and dP = dO V^T
      tc::mbar_wait(landed_qk, j & 1);
      pc.tick(1);
      if (rc) {
        if (a.pos_is_head) {
          const int h = (blockIdx.x + j * gridDim.x) % a.H;
          if (h != cur_h) { load_tab<RC>(tab, tcs, tsn, h, a.rot_dim); cur_h = h; }
        }
        rot_row<RC, false>(smem + (half ? L::K : L::Q), t, rc, tab);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      pc.tick(2);
      tc::bar_sync(BAR_C, NC);
      pc.tick(3);
      if (tid == ISSUER) {
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
        tc::mbar_wait(landed_dv, j & 1);                            // dO and V follow Q and K through the loader
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_oper_k(sdo, ks), desc_oper_k(sv, ks), idesc, ks > 0);
        tc::mma_commit(sdp_full);
      }
      tc::mbar_wait(aux_ready, j & 1);
      if (j > 0) tc::mbar_wait(stg_free, (j - 1) & 1);              // P / dS double as the previous item's staging
      tc::mbar_wait(sdp_full, j & 1);
      tc::fence_after_sync();
      pc.tick(4);
      const float lse2 = aux_lse2[t];
      uint32_t pk[32];                                              // this thread's 64 columns of P, bf16 pairs
      float delta = 0.f;                                            // rowsum(P * dP) == rowsum(dO * O)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
          const int lim = !qrow ? -1 : ((a.causal && blk == rg) ? lane : 31);
#pragma unroll
          for (int hb = 0; hb < 2; ++hb) {                          // 16 columns at a time keeps the live set small
            float sv16[16], dp[16];
            tc::tmem_ld16(lane_base + T_S + blk * 32 + hb * 16, sv16);
            tc::tmem_ld16(lane_base + T_DP + blk * 32 + hb * 16, dp);
            tc::tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; e += 4) {
              const float4 b = b4[(blk * 32 + hb * 16 + e) >> 2];
              const float bb[4] = {b.x, b.y, b.z, b.w};
              float p[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                p[u] = tc::ex2(fmaf(sv16[e + u], scale2, bb[u]) - lse2);
                if (hb * 16 + e + u > lim) p[u] = 0.f;
                delta = fmaf(p[u], dp[e + u], delta);
              }
              const __nv_bfloat162 p01 = __floats2bfloat162_rn(p[0], p[1]), p23 = __floats2bfloat162_rn(p[2], p[3]);
              pk[i * 16 + hb * 8 + e / 2] = *reinterpret_cast<const uint32_t*>(&p01);
              pk[i * 16 + hb * 8 + e / 2 + 1] = *reinterpret_cast<const uint32_t*>(&p23);
            }
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) pk[i * 16 + e] = 0u;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(smem + L::P + pbuf_off(t, blk * 4 + q)) =
              make_uint4(pk[i * 16 + q * 4], pk[i * 16 + q * 4 + 1], pk[i * 16 + q * 4 + 2], pk[i * 16 + q * 4 + 3]);
      }
      red_delta[half * 128 + t] = delta;
      pc.tick(5);
      tc::bar_sync(BAR_C, NC);
      pc.tick(6);
      delta += red_delta[(half ^ 1) * 128 + t];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
#pragma unroll
        for (int hb = 0; hb < 2; ++hb) {
          uint32_t dsk[8];
          if (blk < nblk) {
            float dp[16];
            tc::tmem_ld16(lane_base + T_DP + blk * 32 + hb * 16, dp);
            tc::tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 8; ++e) {
              const float2 p = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk[i * 16 + hb * 8 + e]));
              const __nv_bfloat162 d2 = __floats2bfloat162_rn(p.x * (dp[2 * e] - delta) * scale, p.y * (dp[2 * e + 1] - delta) * scale);
              dsk[e] = *reinterpret_cast<const uint32_t*>(&d2);
            }
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) dsk[e] = 0u;
          }
#pragma unroll
          for (int q = 0; q < 2; ++q)
            *reinterpret_cast<uint4*>(smem + L::DS + pbuf_off(t, blk * 4 + hb * 2 + q)) =
                make_uint4(dsk[q * 4], dsk[q * 4 + 1], dsk[q * 4 + 2], dsk[q * 4 + 3]);
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      pc.tick(7);
      tc::bar_sync(BAR_C, NC);
      pc.tick(8);
      if (tid == ISSUER) {
        tc::fence_after_sync();
        const int kq = (a.Sq + 15) >> 4, kk = (a.Sk + 15) >> 4;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kq) mma_oper_mn<DH>(tmem + T_DV, desc_p_mn(sp, ks), sdo, ks, 1, ks > 0);   // dV = P^T dO
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kq) mma_oper_mn<DH>(tmem + T_DK, desc_p_mn(sds, ks), sq, ks, 1, ks > 0);   // dK = dS^T Q
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kk) mma_oper_mn<DH>(tmem + T_DQ, desc_p_k(sds, ks), sk, ks, 0, ks > 0);    // dQ = dS K
        tc::mma_commit(grads_full);
      }
      tc::mbar_wait(grads_full, j & 1);
      tc::fence_after_sync();
      pc.tick(9);
      // half 0 owns the dQ row, half 1 the dK row (each un-rotates with its own table); dV is split by columns
      if (half == 0) stage_acc_unrot<DH, RC>(lane_base + T_DQ, stg_x + t * G::PITCH, rc, tab);
      else           stage_acc_unrot<DH, RC>(lane_base + T_DK, stg_z + t * G::PITCH, rc, tab);
      stage_acc_cols<HALF>(lane_base + T_DV + half * HALF, stg_y + t * G::PITCH + half * HALF * 2, 1.0f);
      tc::fence_before_sync();
      tc::mbar_arrive(staged);
      pc.tick(10);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

}  // namespace t2

// ------------------------------------------------------------------------------------ host side
static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

// single-tile problems the persistent kernels take (the caller has already checked dtype, views, Dh)
int attn_tc2_supported(const AttnArgs& a, int backward) {
  if (a.Sq > 128 || a.Sk > 128) return 0;
  if (a.qcos) {
    const int rc_max = a.Dh == 96 ? 6 : 8;
    if (a.rot_dim / 8 > rc_max) return 0;
    const uintptr_t al = reinterpret_cast<uintptr_t>(a.qcos) | reinterpret_cast<uintptr_t>(a.qsin) |
                         reinterpret_cast<uintptr_t>(a.kcos) | reinterpret_cast<uintptr_t>(a.ksin);
    if (al % 8 != 0) return 0;
  }
  if (backward && !a.lse) return 0;
  const HeadView* vs[7] = {&a.q, &a.k, &a.v, &a.o, &a.dq, &a.dk, &a.dv};
  for (int i = 0; i < (backward ? 7 : 4); ++i)
    if (vs[i]->ss <= 0 || vs[i]->ss >= (1 << 23)) return 0;   // row offsets are computed in 32 bits
  return 1;
}

template <typename Kern>
static int launch2(Kern kern, size_t smem, const AttnArgs& a, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  const int64_t items = (int64_t)a.N * a.H;
  const int grid = (int)(items < num_sms() ? items : num_sms());
  kern<<<grid, t2::NT, smem, st>>>(a, (int)items);
  MEANT_LAUNCH_CHECK();
  return 0;
}

}  // namespace meant
extern "C" int meant_debug_prof(unsigned long long* out64, int reset) {
#ifdef MEANT_PROF
  cudaDeviceSynchronize();
  if (out64 && cudaMemcpyFromSymbol(out64, meant::t2::g_prof, sizeof(unsigned long long) * 64) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[64] = {0};
    cudaMemcpyToSymbol(meant::t2::g_prof, z, sizeof(z));
  }
  return 1;
#else
  (void)out64; (void)reset;
  return 0;
#endif
}
namespace meant {

int attn_tc2_fwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch2(t2::attn2_fwd_kernel<96, 6>, t2::FwdL<96>::TOTAL, a, st);
  return launch2(t2::attn2_fwd_kernel<64, 8>, t2::FwdL<64>::TOTAL, a, st);
}

int attn_tc2_bwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch2(t2::attn2_bwd_kernel<96, 6>, t2::BwdL<96>::TOTAL, a, st);
  return launch2(t2::attn2_bwd_kernel<64, 8>, t2::BwdL<64>::TOTAL, a, st);
}

}  // namespace meant
