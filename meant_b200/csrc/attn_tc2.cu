// Persistent, warp-specialised fused rotary/xPos attention for sequences that fit one 128-row
// tile (Sq, Sk <= 128: the MEANT text path S = 128, TimeSformer time groups), forward and
// backward, on tcgen05 tensor cores with TMEM accumulators.  bf16 in, fp32 accumulate.
//
// One CTA per SM walks (sequence, head) work items blockIdx.x, blockIdx.x + gridDim.x, ...
// (consecutive CTAs take the heads of one sequence, so the packed `b s (h d)` rows they read
// are contiguous in DRAM).  320 threads, three roles:
//   producer warp 8    one thread drives TMA: Q/K/V[/dO] head tiles go from the packed projections
//                      (5-D tensor maps over the head views, rows past S zero-filled) straight into
//                      SWIZZLE_128B / SWIZZLE_64B operand sub-tiles, completion counted in bytes on an
//                      mbarrier; staged results leave through TMA stores.  The warp also turns the
//                      padding-mask row (and, backward, the LSE row) into per-item shared vectors.
//   MMA warp 9         one thread waits for "operands rotated" / "P written" and issues the
//                      tcgen05.mma chains, so no computing warp ever stalls on instruction issue.
//   compute warps 0-7  warps w and w+4 share TMEM lanes (query rows) 32(w%4).. : thread (row t, half)
//                      rotates + xPos-scales row t of Q (half 0) or K (half 1) IN PLACE from
//                      per-thread register tables (fp16 pairs, loaded once), and the halves split
//                      the key columns in interleaved 32-column blocks (balanced under the causal
//                      mask) for the softmax (forward) or the P/dS recompute (backward) straight out
//                      of TMEM, exchanging the row max / row sums through shared memory behind a
//                      64-thread named barrier; accumulators go TMEM -> (un-rotate dQ/dK in fp32
//                      registers) -> bf16 -> swizzled staging tile.
// Every hand-off is an mbarrier (tcgen05.commit for MMA completion, expect_tx for TMA, arrive for
// data), never __syncthreads: loads of the next items, the tensor core, the softmax of item j and
// the stores of item j-1 overlap.  Forward keeps three Q|K|V stages in flight; backward double
// buffers Q|K and recycles V / dO as soon as their last MMA has run.  Both kernels are HBM-bound
// at S = 128 (see DESIGN.md for the byte counts).
#include "attn_common.cuh"
#include "tc_utils.cuh"
#include "tma_utils.cuh"
#include <cuda_fp16.h>
#include <string.h>
#include <math_constants.h>

namespace meant {
namespace t2 {

constexpr uint32_t SUB_A = 128 * 128;     // 128 rows x 64 columns, SWIZZLE_128B sub-tile
constexpr uint32_t SUB_B = 128 * 64;      // 128 rows x 32 columns, SWIZZLE_64B sub-tile
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr int NC = 256;                  // compute threads (warps 0-7)

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

// the same with 16-byte loads (tables 16-byte aligned, rot_dim % 4 == 0): a quarter of the load instructions -- the
// multi-step kernel reloads the key-side table at every step
template <int RC>
__device__ __forceinline__ void load_tab4(RotTab<RC>& T, const float* cs, const float* sn, int pos, int rot_dim) {
  const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * rot_dim);
  const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * rot_dim);
#pragma unroll
  for (int i = 0; i < RC * 2; ++i) {
    float4 a = make_float4(1.f, 1.f, 1.f, 1.f), b = make_float4(0.f, 0.f, 0.f, 0.f);
    if (4 * i < rot_dim) { a = __ldg(c4 + i); b = __ldg(s4 + i); }
    T.c[2 * i] = __floats2half2_rn(a.x, a.y);
    T.c[2 * i + 1] = __floats2half2_rn(a.z, a.w);
    T.s[2 * i] = __floats2half2_rn(b.x, b.y);
    T.s[2 * i + 1] = __floats2half2_rn(b.z, b.w);
  }
}

// Rotate (INV = false) or un-rotate (transpose rotation, INV = true) the first rc chunks of one
// row r of an operand tile in place (RC <= 8: the rotated chunks all live in the SWIZZLE_128B sub-tile).
template <int RC, bool INV>
__device__ __forceinline__ void rot_chunk(uint4& raw, int ch, const RotTab<RC>& T) {
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
}

template <int RC, bool INV>
__device__ __forceinline__ void rot_row(uint8_t* tile, int r, int rc, const RotTab<RC>& T) {
  static_assert(RC <= 8, "rotated chunks must stay in the first sub-tile");
  if (rc == RC) {          // the usual case (rot_dim = 48 of Dh = 96, full Dh = 64): all loads first, no guards
    uint4 raw[RC];
#pragma unroll
    for (int ch = 0; ch < RC; ++ch) raw[ch] = *reinterpret_cast<const uint4*>(tile + tc::sw128_off(r, ch));
#pragma unroll
    for (int ch = 0; ch < RC; ++ch) {
      rot_chunk<RC, INV>(raw[ch], ch, T);
      *reinterpret_cast<uint4*>(tile + tc::sw128_off(r, ch)) = raw[ch];
    }
    return;
  }
#pragma unroll
  for (int ch = 0; ch < RC; ++ch) {
    if (ch < rc) {
      uint4* p = reinterpret_cast<uint4*>(tile + tc::sw128_off(r, ch));
      uint4 raw = *p;
      rot_chunk<RC, INV>(raw, ch, T);
      *p = raw;
    }
  }
}

// additive key bias in the exp2 domain: padding mask (meant/xPosAttention.py:54-56) and keys past Sk
__device__ __forceinline__ float key_bias2(float mval, int k, int Sk) {
  return k < Sk ? (1.0f - mval) * -1e9f * LOG2E : -CUDART_INF_F;
}

// Packed groups (GRP kernels): the TimeSformer '(b n) f d' / '(b f) n d' groups (src/meant/timesformer_pytorch.py:122)
// are much shorter than a 128-row tile (a time group is f <= 24 rows), so G consecutive groups n2 = blk*G .. of one
// n1 share a tile behind a block-diagonal mask: tile row r = g * gl + s.  One TMA box (columns, 1 head, gl rows,
// G groups) over the in-place head view lands / leaves exactly those rows, so nothing is rearranged in memory.
// The optional shared extra key (meant_attn_args.kx / vx: the cls token) is tile row xr = G * gl, written by the
// producer warp; every query row sees its own group's gl keys and that row.
struct Grp {
  int gl;      // rows per group (Sq == Sk - has_x)
  int G;       // groups per tile
  int nb;      // tiles per n1: ceil(n2_count / G)
  int xr;      // G * gl: query rows / body key rows per tile; the tile row of the extra key
  int has_x;
};

// work item -> (n1, first group n2, head); heads fastest, as in the un-grouped kernels
__device__ __forceinline__ void grp_coords(int item, int H, const Grp& g, int& n1, int& n2, int& h) {
  h = item % H;
  const int rest = item / H;
  n2 = (rest % g.nb) * g.G;
  n1 = rest / g.nb;
}

// 32-column blocks that hold a visible key for any of the rows [t0, t0 + 32): bit b <-> columns [32 b, 32 b + 32)
__device__ __forceinline__ uint32_t grp_block_mask(int t0, const Grp& g) {
  const int lo = min(t0, g.xr - 1) / g.gl * g.gl, hi = min((min(t0 + 31, g.xr - 1) / g.gl + 1) * g.gl, g.xr);
  uint32_t m = 0;
  for (int b = lo >> 5; b <= (hi - 1) >> 5; ++b) m |= 1u << b;
  if (g.has_x) m |= 1u << (g.xr >> 5);
  return m;
}

// zero row r of an operand tile (rows a packed-group box does not cover)
template <int CPR>
__device__ __forceinline__ void zero_row(uint8_t* tile, int r) {
#pragma unroll
  for (int ch = 0; ch < CPR; ++ch) *reinterpret_cast<uint4*>(tile + oper_off(r, ch)) = make_uint4(0u, 0u, 0u, 0u);
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
  static_assert(G::PBUF <= 2 * G::OPER, "P must fit in Q | K");
};

// thread roles of the forward kernel: 8 compute warps, one producer warp (TMA loads / stores, key bias), one
// MMA-issue warp -- nobody who computes ever waits for an instruction issue, nobody who issues ever computes
constexpr int F_PROD = NC / 32, F_MMA = F_PROD + 1, F_NT = NC + 64;
constexpr int BAR_PAIR = 3;              // named barriers 3..6: the two warps that share a TMEM lane quarter

struct FwdMaps { CUtensorMap m[8]; };    // q, k, v, o  x  (columns [0,64) SWIZZLE_128B, columns [64,96) SWIZZLE_64B)

// this thread's accumulator columns [c0, c0 + 8 * NCH) of row t (NCH = 4 or 6 chunks) -> bf16 -> swizzled operand tile
template <int NCH>
__device__ __forceinline__ void stage_acc_sw(uint32_t taddr, uint8_t* tile, int t, int ch0, float mul) {
  float o[32];
  tc::tmem_ld32(taddr, o);
  tc::tmem_ld_wait();
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    uint4 pk;
    __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
    for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(o[q * 8 + 2 * i] * mul, o[q * 8 + 2 * i + 1] * mul);
    *reinterpret_cast<uint4*>(tile + oper_off(t, ch0 + q)) = pk;
  }
  if (NCH > 4) {
    tc::tmem_ld16(taddr + 32, o);
    tc::tmem_ld_wait();
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      uint4 pk;
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(o[q * 8 + 2 * i] * mul, o[q * 8 + 2 * i + 1] * mul);
      *reinterpret_cast<uint4*>(tile + oper_off(t, ch0 + 4 + q)) = pk;
    }
  }
}

template <int DH, int RC, bool GRP>
__global__ void __launch_bounds__(F_NT, 1) attn2_fwd_kernel(AttnArgs a, int n_items, const __grid_constant__ FwdMaps maps,
                                                            const Grp grp) {
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
           *aux_ready = bars + 4 * NST, *rot_done = bars + 5 * NST, *p_ready = bars + 6 * NST;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (GRP) {
    // a packed-group box covers grp.xr < 128 rows: the rows below it are never written by TMA, so every operand
    // row starts out as zero (V rows stay zero for good; Q / K rows are re-zeroed by their compute threads)
    for (uint32_t i = tid * 16; i < L::AUX; i += F_NT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0u, 0u, 0u, 0u);
    tc::fence_async_smem();
  }
  if (tid == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&landed[i], 1);
      tc::mbar_init(&aux_ready[i], 32);
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&o_full[i], 1);
      tc::mbar_init(&o_staged[i], NC);
      tc::mbar_init(&rot_done[i], NC);
      tc::mbar_init(&p_ready[i], NC);
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
  // per-tile extents: query rows, keys (packed groups: G groups of gl rows, plus the shared extra key)
  const int Tq = GRP ? grp.xr : a.Sq, Tk = GRP ? grp.xr + grp.has_x : a.Sk;
  const int n_s = (Tk + 15) & ~15;                              // PV K extent (keys, rounded to one MMA K step)
  const uint32_t tile_bytes = GRP ? (uint32_t)grp.xr * DH * 2u : G::OPER;   // what one operand's boxes deliver

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    Prof pf;
    pf.init(blockIdx.x == 0 && lane == 0, 16);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    auto coords = [&](int item, int& n, int& h, int& n2, int& n1) {
      if (GRP) {
        grp_coords(item, a.H, grp, n1, n2, h);
        n = n1 * a.n2_count + n2;
      } else {
        n = item / a.H;
        h = item - n * a.H;
        n1 = n / a.n2_count;
        n2 = n - n1 * a.n2_count;
      }
    };
    auto store_out = [&](int item, int st) {                      // staged O of `item` -> global (lane 0)
      int n, h, n2, n1;
      coords(item, n, h, n2, n1);
      const uint32_t sb = s0 + st * L::STAGE;
      tma::store_box(&maps.m[6], sb, 0, h, 0, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[7], sb + SUB_A, 64, h, 0, n2, n1);
      tma::store_commit();
    };
    // the shared extra key / value row of an item: lanes [0, CPR) fetch K's 16-byte chunks, [CPR, 2 CPR) V's
    auto load_x = [&](int item) -> uint4 {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (GRP && grp.has_x && lane < 2 * G::CPR) {
        int n, h, n2, n1;
        coords(item, n, h, n2, n1);
        const HeadView& xv = lane < G::CPR ? a.kx : a.vx;
        v = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(xv.p) + (int64_t)n1 * xv.sn1 + (int64_t)h * xv.sh +
                                            (lane % G::CPR) * 8);
      }
      return v;
    };
    uint4 xnext = n_my > 0 ? load_x(blockIdx.x) : make_uint4(0u, 0u, 0u, 0u);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, st = j % NST;
      int n, h, n2, n1;
      coords(item, n, h, n2, n1);
      pf.tick(0);
      if (j >= NST) {   // the stage still holds item j - NST: its staged output goes out first
        tc::mbar_wait(&o_staged[st], ((j - NST) / NST) & 1);
        pf.tick(1);
        if (lane == 0) {
          store_out(item - NST * (int)gridDim.x, st);
          tma::store_wait_read<0>();                              // the stage may be overwritten now
        }
        pf.tick(2);
      }
      if (GRP && grp.has_x) {
        __syncwarp();                                             // lane 0's store has read the stage out
        const uint32_t sb = st * L::STAGE;
        if (lane < 2 * G::CPR)
          *reinterpret_cast<uint4*>(smem + sb + (lane < G::CPR ? 1 : 2) * G::OPER + oper_off(grp.xr, lane % G::CPR)) = xnext;
        tc::fence_async_smem();
        __syncwarp();                                             // ordered before lane 0's expect_tx arrive (release)
      }
      if (lane == 0) {
        const uint32_t sb = s0 + st * L::STAGE;
        tma::mbar_expect_tx(&landed[st], 3 * tile_bytes);
#pragma unroll
        for (int v = 0; v < 3; ++v) {
          tma::load_box(sb + v * G::OPER, &maps.m[2 * v], &landed[st], 0, h, 0, n2, n1);
          if (G::TAIL) tma::load_box(sb + v * G::OPER + SUB_A, &maps.m[2 * v + 1], &landed[st], 64, h, 0, n2, n1);
        }
      }
      if (j + 1 < n_my) xnext = load_x(item + (int)gridDim.x);    // in flight while this item's boxes are issued
      pf.tick(3);
      float4 mv = make_float4(1.f, 1.f, 1.f, 1.f);
      const int k0 = lane * 4;
      if (!GRP && a.key_mask) {
        const float* mrow = a.key_mask + (int64_t)n * a.Sk;
        if (k0 + 3 < a.Sk && (a.Sk & 3) == 0) mv = __ldg(reinterpret_cast<const float4*>(mrow + k0));
        else {
          if (k0 < a.Sk) mv.x = __ldg(mrow + k0);
          if (k0 + 1 < a.Sk) mv.y = __ldg(mrow + k0 + 1);
          if (k0 + 2 < a.Sk) mv.z = __ldg(mrow + k0 + 2);
          if (k0 + 3 < a.Sk) mv.w = __ldg(mrow + k0 + 3);
        }
      }
      reinterpret_cast<float4*>(smem + L::AUX + st * 512)[lane] =
          make_float4(key_bias2(mv.x, k0, Tk), key_bias2(mv.y, k0 + 1, Tk), key_bias2(mv.z, k0 + 2, Tk),
                      key_bias2(mv.w, k0 + 3, Tk));
      tc::mbar_arrive(&aux_ready[st]);
      pf.tick(4);
    }
    for (int j = max(0, n_my - NST); j < n_my; ++j) {
      const int st = j % NST;
      tc::mbar_wait(&o_staged[st], (j / NST) & 1);
      if (lane == 0) store_out(blockIdx.x + j * gridDim.x, st);
    }
    if (lane == 0) tma::store_wait_all();
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      auto issue_s = [&](int j) {
        const int st = j % NST;
        tc::mbar_wait(&rot_done[st], (j / NST) & 1);
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + st * 128, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
        tc::mma_commit(&s_full[st]);
      };
      if (n_my > 0) issue_s(0);
      for (int j = 0; j < n_my; ++j) {
        const int st = j % NST;
        tc::mbar_wait(&p_ready[st], (j / NST) & 1);
        tc::fence_after_sync();
        const uint32_t sp = s0 + st * L::STAGE, sv = sp + 2 * G::OPER;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks * 16 < n_s) mma_oper_mn<DH>(tmem + st * 128, desc_p_k(sp, ks), sv, ks, 0, ks > 0);
        tc::mma_commit(&o_full[st]);
        if (j + 1 < n_my) issue_s(j + 1);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;                    // TMEM lane quarter / column half
    const int t = rg * 32 + lane;                                 // query row (and the Q or K row this thread rotates)
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    float* red_max = reinterpret_cast<float*>(smem + L::RED);     // [2][128]
    float* red_sum = red_max + 256;                               // [2][128]
    int nblk = (Tk + 31) >> 5;                                    // 32-key blocks with any unmasked key for this warp's rows
    if (a.causal) nblk = min(nblk, rg + 1);
    // packed groups: this row's visible keys are [glo, ghi) and the extra key gx; bmask = blocks its warp touches
    const int glo = GRP ? min(t, grp.xr - 1) / grp.gl * grp.gl : 0, ghi = GRP ? glo + grp.gl : 0;
    const int gx = GRP && grp.has_x ? grp.xr : -1;
    const uint32_t bmask = GRP ? grp_block_mask(rg * 32, grp) : 0xffffffffu;
    RotTab<RC> tab;                                               // half 0: q tables, half 1: k tables
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_h = -1;
    if (rc && !a.pos_is_head) load_tab<RC>(tab, tcs, tsn, GRP ? min(t, grp.xr - 1) % grp.gl : min(t, (half ? a.Sk : a.Sq) - 1), a.rot_dim);
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    // rotate item j's Q / K rows in place and hand the stage to the MMA warp
    auto prepare = [&](int j) {
      const int st = j % NST;
      pf.tick(8);
      tc::mbar_wait(&landed[st], (j / NST) & 1);
      pf.tick(9);
      uint8_t* mine = smem + st * L::STAGE + half * G::OPER;
      if (GRP && t >= grp.xr) {           // rows the boxes do not cover hold the previous item's P / O staging
        if (half == 0 || t > grp.xr || !grp.has_x) zero_row<G::CPR>(mine, t);   // (the extra key row is the producer's)
      } else if (rc) {
        if (a.pos_is_head) {
          const int h = (blockIdx.x + j * gridDim.x) % a.H;
          if (h != cur_h) { load_tab<RC>(tab, tcs, tsn, h, a.rot_dim); cur_h = h; }
        }
        rot_row<RC, false>(mine, t, rc, tab);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&rot_done[st]);
      pf.tick(10);
    };
    if (n_my > 0) prepare(0);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, st = j % NST;
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
        if (blk < nblk && ((bmask >> blk) & 1u)) {                  // warp-uniform
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const bool diag = a.causal && blk == rg;                  // diagonal block: keys after the query are masked
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 b = b4[(blk * 32 + e) >> 2];
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) sv[e + u] = fmaf(sv[e + u], scale2, bb[u]);
          }
          if (GRP) {                                                // block-diagonal mask: own group + the extra key
#pragma unroll
            for (int e = 0; e < 32; ++e) {
              const int c = blk * 32 + e;
              m = fmaxf(m, ((c >= glo && c < ghi) || c == gx) ? sv[e] : -CUDART_INF_F);
            }
          } else if (diag) {                                        // warp-uniform
#pragma unroll
            for (int e = 0; e < 32; ++e) m = fmaxf(m, e > lane ? -CUDART_INF_F : sv[e]);
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) m = fmaxf(m, sv[e]);
          }
        }
      }
      red_max[half * 128 + t] = m;
      pf.tick(2);
      tc::bar_sync(BAR_PAIR + rg, 64);                              // the two halves of this row group only
      pf.tick(3);
      m = fmaxf(m, red_max[(half ^ 1) * 128 + t]);
      const float m_use = m == -CUDART_INF_F ? 0.f : m;
      float l0 = 0.f, l1 = 0.f;
      uint8_t* pbuf = smem + st * L::STAGE;                         // P over the consumed Q | K
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk && ((bmask >> blk) & 1u)) {
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const bool diag = a.causal && blk == rg;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 ba = b4[(blk * 32 + q * 8) >> 2], bc = b4[((blk * 32 + q * 8) >> 2) + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u] - m_use)), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1] - m_use));
              if (GRP) {
                const int c = blk * 32 + e;
                if (!((c >= glo && c < ghi) || c == gx)) p0 = 0.f;
                if (!((c + 1 >= glo && c + 1 < ghi) || c + 1 == gx)) p1 = 0.f;
              } else if (diag) {                                    // warp-uniform: off-diagonal blocks skip the masking
                if (e > lane) p0 = 0.f;
                if (e + 1 > lane) p1 = 0.f;
              }
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
      tc::mbar_arrive(&p_ready[st]);                                // the MMA warp starts O = P V
      pf.tick(4);
      if (j + 1 < n_my) prepare(j + 1);                             // next item's rotation hides the PV latency
      tc::bar_sync(BAR_PAIR + rg, 64);
      l += red_sum[(half ^ 1) * 128 + t];
      pf.tick(5);
      tc::mbar_wait(&o_full[st], par);
      tc::fence_after_sync();
      pf.tick(6);
      stage_acc_sw<G::CPR / 2>(ts + half * HALF, smem + st * L::STAGE, t, half * (G::CPR / 2), l > 0.f ? 1.0f / l : 0.f);
      if (half == 0 && a.lse) {
        if (GRP) {
          int n1, n2, h;
          grp_coords(item, a.H, grp, n1, n2, h);
          const int g = t / grp.gl;
          if (t < grp.xr && n2 + g < a.n2_count)
            a.lse[((int64_t)(n1 * a.n2_count + n2 + g) * a.H + h) * grp.gl + (t - g * grp.gl)] = m_use * LN2 + logf(l);
        } else if (t < a.Sq) {
          const int n = item / a.H, h = item - n * a.H;
          a.lse[((int64_t)n * a.H + h) * a.Sq + t] = m_use * LN2 + logf(l);
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&o_staged[st]);
      pf.tick(7);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== forward, several key tiles per query tile
// The same three-role pipeline for Sq or Sk > 128 (the 196-patch vision path, S = 512 text): a work item is one
// (sequence, head, 128-query tile) and runs one STEP per 128-key tile, each step being exactly a single-tile pass --
// Q | K | V boxes into a stage (Q is simply fetched again: it comes from L2), rotation in place, S = Q K^T into the
// stage's TMEM buffer, softmax out of TMEM, P over the consumed Q | K, O_j = P V over the consumed S -- except that
// the softmax is the online one: the running row max / row sum and the output accumulator (this thread's half of the
// head dimension) live in registers across the steps, O_j is folded in with the usual exp2(m_old - m_new) correction,
// and only the item's last step normalises, stages and stores.  Steps of consecutive items are pipelined exactly
// like single-tile items (NST stages in flight), so the tensor core, the TMA loads and the softmax of neighbouring
// steps overlap and nobody who computes waits for an instruction issue.  Rotation happens in the kernel: no
// pre-rotated copies of Q and K exist.  Causal problems (Sq == Sk) skip the key tiles above the diagonal; the query
// tiles are dealt so that every CTA sees a mix of short and long items (ms_item).
struct MsDims { int nq, nk; };                    // query tiles / key tiles per (sequence, head)

template <int DH>
struct FwdMsL {
  using G = Geo<DH>;
  static constexpr int NST = 3;
  static constexpr uint32_t STAGE = 3 * G::OPER;
  static constexpr uint32_t AUX = NST * STAGE;                 // [NST][128] fp32 additive key bias (exp2 domain)
  static constexpr uint32_t RED = AUX + NST * 512;             // [2][2][128] row max (double buffered) | [2][128] row sum
  static constexpr uint32_t BARS = RED + 3072;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;
  static_assert(G::PBUF <= 2 * G::OPER, "P must fit in Q | K");
};

// item id -> sequence, head, query tile.  The query tiles of one (sequence, head) are permuted by a rotation that
// depends on that pair only (a bijection per pair), pseudo-randomly, so that a CTA whose ids are congruent mod nq
// (gridDim.x a multiple of nq) still sees every query tile -- causal items differ in length.
__device__ __forceinline__ void ms_item(int id, int j, int H, int nq, int& n, int& h, int& qt) {
  (void)j;
  const int nh = id / nq;
  qt = (id - nh * nq + __popc(nh)) % nq;
  n = nh / H;
  h = nh - n * H;
}
// number of key tiles item (.., qt) visits
__device__ __forceinline__ int ms_steps(const AttnArgs& a, int qt, int nk) {
  if (!a.causal) return nk;
  const int last_q = min(qt * 128 + 128, a.Sq) - 1;
  return min(nk, last_q / 128 + 1);
}

template <int DH, int RC>
__global__ void __launch_bounds__(F_NT, 1) attn2_fwd_ms_kernel(AttnArgs a, int n_items, const __grid_constant__ FwdMaps maps,
                                                               const MsDims dm) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  using L = FwdMsL<DH>;
  using G = Geo<DH>;
  constexpr int NST = L::NST;
  constexpr int HALF = DH / 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed = bars, *s_full = bars + NST, *o_full = bars + 2 * NST, *o_staged = bars + 3 * NST,
           *aux_ready = bars + 4 * NST, *rot_done = bars + 5 * NST, *p_ready = bars + 6 * NST;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&landed[i], 1);
      tc::mbar_init(&aux_ready[i], 32);
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&o_full[i], 1);
      tc::mbar_init(&o_staged[i], NC);
      tc::mbar_init(&rot_done[i], NC);
      tc::mbar_init(&p_ready[i], NC);
    }
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = (a.qcos && !a.q_rot.p) ? a.rot_dim / 8 : 0;      // q_rot / k_rot given: the maps read those, nothing to rotate
  const uint32_t s0 = tc::smem_u32(smem);

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    // what stage st holds when it is about to be refilled: the staged output of a finished item, or nothing
    int held_n[NST], held_h[NST], held_q[NST];
#pragma unroll
    for (int i = 0; i < NST; ++i) { held_q[i] = -1; held_n[i] = 0; held_h[i] = 0; }
    auto store_out = [&](int st) {                                // lane 0
      const uint32_t sb = s0 + st * L::STAGE;
      const int n1 = held_n[st] / a.n2_count, n2 = held_n[st] - n1 * a.n2_count;
      tma::store_box(&maps.m[6], sb, 0, held_h[st], held_q[st] * 128, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[7], sb + SUB_A, 64, held_h[st], held_q[st] * 128, n2, n1);
      tma::store_commit();
    };
    int s = 0;
    for (int j = 0; j < n_my; ++j) {
      int n, h, qt;
      ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, n, h, qt);
      const int n1 = n / a.n2_count, n2 = n - n1 * a.n2_count;
      const int steps = ms_steps(a, qt, dm.nk);
      for (int kt = 0; kt < steps; ++kt, ++s) {
        const int st = s % NST;
        if (s >= NST) {
          tc::mbar_wait(&o_staged[st], ((s - NST) / NST) & 1);
          if (lane == 0 && held_q[st] >= 0) {
            store_out(st);
            tma::store_wait_read<0>();                            // the stage may be overwritten now
          }
        }
        // static indexing of the held_* arrays (they stay in registers)
#pragma unroll
        for (int i = 0; i < NST; ++i)
          if (i == st) {
            held_q[i] = kt == steps - 1 ? qt : -1;
            held_n[i] = n;
            held_h[i] = h;
          }
        if (lane == 0) {
          const uint32_t sb = s0 + st * L::STAGE;
          tma::mbar_expect_tx(&landed[st], 3 * G::OPER);
#pragma unroll
          for (int v = 0; v < 3; ++v) {
            const int r0 = v == 0 ? qt * 128 : kt * 128;
            tma::load_box(sb + v * G::OPER, &maps.m[2 * v], &landed[st], 0, h, r0, n2, n1);
            if (G::TAIL) tma::load_box(sb + v * G::OPER + SUB_A, &maps.m[2 * v + 1], &landed[st], 64, h, r0, n2, n1);
          }
        }
        float4 mv = make_float4(1.f, 1.f, 1.f, 1.f);
        const int k0 = kt * 128 + lane * 4;
        if (a.key_mask) {
          const float* mrow = a.key_mask + (int64_t)n * a.Sk;
          if (k0 + 3 < a.Sk && (a.Sk & 3) == 0) mv = __ldg(reinterpret_cast<const float4*>(mrow + k0));
          else {
            if (k0 < a.Sk) mv.x = __ldg(mrow + k0);
            if (k0 + 1 < a.Sk) mv.y = __ldg(mrow + k0 + 1);
            if (k0 + 2 < a.Sk) mv.z = __ldg(mrow + k0 + 2);
            if (k0 + 3 < a.Sk) mv.w = __ldg(mrow + k0 + 3);
          }
        }
        reinterpret_cast<float4*>(smem + L::AUX + st * 512)[lane] =
            make_float4(key_bias2(mv.x, k0, a.Sk), key_bias2(mv.y, k0 + 1, a.Sk), key_bias2(mv.z, k0 + 2, a.Sk),
                        key_bias2(mv.w, k0 + 3, a.Sk));
        tc::mbar_arrive(&aux_ready[st]);
      }
    }
    // drain: the last min(s, NST) steps' stages, oldest first
    for (int u = max(0, s - NST); u < s; ++u) {
      const int st = u % NST;
      tc::mbar_wait(&o_staged[st], (u / NST) & 1);
#pragma unroll
      for (int i = 0; i < NST; ++i)
        if (i == st && lane == 0 && held_q[i] >= 0) store_out(i);
    }
    if (lane == 0) tma::store_wait_all();
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      int total = 0;
      for (int j = 0; j < n_my; ++j) {
        int n, h, qt;
        ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, n, h, qt);
        total += ms_steps(a, qt, dm.nk);
      }
      auto issue_s = [&](int s) {
        const int st = s % NST;
        tc::mbar_wait(&rot_done[st], (s / NST) & 1);
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + st * 128, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
        tc::mma_commit(&s_full[st]);
      };
      // S runs two steps ahead of P V: S(s + 1) is complete by the time the softmax of step s is, so the compute
      // warps go from one step's softmax straight into the next one's
      if (total > 0) issue_s(0);
      if (total > 1) issue_s(1);
      for (int s = 0; s < total; ++s) {
        const int st = s % NST;
        tc::mbar_wait(&p_ready[st], (s / NST) & 1);
        tc::fence_after_sync();
        const uint32_t sp = s0 + st * L::STAGE, sv = sp + 2 * G::OPER;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) mma_oper_mn<DH>(tmem + st * 128, desc_p_k(sp, ks), sv, ks, 0, ks > 0);
        tc::mma_commit(&o_full[st]);
        if (s + 2 < total) issue_s(s + 2);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    float* red_max = reinterpret_cast<float*>(smem + L::RED);     // [2 (step parity)][2 (half)][128]
    float* red_sum = red_max + 512;                               // [2][128]
    RotTab<RC> tab;
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_pos = -1;
    const bool vec_tab = ((reinterpret_cast<uintptr_t>(tcs) | reinterpret_cast<uintptr_t>(tsn)) & 15) == 0 && (a.rot_dim & 3) == 0;
    // the step sequence of this CTA, walked twice: one step ahead (prepare) and in place (softmax)
    struct Walk {
      int j, kt, steps, n, h, qt;
    };
    auto start = [&](Walk& w, int j) {
      w.j = j;
      w.kt = 0;
      if (j < n_my) {
        ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, w.n, w.h, w.qt);
        w.steps = ms_steps(a, w.qt, dm.nk);
      }
    };
    auto advance = [&](Walk& w) {
      if (++w.kt == w.steps) start(w, w.j + 1);
    };
    // rotate the step's Q (half 0) / K (half 1) rows in place and hand the stage to the MMA warp
    auto prepare = [&](const Walk& w, int s) {
      const int st = s % NST;
      if (rc) {
        const int pos = a.pos_is_head ? w.h : (half ? min(w.kt * 128 + t, a.Sk - 1) : min(w.qt * 128 + t, a.Sq - 1));
        if (pos != cur_pos) {
          if (vec_tab) load_tab4<RC>(tab, tcs, tsn, pos, a.rot_dim);
          else load_tab<RC>(tab, tcs, tsn, pos, a.rot_dim);
          cur_pos = pos;
        }
      }
      tc::mbar_wait(&landed[st], (s / NST) & 1);
      if (rc) rot_row<RC, false>(smem + st * L::STAGE + half * G::OPER, t, rc, tab);
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&rot_done[st]);
    };
    Walk wp, wc;
    start(wp, 0);
    start(wc, 0);
    for (int u = 0; u < 2; ++u)                                   // two steps ahead of the softmax
      if (wp.j < n_my) {
        prepare(wp, u);
        advance(wp);
      }
    float m_run = -CUDART_INF_F, l_run = 0.f;
    float o_acc[HALF];
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    for (int s = 0; wc.j < n_my; ++s) {
      const int st = s % NST;
      const uint32_t par = (s / NST) & 1;
      pf.tick(0);
      const bool first = wc.kt == 0, last = wc.kt == wc.steps - 1;
      if (first) {
        m_run = -CUDART_INF_F;
        l_run = 0.f;
#pragma unroll
        for (int i = 0; i < HALF; ++i) o_acc[i] = 0.f;
      }
      const int k0 = wc.kt * 128, q0 = wc.qt * 128;
      const bool dtile = a.causal && k0 == q0;                      // the diagonal tile: keys after the query are masked
      int nblk = (min(a.Sk - k0, 128) + 31) >> 5;
      if (dtile) nblk = min(nblk, rg + 1);
      tc::mbar_wait(&aux_ready[st], par);
      tc::mbar_wait(&s_full[st], par);
      tc::fence_after_sync();
      pf.tick(1);
      const float4* b4 = reinterpret_cast<const float4*>(smem + L::AUX + st * 512);
      const uint32_t ts = lane_base + st * 128;
      float m = -CUDART_INF_F;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const bool diag = dtile && blk == rg;
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 b = b4[(blk * 32 + e) >> 2];
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) sv[e + u] = fmaf(sv[e + u], scale2, bb[u]);
          }
          if (diag) {
#pragma unroll
            for (int e = 0; e < 32; ++e) m = fmaxf(m, e > lane ? -CUDART_INF_F : sv[e]);
          } else {
#pragma unroll
            for (int e = 0; e < 32; ++e) m = fmaxf(m, sv[e]);
          }
        }
      }
      float* rmx = red_max + (s & 1) * 256;
      rmx[half * 128 + t] = m;
      pf.tick(2);
      tc::bar_sync(BAR_PAIR + rg, 64);
      pf.tick(3);
      m = fmaxf(m, rmx[(half ^ 1) * 128 + t]);
      const float m_new = fmaxf(m_run, m);
      const float m_use = m_new == -CUDART_INF_F ? 0.f : m_new;
      const float corr = tc::ex2(m_run - m_use);                    // m_run = -inf -> 0
      m_run = m_new;
      float l0 = 0.f, l1 = 0.f;
      uint8_t* pbuf = smem + st * L::STAGE;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const bool diag = dtile && blk == rg;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 ba = b4[(blk * 32 + q * 8) >> 2], bc = b4[((blk * 32 + q * 8) >> 2) + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u] - m_use)), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1] - m_use));
              if (diag) {
                if (e > lane) p0 = 0.f;
                if (e + 1 > lane) p1 = 0.f;
              }
              l0 += p0;
              l1 += p1;
              ph[u] = __floats2bfloat162_rn(p0, p1);
            }
            *reinterpret_cast<uint4*>(pbuf + pbuf_off(t, blk * 4 + q)) = pk;
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(pbuf + pbuf_off(t, blk * 4 + q)) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      l_run = fmaf(l_run, corr, l0 + l1);
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&p_ready[st]);                                // the MMA warp starts O_j = P V
      pf.tick(4);
      if (wp.j < n_my) {                                            // rotation of step s + 2 hides the P V latency of step s
        prepare(wp, s + 2);
        advance(wp);
      }
      pf.tick(5);
      float l = l_run;
      if (last) {
        red_sum[half * 128 + t] = l_run;
        tc::bar_sync(BAR_PAIR + rg, 64);
        l += red_sum[(half ^ 1) * 128 + t];
      }
      tc::mbar_wait(&o_full[st], par);
      tc::fence_after_sync();
      pf.tick(6);
      {
        float pv[HALF];
        tc::tmem_ld32(ts + half * HALF, pv);
        if constexpr (HALF > 32) tc::tmem_ld16(ts + half * HALF + 32, pv + 32);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < HALF; ++i) o_acc[i] = fmaf(o_acc[i], corr, pv[i]);
      }
      if (last) {
        const float inv = l > 0.f ? 1.0f / l : 0.f;
        uint8_t* tile = smem + st * L::STAGE;
#pragma unroll
        for (int q = 0; q < HALF / 8; ++q) {
          uint4 pk;
          __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(o_acc[q * 8 + 2 * i] * inv, o_acc[q * 8 + 2 * i + 1] * inv);
          *reinterpret_cast<uint4*>(tile + oper_off(t, half * (HALF / 8) + q)) = pk;
        }
        if (half == 0 && a.lse && q0 + t < a.Sq)
          a.lse[((int64_t)wc.n * a.H + wc.h) * a.Sq + q0 + t] = m_use * LN2 + logf(l);
        tc::fence_async_smem();
      }
      tc::fence_before_sync();
      tc::mbar_arrive(&o_staged[st]);
      pf.tick(7);
      advance(wc);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== forward, one wide key tile
// 128 < Sk <= 208 (the 196-patch vision path): every key of a (sequence, head) fits ONE tile of 208 rows, so a work
// item -- one (sequence, head, 128-query tile) -- is a single-tile pass again: Q | K | V boxes into one of two stages,
// ONE S = Q K^T of N = Sk rounded up to whole 32-column blocks (at most 208) into the stage's 256 TMEM columns, a
// plain two-pass softmax out of TMEM (no running max, no correction of the output accumulator, no fold through
// registers), P [128 x 224] over the
// consumed Q | K, O = P V over the consumed S, normalise, stage, TMA-store.  Against the multi-step kernel on the
// same problem this halves the steps, never touches the padded second key tile's 60 empty columns and drops the
// online-softmax arithmetic.
// Rotation is not done here: q_rot / k_rot (meant_attn_prerotate, or in place) or no tables at all.
template <int DH>
struct FwdWkL {
  using G = Geo<DH>;
  static constexpr int KW = 208;                                // key rows of a K / V tile: 13 MMA K steps
  static constexpr int NST = 2;
  static constexpr uint32_t KA = KW * 128u;                     // columns [0, 64): SWIZZLE_128B rows
  static constexpr uint32_t KB = G::TAIL ? KW * 64u : 0u;       // columns [64, 96): SWIZZLE_64B rows
  static constexpr uint32_t OPERK = KA + KB;
  static constexpr uint32_t PW = 3 * SUB_A + SUB_B;             // P: key columns [0, 192) in three sub-tiles, [192, 224) in a fourth
  static constexpr bool P_ALIAS = PW <= G::OPER + OPERK;        // P over the consumed Q | K; else one separate P tile
  static constexpr uint32_t STAGE = G::OPER + 2 * OPERK;        // Q | K | V
  static constexpr uint32_t PSEP = NST * STAGE;
  static constexpr uint32_t AUX = PSEP + (P_ALIAS ? 0u : PW);   // [NST][256] fp32 additive key bias (exp2 domain)
  static constexpr uint32_t RED = AUX + NST * 1024;             // [2 (item parity)][2 (half)][128] row max | the same for the row sum
  static constexpr uint32_t BARS = RED + 4096;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;
  static_assert(TOTAL <= 227 * 1024, "shared memory");
  static_assert(KA % 1024 == 0 && KB % 1024 == 0 && STAGE % 1024 == 0, "swizzle atoms are 1024-byte aligned");
};

// chunk c (8 key columns) of row r of the wide P tile
__device__ __forceinline__ uint32_t pw_off(int r, int c) {
  return c < 24 ? (c >> 3) * SUB_A + tc::sw128_off(r, c & 7) : 3 * SUB_A + tc::sw64_off(r, c - 24);
}
// K-major view of the wide P tile, K step ks (16 key columns)
__device__ __forceinline__ uint64_t desc_pw_k(uint32_t base, int ks) {
  return ks < 12 ? tc::make_desc_sw(base + (ks >> 2) * SUB_A + (ks & 3) * 32, 16, 1024, 2) : tc::make_desc_sw(base + 3 * SUB_A, 16, 512, 4);
}

template <int DH>
__global__ void __launch_bounds__(F_NT, 1) attn2_fwd_wk_kernel(AttnArgs a, int n_items, const __grid_constant__ FwdMaps maps,
                                                               const MsDims dm) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  using L = FwdWkL<DH>;
  using G = Geo<DH>;
  constexpr int NST = L::NST;
  constexpr int HALF = DH / 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed = bars, *s_full = bars + NST, *o_full = bars + 2 * NST, *o_staged = bars + 3 * NST,
           *aux_ready = bars + 4 * NST, *p_ready = bars + 5 * NST;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // V rows in [Sk, ceil16(Sk)) are read by the last P V step and never written by a box: everything starts as zero
  for (uint32_t i = tid * 16; i < L::AUX; i += F_NT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0u, 0u, 0u, 0u);
  tc::fence_async_smem();
  if (tid == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&landed[i], 1);
      tc::mbar_init(&aux_ready[i], 32);
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&o_full[i], 1);
      tc::mbar_init(&o_staged[i], NC);
      tc::mbar_init(&p_ready[i], NC);
    }
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const uint32_t s0 = tc::smem_u32(smem);
  const int n_s = min((a.Sk + 31) & ~31, L::KW);                   // S columns / P V K extent: whole 32-column blocks (the softmax reads them)
  // item -> (sequence, head, query tile): the query tiles of one (sequence, head) are neighbours (K, V come from L2)
  auto coords = [&](int item, int& n1, int& n2, int& h, int& qt) {
    qt = item % dm.nq;
    const int nh = item / dm.nq;
    const int n = nh / a.H;
    h = nh - n * a.H;
    n1 = n / a.n2_count;
    n2 = n - n1 * a.n2_count;
  };

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    auto store_out = [&](int item, int st) {                       // lane 0
      int n1, n2, h, qt;
      coords(item, n1, n2, h, qt);
      const uint32_t sb = s0 + st * L::STAGE;
      tma::store_box(&maps.m[6], sb, 0, h, qt * 128, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[7], sb + SUB_A, 64, h, qt * 128, n2, n1);
      tma::store_commit();
    };
    const uint32_t kv_bytes = (uint32_t)a.Sk * DH * 2u;
    Prof pf;
    pf.init(blockIdx.x == 0 && lane == 0, 16);
    for (int j = 0; j < n_my; ++j) {
      pf.tick(0);
      const int item = blockIdx.x + j * gridDim.x, st = j % NST;
      int n1, n2, h, qt;
      coords(item, n1, n2, h, qt);
      const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER, sv = sk + L::OPERK;
      if (j >= NST) {                                              // the stage still holds item j - NST's staged output
        tc::mbar_wait(&o_staged[st], ((j - NST) / NST) & 1);
        pf.tick(1);
        if (lane == 0) {
          store_out(item - NST * (int)gridDim.x, st);
          tma::store_wait_read<0>();
        }
      }
      pf.tick(2);
      // (issuing the K | V boxes earlier -- their half of the stage is dead once P V has run -- measured slower: the
      // output store then queues behind them)
      if (lane == 0) {
        tma::mbar_expect_tx(&landed[st], G::OPER + 2 * kv_bytes);
        tma::load_box(sq, &maps.m[0], &landed[st], 0, h, qt * 128, n2, n1);
        if (G::TAIL) tma::load_box(sq + SUB_A, &maps.m[1], &landed[st], 64, h, qt * 128, n2, n1);
        tma::load_box(sk, &maps.m[2], &landed[st], 0, h, 0, n2, n1);
        if (G::TAIL) tma::load_box(sk + L::KA, &maps.m[3], &landed[st], 64, h, 0, n2, n1);
        tma::load_box(sv, &maps.m[4], &landed[st], 0, h, 0, n2, n1);
        if (G::TAIL) tma::load_box(sv + L::KA, &maps.m[5], &landed[st], 64, h, 0, n2, n1);
      }
      float* bias = reinterpret_cast<float*>(smem + L::AUX + st * 1024);
      const float* mrow = a.key_mask ? a.key_mask + (int64_t)(n1 * a.n2_count + n2) * a.Sk : nullptr;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = u * 32 + lane;
        bias[k] = key_bias2(mrow && k < a.Sk ? __ldg(mrow + k) : 1.f, k, a.Sk);
      }
      tc::mbar_arrive(&aux_ready[st]);
      pf.tick(3);
    }
    for (int j = max(0, n_my - NST); j < n_my; ++j) {
      const int st = j % NST;
      tc::mbar_wait(&o_staged[st], (j / NST) & 1);
      if (lane == 0) store_out(blockIdx.x + j * gridDim.x, st);
    }
    if (lane == 0) tma::store_wait_all();
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      const uint32_t idesc_s = tc::make_idesc(128, n_s, 0, 0);
      auto issue_s = [&](int j) {
        const int st = j % NST;
        tc::mbar_wait(&landed[st], (j / NST) & 1);
        tc::fence_after_sync();
        const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) {
          const uint64_t bd = ks < 4 ? tc::make_desc_sw(sk + ks * 32, 16, 1024, 2) : tc::make_desc_sw(sk + L::KA + (ks - 4) * 32, 16, 512, 4);
          tc::mma_ss(tmem + st * 256, desc_oper_k(sq, ks), bd, idesc_s, ks > 0);
        }
        tc::mma_commit(&s_full[st]);
      };
      Prof pf;
      pf.init(blockIdx.x == 0, 32);
      if (n_my > 0) issue_s(0);
      for (int j = 0; j < n_my; ++j) {
        const int st = j % NST;
        pf.tick(0);
        tc::mbar_wait(&p_ready[st], (j / NST) & 1);
        pf.tick(1);
        tc::fence_after_sync();
        const uint32_t sp = L::P_ALIAS ? s0 + st * L::STAGE : s0 + L::PSEP;
        const uint32_t sv = s0 + st * L::STAGE + G::OPER + L::OPERK;
        for (int ks = 0; ks * 16 < n_s; ++ks) {
          const uint64_t ad = desc_pw_k(sp, ks);
          tc::mma_ss(tmem + st * 256, ad, tc::make_desc_sw(sv + ks * 2048, SUB_A, 1024, 2), tc::make_idesc(128, 64, 0, 1), ks > 0);
          if (G::TAIL)
            tc::mma_ss(tmem + st * 256 + 64, ad, tc::make_desc_sw(sv + L::KA + ks * 1024, SUB_B, 512, 4), tc::make_idesc(128, 32, 0, 1), ks > 0);
        }
        tc::mma_commit(&o_full[st]);
        pf.tick(3);
        // S of the next item behind P V: it is complete by the time the compute warps have staged this item's output.
        // (Issued in front of P V -- blocking, or polled together with p_ready -- it measured 111 / 92 us against 89:
        // P is usually ready before the next boxes are, and a polling MMA thread takes issue slots from two softmax warps)
        if (j + 1 < n_my) issue_s(j + 1);
        pf.tick(4);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;                     // TMEM lane quarter / which 32-column blocks
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    float* red_max = reinterpret_cast<float*>(smem + L::RED);      // [2][2][128]
    float* red_sum = red_max + 512;                                // [2][2][128]
    const int nblk = (a.Sk + 31) >> 5;                             // <= 7; block 6 holds 16 columns
    // scores of block blk (32 key columns; block 6: 16 + 16 masked), scaled, biased
    auto load_blk = [&](uint32_t ts, const float* bias, int blk, float* sv) {
      if (blk < 6) tc::tmem_ld32(ts + blk * 32, sv);
      else {
        tc::tmem_ld16(ts + 192, sv);
#pragma unroll
        for (int e = 16; e < 32; ++e) sv[e] = 0.f;
      }
      tc::tmem_ld_wait();
      const float4* b4 = reinterpret_cast<const float4*>(bias + blk * 32);
#pragma unroll
      for (int e = 0; e < 32; e += 4) {
        const float4 b = b4[e >> 2];
        sv[e] = fmaf(sv[e], scale2, b.x);
        sv[e + 1] = fmaf(sv[e + 1], scale2, b.y);
        sv[e + 2] = fmaf(sv[e + 2], scale2, b.z);
        sv[e + 3] = fmaf(sv[e + 3], scale2, b.w);
      }
    };
    // row max of item j (both halves)
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    auto row_max = [&](int j) -> float {
      const int st = j % NST;
      const uint32_t par = (j / NST) & 1;
      pf.tick(0);
      tc::mbar_wait(&aux_ready[st], par);
      tc::mbar_wait(&s_full[st], par);
      pf.tick(1);
      tc::fence_after_sync();
      const float* bias = reinterpret_cast<const float*>(smem + L::AUX + st * 1024);
      const uint32_t ts = lane_base + st * 256;
      float m = -CUDART_INF_F;
      // (letting the warps whose 32 query rows all lie past Sq -- one of four in the last tile at Sq = 196 -- skip the
      // arithmetic measured slower, not faster)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32];
          load_blk(ts, bias, blk, sv);
#pragma unroll
          for (int e = 0; e < 32; ++e) m = fmaxf(m, sv[e]);
        }
      }
      float* rmx = red_max + (j & 1) * 256;
      rmx[half * 128 + t] = m;
      pf.tick(2);
      tc::bar_sync(BAR_PAIR + rg, 64);
      pf.tick(3);
      return fmaxf(m, rmx[(half ^ 1) * 128 + t]);
    };
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, st = j % NST;
      const float m_cur = row_max(j);
      const uint32_t par = (j / NST) & 1;
      const float* bias = reinterpret_cast<const float*>(smem + L::AUX + st * 1024);
      const uint32_t ts = lane_base + st * 256;
      const float m_use = m_cur == -CUDART_INF_F ? 0.f : m_cur;
      float l0 = 0.f, l1 = 0.f;
      uint8_t* pbuf = L::P_ALIAS ? smem + st * L::STAGE : smem + L::PSEP;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
          float sv[32];
          load_blk(ts, bias, blk, sv);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (blk == 6 && q >= 2) break;                          // columns [208, 224) do not exist
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              const float p0 = tc::ex2(sv[e] - m_use), p1 = tc::ex2(sv[e + 1] - m_use);
              l0 += p0;
              l1 += p1;
              ph[u] = __floats2bfloat162_rn(p0, p1);
            }
            *reinterpret_cast<uint4*>(pbuf + pw_off(t, blk * 4 + q)) = pk;
          }
        }
      }
      float l = l0 + l1;
      float* rsm = red_sum + (j & 1) * 256;
      rsm[half * 128 + t] = l;
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&p_ready[st]);                                // the MMA warp starts O = P V
      pf.tick(4);
      tc::bar_sync(BAR_PAIR + rg, 64);
      l += rsm[(half ^ 1) * 128 + t];
      pf.tick(5);
      tc::mbar_wait(&o_full[st], par);
      pf.tick(6);
      tc::fence_after_sync();
      stage_acc_sw<G::CPR / 2>(ts + half * HALF, smem + st * L::STAGE, t, half * (G::CPR / 2), l > 0.f ? 1.0f / l : 0.f);
      if (half == 0 && a.lse) {
        const int qt = item % dm.nq, nh = item / dm.nq;
        if (qt * 128 + t < a.Sq) a.lse[(int64_t)nh * a.Sq + qt * 128 + t] = m_use * LN2 + logf(l);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&o_staged[st]);
      pf.tick(7);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== backward, several tiles: dQ pass
// The multi-step pipeline again, for the gradient of the queries when Sq or Sk > 128.  An item is a (sequence, head,
// 128-query tile), a step one 128-key tile: Q | dO | K | V boxes into one of two stages, S = Q K^T and dP = dO V^T into
// the stage's two TMEM buffers (issued one step ahead, under the previous step's arithmetic), P = exp2(S * scale - LSE)
// and dS = P (dP - delta) * scale out of TMEM in ONE pass (the row statistics are known: no max, no exchange between
// the two threads of a row), dS (bf16) into its own buffer, dQ_j = dS K over the consumed S, folded into this thread's
// half of the dQ row in registers.  Q, dO and V of a stage are dead as soon as S | dP have run, so their next boxes are
// issued then -- a whole step before K's, which has to wait for dQ_j.  The item's last step un-rotates (transpose
// rotation, table fetched under the last MMA), stages over the consumed K and TMA-stores.
// delta = rowsum(dO * O) comes from attn_delta_kernel (below), which runs first.
template <int DH>
struct BwdDqL {
  using G = Geo<DH>;
  // Buffers are recycled as early as their last reader allows: Q | dO (two pairs) and V (one) are dead once S | dP of
  // their step have run -- a whole step before that step's arithmetic -- while K lives until dQ_j = dS K at the end of
  // the step, so K gets three buffers: the boxes of step s + 2 are in flight while step s computes.
  static constexpr uint32_t QDO = 0;                           // [2][Q | dO]
  static constexpr uint32_t V = 4 * G::OPER;                   // [1]
  static constexpr uint32_t K = 5 * G::OPER;                   // [3]   (the dQ staging of an item's last step aliases its K)
  static constexpr uint32_t DS = 8 * G::OPER;                  // one dS tile: dQ_j has consumed it before the next one is written
  static constexpr uint32_t AUX = DS + G::PBUF;                // [2][128] fp32 additive key bias (exp2 domain)
  static constexpr uint32_t BARS = AUX + 2 * 512;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;
  static_assert(TOTAL <= 232448, "exceeds the 227 KB of shared memory a CTA can opt in to");
};

struct BwdMsMaps { CUtensorMap m[14]; };   // q, k, v, dO (loads), dq, dk, dv (stores)  x  (SWIZZLE_128B part, SWIZZLE_64B part)

// un-rotate (transpose rotation) 8 fp32 gradient values of chunk ch in place
template <int RC>
__device__ __forceinline__ void unrot8(float* g, int ch, const RotTab<RC>& T) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 c = __half22float2(T.c[ch * 4 + i]), s = __half22float2(T.s[ch * 4 + i]);
    const float y0 = g[2 * i] * c.x + g[2 * i + 1] * s.y, y1 = g[2 * i + 1] * c.y - g[2 * i] * s.x;
    g[2 * i] = y0;
    g[2 * i + 1] = y1;
  }
}
__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];\n" ::"l"(p)); }

// ROT_IN: rotate Q / K on the way in (no pre-rotated copies were passed); the table registers exist only then
template <int DH, int RC, bool ROT_IN>
__global__ void __launch_bounds__(F_NT, 1) attn2_bwd_dq_ms_kernel(AttnArgs a, int n_items, const __grid_constant__ BwdMsMaps maps,
                                                                  const MsDims dm) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  using L = BwdDqL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed_qdo = bars, *landed_k = bars + 2, *landed_v = bars + 5, *sdp_full = bars + 6, *dq_full = bars + 8,
           *freed = bars + 10, *aux_ready = bars + 13, *rot_done = bars + 15, *ds_ready = bars + 17;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&landed_qdo[i], 1);
      tc::mbar_init(&aux_ready[i], 32);
      tc::mbar_init(&sdp_full[i], 1);
      tc::mbar_init(&dq_full[i], 1);
      tc::mbar_init(&rot_done[i], NC);
      tc::mbar_init(&ds_ready[i], NC);
    }
    for (int i = 0; i < 3; ++i) {
      tc::mbar_init(&landed_k[i], 1);
      tc::mbar_init(&freed[i], NC);
    }
    tc::mbar_init(landed_v, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = a.qcos ? a.rot_dim / 8 : 0;                    // chunks to un-rotate in dQ
  const int rc_in = ROT_IN ? rc : 0;                            // chunks to rotate on the way in
  const uint32_t s0 = tc::smem_u32(smem);

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 10; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    int held_n[3], held_h[3], held_q[3];                          // what K buffer kb stages: a finished item's dQ, or nothing
#pragma unroll
    for (int i = 0; i < 3; ++i) { held_q[i] = -1; held_n[i] = 0; held_h[i] = 0; }
    auto store_out = [&](int kb) {                                // lane 0
      const uint32_t sb = s0 + L::K + kb * G::OPER;
      const int n1 = held_n[kb] / a.n2_count, n2 = held_n[kb] - n1 * a.n2_count;
      tma::store_box(&maps.m[8], sb, 0, held_h[kb], held_q[kb] * 128, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[9], sb + SUB_A, 64, held_h[kb], held_q[kb] * 128, n2, n1);
      tma::store_commit();
    };
    int s = 0;
    for (int j = 0; j < n_my; ++j) {
      int n, h, qt;
      ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, n, h, qt);
      const int n1 = n / a.n2_count, n2 = n - n1 * a.n2_count;
      const int steps = ms_steps(a, qt, dm.nk);
      for (int kt = 0; kt < steps; ++kt, ++s) {
        const int b2 = s & 1, kb = s % 3;
        auto load_tile = [&](uint32_t dst, int view, int r0, uint64_t* bar) {
          tma::load_box(dst, &maps.m[2 * view], bar, 0, h, r0, n2, n1);
          if (G::TAIL) tma::load_box(dst + SUB_A, &maps.m[2 * view + 1], bar, 64, h, r0, n2, n1);
        };
        // Q | dO pair b2: free once S | dP of step s - 2 have run
        if (s >= 2) tc::mbar_wait(&sdp_full[b2], ((s - 2) >> 1) & 1);
        if (lane == 0) {
          tma::mbar_expect_tx(&landed_qdo[b2], 2 * G::OPER);
          load_tile(s0 + L::QDO + b2 * 2 * G::OPER, 0, qt * 128, &landed_qdo[b2]);
          load_tile(s0 + L::QDO + b2 * 2 * G::OPER + G::OPER, 3, qt * 128, &landed_qdo[b2]);
        }
        // V: free once dP of step s - 1 has run
        if (s >= 1) tc::mbar_wait(&sdp_full[b2 ^ 1], ((s - 1) >> 1) & 1);
        if (lane == 0) {
          tma::mbar_expect_tx(landed_v, G::OPER);
          load_tile(s0 + L::V, 2, kt * 128, landed_v);
        }
        // K buffer kb: free once step s - 3 is over (dQ_j read it; a last step staged its dQ there: store it first)
        if (s >= 3) {
          tc::mbar_wait(&freed[kb], ((s - 3) / 3) & 1);
          if (lane == 0 && held_q[kb] >= 0) {
            store_out(kb);
            tma::store_wait_read<0>();
          }
        }
        held_q[kb] = kt == steps - 1 ? qt : -1;
        held_n[kb] = n;
        held_h[kb] = h;
        if (lane == 0) {
          tma::mbar_expect_tx(&landed_k[kb], G::OPER);
          load_tile(s0 + L::K + kb * G::OPER, 1, kt * 128, &landed_k[kb]);
        }
        // key bias slot b2: the arithmetic of step s - 2 has read it
        if (s >= 2) tc::mbar_wait(&freed[(s - 2) % 3], ((s - 2) / 3) & 1);
        float4 mv = make_float4(1.f, 1.f, 1.f, 1.f);
        const int k0 = kt * 128 + lane * 4;
        if (a.key_mask) {
          const float* mrow = a.key_mask + (int64_t)n * a.Sk;
          if (k0 + 3 < a.Sk && (a.Sk & 3) == 0) mv = __ldg(reinterpret_cast<const float4*>(mrow + k0));
          else {
            if (k0 < a.Sk) mv.x = __ldg(mrow + k0);
            if (k0 + 1 < a.Sk) mv.y = __ldg(mrow + k0 + 1);
            if (k0 + 2 < a.Sk) mv.z = __ldg(mrow + k0 + 2);
            if (k0 + 3 < a.Sk) mv.w = __ldg(mrow + k0 + 3);
          }
        }
        reinterpret_cast<float4*>(smem + L::AUX + b2 * 512)[lane] =
            make_float4(key_bias2(mv.x, k0, a.Sk), key_bias2(mv.y, k0 + 1, a.Sk), key_bias2(mv.z, k0 + 2, a.Sk),
                        key_bias2(mv.w, k0 + 3, a.Sk));
        tc::mbar_arrive(&aux_ready[b2]);
      }
    }
    for (int u = max(0, s - 3); u < s; ++u) {
      const int kb = u % 3;
      tc::mbar_wait(&freed[kb], (u / 3) & 1);
      if (lane == 0 && held_q[kb] >= 0) store_out(kb);
    }
    if (lane == 0) tma::store_wait_all();
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      int total = 0;
      for (int j = 0; j < n_my; ++j) {
        int n, h, qt;
        ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, n, h, qt);
        total += ms_steps(a, qt, dm.nk);
      }
      auto issue_sdp = [&](int s) {
        const int b2 = s & 1;
        tc::mbar_wait(&rot_done[b2], (s >> 1) & 1);
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER, sdo = sq + G::OPER, sk = s0 + L::K + (s % 3) * G::OPER, sv = s0 + L::V;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + b2 * 256, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + b2 * 256 + 128, desc_oper_k(sdo, ks), desc_oper_k(sv, ks), idesc, ks > 0);
        tc::mma_commit(&sdp_full[b2]);
      };
      if (total > 0) issue_sdp(0);
      for (int s = 0; s < total; ++s) {
        const int b2 = s & 1;
        if (s + 1 < total) issue_sdp(s + 1);                      // one step ahead: runs under this step's dS arithmetic
        tc::mbar_wait(&ds_ready[b2], (s >> 1) & 1);
        tc::fence_after_sync();
        const uint32_t sds = s0 + L::DS, sk = s0 + L::K + (s % 3) * G::OPER;
#pragma unroll
        for (int ks = 0; ks < 8; ++ks) mma_oper_mn<DH>(tmem + b2 * 256, desc_p_k(sds, ks), sk, ks, 0, ks > 0);
        tc::mma_commit(&dq_full[b2]);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    RotTab<RC> tab;                                               // rotation on the way in: half 0 q tables, half 1 k tables
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_pos = -1;
    const bool vec_tab = ((reinterpret_cast<uintptr_t>(a.qcos) | reinterpret_cast<uintptr_t>(a.qsin) | reinterpret_cast<uintptr_t>(a.kcos) |
                           reinterpret_cast<uintptr_t>(a.ksin)) & 15) == 0 && (a.rot_dim & 3) == 0;
    // the two threads of a row take the 8-column chunks of dQ alternately (c = 2 i + half): both share the rotated ones
    constexpr int NCH = G::CPR / 2, RCT = (RC + 1) / 2;           // chunks per thread / of them at most rotated
    struct Walk {
      int j, kt, steps, n, h, qt;
    };
    auto start = [&](Walk& w, int j) {
      w.j = j;
      w.kt = 0;
      if (j < n_my) {
        ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nq, w.n, w.h, w.qt);
        w.steps = ms_steps(a, w.qt, dm.nk);
      }
    };
    auto advance = [&](Walk& w) {
      if (++w.kt == w.steps) start(w, w.j + 1);
    };
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    float lse2_n = 0.f, dlt_n = 0.f;                              // row statistics of the walker's item, fetched one step ahead
    auto prepare = [&](const Walk& w, int s) {
      const int b2 = s & 1, kb = s % 3;
      if (w.kt == 0) {
        const int row = min(w.qt * 128 + t, a.Sq - 1);
        const int64_t si = ((int64_t)w.n * a.H + w.h) * a.Sq + row;
        lse2_n = __ldg(a.lse + si);
        dlt_n = __ldg(a.delta + si);
      }
      if (ROT_IN && rc_in) {
        const int pos = a.pos_is_head ? w.h : (half ? min(w.kt * 128 + t, a.Sk - 1) : min(w.qt * 128 + t, a.Sq - 1));
        if (pos != cur_pos) {
          if (vec_tab) load_tab4<RC>(tab, tcs, tsn, pos, a.rot_dim);
          else load_tab<RC>(tab, tcs, tsn, pos, a.rot_dim);
          cur_pos = pos;
        }
      }
      pf.tick(6);
      tc::mbar_wait(&landed_qdo[b2], (s >> 1) & 1);
      pf.tick(7);
      tc::mbar_wait(&landed_k[kb], (s / 3) & 1);
      pf.tick(8);
      tc::mbar_wait(landed_v, s & 1);
      pf.tick(9);
      if (ROT_IN && rc_in) rot_row<RC, false>(smem + (half ? L::K + kb * G::OPER : L::QDO + b2 * 2 * G::OPER), t, rc_in, tab);
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&rot_done[b2]);
    };
    Walk wp, wc;
    start(wp, 0);
    start(wc, 0);
    if (wp.j < n_my) {
      prepare(wp, 0);
      advance(wp);
    }
    float lse2 = 0.f, dlt = 0.f;
    float dq_acc[NCH][8];
    for (int s = 0; wc.j < n_my; ++s) {
      const int b2 = s & 1, kb = s % 3;
      const uint32_t par = (s >> 1) & 1;
      const bool first = wc.kt == 0, last = wc.kt == wc.steps - 1;
      const int k0 = wc.kt * 128, q0 = wc.qt * 128;
      if (first) {
#pragma unroll
        for (int i = 0; i < NCH; ++i)
#pragma unroll
          for (int e = 0; e < 8; ++e) dq_acc[i][e] = 0.f;
        lse2 = lse2_n * LOG2E;                                      // fetched when the walker entered this item
        dlt = dlt_n;
      }
      if (wp.j < n_my) {                                            // next step's operands: S | dP (s + 1) run under this step
        prepare(wp, s + 1);
        advance(wp);
      }
      const bool dtile = a.causal && k0 == q0;
      int nblk = (min(a.Sk - k0, 128) + 31) >> 5;
      if (dtile) nblk = min(nblk, rg + 1);
      pf.tick(0);
      tc::mbar_wait(&aux_ready[b2], par);
      tc::mbar_wait(&sdp_full[b2], par);
      tc::fence_after_sync();
      pf.tick(1);
      const float4* b4 = reinterpret_cast<const float4*>(smem + L::AUX + b2 * 512);
      const uint32_t ts = lane_base + b2 * 256;
      uint8_t* dsbuf = smem + L::DS;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32], dp[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld32(ts + 128 + blk * 32, dp);
          tc::tmem_ld_wait();
          const bool diag = dtile && blk == rg;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 ba = b4[(blk * 32 + q * 8) >> 2], bc = b4[((blk * 32 + q * 8) >> 2) + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u] - lse2)), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1] - lse2));
              if (diag) {
                if (e > lane) p0 = 0.f;
                if (e + 1 > lane) p1 = 0.f;
              }
              ph[u] = __floats2bfloat162_rn(p0 * (dp[e] - dlt) * a.scale, p1 * (dp[e + 1] - dlt) * a.scale);
            }
            *reinterpret_cast<uint4*>(dsbuf + pbuf_off(t, blk * 4 + q)) = pk;
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(dsbuf + pbuf_off(t, blk * 4 + q)) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&ds_ready[b2]);                               // the MMA warp starts dQ_j = dS K
      pf.tick(2);
      // the un-rotation tables of this thread's chunks: fetched under the item's last dQ_j MMAs (fetched after them, by
      // the thread that held every rotated column, they cost as much as a step)
      float cv[RCT][8], sn[RCT][8];
      if (last) {
        const int pos = a.pos_is_head ? wc.h : min(q0 + t, a.Sq - 1);
        const float* cb = a.qcos + (int64_t)pos * a.rot_dim;
        const float* sb = a.qsin + (int64_t)pos * a.rot_dim;
#pragma unroll
        for (int i = 0; i < RCT; ++i) {
          const int c = 2 * i + half;
          if (c < rc) {
            const float4 c0 = __ldg(reinterpret_cast<const float4*>(cb + c * 8)), c1 = __ldg(reinterpret_cast<const float4*>(cb + c * 8 + 4));
            const float4 z0 = __ldg(reinterpret_cast<const float4*>(sb + c * 8)), z1 = __ldg(reinterpret_cast<const float4*>(sb + c * 8 + 4));
            cv[i][0] = c0.x; cv[i][1] = c0.y; cv[i][2] = c0.z; cv[i][3] = c0.w; cv[i][4] = c1.x; cv[i][5] = c1.y; cv[i][6] = c1.z; cv[i][7] = c1.w;
            sn[i][0] = z0.x; sn[i][1] = z0.y; sn[i][2] = z0.z; sn[i][3] = z0.w; sn[i][4] = z1.x; sn[i][5] = z1.y; sn[i][6] = z1.z; sn[i][7] = z1.w;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) { cv[i][e] = 1.f; sn[i][e] = 0.f; }
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < RCT; ++i)
#pragma unroll
          for (int e = 0; e < 8; ++e) { cv[i][e] = 1.f; sn[i][e] = 0.f; }
      }
      tc::mbar_wait(&dq_full[b2], par);
      tc::fence_after_sync();
      pf.tick(3);
      {
        float g[NCH][8];
#pragma unroll
        for (int i = 0; i < NCH; ++i) tc::tmem_ld8(ts + (2 * i + half) * 8, g[i]);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < NCH; ++i)
#pragma unroll
          for (int e = 0; e < 8; ++e) dq_acc[i][e] += g[i][e];
      }
      pf.tick(4);
      if (last) {
        // un-rotate (transpose rotation; identity tables past rc), bf16, staged over the consumed K
#pragma unroll
        for (int i = 0; i < RCT; ++i) {
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float y0 = dq_acc[i][e] * cv[i][e] + dq_acc[i][e + 1] * sn[i][e + 1], y1 = dq_acc[i][e + 1] * cv[i][e + 1] - dq_acc[i][e] * sn[i][e];
            dq_acc[i][e] = y0;
            dq_acc[i][e + 1] = y1;
          }
        }
        uint8_t* tile = smem + L::K + kb * G::OPER;
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          uint4 pk;
          __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(dq_acc[i][2 * e], dq_acc[i][2 * e + 1]);
          *reinterpret_cast<uint4*>(tile + oper_off(t, 2 * i + half)) = pk;
        }
        tc::fence_async_smem();
      }
      tc::fence_before_sync();
      tc::mbar_arrive(&freed[kb]);
      pf.tick(5);
      advance(wc);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== backward, one wide key tile: dQ pass
// The dQ pass of the multi-step backward for the shapes of the wide-key forward (128 < Sk <= 208, not causal,
// rotation already applied): an item is one (sequence, head, 128-query tile) and sees every key at once -- Q | dO
// boxes into one of two pairs, the WHOLE K into one of two 208-row tiles, the whole V into a single one (it is dead as
// soon as dP has run), S = Q K^T and dP = dO V^T with N = Sk rounded up into 2 x 208 TMEM columns, P = exp2(S * scale -
// LSE) and dS = P (dP - delta) * scale in one pass out of TMEM, dS [128 x 224] over the consumed Q | dO (its last 32
// columns in a sub-tile of their own), dQ = dS K into the remaining 96 TMEM columns, un-rotate, stage over the Q tile,
// TMA-store.  One step per item: no fold of dQ_j through registers, half the S | dP launches, no padded key tile.
template <int DH>
struct BwdDqWkL {
  using G = Geo<DH>;
  using W = FwdWkL<DH>;
  static constexpr bool DS_ALIAS = 3 * SUB_A <= 2 * G::OPER;   // dS columns [0, 192) over the consumed Q | dO (Dh = 96)
  static constexpr uint32_t QDO = 0;                           // [2][Q | dO]
  static constexpr uint32_t K = 4 * G::OPER;                   // [2] wide K tiles
  static constexpr uint32_t V = K + 2 * W::OPERK;              // one wide V tile
  static constexpr uint32_t DST = V + W::OPERK;                // dS columns [192, 224): SWIZZLE_64B sub-tile
  static constexpr uint32_t DSM = DST + SUB_B;                 // dS columns [0, 192) when they have their own buffer
  static constexpr uint32_t AUX = DSM + (DS_ALIAS ? 0u : 3 * SUB_A);   // [2][256] fp32 additive key bias
  static constexpr uint32_t BARS = AUX + 2048;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;
  static_assert(TOTAL <= 227 * 1024, "shared memory");
};

template <int DH, int RC>
__global__ void __launch_bounds__(F_NT, 1) attn2_bwd_dq_wk_kernel(AttnArgs a, int n_items, const __grid_constant__ BwdMsMaps maps,
                                                                  const MsDims dm) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  using L = BwdDqWkL<DH>;
  using W = FwdWkL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  constexpr uint32_t T_S = 0, T_DP = W::KW, T_DQ = 2 * W::KW;   // TMEM columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed_qdo = bars, *landed_k = bars + 2, *landed_v = bars + 4, *sdp_full = bars + 5, *dq_full = bars + 6,
           *ds_ready = bars + 7, *staged = bars + 8, *aux_ready = bars + 10;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // K / V rows in [Sk, n_s) are read by the MMAs and never written by a box: everything starts as zero
  for (uint32_t i = tid * 16; i < L::AUX; i += F_NT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0u, 0u, 0u, 0u);
  tc::fence_async_smem();
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&landed_qdo[i], 1);
      tc::mbar_init(&landed_k[i], 1);
      tc::mbar_init(&staged[i], NC);
      tc::mbar_init(&aux_ready[i], 32);
    }
    tc::mbar_init(landed_v, 1);
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(dq_full, 1);
    tc::mbar_init(ds_ready, NC);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = a.qcos ? a.rot_dim / 8 : 0;                    // chunks to un-rotate in dQ
  const uint32_t s0 = tc::smem_u32(smem);
  const int n_s = min((a.Sk + 31) & ~31, W::KW);                // S | dP columns, dQ K extent (whole 32-column blocks)
  auto coords = [&](int item, int& n, int& h, int& qt) {
    qt = item % dm.nq;
    const int nh = item / dm.nq;
    n = nh / a.H;
    h = nh - n * a.H;
  };
  // dS: key columns [0, 192) in three SWIZZLE_128B sub-tiles at `dsm`, [192, 224) in a SWIZZLE_64B sub-tile at DST
  auto ds_main = [&](int b2) -> uint32_t { return L::DS_ALIAS ? L::QDO + b2 * 2 * G::OPER : L::DSM; };

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 10; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    const uint32_t kv_bytes = (uint32_t)a.Sk * DH * 2u;
    Prof pf;
    pf.init(blockIdx.x == 0 && lane == 0, 16);
    for (int j = 0; j < n_my; ++j) {
      pf.tick(0);
      const int item = blockIdx.x + j * gridDim.x, b2 = j & 1;
      int n, h, qt;
      coords(item, n, h, qt);
      const int n1 = n / a.n2_count, n2 = n - n1 * a.n2_count;
      const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER, sk = s0 + L::K + b2 * W::OPERK, sv = s0 + L::V;
      if (j >= 2) {                   // pair / K tile b2 belong to item j - 2: its dQ, staged over the Q tile, goes out first
        tc::mbar_wait(&staged[b2], ((j - 2) >> 1) & 1);
        pf.tick(1);
        if (lane == 0) {
          int pn, ph, pq;
          coords(item - 2 * (int)gridDim.x, pn, ph, pq);
          const int p1 = pn / a.n2_count, p2 = pn - p1 * a.n2_count;
          tma::store_box(&maps.m[8], sq, 0, ph, pq * 128, p2, p1);
          if (G::TAIL) tma::store_box(&maps.m[9], sq + SUB_A, 64, ph, pq * 128, p2, p1);
          tma::store_commit();
          tma::store_wait_read<0>();
        }
      }
      if (lane == 0) {
        tma::mbar_expect_tx(&landed_qdo[b2], 2 * G::OPER);
        tma::load_box(sq, &maps.m[0], &landed_qdo[b2], 0, h, qt * 128, n2, n1);
        if (G::TAIL) tma::load_box(sq + SUB_A, &maps.m[1], &landed_qdo[b2], 64, h, qt * 128, n2, n1);
        tma::load_box(sq + G::OPER, &maps.m[6], &landed_qdo[b2], 0, h, qt * 128, n2, n1);
        if (G::TAIL) tma::load_box(sq + G::OPER + SUB_A, &maps.m[7], &landed_qdo[b2], 64, h, qt * 128, n2, n1);
        tma::mbar_expect_tx(&landed_k[b2], kv_bytes);
        tma::load_box(sk, &maps.m[2], &landed_k[b2], 0, h, 0, n2, n1);
        if (G::TAIL) tma::load_box(sk + W::KA, &maps.m[3], &landed_k[b2], 64, h, 0, n2, n1);
      }
      pf.tick(2);
      // V: free once dP of item j - 1 has run (every phase of sdp_full is waited for, in order)
      if (j >= 1) tc::mbar_wait(sdp_full, (j - 1) & 1);
      pf.tick(3);
      if (lane == 0) {
        tma::mbar_expect_tx(landed_v, kv_bytes);
        tma::load_box(sv, &maps.m[4], landed_v, 0, h, 0, n2, n1);
        if (G::TAIL) tma::load_box(sv + W::KA, &maps.m[5], landed_v, 64, h, 0, n2, n1);
      }
      float* bias = reinterpret_cast<float*>(smem + L::AUX + b2 * 1024);   // item j - 2 has read it (staged above)
      const float* mrow = a.key_mask ? a.key_mask + (int64_t)n * a.Sk : nullptr;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int k = u * 32 + lane;
        bias[k] = key_bias2(mrow && k < a.Sk ? __ldg(mrow + k) : 1.f, k, a.Sk);
      }
      tc::mbar_arrive(&aux_ready[b2]);
      pf.tick(4);
    }
    for (int j = max(0, n_my - 2); j < n_my; ++j) {
      const int b2 = j & 1;
      tc::mbar_wait(&staged[b2], (j >> 1) & 1);
      if (lane == 0) {
        int pn, ph, pq;
        coords(blockIdx.x + j * gridDim.x, pn, ph, pq);
        const int p1 = pn / a.n2_count, p2 = pn - p1 * a.n2_count;
        const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER;
        tma::store_box(&maps.m[8], sq, 0, ph, pq * 128, p2, p1);
        if (G::TAIL) tma::store_box(&maps.m[9], sq + SUB_A, 64, ph, pq * 128, p2, p1);
        tma::store_commit();
      }
    }
    if (lane == 0) tma::store_wait_all();
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      const uint32_t idesc_s = tc::make_idesc(128, n_s, 0, 0);
      auto wide_k = [&](uint32_t base, int ks) -> uint64_t {      // K-major view of a wide K / V tile, K step ks of the head dimension
        return ks < 4 ? tc::make_desc_sw(base + ks * 32, 16, 1024, 2) : tc::make_desc_sw(base + W::KA + (ks - 4) * 32, 16, 512, 4);
      };
      auto issue_sdp = [&](int j) {
        const int b2 = j & 1;
        tc::mbar_wait(&landed_qdo[b2], (j >> 1) & 1);
        tc::mbar_wait(&landed_k[b2], (j >> 1) & 1);
        tc::mbar_wait(landed_v, j & 1);
        tc::fence_after_sync();
        const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER, sdo = sq + G::OPER, sk = s0 + L::K + b2 * W::OPERK, sv = s0 + L::V;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_oper_k(sq, ks), wide_k(sk, ks), idesc_s, ks > 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_oper_k(sdo, ks), wide_k(sv, ks), idesc_s, ks > 0);
        tc::mma_commit(sdp_full);
      };
      Prof pf;
      pf.init(blockIdx.x == 0, 32);
      if (n_my > 0) issue_sdp(0);
      for (int j = 0; j < n_my; ++j) {
        const int b2 = j & 1;
        pf.tick(0);
        tc::mbar_wait(ds_ready, j & 1);                           // dS written; S | dP (and the previous dQ) read
        pf.tick(1);
        tc::fence_after_sync();
        const uint32_t dsm = s0 + ds_main(b2), dst = s0 + L::DST, sk = s0 + L::K + b2 * W::OPERK;
        for (int ks = 0; ks * 16 < n_s; ++ks) {
          const uint64_t ad = ks < 12 ? tc::make_desc_sw(dsm + (ks >> 2) * SUB_A + (ks & 3) * 32, 16, 1024, 2)
                                      : tc::make_desc_sw(dst + (ks - 12) * 32, 16, 512, 4);
          tc::mma_ss(tmem + T_DQ, ad, tc::make_desc_sw(sk + ks * 2048, SUB_A, 1024, 2), tc::make_idesc(128, 64, 0, 1), ks > 0);
          if (G::TAIL)
            tc::mma_ss(tmem + T_DQ + 64, ad, tc::make_desc_sw(sk + W::KA + ks * 1024, SUB_B, 512, 4), tc::make_idesc(128, 32, 0, 1), ks > 0);
        }
        tc::mma_commit(dq_full);
        pf.tick(2);
        if (j + 1 < n_my) issue_sdp(j + 1);                       // behind dQ: runs while the compute warps drain and stage it
        pf.tick(3);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    const int nblk = (a.Sk + 31) >> 5;                            // <= 7; block 6 holds 16 columns
    constexpr int NCH = G::CPR / 2, RCT = (RC + 1) / 2;           // chunks of the dQ row per thread / of them at most rotated
    auto row_stats = [&](int j, float& lse2, float& dlt) {
      int n, h, qt;
      coords(blockIdx.x + j * gridDim.x, n, h, qt);
      const int row = min(qt * 128 + t, a.Sq - 1);
      const int64_t si = ((int64_t)n * a.H + h) * a.Sq + row;
      lse2 = __ldg(a.lse + si) * LOG2E;
      dlt = __ldg(a.delta + si);
    };
    float lse2 = 0.f, dlt = 0.f;
    if (n_my > 0) row_stats(0, lse2, dlt);
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, b2 = j & 1;
      int n, h, qt;
      coords(item, n, h, qt);
      pf.tick(0);
      tc::mbar_wait(&aux_ready[b2], (j >> 1) & 1);
      tc::mbar_wait(sdp_full, j & 1);
      pf.tick(1);
      tc::fence_after_sync();
      const float* bias = reinterpret_cast<const float*>(smem + L::AUX + b2 * 1024);
      uint8_t* dsm = smem + ds_main(b2);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32], dp[32];
          if (blk < 6) {
            tc::tmem_ld32(lane_base + T_S + blk * 32, sv);
            tc::tmem_ld32(lane_base + T_DP + blk * 32, dp);
          } else {
            tc::tmem_ld16(lane_base + T_S + 192, sv);
            tc::tmem_ld16(lane_base + T_DP + 192, dp);
#pragma unroll
            for (int e = 16; e < 32; ++e) sv[e] = dp[e] = 0.f;
          }
          tc::tmem_ld_wait();
          const float4* b4 = reinterpret_cast<const float4*>(bias + blk * 32);
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (blk == 6 && q >= 2) break;                          // columns [208, 224) do not exist
            const float4 ba = b4[2 * q], bc = b4[2 * q + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              const float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u] - lse2)), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1] - lse2));
              ph[u] = __floats2bfloat162_rn(p0 * (dp[e] - dlt) * a.scale, p1 * (dp[e + 1] - dlt) * a.scale);
            }
            const int c = blk * 4 + q;
            uint8_t* dstp = c < 24 ? dsm + (c >> 3) * SUB_A + tc::sw128_off(t, c & 7) : smem + L::DST + tc::sw64_off(t, c - 24);
            *reinterpret_cast<uint4*>(dstp) = pk;
          }
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(ds_ready);                                    // the MMA warp starts dQ = dS K
      pf.tick(2);
      float lse2_n = 0.f, dlt_n = 0.f;
      if (j + 1 < n_my) row_stats(j + 1, lse2_n, dlt_n);           // in flight under dQ
      // The two threads of a row take the 8-column chunks of dQ alternately (c = 2 i + half), so both share the rotated
      // ones; the un-rotation tables of this thread's chunks are fetched NOW, under the dQ MMAs (fetched after them,
      // by the thread that held all the rotated columns, they were a third of the item's time)
      float cv[RCT][8], sn[RCT][8];
      {
        const int pos = a.pos_is_head ? h : min(qt * 128 + t, a.Sq - 1);
        const float* cb = a.qcos + (int64_t)pos * a.rot_dim;
        const float* sb = a.qsin + (int64_t)pos * a.rot_dim;
#pragma unroll
        for (int i = 0; i < RCT; ++i) {
          const int c = 2 * i + half;
          if (c < rc) {
            const float4 c0 = __ldg(reinterpret_cast<const float4*>(cb + c * 8)), c1 = __ldg(reinterpret_cast<const float4*>(cb + c * 8 + 4));
            const float4 z0 = __ldg(reinterpret_cast<const float4*>(sb + c * 8)), z1 = __ldg(reinterpret_cast<const float4*>(sb + c * 8 + 4));
            cv[i][0] = c0.x; cv[i][1] = c0.y; cv[i][2] = c0.z; cv[i][3] = c0.w; cv[i][4] = c1.x; cv[i][5] = c1.y; cv[i][6] = c1.z; cv[i][7] = c1.w;
            sn[i][0] = z0.x; sn[i][1] = z0.y; sn[i][2] = z0.z; sn[i][3] = z0.w; sn[i][4] = z1.x; sn[i][5] = z1.y; sn[i][6] = z1.z; sn[i][7] = z1.w;
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) { cv[i][e] = 1.f; sn[i][e] = 0.f; }
          }
        }
      }
      pf.tick(3);
      tc::mbar_wait(dq_full, j & 1);
      pf.tick(4);
      tc::fence_after_sync();
      float g[NCH][8];
#pragma unroll
      for (int i = 0; i < NCH; ++i) tc::tmem_ld8(lane_base + T_DQ + (2 * i + half) * 8, g[i]);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < RCT; ++i) {                               // transpose rotation (identity tables past rc)
#pragma unroll
        for (int e = 0; e < 8; e += 2) {
          const float y0 = g[i][e] * cv[i][e] + g[i][e + 1] * sn[i][e + 1], y1 = g[i][e + 1] * cv[i][e + 1] - g[i][e] * sn[i][e];
          g[i][e] = y0;
          g[i][e + 1] = y1;
        }
      }
      pf.tick(5);
      uint8_t* tile = smem + L::QDO + b2 * 2 * G::OPER;             // over the consumed Q (and the dS that followed it)
#pragma unroll
      for (int i = 0; i < NCH; ++i) {
        uint4 pk;
        __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
        for (int e = 0; e < 4; ++e) ph[e] = __floats2bfloat162_rn(g[i][2 * e], g[i][2 * e + 1]);
        *reinterpret_cast<uint4*>(tile + oper_off(t, 2 * i + half)) = pk;
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&staged[b2]);
      pf.tick(6);
      lse2 = lse2_n;
      dlt = dlt_n;
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ============================================================================== backward, several tiles: dK | dV pass
// An item is a (sequence, head, 128-key tile) whose K | V stay in shared memory; a step is one 128-query tile (causal:
// the tiles on and below the diagonal): Q | dO boxes into one of two pairs, S = Q K^T and dP = dO V^T into TMEM, P and
// dS recomputed in one pass as in the dQ kernel and written as bf16 tiles, dV += P^T dO and dK += dS^T Q accumulated
// in TMEM across the item's steps (both operands MN-major: nothing is transposed).  The MMA thread issues S | dP of
// step s + 1 BEFORE dV | dK of step s, so the next step's arithmetic (kept in registers until the P / dS tiles are
// free) runs under them.  K | V of the next item are fetched as soon as the last S | dP of this one have run.  The
// item's last step drains the accumulators, un-rotates dK (transpose rotation), stages over P / dS and TMA-stores.
template <int DH>
struct BwdDkvL {
  using G = Geo<DH>;
  static constexpr uint32_t KV = 0;                            // K | V of the item
  static constexpr uint32_t QDO = 2 * G::OPER;                 // [2][Q | dO]
  static constexpr uint32_t P = 6 * G::OPER, DS = P + G::PBUF; // one tile each (dV / dK staging at the item's end)
  static constexpr uint32_t AUX = DS + G::PBUF;                // [2][128] fp32 additive key bias (exp2 domain), by item parity
  static constexpr uint32_t BARS = AUX + 2 * 512;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;
  static_assert(G::OPER <= G::PBUF, "gradient staging must fit in the P / dS tiles");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB of shared memory a CTA can opt in to");
};

// first query tile of key tile kt's item, and its number of steps
__device__ __forceinline__ int dkv_q0(const AttnArgs& a, int kt) { return a.causal ? kt : 0; }

template <int DH, int RC, bool ROT_IN>
__global__ void __launch_bounds__(F_NT, 1) attn2_bwd_dkv_ms_kernel(AttnArgs a, int n_items, const __grid_constant__ BwdMsMaps maps,
                                                                   const MsDims dm) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  using L = BwdDkvL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 352;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed_kv = bars, *landed_qdo = bars + 1, *aux_ready = bars + 3, *rot_done = bars + 5, *sdp_full = bars + 7,
           *pds_ready = bars + 8, *mma_done = bars + 9, *drained = bars + 11, *staged = bars + 12, *stg_free = bars + 13;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    tc::mbar_init(landed_kv, 1);
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(pds_ready, NC);
    tc::mbar_init(drained, NC);
    tc::mbar_init(staged, NC);
    tc::mbar_init(stg_free, 1);
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&landed_qdo[i], 1);
      tc::mbar_init(&aux_ready[i], 32);          // by item parity: the producer can never be two phases ahead of a reader
      tc::mbar_init(&rot_done[i], NC);
      tc::mbar_init(&mma_done[i], 1);
    }
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const int n_my = (n_items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  const int rc = a.qcos ? a.rot_dim / 8 : 0;                    // chunks to un-rotate in dK
  const int rc_in = ROT_IN ? rc : 0;
  const uint32_t s0 = tc::smem_u32(smem);
  // item j of this CTA -> (sequence, head, key tile); steps = query tiles q0 .. nq - 1
  auto item_of = [&](int j, int& n, int& h, int& kt, int& q0, int& steps) {
    ms_item(blockIdx.x + j * gridDim.x, j, a.H, dm.nk, n, h, kt);
    q0 = dkv_q0(a, kt);
    steps = dm.nq - q0;
  };

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 14; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    auto load_tile = [&](uint32_t dst, int view, int r0, uint64_t* bar, int h, int n2, int n1) {
      tma::load_box(dst, &maps.m[2 * view], bar, 0, h, r0, n2, n1);
      if (G::TAIL) tma::load_box(dst + SUB_A, &maps.m[2 * view + 1], bar, 64, h, r0, n2, n1);
    };
    auto store_tile = [&](int view, uint32_t src, int r0, int h, int n2, int n1) {
      tma::store_box(&maps.m[2 * view], src, 0, h, r0, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[2 * view + 1], src + SUB_A, 64, h, r0, n2, n1);
    };
    int s = 0, pn = 0, ph = 0, pkt = 0;                           // previous item's coordinates (its staged gradients)
    for (int j = 0; j < n_my; ++j) {
      int n, h, kt, q0, steps;
      item_of(j, n, h, kt, q0, steps);
      const int n1 = n / a.n2_count, n2 = n - n1 * a.n2_count;
      for (int u = 0; u < steps; ++u, ++s) {
        const int b2 = s & 1;
        // every phase of sdp_full is waited for, in order (a wait that skipped phases could be satisfied by an older
        // phase of the same parity); it completes long before mma_done(s - 2) below, so nothing is delayed by it
        if (s >= 1) tc::mbar_wait(sdp_full, (s - 1) & 1);
        if (u == 0) {                                             // K | V of the previous item are dead
          if (lane == 0) {
            tma::mbar_expect_tx(landed_kv, 2 * G::OPER);
            load_tile(s0 + L::KV, 1, kt * 128, landed_kv, h, n2, n1);
            load_tile(s0 + L::KV + G::OPER, 2, kt * 128, landed_kv, h, n2, n1);
          }
          float4 mv = make_float4(1.f, 1.f, 1.f, 1.f);
          const int k0 = kt * 128 + lane * 4;
          if (a.key_mask) {
            const float* mrow = a.key_mask + (int64_t)n * a.Sk;
            if (k0 + 3 < a.Sk && (a.Sk & 3) == 0) mv = __ldg(reinterpret_cast<const float4*>(mrow + k0));
            else {
              if (k0 < a.Sk) mv.x = __ldg(mrow + k0);
              if (k0 + 1 < a.Sk) mv.y = __ldg(mrow + k0 + 1);
              if (k0 + 2 < a.Sk) mv.z = __ldg(mrow + k0 + 2);
              if (k0 + 3 < a.Sk) mv.w = __ldg(mrow + k0 + 3);
            }
          }
          reinterpret_cast<float4*>(smem + L::AUX + (j & 1) * 512)[lane] =
              make_float4(key_bias2(mv.x, k0, a.Sk), key_bias2(mv.y, k0 + 1, a.Sk), key_bias2(mv.z, k0 + 2, a.Sk),
                          key_bias2(mv.w, k0 + 3, a.Sk));
          tc::mbar_arrive(&aux_ready[j & 1]);
        }
        if (s >= 2) tc::mbar_wait(&mma_done[b2], ((s - 2) >> 1) & 1);   // dV | dK of step s - 2 have read this Q | dO pair
        if (lane == 0) {
          tma::mbar_expect_tx(&landed_qdo[b2], 2 * G::OPER);
          load_tile(s0 + L::QDO + b2 * 2 * G::OPER, 0, (q0 + u) * 128, &landed_qdo[b2], h, n2, n1);
          load_tile(s0 + L::QDO + b2 * 2 * G::OPER + G::OPER, 3, (q0 + u) * 128, &landed_qdo[b2], h, n2, n1);
        }
        if (u == 0 && j > 0) {                                    // the previous item's staged dV (over P) and dK (over dS)
          tc::mbar_wait(staged, (j - 1) & 1);
          if (lane == 0) {
            const int p1 = pn / a.n2_count, p2 = pn - p1 * a.n2_count;
            store_tile(6, s0 + L::P, pkt * 128, ph, p2, p1);
            store_tile(5, s0 + L::DS, pkt * 128, ph, p2, p1);
            tma::store_commit();
            tma::store_wait_read<0>();
            tc::mbar_arrive(stg_free);
          }
        }
      }
      pn = n; ph = h; pkt = kt;
    }
    if (n_my > 0) {
      tc::mbar_wait(staged, (n_my - 1) & 1);
      if (lane == 0) {
        const int p1 = pn / a.n2_count, p2 = pn - p1 * a.n2_count;
        store_tile(6, s0 + L::P, pkt * 128, ph, p2, p1);
        store_tile(5, s0 + L::DS, pkt * 128, ph, p2, p1);
        tma::store_commit();
        tma::store_wait_all();
      }
    }
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      const uint32_t sk = s0 + L::KV, sv = sk + G::OPER, sp = s0 + L::P, sds = s0 + L::DS;
      auto issue_sdp = [&](int s) {
        const int b2 = s & 1;
        tc::mbar_wait(&rot_done[b2], (s >> 1) & 1);
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER, sdo = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_oper_k(sdo, ks), desc_oper_k(sv, ks), idesc, ks > 0);
        tc::mma_commit(sdp_full);
      };
      int total = 0;
      for (int j = 0; j < n_my; ++j) {
        int n, h, kt, q0, steps;
        item_of(j, n, h, kt, q0, steps);
        total += steps;
      }
      if (total > 0) issue_sdp(0);
      int s = 0;
      for (int j = 0; j < n_my; ++j) {
        int n, h, kt, q0, steps;
        item_of(j, n, h, kt, q0, steps);
        for (int u = 0; u < steps; ++u, ++s) {
          const int b2 = s & 1;
          tc::mbar_wait(pds_ready, s & 1);                        // P, dS written; S | dP read
          tc::fence_after_sync();
          // the next step's S | dP first (its arithmetic then runs under dV | dK) -- except at the item's last step,
          // where the compute warps wait for the accumulators: there dV | dK go first and the next item's S | dP run
          // while the gradients are drained, un-rotated and staged
          const bool last_u = u == steps - 1;
          if (s + 1 < total && !last_u) issue_sdp(s + 1);
          if (u == 0 && j > 0) {
            tc::mbar_wait(drained, (j - 1) & 1);                  // the previous item's accumulators have been read
            tc::fence_after_sync();
          }
          const uint32_t sq = s0 + L::QDO + b2 * 2 * G::OPER, sdo = sq + G::OPER;
#pragma unroll
          for (int ks = 0; ks < 8; ++ks) {                        // two independent accumulation chains, interleaved
            mma_oper_mn<DH>(tmem + T_DV, desc_p_mn(sp, ks), sdo, ks, 1, u > 0 || ks > 0);    // dV += P^T dO
            mma_oper_mn<DH>(tmem + T_DK, desc_p_mn(sds, ks), sq, ks, 1, u > 0 || ks > 0);    // dK += dS^T Q
          }
          tc::mma_commit(&mma_done[b2]);
          if (s + 1 < total && last_u) issue_sdp(s + 1);
        }
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    RotTab<RC> tab;                                               // rotation on the way in: half 0 q tables, half 1 k tables
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_pos = -1;
    const bool vec_tab = ((reinterpret_cast<uintptr_t>(a.qcos) | reinterpret_cast<uintptr_t>(a.qsin) | reinterpret_cast<uintptr_t>(a.kcos) |
                           reinterpret_cast<uintptr_t>(a.ksin)) & 15) == 0 && (a.rot_dim & 3) == 0;
    struct Walk {
      int j, u, steps, n, h, kt, q0;
    };
    auto start = [&](Walk& w, int j) {
      w.j = j;
      w.u = 0;
      if (j < n_my) item_of(j, w.n, w.h, w.kt, w.q0, w.steps);
    };
    auto advance = [&](Walk& w) {
      if (++w.u == w.steps) start(w, w.j + 1);
    };
    float lse_n = 0.f, dlt_n = 0.f;                               // row statistics of the walker's query tile
    auto prepare = [&](const Walk& w, int s) {
      const int b2 = s & 1;
      {
        const int row = min((w.q0 + w.u) * 128 + t, a.Sq - 1);
        const int64_t si = ((int64_t)w.n * a.H + w.h) * a.Sq + row;
        lse_n = __ldg(a.lse + si);
        dlt_n = __ldg(a.delta + si);
      }
      if (ROT_IN && rc_in) {
        const int pos = a.pos_is_head ? w.h : (half ? min(w.kt * 128 + t, a.Sk - 1) : min((w.q0 + w.u) * 128 + t, a.Sq - 1));
        if (pos != cur_pos) {
          if (vec_tab) load_tab4<RC>(tab, tcs, tsn, pos, a.rot_dim);
          else load_tab<RC>(tab, tcs, tsn, pos, a.rot_dim);
          cur_pos = pos;
        }
      }
      tc::mbar_wait(&landed_qdo[b2], (s >> 1) & 1);
      if (w.u == 0) tc::mbar_wait(landed_kv, w.j & 1);
      if (ROT_IN && rc_in) {
        if (half == 0) rot_row<RC, false>(smem + L::QDO + b2 * 2 * G::OPER, t, rc_in, tab);
        else if (w.u == 0) rot_row<RC, false>(smem + L::KV, t, rc_in, tab);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&rot_done[b2]);
    };
    Walk wp, wc;
    start(wp, 0);
    start(wc, 0);
    if (wp.j < n_my) {
      prepare(wp, 0);
      advance(wp);
    }
    Prof pf;
    pf.init(blockIdx.x == 0 && tid == 96, 0);
    for (int s = 0; wc.j < n_my; ++s) {
      const bool first = wc.u == 0, last = wc.u == wc.steps - 1;
      const int qt = wc.q0 + wc.u, k0 = wc.kt * 128;
      pf.tick(0);
      const float lse2 = lse_n * LOG2E, dlt = dlt_n;                // fetched by the walker one step ahead
      const bool dtile = a.causal && qt == wc.kt;
      int nblk = (min(a.Sk - k0, 128) + 31) >> 5;
      if (dtile) nblk = min(nblk, rg + 1);
      if (first) tc::mbar_wait(&aux_ready[wc.j & 1], (wc.j >> 1) & 1);
      tc::mbar_wait(sdp_full, s & 1);
      tc::fence_after_sync();
      pf.tick(1);
      const float4* b4 = reinterpret_cast<const float4*>(smem + L::AUX + (wc.j & 1) * 512);
      // arithmetic first, into registers: the P / dS tiles may still be read by dV | dK of the previous step
      uint4 pp[2][4], dd[2][4];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32], dp[32];
          tc::tmem_ld32(lane_base + T_S + blk * 32, sv);
          tc::tmem_ld32(lane_base + T_DP + blk * 32, dp);
          tc::tmem_ld_wait();
          const bool diag = dtile && blk == rg;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const float4 ba = b4[(blk * 32 + q * 8) >> 2], bc = b4[((blk * 32 + q * 8) >> 2) + 1];
            const float bb[8] = {ba.x, ba.y, ba.z, ba.w, bc.x, bc.y, bc.z, bc.w};
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pp[i][q]);
            __nv_bfloat162* dh = reinterpret_cast<__nv_bfloat162*>(&dd[i][q]);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const int e = q * 8 + 2 * u;
              float p0 = tc::ex2(fmaf(sv[e], scale2, bb[2 * u] - lse2)), p1 = tc::ex2(fmaf(sv[e + 1], scale2, bb[2 * u + 1] - lse2));
              if (diag) {
                if (e > lane) p0 = 0.f;
                if (e + 1 > lane) p1 = 0.f;
              }
              ph[u] = __floats2bfloat162_rn(p0, p1);
              dh[u] = __floats2bfloat162_rn(p0 * (dp[e] - dlt) * a.scale, p1 * (dp[e + 1] - dlt) * a.scale);
            }
          }
        } else {
#pragma unroll
          for (int q = 0; q < 4; ++q) pp[i][q] = dd[i][q] = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      tc::fence_before_sync();                                      // the TMEM reads above are complete (tcgen05.wait::ld)
      pf.tick(2);
      if (s >= 1) tc::mbar_wait(&mma_done[(s - 1) & 1], ((s - 1) >> 1) & 1);   // P / dS tiles are free
      if (first && wc.j > 0) tc::mbar_wait(stg_free, (wc.j - 1) & 1);          // ... and the staged gradients have left them
      pf.tick(3);
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          *reinterpret_cast<uint4*>(smem + L::P + pbuf_off(t, blk * 4 + q)) = pp[i][q];
          *reinterpret_cast<uint4*>(smem + L::DS + pbuf_off(t, blk * 4 + q)) = dd[i][q];
        }
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(pds_ready);
      pf.tick(4);
      if (wp.j < n_my) {
        prepare(wp, s + 1);
        advance(wp);
      }
      pf.tick(5);
      if (last) {
        // The two threads of a row take the 8-column chunks of dV and dK alternately (c = 2 i + half), so both share dK's
        // rotated chunks; their un-rotation tables are fetched first, under the item's last dV | dK MMAs (fetched after
        // them, by the thread that held every rotated column, they cost as much as a step)
        constexpr int NCH = G::CPR / 2, RCT = (RC + 1) / 2;
        float cv[RCT][8], sn[RCT][8];
        {
          const int pos = a.pos_is_head ? wc.h : min(k0 + t, a.Sk - 1);
          const float* cb = a.kcos + (int64_t)pos * a.rot_dim;
          const float* sb = a.ksin + (int64_t)pos * a.rot_dim;
#pragma unroll
          for (int i = 0; i < RCT; ++i) {
            const int c = 2 * i + half;
            if (c < rc) {
              const float4 c0 = __ldg(reinterpret_cast<const float4*>(cb + c * 8)), c1 = __ldg(reinterpret_cast<const float4*>(cb + c * 8 + 4));
              const float4 z0 = __ldg(reinterpret_cast<const float4*>(sb + c * 8)), z1 = __ldg(reinterpret_cast<const float4*>(sb + c * 8 + 4));
              cv[i][0] = c0.x; cv[i][1] = c0.y; cv[i][2] = c0.z; cv[i][3] = c0.w; cv[i][4] = c1.x; cv[i][5] = c1.y; cv[i][6] = c1.z; cv[i][7] = c1.w;
              sn[i][0] = z0.x; sn[i][1] = z0.y; sn[i][2] = z0.z; sn[i][3] = z0.w; sn[i][4] = z1.x; sn[i][5] = z1.y; sn[i][6] = z1.z; sn[i][7] = z1.w;
            } else {
#pragma unroll
              for (int e = 0; e < 8; ++e) { cv[i][e] = 1.f; sn[i][e] = 0.f; }
            }
          }
        }
        tc::mbar_wait(&mma_done[s & 1], (s >> 1) & 1);              // the item's accumulators are complete
        tc::fence_after_sync();
        pf.tick(6);
        float gv[NCH][8], gk[NCH][8];
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          tc::tmem_ld8(lane_base + T_DV + (2 * i + half) * 8, gv[i]);
          tc::tmem_ld8(lane_base + T_DK + (2 * i + half) * 8, gk[i]);
        }
        tc::tmem_ld_wait();
        tc::fence_before_sync();
        tc::mbar_arrive(drained);                                   // the next item's dV | dK may overwrite them
#pragma unroll
        for (int i = 0; i < RCT; ++i) {                             // transpose rotation (identity tables past rc)
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            const float y0 = gk[i][e] * cv[i][e] + gk[i][e + 1] * sn[i][e + 1], y1 = gk[i][e + 1] * cv[i][e + 1] - gk[i][e] * sn[i][e];
            gk[i][e] = y0;
            gk[i][e + 1] = y1;
          }
        }
#pragma unroll
        for (int i = 0; i < NCH; ++i) {
          uint4 pv, pk;
          __nv_bfloat162* hv = reinterpret_cast<__nv_bfloat162*>(&pv);
          __nv_bfloat162* hk = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            hv[e] = __floats2bfloat162_rn(gv[i][2 * e], gv[i][2 * e + 1]);
            hk[e] = __floats2bfloat162_rn(gk[i][2 * e], gk[i][2 * e + 1]);
          }
          *reinterpret_cast<uint4*>(smem + L::P + oper_off(t, 2 * i + half)) = pv;
          *reinterpret_cast<uint4*>(smem + L::DS + oper_off(t, 2 * i + half)) = pk;
        }
        tc::fence_async_smem();
        tc::mbar_arrive(staged);
        pf.tick(7);
      }
      advance(wc);
    }
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// delta[n, h, s] = sum_d dO[n, s, h, d] * O[n, s, h, d]  (args: o = dO, out_saved = O).  Four lanes per (token, head):
// a lane multiplies Dh / 32 consecutive 16-byte chunks of both rows (all loads issued before the first multiply), two
// shuffles add the four partial sums.  A warp covers eight consecutive (token, head) pairs per iteration -- with eight
// heads exactly one token row of dO and one of O, read as contiguous 16-byte chunks.  HBM-bound: reads dO and O once.
constexpr int DROWS = 8;      // (kept for the host-side grid computation: token rows per CTA iteration)
template <int CPL>            // chunks per lane: Dh / 32
__global__ void __launch_bounds__(256) attn_delta_kernel(AttnArgs a) {
  const int lane = threadIdx.x & 31, part = lane & 3;
  const int64_t units = (int64_t)a.N * a.Sq * a.H;                 // (token, head) pairs, heads fastest
  const int64_t wstride = (int64_t)gridDim.x * 8 * 8;
  for (int64_t u0 = ((int64_t)blockIdx.x * 8 + (threadIdx.x >> 5)) * 8; u0 < units; u0 += wstride) {
    const int64_t u = u0 + (lane >> 2);
    float acc = 0.f;
    int64_t out_idx = -1;
    if (u < units) {
      const int h = (int)(u % a.H);
      const int64_t row = u / a.H;
      const int s = (int)(row % a.Sq), n = (int)(row / a.Sq);
      const uint4* xp = reinterpret_cast<const uint4*>(view_ptr<__nv_bfloat16>(a.o, n, a.n2_count, s, h)) + part * CPL;
      const uint4* yp = reinterpret_cast<const uint4*>(view_ptr<__nv_bfloat16>(a.out_saved, n, a.n2_count, s, h)) + part * CPL;
      uint4 x[CPL], y[CPL];
#pragma unroll
      for (int c = 0; c < CPL; ++c) { x[c] = tc::ldg_stream16(xp + c); y[c] = tc::ldg_stream16(yp + c); }
#pragma unroll
      for (int c = 0; c < CPL; ++c) {
        const __nv_bfloat162* xh = reinterpret_cast<const __nv_bfloat162*>(&x[c]);
        const __nv_bfloat162* yh = reinterpret_cast<const __nv_bfloat162*>(&y[c]);
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float2 p = __bfloat1622float2(xh[e]), q = __bfloat1622float2(yh[e]);
          acc = fmaf(p.x, q.x, fmaf(p.y, q.y, acc));
        }
      }
      out_idx = ((int64_t)n * a.H + h) * a.Sq + s;
    }
    acc += __shfl_xor_sync(0xffffffffu, acc, 1);
    acc += __shfl_xor_sync(0xffffffffu, acc, 2);
    if (part == 0 && out_idx >= 0) a.delta[out_idx] = acc;
  }
}

// ============================================================================== backward
// Shared memory: the Q | K pair is double buffered (item j+1's pair lands while item j computes).  V and dO swap two
// slots: V(j)'s slot is free as soon as dP = dO V^T has run and takes dO(j+1); dO(j)'s slot is free as soon as
// dV = P^T dO has run (issued right after pass 1, under pass 2) and takes V(j+1).  P and dS are single.  The three
// gradient tiles are staged (swizzled operand layout, what the TMA store expects) over buffers that are dead by
// then: dQ over this item's K (dead once dQ = dS K has run), dK over its Q, dV over dS -- the Q | K pair is not
// needed again for two items, so only dV's store stands between this item and the next one's dS.
template <int DH>
struct BwdL {
  using G = Geo<DH>;
  static constexpr uint32_t QK = 0;                            // [2][Q | K]
  static constexpr uint32_t VD = 4 * G::OPER;                  // [2]: V(j) in slot j & 1, dO(j) in the other one
  static constexpr uint32_t P = 6 * G::OPER, DS = P + G::PBUF;
  static constexpr uint32_t AUX = DS + G::PBUF;                // [2][bias2[128] | lse*log2e[128]], then delta exchange [2][128]
  static constexpr uint32_t RED = AUX + 2048;
  static constexpr uint32_t BARS = RED + 1024;
  static constexpr uint32_t TOTAL = BARS + 256 + 1024;          // + alignment slack
  static_assert(G::OPER <= G::PBUF, "gradient staging must fit in the P / dS buffers");
  static_assert(TOTAL <= 232448, "exceeds the 227 KB of shared memory a CTA can opt in to");
};

struct BwdMaps { CUtensorMap m[14]; };   // q, k, v, dO (loads), dq, dk, dv (stores)  x  (SWIZZLE_128B part, SWIZZLE_64B part)

// Full accumulator row (DH fp32 columns) -> un-rotate the first rc chunks in fp32 registers -> bf16 -> swizzled tile row t
template <int DH, int RC>
__device__ __forceinline__ void stage_acc_unrot_sw(uint32_t taddr, uint8_t* tile, int t, int rc, const RotTab<RC>& T) {
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
      *reinterpret_cast<uint4*>(tile + oper_off(t, ch)) = pk;
    }
  }
}

// fp32 accumulator row `row` (TMEM lanes of this warp, columns [taddr, taddr + ncols)) added atomically to dst[0..ncols):
// the gradient of the shared extra key / value row, which sums over every tile that reads it.  Whole warp calls
// (tcgen05.ld is warp-collective); only the lane that owns the row adds.
__device__ __forceinline__ void x_row_atomic(uint32_t taddr, int ncols, bool mine, float* dst) {
  for (int c0 = 0; c0 < ncols; c0 += 16) {
    float g[16];
    tc::tmem_ld16(taddr + c0, g);
    tc::tmem_ld_wait();
    if (mine) {
#pragma unroll
      for (int i = 0; i < 16; ++i) atomicAdd(dst + c0 + i, g[i]);
    }
  }
}

template <int DH, int RC, bool GRP>
__global__ void __launch_bounds__(F_NT, 1) attn2_bwd_kernel(AttnArgs a, int n_items, const __grid_constant__ BwdMaps maps,
                                                            const Grp grp) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzle atoms are 1024-byte aligned
  using L = BwdL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 352, T_DQ = 0;   // dQ reuses the consumed S columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *landed_qk = bars, *aux_ready = bars + 2, *rot_done = bars + 4, *landed_v = bars + 6, *landed_do = bars + 7,
           *sdp_full = bars + 8, *pds_ready = bars + 9, *grads_full = bars + 10, *staged = bars + 11, *stg_free = bars + 12,
           *dq_full = bars + 13, *dq_staged = bars + 14, *p_ready = bars + 15, *dv_done = bars + 16;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (GRP) {   // rows a packed-group box does not cover start out as zero (see the forward kernel)
    for (uint32_t i = tid * 16; i < L::AUX; i += F_NT * 16) *reinterpret_cast<uint4*>(smem + i) = make_uint4(0u, 0u, 0u, 0u);
    tc::fence_async_smem();
  }
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      tc::mbar_init(&landed_qk[i], 1);
      tc::mbar_init(&aux_ready[i], 32);
      tc::mbar_init(&rot_done[i], NC);
    }
    tc::mbar_init(landed_v, 1);
    tc::mbar_init(landed_do, 1);
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(pds_ready, NC);
    tc::mbar_init(grads_full, 1);
    tc::mbar_init(staged, NC);
    tc::mbar_init(stg_free, 1);
    tc::mbar_init(dq_full, 1);
    tc::mbar_init(dq_staged, NC / 2);
    tc::mbar_init(p_ready, NC);
    tc::mbar_init(dv_done, 1);
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
  const uint32_t svd = s0 + L::VD, sp = s0 + L::P, sds = s0 + L::DS;
  const int Tq = GRP ? grp.xr : a.Sq, Tk = GRP ? grp.xr + grp.has_x : a.Sk;   // per-tile query rows / keys
  const uint32_t tile_bytes = GRP ? (uint32_t)grp.xr * DH * 2u : G::OPER;

  if (warp == F_PROD) {
    // ------------------------------------------------------------------ producer warp
    Prof pf;
    pf.init(blockIdx.x == 0 && lane == 0, 16);
    if (lane == 0) {
#pragma unroll
      for (int i = 0; i < 14; ++i)
        if (G::TAIL || !(i & 1)) tma::prefetch_map(&maps.m[i]);
    }
    auto coords = [&](int j, int& h, int& n2, int& n1) {
      const int item = blockIdx.x + j * gridDim.x;
      if (GRP) {
        grp_coords(item, a.H, grp, n1, n2, h);
      } else {
        const int n = item / a.H;
        h = item - n * a.H;
        n1 = n / a.n2_count;
        n2 = n - n1 * a.n2_count;
      }
    };
    auto load_tile = [&](int view, uint32_t dst, uint64_t* bar, int h, int n2, int n1) {
      tma::load_box(dst, &maps.m[2 * view], bar, 0, h, 0, n2, n1);
      if (G::TAIL) tma::load_box(dst + SUB_A, &maps.m[2 * view + 1], bar, 64, h, 0, n2, n1);
    };
    auto store_tile = [&](int view, uint32_t src, int h, int n2, int n1) {
      tma::store_box(&maps.m[2 * view], src, 0, h, 0, n2, n1);
      if (G::TAIL) tma::store_box(&maps.m[2 * view + 1], src + SUB_A, 64, h, 0, n2, n1);
    };
    auto store_grads = [&](int j) {                               // staged gradients of item j -> global (lane 0)
      int h, n2, n1;
      coords(j, h, n2, n1);
      const uint32_t sq = s0 + L::QK + (j & 1) * 2 * G::OPER;
      store_tile(6, sds, h, n2, n1);                              // dv (over dS): its own group, released first
      tma::store_commit();
      store_tile(4, sq + G::OPER, h, n2, n1);                     // dq (over this item's K)
      store_tile(5, sq, h, n2, n1);                               // dk (over this item's Q)
      tma::store_commit();
    };
    // shared extra key / value row of item j: lanes [0, CPR) hold K's 16-byte chunks, [CPR, 2 CPR) V's
    auto load_x = [&](int j) -> uint4 {
      uint4 v = make_uint4(0u, 0u, 0u, 0u);
      if (GRP && grp.has_x && lane < 2 * G::CPR) {
        int h, n2, n1;
        coords(j, h, n2, n1);
        const HeadView& xv = lane < G::CPR ? a.kx : a.vx;
        v = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(xv.p) + (int64_t)n1 * xv.sn1 + (int64_t)h * xv.sh +
                                            (lane % G::CPR) * 8);
      }
      return v;
    };
    uint4 xnext = n_my > 0 ? load_x(0) : make_uint4(0u, 0u, 0u, 0u);
    for (int j = 0; j < n_my; ++j) {
      const int b = j & 1;
      int h, n2, n1;
      coords(j, h, n2, n1);
      const int n = n1 * a.n2_count + n2;
      const uint4 xcur = xnext;
      pf.tick(0);
      if (lane == 0 && j >= 2) tma::store_wait_read<0>();   // Q | K pair b still holds item j-2's dQ / dK staging until read
      if (GRP && grp.has_x) {
        __syncwarp();
        if (lane < G::CPR)                                        // K's extra row, ahead of the boxes of this pair
          *reinterpret_cast<uint4*>(smem + L::QK + (b * 2 + 1) * G::OPER + oper_off(grp.xr, lane)) = xcur;
        tc::fence_async_smem();
        __syncwarp();
      }
      if (lane == 0) {
        const uint32_t sq = s0 + L::QK + b * 2 * G::OPER;
        tma::mbar_expect_tx(&landed_qk[b], 2 * tile_bytes);
        load_tile(0, sq, &landed_qk[b], h, n2, n1);
        load_tile(1, sq + G::OPER, &landed_qk[b], h, n2, n1);
      }
      if (j + 1 < n_my) xnext = load_x(j + 1);
      pf.tick(1);
      if (j > 0) tc::mbar_wait(sdp_full, (j - 1) & 1);            // V(j-1)'s slot is free once dP of that item has run
      pf.tick(2);
      if (lane == 0) {
        tma::mbar_expect_tx(landed_do, tile_bytes);
        load_tile(3, svd + (b ^ 1) * G::OPER, landed_do, h, n2, n1);
      }
      {
        const int k0 = lane * 4;
        float mv[4] = {1.f, 1.f, 1.f, 1.f}, ls[4] = {0.f, 0.f, 0.f, 0.f};
        const float* mrow = !GRP && a.key_mask ? a.key_mask + (int64_t)n * a.Sk : nullptr;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int r = k0 + u;
          if (mrow && r < a.Sk) mv[u] = __ldg(mrow + r);
          if (GRP) {                                              // tile row r = group g, row r - g gl of that group
            const int g = r / grp.gl;
            if (r < grp.xr && n2 + g < a.n2_count)
              ls[u] = __ldg(a.lse + ((int64_t)(n + g) * a.H + h) * grp.gl + (r - g * grp.gl));
          } else if (r < a.Sq) {
            ls[u] = __ldg(a.lse + ((int64_t)n * a.H + h) * a.Sq + r);
          }
        }
        float* aux = reinterpret_cast<float*>(smem + L::AUX + b * 1024);
        reinterpret_cast<float4*>(aux)[lane] = make_float4(key_bias2(mv[0], k0, Tk), key_bias2(mv[1], k0 + 1, Tk),
                                                           key_bias2(mv[2], k0 + 2, Tk), key_bias2(mv[3], k0 + 3, Tk));
        reinterpret_cast<float4*>(aux + 128)[lane] = make_float4(ls[0] * LOG2E, ls[1] * LOG2E, ls[2] * LOG2E, ls[3] * LOG2E);
        tc::mbar_arrive(&aux_ready[b]);
      }
      pf.tick(3);
      if (j > 0) tc::mbar_wait(dv_done, (j - 1) & 1);             // dO(j-1)'s slot is free once dV of that item has run
      pf.tick(4);
      if (GRP && grp.has_x) {
        if (lane >= G::CPR && lane < 2 * G::CPR)                  // V's extra row, ahead of V's boxes
          *reinterpret_cast<uint4*>(smem + L::VD + b * G::OPER + oper_off(grp.xr, lane - G::CPR)) = xcur;
        tc::fence_async_smem();
        __syncwarp();
      }
      if (lane == 0) {
        tma::mbar_expect_tx(landed_v, tile_bytes);
        load_tile(2, svd + b * G::OPER, landed_v, h, n2, n1);
      }
      if (j > 0) {
        tc::mbar_wait(staged, (j - 1) & 1);
        pf.tick(5);
        if (lane == 0) {
          store_grads(j - 1);
          tma::store_wait_read<1>();                              // the dV group has left dS
          tc::mbar_arrive(stg_free);
        }
        pf.tick(6);
      }
    }
    if (n_my > 0) {
      tc::mbar_wait(staged, (n_my - 1) & 1);
      if (lane == 0) {
        store_grads(n_my - 1);
        tma::store_wait_all();
      }
    }
  } else if (warp == F_MMA) {
    // ------------------------------------------------------------------ MMA-issue warp (one thread)
    if (lane == 0) {
      for (int j = 0; j < n_my; ++j) {
        const int b = j & 1;
        const uint32_t sq = s0 + L::QK + b * 2 * G::OPER, sk = sq + G::OPER;
        const uint32_t sv = svd + b * G::OPER, sdo = svd + (b ^ 1) * G::OPER;
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        tc::mbar_wait(&rot_done[b], (j >> 1) & 1);
        if (j > 0) tc::mbar_wait(dq_staged, (j - 1) & 1);         // dQ of the previous item has left the S columns
        tc::fence_after_sync();
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_oper_k(sq, ks), desc_oper_k(sk, ks), idesc, ks > 0);
        tc::mbar_wait(landed_v, j & 1);
        tc::mbar_wait(landed_do, j & 1);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_oper_k(sdo, ks), desc_oper_k(sv, ks), idesc, ks > 0);
        tc::mma_commit(sdp_full);
        const int kq = (Tq + 15) >> 4, kk = (Tk + 15) >> 4;
        tc::mbar_wait(p_ready, j & 1);                            // P is complete after pass 1: dV runs under pass 2
        tc::fence_after_sync();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kq) mma_oper_mn<DH>(tmem + T_DV, desc_p_mn(sp, ks), sdo, ks, 1, ks > 0);   // dV = P^T dO
        tc::mma_commit(dv_done);
        tc::mbar_wait(pds_ready, j & 1);
        tc::fence_after_sync();
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kk) mma_oper_mn<DH>(tmem + T_DQ, desc_p_k(sds, ks), sk, ks, 0, ks > 0);    // dQ = dS K
        tc::mma_commit(dq_full);                                  // dQ can be staged while dK, dV still run
#pragma unroll
        for (int ks = 0; ks < 8; ++ks)
          if (ks < kq) mma_oper_mn<DH>(tmem + T_DK, desc_p_mn(sds, ks), sq, ks, 1, ks > 0);   // dK = dS^T Q
        tc::mma_commit(grads_full);
      }
    }
  } else {
    // ------------------------------------------------------------------ compute group
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale = a.scale, scale2 = a.scale * LOG2E;
    float* red_delta = reinterpret_cast<float*>(smem + L::RED);
    const bool qrow = t < Tq;
    int nblk = (Tk + 31) >> 5;
    if (a.causal) nblk = min(nblk, rg + 1);
    const int glo = GRP ? min(t, grp.xr - 1) / grp.gl * grp.gl : 0, ghi = GRP ? glo + grp.gl : 0;
    const int gx = GRP && grp.has_x ? grp.xr : -1;
    const uint32_t bmask = GRP ? grp_block_mask(rg * 32, grp) : 0xffffffffu;
    RotTab<RC> tab;                                               // half 0: q tables (Q, dQ), half 1: k tables (K, dK)
    const float* tcs = half ? a.kcos : a.qcos;
    const float* tsn = half ? a.ksin : a.qsin;
    int cur_h = -1;
    if (rc && !a.pos_is_head) load_tab<RC>(tab, tcs, tsn, GRP ? min(t, grp.xr - 1) % grp.gl : min(t, (half ? a.Sk : a.Sq) - 1), a.rot_dim);
    Prof pc;
    pc.init(blockIdx.x == 0 && tid == 96, 0);
    // rotate item j's Q / K rows in place (buffer j & 1) and hand the pair to the MMA warp
    auto prepare = [&](int j) {
      const int b = j & 1;
      pc.tick(8);
      tc::mbar_wait(&landed_qk[b], (j >> 1) & 1);
      pc.tick(9);
      uint8_t* mine = smem + L::QK + (b * 2 + half) * G::OPER;
      if (GRP && t >= grp.xr) {           // rows the boxes do not cover hold an older item's dQ / dK staging
        if (half == 0 || t > grp.xr || !grp.has_x) zero_row<G::CPR>(mine, t);   // (the extra key row is the producer's)
      } else if (rc) {
        if (a.pos_is_head) {
          const int h = (blockIdx.x + j * gridDim.x) % a.H;
          if (h != cur_h) { load_tab<RC>(tab, tcs, tsn, h, a.rot_dim); cur_h = h; }
        }
        rot_row<RC, false>(mine, t, rc, tab);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(&rot_done[b]);
      pc.tick(10);
    };
    if (n_my > 0) prepare(0);
    for (int j = 0; j < n_my; ++j) {
      const int b = j & 1;
      const float* aux = reinterpret_cast<const float*>(smem + L::AUX + b * 1024);
      const float4* b4 = reinterpret_cast<const float4*>(aux);
      pc.tick(0);
      tc::mbar_wait(&aux_ready[b], (j >> 1) & 1);
      pc.tick(1);
      tc::mbar_wait(sdp_full, j & 1);
      tc::fence_after_sync();
      pc.tick(2);
      const float lse2 = aux[128 + t];
      uint32_t pk[32];                                              // this thread's 64 columns of P, bf16 pairs
      float delta = 0.f;                                            // rowsum(P * dP) == rowsum(dO * O)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk && ((bmask >> blk) & 1u)) {
          const int lim = !qrow ? -1 : ((a.causal && blk == rg) ? lane : 31);
          const bool masked = (a.causal && blk == rg) || rg * 32 + 31 >= Tq;   // warp-uniform: most blocks skip the test
#pragma unroll
          for (int hb = 0; hb < 2; ++hb) {                          // 16 columns at a time keeps the live set small
            float sv16[16], dp[16];
            tc::tmem_ld16(lane_base + T_S + blk * 32 + hb * 16, sv16);
            tc::tmem_ld16(lane_base + T_DP + blk * 32 + hb * 16, dp);
            tc::tmem_ld_wait();
#pragma unroll
            for (int e = 0; e < 16; e += 4) {
              const float4 bq = b4[(blk * 32 + hb * 16 + e) >> 2];
              const float bb[4] = {bq.x, bq.y, bq.z, bq.w};
              float p[4];
#pragma unroll
              for (int u = 0; u < 4; ++u) {
                p[u] = tc::ex2(fmaf(sv16[e + u], scale2, bb[u]) - lse2);
                if (GRP) {                                          // block-diagonal mask: own group + the extra key
                  const int c = blk * 32 + hb * 16 + e + u;
                  if (!qrow || !((c >= glo && c < ghi) || c == gx)) p[u] = 0.f;
                } else if (masked && hb * 16 + e + u > lim) p[u] = 0.f;
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
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(p_ready);                                     // the MMA warp starts dV = P^T dO
      pc.tick(3);
      tc::bar_sync(BAR_PAIR + rg, 64);                              // the two halves of this row group only
      pc.tick(4);
      delta += red_delta[(half ^ 1) * 128 + t];
      if (j > 0) tc::mbar_wait(stg_free, (j - 1) & 1);              // dS doubles as the previous item's dV staging
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
#pragma unroll
        for (int hb = 0; hb < 2; ++hb) {
          uint32_t dsk[8];
          if (blk < nblk && ((bmask >> blk) & 1u)) {
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
      tc::mbar_arrive(pds_ready);                                   // the MMA warp starts dV, dK, dQ
      pc.tick(5);
      // next item's rotation hides the gradient MMAs (head-indexed tables are still needed to un-rotate this item)
      if (j + 1 < n_my && !a.pos_is_head) prepare(j + 1);
      pc.tick(6);
      // half 0 owns the dQ row, half 1 the dK row (each un-rotates with its own table); dV is split by columns
      uint8_t* pair = smem + L::QK + b * 2 * G::OPER;
      // the shared extra key's gradients (tile row grp.xr) are summed over tiles in fp32 instead of being stored
      const bool xwarp = GRP && grp.has_x && rg == (grp.xr >> 5);   // warp-uniform
      int xn1 = 0, xh = 0;
      if (xwarp) {
        int n2;
        grp_coords(blockIdx.x + j * gridDim.x, a.H, grp, xn1, n2, xh);
      }
      if (half == 0) {
        tc::mbar_wait(dq_full, j & 1);
        tc::fence_after_sync();
        pc.tick(7);
        stage_acc_unrot_sw<DH, RC>(lane_base + T_DQ, pair + G::OPER, t, rc, tab);      // over K
        tc::fence_before_sync();
        tc::mbar_arrive(dq_staged);                                 // the S columns are free for the next item
        tc::mbar_wait(grads_full, j & 1);
        tc::fence_after_sync();
      } else {
        tc::mbar_wait(grads_full, j & 1);
        tc::fence_after_sync();
        pc.tick(7);
        stage_acc_unrot_sw<DH, RC>(lane_base + T_DK, pair, t, rc, tab);                 // over Q
        if (xwarp) x_row_atomic(lane_base + T_DK, DH, lane == (grp.xr & 31), a.dkx + ((int64_t)xn1 * a.H + xh) * DH);
      }
      stage_acc_sw<G::CPR / 2>(lane_base + T_DV + half * HALF, smem + L::DS, t, half * (G::CPR / 2), 1.0f);
      if (xwarp) x_row_atomic(lane_base + T_DV + half * HALF, HALF, lane == (grp.xr & 31),
                              a.dvx + ((int64_t)xn1 * a.H + xh) * DH + half * HALF);
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::mbar_arrive(staged);
      if (j + 1 < n_my && a.pos_is_head) prepare(j + 1);
      pc.tick(11);
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

// Packed-group geometry of a problem (t2::Grp; G = 0: not a packed-group problem).  Taken when the problem is plain
// self-attention inside short groups -- Sq == Sk (- the shared extra key), no causal / padding mask, token-position
// tables -- and either carries the extra key (any group count) or at least two groups of one n1 fit a tile.
static t2::Grp grp_of(const AttnArgs& a) {
  t2::Grp g = {0, 0, 0, 0, 0};
  const int has_x = a.kx.p != nullptr ? 1 : 0;
  const int gl = a.Sk - has_x;
  if (gl <= 0 || a.Sq != gl || gl + has_x > 128 || a.causal || a.key_mask || a.pos_is_head) return g;
  int G = (128 - has_x) / gl;
  if (G > a.n2_count) G = a.n2_count;
  if (G < 1 || (!has_x && G < 2)) return g;
  g.gl = gl;
  g.G = G;
  g.nb = (a.n2_count + G - 1) / G;
  g.xr = G * gl;
  g.has_x = has_x;
  return g;
}

// single-tile problems the persistent kernels take (the caller has already checked dtype, views, Dh)
int attn_tc2_supported(const AttnArgs& a, int backward) {
  if (a.Sq > 128 || a.Sk > 128) return 0;
  if (a.kx.p != nullptr && grp_of(a).G == 0) return 0;      // the extra key exists in the packed-group kernels only
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

// tensor maps for `count` head views (a 64-column SWIZZLE_128B map and, for Dh = 96, a 32-column SWIZZLE_64B map each);
// packed groups: boxes of gl rows x G groups instead of one 128-row tile
template <int DH>
static int build_maps(CUtensorMap* m, const AttnArgs& a, const HeadView* const* vs, const int* seq, int count, const t2::Grp& g) {
  const int rows = g.G ? g.gl : 128, ngrp = g.G ? g.G : 1;
  for (int i = 0; i < count; ++i) {
    int rc = tma::make_head_map(&m[2 * i], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 64, rows, ngrp);
    if (rc == 0 && DH > 64) rc = tma::make_head_map(&m[2 * i + 1], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 32, rows, ngrp);
    if (rc) return rc;
  }
  return 0;
}

template <typename Kern, typename Maps>
static int launch2(Kern kern, size_t smem, const AttnArgs& a, const Maps& maps, const t2::Grp& g, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  // work items: (sequence, head), or with packed groups (n1, tile of G groups, head)
  const int64_t items = g.G ? (int64_t)(a.N / a.n2_count) * g.nb * a.H : (int64_t)a.N * a.H;
  if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
  const int grid = (int)(items < num_sms() ? items : num_sms());
  kern<<<grid, t2::F_NT, smem, st>>>(a, (int)items, maps, g);
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

// forward: tensor maps for the q, k, v loads and the o store, then one persistent CTA per SM
template <int DH, int RC>
static int launch_fwd(const AttnArgs& a, cudaStream_t st) {
  t2::FwdMaps maps;
  memset(&maps, 0, sizeof(maps));
  const t2::Grp g = grp_of(a);
  const int skb = a.Sk - g.has_x;                   // rows k / v hold (the extra key lives in kx / vx)
  const HeadView* vs[4] = {&a.q, &a.k, &a.v, &a.o};
  const int seq[4] = {a.Sq, skb, skb, a.Sq};
  if (int rc = build_maps<DH>(maps.m, a, vs, seq, 4, g)) return rc;
  if (g.G) return launch2(t2::attn2_fwd_kernel<DH, RC, true>, t2::FwdL<DH>::TOTAL, a, maps, g, st);
  return launch2(t2::attn2_fwd_kernel<DH, RC, false>, t2::FwdL<DH>::TOTAL, a, maps, g, st);
}

template <int DH, int RC>
static int launch_bwd(const AttnArgs& a, cudaStream_t st) {
  t2::BwdMaps maps;
  memset(&maps, 0, sizeof(maps));
  const t2::Grp g = grp_of(a);
  const int skb = a.Sk - g.has_x;
  if (g.has_x && (!a.dkx || !a.dvx)) return MEANT_EARG;
  const HeadView* vs[7] = {&a.q, &a.k, &a.v, &a.o, &a.dq, &a.dk, &a.dv};
  const int seq[7] = {a.Sq, skb, skb, a.Sq, a.Sq, skb, skb};
  if (int rc = build_maps<DH>(maps.m, a, vs, seq, 7, g)) return rc;
  if (g.G) return launch2(t2::attn2_bwd_kernel<DH, RC, true>, t2::BwdL<DH>::TOTAL, a, maps, g, st);
  return launch2(t2::attn2_bwd_kernel<DH, RC, false>, t2::BwdL<DH>::TOTAL, a, maps, g, st);
}

// multi-tile problems the multi-step forward takes (the caller has already checked dtype, views, Dh)
int attn_tc2_ms_supported(const AttnArgs& a) {
  if (a.Sq <= 128 && a.Sk <= 128) return 0;
  if (a.kx.p != nullptr) return 0;
  if (a.causal && a.Sq != a.Sk) return 0;
  if (a.qcos && !a.q_rot.p) {
    const int rc_max = a.Dh == 96 ? 6 : 8;
    if (a.rot_dim / 8 > rc_max) return 0;
    const uintptr_t al = reinterpret_cast<uintptr_t>(a.qcos) | reinterpret_cast<uintptr_t>(a.qsin) |
                         reinterpret_cast<uintptr_t>(a.kcos) | reinterpret_cast<uintptr_t>(a.ksin);
    if (al % 8 != 0) return 0;
  }
  const bool pre = a.q_rot.p != nullptr;
  const HeadView* vs[4] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o};
  for (int i = 0; i < 4; ++i)
    if (vs[i]->ss <= 0 || vs[i]->ss >= (1 << 23)) return 0;
  return 1;
}
// ... and the multi-step backward (q_rot / k_rot, when given, are what forward saved)
int attn_tc2_ms_bwd_supported(const AttnArgs& a) {
  if (!attn_tc2_ms_supported(a) || !a.lse || !a.delta || !a.out_saved.p) return 0;
  if (a.qcos) {                                     // the un-rotation of dQ / dK always runs in the kernel (16-byte table loads)
    const int rc_max = a.Dh == 96 ? 6 : 8;
    if (a.rot_dim / 8 > rc_max || a.rot_dim % 8) return 0;
    const uintptr_t al = reinterpret_cast<uintptr_t>(a.qcos) | reinterpret_cast<uintptr_t>(a.qsin) |
                         reinterpret_cast<uintptr_t>(a.kcos) | reinterpret_cast<uintptr_t>(a.ksin);
    if (al % 16 != 0) return 0;
  }
  const HeadView* gs[3] = {&a.dq, &a.dk, &a.dv};
  for (int i = 0; i < 3; ++i)
    if (!gs[i]->p || gs[i]->ss <= 0 || gs[i]->ss >= (1 << 23)) return 0;
  return 1;
}

template <int DH, int RC>
static int launch_fwd_ms(const AttnArgs& a, cudaStream_t st) {
  t2::FwdMaps maps;
  memset(&maps, 0, sizeof(maps));
  const t2::Grp g = {0, 0, 0, 0, 0};
  const bool pre = a.q_rot.p != nullptr;
  const HeadView* vs[4] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o};
  const int seq[4] = {a.Sq, a.Sk, a.Sk, a.Sq};
  if (int rc = build_maps<DH>(maps.m, a, vs, seq, 4, g)) return rc;
  const t2::MsDims dm = {(a.Sq + 127) / 128, (a.Sk + 127) / 128};
  const int64_t items = (int64_t)a.N * a.H * dm.nq;
  if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
  auto kern = t2::attn2_fwd_ms_kernel<DH, RC>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t2::FwdMsL<DH>::TOTAL);
  if (e != cudaSuccess) return (int)e;
  const int grid = (int)(items < num_sms() ? items : num_sms());
  kern<<<grid, t2::F_NT, t2::FwdMsL<DH>::TOTAL, st>>>(a, (int)items, maps, dm);
  MEANT_LAUNCH_CHECK();
  return 0;
}

int attn_tc2_wk_supported(const AttnArgs& a);

// multi-step backward: delta, then the dQ pass (bit 1 of `parts`) and / or the dK | dV pass (bit 0)
template <int DH, int RC>
static int launch_bwd_ms(const AttnArgs& a, int parts, cudaStream_t st) {
  if (!a.lse || !a.delta || !a.out_saved.p) return MEANT_EARG;
  t2::BwdMsMaps maps;
  memset(&maps, 0, sizeof(maps));
  const t2::Grp g = {0, 0, 0, 0, 0};
  const bool pre = a.q_rot.p != nullptr;
  const HeadView* vs[7] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o, &a.dq, &a.dk, &a.dv};
  const int seq[7] = {a.Sq, a.Sk, a.Sk, a.Sq, a.Sq, a.Sk, a.Sk};
  if (int rc = build_maps<DH>(maps.m, a, vs, seq, 7, g)) return rc;
  const t2::MsDims dm = {(a.Sq + 127) / 128, (a.Sk + 127) / 128};
  {
    const int64_t units = (int64_t)a.N * a.Sq * a.H;
    int64_t blocks = (units + 63) / 64;                      // 8 warps x 8 (token, head) pairs per CTA iteration
    if (blocks > num_sms() * 8) blocks = num_sms() * 8;
    if (DH == 96) t2::attn_delta_kernel<3><<<(unsigned)blocks, 256, 0, st>>>(a);
    else t2::attn_delta_kernel<2><<<(unsigned)blocks, 256, 0, st>>>(a);
    MEANT_LAUNCH_CHECK();
  }
  static const bool no_wk = getenv("MEANT_ATTN_NO_WK") != nullptr;    // read once: A/B against the multi-step dQ pass
  if ((parts & 2) && !no_wk && attn_tc2_wk_supported(a)) {
    // every key in one tile: the wide-key dQ pass (K and V boxes hold all Sk rows)
    t2::BwdMsMaps wm;
    memset(&wm, 0, sizeof(wm));
    const int rows[5] = {128, a.Sk, a.Sk, 128, 128};
    int rcw = 0;
    for (int i = 0; i < 5 && rcw == 0; ++i) {
      rcw = tma::make_head_map(&wm.m[2 * i], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 64, rows[i], 1);
      if (rcw == 0 && DH > 64) rcw = tma::make_head_map(&wm.m[2 * i + 1], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 32, rows[i], 1);
    }
    if (rcw == 0) {
      const int64_t items = (int64_t)a.N * a.H * dm.nq;
      if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
      auto kern = t2::attn2_bwd_dq_wk_kernel<DH, RC>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t2::BwdDqWkL<DH>::TOTAL);
      if (e != cudaSuccess) return (int)e;
      const int grid = (int)(items < num_sms() ? items : num_sms());
      kern<<<grid, t2::F_NT, t2::BwdDqWkL<DH>::TOTAL, st>>>(a, (int)items, wm, dm);
      MEANT_LAUNCH_CHECK();
      parts &= ~2;
    }
  }
  if (parts & 2) {
    const int64_t items = (int64_t)a.N * a.H * dm.nq;
    if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
    auto kern = pre ? t2::attn2_bwd_dq_ms_kernel<DH, RC, false> : t2::attn2_bwd_dq_ms_kernel<DH, RC, true>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t2::BwdDqL<DH>::TOTAL);
    if (e != cudaSuccess) return (int)e;
    const int grid = (int)(items < num_sms() ? items : num_sms());
    kern<<<grid, t2::F_NT, t2::BwdDqL<DH>::TOTAL, st>>>(a, (int)items, maps, dm);
    MEANT_LAUNCH_CHECK();
  }
  if (parts & 1) {
    const int64_t items = (int64_t)a.N * a.H * dm.nk;
    if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
    auto kern = pre ? t2::attn2_bwd_dkv_ms_kernel<DH, RC, false> : t2::attn2_bwd_dkv_ms_kernel<DH, RC, true>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t2::BwdDkvL<DH>::TOTAL);
    if (e != cudaSuccess) return (int)e;
    const int grid = (int)(items < num_sms() ? items : num_sms());
    kern<<<grid, t2::F_NT, t2::BwdDkvL<DH>::TOTAL, st>>>(a, (int)items, maps, dm);
    MEANT_LAUNCH_CHECK();
  }
  return 0;
}

// one-wide-key-tile forward: 128 < Sk <= 208, not causal, rotation already applied (or none)
int attn_tc2_wk_supported(const AttnArgs& a) {
  if (a.Sk <= 128 || a.Sk > t2::FwdWkL<96>::KW || a.causal || a.kx.p != nullptr) return 0;
  if (a.qcos && !a.q_rot.p) return 0;
  const bool pre = a.q_rot.p != nullptr;
  const HeadView* vs[4] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o};
  for (int i = 0; i < 4; ++i)
    if (vs[i]->ss <= 0 || vs[i]->ss >= (1 << 23)) return 0;
  return 1;
}

template <int DH>
static int launch_fwd_wk(const AttnArgs& a, cudaStream_t st) {
  t2::FwdMaps maps;
  memset(&maps, 0, sizeof(maps));
  const bool pre = a.q_rot.p != nullptr;
  const HeadView* vs[4] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o};
  const int seq[4] = {a.Sq, a.Sk, a.Sk, a.Sq};
  const int rows[4] = {128, a.Sk, a.Sk, 128};                 // K and V: one box holds every key
  for (int i = 0; i < 4; ++i) {
    int rc = tma::make_head_map(&maps.m[2 * i], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 64, rows[i], 1);
    if (rc == 0 && DH > 64) rc = tma::make_head_map(&maps.m[2 * i + 1], *vs[i], a.N, a.n2_count, a.H, seq[i], DH, 32, rows[i], 1);
    if (rc) return rc;
  }
  const t2::MsDims dm = {(a.Sq + 127) / 128, 1};
  const int64_t items = (int64_t)a.N * a.H * dm.nq;
  if (items > 0x7fffffff) return MEANT_EUNSUPPORTED;
  auto kern = t2::attn2_fwd_wk_kernel<DH>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)t2::FwdWkL<DH>::TOTAL);
  if (e != cudaSuccess) return (int)e;
  const int grid = (int)(items < num_sms() ? items : num_sms());
  kern<<<grid, t2::F_NT, t2::FwdWkL<DH>::TOTAL, st>>>(a, (int)items, maps, dm);
  MEANT_LAUNCH_CHECK();
  return 0;
}

int attn_tc2_fwd_wk(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch_fwd_wk<96>(a, st);
  return launch_fwd_wk<64>(a, st);
}

int attn_tc2_bwd_ms(const AttnArgs& a, int parts, cudaStream_t st) {
  if (a.Dh == 96) return launch_bwd_ms<96, 6>(a, parts, st);
  return launch_bwd_ms<64, 8>(a, parts, st);
}

int attn_tc2_fwd_ms(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch_fwd_ms<96, 6>(a, st);
  return launch_fwd_ms<64, 8>(a, st);
}

// MEANT_EUNSUPPORTED: a view the tensor maps cannot describe (the caller falls back to the multi-tile kernels)
int attn_tc2_fwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch_fwd<96, 6>(a, st);
  return launch_fwd<64, 8>(a, st);
}

int attn_tc2_bwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch_bwd<96, 6>(a, st);
  return launch_bwd<64, 8>(a, st);
}

}  // namespace meant
