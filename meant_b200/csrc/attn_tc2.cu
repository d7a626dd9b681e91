// Persistent, warp-specialised fused rotary/xPos attention for sequences that fit one 128-row
// tile (Sq, Sk <= 128: the MEANT text path S = 128, TimeSformer time groups), forward and
// backward, on tcgen05 tensor cores with TMEM accumulators.  bf16 in, fp32 accumulate.
//
// One CTA per SM walks (sequence, head) work items blockIdx.x, blockIdx.x + gridDim.x, ...
// (consecutive CTAs take the heads of one sequence, so the packed `b s (h d)` rows they read
// are contiguous in DRAM).  384 threads, two roles:
//   loader warps 8-11  cp.async (LDGSTS, 16 B, zero-filled past S) Q/K/V[/dO] tiles straight from
//                      the packed projections into the shared "chunk layout" (tc_utils.cuh),
//                      several items ahead; rotate + xPos-scale Q and K rows IN PLACE from
//                      per-thread register tables (thread <-> row, fp16 pairs, loaded once);
//                      issue the first MMA(s) of the item; later drain the staged results to
//                      global memory with coalesced 16-byte stores (un-rotating dQ/dK first).
//   compute warps 0-7  warps w and w+4 share TMEM lanes (query rows) 32(w%4).. and split the key
//                      columns in interleaved 32-column blocks (balanced under the causal mask);
//                      softmax (forward) or P/dS recompute (backward) straight out of TMEM with
//                      one shared-memory exchange of the row max / row sums, bf16 P/dS to shared
//                      memory, issue the dependent MMAs, move accumulators TMEM -> bf16 -> staging.
// Hand-offs are mbarriers (tcgen05.commit for MMA completion, mbarrier.arrive for data), never
// __syncthreads, so loads of item j+1/j+2, the tensor core on item j and the stores of item
// j-1 overlap.  Both kernels are HBM-bound at S = 128 (see DESIGN.md for the byte counts).
#include "attn_common.cuh"
#include "tc_utils.cuh"
#include <cuda_fp16.h>
#include <math_constants.h>

namespace meant {
namespace t2 {

constexpr uint32_t CS = 128 * 16 + 16;   // chunk stride of a 128-row chunk-layout buffer
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr int BAR_C = 1, BAR_L = 2;      // named barriers of the compute (256 threads) / loader (128) groups
constexpr int NC = 256, NL = 128, NT = NC + NL;

template <int DH>
struct Geo {
  static constexpr int CPR = DH / 8;                 // 16-byte chunks per row
  static constexpr uint32_t OPER = CPR * CS;         // one 128 x DH operand tile
  static constexpr uint32_t PBUF = 16 * CS;          // one 128 x 128 P / dS tile
  static constexpr uint32_t PITCH = DH * 2 + 16;     // staging row pitch (bank-conflict free)
};

__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int ks) { return tc::make_desc(base + ks * 2 * CS, CS, 128); }
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int ks) { return tc::make_desc(base + ks * 256, 128, CS); }

__device__ __forceinline__ const __nv_bfloat16* head_base(const HeadView& v, int n, int n2c, int h) {
  const int n1 = n / n2c, n2 = n - n1 * n2c;
  return static_cast<const __nv_bfloat16*>(v.p) + (int64_t)n1 * v.sn1 + (int64_t)n2 * v.sn2 + (int64_t)h * v.sh;
}

// 128 threads: rows [0,128) of head (n, h) -> chunk layout at shared address sbase; rows >= S are zero filled.
template <int DH>
__device__ __forceinline__ void issue_tile(uint32_t sbase, const HeadView& v, int n, int n2c, int h, int S, int t) {
  constexpr int CPR = DH / 8;
  const __nv_bfloat16* base = head_base(v, n, n2c, h);
#pragma unroll 4
  for (int k = 0; k < CPR; ++k) {
    const int idx = t + k * 128;
    const int r = idx / CPR, ch = idx - r * CPR;
    const bool ok = r < S;
    tc::cp_async16(sbase + ch * CS + r * 16, base + (int64_t)(ok ? r : 0) * v.ss + ch * 8, ok ? 16u : 0u);
  }
}

// 128 threads: staged rows (row-major, PITCH bytes apart) -> head (n, h) of a global view, coalesced.
template <int DH>
__device__ __forceinline__ void drain_rows(const uint8_t* stg, const HeadView& v, int n, int n2c, int h, int S, int t) {
  constexpr int CPR = DH / 8;
  __nv_bfloat16* base = const_cast<__nv_bfloat16*>(head_base(v, n, n2c, h));
#pragma unroll 4
  for (int k = 0; k < CPR; ++k) {
    const int idx = t + k * 128;
    const int r = idx / CPR, ch = idx - r * CPR;
    if (r < S) {
      const uint4 val = *reinterpret_cast<const uint4*>(stg + r * Geo<DH>::PITCH + ch * 16);
      *reinterpret_cast<uint4*>(base + (int64_t)r * v.ss + ch * 8) = val;
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
// row in place; chunk ch of the row lives at row_ptr + ch * chunk_stride.
template <int RC, bool INV>
__device__ __forceinline__ void rot_row(uint8_t* row_ptr, uint32_t chunk_stride, int rc, const RotTab<RC>& T) {
#pragma unroll
  for (int ch = 0; ch < RC; ++ch) {
    if (ch < rc) {
      uint4* p = reinterpret_cast<uint4*>(row_ptr + ch * chunk_stride);
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
  static constexpr uint32_t TOTAL = BARS + 128;
  static_assert(G::PBUF <= 2 * G::OPER && 128 * G::PITCH <= 2 * G::OPER, "P / staging must fit in Q | K");
};

template <int DH, int RC>
__global__ void __launch_bounds__(NT, 1) attn2_fwd_kernel(AttnArgs a, int n_items) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  using L = FwdL<DH>;
  using G = Geo<DH>;
  constexpr int NST = L::NST;
  constexpr int HALF = DH / 2;                                 // accumulator columns per compute half
  // TMEM: one 128-column buffer per stage: S, then (once P is in shared memory) O over its first DH columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *s_full = bars, *aux_ready = bars + NST, *o_staged = bars + 2 * NST, *o_full = bars + 3 * NST;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&s_full[i], 1);
      tc::mbar_init(&aux_ready[i], NL);
      tc::mbar_init(&o_staged[i], NC);
      tc::mbar_init(&o_full[i], 1);
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
    tc::reg_inc<216>();   // 96 table registers live here; the compute warps give theirs up (no local-memory spills)
    const int t = tid - NC;
    RotTab<RC> tq, tk;
    int cur_h = -1;
    if (rc && !a.pos_is_head) {
      load_tab<RC>(tq, a.qcos, a.qsin, min(t, a.Sq - 1), a.rot_dim);
      load_tab<RC>(tk, a.kcos, a.ksin, min(t, a.Sk - 1), a.rot_dim);
    }
    // cp.async the three tiles of item j; returns this thread's padding-mask value (consumed later by post_aux)
    auto issue = [&](int j) -> float {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      float mval = 1.f;
      if (a.key_mask && t < a.Sk) mval = __ldg(a.key_mask + (int64_t)n * a.Sk + t);
      const uint32_t sb = s0 + st * L::STAGE;
      issue_tile<DH>(sb, a.q, n, a.n2_count, h, a.Sq, t);
      issue_tile<DH>(sb + G::OPER, a.k, n, a.n2_count, h, a.Sk, t);
      issue_tile<DH>(sb + 2 * G::OPER, a.v, n, a.n2_count, h, a.Sk, t);
      tc::cp_async_commit();
      return mval;
    };
    auto post_aux = [&](int j, float mval) {
      const int st = j % NST;
      reinterpret_cast<float*>(smem + L::AUX + st * 512)[t] = key_bias2(mval, t, a.Sk);
      tc::mbar_arrive(&aux_ready[st]);
    };
    auto copy_out = [&](int j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      tc::mbar_wait(&o_staged[st], (j / NST) & 1);
      drain_rows<DH>(smem + st * L::STAGE, a.o, n, a.n2_count, h, a.Sq, t);
      tc::bar_sync(BAR_L, NL);    // every staged row has been read before the stage is refilled
    };
    float mnext = issue(0);
    post_aux(0, mnext);
    for (int j = 0; j < n_my; ++j) {
      const int st = j % NST;
      const bool more = j + 1 < n_my;
      if (more) {
        if (j + 1 >= NST) copy_out(j + 1 - NST);
        mnext = issue(j + 1);
        tc::cp_async_wait<1>();
      } else {
        tc::cp_async_wait<0>();
      }
      tc::bar_sync(BAR_L, NL);      // item j's tiles have landed for every loader thread
      if (rc) {
        if (a.pos_is_head) {
          const int item = blockIdx.x + j * gridDim.x, h = item % a.H;
          if (h != cur_h) {
            load_tab<RC>(tq, a.qcos, a.qsin, h, a.rot_dim);
            load_tab<RC>(tk, a.kcos, a.ksin, h, a.rot_dim);
            cur_h = h;
          }
        }
        rot_row<RC, false>(smem + st * L::STAGE + t * 16, CS, rc, tq);
        rot_row<RC, false>(smem + st * L::STAGE + G::OPER + t * 16, CS, rc, tk);
      }
      tc::fence_async_smem();
      tc::bar_sync(BAR_L, NL);
      if (t == 0) {
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
        const uint32_t sq = s0 + st * L::STAGE, sk = sq + G::OPER;
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + st * 128, desc_kmajor(sq, ks), desc_kmajor(sk, ks), idesc, ks > 0);
        tc::mma_commit(&s_full[st]);
      }
      if (more) post_aux(j + 1, mnext);
    }
    for (int j = max(0, n_my - NST); j < n_my; ++j) copy_out(j);
  } else {
    // ------------------------------------------------------------------ compute group
    tc::reg_dec<144>();
    const int rg = warp & 3, half = warp >> 2;                    // TMEM lane quarter / column half
    const int t = rg * 32 + lane;                                 // query row
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale2 = a.scale * LOG2E;
    float* red_max = reinterpret_cast<float*>(smem + L::RED);     // [2][128]
    float* red_sum = red_max + 256;                               // [2][128]
    int nblk = (a.Sk + 31) >> 5;                                  // 32-key blocks with any unmasked key for this warp's rows
    if (a.causal) nblk = min(nblk, rg + 1);
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H, st = j % NST;
      const uint32_t par = (j / NST) & 1;
      tc::mbar_wait(&aux_ready[st], par);
      tc::mbar_wait(&s_full[st], par);
      tc::fence_after_sync();
      const float4* b4 = reinterpret_cast<const float4*>(smem + L::AUX + st * 512);
      const uint32_t ts = lane_base + st * 128;
      float x[64];
      float m = -CUDART_INF_F;
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {                                           // warp-uniform
          float sv[32];
          tc::tmem_ld32(ts + blk * 32, sv);
          tc::tmem_ld_wait();
          const bool diag = a.causal && blk == rg;
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 b = b4[(blk * 32 + e) >> 2];
            const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              float v = fmaf(sv[e + u], scale2, bb[u]);
              if (diag && e + u > lane) v = -CUDART_INF_F;
              x[i * 32 + e + u] = v;
              m = fmaxf(m, v);
            }
          }
        }
      }
      red_max[half * 128 + t] = m;
      tc::bar_sync(BAR_C, NC);
      m = fmaxf(m, red_max[(half ^ 1) * 128 + t]);
      const float m_use = m == -CUDART_INF_F ? 0.f : m;
      float l = 0.f;
      uint8_t* pbuf = smem + st * L::STAGE;                         // P over the consumed Q | K
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            uint4 pk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              const float p0 = tc::ex2(x[i * 32 + q * 8 + 2 * u] - m_use), p1 = tc::ex2(x[i * 32 + q * 8 + 2 * u + 1] - m_use);
              l += p0 + p1;
              ph[u] = __floats2bfloat162_rn(p0, p1);
            }
            *reinterpret_cast<uint4*>(pbuf + (blk * 4 + q) * CS + t * 16) = pk;
          }
        } else if (blk * 32 < n_s) {
#pragma unroll
          for (int q = 0; q < 4; ++q) *reinterpret_cast<uint4*>(pbuf + (blk * 4 + q) * CS + t * 16) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      red_sum[half * 128 + t] = l;
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::bar_sync(BAR_C, NC);
      if (tid == 0) {
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, DH, 0, 1);
        const uint32_t sp = s0 + st * L::STAGE, sv = sp + 2 * G::OPER;
        for (int ks = 0; ks < n_s / 16; ++ks) tc::mma_ss(tmem + st * 128, desc_kmajor(sp, ks), desc_mnmajor(sv, ks), idesc, ks > 0);
        tc::mma_commit(&o_full[st]);
      }
      l += red_sum[(half ^ 1) * 128 + t];
      tc::mbar_wait(&o_full[st], par);
      tc::fence_after_sync();
      stage_acc_cols<HALF>(ts + half * HALF, smem + st * L::STAGE + t * G::PITCH + half * HALF * 2, l > 0.f ? 1.0f / l : 0.f);
      if (half == 0 && t < a.Sq && a.lse) a.lse[((int64_t)n * a.H + h) * a.Sq + t] = m_use * LN2 + logf(l);
      tc::fence_before_sync();
      tc::mbar_arrive(&o_staged[st]);
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
  static constexpr uint32_t AUX = DS + G::PBUF;                // bias2[128] | lse*log2e[128] | delta exchange [2][128]
  static constexpr uint32_t BARS = AUX + 2048;
  static constexpr uint32_t TOTAL = BARS + 128;
  static_assert(128 * G::PITCH <= G::PBUF, "gradient staging must fit in the P / dS buffers");
};

template <int DH, int RC>
__global__ void __launch_bounds__(NT, 1) attn2_bwd_kernel(AttnArgs a, int n_items) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t tmem_slot;
  using L = BwdL<DH>;
  using G = Geo<DH>;
  constexpr int HALF = DH / 2;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 352, T_DQ = 0;   // dQ reuses the consumed S columns
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + L::BARS);
  uint64_t *sdp_full = bars, *grads_full = bars + 1, *aux_ready = bars + 2, *staged_x = bars + 3, *staged_y = bars + 4,
           *x_free = bars + 5, *y_free = bars + 6;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    tc::mbar_init(sdp_full, 1);
    tc::mbar_init(grads_full, 1);
    tc::mbar_init(aux_ready, NL);
    tc::mbar_init(staged_x, NC);
    tc::mbar_init(staged_y, NC);
    tc::mbar_init(x_free, NL);
    tc::mbar_init(y_free, NL);
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
  uint8_t* stg_x = smem + L::P;     // gradient staging areas (free once the gradient MMAs are done)
  uint8_t* stg_y = smem + L::DS;
  float* aux_bias = reinterpret_cast<float*>(smem + L::AUX);
  float* aux_lse2 = aux_bias + 128;
  float* red_delta = aux_bias + 256;                                // [2][128]

  if (tid >= NC) {
    // ------------------------------------------------------------------ loader group
    tc::reg_inc<216>();   // 96 table registers live here; the compute warps give theirs up (no local-memory spills)
    const int t = tid - NC;
    RotTab<RC> tq, tk;
    int cur_h = -1;
    if (rc && !a.pos_is_head) {
      load_tab<RC>(tq, a.qcos, a.qsin, min(t, a.Sq - 1), a.rot_dim);
      load_tab<RC>(tk, a.kcos, a.ksin, min(t, a.Sk - 1), a.rot_dim);
    }
    auto tabs_for = [&](int h) {
      if (rc && a.pos_is_head && h != cur_h) {
        load_tab<RC>(tq, a.qcos, a.qsin, h, a.rot_dim);
        load_tab<RC>(tk, a.kcos, a.ksin, h, a.rot_dim);
        cur_h = h;
      }
    };
    uint32_t xs = 0, ys = 0;
    // drain the three gradients of item j as the compute group stages them: dQ (X), dV (Y), dK (X)
    auto service = [&](int j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H;
      tabs_for(h);
      tc::mbar_wait(staged_x, xs & 1); ++xs;
      if (rc) { rot_row<RC, true>(stg_x + t * G::PITCH, 16, rc, tq); tc::bar_sync(BAR_L, NL); }
      drain_rows<DH>(stg_x, a.dq, n, a.n2_count, h, a.Sq, t);
      tc::mbar_arrive(x_free);
      tc::mbar_wait(staged_y, ys & 1); ++ys;
      drain_rows<DH>(stg_y, a.dv, n, a.n2_count, h, a.Sk, t);
      tc::mbar_arrive(y_free);
      tc::mbar_wait(staged_x, xs & 1); ++xs;
      if (rc) { rot_row<RC, true>(stg_x + t * G::PITCH, 16, rc, tk); tc::bar_sync(BAR_L, NL); }
      drain_rows<DH>(stg_x, a.dk, n, a.n2_count, h, a.Sk, t);
      tc::mbar_arrive(x_free);
    };
    for (int j = 0; j < n_my; ++j) {
      const int item = blockIdx.x + j * gridDim.x, n = item / a.H, h = item - n * a.H;
      if (j > 0) tc::mbar_wait(grads_full, (j - 1) & 1);           // operand tiles of item j-1 are consumed
      float mval = 1.f, lse = 0.f;
      if (a.key_mask && t < a.Sk) mval = __ldg(a.key_mask + (int64_t)n * a.Sk + t);
      if (t < a.Sq) lse = __ldg(a.lse + ((int64_t)n * a.H + h) * a.Sq + t);
      issue_tile<DH>(sq, a.q, n, a.n2_count, h, a.Sq, t);
      issue_tile<DH>(sk, a.k, n, a.n2_count, h, a.Sk, t);
      issue_tile<DH>(sv, a.v, n, a.n2_count, h, a.Sk, t);
      issue_tile<DH>(sdo, a.o, n, a.n2_count, h, a.Sq, t);
      tc::cp_async_commit();
      if (j > 0) service(j - 1);
      aux_bias[t] = key_bias2(mval, t, a.Sk);
      aux_lse2[t] = lse * LOG2E;
      tc::mbar_arrive(aux_ready);
      tc::cp_async_wait<0>();
      tc::bar_sync(BAR_L, NL);
      if (rc) {
        tabs_for(h);
        rot_row<RC, false>(smem + L::Q + t * 16, CS, rc, tq);
        rot_row<RC, false>(smem + L::K + t * 16, CS, rc, tk);
      }
      tc::fence_async_smem();
      tc::bar_sync(BAR_L, NL);
      if (t == 0) {
        tc::fence_after_sync();
        constexpr uint32_t idesc = tc::make_idesc(128, 128, 0, 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_kmajor(sq, ks), desc_kmajor(sk, ks), idesc, ks > 0);
#pragma unroll
        for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_kmajor(sdo, ks), desc_kmajor(sv, ks), idesc, ks > 0);
        tc::mma_commit(sdp_full);
      }
    }
    if (n_my > 0) service(n_my - 1);
  } else {
    // ------------------------------------------------------------------ compute group
    tc::reg_dec<144>();
    const int rg = warp & 3, half = warp >> 2;
    const int t = rg * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(rg * 32) << 16);
    const float scale = a.scale, scale2 = a.scale * LOG2E;
    const float4* b4 = reinterpret_cast<const float4*>(aux_bias);
    const bool qrow = t < a.Sq;
    int nblk = (a.Sk + 31) >> 5;
    if (a.causal) nblk = min(nblk, rg + 1);
    uint8_t* my_x = stg_x + t * G::PITCH + half * HALF * 2;
    uint8_t* my_y = stg_y + t * G::PITCH + half * HALF * 2;
    uint32_t xu = 0, yu = 0;
    for (int j = 0; j < n_my; ++j) {
      tc::mbar_wait(aux_ready, j & 1);
      tc::mbar_wait(sdp_full, j & 1);
      tc::fence_after_sync();
      const float lse2 = aux_lse2[t];
      uint32_t pk[32];                                              // this thread's 64 columns of P, bf16 pairs
      float delta = 0.f;                                            // rowsum(P * dP) == rowsum(dO * O)
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        if (blk < nblk) {
          float sv32[32], dp[32];
          tc::tmem_ld32(lane_base + T_S + blk * 32, sv32);
          tc::tmem_ld32(lane_base + T_DP + blk * 32, dp);
          tc::tmem_ld_wait();
          const bool diag = a.causal && blk == rg;
#pragma unroll
          for (int e = 0; e < 32; e += 4) {
            const float4 b = b4[(blk * 32 + e) >> 2];
            const float bb[4] = {b.x, b.y, b.z, b.w};
            float p[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
              p[u] = tc::ex2(fmaf(sv32[e + u], scale2, bb[u]) - lse2);
              if (!qrow || (diag && e + u > lane)) p[u] = 0.f;
              delta = fmaf(p[u], dp[e + u], delta);
            }
            const __nv_bfloat162 p01 = __floats2bfloat162_rn(p[0], p[1]), p23 = __floats2bfloat162_rn(p[2], p[3]);
            pk[i * 16 + e / 2] = *reinterpret_cast<const uint32_t*>(&p01);
            pk[i * 16 + e / 2 + 1] = *reinterpret_cast<const uint32_t*>(&p23);
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) pk[i * 16 + e] = 0u;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(smem + L::P + (blk * 4 + q) * CS + t * 16) =
              make_uint4(pk[i * 16 + q * 4], pk[i * 16 + q * 4 + 1], pk[i * 16 + q * 4 + 2], pk[i * 16 + q * 4 + 3]);
      }
      red_delta[half * 128 + t] = delta;
      tc::bar_sync(BAR_C, NC);
      delta += red_delta[(half ^ 1) * 128 + t];
#pragma unroll
      for (int i = 0; i < 2; ++i) {
        const int blk = 2 * i + half;
        uint32_t dsk[16];
        if (blk < nblk) {
          float dp[32];
          tc::tmem_ld32(lane_base + T_DP + blk * 32, dp);
          tc::tmem_ld_wait();
#pragma unroll
          for (int e = 0; e < 16; ++e) {
            const float2 p = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk[i * 16 + e]));
            const __nv_bfloat162 d2 = __floats2bfloat162_rn(p.x * (dp[2 * e] - delta) * scale, p.y * (dp[2 * e + 1] - delta) * scale);
            dsk[e] = *reinterpret_cast<const uint32_t*>(&d2);
          }
        } else {
#pragma unroll
          for (int e = 0; e < 16; ++e) dsk[e] = 0u;
        }
#pragma unroll
        for (int q = 0; q < 4; ++q)
          *reinterpret_cast<uint4*>(smem + L::DS + (blk * 4 + q) * CS + t * 16) =
              make_uint4(dsk[q * 4], dsk[q * 4 + 1], dsk[q * 4 + 2], dsk[q * 4 + 3]);
      }
      tc::fence_async_smem();
      tc::fence_before_sync();
      tc::bar_sync(BAR_C, NC);
      if (tid == 0) {
        tc::fence_after_sync();
        constexpr uint32_t id_t = tc::make_idesc(128, DH, 1, 1);     // A^T (MN-major) x B (MN-major)
        constexpr uint32_t id_n = tc::make_idesc(128, DH, 0, 1);     // A (K-major)    x B (MN-major)
        const int kq = (a.Sq + 15) >> 4, kk = (a.Sk + 15) >> 4;
        for (int ks = 0; ks < kq; ++ks) tc::mma_ss(tmem + T_DV, desc_mnmajor(sp, ks), desc_mnmajor(sdo, ks), id_t, ks > 0);   // dV = P^T dO
        for (int ks = 0; ks < kq; ++ks) tc::mma_ss(tmem + T_DK, desc_mnmajor(sds, ks), desc_mnmajor(sq, ks), id_t, ks > 0);   // dK = dS^T Q
        for (int ks = 0; ks < kk; ++ks) tc::mma_ss(tmem + T_DQ, desc_kmajor(sds, ks), desc_mnmajor(sk, ks), id_n, ks > 0);    // dQ = dS K
        tc::mma_commit(grads_full);
      }
      tc::mbar_wait(grads_full, j & 1);
      tc::fence_after_sync();
      // dQ first: it frees the S columns for the next item's score MMA
      if (xu >= 1) tc::mbar_wait(x_free, (xu - 1) & 1);
      ++xu;
      stage_acc_cols<HALF>(lane_base + T_DQ + half * HALF, my_x, 1.0f);
      tc::fence_before_sync();
      tc::mbar_arrive(staged_x);
      if (yu >= 1) tc::mbar_wait(y_free, (yu - 1) & 1);
      ++yu;
      stage_acc_cols<HALF>(lane_base + T_DV + half * HALF, my_y, 1.0f);
      tc::fence_before_sync();
      tc::mbar_arrive(staged_y);
      tc::mbar_wait(x_free, (xu - 1) & 1);
      ++xu;
      stage_acc_cols<HALF>(lane_base + T_DK + half * HALF, my_x, 1.0f);
      tc::fence_before_sync();
      tc::mbar_arrive(staged_x);
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

int attn_tc2_fwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch2(t2::attn2_fwd_kernel<96, 6>, t2::FwdL<96>::TOTAL, a, st);
  return launch2(t2::attn2_fwd_kernel<64, 8>, t2::FwdL<64>::TOTAL, a, st);
}

int attn_tc2_bwd(const AttnArgs& a, cudaStream_t st) {
  if (a.Dh == 96) return launch2(t2::attn2_bwd_kernel<96, 6>, t2::BwdL<96>::TOTAL, a, st);
  return launch2(t2::attn2_bwd_kernel<64, 8>, t2::BwdL<64>::TOTAL, a, st);
}

}  // namespace meant
