// Fused xPos/rotary flash attention on tcgen05 tensor cores (sm_100a), bf16 in / fp32 accumulate.
//
// Forward, one CTA per (sequence, head, 128-query tile), 128 threads, thread t <-> query row t
// <-> TMEM lane t (softmax is thread-local, no shuffles):
//   stage   Q, K (rotated + xPos-scaled in registers on the way in) and V as 16-byte chunks into
//           the shared "chunk layout" (tc_utils.cuh) straight from the packed `b s (h d)` projections
//   MMA 1   S = Q K^T            (A, B K-major)                    -> TMEM cols [0,128)
//   softmax scale, causal / padding mask, online max/sum in fp32; P (bf16) -> shared memory
//   MMA 2   PV = P V             (A K-major, B = V MN-major: no transpose copy) -> TMEM cols [128,128+DH)
//   O accumulates in registers across key tiles; epilogue writes `b s (h d)` and the LSE.
// Backward (Sq, Sk <= 128: the text path), one CTA per (sequence, head), recompute based:
//   S = Q K^T, dP = dO V^T -> P, dS (bf16, shared) -> dV = P^T dO, dK = dS^T Q, dQ = dS K,
//   every transposed operand read MN-major from the buffer that was staged once; dQ/dK are
//   un-rotated (transpose rotation) in the epilogue.
// Backward, multi-tile (Sq or Sk > 128: the 196-patch vision path, S = 512 text), two launches of one kernel,
// 256 threads (threads t and t+128 share query row / TMEM lane t and split the key columns):
//   MODE 1  one CTA per (sequence, head, 128-query tile): Q, dO staged once, loop over key tiles,
//           dQ += dS K accumulates in TMEM; also writes delta = rowsum(dO * O)           (runs first)
//   MODE 0  one CTA per (sequence, head, 128-key tile): K, V staged once, loop over query tiles,
//           dV += P^T dO and dK += dS^T Q accumulate in TMEM across the loop
//   Operands are staged with cp.async (all 16-byte copies of a tile in flight at once, the streamed pair
//   double-buffered so tile j+1 loads under tile j's MMAs and softmax recompute) and rotated in place.
//   S and dP are recomputed in both (8 MMA chains per tile pair instead of 5): no atomics, no fp32
//   workspace, bit-reproducible gradients.  Causal problems skip the tile pairs above the diagonal.
// Work per sequence and the roofline that bounds it: DESIGN.md.
#include <type_traits>
#include "attn_common.cuh"
#include "tc_utils.cuh"
#include "tma_utils.cuh"
#include <math_constants.h>
#include <atomic>
#include <string.h>
#include <stdlib.h>

namespace meant {

constexpr int TQ = 128;                       // query rows per CTA == TMEM lanes
constexpr int TK = 128;                       // keys per tile
constexpr uint32_t CS128 = 128 * 16;          // chunk stride of a 128-row chunk-layout buffer (unpadded: a TMA box lands in it)
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

// Optional phase timing of the multi-tile forward (build with MEANT_PROF=1): thread 0 of CTA (0,0,0) adds
// clock64 deltas per phase; read with meant_debug_prof_mt().  Never compiled into product builds.
#ifdef MEANT_PROF
__device__ unsigned long long g_prof_mt[32];
#define MT_PROF_INIT() const bool prof_on = threadIdx.x == 0 && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0; long long prof_t = clock64()
#define MT_PROF(slot) do { if (prof_on) { const long long now_ = clock64(); atomicAdd(&g_prof_mt[slot], (unsigned long long)(now_ - prof_t)); prof_t = now_; } } while (0)
#else
#define MT_PROF_INIT() do {} while (0)
#define MT_PROF(slot) do {} while (0)
#endif

// Stage rows [s0, s0+128) of a head view into chunk layout, optionally rotating pairs with the
// per-element tables (cos/sin already carry the xPos scale).  Lanes walk 16-byte chunks of a
// row first, so global reads are contiguous DH*2-byte segments.
template <int DH, int NT = 128>
__device__ __forceinline__ void stage_rows(uint8_t* smem, const HeadView& v, int n, int n2c, int h, int s0, int S,
                                           const float* __restrict__ cs, const float* __restrict__ sn, int rot_dim,
                                           int pos_is_head) {
  constexpr int CPR = DH / 8;
  for (int idx = threadIdx.x; idx < 128 * CPR; idx += NT) {
    const int r = idx / CPR, ch = idx - r * CPR;
    const int srow = s0 + r;
    uint4 raw = make_uint4(0u, 0u, 0u, 0u);
    if (srow < S) {
      raw = *reinterpret_cast<const uint4*>(view_ptr<__nv_bfloat16>(v, n, n2c, srow, h) + ch * 8);
      if (cs != nullptr && ch * 8 < rot_dim) {
        const int pos = pos_is_head ? h : srow;
        const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * rot_dim + ch * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * rot_dim + ch * 8);
        const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
        __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
        float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
        float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
        h2[0] = __floats2bfloat162_rn(p0.x * ca.x - p0.y * sa.x, p0.y * ca.y + p0.x * sa.y);
        h2[1] = __floats2bfloat162_rn(p1.x * ca.z - p1.y * sa.z, p1.y * ca.w + p1.x * sa.w);
        h2[2] = __floats2bfloat162_rn(p2.x * cb.x - p2.y * sb.x, p2.y * cb.y + p2.x * sb.y);
        h2[3] = __floats2bfloat162_rn(p3.x * cb.z - p3.y * sb.z, p3.y * cb.w + p3.x * sb.w);
      }
    }
    *reinterpret_cast<uint4*>(smem + ch * CS128 + r * 16) = raw;
  }
}

// K-major operand (rows = M/N, cols = K) / MN-major operand (rows = K, cols = M/N) descriptors of
// a 128-row chunk-layout buffer at K-step ks (16 K-elements per MMA).
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t base, int ks) {
  return tc::make_desc(base + ks * 2 * CS128, CS128, 128);
}
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t base, int ks) {
  return tc::make_desc(base + ks * 256, 128, CS128);
}

__device__ __forceinline__ float masked_score(float s, float scale, const float* mrow, int kj, bool valid) {
  float x = s * scale;
  if (mrow != nullptr && valid) x += (1.0f - __ldg(mrow + kj)) * -1e9f;   // meant/xPosAttention.py:54-56
  return valid ? x * LOG2E : -CUDART_INF_F;
}

template <int DH>
struct FwdSmem {
  static constexpr uint32_t OPER = (DH / 8) * CS128;       // Q, K or V tile
  static constexpr uint32_t PBUF = (TK / 8) * CS128;       // P tile
  static constexpr uint32_t Q = 0, K = OPER, V = 2 * OPER, P = 3 * OPER, TOTAL = 3 * OPER + PBUF + 1024;   // + alignment slack
};

// Asynchronous staging (multi-tile kernels, 256 threads): rows [s0, s0+128) of a head view go to the chunk layout
// as 16-byte cp.async copies (rows past S zero-filled), all in flight at once; the caller commits / waits the
// group.  Thread -> (row tid/2, chunk 2*round + tid%2): a lane pair reads one whole 32-byte sector, and the chunk
// index is warp-uniform up to its parity, so "is this chunk rotated" never diverges (rot_dim % 16 == 0).
// rotate_staged then rotates the SAME chunks this thread copied (identical walk), in place, with all table
// loads of a tile issued before the first use.
template <int DH, int NT>
__device__ __forceinline__ void stage_async(uint8_t* smem, const HeadView& v, int n, int n2c, int h, int s0, int S) {
  static_assert(NT == 256, "thread -> (row, chunk parity) mapping");
  constexpr int CPR = DH / 8;
  const int r = threadIdx.x >> 1, par = threadIdx.x & 1;
  const bool ok = s0 + r < S;
  const __nv_bfloat16* src = view_ptr<__nv_bfloat16>(v, n, n2c, ok ? s0 + r : 0, h) + par * 8;
  const uint32_t dst = tc::smem_u32(smem) + par * CS128 + r * 16;
  const uint32_t bytes = ok ? 16u : 0u;
#pragma unroll
  for (int it = 0; it < CPR / 2; ++it) tc::cp_async16(dst + it * 2 * CS128, src + it * 16, bytes);
}

// RR = rotated rounds (rot_dim / 16)
template <int DH, int RR>
__device__ __forceinline__ void rotate_rounds(uint8_t* smem, int pos, const float* __restrict__ cs,
                                              const float* __restrict__ sn, int rot_dim) {
  const int r = threadIdx.x >> 1, par = threadIdx.x & 1;
  const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * rot_dim + par * 8);
  const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * rot_dim + par * 8);
  float4 ca[RR], cb[RR], sa[RR], sb[RR];
#pragma unroll
  for (int it = 0; it < RR; ++it) {
    ca[it] = __ldg(c4 + it * 4);
    cb[it] = __ldg(c4 + it * 4 + 1);
    sa[it] = __ldg(s4 + it * 4);
    sb[it] = __ldg(s4 + it * 4 + 1);
  }
  uint8_t* slot0 = smem + par * CS128 + r * 16;
#pragma unroll
  for (int it = 0; it < RR; ++it) {
    uint4* slot = reinterpret_cast<uint4*>(slot0 + it * 2 * CS128);
    uint4 raw = *slot;
    __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
    const float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
    const float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
    h2[0] = __floats2bfloat162_rn(p0.x * ca[it].x - p0.y * sa[it].x, p0.y * ca[it].y + p0.x * sa[it].y);
    h2[1] = __floats2bfloat162_rn(p1.x * ca[it].z - p1.y * sa[it].z, p1.y * ca[it].w + p1.x * sa[it].w);
    h2[2] = __floats2bfloat162_rn(p2.x * cb[it].x - p2.y * sb[it].x, p2.y * cb[it].y + p2.x * sb[it].y);
    h2[3] = __floats2bfloat162_rn(p3.x * cb[it].z - p3.y * sb[it].z, p3.y * cb[it].w + p3.x * sb[it].w);
    *slot = raw;
  }
}

template <int DH, int NT>
__device__ __forceinline__ void rotate_staged(uint8_t* smem, int h, int s0, int S, const float* __restrict__ cs,
                                              const float* __restrict__ sn, int rot_dim, int pos_is_head) {
  static_assert(NT == 256, "same walk as stage_async");
  if (cs == nullptr) return;
  // rows past S were zero-filled and rotate to zero: no row predicate, the table row is clamped instead
  const int pos = pos_is_head ? h : min(s0 + (int)(threadIdx.x >> 1), S - 1);
  if (rot_dim == 48 && DH >= 48) rotate_rounds<DH, 3>(smem, pos, cs, sn, rot_dim);
  else if (rot_dim == 64 && DH >= 64) rotate_rounds<DH, 4>(smem, pos, cs, sn, rot_dim);
  else {                                    // any rot_dim % 8 == 0: per-chunk predicate
    const int r = threadIdx.x >> 1, par = threadIdx.x & 1;
    for (int ch = par; ch * 8 < rot_dim; ch += 2) {
      uint4* slot = reinterpret_cast<uint4*>(smem + ch * CS128 + r * 16);
      uint4 raw = *slot;
      const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * rot_dim + ch * 8);
      const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * rot_dim + ch * 8);
      const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
      __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
      const float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
      const float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
      h2[0] = __floats2bfloat162_rn(p0.x * ca.x - p0.y * sa.x, p0.y * ca.y + p0.x * sa.y);
      h2[1] = __floats2bfloat162_rn(p1.x * ca.z - p1.y * sa.z, p1.y * ca.w + p1.x * sa.w);
      h2[2] = __floats2bfloat162_rn(p2.x * cb.x - p2.y * sb.x, p2.y * cb.y + p2.x * sb.y);
      h2[3] = __floats2bfloat162_rn(p3.x * cb.z - p3.y * sb.z, p3.y * cb.w + p3.x * sb.w);
      *slot = raw;
    }
  }
}

// additive score term of key kj in the log2 domain: (1 - m) * -1e9 * log2(e), -inf past the last key
__device__ __forceinline__ float key_term(const float* mrow, int kj, int Sk) {
  if (kj >= Sk) return -CUDART_INF_F;
  return mrow != nullptr ? (1.0f - __ldg(mrow + kj)) * (-1e9f * LOG2E) : 0.f;
}

// Tensor maps of the multi-tile forward: q (rotated copy when there is one), k, v as chunk-layout boxes.
struct FwdMtMaps {
  CUtensorMap m[3];
  CUtensorMap o[2];     // output as head-view boxes (64 columns SWIZZLE_128B [+ 32 columns SWIZZLE_64B])
  int use_tma;
};

template <int DH>
__global__ void __launch_bounds__(256, 2) attn_tc_fwd_kernel(AttnArgs a, const __grid_constant__ FwdMtMaps maps) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzled output boxes: 1024-byte atoms
  __shared__ uint64_t bar_s, bar_o, bar_k, bar_v;   // MMA chains done; TMA: Q|K and V tiles landed
  __shared__ uint32_t tmem_slot;
  __shared__ float kterm[TK];
  __shared__ float xch[2][TQ];          // row max / row sum exchange between the two column halves
  using L = FwdSmem<DH>;
  constexpr int NT = 256, HC = DH / 2, HK = TK / 2;   // O columns / key columns per thread
  const int tid = threadIdx.x, warp = tid >> 5, t = tid & 127, hf = tid >> 7;
  const int q0 = blockIdx.x * TQ, h = blockIdx.y, n = blockIdx.z;
  const int n1 = n / a.n2_count, n2 = n - n1 * a.n2_count, hc = h * (DH / 8);   // chunk-map coordinates of this head
  if (tid == 0) {
    tc::mbar_init(&bar_s, 1);
    tc::mbar_init(&bar_o, 1);
    tc::mbar_init(&bar_k, 1);
    tc::mbar_init(&bar_v, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  // Shared extra key / value row (meant_attn_args.kx / vx): k and v hold Skb = Sk - 1 rows, the tile that contains key
  // Sk - 1 gets that row written in after it has landed (threads [0, CPR): K's 16-byte chunks, [CPR, 2 CPR): V's).
  const bool has_x = a.kx.p != nullptr;
  const int Skb = a.Sk - (has_x ? 1 : 0);
  uint4 xreg = make_uint4(0u, 0u, 0u, 0u);
  if (has_x && tid < 2 * (DH / 8)) {
    const HeadView& xv = tid < DH / 8 ? a.kx : a.vx;
    xreg = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(xv.p) + (int64_t)n1 * xv.sn1 + (int64_t)h * xv.sh +
                                           (tid % (DH / 8)) * 8);
  }
  // cp.async groups, oldest first: [Q + K(0)] [V(0)] then per tile [K(kt+1)] [V(kt+1)] (empty past the end)
  const bool pre = a.q_rot.p != nullptr;          // Q, K already rotated by meant_attn_prerotate
  const HeadView& qv = pre ? a.q_rot : a.q;
  const HeadView& kv = pre ? a.k_rot : a.k;
  const float* qcos = pre ? nullptr : a.qcos;
  const float* kcos = pre ? nullptr : a.kcos;
  // Operand tiles: one TMA box each, straight into the chunk layout (make_chunk_map), issued by thread 0 and counted
  // in bytes on bar_k (Q with K(0), then K(kt+1)) / bar_v; views TMA cannot describe keep the cp.async groups.
  const bool use_tma = maps.use_tma != 0;
  uint32_t par_k = 0, par_v = 0;
  if (use_tma) {
    __syncthreads();                    // the barriers exist before anybody polls them
    if (tid == 0) {
      tma::mbar_expect_tx(&bar_k, 2 * L::OPER);
      tma::load_box(tc::smem_u32(smem + L::Q), &maps.m[0], &bar_k, 0, q0, hc, n2, n1);
      tma::load_box(tc::smem_u32(smem + L::K), &maps.m[1], &bar_k, 0, 0, hc, n2, n1);
      tma::mbar_expect_tx(&bar_v, L::OPER);
      tma::load_box(tc::smem_u32(smem + L::V), &maps.m[2], &bar_v, 0, 0, hc, n2, n1);
    }
  } else {
    stage_async<DH, NT>(smem + L::Q, qv, n, a.n2_count, h, q0, a.Sq);
    stage_async<DH, NT>(smem + L::K, kv, n, a.n2_count, h, 0, Skb);
    tc::cp_async_commit();
    stage_async<DH, NT>(smem + L::V, a.v, n, a.n2_count, h, 0, Skb);
    tc::cp_async_commit();
  }

  const int qi = q0 + t;
  const float* mrow = a.key_mask ? a.key_mask + (int64_t)n * a.Sk : nullptr;
  const float sc2 = a.scale * LOG2E;
  MT_PROF_INIT();
  float m_run = -CUDART_INF_F, l_run = 0.f;    // l_run: this half's columns only, combined at the end
  float o_acc[HC];
#pragma unroll
  for (int i = 0; i < HC; ++i) o_acc[i] = 0.f;
  const int last_q = min(q0 + TQ, a.Sq) - 1;
  int n_tiles = (a.Sk + TK - 1) / TK;
  if (a.causal) n_tiles = min(n_tiles, last_q / TK + 1);
  uint32_t phase = 0, tmem = 0;
  const uint32_t sq = tc::smem_u32(smem + L::Q), sk = tc::smem_u32(smem + L::K), sv = tc::smem_u32(smem + L::V),
                 sp = tc::smem_u32(smem + L::P);
#pragma unroll 1
  for (int kt = 0; kt < n_tiles; ++kt) {
    const int k0 = kt * TK;
    const bool more = kt + 1 < n_tiles;
    const bool diag = a.causal && k0 + TK - 1 > q0;       // some (row, key) of this tile pair is above the diagonal
    const bool plain = !diag && mrow == nullptr && k0 + TK <= a.Sk && sc2 > 0.f;   // no additive term, no masking
    MT_PROF(10);
    if (use_tma) {
      tc::mbar_wait(&bar_k, par_k);
      par_k ^= 1;
    } else {
      tc::cp_async_wait<1>();          // K(kt) (and Q) landed; V(kt) may still be in flight
    }
    MT_PROF(0);
    if (kt == 0) rotate_staged<DH, NT>(smem + L::Q, h, q0, a.Sq, qcos, a.qsin, a.rot_dim, a.pos_is_head);
    rotate_staged<DH, NT>(smem + L::K, h, k0, a.Sk, kcos, a.ksin, a.rot_dim, a.pos_is_head);
    const int xr = Skb - k0;                                  // tile row of the extra key, if it is in this tile
    if (has_x && xr >= 0 && xr < TK) {
      __syncthreads();                                        // row xr was zero-filled and "rotated" by another thread
      if (tid < DH / 8) *reinterpret_cast<uint4*>(smem + L::K + tid * CS128 + xr * 16) = xreg;
    }
    if (!plain && tid < TK) kterm[tid] = key_term(mrow, k0 + tid, a.Sk);
    MT_PROF(1);
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    tmem = tmem_slot;
    MT_PROF(2);
    if (tid == 0) {
      constexpr uint32_t idesc = tc::make_idesc(TQ, TK, 0, 0);
#pragma unroll
      for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem, desc_kmajor(sq, ks), desc_kmajor(sk, ks), idesc, ks > 0);
      tc::mma_commit(&bar_s);
    }
    tc::mbar_wait(&bar_s, phase);
    tc::fence_after_sync();
    MT_PROF(3);
    if (use_tma) {                                                                      // K is free: next tile
      if (more && tid == 0) {
        tma::mbar_expect_tx(&bar_k, L::OPER);
        tma::load_box(tc::smem_u32(smem + L::K), &maps.m[1], &bar_k, 0, k0 + TK, hc, n2, n1);
      }
    } else {
      if (more) stage_async<DH, NT>(smem + L::K, kv, n, a.n2_count, h, k0 + TK, Skb);
      tc::cp_async_commit();
    }
    MT_PROF(4);
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int cb = hf * HK;            // this thread's key columns: [cb, cb + HK)
    // pass 1: row max over my columns, then across the two halves
    float m_tile = -CUDART_INF_F, m_raw = -CUDART_INF_F;
#pragma unroll 1
    for (int c0 = cb; c0 < cb + HK; c0 += 32) {
      float s[32];
      tc::tmem_ld32(lane_base + c0, s);
      tc::tmem_ld_wait();
      if (plain) {              // sc2 > 0: the max of the raw scores, scaled once below
#pragma unroll
        for (int i = 0; i < 32; ++i) m_raw = fmaxf(m_raw, s[i]);
      } else {
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          float x = fmaf(s[i], sc2, kterm[c0 + i]);
          if (diag && k0 + c0 + i > qi) x = -CUDART_INF_F;
          m_tile = fmaxf(m_tile, x);
        }
      }
    }
    if (plain) m_tile = m_raw * sc2;
    xch[hf][t] = m_tile;
    __syncthreads();
    m_tile = fmaxf(m_tile, xch[hf ^ 1][t]);
    MT_PROF(5);
    const float m_new = fmaxf(m_run, m_tile);
    const float m_use = m_new == -CUDART_INF_F ? 0.f : m_new;
    const float corr = tc::ex2(m_run - m_use);   // m_run = -inf -> 0
    // pass 2: P = exp2(x - m), bf16 -> shared (row t, 16-byte chunks), row sum
    float l_tile = 0.f;
#pragma unroll 1
    for (int c0 = cb; c0 < cb + HK; c0 += 32) {
      float s[32];
      tc::tmem_ld32(lane_base + c0, s);
      tc::tmem_ld_wait();
      uint4 pk[4];
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
      for (int i = 0; i < 32; i += 2) {
        float x0, x1;
        if (plain) {
          x0 = fmaf(s[i], sc2, -m_use);
          x1 = fmaf(s[i + 1], sc2, -m_use);
        } else {
          x0 = fmaf(s[i], sc2, kterm[c0 + i]) - m_use;
          x1 = fmaf(s[i + 1], sc2, kterm[c0 + i + 1]) - m_use;
          if (diag && k0 + c0 + i > qi) x0 = -CUDART_INF_F;
          if (diag && k0 + c0 + i + 1 > qi) x1 = -CUDART_INF_F;
        }
        const float e0 = tc::ex2(x0), e1 = tc::ex2(x1);
        ph[i >> 1] = __floats2bfloat162_rn(e0, e1);
        l_tile += e0 + e1;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) *reinterpret_cast<uint4*>(smem + L::P + (c0 / 8 + j) * CS128 + t * 16) = pk[j];
    }
    l_run = l_run * corr + l_tile;
    m_run = m_new;
    MT_PROF(6);
    if (use_tma) {
      tc::mbar_wait(&bar_v, par_v);
      par_v ^= 1;
    } else {
      tc::cp_async_wait<1>();          // V(kt) landed (K(kt+1) may still be in flight)
    }
    if (has_x && xr >= 0 && xr < TK) {
      if (!use_tma) __syncthreads();   // cp.async: the zero fill of row xr belongs to other threads' groups
      if (tid >= DH / 8 && tid < 2 * (DH / 8)) *reinterpret_cast<uint4*>(smem + L::V + (tid - DH / 8) * CS128 + xr * 16) = xreg;
    }
    tc::fence_async_smem();
    tc::fence_before_sync();
    __syncthreads();
    tc::fence_after_sync();
    MT_PROF(7);
    if (tid == 0) {
      constexpr uint32_t idesc = tc::make_idesc(TQ, DH, 0, 1);
#pragma unroll
      for (int ks = 0; ks < TK / 16; ++ks)
        tc::mma_ss(tmem + 128, desc_kmajor(sp, ks), desc_mnmajor(sv, ks), idesc, ks > 0);
      tc::mma_commit(&bar_o);
    }
    tc::mbar_wait(&bar_o, phase);
    tc::fence_after_sync();
    MT_PROF(8);
    if (use_tma) {                                                                      // V is free: next tile
      if (more && tid == 0) {
        tma::mbar_expect_tx(&bar_v, L::OPER);
        tma::load_box(tc::smem_u32(smem + L::V), &maps.m[2], &bar_v, 0, k0 + TK, hc, n2, n1);
      }
    } else {
      if (more) stage_async<DH, NT>(smem + L::V, a.v, n, a.n2_count, h, k0 + TK, Skb);
      tc::cp_async_commit();
    }
    {
      float pv[HC];
#pragma unroll
      for (int c0 = 0; c0 < HC; c0 += 16) tc::tmem_ld16(lane_base + 128 + hf * HC + c0, pv + c0);
      tc::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < HC; ++i) o_acc[i] = fmaf(o_acc[i], corr, pv[i]);
    }
    phase ^= 1;
    MT_PROF(9);
  }
  tc::cp_async_wait<0>();
  xch[hf][t] = l_run;
  __syncthreads();
  l_run += xch[hf ^ 1][t];
  if (use_tma) {
    // O leaves through the free P buffer as swizzled sub-tiles and one or two TMA boxes (rows past Sq clipped)
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
#pragma unroll
    for (int c0 = 0; c0 < HC; c0 += 8) {
      uint4 pk;
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 8; i += 2) ph[i >> 1] = __floats2bfloat162_rn(o_acc[c0 + i] * inv, o_acc[c0 + i + 1] * inv);
      const int ch = (hf * HC + c0) / 8;
      const uint32_t off = ch < 8 ? tc::sw128_off(t, ch) : 128u * 128u + tc::sw64_off(t, ch - 8);
      *reinterpret_cast<uint4*>(smem + L::P + off) = pk;
    }
    if (qi < a.Sq && a.lse && hf == 0) a.lse[((int64_t)n * a.H + h) * a.Sq + qi] = m_run * LN2 + logf(l_run);
    tc::fence_async_smem();
    __syncthreads();
    if (tid == 0) {
      tma::store_box(&maps.o[0], tc::smem_u32(smem + L::P), 0, h, q0, n2, n1);
      if (DH > 64) tma::store_box(&maps.o[1], tc::smem_u32(smem + L::P) + 128u * 128u, 64, h, q0, n2, n1);
      tma::store_commit();
      tma::store_wait_read<0>();
    }
  } else if (qi < a.Sq) {
    const float inv = l_run > 0.f ? 1.0f / l_run : 0.f;
    __nv_bfloat16* op = view_ptr<__nv_bfloat16>(a.o, n, a.n2_count, qi, h) + hf * HC;
#pragma unroll
    for (int c0 = 0; c0 < HC; c0 += 8) {
      uint4 pk;
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
      for (int i = 0; i < 8; i += 2) ph[i >> 1] = __floats2bfloat162_rn(o_acc[c0 + i] * inv, o_acc[c0 + i + 1] * inv);
      *reinterpret_cast<uint4*>(op + c0) = pk;
    }
    if (a.lse && hf == 0) a.lse[((int64_t)n * a.H + h) * a.Sq + qi] = m_run * LN2 + logf(l_run);
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_slot, 256);
  MT_PROF(11);
}

// ------------------------------------------------------------------------------------ backward
template <int DH>
struct BwdSmem {
  static constexpr uint32_t OPER = (DH / 8) * CS128;
  static constexpr uint32_t PBUF = (TK / 8) * CS128;
  static constexpr uint32_t Q = 0, K = OPER, V = 2 * OPER, DO = 3 * OPER, P = 4 * OPER, DS = 4 * OPER + PBUF,
                            TOTAL = 4 * OPER + 2 * PBUF;
};

// un-rotate (transpose rotation) 16 consecutive features [c0, c0+16) of a gradient row in place
__device__ __forceinline__ void unrot16(float* g, const float* cs, const float* sn, int pos, int rot_dim, int c0) {
  if (c0 >= rot_dim) return;
  if (c0 + 16 <= rot_dim) {   // whole 16-feature block rotated: 128-bit table loads
    const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * rot_dim + c0);
    const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * rot_dim + c0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 c = __ldg(c4 + q), s = __ldg(s4 + q);
      const float a0 = g[4 * q], a1 = g[4 * q + 1], b0 = g[4 * q + 2], b1 = g[4 * q + 3];
      g[4 * q] = a0 * c.x + a1 * s.y;
      g[4 * q + 1] = a1 * c.y - a0 * s.x;
      g[4 * q + 2] = b0 * c.z + b1 * s.w;
      g[4 * q + 3] = b1 * c.w - b0 * s.z;
    }
    return;
  }
#pragma unroll
  for (int i = 0; i < 16; i += 2)
    if (c0 + i < rot_dim) unrot_pair(g[i], g[i + 1], cs, sn, pos, rot_dim, c0 + i);
}

template <int DH>
__global__ void __launch_bounds__(128, 1) attn_tc_bwd_kernel(AttnArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar1, bar2;
  __shared__ uint32_t tmem_slot;
  using L = BwdSmem<DH>;
  constexpr uint32_t T_S = 0, T_DP = 128, T_DV = 256, T_DK = 384, T_DQ = 0;
  const int t = threadIdx.x, warp = t >> 5;
  const int h = blockIdx.y, n = blockIdx.z;
  if (t == 0) {
    tc::mbar_init(&bar1, 1);
    tc::mbar_init(&bar2, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  stage_rows<DH>(smem + L::Q, a.q, n, a.n2_count, h, 0, a.Sq, a.qcos, a.qsin, a.rot_dim, a.pos_is_head);
  stage_rows<DH>(smem + L::K, a.k, n, a.n2_count, h, 0, a.Sk, a.kcos, a.ksin, a.rot_dim, a.pos_is_head);
  stage_rows<DH>(smem + L::V, a.v, n, a.n2_count, h, 0, a.Sk, nullptr, nullptr, 0, 0);
  stage_rows<DH>(smem + L::DO, a.o, n, a.n2_count, h, 0, a.Sq, nullptr, nullptr, 0, 0);
  // delta_t = <dO_t, O_t>, lse_t (thread t owns query row t)
  float delta = 0.f, lse2 = 0.f;
  if (t < a.Sq) {
    const __nv_bfloat16* op = view_ptr<__nv_bfloat16>(a.out_saved, n, a.n2_count, t, h);
    const __nv_bfloat16* gp = view_ptr<__nv_bfloat16>(a.o, n, a.n2_count, t, h);
#pragma unroll
    for (int c0 = 0; c0 < DH; c0 += 8) {
      const uint4 ov = *reinterpret_cast<const uint4*>(op + c0), gv = *reinterpret_cast<const uint4*>(gp + c0);
      const __nv_bfloat162* o2 = reinterpret_cast<const __nv_bfloat162*>(&ov);
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&gv);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 x = __bfloat1622float2(o2[i]), y = __bfloat1622float2(g2[i]);
        delta += x.x * y.x + x.y * y.y;
      }
    }
    const int64_t li = ((int64_t)n * a.H + h) * a.Sq + t;
    lse2 = a.lse[li] * LOG2E;
    a.delta[li] = delta;
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  const uint32_t sq = tc::smem_u32(smem + L::Q), sk = tc::smem_u32(smem + L::K), sv = tc::smem_u32(smem + L::V),
                 sdo = tc::smem_u32(smem + L::DO), sp = tc::smem_u32(smem + L::P), sds = tc::smem_u32(smem + L::DS);
  if (t == 0) {
    constexpr uint32_t idesc = tc::make_idesc(TQ, TK, 0, 0);
#pragma unroll
    for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_kmajor(sq, ks), desc_kmajor(sk, ks), idesc, ks > 0);
#pragma unroll
    for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_kmajor(sdo, ks), desc_kmajor(sv, ks), idesc, ks > 0);
    tc::mma_commit(&bar1);
  }
  tc::mbar_wait(&bar1, 0);
  tc::fence_after_sync();
  const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
  const float* mrow = a.key_mask ? a.key_mask + (int64_t)n * a.Sk : nullptr;
#pragma unroll 1
  for (int c0 = 0; c0 < TK; c0 += 16) {
    float s[16], dp[16];
    tc::tmem_ld16(lane_base + T_S + c0, s);
    tc::tmem_ld16(lane_base + T_DP + c0, dp);
    tc::tmem_ld_wait();
    uint4 pk[2], dk[2];
    __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(pk);
    __nv_bfloat162* dh = reinterpret_cast<__nv_bfloat162*>(dk);
#pragma unroll
    for (int i = 0; i < 16; i += 2) {
      float p[2], ds[2];
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int kj = c0 + i + j;
        const bool valid = t < a.Sq && kj < a.Sk && (!a.causal || kj <= t);
        p[j] = valid ? exp2f(masked_score(s[i + j], a.scale, mrow, kj, true) - lse2) : 0.f;
        ds[j] = p[j] * (dp[i + j] - delta) * a.scale;
      }
      ph[i >> 1] = __floats2bfloat162_rn(p[0], p[1]);
      dh[i >> 1] = __floats2bfloat162_rn(ds[0], ds[1]);
    }
    *reinterpret_cast<uint4*>(smem + L::P + (c0 / 8) * CS128 + t * 16) = pk[0];
    *reinterpret_cast<uint4*>(smem + L::P + (c0 / 8 + 1) * CS128 + t * 16) = pk[1];
    *reinterpret_cast<uint4*>(smem + L::DS + (c0 / 8) * CS128 + t * 16) = dk[0];
    *reinterpret_cast<uint4*>(smem + L::DS + (c0 / 8 + 1) * CS128 + t * 16) = dk[1];
  }
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  if (t == 0) {
    constexpr uint32_t id_t = tc::make_idesc(TK, DH, 1, 1);   // A^T (MN-major) x B (MN-major)
    constexpr uint32_t id_n = tc::make_idesc(TQ, DH, 0, 1);   // A (K-major)    x B (MN-major)
#pragma unroll
    for (int ks = 0; ks < TQ / 16; ++ks)      // dV = P^T dO   (K = queries)
      tc::mma_ss(tmem + T_DV, desc_mnmajor(sp, ks), desc_mnmajor(sdo, ks), id_t, ks > 0);
#pragma unroll
    for (int ks = 0; ks < TQ / 16; ++ks)      // dK = dS^T Q
      tc::mma_ss(tmem + T_DK, desc_mnmajor(sds, ks), desc_mnmajor(sq, ks), id_t, ks > 0);
#pragma unroll
    for (int ks = 0; ks < TK / 16; ++ks)      // dQ = dS K     (K = keys)
      tc::mma_ss(tmem + T_DQ, desc_kmajor(sds, ks), desc_mnmajor(sk, ks), id_n, ks > 0);
    tc::mma_commit(&bar2);
  }
  tc::mbar_wait(&bar2, 0);
  tc::fence_after_sync();
  // epilogue: row t of dV / dK is key t, row t of dQ is query t
  auto store_row = [&](uint32_t tcol, __nv_bfloat16* dst, const float* cs, const float* sn, int pos) {
#pragma unroll
    for (int c0 = 0; c0 < DH; c0 += 16) {
      float g[16];
      tc::tmem_ld16(lane_base + tcol + c0, g);
      tc::tmem_ld_wait();
      if (cs != nullptr) unrot16(g, cs, sn, pos, a.rot_dim, c0);
      uint4 pk[2];
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
      for (int i = 0; i < 16; i += 2) ph[i >> 1] = __floats2bfloat162_rn(g[i], g[i + 1]);
      if (dst != nullptr) {
        *reinterpret_cast<uint4*>(dst + c0) = pk[0];
        *reinterpret_cast<uint4*>(dst + c0 + 8) = pk[1];
      }
    }
  };
  const bool krow = t < a.Sk, qrow = t < a.Sq;
  store_row(T_DV, krow ? view_ptr<__nv_bfloat16>(a.dv, n, a.n2_count, t, h) : nullptr, nullptr, nullptr, 0);
  store_row(T_DK, krow ? view_ptr<__nv_bfloat16>(a.dk, n, a.n2_count, t, h) : nullptr, a.kcos, a.ksin,
            a.pos_is_head ? h : min(t, a.Sk - 1));
  store_row(T_DQ, qrow ? view_ptr<__nv_bfloat16>(a.dq, n, a.n2_count, t, h) : nullptr, a.qcos, a.qsin,
            a.pos_is_head ? h : min(t, a.Sq - 1));
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ---------------------------------------------------------------------- backward, multi-tile
// Shared memory: the fixed pair (MODE 0: K|V, MODE 1: Q|dO), the streamed pair double-buffered, P, dS.
template <int DH>
struct MtSmem {
  static constexpr uint32_t OPER = (DH / 8) * CS128;
  static constexpr uint32_t PBUF = (TK / 8) * CS128;
  static constexpr uint32_t FIX0 = 0, FIX1 = OPER, STR = 2 * OPER /* [2][2] */, P = 6 * OPER, DS = 6 * OPER + PBUF,
                            TOTAL = 6 * OPER + 2 * PBUF + 1024;   // + alignment slack
};

constexpr int MT_LOADERS = 128, MT_THREADS = 256 + MT_LOADERS;   // multi-tile backward: 8 compute warps + a loader warpgroup

// stage_async's walk for the 128 loader threads: lane pairs on one row's 32-byte sectors, two passes of 64 rows.
template <int DH>
__device__ __forceinline__ void stage_async_half(uint8_t* smem, const HeadView& v, int n, int n2c, int h, int s0, int S, int lt) {
  constexpr int CPR = DH / 8;
  const int par = lt & 1;
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    const int r = (lt >> 1) + pass * 64;
    const bool ok = s0 + r < S;
    const __nv_bfloat16* src = view_ptr<__nv_bfloat16>(v, n, n2c, ok ? s0 + r : 0, h) + par * 8;
    const uint32_t dst = tc::smem_u32(smem) + par * CS128 + r * 16;
    const uint32_t bytes = ok ? 16u : 0u;
#pragma unroll
    for (int it = 0; it < CPR / 2; ++it) tc::cp_async16(dst + it * 2 * CS128, src + it * 16, bytes);
  }
}

// Tensor maps of the multi-tile backward (chunk-layout boxes, tma_utils.cuh::make_chunk_map); use_tma = 0 keeps the
// cp.async loader warpgroup and the sector-wise stores (views TMA cannot describe).
// Gradients leave as the head-view boxes of tma_utils.cuh::make_head_map (64 columns SWIZZLE_128B [+ 32 columns
// SWIZZLE_64B]): 128- / 64-byte inner rows store several times faster than 16-byte chunk boxes would.
enum { MT_Q = 0, MT_K, MT_V, MT_DO, MT_NMAPS };
enum { MT_DQ = 0, MT_DK, MT_DV };
struct MtMaps {
  CUtensorMap m[MT_NMAPS];
  CUtensorMap o[3][2];
  int use_tma;
};

template <int DH, int MODE>
__global__ void __launch_bounds__(MT_THREADS, 1) attn_tc_bwd_mt_kernel(AttnArgs a, const __grid_constant__ MtMaps maps) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);   // swizzled staging boxes: 1024-byte atoms
  // bar1 / bar2: MMA chains done.  full_* / empty_*: operand buffers between the loader warpgroup and the compute warps.
  __shared__ uint64_t bar1, bar2, full_fix, empty_fix, full_str[2], empty_str[2];
  __shared__ uint32_t tmem_slot;
  __shared__ float kterm[TK];
  __shared__ float sdelta[TQ];
  using L = MtSmem<DH>;
  static_assert(DH / 8 <= 16, "delta walk: one 16-lane group per row");
  constexpr int NT = 256;
  constexpr uint32_t T_S = 0, T_DP = 128, T_A = 256, T_B = 384;   // MODE 0: A = dV, B = dK; MODE 1: A = dQ
  const int tid = threadIdx.x, warp = tid >> 5, t = tid & 127, half = tid >> 7;
  // Persistent: CTA c walks work items (tile, head, sequence) k * G + (c + k) % G, so that its consecutive items
  // have different tile indices (causal tiles differ in length) and neighbouring CTAs share one head's operands in
  // L2.  Warps 0-7 compute; warps 8-11 (one warpgroup, registers handed to the compute warps) only issue the 16-byte
  // cp.async copies - their issue blocks on the memory pipeline - and run ahead as far as the buffers allow, so the
  // next item's operands land while the finished item's gradients are staged out.
  int tile = 0, h = 0, n = 0, n1 = 0, n2 = 0, hc = 0;   // (n1, n2, hc): chunk-map coordinates of (n, h)
  MT_PROF_INIT();
  constexpr int PB = 12 + MODE * 10;   // phase-timer slots of this pass
  // shared extra key / value row (meant_attn_args.kx / vx): see the forward kernel
  const bool has_x = a.kx.p != nullptr;
  const int Skb = a.Sk - (has_x ? 1 : 0);
  if (tid == 0) {
    tc::mbar_init(&bar1, 1);
    tc::mbar_init(&bar2, 1);
    const uint32_t producers = maps.use_tma ? 1u : (uint32_t)MT_LOADERS;   // one expect_tx arrival, or one pre-counted
    tc::mbar_init(&full_fix, producers);                                     // cp.async arrival per loader thread
    tc::mbar_init(&full_str[0], producers);
    tc::mbar_init(&full_str[1], producers);
    tc::mbar_init(&empty_fix, 1);              // tcgen05.commit of the last MMA chain that read the buffer
    tc::mbar_init(&empty_str[0], 1);
    tc::mbar_init(&empty_str[1], 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();                             // all 384 threads: barriers initialised, TMEM address published
  tc::fence_after_sync();
  const int nq = (a.Sq + TQ - 1) / TQ, nk = (a.Sk + TK - 1) / TK;
  const int nt = MODE == 0 ? nk : nq;
  const int64_t n_items = (int64_t)nt * a.H * a.N;
  const float* mrow = nullptr;
  const float sc2 = a.scale * LOG2E;

  const bool pre = a.q_rot.p != nullptr;          // Q, K already rotated by meant_attn_prerotate
  const HeadView& qv = pre ? a.q_rot : a.q;
  const HeadView& kv = pre ? a.k_rot : a.k;
  const float* qcos = pre ? nullptr : a.qcos;
  const float* kcos = pre ? nullptr : a.kcos;
  // (measured and not kept, none faster: walks with 8 or 12 consecutive lanes on one row's chunks, an L2 bulk
  // prefetch of the next item, un-rotation tables fetched ahead of the next item's copies - the 16-byte cp.async
  // stream itself, ~22 B/clk per SM while it is being issued, is what bounds the item)
  // work item kth of this CTA -> (tile, h, n, j0, j1); false past the end.  Both roles walk the same sequence.
  int j0 = 0, j1 = 0;   // tiles of the other axis; causal problems skip the pairs above the diagonal
  auto decode_item = [&](int64_t kth) -> bool {
    const int64_t G = gridDim.x, idx = kth * G + (blockIdx.x + kth) % G;
    if (idx >= n_items) return false;
    tile = (int)(idx % nt);
    h = (int)((idx / nt) % a.H);
    n = (int)(idx / ((int64_t)nt * a.H));
    n1 = n / a.n2_count;
    n2 = n - n1 * a.n2_count;
    hc = h * (DH / 8);
    if (MODE == 0) {
      j0 = a.causal ? tile : 0;
      j1 = nq;
    } else {
      j0 = 0;
      j1 = a.causal ? min(nk, tile + 1) : nk;
    }
    return true;
  };

  if (tid >= NT) {   // ------------------------------------------------------------------ loader warpgroup
    tc::reg_dec<40>();
    const int lt = tid - NT;
    if (maps.use_tma) {   // one thread drives TMA: one box per operand tile, straight into the chunk layout
      if (lt != 0) return;
      const CUtensorMap* mq = &maps.m[MT_Q];
      const CUtensorMap* mk = &maps.m[MT_K];
      const CUtensorMap* mv = &maps.m[MT_V];
      const CUtensorMap* mo = &maps.m[MT_DO];
      tma::prefetch_map(mq);
      tma::prefetch_map(mk);
      tma::prefetch_map(mv);
      tma::prefetch_map(mo);
      const CUtensorMap* fa = MODE == 0 ? mk : mq;   // fixed pair
      const CUtensorMap* fb = MODE == 0 ? mv : mo;
      const CUtensorMap* sa = MODE == 0 ? mq : mk;   // streamed pair
      const CUtensorMap* sb = MODE == 0 ? mo : mv;
      uint32_t pe_fix = 1, pe_str0 = 1, pe_str1 = 1;
#pragma unroll 1
      for (int64_t kth = 0; decode_item(kth); ++kth) {
        if (j0 >= j1) continue;
        tc::mbar_wait(&empty_fix, pe_fix);
        pe_fix ^= 1;
        tma::mbar_expect_tx(&full_fix, 2 * L::OPER);
        tma::load_box(tc::smem_u32(smem + L::FIX0), fa, &full_fix, 0, tile * 128, hc, n2, n1);
        tma::load_box(tc::smem_u32(smem + L::FIX1), fb, &full_fix, 0, tile * 128, hc, n2, n1);
#pragma unroll 1
        for (int j = j0; j < j1; ++j) {
          const int b = (j - j0) & 1;
          if (b == 0) {
            tc::mbar_wait(&empty_str[0], pe_str0);
            pe_str0 ^= 1;
          } else {
            tc::mbar_wait(&empty_str[1], pe_str1);
            pe_str1 ^= 1;
          }
          const uint32_t x = tc::smem_u32(smem + L::STR + (uint32_t)b * 2 * L::OPER);
          tma::mbar_expect_tx(&full_str[b], 2 * L::OPER);
          tma::load_box(x, sa, &full_str[b], 0, j * 128, hc, n2, n1);
          tma::load_box(x + L::OPER, sb, &full_str[b], 0, j * 128, hc, n2, n1);
        }
      }
      return;
    }
    auto stage = [&](uint8_t* dst, const HeadView& v, int s0, int S) {
      stage_async_half<DH>(dst, v, n, a.n2_count, h, s0, S, lt);
    };
    uint32_t pe_fix = 1, pe_str0 = 1, pe_str1 = 1;   // a fresh barrier's "previous phase" counts as complete
#pragma unroll 1
    for (int64_t kth = 0; decode_item(kth); ++kth) {
      if (j0 >= j1) continue;                  // empty item: nothing is loaded, nothing is waited for
      tc::mbar_wait(&empty_fix, pe_fix);
      pe_fix ^= 1;
      if (MODE == 0) {
        stage(smem + L::FIX0, kv, tile * TK, Skb);
        stage(smem + L::FIX1, a.v, tile * TK, Skb);
      } else {
        stage(smem + L::FIX0, qv, tile * TQ, a.Sq);
        stage(smem + L::FIX1, a.o, tile * TQ, a.Sq);
      }
      tc::cp_async_arrive(&full_fix);
#pragma unroll 1
      for (int j = j0; j < j1; ++j) {
        const int b = (j - j0) & 1;
        if (b == 0) {
          tc::mbar_wait(&empty_str[0], pe_str0);
          pe_str0 ^= 1;
        } else {
          tc::mbar_wait(&empty_str[1], pe_str1);
          pe_str1 ^= 1;
        }
        uint8_t* x = smem + L::STR + (uint32_t)b * 2 * L::OPER;   // MODE 0: Q(j) | dO(j),  MODE 1: K(j) | V(j)
        if (MODE == 0) {
          stage(x, qv, j * TQ, a.Sq);
          stage(x + L::OPER, a.o, j * TQ, a.Sq);
        } else {
          stage(x, kv, j * TK, Skb);
          stage(x + L::OPER, a.v, j * TK, Skb);
        }
        tc::cp_async_arrive(&full_str[b]);
      }
    }
    tc::cp_async_commit();
    tc::cp_async_wait<0>();
    return;
  }
  // ------------------------------------------------------------------------------------- compute warps
  tc::reg_inc<232>();
  // delta = <dO, O>: computed once per query tile by the dQ pass (which runs first) and kept in a.delta for the
  // dK|dV pass.  Coalesced: a 16-lane group reads one row of O and dO (DH*2 contiguous bytes each), two rows per
  // warp instruction, all loads of the warp's 16 rows in flight before the first use.
  // Rows past Sq keep lse = delta = 0: their Q and dO rows are zero-filled, so P is finite and dS, P^T dO vanish.
  float delta = 0.f, lse_raw = 0.f;
  auto row_stats = [&](int qi) {               // per-thread statistics of query row qi (coalesced fp32 reads);
    delta = 0.f;                               // nothing consumes the loads here: the latency hides behind what follows
    lse_raw = 0.f;
    if (qi >= a.Sq) return;
    const int64_t li = ((int64_t)n * a.H + h) * a.Sq + qi;
    lse_raw = __ldg(a.lse + li);
    if (MODE == 0) delta = a.delta[li];
  };
  uint4 ov[MODE == 1 ? 8 : 1], gv[MODE == 1 ? 8 : 1];
  auto tile_delta_issue = [&](int q0) {
    const int lane = tid & 31, sub = lane >> 4, c = lane & 15;
    const __nv_bfloat16* ob = view_ptr<__nv_bfloat16>(a.out_saved, n, a.n2_count, 0, h) + c * 8;   // row 0 of this head
    const __nv_bfloat16* gb = view_ptr<__nv_bfloat16>(a.o, n, a.n2_count, 0, h) + c * 8;
    const int64_t oss = a.out_saved.ss, gss = a.o.ss;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const int qi = q0 + warp * 16 + it * 2 + sub;
      ov[it] = make_uint4(0u, 0u, 0u, 0u);
      gv[it] = ov[it];
      if (qi < a.Sq && c < DH / 8) {
        ov[it] = *reinterpret_cast<const uint4*>(ob + qi * oss);
        gv[it] = *reinterpret_cast<const uint4*>(gb + qi * gss);
      }
    }
  };
  auto tile_delta_reduce = [&](int q0) {       // same item as the matching tile_delta_issue (n, h unchanged)
    const int lane = tid & 31, sub = lane >> 4, c = lane & 15;
    float d[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
      const __nv_bfloat162* o2 = reinterpret_cast<const __nv_bfloat162*>(&ov[it]);
      const __nv_bfloat162* g2 = reinterpret_cast<const __nv_bfloat162*>(&gv[it]);
      float acc = 0.f;
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 x = __bfloat1622float2(o2[i]), y = __bfloat1622float2(g2[i]);
        acc += x.x * y.x + x.y * y.y;
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      d[it] = acc;
    }
    if (c == 0) {
#pragma unroll
      for (int it = 0; it < 8; ++it) {
        const int r = warp * 16 + it * 2 + sub, qi = q0 + r;
        sdelta[r] = d[it];
        if (qi < a.Sq) a.delta[((int64_t)n * a.H + h) * a.Sq + qi] = d[it];
      }
    }
  };

  uint4 xreg = make_uint4(0u, 0u, 0u, 0u);      // threads [0, CPR): K's chunks of the extra row, [CPR, 2 CPR): V's
  auto begin_item = [&](int64_t kth) -> bool {   // the compute side of an item: its statistics
    if (!decode_item(kth)) return false;
    mrow = a.key_mask ? a.key_mask + (int64_t)n * a.Sk : nullptr;
    if (has_x && tid < 2 * (DH / 8)) {
      const HeadView& xv = tid < DH / 8 ? a.kx : a.vx;
      xreg = *reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(xv.p) + (int64_t)n1 * xv.sn1 + (int64_t)h * xv.sh +
                                             (tid % (DH / 8)) * 8);
    }
    if (MODE == 1) {
      row_stats(tile * TQ + t);
      tile_delta_issue(tile * TQ);            // reduced after the previous item's gradients are staged out
    }
    if (MODE == 0 && tid < TK) kterm[tid] = key_term(mrow, tile * TK + tid, a.Sk);
    return true;
  };
  begin_item(0);
  if (MODE == 1) tile_delta_reduce(tile * TQ);   // visible to the softmax after the loop's first barrier

  uint32_t phase = 0, pf_fix = 0, pf_str0 = 0, pf_str1 = 0;
  const uint32_t tmem = tmem_slot;
  bool first = true;
  int64_t kth = 0;
#pragma unroll 1
  for (;;) {
#pragma unroll 1
  for (int j = j0; j < j1; ++j) {
    const int b = (j - j0) & 1;
    const int q0 = (MODE == 0 ? j : tile) * TQ, k0 = (MODE == 0 ? tile : j) * TK;
    MT_PROF(PB + 0);
    if (MODE == 0) row_stats(q0 + t);           // latency hidden behind the wait
    if (first) {
      tc::mbar_wait(&full_fix, pf_fix);         // the fixed pair landed
      pf_fix ^= 1;
    }
    if (b == 0) {                               // streamed pair j landed
      tc::mbar_wait(&full_str[0], pf_str0);
      pf_str0 ^= 1;
    } else {
      tc::mbar_wait(&full_str[1], pf_str1);
      pf_str1 ^= 1;
    }
    MT_PROF(PB + 1);
    uint8_t* xs = smem + L::STR + (uint32_t)b * 2 * L::OPER;
    if (MODE == 0) {
      rotate_staged<DH, NT>(xs, h, q0, a.Sq, qcos, a.qsin, a.rot_dim, a.pos_is_head);
      if (first) rotate_staged<DH, NT>(smem + L::FIX0, h, k0, a.Sk, kcos, a.ksin, a.rot_dim, a.pos_is_head);
    } else {
      rotate_staged<DH, NT>(xs, h, k0, a.Sk, kcos, a.ksin, a.rot_dim, a.pos_is_head);
      if (first) rotate_staged<DH, NT>(smem + L::FIX0, h, q0, a.Sq, qcos, a.qsin, a.rot_dim, a.pos_is_head);
      if (tid < TK) kterm[tid] = key_term(mrow, k0 + tid, a.Sk);
    }
    {   // the extra key / value row goes into the K | V pair (MODE 0: fixed, once per item; MODE 1: streamed tile j)
      const int xr = Skb - k0;
      if (has_x && xr >= 0 && xr < TK && (MODE == 1 || first)) {
        tc::bar_sync(1, NT);                     // row xr was zero-filled (and "rotated") by other threads
        uint8_t* kd = MODE == 0 ? smem + L::FIX0 : xs;
        if (tid < DH / 8) *reinterpret_cast<uint4*>(kd + tid * CS128 + xr * 16) = xreg;
        else if (tid < 2 * (DH / 8)) *reinterpret_cast<uint4*>(kd + L::OPER + (tid - DH / 8) * CS128 + xr * 16) = xreg;
      }
    }
    const uint32_t sq = tc::smem_u32(MODE == 0 ? xs : smem + L::FIX0), sdo = sq + L::OPER;
    const uint32_t sk = tc::smem_u32(MODE == 0 ? smem + L::FIX0 : xs), sv = sk + L::OPER;
    const uint32_t sp = tc::smem_u32(smem + L::P), sds = tc::smem_u32(smem + L::DS);
    MT_PROF(PB + 2);
    tc::fence_async_smem();
    tc::fence_before_sync();
    tc::bar_sync(1, NT);
    tc::fence_after_sync();
    MT_PROF(PB + 3);
    if (MODE == 1 && first) delta = sdelta[t];
    if (tid == 0) {
      constexpr uint32_t idesc = tc::make_idesc(TQ, TK, 0, 0);
#pragma unroll
      for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_S, desc_kmajor(sq, ks), desc_kmajor(sk, ks), idesc, ks > 0);
#pragma unroll
      for (int ks = 0; ks < DH / 16; ++ks) tc::mma_ss(tmem + T_DP, desc_kmajor(sdo, ks), desc_kmajor(sv, ks), idesc, ks > 0);
      tc::mma_commit(&bar1);
    }
    tc::mbar_wait(&bar1, phase);
    tc::fence_after_sync();
    if (first && maps.use_tma) {   // P / dS are about to be rewritten: the previous item's gradient boxes must have
      if (tid == 0) tma::store_wait_read<0>();   // been read out (the read-out ran under this item's S, dP MMAs)
      tc::bar_sync(1, NT);
    }
    MT_PROF(PB + 4);
    const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
    const int qi = q0 + t;
    const bool diag = a.causal && k0 + TK - 1 > q0;
    const bool plain = !diag && mrow == nullptr && k0 + TK <= a.Sk;   // no additive key term, nothing above the diagonal
    const float lse2 = lse_raw * LOG2E;
    // This thread's 64 key columns: all four TMEM loads in flight, then 64 independent element chains.
    {
      const int cb = half * (TK / 2);
      float s[2][32], dp[2][32];
      tc::tmem_ld32(lane_base + T_S + cb, s[0]);
      tc::tmem_ld32(lane_base + T_DP + cb, dp[0]);
      tc::tmem_ld32(lane_base + T_S + cb + 32, s[1]);
      tc::tmem_ld32(lane_base + T_DP + cb + 32, dp[1]);
      tc::tmem_ld_wait();
      const float dsc = a.scale;
      auto emit = [&](auto variant) {            // 0: bare scores, 1: + key term (padding, keys past Sk), 2: + causal diagonal
        constexpr int VAR = decltype(variant)::value;
#pragma unroll
        for (int blk = 0; blk < 2; ++blk) {
#pragma unroll
          for (int i0 = 0; i0 < 32; i0 += 8) {
            const int c0 = cb + blk * 32 + i0;
            uint4 pk, dk;
            __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
            __nv_bfloat162* dh = reinterpret_cast<__nv_bfloat162*>(&dk);
#pragma unroll
            for (int i = 0; i < 8; i += 2) {
              float p[2], ds[2];
#pragma unroll
              for (int jj = 0; jj < 2; ++jj) {
                const int c = c0 + i + jj;
                float x;
                if (VAR == 0) {
                  x = fmaf(s[blk][i0 + i + jj], sc2, -lse2);
                } else {
                  x = fmaf(s[blk][i0 + i + jj], sc2, kterm[c]) - lse2;           // -inf past the last key
                  if (VAR == 2 && k0 + c > qi) x = -CUDART_INF_F;
                }
                p[jj] = tc::ex2(x);
                ds[jj] = p[jj] * (dp[blk][i0 + i + jj] - delta) * dsc;
              }
              ph[i >> 1] = __floats2bfloat162_rn(p[0], p[1]);
              dh[i >> 1] = __floats2bfloat162_rn(ds[0], ds[1]);
            }
            if (MODE == 0) *reinterpret_cast<uint4*>(smem + L::P + (c0 / 8) * CS128 + t * 16) = pk;
            *reinterpret_cast<uint4*>(smem + L::DS + (c0 / 8) * CS128 + t * 16) = dk;
          }
        }
      };
      if (plain) emit(std::integral_constant<int, 0>{});
      else if (!diag) emit(std::integral_constant<int, 1>{});
      else emit(std::integral_constant<int, 2>{});
    }
    MT_PROF(PB + 5);
    tc::fence_async_smem();
    tc::fence_before_sync();
    tc::bar_sync(1, NT);
    tc::fence_after_sync();
    MT_PROF(PB + 6);
    if (tid == 0) {
      constexpr uint32_t id_t = tc::make_idesc(TK, DH, 1, 1);   // A^T (MN-major) x B (MN-major)
      constexpr uint32_t id_n = tc::make_idesc(TQ, DH, 0, 1);   // A (K-major)    x B (MN-major)
      const uint32_t acc = first ? 0u : 1u;
      if (MODE == 0) {
#pragma unroll
        for (int ks = 0; ks < TQ / 16; ++ks)      // dV += P^T dO   (K = queries)
          tc::mma_ss(tmem + T_A, desc_mnmajor(sp, ks), desc_mnmajor(sdo, ks), id_t, acc | (ks > 0));
#pragma unroll
        for (int ks = 0; ks < TQ / 16; ++ks)      // dK += dS^T Q
          tc::mma_ss(tmem + T_B, desc_mnmajor(sds, ks), desc_mnmajor(sq, ks), id_t, acc | (ks > 0));
      } else {
#pragma unroll
        for (int ks = 0; ks < TK / 16; ++ks)      // dQ += dS K     (K = keys)
          tc::mma_ss(tmem + T_A, desc_kmajor(sds, ks), desc_mnmajor(sk, ks), id_n, acc | (ks > 0));
      }
      tc::mma_commit(&bar2);
      tc::mma_commit(&empty_str[b]);            // the loader may refill this streamed pair ...
      if (j + 1 == j1) tc::mma_commit(&empty_fix);   // ... and, after the item's last chain, the fixed pair
    }
    tc::mbar_wait(&bar2, phase);   // P and dS may be rewritten
    tc::fence_after_sync();
    MT_PROF(PB + 7);
    first = false;
    phase ^= 1;
  }
  const int etile = tile, eh = h, en = n;   // the finished item: its gradients are staged out below
  const bool zero = first;
  // (zero: no tile pair at all - causal with Sk > Sq - the gradients of this tile are zero)
  const bool more = begin_item(++kth);      // every buffer the loads target was released by the last bar2
  MT_PROF(PB + 8);
  const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
  // Accumulators leave through the (now free) P / dS buffers: thread (row t, column half) un-rotates and packs its
  // row into the chunk layout, then lane pairs write whole 32-byte sectors of the head view (stage_async's walk).
  auto stage_out = [&](uint32_t tcol, uint8_t* buf, const float* cs, const float* sn, int pos, float* xdst) {
    constexpr int NB = DH / 32;                // 16-column blocks per thread
    float g[NB][16];
    if (!zero) {
#pragma unroll
      for (int bi = 0; bi < NB; ++bi) tc::tmem_ld16(lane_base + tcol + half * (DH / 2) + bi * 16, g[bi]);
      tc::tmem_ld_wait();
    } else {
#pragma unroll
      for (int bi = 0; bi < NB; ++bi)
#pragma unroll
        for (int i = 0; i < 16; ++i) g[bi][i] = 0.f;
    }
    if (xdst != nullptr) {   // this row is the shared extra key: its gradient sums over the sequences that read it
#pragma unroll
      for (int bi = 0; bi < NB; ++bi)
#pragma unroll
        for (int i = 0; i < 16; ++i) atomicAdd(xdst + half * (DH / 2) + bi * 16 + i, g[bi][i]);
    }
    if (cs != nullptr) {
#pragma unroll
      for (int bi = 0; bi < NB; ++bi) unrot16(g[bi], cs, sn, pos, a.rot_dim, half * (DH / 2) + bi * 16);
    }
#pragma unroll
    for (int bi = 0; bi < NB; ++bi) {
      const int c0 = half * (DH / 2) + bi * 16;
      uint4 pk[2];
      __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(pk);
#pragma unroll
      for (int i = 0; i < 16; i += 2) ph[i >> 1] = __floats2bfloat162_rn(g[bi][i], g[bi][i + 1]);
      if (maps.use_tma) {   // columns [0, 64): 128-byte rows, SWIZZLE_128B; [64, 96): 64-byte rows, SWIZZLE_64B
#pragma unroll
        for (int q = 0; q < 2; ++q) {
          const int ch = c0 / 8 + q;
          const uint32_t off = ch < 8 ? tc::sw128_off(t, ch) : 128u * 128u + tc::sw64_off(t, ch - 8);
          *reinterpret_cast<uint4*>(buf + off) = pk[q];
        }
      } else {
        *reinterpret_cast<uint4*>(buf + (c0 / 8) * CS128 + t * 16) = pk[0];
        *reinterpret_cast<uint4*>(buf + (c0 / 8 + 1) * CS128 + t * 16) = pk[1];
      }
    }
  };
  auto store_out = [&](int which, const uint8_t* buf) {   // one thread: the staged tile as one or two TMA boxes
    const uint32_t src = tc::smem_u32(buf);
    const int en1 = en / a.n2_count, en2 = en - en1 * a.n2_count;
    tma::store_box(&maps.o[which][0], src, 0, eh, etile * 128, en2, en1);
    if (DH > 64) tma::store_box(&maps.o[which][1], src + 128u * 128u, 64, eh, etile * 128, en2, en1);
  };
  auto flush_out = [&](const uint8_t* buf, const HeadView& dst, int S) {
    const int r = tid >> 1, par = tid & 1, srow = etile * 128 + r;
    if (srow >= S) return;
    __nv_bfloat16* gp = view_ptr<__nv_bfloat16>(dst, en, a.n2_count, srow, eh) + par * 8;
    const uint8_t* sp = buf + par * CS128 + r * 16;
#pragma unroll
    for (int it = 0; it < DH / 16; ++it)
      *reinterpret_cast<uint4*>(gp + it * 16) = *reinterpret_cast<const uint4*>(sp + it * 2 * CS128);
  };
  const int row = etile * 128 + t;
  if (MODE == 0) {
    const bool xrow = has_x && row == Skb && !zero;
    const int64_t xoff = ((int64_t)(en / a.n2_count) * a.H + eh) * DH;
    stage_out(T_A, smem + L::P, nullptr, nullptr, 0, xrow ? a.dvx + xoff : nullptr);
    stage_out(T_B, smem + L::DS, xrow ? nullptr : a.kcos, a.ksin, a.pos_is_head ? eh : min(row, a.Sk - 1), xrow ? a.dkx + xoff : nullptr);
    tc::fence_async_smem();
    tc::bar_sync(1, NT);
    if (maps.use_tma) {
      if (tid == 0) {
        store_out(MT_DV, smem + L::P);
        store_out(MT_DK, smem + L::DS);
        tma::store_commit();
      }
    } else {
      flush_out(smem + L::P, a.dv, Skb);
      flush_out(smem + L::DS, a.dk, Skb);
    }
  } else {
    stage_out(T_A, smem + L::DS, a.qcos, a.qsin, a.pos_is_head ? eh : min(row, a.Sq - 1), nullptr);
    tc::fence_async_smem();
    tc::bar_sync(1, NT);
    if (maps.use_tma) {
      if (tid == 0) {
        store_out(MT_DQ, smem + L::DS);
        tma::store_commit();
      }
    } else {
      flush_out(smem + L::DS, a.dq, a.Sq);
    }
    if (more) tile_delta_reduce(tile * TQ);
  }
  MT_PROF(PB + 9);
  if (!more) break;
  first = true;
  }   // items
  if (tid == 0 && maps.use_tma) tma::store_wait_all();
  tc::fence_before_sync();
  tc::bar_sync(1, NT);
  if (warp == 0) tc::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------ pre-rotation
// q -> q_rot (blockIdx.y = 0) and k -> k_rot (blockIdx.y = 1) in 16-byte chunks, chunks of a head's row contiguous
// across the warp.  HBM-bound: reads and writes Q and K once.
__global__ void __launch_bounds__(256) attn_prerotate_kernel(AttnArgs a) {
  const bool isk = blockIdx.y != 0;
  const HeadView& src = isk ? a.k : a.q;
  const HeadView& dst = isk ? a.k_rot : a.q_rot;
  const float* __restrict__ cs = isk ? a.kcos : a.qcos;
  const float* __restrict__ sn = isk ? a.ksin : a.qsin;
  const int S = isk ? a.Sk : a.Sq;
  // in place (q_rot aliases q): only the rotated chunks of a head row are touched -- half the traffic at rot_dim = Dh / 2
  const bool inplace = src.p == dst.p;
  const int cpr = inplace ? (a.rot_dim >> 3) : (a.Dh >> 3);
  if (!a.pos_is_head) {
    // Token-position tables do not depend on the head: a thread owns (row, chunk), fetches its 64 bytes of table
    // values ONCE and walks the heads (table traffic through L1 was 4x the payload with a thread per head chunk).
    const int64_t total = (int64_t)a.N * S * cpr;
    for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
      const int ch = (int)(idx % cpr);
      const int64_t row = idx / cpr;
      const int sr = (int)(row % S), n = (int)(row / S);
      const __nv_bfloat16* sp = view_ptr<__nv_bfloat16>(src, n, a.n2_count, sr, 0) + ch * 8;
      __nv_bfloat16* dp = view_ptr<__nv_bfloat16>(dst, n, a.n2_count, sr, 0) + ch * 8;
      const bool rot = ch * 8 < a.rot_dim;
      float4 ca, cb, sa, sb;
      if (rot) {
        const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)sr * a.rot_dim + ch * 8);
        const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)sr * a.rot_dim + ch * 8);
        ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
      }
      if (a.H == 8) {          // the models' head count: all eight 16-byte loads in flight before the first store
        uint4 raw[8];
#pragma unroll
        for (int h = 0; h < 8; ++h) raw[h] = tc::ldg_stream16(sp + (int64_t)h * src.sh);
#pragma unroll
        for (int h = 0; h < 8; ++h) {
          if (rot) {
            __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw[h]);
            const float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
            const float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
            h2[0] = __floats2bfloat162_rn(p0.x * ca.x - p0.y * sa.x, p0.y * ca.y + p0.x * sa.y);
            h2[1] = __floats2bfloat162_rn(p1.x * ca.z - p1.y * sa.z, p1.y * ca.w + p1.x * sa.w);
            h2[2] = __floats2bfloat162_rn(p2.x * cb.x - p2.y * sb.x, p2.y * cb.y + p2.x * sb.y);
            h2[3] = __floats2bfloat162_rn(p3.x * cb.z - p3.y * sb.z, p3.y * cb.w + p3.x * sb.w);
          }
          *reinterpret_cast<uint4*>(dp + (int64_t)h * dst.sh) = raw[h];
        }
        continue;
      }
#pragma unroll 4
      for (int h = 0; h < a.H; ++h) {
        uint4 raw = *reinterpret_cast<const uint4*>(sp + (int64_t)h * src.sh);
        if (rot) {
          __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
          const float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
          const float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
          h2[0] = __floats2bfloat162_rn(p0.x * ca.x - p0.y * sa.x, p0.y * ca.y + p0.x * sa.y);
          h2[1] = __floats2bfloat162_rn(p1.x * ca.z - p1.y * sa.z, p1.y * ca.w + p1.x * sa.w);
          h2[2] = __floats2bfloat162_rn(p2.x * cb.x - p2.y * sb.x, p2.y * cb.y + p2.x * sb.y);
          h2[3] = __floats2bfloat162_rn(p3.x * cb.z - p3.y * sb.z, p3.y * cb.w + p3.x * sb.w);
        }
        *reinterpret_cast<uint4*>(dp + (int64_t)h * dst.sh) = raw;
      }
    }
    return;
  }
  const int64_t total = (int64_t)a.N * S * a.H * cpr;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * blockDim.x) {
    const int ch = (int)(idx % cpr);
    int64_t rest = idx / cpr;
    const int h = (int)(rest % a.H);
    rest /= a.H;
    const int sr = (int)(rest % S), n = (int)(rest / S);
    uint4 raw = *reinterpret_cast<const uint4*>(view_ptr<__nv_bfloat16>(src, n, a.n2_count, sr, h) + ch * 8);
    if (ch * 8 < a.rot_dim) {
      const int pos = a.pos_is_head ? h : sr;
      const float4* c4 = reinterpret_cast<const float4*>(cs + (int64_t)pos * a.rot_dim + ch * 8);
      const float4* s4 = reinterpret_cast<const float4*>(sn + (int64_t)pos * a.rot_dim + ch * 8);
      const float4 ca = __ldg(c4), cb = __ldg(c4 + 1), sa = __ldg(s4), sb = __ldg(s4 + 1);
      __nv_bfloat162* h2 = reinterpret_cast<__nv_bfloat162*>(&raw);
      const float2 p0 = __bfloat1622float2(h2[0]), p1 = __bfloat1622float2(h2[1]);
      const float2 p2 = __bfloat1622float2(h2[2]), p3 = __bfloat1622float2(h2[3]);
      h2[0] = __floats2bfloat162_rn(p0.x * ca.x - p0.y * sa.x, p0.y * ca.y + p0.x * sa.y);
      h2[1] = __floats2bfloat162_rn(p1.x * ca.z - p1.y * sa.z, p1.y * ca.w + p1.x * sa.w);
      h2[2] = __floats2bfloat162_rn(p2.x * cb.x - p2.y * sb.x, p2.y * cb.y + p2.x * sb.y);
      h2[3] = __floats2bfloat162_rn(p3.x * cb.z - p3.y * sb.z, p3.y * cb.w + p3.x * sb.w);
    }
    *reinterpret_cast<uint4*>(view_ptr<__nv_bfloat16>(dst, n, a.n2_count, sr, h) + ch * 8) = raw;
  }
}

// ------------------------------------------------------------------------------------ host side
// attn_tc2.cu: persistent warp-specialised kernels for single-tile problems (Sq, Sk <= 128)
int attn_tc2_supported(const AttnArgs& a, int backward);
int attn_tc2_fwd(const AttnArgs& a, cudaStream_t st);
int attn_tc2_bwd(const AttnArgs& a, cudaStream_t st);
int attn_tc2_ms_supported(const AttnArgs& a);
int attn_tc2_fwd_ms(const AttnArgs& a, cudaStream_t st);
int attn_tc2_wk_supported(const AttnArgs& a);
int attn_tc2_fwd_wk(const AttnArgs& a, cudaStream_t st);
int attn_tc2_ms_bwd_supported(const AttnArgs& a);
int attn_tc2_bwd_ms(const AttnArgs& a, int parts, cudaStream_t st);

static bool view_ok(const HeadView& v) {
  return v.p != nullptr && (reinterpret_cast<uintptr_t>(v.p) % 16 == 0) && v.sn1 % 8 == 0 && v.sn2 % 8 == 0 &&
         v.ss % 8 == 0 && v.sh % 8 == 0;
}

static bool xview_ok(const HeadView& v) {
  return v.p != nullptr && (reinterpret_cast<uintptr_t>(v.p) % 16 == 0) && v.sn1 % 8 == 0 && v.sh % 8 == 0;
}

int attn_tc_supported(const AttnArgs& a, int dtype, int backward) {
  if (dtype != MEANT_DT_BF16) return 0;
  if (a.kx.p != nullptr) {   // shared extra key: plain non-causal attention only, Sk counts it
    if (!xview_ok(a.kx) || !xview_ok(a.vx) || a.causal || a.key_mask || a.pos_is_head || a.q_rot.p || a.Sk < 2) return 0;
    if (backward && (!a.dkx || !a.dvx)) return 0;
  }
  if (a.Dh != 64 && a.Dh != 96) return 0;
  if (a.N <= 0 || a.H <= 0 || a.Sq <= 0 || a.Sk <= 0 || a.n2_count <= 0 || a.N > 65535 || a.H > 65535) return 0;
  if (a.rot_dim % 8 != 0 || a.rot_dim > a.Dh) return 0;
  if (a.qcos && (!a.qsin || !a.kcos || !a.ksin)) return 0;
  if (!view_ok(a.q) || !view_ok(a.k) || !view_ok(a.v) || !view_ok(a.o)) return 0;
  if (a.q_rot.p != nullptr && (!view_ok(a.q_rot) || !view_ok(a.k_rot))) return 0;
  if (backward) {
    if (!view_ok(a.dq) || !view_ok(a.dk) || !view_ok(a.dv) || !view_ok(a.out_saved) || !a.lse || !a.delta) return 0;
  }
  return 1;
}

template <typename Kern>
static int launch_tc(Kern kern, size_t smem, dim3 grid, const AttnArgs& a, cudaStream_t st, int threads = 128) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  kern<<<grid, threads, smem, st>>>(a);
  MEANT_LAUNCH_CHECK();
  return 0;
}

// MEANT_MT_NO_TMA=1 keeps the multi-tile kernels on their cp.async loaders (the GPU suite runs both); read once
static bool mt_no_tma() {
  static const bool v = getenv("MEANT_MT_NO_TMA") != nullptr;
  return v;
}

template <int DH>
static int launch_bwd_mt(const AttnArgs& a, cudaStream_t st, int parts = 3) {
  // persistent kernels: one CTA per SM, work items (tile, head, sequence) strided inside the kernel
  const int64_t ik = (int64_t)((a.Sk + TK - 1) / TK) * a.H * a.N, iq = (int64_t)((a.Sq + TQ - 1) / TQ) * a.H * a.N;
  const dim3 gk((unsigned)(ik < 148 ? ik : 148)), gq((unsigned)(iq < 148 ? iq : 148));
  // the dQ pass goes first: it also writes delta = rowsum(dO * O), which the dK|dV pass reads
  MtMaps maps;
  memset(&maps, 0, sizeof(maps));
  {
    const bool pre = a.q_rot.p != nullptr;
    const int skb = a.Sk - (a.kx.p != nullptr ? 1 : 0);   // rows k / v / dk / dv really hold (meant_attn_args.kx)
    const HeadView* vs[MT_NMAPS] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v, &a.o};
    const int seq[MT_NMAPS] = {a.Sq, skb, skb, a.Sq};
    int ok = !mt_no_tma();
    for (int i = 0; i < MT_NMAPS && ok; ++i) ok = tma::make_chunk_map(&maps.m[i], *vs[i], a.N, a.n2_count, a.H, seq[i], DH) == 0;
    const HeadView* os[3] = {&a.dq, &a.dk, &a.dv};
    const int oseq[3] = {a.Sq, skb, skb};
    for (int i = 0; i < 3 && ok; ++i) {
      ok = tma::make_head_map(&maps.o[i][0], *os[i], a.N, a.n2_count, a.H, oseq[i], DH, 64) == 0;
      if (ok && DH > 64) ok = tma::make_head_map(&maps.o[i][1], *os[i], a.N, a.n2_count, a.H, oseq[i], DH, 32) == 0;
    }
    maps.use_tma = ok;
  }
  auto launch = [&](auto kern, dim3 grid) -> int {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MtSmem<DH>::TOTAL);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, MT_THREADS, MtSmem<DH>::TOTAL, st>>>(a, maps);
    MEANT_LAUNCH_CHECK();
    return 0;
  };
  if (parts & 2) {
    const int r = launch(attn_tc_bwd_mt_kernel<DH, 1>, gq);
    if (r != 0) return r;
  }
  return (parts & 1) ? launch(attn_tc_bwd_mt_kernel<DH, 0>, gk) : 0;
}

// which parts of the multi-tile backward the multi-step kernels (attn_tc2.cu) take: bit 1 the dQ pass, bit 0 the
// dK | dV pass.  MEANT_ATTN_BWD_MS (0..3, read once) overrides for A/B runs; MEANT_ATTN_OLD_MT forces 0.
static int bwd_ms_parts() {
  static const int v = [] {
    if (getenv("MEANT_ATTN_OLD_MT")) return 0;
    const char* e = getenv("MEANT_ATTN_BWD_MS");
    return e ? atoi(e) & 3 : 3;
  }();
  return v;
}

static std::atomic<long long> g_tma_fallbacks{0};   // single-tile problems that had to leave the TMA kernels

int attn_tc_fwd(const AttnArgs& a, cudaStream_t st) {
  if (attn_tc2_supported(a, 0)) {
    const int r2 = attn_tc2_fwd(a, st);
    if (r2 != MEANT_EUNSUPPORTED) return r2;     // views TMA cannot describe fall through to the multi-tile kernel
    ++g_tma_fallbacks;
  }
  static const bool old_mt = getenv("MEANT_ATTN_OLD_MT") != nullptr;   // read once: A/B against the per-tile-pair kernel below
  static const bool no_wk = getenv("MEANT_ATTN_NO_WK") != nullptr;      // read once: A/B against the multi-step kernel
  if (!old_mt && !no_wk && attn_tc2_wk_supported(a)) {
    const int r2 = attn_tc2_fwd_wk(a, st);
    if (r2 != MEANT_EUNSUPPORTED) return r2;
  }
  if (!old_mt && attn_tc2_ms_supported(a)) {
    const int r2 = attn_tc2_fwd_ms(a, st);
    if (r2 != MEANT_EUNSUPPORTED) return r2;
  }
  dim3 grid((a.Sq + TQ - 1) / TQ, a.H, a.N);
  FwdMtMaps maps;
  memset(&maps, 0, sizeof(maps));
  {
    const bool pre = a.q_rot.p != nullptr;
    const int skb = a.Sk - (a.kx.p != nullptr ? 1 : 0);
    const HeadView* vs[3] = {pre ? &a.q_rot : &a.q, pre ? &a.k_rot : &a.k, &a.v};
    const int seq[3] = {a.Sq, skb, skb};
    int ok = !mt_no_tma();
    for (int i = 0; i < 3 && ok; ++i) ok = tma::make_chunk_map(&maps.m[i], *vs[i], a.N, a.n2_count, a.H, seq[i], a.Dh) == 0;
    if (ok) ok = tma::make_head_map(&maps.o[0], a.o, a.N, a.n2_count, a.H, a.Sq, a.Dh, 64) == 0;
    if (ok && a.Dh > 64) ok = tma::make_head_map(&maps.o[1], a.o, a.N, a.n2_count, a.H, a.Sq, a.Dh, 32) == 0;
    maps.use_tma = ok;
  }
  auto launch = [&](auto kern, size_t smem) -> int {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<grid, 256, smem, st>>>(a, maps);
    MEANT_LAUNCH_CHECK();
    return 0;
  };
  if (a.Dh == 96) return launch(attn_tc_fwd_kernel<96>, FwdSmem<96>::TOTAL);
  return launch(attn_tc_fwd_kernel<64>, FwdSmem<64>::TOTAL);
}

int attn_tc_bwd(const AttnArgs& a, cudaStream_t st) {
  if (attn_tc2_supported(a, 1)) {
    const int r2 = attn_tc2_bwd(a, st);
    if (r2 != MEANT_EUNSUPPORTED) return r2;
    ++g_tma_fallbacks;
  }
  if (a.Sq > TQ || a.Sk > TK || a.kx.p != nullptr) {
    int ms = attn_tc2_ms_bwd_supported(a) ? bwd_ms_parts() : 0;
    if (ms) {
      const int r2 = attn_tc2_bwd_ms(a, ms, st);           // delta + the selected passes
      if (r2 == MEANT_EUNSUPPORTED) ms = 0;
      else if (r2 != 0) return r2;
    }
    if (ms == 3) return 0;
    return a.Dh == 96 ? launch_bwd_mt<96>(a, st, 3 & ~ms) : launch_bwd_mt<64>(a, st, 3 & ~ms);
  }
  dim3 grid(1, a.H, a.N);
  if (a.Dh == 96) return launch_tc(attn_tc_bwd_kernel<96>, BwdSmem<96>::TOTAL, grid, a, st);
  return launch_tc(attn_tc_bwd_kernel<64>, BwdSmem<64>::TOTAL, grid, a, st);
}

long long attn_tma_fallbacks() { return g_tma_fallbacks.load(); }

int attn_prerotate(const AttnArgs& a, int dtype, cudaStream_t st) {
  if (dtype != MEANT_DT_BF16 || !a.qcos || !a.qsin || !a.kcos || !a.ksin || a.rot_dim <= 0) return MEANT_EARG;
  if ((a.q_rot.p == a.q.p) != (a.k_rot.p == a.k.p)) return MEANT_EARG;      // in place for both or for neither
  if (!view_ok(a.q) || !view_ok(a.k) || !view_ok(a.q_rot) || !view_ok(a.k_rot)) return MEANT_EUNSUPPORTED;
  if (a.N <= 0 || a.H <= 0 || a.Sq <= 0 || a.Sk <= 0 || a.n2_count <= 0 || a.Dh % 8 != 0 || a.rot_dim % 8 != 0 ||
      a.rot_dim > a.Dh)
    return MEANT_EARG;
  const int64_t total = (int64_t)a.N * (a.Sq > a.Sk ? a.Sq : a.Sk) * a.H * (a.Dh / 8);
  int64_t blocks = (total + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  attn_prerotate_kernel<<<dim3((unsigned)blocks, 2), 256, 0, st>>>(a);
  MEANT_LAUNCH_CHECK();
  return 0;
}

}  // namespace meant

extern "C" int meant_debug_prof_mt(unsigned long long* out32, int reset) {
#ifdef MEANT_PROF
  cudaDeviceSynchronize();
  if (out32 && cudaMemcpyFromSymbol(out32, meant::g_prof_mt, sizeof(unsigned long long) * 32) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[32] = {0};
    cudaMemcpyToSymbol(meant::g_prof_mt, z, sizeof(z));
  }
  return 1;
#else
  (void)out32; (void)reset;
  return 0;
#endif
}
