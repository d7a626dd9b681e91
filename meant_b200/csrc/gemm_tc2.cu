// The other GEMM forms of the Linear layers, same tcgen05 / TMA / TMEM pipeline as gemm_tc.cu:
//   meant_gemm_bf16        C[M,N] (bf16) = op(A) op(B) (+ bias[N]) (+ R[M,N]); either operand K-major or MN-major:
//                            forward   y  = x W^T + b (+ residual)      A = x  [M,K],  B = W  [N,K]   (a_mn 0, b_mn 0)
//                            dgrad     dX = dY W                        A = dY [M,K'], B = W  [K',N]  (a_mn 0, b_mn 1)
//   meant_gemm_bf16_wgrad  dW[N,K] (fp32) += dY[T,N]^T X[T,K],  db[N] (fp32) += column sums of dY
//                          (meant/meant.py:101-107, meant/xPosAttention.py:24-32: the backward of every nn.Linear)
// MN-major operands are fetched as 64-column x 64-row SWIZZLE_128B boxes of the row-major tensor (the reduction index
// runs down the rows) and described to the MMA with the MN-major descriptor convention tests/test_tc_probe_gpu.py pins.
// Nothing is transposed in memory.
//
// Weight gradient: the reduction runs over T = tokens (98304) while the output has only (N/128) x (K/256) = 18..54
// tiles, so the reduction is cut into `splits` ranges (chosen on the host so that tiles x splits fills whole waves of
// SMs) and every (split, tile) item flushes its fp32 partial tile with TMA reduce-add (cp.reduce.async.bulk.tensor
// ... add).  Items are dealt split-major, so the CTAs of a wave walk the SAME token rows of different tiles at the same
// pace and dY / X boxes are shared through L2.  The bias gradient costs no pass over dY: items of the first tile
// column issue one extra N = 16 MMA per K step against a constant tile of ones (16 extra TMEM columns), i.e.
// sum_t dY[t, n] comes out of the tensor core.
// Summation order of the partial tiles is not fixed: fp32 results can differ in the last bits from run to run.
#include "common.cuh"
#include "tc_utils.cuh"
#include "tma_utils.cuh"
#include <string.h>
#include <stdio.h>
#include <stdlib.h>

namespace meant {
namespace g2 {

constexpr int BM = 128, BN = 256, BK = 64, NSTAGE = 4, NT = 192, BAR_EPI = 1;
constexpr uint32_t A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2, STAGE = A_BYTES + B_BYTES;
constexpr uint32_t SUB = BK * 128;                           // one MN-major sub-tile: BK rows x 64 columns
constexpr uint32_t OUT_BOX = 128 * 128;                      // 128 rows x 128 bytes (64 bf16 | 32 fp32), SWIZZLE_128B
constexpr uint32_t OFF_OUT = NSTAGE * STAGE, OFF_ONES = OFF_OUT + 2 * OUT_BOX, OFF_BARS = OFF_ONES + 512;
constexpr uint32_t SMEM_TOTAL = OFF_BARS + 256 + 1024;
static_assert(SMEM_TOTAL <= 232448, "exceeds the shared memory a CTA can opt in to");

struct Maps { CUtensorMap a, b, c, r; };
struct Params {
  const float* bias;     // MODE 0: [N] or null
  float* db;             // MODE 1: [M] or null
  int M, N, K;           // output rows, output columns, reduction length
  int has_res;           // MODE 0: add the tile of maps.r
  int bsplit;            // MN-major B: four N = 64 MMAs per K step instead of one N = 256 MMA (measured 1.5x slower; experiments)
  int splits;            // MODE 1: number of K splits (<= K steps)
};

__device__ __forceinline__ void load_box2(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
               ::"r"(dst), "l"(map), "r"(tc::smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void store_box2(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%2, %3}], [%1];\n"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void reduce_add_box2(const CUtensorMap* map, uint32_t src, int c0, int c1) {
  asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];\n"
               ::"l"(map), "r"(src), "r"(c0), "r"(c1) : "memory");
}

// one unit of work: an output tile and a range of K steps
struct Work { int m0, n0, kb, ke, tn; };

// MODE 0: tiles dealt round-robin, whole reduction each.  MODE 1: items = (K split, tile), split-major, dealt round-robin:
// the CTAs of one wave sweep the SAME rows of the two operands at the same pace, so a row block comes from HBM once and
// from L2 for every other tile (a tile-major cut of the (tile, k) space measured 1.6x slower: every CTA streamed its own rows).
template <int MODE>
struct Sched {
  int tiles_n, kt, t, tiles, splits;
  __device__ Sched(int M, int N, int K, int splits_) {
    tiles_n = (N + BN - 1) / BN;
    tiles = ((M + BM - 1) / BM) * tiles_n;
    kt = (K + BK - 1) / BK;
    splits = MODE == 0 ? 1 : splits_;
    t = blockIdx.x;
  }
  __device__ bool next(Work& w) {
    if (t >= tiles * splits) return false;
    const int sp = t / tiles, tile = t - sp * tiles;
    w.m0 = (tile / tiles_n) * BM; w.tn = tile % tiles_n; w.n0 = w.tn * BN;
    w.kb = (int)((long long)sp * kt / splits);
    w.ke = (int)((long long)(sp + 1) * kt / splits);
    t += gridDim.x;
    return true;
  }
};

template <bool A_MN, bool B_MN, int MODE>
__global__ void __launch_bounds__(NT, 1) gemm_gen_kernel(const __grid_constant__ Maps maps, const Params p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  constexpr int NACC = MODE == 0 ? 2 : 1;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BARS);
  uint64_t *full = bars, *empty = bars + NSTAGE, *acc_full = bars + 2 * NSTAGE, *acc_empty = acc_full + 2, *res_bar = acc_empty + 2;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    for (int i = 0; i < NSTAGE; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_empty[i], 128); }
    tc::mbar_init(res_bar, 1);
    tc::fence_mbar_init();
  }
  if (MODE == 1) {                                         // the constant tile of ones (bf16 1.0 = 0x3F80)
    for (int i = tid; i < 512 / 4; i += NT) reinterpret_cast<uint32_t*>(smem + OFF_ONES)[i] = 0x3F803F80u;
    tc::fence_async_smem();
  }
  if (warp == 1) tc::tmem_alloc(&tmem_slot, 512);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot, s0 = tc::smem_u32(smem);
  const bool want_db = MODE == 1 && p.db != nullptr;

  if (warp == 0) {
    if (lane == 0) {
      tma::prefetch_map(&maps.a);
      tma::prefetch_map(&maps.b);
      Sched<MODE> sch(p.M, p.N, p.K, p.splits);
      Work w;
      int it = 0;
      while (sch.next(w)) {
        for (int k = w.kb; k < w.ke; ++k, ++it) {
          const int s = it % NSTAGE;
          if (it >= NSTAGE) tc::mbar_wait(&empty[s], ((it / NSTAGE) - 1) & 1);
          tma::mbar_expect_tx(&full[s], STAGE);
          const uint32_t sa = s0 + s * STAGE, sb = sa + A_BYTES;
          if (!A_MN) load_box2(sa, &maps.a, &full[s], k * BK, w.m0);
          else {
#pragma unroll
            for (int j = 0; j < BM / 64; ++j) load_box2(sa + j * SUB, &maps.a, &full[s], w.m0 + 64 * j, k * BK);
          }
          if (!B_MN) load_box2(sb, &maps.b, &full[s], k * BK, w.n0);
          else {
#pragma unroll
            for (int j = 0; j < BN / 64; ++j) load_box2(sb + j * SUB, &maps.b, &full[s], w.n0 + 64 * j, k * BK);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      constexpr uint32_t idesc = tc::make_idesc(128, BN, A_MN, B_MN), idesc64 = tc::make_idesc(128, 64, A_MN, B_MN);
      constexpr uint32_t idesc_ones = tc::make_idesc(128, 16, A_MN, 0);
      const uint64_t ones_desc = tc::make_desc(s0 + OFF_ONES, 256, 128);   // K-major, no swizzle: 2 K chunks of 16 rows x 16 B
      Sched<MODE> sch(p.M, p.N, p.K, p.splits);
      Work w;
      int it = 0, ti = 0;
      while (sch.next(w)) {
        const int acc = ti % NACC, use = ti / NACC;
        if (use >= 1) tc::mbar_wait(&acc_empty[acc], (use - 1) & 1);
        tc::fence_after_sync();
        const uint32_t tacc = tmem + acc * BN;
        const bool bias_col = want_db && w.tn == 0;
        for (int k = w.kb; k < w.ke; ++k, ++it) {
          const int s = it % NSTAGE;
          tc::mbar_wait(&full[s], (it / NSTAGE) & 1);
          tc::fence_after_sync();
          const uint32_t sa = s0 + s * STAGE, sb = sa + A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint32_t on = (k > w.kb || kk > 0) ? 1u : 0u;
            const uint64_t ad = A_MN ? tc::make_desc_sw(sa + kk * 2048, SUB, 1024, 2) : tc::make_desc_sw(sa + kk * 32, 16, 1024, 2);
            if (!B_MN) tc::mma_ss(tacc, ad, tc::make_desc_sw(sb + kk * 32, 16, 1024, 2), idesc, on);
            else if (!p.bsplit) tc::mma_ss(tacc, ad, tc::make_desc_sw(sb + kk * 2048, SUB, 1024, 2), idesc, on);
            else {
#pragma unroll
              for (int j = 0; j < BN / 64; ++j)
                tc::mma_ss(tacc + 64 * j, ad, tc::make_desc_sw(sb + j * SUB + kk * 2048, SUB, 1024, 2), idesc64, on);
            }
            if (MODE == 1 && bias_col) tc::mma_ss(tmem + BN, ad, ones_desc, idesc_ones, on);
          }
          tc::mma_commit(&empty[s]);                   // the stage may be refilled once these MMAs have read it
        }
        tc::mma_commit(&acc_full[acc]);
        ++ti;
      }
    }
  } else {
    const int q = warp & 3;                            // TMEM lane quarter this warp may read
    const int row = q * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    const bool leader = warp == 2 && lane == 0;
    Sched<MODE> sch(p.M, p.N, p.K, p.splits);
    Work w;
    int ti = 0;
    uint32_t rphase = 0;
    while (sch.next(w)) {
      const int acc = ti % NACC, use = ti / NACC;
      ++ti;
      tc::mbar_wait(&acc_full[acc], use & 1);
      tc::fence_after_sync();
      constexpr int ROUNDS = MODE == 0 ? 2 : 4, BOXC = MODE == 0 ? 64 : 32;
#pragma unroll 1
      for (int r = 0; r < ROUNDS; ++r) {
        int nvalid = 0;
#pragma unroll
        for (int bi = 0; bi < 2; ++bi) nvalid += (w.n0 + (r * 2 + bi) * BOXC < p.N) ? 1 : 0;
        const bool res = MODE == 0 && p.has_res && nvalid > 0;
        if (leader) {
          tma::store_wait_read<0>();                   // the previous stores have read the staging boxes
          if (res) {
            tma::mbar_expect_tx(res_bar, nvalid * OUT_BOX);
            for (int bi = 0; bi < nvalid; ++bi) load_box2(s0 + OFF_OUT + bi * OUT_BOX, &maps.r, res_bar, w.n0 + (r * 2 + bi) * BOXC, w.m0);
          }
        }
        tc::bar_sync(BAR_EPI, 128);
        if (res) {
          tc::mbar_wait(res_bar, rphase);
          rphase ^= 1;
        }
#pragma unroll
        for (int bi = 0; bi < 2; ++bi) {
          const int g = r * 2 + bi;
          uint8_t* box = smem + OFF_OUT + bi * OUT_BOX;
          if (MODE == 0) {
#pragma unroll
            for (int c = 0; c < 2; ++c) {
              float v[32];
              tc::tmem_ld32(lane_base + acc * BN + g * 64 + c * 32, v);
              tc::tmem_ld_wait();
#pragma unroll
              for (int gq = 0; gq < 4; ++gq) {
                float b[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
                const int col = w.n0 + g * 64 + c * 32 + gq * 8;
                if (p.bias && col < p.N) {             // N is a multiple of 8: a chunk is in or out as a whole
                  const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col)), b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + 1);
                  b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
                }
                uint4* slot = reinterpret_cast<uint4*>(box + tc::sw128_off(row, c * 4 + gq));
                if (res && bi < nvalid) {
                  const uint4 rv = *slot;
                  const __nv_bfloat162* rh = reinterpret_cast<const __nv_bfloat162*>(&rv);
#pragma unroll
                  for (int i = 0; i < 4; ++i) {
                    const float2 f = __bfloat1622float2(rh[i]);
                    b[2 * i] += f.x;
                    b[2 * i + 1] += f.y;
                  }
                }
                uint4 pk;
                __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
                for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(v[gq * 8 + 2 * i] + b[2 * i], v[gq * 8 + 2 * i + 1] + b[2 * i + 1]);
                *slot = pk;
              }
            }
          } else {
            float v[32];
            tc::tmem_ld32(lane_base + g * 32, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int gq = 0; gq < 8; ++gq)
              *reinterpret_cast<float4*>(box + tc::sw128_off(row, gq)) = make_float4(v[4 * gq], v[4 * gq + 1], v[4 * gq + 2], v[4 * gq + 3]);
          }
        }
        if (r == ROUNDS - 1) {
          if (MODE == 1 && want_db && w.tn == 0) {     // the ones column: sum over this item's K range of A[:, row]
            float v[16];
            tc::tmem_ld16(lane_base + BN, v);
            tc::tmem_ld_wait();
            if (w.m0 + row < p.M) atomicAdd(p.db + w.m0 + row, v[0]);
          }
          tc::fence_before_sync();
          tc::mbar_arrive(&acc_empty[acc]);            // the accumulator may be overwritten
        }
        tc::fence_async_smem();
        tc::bar_sync(BAR_EPI, 128);
        if (leader) {
#pragma unroll
          for (int bi = 0; bi < 2; ++bi) {
            const int c0 = w.n0 + (r * 2 + bi) * BOXC;
            if (c0 < p.N && w.m0 < p.M) {
              if (MODE == 0) store_box2(&maps.c, s0 + OFF_OUT + bi * OUT_BOX, c0, w.m0);
              else reduce_add_box2(&maps.c, s0 + OFF_OUT + bi * OUT_BOX, c0, w.m0);
            }
          }
          tma::store_commit();
        }
      }
    }
    if (leader) tma::store_wait_all();
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tmem, 512);
}

// ------------------------------------------------------------------------------------------ CTA pairs (cta_group::2)
// The same GEMM with two CTAs (one cluster, neighbouring SMs) on one 256 x 256 tile: each CTA loads ITS 128 rows of A
// and HALF of the B tile (128 of the 256 output columns) per K step, the leader CTA issues one
// tcgen05.mma.cta_group::2 (M = 256) that feeds both tensor cores -- each reads its own A rows and both halves of B --
// and each CTA's TMEM receives its 128 x 256 half of the accumulator.  Per CTA and K step the TMA writes 32 KB instead
// of 48 KB and the MMA reads a third less shared memory.  Correct, but measured SLOWER than the single-CTA kernel on
// the block's shapes (180 vs 102 us at 98304 x 768 x 768; 103 us with the loads switched off, i.e. MMA + epilogue alone
// already sit at the single-CTA kernel's time: with K = 768 a tile is 12 K steps against a 64 KB epilogue), so it is
// opt-in: MEANT_GEMM_PAIR=1.
//   full[s]      lives in the LEADER: both producers arrive on it (the peer remotely) with their own byte counts, and
//                both CTAs' boxes complete on it (cp.async.bulk.tensor ... cta_group::2 may signal the peer's barrier)
//   empty[s]     in each CTA, signalled by the leader's tcgen05.commit multicast: the stage may be refilled
//   acc_full[a]  in each CTA, same multicast: each epilogue drains its own TMEM half
//   acc_empty[a] in the leader, two arrivals: one elected epilogue thread per CTA (the peer's remotely)
// A is K-major; B K-major (forward) or MN-major (dgrad).  Bias epilogue only.
constexpr int NSTAGE2 = 6;
constexpr uint32_t HALF_B = 128 * BK * 2, STAGE2 = A_BYTES + HALF_B;
constexpr uint32_t OFF_OUT2 = NSTAGE2 * STAGE2, OFF_BARS2 = OFF_OUT2 + 2 * OUT_BOX, SMEM_TOTAL2 = OFF_BARS2 + 256 + 1024;
static_assert(SMEM_TOTAL2 <= 232448, "exceeds the shared memory a CTA can opt in to");

__device__ __forceinline__ uint32_t cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address) in CTA `rank` of this cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void remote_arrive_expect_tx(uint32_t bar_cluster_addr, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [%0], %1;\n" ::"r"(bar_cluster_addr), "r"(bytes) : "memory");
}
__device__ __forceinline__ void remote_arrive(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];\n" ::"r"(bar_cluster_addr) : "memory");
}
// TMA box into this CTA's shared memory, completion counted on a barrier that may live in the peer CTA
__device__ __forceinline__ void load_box2_pair(uint32_t dst, const CUtensorMap* map, uint32_t bar_cluster_addr, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];\n"
               ::"r"(dst), "l"(map), "r"(bar_cluster_addr), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void mma_ss_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_commit_pair(uint64_t* bar) {           // arrives on `bar` in BOTH CTAs
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n"
               ::"r"(tc::smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

template <bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NT, 1) gemm_pair_kernel(const __grid_constant__ Maps maps, const Params p) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + OFF_BARS2);
  uint64_t *full = bars, *empty = bars + NSTAGE2, *acc_full = bars + 2 * NSTAGE2, *acc_empty = acc_full + 2;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = cluster_rank();
  const bool leader = rank == 0;
  if (tid == 0) {
    for (int i = 0; i < NSTAGE2; ++i) { tc::mbar_init(&full[i], 2); tc::mbar_init(&empty[i], 1); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&acc_full[i], 1); tc::mbar_init(&acc_empty[i], 2); }
    tc::fence_mbar_init();
  }
  if (warp == 1) {                                         // one warp of EACH CTA of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(tc::smem_u32(&tmem_slot)), "r"(512u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;\n" ::: "memory");
  }
  tc::fence_before_sync();
  __syncthreads();
  cluster_sync_all();                                      // both CTAs' barriers and TMEM exist before anything remote happens
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot, s0 = tc::smem_u32(smem);
  const int tiles_n = (p.N + BN - 1) / BN, tiles_m = (p.M + 255) / 256, tiles = tiles_m * tiles_n, kt = (p.K + BK - 1) / BK;
  const int cluster = blockIdx.x >> 1, n_clusters = gridDim.x >> 1;

  if (warp == 0) {
    if (lane == 0) {
      tma::prefetch_map(&maps.a);
      tma::prefetch_map(&maps.b);
      int it = 0;
      for (int t = cluster; t < tiles; t += n_clusters) {
        const int m0 = (t / tiles_n) * 256 + (int)rank * 128, n0 = (t % tiles_n) * BN + (int)rank * 128;
        for (int k = 0; k < kt; ++k, ++it) {
          const int s = it % NSTAGE2;
          if (it >= NSTAGE2) tc::mbar_wait(&empty[s], ((it / NSTAGE2) - 1) & 1);
          const uint32_t fb = map_to_cta(tc::smem_u32(&full[s]), 0);       // the leader's barrier
          remote_arrive_expect_tx(fb, STAGE2);
          const uint32_t sa = s0 + s * STAGE2, sb = sa + A_BYTES;
          load_box2_pair(sa, &maps.a, fb, k * BK, m0);
          if (!B_MN) load_box2_pair(sb, &maps.b, fb, k * BK, n0);
          else {
#pragma unroll
            for (int j = 0; j < 2; ++j) load_box2_pair(sb + j * SUB, &maps.b, fb, n0 + 64 * j, k * BK);
          }
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && leader) {
      constexpr uint32_t idesc = tc::make_idesc(256, BN, 0, B_MN);
      int it = 0, ti = 0;
      for (int t = cluster; t < tiles; t += n_clusters, ++ti) {
        const int acc = ti & 1, use = ti >> 1;
        if (use >= 1) tc::mbar_wait(&acc_empty[acc], (use - 1) & 1);
        tc::fence_after_sync();
        for (int k = 0; k < kt; ++k, ++it) {
          const int s = it % NSTAGE2;
          tc::mbar_wait(&full[s], (it / NSTAGE2) & 1);
          tc::fence_after_sync();
          const uint32_t sa = s0 + s * STAGE2, sb = sa + A_BYTES;
#pragma unroll
          for (int kk = 0; kk < BK / 16; ++kk) {
            const uint64_t ad = tc::make_desc_sw(sa + kk * 32, 16, 1024, 2);
            const uint64_t bd = B_MN ? tc::make_desc_sw(sb + kk * 2048, SUB, 1024, 2) : tc::make_desc_sw(sb + kk * 32, 16, 1024, 2);
            mma_ss_pair(tmem + acc * BN, ad, bd, idesc, (k > 0 || kk > 0) ? 1u : 0u);
          }
          mma_commit_pair(&empty[s]);                      // both CTAs may refill the stage
        }
        mma_commit_pair(&acc_full[acc]);
      }
    }
  } else {
    const int q = warp & 3;
    const int row = q * 32 + lane;
    const uint32_t lane_base = tmem + ((uint32_t)(q * 32) << 16);
    const bool elected = warp == 2 && lane == 0;
    int ti = 0;
    for (int t = cluster; t < tiles; t += n_clusters, ++ti) {
      const int m0 = (t / tiles_n) * 256 + (int)rank * 128, n0 = (t % tiles_n) * BN, acc = ti & 1, use = ti >> 1;
      tc::mbar_wait(&acc_full[acc], use & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int r = 0; r < 2; ++r) {
        if (elected) tma::store_wait_read<0>();
        tc::bar_sync(BAR_EPI, 128);
#pragma unroll
        for (int bi = 0; bi < 2; ++bi) {
          const int g = r * 2 + bi;
          uint8_t* box = smem + OFF_OUT2 + bi * OUT_BOX;
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            float v[32];
            tc::tmem_ld32(lane_base + acc * BN + g * 64 + c * 32, v);
            tc::tmem_ld_wait();
#pragma unroll
            for (int gq = 0; gq < 4; ++gq) {
              float b[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
              const int col = n0 + g * 64 + c * 32 + gq * 8;
              if (p.bias && col < p.N) {
                const float4 b0 = __ldg(reinterpret_cast<const float4*>(p.bias + col)), b1 = __ldg(reinterpret_cast<const float4*>(p.bias + col) + 1);
                b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
              }
              uint4 pk;
              __nv_bfloat162* ph = reinterpret_cast<__nv_bfloat162*>(&pk);
#pragma unroll
              for (int i = 0; i < 4; ++i) ph[i] = __floats2bfloat162_rn(v[gq * 8 + 2 * i] + b[2 * i], v[gq * 8 + 2 * i + 1] + b[2 * i + 1]);
              *reinterpret_cast<uint4*>(box + tc::sw128_off(row, c * 4 + gq)) = pk;
            }
          }
        }
        tc::fence_async_smem();
        tc::fence_before_sync();
        tc::bar_sync(BAR_EPI, 128);
        if (elected) {
          if (r == 1) remote_arrive(map_to_cta(tc::smem_u32(&acc_empty[acc]), 0));   // this CTA's half of the accumulator is drained
#pragma unroll
          for (int bi = 0; bi < 2; ++bi) {
            const int c0 = n0 + (r * 2 + bi) * 64;
            if (c0 < p.N && m0 < p.M) store_box2(&maps.c, s0 + OFF_OUT2 + bi * OUT_BOX, c0, m0);
          }
          tma::store_commit();
        }
      }
    }
    if (elected) tma::store_wait_all();
  }
  tc::fence_before_sync();
  __syncthreads();
  cluster_sync_all();                                      // nobody leaves (or frees TMEM) while the peer may still signal it
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;\n" ::"r"(tmem), "r"(512u) : "memory");
}

// 2-D tensor map over a row-major tensor: `cols` contiguous elements per row, row pitch ld elements, SWIZZLE_128B
static int make_map2(CUtensorMap* m, const void* p, int64_t rows, int64_t cols, int64_t ld, int box_cols, int box_rows, bool f32) {
  tma::EncodeTiledFn enc = tma::encode_fn();
  const int es = f32 ? 4 : 2;
  if (!enc || (reinterpret_cast<uintptr_t>(p) & 15) || (ld * es) % 16 || cols <= 0 || rows <= 0) return MEANT_EUNSUPPORTED;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)(ld * es)};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1u, 1u};
  CUresult r = CUDA_ERROR_INVALID_CONTEXT;
  for (int attempt = 0; attempt < 2 && r == CUDA_ERROR_INVALID_CONTEXT; ++attempt) {
    if (attempt) {                                     // see tma_utils.cuh: threads without a current driver context
      int dev = 0;
      if (cudaGetDevice(&dev) != cudaSuccess || cudaSetDevice(dev) != cudaSuccess) break;
    }
    r = enc(m, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(p), dims, strides,
            box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  return r == CUDA_SUCCESS ? 0 : MEANT_EUNSUPPORTED;
}

static int sm_count() {
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  return sms;
}

template <typename Kern>
static int launch(Kern kern, int grid, const Maps& maps, const Params& p, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_TOTAL);
  if (e != cudaSuccess) return (int)e;
  kern<<<grid, NT, SMEM_TOTAL, st>>>(maps, p);
  MEANT_LAUNCH_CHECK();
  return 0;
}

}  // namespace g2
}  // namespace meant

using namespace meant;

extern "C" int meant_gemm_bf16(const void* A, const void* B, void* C, int64_t M, int N, int K, int64_t lda, int64_t ldb,
                               int64_t ldc, int a_mn, int b_mn, const float* bias, const void* residual, int64_t ldr,
                               meant_stream_t stream) {
  if (!A || !B || !C || M < 0 || N <= 0 || K <= 0 || ldc < N || (residual && ldr < N)) return MEANT_EARG;
  if (lda < (a_mn ? M : K) || ldb < (b_mn ? N : K)) return MEANT_EARG;
  if (N % 8 || K % 8 || (a_mn && M % 8) || (bias && (reinterpret_cast<uintptr_t>(bias) & 15)) || M > 0x7fffffff) return MEANT_EUNSUPPORTED;
  if (M == 0) return 0;
  g2::Maps maps;
  memset(&maps, 0, sizeof(maps));
  int rc = a_mn ? g2::make_map2(&maps.a, A, K, M, lda, 64, g2::BK, false) : g2::make_map2(&maps.a, A, M, K, lda, g2::BK, 128, false);
  if (!rc) rc = b_mn ? g2::make_map2(&maps.b, B, K, N, ldb, 64, g2::BK, false) : g2::make_map2(&maps.b, B, N, K, ldb, g2::BK, g2::BN, false);
  if (!rc) rc = g2::make_map2(&maps.c, C, M, N, ldc, 64, 128, false);
  if (!rc && residual) rc = g2::make_map2(&maps.r, residual, M, N, ldr, 64, 128, false);
  if (rc) return rc;
  static const bool env_split = getenv("MEANT_GEMM_BSPLIT") != nullptr;     // read once
  g2::Params p{bias, nullptr, (int)M, N, K, residual ? 1 : 0, env_split ? 1 : 0, 1};
  const int sms = g2::sm_count();
  const int64_t tiles = ((M + g2::BM - 1) / g2::BM) * ((N + g2::BN - 1) / g2::BN);
  const int grid = (int)(tiles < sms ? tiles : sms);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  // CTA pairs (cta_group::2) for the large forward / dgrad problems: MEANT_GEMM_PAIR=1 (read once; experiments)
  static const bool env_pair = getenv("MEANT_GEMM_PAIR") != nullptr;
  if (env_pair && !a_mn && !residual && sms % 2 == 0 && ((M + 255) / 256) * ((N + g2::BN - 1) / g2::BN) >= sms / 2) {
    g2::Maps m2 = maps;
    int rc2 = b_mn ? 0 : g2::make_map2(&m2.b, B, N, K, ldb, g2::BK, 128, false);     // half a B tile per CTA
    if (rc2) return rc2;
    auto kern = b_mn ? g2::gemm_pair_kernel<true> : g2::gemm_pair_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)g2::SMEM_TOTAL2);
    if (e != cudaSuccess) return (int)e;
    // A pair needs both SMs of one TPC; not every TPC of a B200 has two SMs enabled, so fewer than sms / 2 pairs can
    // be resident.  The persistent grid must not exceed that number: a pair that starts late starts after the early
    // ones have taken its tiles' time (a grid of sms CTAs measured exactly 2x slower).
    static int pairs_cached[2] = {0, 0};
    int& pairs = pairs_cached[b_mn ? 1 : 0];
    if (pairs == 0) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3((unsigned)sms);
      cfg.blockDim = dim3(g2::NT);
      cfg.dynamicSmemBytes = g2::SMEM_TOTAL2;
      cudaLaunchAttribute at;
      at.id = cudaLaunchAttributeClusterDimension;
      at.val.clusterDim.x = 2;
      at.val.clusterDim.y = 1;
      at.val.clusterDim.z = 1;
      cfg.attrs = &at;
      cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n <= 0) n = sms / 2;
      if (n > sms / 2) n = sms / 2;
      pairs = n;
      if (getenv("MEANT_DEBUG")) fprintf(stderr, "meant_gemm_bf16: %d resident CTA pairs of %d SMs\n", n, sms);
    }
    kern<<<2 * pairs, g2::NT, g2::SMEM_TOTAL2, st>>>(m2, p);
    MEANT_LAUNCH_CHECK();
    return 0;
  }
  if (!a_mn && !b_mn) return g2::launch(g2::gemm_gen_kernel<false, false, 0>, grid, maps, p, st);
  if (!a_mn && b_mn) return g2::launch(g2::gemm_gen_kernel<false, true, 0>, grid, maps, p, st);
  if (a_mn && !b_mn) return g2::launch(g2::gemm_gen_kernel<true, false, 0>, grid, maps, p, st);
  return g2::launch(g2::gemm_gen_kernel<true, true, 0>, grid, maps, p, st);
}

extern "C" int meant_gemm_bf16_wgrad(const void* dY, const void* X, float* dW, float* db, int64_t T, int N, int K,
                                     int64_t lddy, int64_t ldx, int64_t lddw, meant_stream_t stream) {
  if (!dY || !X || !dW || T < 0 || N <= 0 || K <= 0 || lddy < N || ldx < K || lddw < K) return MEANT_EARG;
  if (N % 8 || K % 8 || T > 0x7fffffff) return MEANT_EUNSUPPORTED;
  if (T == 0) return 0;
  g2::Maps maps;
  memset(&maps, 0, sizeof(maps));
  int rc = g2::make_map2(&maps.a, dY, T, N, lddy, 64, g2::BK, false);          // A[m = n, k = t] = dY[t, n]: MN-major
  if (!rc) rc = g2::make_map2(&maps.b, X, T, K, ldx, 64, g2::BK, false);       // B[n = k, k = t] = X[t, k]:  MN-major
  if (!rc) rc = g2::make_map2(&maps.c, dW, N, K, lddw, 32, 128, true);
  if (rc) return rc;
  static const bool env_split = getenv("MEANT_GEMM_BSPLIT") != nullptr;     // read once
  const int sms = g2::sm_count();
  const int64_t tiles = (int64_t)((N + g2::BM - 1) / g2::BM) * ((K + g2::BN - 1) / g2::BN), kt = (T + g2::BK - 1) / g2::BK;
  // K splits: minimise waves x (K steps per item + the un-overlapped flush, ~14 K steps' worth of time)
  int splits = 1;
  double best = 1e30;
  for (int s = 1; s <= 32 && s <= kt; ++s) {
    const int64_t waves = (tiles * s + sms - 1) / sms;
    const double cost = (double)waves * ((double)kt / s + 14.0);
    if (cost < best) { best = cost; splits = s; }
  }
  g2::Params p{nullptr, db, N, K, (int)T, 0, env_split ? 1 : 0, splits};
  const int64_t total = tiles * splits;
  const int grid = (int)(total < sms ? total : sms);
  return g2::launch(g2::gemm_gen_kernel<true, true, 1>, grid, maps, p, reinterpret_cast<cudaStream_t>(stream));
}
