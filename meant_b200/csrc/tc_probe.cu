// Bring-up / self-test kernel for the tcgen05 building blocks: one 128 x N x K bf16 GEMM tile
// D = op(A) op(B)^T with either operand K-major or MN-major from the shared "chunk layout"
// (tc_utils.cuh).  tests/test_tc_probe_gpu.py checks it against torch for every major-ness
// the attention kernels rely on.
#include "common.cuh"
#include "tc_utils.cuh"

namespace meant {

// X[R][C] (row stride ld elements, C % 8 == 0) -> chunk layout at smem
__device__ __forceinline__ void stage_chunks(const __nv_bfloat16* __restrict__ g, int R, int C, int64_t ld,
                                             uint8_t* smem) {
  const int cpr = C >> 3;
  const uint32_t CS = R * 16 + 16;
  for (int idx = threadIdx.x; idx < R * cpr; idx += blockDim.x) {
    const int r = idx / cpr, ch = idx - r * cpr;
    const uint4 v = *reinterpret_cast<const uint4*>(g + (int64_t)r * ld + ch * 8);
    *reinterpret_cast<uint4*>(smem + ch * CS + r * 16) = v;
  }
}

__global__ void __launch_bounds__(128) tc_probe_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D,
                                                       int N, int K, int a_mn, int b_mn) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int M = 128;
  const int warp = threadIdx.x >> 5;
  // operand staging
  const int Ra = a_mn ? K : M, Ca = a_mn ? M : K;
  const int Rb = b_mn ? K : N, Cb = b_mn ? N : K;
  const uint32_t CSa = Ra * 16 + 16, CSb = Rb * 16 + 16;
  uint8_t* sA = smem;
  uint8_t* sB = smem + ((CSa * (Ca >> 3) + 127) & ~127u);
  stage_chunks(A, Ra, Ca, Ca, sA);
  stage_chunks(B, Rb, Cb, Cb, sB);
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    const uint32_t idesc = tc::make_idesc(M, N, a_mn, b_mn);
    const uint32_t a0 = tc::smem_u32(sA), b0 = tc::smem_u32(sB);
    for (int ks = 0; ks < K / 16; ++ks) {
      const uint64_t da = a_mn ? tc::make_desc(a0 + ks * 256, 128, CSa) : tc::make_desc(a0 + ks * 2 * CSa, CSa, 128);
      const uint64_t db = b_mn ? tc::make_desc(b0 + ks * 256, 128, CSb) : tc::make_desc(b0 + ks * 2 * CSb, CSb, 128);
      tc::mma_ss(tmem, da, db, idesc, ks > 0);
    }
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  const int row = threadIdx.x;
  for (int c0 = 0; c0 < N; c0 += 16) {
    float v[16];
    tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tc::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) D[(int64_t)row * N + c0 + i] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

// Same GEMM tile with SWIZZLED operand sub-tiles (tc_utils.cuh): the contiguous axis of an operand is cut
// into 64-column SWIZZLE_128B sub-tiles (or one 32-column SWIZZLE_64B sub-tile when it is 32 wide).
struct SwTile {
  int R, C, w, rowb, nsub;       // rows, cols, sub-tile width (64 | 32), row bytes, number of sub-tiles
  uint32_t type, sub_stride;     // layout type (2 | 4), bytes between sub-tiles
};
__device__ __forceinline__ SwTile sw_tile(int R, int C) {
  SwTile t;
  t.R = R; t.C = C;
  t.w = (C % 64 == 0) ? 64 : 32;
  t.rowb = t.w * 2;
  t.nsub = C / t.w;
  t.type = t.w == 64 ? 2u : 4u;
  t.sub_stride = (uint32_t)(R * t.rowb);
  return t;
}
__device__ __forceinline__ void stage_sw(const __nv_bfloat16* __restrict__ g, const SwTile& t, uint8_t* smem) {
  const int cpr = t.C >> 3, cps = t.w >> 3;
  for (int idx = threadIdx.x; idx < t.R * cpr; idx += blockDim.x) {
    const int r = idx / cpr, ch = idx - r * cpr, sub = ch / cps, c = ch - sub * cps;
    const uint4 v = *reinterpret_cast<const uint4*>(g + (int64_t)r * t.C + ch * 8);
    *reinterpret_cast<uint4*>(smem + sub * t.sub_stride + (t.w == 64 ? tc::sw128_off(r, c) : tc::sw64_off(r, c))) = v;
  }
}
// descriptor of K step ks (16 K elements); kmajor: K runs along the columns, else along the rows
__device__ __forceinline__ uint64_t sw_desc(uint32_t base, const SwTile& t, int ks, bool kmajor) {
  if (kmajor) {
    const int col = ks * 16, sub = col / t.w, within = col - sub * t.w;
    return tc::make_desc_sw(base + sub * t.sub_stride + within * 2, 16, 8 * t.rowb, t.type);
  }
  return tc::make_desc_sw(base + ks * 16 * t.rowb, t.sub_stride, 8 * t.rowb, t.type);
}

__global__ void __launch_bounds__(128) tc_probe_sw_kernel(const __nv_bfloat16* A, const __nv_bfloat16* B, float* D,
                                                          int N, int K, int a_mn, int b_mn) {
  extern __shared__ __align__(128) uint8_t smem_raw[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  uint8_t* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
  const int M = 128, warp = threadIdx.x >> 5;
  const SwTile ta = sw_tile(a_mn ? K : M, a_mn ? M : K), tb = sw_tile(b_mn ? K : N, b_mn ? N : K);
  uint8_t* sA = smem;
  uint8_t* sB = smem + ((ta.nsub * ta.sub_stride + 1023) & ~1023u);
  stage_sw(A, ta, sA);
  stage_sw(B, tb, sB);
  if (threadIdx.x == 0) {
    tc::mbar_init(&bar, 1);
    tc::fence_mbar_init();
  }
  if (warp == 0) tc::tmem_alloc(&tmem_slot, 256);
  tc::fence_async_smem();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = tmem_slot;
  if (threadIdx.x == 0) {
    // an MN-major B operand wider than one sub-tile is issued as one MMA per sub-tile (N = 64 or 32 columns of D each)
    const uint32_t a0 = tc::smem_u32(sA), b0 = tc::smem_u32(sB);
    const int nb = b_mn ? tb.nsub : 1, nw = b_mn ? tb.w : N;
    for (int j = 0; j < nb; ++j) {
      const uint32_t idesc = tc::make_idesc(M, nw, a_mn, b_mn);
      for (int ks = 0; ks < K / 16; ++ks) {
        uint64_t db = sw_desc(b0 + (b_mn ? j * tb.sub_stride : 0), tb, ks, !b_mn);
        tc::mma_ss(tmem + j * nw, sw_desc(a0, ta, ks, !a_mn), db, idesc, ks > 0);
      }
    }
    tc::mma_commit(&bar);
  }
  tc::mbar_wait(&bar, 0);
  tc::fence_after_sync();
  const int row = threadIdx.x;
  for (int c0 = 0; c0 < N; c0 += 16) {
    float v[16];
    tc::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    tc::tmem_ld_wait();
#pragma unroll
    for (int i = 0; i < 16; ++i) D[(int64_t)row * N + c0 + i] = v[i];
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 256);
}

}  // namespace meant

extern "C" int meant_tc_probe_sw(const void* A, const void* B, float* D, int N, int K, int a_mn, int b_mn,
                                 meant_stream_t stream) {
  if (!A || !B || !D || N < 16 || N > 256 || N % 16 || K < 16 || K % 16 || K > 256) return MEANT_EARG;
  auto ok = [](int c) { return c == 32 || c % 64 == 0; };
  if (!ok(a_mn ? 128 : K) || !ok(b_mn ? N : K)) return MEANT_EUNSUPPORTED;
  const size_t smem = (size_t)(128 * K * 2 + 1024) + (size_t)(N * K * 2 + 1024) + 2048;
  cudaError_t e = cudaFuncSetAttribute(meant::tc_probe_sw_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  meant::tc_probe_sw_kernel<<<1, 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      (const __nv_bfloat16*)A, (const __nv_bfloat16*)B, D, N, K, a_mn, b_mn);
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_tc_probe(const void* A, const void* B, float* D, int N, int K, int a_mn, int b_mn,
                              meant_stream_t stream) {
  if (!A || !B || !D || N < 16 || N > 256 || N % 16 || K < 16 || K % 16 || K > 256) return MEANT_EARG;
  const int Ra = a_mn ? K : 128, Ca = a_mn ? 128 : K, Rb = b_mn ? K : N, Cb = b_mn ? N : K;
  const size_t smem = (((size_t)(Ra * 16 + 16) * (Ca / 8) + 127) & ~(size_t)127) + (size_t)(Rb * 16 + 16) * (Cb / 8) + 128;
  cudaError_t e = cudaFuncSetAttribute(meant::tc_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return (int)e;
  meant::tc_probe_kernel<<<1, 128, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
      (const __nv_bfloat16*)A, (const __nv_bfloat16*)B, D, N, K, a_mn, b_mn);
  MEANT_LAUNCH_CHECK();
  return 0;
}
