// TMA (cp.async.bulk.tensor) helpers: tensor maps over head views, built on the host through the driver
// entry point (no libcuda link), and the load / store / mbarrier instructions (sm_100a only).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "attn_common.cuh"

namespace meant {
namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// Tensor map over a bf16 head view (attn_common.cuh): element (n, s, h, d) at p + n1 sn1 + n2 sn2 + s ss + h sh + d.
// Dimensions, fastest first: (d, h, s, n2, n1); the box is `box_cols` columns x 128 rows of ONE head, written to
// shared memory as 128 rows of box_cols * 2 bytes with the matching swizzle (128 B rows: SWIZZLE_128B, 64 B rows:
// SWIZZLE_64B) -- the operand sub-tile layout of attn_tc2.cu.  Rows >= S are out of bounds: zero-filled on load,
// dropped on store.  Returns 0, or MEANT_EUNSUPPORTED when the view cannot be described.
// box_rows x box_n2: rows of one sequence x consecutive n2 sequences per box (default: one 128-row tile of one sequence;
// the packed-group kernels of attn_tc2.cu take gl rows of G groups, which land as tile rows g * gl + s; the wide-key
// forward takes every key of a sequence, up to 208 rows, in one box).
inline int make_head_map(CUtensorMap* m, const HeadView& v, int N, int n2_count, int H, int S, int Dh, int box_cols,
                         int box_rows = 128, int box_n2 = 1) {
  EncodeTiledFn enc = encode_fn();
  if (!enc || n2_count <= 0 || N % n2_count) return MEANT_EUNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(v.p) & 15) return MEANT_EUNSUPPORTED;
  const cuuint64_t dims[5] = {(cuuint64_t)Dh, (cuuint64_t)H, (cuuint64_t)S, (cuuint64_t)n2_count, (cuuint64_t)(N / n2_count)};
  int64_t st[4] = {v.sh * 2, v.ss * 2, v.sn2 * 2, v.sn1 * 2};          // bytes, dims 1..4
  cuuint64_t strides[4];
  for (int i = 0; i < 4; ++i) {
    if (dims[i + 1] == 1 && st[i] <= 0) st[i] = 16;                       // never stepped: any legal value
    if (st[i] <= 0 || st[i] % 16 || st[i] >= ((int64_t)1 << 40)) return MEANT_EUNSUPPORTED;
    strides[i] = (cuuint64_t)st[i];
  }
  if (box_rows <= 0 || box_rows > 256 || box_n2 <= 0 || box_n2 > 256 || box_rows * box_n2 > 256) return MEANT_EUNSUPPORTED;
  const cuuint32_t box[5] = {(cuuint32_t)box_cols, 1u, (cuuint32_t)box_rows, (cuuint32_t)box_n2, 1u};
  const cuuint32_t estr[5] = {1u, 1u, 1u, 1u, 1u};
  const CUtensorMapSwizzle sw = box_cols * 2 == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                              : box_cols * 2 == 64  ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(v.p), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_ERROR_INVALID_CONTEXT) {
    // a thread that has only used the runtime lazily (e.g. an autograd worker before its first launch) has no
    // current driver context yet: bind the device's primary context and encode again
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && cudaSetDevice(dev) == cudaSuccess)
      r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(v.p), dims, strides, box, estr,
              CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS && getenv("MEANT_DEBUG"))
    fprintf(stderr, "meant_b200: cuTensorMapEncodeTiled failed (%d): p=%p dims=(%llu,%llu,%llu,%llu,%llu) strides=(%llu,%llu,%llu,%llu) box=%d\n",
            (int)r, v.p, (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
            (unsigned long long)dims[3], (unsigned long long)dims[4], (unsigned long long)strides[0],
            (unsigned long long)strides[1], (unsigned long long)strides[2], (unsigned long long)strides[3], box_cols);
  return r == CUDA_SUCCESS ? 0 : MEANT_EUNSUPPORTED;
}

// Tensor map that lands a head tile in the unswizzled CHUNK layout of attn_tc.cu ([chunk][row][16 B], chunk stride
// 128 * 16 B): dimensions, fastest first, (8 elements of a chunk, row s, chunk of the whole row, n2, n1); box =
// 8 x 128 x Dh/8 x 1 x 1 at chunk coordinate h * Dh/8.  The heads of a row are contiguous (sh == Dh in every view the
// models build), so head and chunk fold into ONE dimension of H * Dh/8 chunks 16 bytes apart, which leaves room for
// the split sequence index n = n1 * n2_count + n2 of the TimeSformer groups.  One copy per operand tile; rows >= S
// zero-filled on load, dropped on store.  Views whose heads are not contiguous: MEANT_EUNSUPPORTED (cp.async path).
inline int make_chunk_map(CUtensorMap* m, const HeadView& v, int N, int n2_count, int H, int S, int Dh) {
  EncodeTiledFn enc = encode_fn();
  if (!enc || n2_count <= 0 || N % n2_count || Dh % 8 || v.sh != Dh) return MEANT_EUNSUPPORTED;
  if (reinterpret_cast<uintptr_t>(v.p) & 15) return MEANT_EUNSUPPORTED;
  const cuuint64_t dims[5] = {8u, (cuuint64_t)S, (cuuint64_t)(H * (Dh / 8)), (cuuint64_t)n2_count, (cuuint64_t)(N / n2_count)};
  int64_t st[4] = {v.ss * 2, 16, v.sn2 * 2, v.sn1 * 2};                 // bytes, dims 1..4
  cuuint64_t strides[4];
  for (int i = 0; i < 4; ++i) {
    if (dims[i + 1] == 1 && st[i] <= 0) st[i] = 16;
    if (st[i] <= 0 || st[i] % 16 || st[i] >= ((int64_t)1 << 40)) return MEANT_EUNSUPPORTED;
    strides[i] = (cuuint64_t)st[i];
  }
  const cuuint32_t box[5] = {8u, 128u, (cuuint32_t)(Dh / 8), 1u, 1u};
  const cuuint32_t estr[5] = {1u, 1u, 1u, 1u, 1u};
  CUresult r = CUDA_ERROR_UNKNOWN;
  for (int attempt = 0; attempt < 2 && r != CUDA_SUCCESS; ++attempt) {
    if (attempt == 1) {   // no current driver context on this thread yet (see make_head_map)
      int dev = 0;
      if (r != CUDA_ERROR_INVALID_CONTEXT || cudaGetDevice(&dev) != cudaSuccess || cudaSetDevice(dev) != cudaSuccess) break;
    }
    r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(v.p), dims, strides, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS && getenv("MEANT_DEBUG"))
    fprintf(stderr, "meant_b200: chunk map encode failed (%d): p=%p S=%d Dh=%d strides=(%lld,%lld,%lld,%lld)\n", (int)r, v.p, S,
            Dh, (long long)st[0], (long long)st[1], (long long)st[2], (long long)st[3]);
  return r == CUDA_SUCCESS ? 0 : MEANT_EUNSUPPORTED;
}

#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_addr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// arrive (count 1) and announce `bytes` of asynchronous traffic that will complete on this barrier
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(smem_addr(bar)), "r"(bytes) : "memory");
}
// box at coordinates (c0 = column, h, s0, n2, n1) -> shared memory; completes `box bytes` on bar
__device__ __forceinline__ void load_box(uint32_t dst_smem, const CUtensorMap* map, uint64_t* bar, int c0, int h, int s0,
                                         int n2, int n1) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6, %7}], [%2];\n"
      ::"r"(dst_smem), "l"(map), "r"(smem_addr(bar)), "r"(c0), "r"(h), "r"(s0), "r"(n2), "r"(n1)
      : "memory");
}
// shared memory -> box at the same coordinates (bulk async-group completion)
__device__ __forceinline__ void store_box(const CUtensorMap* map, uint32_t src_smem, int c0, int h, int s0, int n2, int n1) {
  asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];\n"
               ::"l"(map), "r"(src_smem), "r"(c0), "r"(h), "r"(s0), "r"(n2), "r"(n1)
               : "memory");
}
__device__ __forceinline__ void store_commit() { asm volatile("cp.async.bulk.commit_group;\n" ::: "memory"); }
// all but the newest N store groups have finished READING shared memory
template <int N>
__device__ __forceinline__ void store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;\n" ::"n"(N) : "memory"); }
__device__ __forceinline__ void store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;\n" ::: "memory"); }
__device__ __forceinline__ void prefetch_map(const CUtensorMap* map) {
  asm volatile("prefetch.tensormap [%0];\n" ::"l"(map) : "memory");
}
#endif

}  // namespace tma
}  // namespace meant
