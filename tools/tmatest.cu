// TMA check + load-path benchmark for the attention operand tiles (B200):
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I. -o build/tmatest tools/tmatest.cu && build/tmatest
// 1. correctness: a head tile of a packed [N, S, 3*dim] projection (S = 100 < 128) is TMA-loaded as
//    [64 columns, SWIZZLE_128B | 32 columns, SWIZZLE_64B], dumped, compared with the layout the UMMA
//    descriptors of attn_tc2.cu expect (sw128_off / sw64_off), then TMA-stored into a second tensor.
// 2. bandwidth: one persistent CTA per SM streams Q|K|V tiles (72 KB per item, 3 stages) with nothing else.
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <cuda_bf16.h>
#include "meant_b200/csrc/tma_utils.cuh"

using namespace meant;

__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(tma::smem_addr(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0,1,0,p;\n\t}" : "=r"(ok) : "r"(tma::smem_addr(b)), "r"(par) : "memory");
}

struct Maps { CUtensorMap a, b, oa, ob; };
constexpr uint32_t SUB_A = 128 * 128, SUB_B = 128 * 64, OPER = SUB_A + SUB_B;

__global__ void check_kernel(const __grid_constant__ Maps m, uint8_t* dump, int n, int h) {
  extern __shared__ __align__(128) uint8_t raw[];
  uint8_t* smem = raw + ((1024u - (tma::smem_addr(raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  if (threadIdx.x == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  if (threadIdx.x == 0) {
    tma::mbar_expect_tx(&bar, OPER);
    tma::load_box(tma::smem_addr(smem), &m.a, &bar, 0, h, 0, 0, n);
    tma::load_box(tma::smem_addr(smem) + SUB_A, &m.b, &bar, 64, h, 0, 0, n);
  }
  mbar_wait(&bar, 0);
  for (uint32_t i = threadIdx.x; i < OPER; i += blockDim.x) dump[i] = smem[i];
  __syncthreads();
  if (threadIdx.x == 0) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    tma::store_box(&m.oa, tma::smem_addr(smem), 0, h, 0, 0, n);
    tma::store_box(&m.ob, tma::smem_addr(smem) + SUB_A, 64, h, 0, 0, n);
    tma::store_commit();
    tma::store_wait_all();
  }
}

struct Maps6 { CUtensorMap m[6]; };
__global__ void __launch_bounds__(128, 1) stream_kernel(const __grid_constant__ Maps6 m, int items, int H, int* sink) {
  extern __shared__ __align__(128) uint8_t raw[];
  uint8_t* smem = raw + ((1024u - (tma::smem_addr(raw) & 1023u)) & 1023u);
  __shared__ uint64_t full[3], empty[3];
  if (threadIdx.x == 0) {
    for (int i = 0; i < 3; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 96); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int n_my = (items - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  if (threadIdx.x < 32) {
    if (threadIdx.x == 0) {
      for (int j = 0; j < n_my; ++j) {
        const int item = blockIdx.x + j * gridDim.x, n = item / H, h = item % H, st = j % 3;
        if (j >= 3) mbar_wait(&empty[st], ((j - 3) / 3) & 1);
        tma::mbar_expect_tx(&full[st], 3 * OPER);
        const uint32_t sb = tma::smem_addr(smem) + st * 3 * OPER;
        for (int t = 0; t < 3; ++t) {
          tma::load_box(sb + t * OPER, &m.m[2 * t], &full[st], 0, h, 0, 0, n);
          tma::load_box(sb + t * OPER + SUB_A, &m.m[2 * t + 1], &full[st], 64, h, 0, 0, n);
        }
      }
    }
  } else {
    int acc = 0;
    for (int j = 0; j < n_my; ++j) {
      const int st = j % 3;
      mbar_wait(&full[st], (j / 3) & 1);
      acc += smem[st * 3 * OPER + threadIdx.x * 64];
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tma::smem_addr(&empty[st])) : "memory");
    }
    if (acc == 0x7fffffff) *sink = acc;
  }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at line %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)

int main() {
  const int N = 768, S = 100, H = 8, DH = 96, DIM = H * DH, W = 3 * DIM;
  const size_t elems = (size_t)N * S * W;
  std::vector<__nv_bfloat16> host(elems);
  for (size_t i = 0; i < elems; ++i) host[i] = __float2bfloat16((float)((i * 2654435761u >> 7) % 2048) - 1024.f);
  __nv_bfloat16 *src, *dst;
  CK(cudaMalloc(&src, elems * 2));
  CK(cudaMalloc(&dst, elems * 2));
  CK(cudaMemcpy(src, host.data(), elems * 2, cudaMemcpyHostToDevice));
  CK(cudaMemset(dst, 0, elems * 2));
  // the k view of the packed projection
  HeadView kv{src + DIM, (int64_t)S * W, 0, W, DH}, ov{dst + DIM, (int64_t)S * W, 0, W, DH};
  Maps m;
  int rc = tma::make_head_map(&m.a, kv, N, 1, H, S, DH, 64) | tma::make_head_map(&m.b, kv, N, 1, H, S, DH, 32) |
           tma::make_head_map(&m.oa, ov, N, 1, H, S, DH, 64) | tma::make_head_map(&m.ob, ov, N, 1, H, S, DH, 32);
  printf("make_head_map rc=%d\n", rc);
  if (rc) return 1;
  uint8_t* dump;
  CK(cudaMalloc(&dump, OPER));
  CK(cudaFuncSetAttribute(check_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, OPER + 1024));
  const int n = 5, h = 3;
  check_kernel<<<1, 128, OPER + 1024>>>(m, dump, n, h);
  CK(cudaDeviceSynchronize());
  std::vector<uint8_t> d(OPER);
  CK(cudaMemcpy(d.data(), dump, OPER, cudaMemcpyDeviceToHost));
  long bad = 0;
  for (int r = 0; r < 128; ++r)
    for (int c = 0; c < 12; ++c) {
      const uint32_t off = c < 8 ? r * 128 + ((c ^ (r & 7)) << 4) : SUB_A + r * 64 + (((c - 8) ^ ((r >> 1) & 3)) << 4);
      for (int e = 0; e < 8; ++e) {
        const __nv_bfloat16 got = reinterpret_cast<const __nv_bfloat16*>(d.data() + off)[e];
        const float want = r < S ? __bfloat162float(host[((size_t)n * S + r) * W + DIM + h * DH + c * 8 + e]) : 0.f;
        if (__bfloat162float(got) != want) ++bad;
      }
    }
  printf("load layout mismatches: %ld\n", bad);
  std::vector<__nv_bfloat16> back(elems);
  CK(cudaMemcpy(back.data(), dst, elems * 2, cudaMemcpyDeviceToHost));
  long bad2 = 0, nz = 0;
  for (size_t i = 0; i < elems; ++i) {
    const size_t col = i % W, row = (i / W) % S, nn = i / ((size_t)W * S);
    const bool in = nn == (size_t)n && col >= (size_t)(DIM + h * DH) && col < (size_t)(DIM + (h + 1) * DH) && row < (size_t)S;
    const float want = in ? __bfloat162float(host[i]) : 0.f;
    if (__bfloat162float(back[i]) != want) ++bad2;
    if (__bfloat162float(back[i]) != 0.f) ++nz;
  }
  printf("store mismatches: %ld (nonzero %ld, expected about %d)\n", bad2, nz, S * DH);

  // ---- bandwidth (S = 128 view of a second buffer)
  const int S2 = 128;
  __nv_bfloat16* big;
  const size_t e2 = (size_t)N * S2 * W;
  CK(cudaMalloc(&big, e2 * 2));
  CK(cudaMemset(big, 1, e2 * 2));
  Maps6 m6;
  for (int t = 0; t < 3; ++t) {
    HeadView v{big + t * DIM, (int64_t)S2 * W, 0, W, DH};
    if (tma::make_head_map(&m6.m[2 * t], v, N, 1, H, S2, DH, 64) | tma::make_head_map(&m6.m[2 * t + 1], v, N, 1, H, S2, DH, 32)) return 2;
  }
  int* sink;
  CK(cudaMalloc(&sink, 4));
  const size_t sm = 9 * OPER + 1024;
  CK(cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sm));
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  for (int it = 0; it < 3; ++it) stream_kernel<<<148, 128, sm>>>(m6, N * H, H, sink);
  CK(cudaDeviceSynchronize());
  cudaEventRecord(e0);
  for (int it = 0; it < 10; ++it) stream_kernel<<<148, 128, sm>>>(m6, N * H, H, sink);
  cudaEventRecord(e1);
  CK(cudaDeviceSynchronize());
  float ms;
  cudaEventElapsedTime(&ms, e0, e1);
  printf("TMA stream: %.1f us per pass, %.0f GB/s\n", ms * 100, (double)N * H * 3 * OPER / (ms * 1e-4) / 1e9);
  return (bad || bad2) ? 1 : 0;
}
