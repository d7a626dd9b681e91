// Load-path micro-benchmark for the attention tile loader (B200): how fast can ONE persistent CTA per
// SM pull 128-row x 192-byte head tiles (row pitch 4608 B, the packed qkv layout) into shared memory?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldbench tools/ldbench.cu && ./ldbench
// Variants: LDG.128 -> STS (threads, loads in flight), cp.async 16 B (scattered / row-contiguous dst),
// cp.async.bulk per 192-byte row, cp.async.bulk of a contiguous 24 KB tile.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <vector>

constexpr int ROWS = 128, ROWB = 192, PITCH = 4608, TILE = ROWS * ROWB;
constexpr uint32_t CS = 128 * 16 + 16;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t par) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0,1,0,p;\n\t}" : "=r"(ok) : "r"(smem_u32(b)), "r"(par) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes),
               "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ uint4 ldg16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ void cpa16(uint32_t dst, const void* src) { asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory"); }

// tile index -> base pointer (tile t covers head t%24 of row block t/24: 24 x 192 B = 4608 B per row)
__device__ __forceinline__ const uint8_t* tile_base(const uint8_t* g, int t, int contiguous) {
  if (contiguous) return g + (size_t)t * TILE;
  return g + (size_t)(t / 24) * ROWS * PITCH + (t % 24) * ROWB;
}

// mode 0: LDG -> STS chunk layout, NB tiles of 12 loads in flight per thread (128-thread groups)
template <int NB>
__global__ void k_ldg(const uint8_t* g, int ntiles, int* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int nthr = blockDim.x, t = threadIdx.x;
  const int per = ROWS * 12 / nthr;   // chunks per thread per tile (nthr = 128 -> 12, 256 -> 6, 512 -> 3)
  uint4 buf[NB][12];
  int acc = 0;
  int my = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x * NB) {
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int tl = tile + b * gridDim.x;
      if (tl < ntiles) {
        const uint8_t* base = tile_base(g, tl, 0);
#pragma unroll
        for (int k = 0; k < 12; ++k)
          if (k < per) {
            const int idx = t + k * nthr, r = idx / 12, ch = idx % 12;
            buf[b][k] = ldg16(base + (size_t)r * PITCH + ch * 16);
          }
      }
    }
#pragma unroll
    for (int b = 0; b < NB; ++b) {
      const int tl = tile + b * gridDim.x;
      if (tl < ntiles) {
#pragma unroll
        for (int k = 0; k < 12; ++k)
          if (k < per) {
            const int idx = t + k * nthr, r = idx / 12, ch = idx % 12;
            *reinterpret_cast<uint4*>(smem + ((my + b) & 3) * 12 * CS + ch * CS + r * 16) = buf[b][k];
            acc ^= buf[b][k].x;
          }
      }
    }
    my += NB;
  }
  if (acc == 0x12345678) *sink = acc;
}

// mode 1: cp.async 16 B, 4-stage ring; scatter = chunk layout destination, else row-major destination
__global__ void k_cpasync(const uint8_t* g, int ntiles, int scatter, int* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  const int nthr = blockDim.x, t = threadIdx.x;
  const int per = ROWS * 12 / nthr;
  const uint32_t s0 = smem_u32(smem);
  int j = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++j) {
    const uint8_t* base = tile_base(g, tile, 0);
    const uint32_t sb = s0 + (j & 3) * 12 * CS;
    for (int k = 0; k < per; ++k) {
      const int idx = t + k * nthr, r = idx / 12, ch = idx % 12;
      cpa16(sb + (scatter ? ch * CS + r * 16 : r * ROWB + ch * 16), base + (size_t)r * PITCH + ch * 16);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 3;" ::: "memory");
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (smem[t] == 0x7f && smem[t + 1] == 0x11) *sink = 1;
}

// mode 1b: the attention loader's pattern: 8 lanes fill one 128-byte core matrix (rows 8 rg + i, chunk 4 q + g),
// zero-fill form, cp.async.mbarrier.arrive.noinc per stage, 3-stage ring; `spin` extra warps poll an mbarrier
// the way waiting compute warps do.
__global__ void k_cpasync_cm(const uint8_t* g, int ntiles, int spin_mode, int* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar[4], never;
  const int t = threadIdx.x;
  if (t == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 128);
    mbar_init(&never, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (t >= 128) {   // spinning "compute" warps
    if (spin_mode == 1) {
      const long long t0 = clock64();
      uint32_t ok = 0;
      while (!ok) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0,1,0,p;\n\t}" : "=r"(ok) : "r"(smem_u32(&never)), "r"(0) : "memory");
        if (clock64() - t0 > 4000000000LL) break;
        if (*(volatile int*)sink == 77) break;   // loader done
      }
    }
    return;
  }
  const uint32_t CSN = 2048;
  const int w = t >> 5, gq = (t >> 3) & 3, i = t & 7;
  const uint32_t s0 = smem_u32(smem);
  int j = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++j) {
    const uint8_t* base = tile_base(g, tile, 0);
    const uint32_t sb = s0 + (j % 3) * 12 * CSN;
    if (j >= 3) mbar_wait(&bar[j % 3], ((j - 3) / 3) & 1);
#pragma unroll
    for (int k = 0; k < 12; ++k) {
      const int u = k * 4 + w, rg = u / 3, cq = u - rg * 3, r = rg * 8 + i, ch = cq * 4 + gq;
      const uint32_t sz = r < 128 ? 16u : 0u;
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sb + ch * CSN + r * 16), "l"(base + (size_t)r * PITCH + ch * 16), "r"(sz) : "memory");
    }
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&bar[j % 3])) : "memory");
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
  if (t == 0 && blockIdx.x == 0 && spin_mode) *(volatile int*)sink = 77;
}

// mode 1c: cp.async 16 B into UMMA-ready swizzled tiles: columns 0-63 -> [128 rows x 128 B] SWIZZLE_128B,
// columns 64-95 -> [128 rows x 64 B] SWIZZLE_64B; lanes walk the 12 chunks of a row (row-contiguous global reads).
__global__ void k_cpasync_sw(const uint8_t* g, int ntiles, int* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int nthr = blockDim.x, t = threadIdx.x;
  const int per = ROWS * 12 / nthr;
  const uint32_t s0 = smem_u32(smem);
  int j = 0;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++j) {
    const uint8_t* base = tile_base(g, tile, 0);
    const uint32_t sb = s0 + (j & 3) * 24576;
    for (int k = 0; k < per; ++k) {
      const int idx = t + k * nthr, r = idx / 12, ch = idx % 12;
      const uint32_t off = ch < 8 ? r * 128 + ((ch ^ (r & 7)) << 4) : 16384 + r * 64 + ((((ch - 8) ^ ((r >> 1) & 3))) << 4);
      cpa16(sb + off, base + (size_t)r * PITCH + ch * 16);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 3;" ::: "memory");
  }
  asm volatile("cp.async.wait_group 0;" ::: "memory");
  if (smem[t] == 0x7f && smem[t + 1] == 0x11) *sink = 1;
}

// mode 2: cp.async.bulk, rows of 192 B (contiguous = 0) or whole contiguous tiles (contiguous = 1), 4-stage ring
__global__ void k_bulk(const uint8_t* g, int ntiles, int contiguous, int* sink) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint64_t bar[4];
  const int t = threadIdx.x;
  if (t == 0) {
    for (int i = 0; i < 4; ++i) mbar_init(&bar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int j = 0;
  const int first = blockIdx.x;
  // prologue: 3 tiles in flight
  auto issue = [&](int tile, int jj) {
    const uint8_t* base = tile_base(g, tile, contiguous);
    uint8_t* dst = smem + (jj & 3) * TILE;
    if (contiguous) {
      if (t == 0) { mbar_expect(&bar[jj & 3], TILE); bulk_g2s(dst, base, TILE, &bar[jj & 3]); }
    } else {
      if (t == 0) mbar_expect(&bar[jj & 3], TILE);
      __syncwarp();
      if (t < 32)
        for (int r = t; r < ROWS; r += 32) bulk_g2s(dst + r * ROWB, base + (size_t)r * PITCH, ROWB, &bar[jj & 3]);
    }
  };
  for (int p = 0; p < 3; ++p)
    if (first + p * (int)gridDim.x < ntiles && t < 32) issue(first + p * gridDim.x, p);
  for (int tile = first; tile < ntiles; tile += gridDim.x, ++j) {
    const int nxt = tile + 3 * gridDim.x;
    if (t < 32) {
      if (nxt < ntiles) issue(nxt, j + 3);
      mbar_wait(&bar[j & 3], (j >> 2) & 1);
    }
    __syncthreads();   // consumers would read the stage here
  }
  if (smem[t] == 0x7f && smem[t + 1] == 0x11) *sink = 1;
}

template <typename F>
static void run(const char* name, F launch, int ntiles) {
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  launch();
  cudaDeviceSynchronize();
  cudaError_t err = cudaGetLastError();
  if (err != cudaSuccess) { printf("%-44s ERROR %s\n", name, cudaGetErrorString(err)); return; }
  float best = 1e9f;
  for (int i = 0; i < 5; ++i) {
    cudaEventRecord(e0);
    launch();
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    if (ms < best) best = ms;
  }
  printf("%-44s %8.1f us  %7.0f GB/s\n", name, best * 1e3, (double)ntiles * TILE / best / 1e6);
}

int main() {
  const int rowblocks = 1024;                       // 1024 x 128 rows x 4608 B = 604 MB
  const int ntiles = rowblocks * 24;
  uint8_t* g;
  int* sink;
  cudaMalloc(&g, (size_t)rowblocks * ROWS * PITCH);
  cudaMalloc(&sink, 4);
  cudaMemset(g, 1, (size_t)rowblocks * ROWS * PITCH);
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int smem = 4 * 12 * CS + 1024;
  cudaFuncSetAttribute(k_ldg<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_ldg<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_cpasync, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_bulk, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * TILE + 1024);
  for (int nthr : {128, 256, 512}) {
    char nm[96];
    snprintf(nm, 96, "ldg->sts %d thr, 1 tile in flight", nthr);
    run(nm, [&] { k_ldg<1><<<sms, nthr, smem>>>(g, ntiles, sink); }, ntiles);
    snprintf(nm, 96, "ldg->sts %d thr, 2 tiles in flight", nthr);
    run(nm, [&] { k_ldg<2><<<sms, nthr, smem>>>(g, ntiles, sink); }, ntiles);
  }
  for (int nthr : {128, 256, 512})
    for (int sc : {1, 0}) {
      char nm[96];
      snprintf(nm, 96, "cp.async16 %d thr, %s dst, 4 stages", nthr, sc ? "chunk-layout" : "row-major");
      run(nm, [&] { k_cpasync<<<sms, nthr, smem>>>(g, ntiles, sc, sink); }, ntiles);
    }
  cudaFuncSetAttribute(k_cpasync_cm, cudaFuncAttributeMaxDynamicSharedMemorySize, 3 * 12 * 2048 + 1024);
  run("cp.async16 core-matrix map, 128 thr", [&] { k_cpasync_cm<<<sms, 128, 3 * 12 * 2048 + 1024>>>(g, ntiles, 0, sink); }, ntiles);
  run("cp.async16 core-matrix map + 256 idle thr", [&] { k_cpasync_cm<<<sms, 384, 3 * 12 * 2048 + 1024>>>(g, ntiles, 0, sink); }, ntiles);
  run("cp.async16 core-matrix map + 256 spinning", [&] { cudaMemset(sink, 0, 4); k_cpasync_cm<<<sms, 384, 3 * 12 * 2048 + 1024>>>(g, ntiles, 1, sink); }, ntiles);
  cudaFuncSetAttribute(k_cpasync_sw, cudaFuncAttributeMaxDynamicSharedMemorySize, 4 * 24576 + 1024);
  for (int nthr : {128, 256})  {
    char nm[96];
    snprintf(nm, 96, "cp.async16 %d thr, SW128|SW64 tiles", nthr);
    run(nm, [&] { k_cpasync_sw<<<sms, nthr, 4 * 24576 + 1024>>>(g, ntiles, sink); }, ntiles);
  }
  run("cp.async.bulk 192 B rows, 4 stages", [&] { k_bulk<<<sms, 128, 4 * TILE + 1024>>>(g, ntiles, 0, sink); }, ntiles);
  run("cp.async.bulk 24 KB contiguous, 4 stages", [&] { k_bulk<<<sms, 128, 4 * TILE + 1024>>>(g, ntiles, 1, sink); }, ntiles);
  run("cp.async.bulk 192 B rows, 2 CTA/SM", [&] { k_bulk<<<2 * sms, 128, 4 * TILE + 1024>>>(g, ntiles, 0, sink); }, ntiles);
  return 0;
}
