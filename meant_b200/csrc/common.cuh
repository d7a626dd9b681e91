// Shared device helpers for the MEANT B200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include "../../include/meant_b200.h"

#define MEANT_DT_F32 0
#define MEANT_DT_BF16 1

// C-ABI error codes (include/meant_b200.h): 0 ok, >0 cudaError_t, <0 argument error.
#define MEANT_EARG (-1)
#define MEANT_ETOOBIG (-3)
#define MEANT_EUNSUPPORTED (-2)

extern "C" void meant_count_launch_(int n);

#define MEANT_LAUNCH_CHECK()                          \
  do {                                                \
    meant_count_launch_(1);                           \
    cudaError_t e_ = cudaGetLastError();              \
    if (e_ != cudaSuccess) return (int)e_;            \
  } while (0)

namespace meant {

__device__ __forceinline__ float to_f(float v) { return v; }
__device__ __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }
template <> __device__ __forceinline__ __nv_bfloat16 from_f<__nv_bfloat16>(float v) {
  return __float2bfloat16_rn(v);
}

// 16-byte vector of T: 4 floats or 8 bf16.
template <typename T> struct Vec16;
template <> struct Vec16<float> {
  static constexpr int N = 4;
  float4 raw;
  __device__ __forceinline__ void load(const float* p) { raw = *reinterpret_cast<const float4*>(p); }
  __device__ __forceinline__ void store(float* p) const { *reinterpret_cast<float4*>(p) = raw; }
  __device__ __forceinline__ void unpack(float* f) const { f[0] = raw.x; f[1] = raw.y; f[2] = raw.z; f[3] = raw.w; }
  __device__ __forceinline__ void pack(const float* f) { raw = make_float4(f[0], f[1], f[2], f[3]); }
};
template <> struct Vec16<__nv_bfloat16> {
  static constexpr int N = 8;
  uint4 raw;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { raw = *reinterpret_cast<const uint4*>(p); }
  __device__ __forceinline__ void store(__nv_bfloat16* p) const { *reinterpret_cast<uint4*>(p) = raw; }
  __device__ __forceinline__ void unpack(float* f) const {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 t = __bfloat1622float2(h[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
  }
  __device__ __forceinline__ void pack(const float* f) {
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&raw);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// GELU of nn.GELU() (exact erf form, meant/meant.py:64,107) and its derivative.  erf comes from
// Abramowitz-Stegun 7.1.26 (|error| <= 1.5e-7, i.e. fp32 round-off level) on the SFU: one ex2 and one
// rcp per element instead of erff's ~40 instructions, which made the GELU-fused norm kernels
// issue-bound rather than HBM-bound.  exp(-x^2/2) is shared between the erf tail and the Gaussian pdf.
// tail = (1 - erf(|x| / sqrt 2)) / 2 and e = exp(-x^2 / 2): the 1/sqrt 2 and the 1/2 are folded into the constants.
__device__ __forceinline__ void gelu_tail(float x, float& tail, float& e) {
  float t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * x * -0.72134752044448170f));               // exp(-x^2/2)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(fmaf(fabsf(x), 0.3275911f * 0.70710678118654752f, 1.0f)));
  float poly = fmaf(t, 0.5f * 1.061405429f, 0.5f * -1.453152027f);
  poly = fmaf(poly, t, 0.5f * 1.421413741f);
  poly = fmaf(poly, t, 0.5f * -0.284496736f);
  poly = fmaf(poly, t, 0.5f * 0.254829592f);
  tail = poly * t * e;
}
__device__ __forceinline__ void gelu_parts(float x, float& cdf, float& pdf) {
  float tail, e;
  gelu_tail(x, tail, e);
  cdf = x >= 0.f ? 1.0f - tail : tail;
  pdf = 0.39894228040143268f * e;
}
// x * cdf(x) = max(x, 0) - |x| * tail
__device__ __forceinline__ float gelu_f(float x) {
  float tail, e;
  gelu_tail(x, tail, e);
  return fmaf(-fabsf(x), tail, fmaxf(x, 0.f));
}
__device__ __forceinline__ float gelu_grad_f(float x) {
  float cdf, pdf;
  gelu_parts(x, cdf, pdf);
  return fmaf(x, pdf, cdf);
}
// activation and derivative together (the GELU-fused norm backward)
__device__ __forceinline__ void gelu_val_grad(float x, float& val, float& grad) {
  float tail, e;
  gelu_tail(x, tail, e);
  val = fmaf(-fabsf(x), tail, fmaxf(x, 0.f));
  const float cdf = x >= 0.f ? 1.0f - tail : tail;
  grad = fmaf(x * 0.39894228040143268f, e, cdf);
}

// Counter-based dropout mask, regenerated in backward: ONE 32-bit hash per FOUR consecutive elements (quad q = element
// index / 4), 8 bits each; an element is kept iff its byte >= thr8 = round(256 p).  The drop probability is therefore
// realised as thr8 / 256 (exact for p = 0.5; 0.1 becomes 26/256) and the survivors are scaled by 256 / (256 - thr8), so
// the expectation is unbiased.  Per element this costs ~2 integer instructions for the hash and 2 for the select (a
// byte-wise SIMD compare yields all four keep masks at once), against ~14 for the 16-bit-per-element scheme it replaces:
// the GELU + dropout norm kernels were bound by exactly this integer work.
__device__ __forceinline__ uint32_t mix32(uint32_t h) {
  h ^= h >> 16;
  h *= 0x85EBCA6Bu;
  h ^= h >> 13;
  h *= 0xC2B2AE35u;
  h ^= h >> 16;
  return h;
}
// thr4: thr8 replicated into the four bytes (drop_thr4 on the host); returns 0xff in byte j iff element 4 q + j is kept
__device__ __forceinline__ uint32_t drop_mask4(uint64_t seed, uint64_t quad, uint32_t thr4) {
  const uint32_t h = mix32(((uint32_t)quad ^ (uint32_t)seed) + ((uint32_t)(quad >> 32) * 0x9E3779B1u + (uint32_t)(seed >> 32)));
  return __vcmpgeu4(h, thr4);
}
// v if element j of the quad is kept, else +0 (a bit-wise AND with the byte mask spread over 32 bits)
__device__ __forceinline__ float drop_sel(float v, uint32_t mask4, int j) {
  return __uint_as_float(__float_as_uint(v) & __byte_perm(mask4, 0u, 0x1111u * (uint32_t)j));
}
// scalar form for the generic (any d) kernels: multiplier of element idx
__device__ __forceinline__ float drop_keep(uint64_t seed, uint64_t idx, uint32_t thr4, float inv_keep) {
  const uint32_t m = drop_mask4(seed, idx >> 2, thr4);
  return ((m >> (8 * (int)(idx & 3))) & 1u) ? inv_keep : 0.0f;
}

}  // namespace meant
