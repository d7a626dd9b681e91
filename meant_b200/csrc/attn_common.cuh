// Argument block shared by the attention kernels (CUDA-core "exact" path and tcgen05 path).
//
// Semantics restated from the reference (citations relative to /root/reference):
//   scores = (Rq q)(Rk k)^T * scale  [+ causal -inf]  [+ (1 - key_mask) * -1e9]   meant/xPosAttention.py:39-56
//   out    = softmax(scores) v                                                       meant/xPosAttention.py:61-63
// Rq / Rk are the per-position pair rotations of meant/rotary_embedding_torch.py:37-44 with the
// xPos scale folded into the cos/sin tables (q: *scale, k: /scale; :108-109).  The tables are
// per element ([P][rot_dim]) because the reference lays the xPos scale out half-split while
// the angles are interleaved (:124-125 vs :142).
#pragma once
#include "common.cuh"

namespace meant {

// The argument structs are the public C ones (include/meant_b200.h).
using HeadView = ::meant_head_view;
using AttnArgs = ::meant_attn_args;

template <typename T>
__device__ __forceinline__ T* view_ptr(const HeadView& v, int n, int n2_count, int s, int h) {
  const int n1 = n / n2_count, n2 = n - n1 * n2_count;
  return static_cast<T*>(v.p) + (int64_t)n1 * v.sn1 + (int64_t)n2 * v.sn2 + (int64_t)s * v.ss + (int64_t)h * v.sh;
}

// rotate one interleaved pair (x0, x1) at feature offset d0 (even) with tables c/s at row `pos`
__device__ __forceinline__ void rot_pair(float& x0, float& x1, const float* c, const float* s, int pos,
                                         int rot_dim, int d0) {
  const float c0 = c[pos * rot_dim + d0], c1 = c[pos * rot_dim + d0 + 1];
  const float s0 = s[pos * rot_dim + d0], s1 = s[pos * rot_dim + d0 + 1];
  const float y0 = x0 * c0 - x1 * s0;
  const float y1 = x1 * c1 + x0 * s1;
  x0 = y0;
  x1 = y1;
}
// transpose of rot_pair: gradient w.r.t. the un-rotated pair
__device__ __forceinline__ void unrot_pair(float& g0, float& g1, const float* c, const float* s, int pos,
                                           int rot_dim, int d0) {
  const float c0 = c[pos * rot_dim + d0], c1 = c[pos * rot_dim + d0 + 1];
  const float s0 = s[pos * rot_dim + d0], s1 = s[pos * rot_dim + d0 + 1];
  const float y0 = g0 * c0 + g1 * s1;
  const float y1 = g1 * c1 - g0 * s0;
  g0 = y0;
  g1 = y1;
}

}  // namespace meant
