// The step after backward (src/in_loop_train.py:237-244,568: CrossEntropyLoss on the model's sigmoid outputs,
// clip_grad_norm_(1.0), AdamW): multi-tensor kernels over the model's ~300 parameter tensors.
//   meant_multi_sqnorm   sum of squares of every gradient (the clip norm), one atomic per CTA
//   meant_adamw_step     AdamW with the clip coefficient derived in-kernel from that sum (no pass that scales the
//                        gradients), decoupled weight decay, bias correction from a device-resident step counter
//                        (CUDA-graph replays advance it), and an optional bf16 copy of every updated parameter (the
//                        shadow the next forward's GEMMs read: no separate cast pass)
//   meant_prob_ce        mean cross entropy over rows of probabilities [B, C] with its gradient, one launch
// HBM-bound: AdamW reads p, g, m, v and writes p, m, v (+ 2 bytes of shadow) per element.
// A launch takes a BY-VALUE table of up to 48 tensors cut into 65536-element chunks (<= 320 chunks): the host side
// (meant_b200/optim.py) walks the parameter list and issues ceil(total / (320 * 65536)) launches.
#include "common.cuh"

namespace meant {

constexpr int OPT_CHUNK = MEANT_OPT_CHUNK;

// four consecutive gradients (fp32: one 16-byte load, bf16: one 8-byte load) / one gradient, as fp32
__device__ __forceinline__ float4 ld4(const float* g) { return __ldg(reinterpret_cast<const float4*>(g)); }
__device__ __forceinline__ float4 ld4(const __nv_bfloat16* g) {
  const uint2 pk = __ldg(reinterpret_cast<const uint2*>(g));
  const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
  const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}

// TG: gradient storage type (fp32 autograd tensors, or the bf16 slices of the data-parallel reducer's buckets)
template <typename TG>
__global__ void __launch_bounds__(256) multi_sqnorm_kernel(const __grid_constant__ meant_tensor_table tab, float* __restrict__ out) {
  const int ti = tab.block_tensor[blockIdx.x];
  const int64_t n = tab.numel[ti], base = (int64_t)tab.block_chunk[blockIdx.x] * OPT_CHUNK;
  const int64_t end = base + OPT_CHUNK < n ? base + OPT_CHUNK : n;
  const TG* __restrict__ g = static_cast<const TG*>(tab.g[ti]);
  float acc = 0.f;
  // vector loads when the chunk start is aligned (chunks are multiples of 4 elements; the tensor base may not be)
  if ((reinterpret_cast<uintptr_t>(g + base) & (4 * sizeof(TG) - 1)) == 0) {
    const int64_t n4 = (end - base) >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += 256) {
      const float4 v = ld4(g + base + (i << 2));
      acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    for (int64_t i = base + (n4 << 2) + threadIdx.x; i < end; i += 256) { const float x = to_f(g[i]); acc += x * x; }
  } else {
    for (int64_t i = base + threadIdx.x; i < end; i += 256) { const float x = to_f(g[i]); acc += x * x; }
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 8) {
    float v = red[threadIdx.x];
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
    if (threadIdx.x == 0) atomicAdd(out, v);
  }
}

struct AdamwHyper {
  float lr, beta1, beta2, eps, wd, max_norm;
};

template <bool VEC, typename TG>
__device__ __forceinline__ void adamw_range(float* __restrict__ p, const TG* __restrict__ g, float* __restrict__ m,
                                            float* __restrict__ v, __nv_bfloat16* __restrict__ sh, int64_t base, int64_t end,
                                            const AdamwHyper& hp, float coef, float bc1, float rbc2) {
  auto upd = [&](float& pp, float gg, float& mm, float& vv) {
    gg *= coef;
    pp *= 1.0f - hp.lr * hp.wd;                       // decoupled weight decay (torch.optim.AdamW)
    mm = hp.beta1 * mm + (1.0f - hp.beta1) * gg;
    vv = hp.beta2 * vv + (1.0f - hp.beta2) * gg * gg;
    const float denom = sqrtf(vv) * rbc2 + hp.eps;    // sqrt(v) / sqrt(1 - beta2^t) + eps
    pp -= (hp.lr / bc1) * (mm / denom);
  };
  if (VEC) {
    const int64_t n4 = (end - base) >> 2;
    for (int64_t i = threadIdx.x; i < n4; i += 256) {
      const int64_t o = base + (i << 2);
      float4 pp = *reinterpret_cast<const float4*>(p + o), mm = *reinterpret_cast<const float4*>(m + o),
             vv = *reinterpret_cast<const float4*>(v + o);
      const float4 gg = ld4(g + o);
      upd(pp.x, gg.x, mm.x, vv.x);
      upd(pp.y, gg.y, mm.y, vv.y);
      upd(pp.z, gg.z, mm.z, vv.z);
      upd(pp.w, gg.w, mm.w, vv.w);
      *reinterpret_cast<float4*>(p + o) = pp;
      *reinterpret_cast<float4*>(m + o) = mm;
      *reinterpret_cast<float4*>(v + o) = vv;
      if (sh) {
        const __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&lo);
        pk.y = *reinterpret_cast<const uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(sh + o) = pk;
      }
    }
    base += n4 << 2;
  }
  for (int64_t i = base + threadIdx.x; i < end; i += 256) {
    float pp = p[i], mm = m[i], vv = v[i];
    upd(pp, to_f(g[i]), mm, vv);
    p[i] = pp;
    m[i] = mm;
    v[i] = vv;
    if (sh) sh[i] = __float2bfloat16_rn(pp);
  }
}

template <typename TG>
__global__ void __launch_bounds__(256) adamw_kernel(const __grid_constant__ meant_tensor_table tab, const float* __restrict__ step,
                                                    const float* __restrict__ sumsq, const AdamwHyper hp) {
  const int ti = tab.block_tensor[blockIdx.x];
  const int64_t n = tab.numel[ti], base = (int64_t)tab.block_chunk[blockIdx.x] * OPT_CHUNK;
  const int64_t end = base + OPT_CHUNK < n ? base + OPT_CHUNK : n;
  float* p = static_cast<float*>(tab.p[ti]);
  const TG* g = static_cast<const TG*>(tab.g[ti]);
  float* m = static_cast<float*>(tab.m[ti]);
  float* v = static_cast<float*>(tab.v[ti]);
  __nv_bfloat16* sh = static_cast<__nv_bfloat16*>(tab.shadow[ti]);
  const float t = *step;
  const float bc1 = 1.0f - powf(hp.beta1, t), rbc2 = rsqrtf(1.0f - powf(hp.beta2, t));
  // torch.nn.utils.clip_grad_norm_: coef = min(1, max_norm / (total_norm + 1e-6))
  float coef = 1.0f;
  if (sumsq != nullptr && hp.max_norm > 0.f) coef = fminf(1.0f, hp.max_norm / (sqrtf(*sumsq) + 1e-6f));
  // every vector access of the fast path is naturally aligned: 16 bytes for fp32, 8 bytes for the bf16 streams
  const uintptr_t al = reinterpret_cast<uintptr_t>(p + base) | reinterpret_cast<uintptr_t>(m + base) |
                       reinterpret_cast<uintptr_t>(v + base) | (reinterpret_cast<uintptr_t>(g + base) * (16 / (4 * sizeof(TG)))) |
                       (sh ? reinterpret_cast<uintptr_t>(sh + base) << 1 : 0);
  if ((al & 15) == 0) adamw_range<true>(p, g, m, v, sh, base, end, hp, coef, bc1, rbc2);
  else adamw_range<false>(p, g, m, v, sh, base, end, hp, coef, bc1, rbc2);
}

// loss = mean_i ( log sum_c exp(p[i, c]) - p[i, y_i] ),  dp[i, c] = gscale * (softmax(p[i])_c - [c == y_i]) / B.
// One CTA; B * C is tiny (the classifier head: C = 2).
__global__ void __launch_bounds__(256) prob_ce_kernel(const float* __restrict__ p, const int64_t* __restrict__ y,
                                                      float* __restrict__ loss, float* __restrict__ dp, int B, int C, float gscale) {
  float acc = 0.f;
  for (int i = threadIdx.x; i < B; i += 256) {
    const float* row = p + (int64_t)i * C;
    float mx = row[0];
    for (int c = 1; c < C; ++c) mx = fmaxf(mx, row[c]);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += expf(row[c] - mx);
    const int64_t yi = y[i];
    const bool ok = yi >= 0 && yi < C;
    acc += ok ? (mx + logf(s) - row[yi]) : 0.f;
    if (dp != nullptr)
      for (int c = 0; c < C; ++c) dp[(int64_t)i * C + c] = ok ? gscale * (expf(row[c] - mx) / s - (c == yi ? 1.f : 0.f)) / B : 0.f;
  }
  __shared__ float red[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < 8; ++i) t += red[i];
    *loss = t / B;
  }
}

static bool table_ok(const meant_tensor_table* t) {
  if (!t || t->n_tensors <= 0 || t->n_tensors > MEANT_OPT_MAX_TENSORS || t->n_blocks <= 0 || t->n_blocks > MEANT_OPT_MAX_BLOCKS)
    return false;
  for (int i = 0; i < t->n_blocks; ++i) {
    const int ti = t->block_tensor[i];
    if (ti >= t->n_tensors || t->block_chunk[i] < 0 || (int64_t)t->block_chunk[i] * OPT_CHUNK >= t->numel[ti]) return false;
  }
  return true;
}

}  // namespace meant

using namespace meant;

extern "C" int meant_multi_sqnorm(const meant_tensor_table* table, float* sumsq, meant_stream_t stream) {
  if (!table_ok(table) || !sumsq) return MEANT_EARG;
  for (int i = 0; i < table->n_tensors; ++i)
    if (!table->g[i] || table->numel[i] <= 0) return MEANT_EARG;
  if (table->g_bf16) multi_sqnorm_kernel<__nv_bfloat16><<<table->n_blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*table, sumsq);
  else multi_sqnorm_kernel<float><<<table->n_blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*table, sumsq);
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_adamw_step(const meant_tensor_table* table, const float* step, const float* sumsq, float lr, float beta1,
                                float beta2, float eps, float weight_decay, float max_norm, meant_stream_t stream) {
  if (!table_ok(table) || !step || lr < 0.f || beta1 < 0.f || beta1 >= 1.f || beta2 < 0.f || beta2 >= 1.f || eps <= 0.f)
    return MEANT_EARG;
  for (int i = 0; i < table->n_tensors; ++i)
    if (!table->p[i] || !table->g[i] || !table->m[i] || !table->v[i] || table->numel[i] <= 0) return MEANT_EARG;
  const AdamwHyper hp{lr, beta1, beta2, eps, weight_decay, max_norm};
  if (table->g_bf16) adamw_kernel<__nv_bfloat16><<<table->n_blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*table, step, sumsq, hp);
  else adamw_kernel<float><<<table->n_blocks, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(*table, step, sumsq, hp);
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_prob_ce(const float* probs, const int64_t* target, float* loss, float* dprobs, int B, int C,
                             float grad_scale, meant_stream_t stream) {
  if (!probs || !target || !loss || B <= 0 || C <= 0) return MEANT_EARG;
  prob_ce_kernel<<<1, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(probs, target, loss, dprobs, B, C, grad_scale);
  MEANT_LAUNCH_CHECK();
  return 0;
}
