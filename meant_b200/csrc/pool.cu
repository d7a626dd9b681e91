// Heads that sit on top of the encoder stacks in the reference's `src` tree and pre-trainers.
//
// 1. Softmax-weighted pooling (src/meant/meant_tweet.py:173,259-261; src/meant/meant_timesformer.py:274-279,335-345):
//      w = Softmax(dim=2)( Linear(d,1)( GELU( LayerNorm( Linear(d,d)(x) ) ) ) ),   pooled = x^T w
//    The first Linear is a GEMM (h = Linear(d,d)(x)); everything after it is one kernel per direction here: per
//    sequence, LayerNorm + exact GELU + the d -> 1 dot per token, the softmax over the tokens and the weighted sum of
//    the x rows -- h and x are read once, nothing of size [tokens, d] is written in the forward.
//    `degenerate`: the reference's image_prep applies Softmax(dim=2) to a [B, tokens, 1] tensor, i.e. over an axis of
//    length ONE -- every weight is exactly 1 and the "pooling" is a plain sum over the tokens (reproduced, not fixed).
// 2. Token cross entropy for the MLM head (pretrain_mlm.py:154,182: nn.CrossEntropyLoss over [tokens, vocab] logits,
//    ignore_index -100): one row per warp group, online log-sum-exp, forward loss and in-place-able gradient.
// HBM-bound kernels: 16-byte loads, a warp per token row, fp32 math.
#include "common.cuh"
#include <math_constants.h>

namespace meant {

constexpr int PW = 8;          // warps per CTA
// d = 32 E with E = 8, 16, 24 or 32 elements per lane (d = 256 .. 1024; the models use 768)

// this lane's E elements of a row of d = 32 E: 16-byte vectors lane, lane + 32, ... (fully unrolled: registers)
template <typename T, int E>
__device__ __forceinline__ void load_row(const T* __restrict__ p, int lane, float* f) {
  constexpr int VN = Vec16<T>::N;
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv) {
    Vec16<T> x;
    x.load(p + (lane + 32 * nv) * VN);
    x.unpack(f + nv * VN);
  }
}
template <typename T, int E>
__device__ __forceinline__ void store_row(T* __restrict__ p, int lane, const float* f) {
  constexpr int VN = Vec16<T>::N;
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv) {
    Vec16<T> x;
    x.pack(f + nv * VN);
    x.store(p + (lane + 32 * nv) * VN);
  }
}
// column index of element e of this lane's vector nv
template <typename T>
__device__ __forceinline__ int col_of(int lane, int nv, int e) { return (lane + 32 * nv) * Vec16<T>::N + e; }

struct PoolArgs {
  const void* x;        // [N, S, d] the tokens that are pooled
  const void* h;        // [N, S, d] Linear(d, d)(x)
  const float* ln_w;    // [d] LayerNorm weight / bias
  const float* ln_b;
  const float* w2;      // [d] Linear(d, 1) weight
  float b2;             //     and bias (shifts every score equally: no effect on the softmax)
  float ln_eps;
  void* out;            // [N, d]
  float* prob;          // [N, S] softmax weights, kept for the backward pass
  int64_t N;
  int S, d, degenerate;
};

// LayerNorm + GELU + dot of one row held in registers; returns the score (all lanes)
template <typename T, int E>
__device__ __forceinline__ float row_score(const float* hrow, int lane, const PoolArgs& a, float& mean, float& rstd) {
  constexpr int VN = Vec16<T>::N, d = 32 * E;
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < E; ++i) s += hrow[i];
  mean = warp_sum(s) / d;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < E; ++i) { const float c = hrow[i] - mean; q += c * c; }
  rstd = rsqrtf(warp_sum(q) / d + a.ln_eps);
  float dot = 0.f;
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv)
#pragma unroll
    for (int e = 0; e < VN; ++e) {
      const int c = col_of<T>(lane, nv, e);
      const float y = (hrow[nv * VN + e] - mean) * rstd * __ldg(a.ln_w + c) + __ldg(a.ln_b + c);
      dot = fmaf(gelu_f(y), __ldg(a.w2 + c), dot);
    }
  return warp_sum(dot) + a.b2;
}

template <typename T, int E>
__global__ void __launch_bounds__(PW * 32) softmax_pool_fwd_kernel(PoolArgs a) {
  extern __shared__ float sm[];                 // [S] scores -> weights | [PW][d] partial sums
  float* sc = sm;
  float* part = sm + ((a.S + 3) & ~3);
  __shared__ float red[PW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t n = blockIdx.x;
  const T* x = static_cast<const T*>(a.x) + n * a.S * a.d;
  const T* h = static_cast<const T*>(a.h) + n * a.S * a.d;
  constexpr int VN = Vec16<T>::N;
  if (!a.degenerate) {
    for (int t = warp; t < a.S; t += PW) {
      float hrow[E], mean, rstd;
      load_row<T, E>(h + (int64_t)t * a.d, lane, hrow);
      const float s = row_score<T, E>(hrow, lane, a, mean, rstd);
      if (lane == 0) sc[t] = s;
    }
    __syncthreads();
    float m = -CUDART_INF_F;
    for (int t = threadIdx.x; t < a.S; t += PW * 32) m = fmaxf(m, sc[t]);
    m = warp_max(m);
    if (lane == 0) red[warp] = m;
    __syncthreads();
    m = red[0];
    for (int w = 1; w < PW; ++w) m = fmaxf(m, red[w]);
    __syncthreads();
    float l = 0.f;
    for (int t = threadIdx.x; t < a.S; t += PW * 32) {
      const float e = expf(sc[t] - m);
      sc[t] = e;
      l += e;
    }
    l = warp_sum(l);
    if (lane == 0) red[warp] = l;
    __syncthreads();
    l = 0.f;
    for (int w = 0; w < PW; ++w) l += red[w];
    const float inv = 1.0f / l;
    for (int t = threadIdx.x; t < a.S; t += PW * 32) {
      const float p = sc[t] * inv;
      sc[t] = p;
      if (a.prob) a.prob[n * a.S + t] = p;
    }
    __syncthreads();
  }
  // pooled[c] = sum_t w_t x[t, c]: each warp sums its rows, then the warps' partial rows are added
  float acc[E];
#pragma unroll
  for (int i = 0; i < E; ++i) acc[i] = 0.f;
  for (int t = warp; t < a.S; t += PW) {
    float xr[E];
    load_row<T, E>(x + (int64_t)t * a.d, lane, xr);
    const float w = a.degenerate ? 1.0f : sc[t];
#pragma unroll
    for (int i = 0; i < E; ++i) acc[i] = fmaf(w, xr[i], acc[i]);
  }
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv)
#pragma unroll
    for (int e = 0; e < VN; ++e) part[warp * a.d + col_of<T>(lane, nv, e)] = acc[nv * VN + e];
  __syncthreads();
  T* out = static_cast<T*>(a.out) + n * a.d;
  for (int c = threadIdx.x; c < a.d; c += PW * 32) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < PW; ++w) s += part[w * a.d + c];
    out[c] = from_f<T>(s);
  }
}

struct PoolBwdArgs {
  PoolArgs f;           // forward arguments (x, h, parameters, prob)
  const void* dout;     // [N, d]
  void* dx;             // [N, S, d] gradient reaching x through the weighted sum (the Linear(d,d) path adds its own)
  void* dh;             // [N, S, d]
  float* dln_w;         // [d] accumulated atomically (zero on entry)
  float* dln_b;
  float* dw2;
};

template <typename T, int E>
__global__ void __launch_bounds__(PW * 32) softmax_pool_bwd_kernel(PoolBwdArgs b) {
  const PoolArgs& a = b.f;
  extern __shared__ float sm[];                 // [S] dp -> dscore | [d] dout | [3][PW][d] column partials
  float* ds = sm;
  float* go = sm + ((a.S + 3) & ~3);
  float* part = go + a.d;
  __shared__ float red[PW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t n = blockIdx.x;
  constexpr int VN = Vec16<T>::N;
  const T* x = static_cast<const T*>(a.x) + n * a.S * a.d;
  const T* h = static_cast<const T*>(a.h) + n * a.S * a.d;
  T* dx = static_cast<T*>(b.dx) + n * a.S * a.d;
  T* dh = static_cast<T*>(b.dh) + n * a.S * a.d;
  for (int c = threadIdx.x; c < a.d; c += PW * 32) go[c] = to_f(static_cast<const T*>(b.dout)[n * a.d + c]);
  __syncthreads();
  float gl[E];                                  // this lane's columns of dout
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv)
#pragma unroll
    for (int e = 0; e < VN; ++e) gl[nv * VN + e] = go[col_of<T>(lane, nv, e)];
  if (a.degenerate) {                           // weights are the constant 1: dx = dout, nothing flows into h
    float z[E];
#pragma unroll
    for (int i = 0; i < E; ++i) z[i] = 0.f;
    for (int t = warp; t < a.S; t += PW) {
      store_row<T, E>(dx + (int64_t)t * a.d, lane, gl);
      store_row<T, E>(dh + (int64_t)t * a.d, lane, z);
    }
    return;
  }
  const float* prob = a.prob + n * a.S;
  // dp_t = <dout, x_t>, then dscore_t = p_t (dp_t - sum_u p_u dp_u)
  float cpart = 0.f;
  for (int t = warp; t < a.S; t += PW) {
    float xr[E];
    load_row<T, E>(x + (int64_t)t * a.d, lane, xr);
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < E; ++i) dot = fmaf(gl[i], xr[i], dot);
    dot = warp_sum(dot);
    if (lane == 0) ds[t] = dot;
    cpart = fmaf(prob[t], dot, cpart);
  }
  if (lane == 0) red[warp] = cpart;
  __syncthreads();
  float ctot = 0.f;
  for (int w = 0; w < PW; ++w) ctot += red[w];
  __syncthreads();
  for (int t = threadIdx.x; t < a.S; t += PW * 32) ds[t] = prob[t] * (ds[t] - ctot);
  __syncthreads();
  float alw[E], alb[E], aw2[E];                 // per-column partials over this warp's rows
#pragma unroll
  for (int i = 0; i < E; ++i) alw[i] = alb[i] = aw2[i] = 0.f;
  for (int t = warp; t < a.S; t += PW) {
    const float p = prob[t], dst = ds[t];
    float tmp[E];
#pragma unroll
    for (int i = 0; i < E; ++i) tmp[i] = p * gl[i];
    store_row<T, E>(dx + (int64_t)t * a.d, lane, tmp);
    float hrow[E];
    load_row<T, E>(h + (int64_t)t * a.d, lane, hrow);
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < E; ++i) s += hrow[i];
    const float mean = warp_sum(s) / a.d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < E; ++i) { const float c = hrow[i] - mean; q += c * c; }
    const float rstd = rsqrtf(warp_sum(q) / a.d + a.ln_eps);
    float m1 = 0.f, m2 = 0.f;                   // sum dxhat, sum dxhat * xhat
#pragma unroll
    for (int nv = 0; nv < E / VN; ++nv)
#pragma unroll
      for (int e = 0; e < VN; ++e) {
        const int i = nv * VN + e, c = col_of<T>(lane, nv, e);
        const float xhat = (hrow[i] - mean) * rstd, lw = __ldg(a.ln_w + c);
        const float y = xhat * lw + __ldg(a.ln_b + c);
        float gv, gg;
        gelu_val_grad(y, gv, gg);
        aw2[i] = fmaf(dst, gv, aw2[i]);
        const float dy = dst * __ldg(a.w2 + c) * gg;
        alw[i] = fmaf(dy, xhat, alw[i]);
        alb[i] += dy;
        const float dxh = dy * lw;
        m1 += dxh;
        m2 = fmaf(dxh, xhat, m2);
        hrow[i] = xhat;                         // keep xhat, dxhat for the second sweep
        tmp[i] = dxh;
      }
    m1 = warp_sum(m1) / a.d;
    m2 = warp_sum(m2) / a.d;
#pragma unroll
    for (int i = 0; i < E; ++i) tmp[i] = rstd * (tmp[i] - m1 - hrow[i] * m2);
    store_row<T, E>(dh + (int64_t)t * a.d, lane, tmp);
  }
  // reduce the per-column partials over the warps, then one atomic per column per CTA
#pragma unroll
  for (int nv = 0; nv < E / VN; ++nv)
#pragma unroll
    for (int e = 0; e < VN; ++e) {
      const int c = col_of<T>(lane, nv, e), i = nv * VN + e;
      part[(0 * PW + warp) * a.d + c] = alw[i];
      part[(1 * PW + warp) * a.d + c] = alb[i];
      part[(2 * PW + warp) * a.d + c] = aw2[i];
    }
  __syncthreads();
  for (int c = threadIdx.x; c < a.d; c += PW * 32) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int w = 0; w < PW; ++w) {
      s0 += part[(0 * PW + w) * a.d + c];
      s1 += part[(1 * PW + w) * a.d + c];
      s2 += part[(2 * PW + w) * a.d + c];
    }
    atomicAdd(b.dln_w + c, s0);
    atomicAdd(b.dln_b + c, s1);
    atomicAdd(b.dw2 + c, s2);
  }
}

// ------------------------------------------------------------------------------------------ token cross entropy
// One CTA per row of [rows, V] logits.  Forward: loss_sum += lse - logit[target] over rows whose target != ignore;
// lse[row] kept.  Backward: dlogits = gscale * (softmax - onehot) for counted rows, 0 for ignored ones.
template <typename T>
__global__ void __launch_bounds__(256) token_ce_fwd_kernel(const T* __restrict__ logits, const int64_t* __restrict__ target,
                                                           float* __restrict__ lse, float* __restrict__ loss_sum,
                                                           float* __restrict__ count, int64_t V, int64_t ld, int64_t ignore) {
  const int64_t row = blockIdx.x;
  const T* p = logits + row * ld;
  float m = -CUDART_INF_F, l = 0.f;
  for (int64_t c = threadIdx.x; c < V; c += 256) {        // online max / sum
    const float v = to_f(p[c]);
    if (v > m) { l = l * expf(m - v) + 1.0f; m = v; }
    else l += expf(v - m);
  }
  __shared__ float rm[8], rl[8];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const float wm = warp_max(m);
  l = warp_sum(m == -CUDART_INF_F ? 0.f : l * expf(m - wm));
  if (lane == 0) { rm[warp] = wm; rl[warp] = l; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = rm[0];
    for (int w = 1; w < 8; ++w) M = fmaxf(M, rm[w]);
    float L = 0.f;
    for (int w = 0; w < 8; ++w) L += rm[w] == -CUDART_INF_F ? 0.f : rl[w] * expf(rm[w] - M);
    const float z = M + logf(L);
    lse[row] = z;
    const int64_t y = target[row];
    if (y != ignore && y >= 0 && y < V) {
      atomicAdd(loss_sum, z - to_f(p[y]));
      atomicAdd(count, 1.0f);
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256) token_ce_bwd_kernel(const T* __restrict__ logits, const int64_t* __restrict__ target,
                                                           const float* __restrict__ lse, const float* __restrict__ count,
                                                           const float* __restrict__ dloss, T* __restrict__ dlogits, int64_t V,
                                                           int64_t ld, int64_t ignore) {
  const int64_t row = blockIdx.x;
  const T* p = logits + row * ld;
  T* g = dlogits + row * ld;
  const int64_t y = target[row];
  const bool counted = y != ignore && y >= 0 && y < V;
  const float cnt = *count, z = lse[row];
  const float s = counted && cnt > 0.f ? *dloss / cnt : 0.f;
  for (int64_t c = threadIdx.x; c < V; c += 256) {
    const float v = s * (expf(to_f(p[c]) - z) - (c == y ? 1.0f : 0.0f));
    g[c] = from_f<T>(v);
  }
}

}  // namespace meant

using namespace meant;

static bool pool_ok(const PoolArgs& a) {
  return a.x && a.out && a.N > 0 && a.S > 0 && (a.d == 256 || a.d == 512 || a.d == 768 || a.d == 1024) &&
         (a.degenerate || (a.h && a.ln_w && a.ln_b && a.w2));
}

template <typename Kern, typename Args>
static int launch_pool(Kern kern, const Args& args, int64_t N, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return MEANT_EUNSUPPORTED;
  kern<<<(unsigned)N, PW * 32, smem, st>>>(args);
  MEANT_LAUNCH_CHECK();
  return 0;
}

#define POOL_DISPATCH(KERN, ARGS, D, SMEM)                                                          \
  do {                                                                                              \
    if (dtype == MEANT_DT_F32) {                                                                    \
      if (D == 256) return launch_pool(KERN<float, 8>, ARGS, N, SMEM, st);                          \
      if (D == 512) return launch_pool(KERN<float, 16>, ARGS, N, SMEM, st);                         \
      if (D == 768) return launch_pool(KERN<float, 24>, ARGS, N, SMEM, st);                         \
      return launch_pool(KERN<float, 32>, ARGS, N, SMEM, st);                                       \
    }                                                                                               \
    if (dtype == MEANT_DT_BF16) {                                                                   \
      if (D == 256) return launch_pool(KERN<__nv_bfloat16, 8>, ARGS, N, SMEM, st);                  \
      if (D == 512) return launch_pool(KERN<__nv_bfloat16, 16>, ARGS, N, SMEM, st);                 \
      if (D == 768) return launch_pool(KERN<__nv_bfloat16, 24>, ARGS, N, SMEM, st);                 \
      return launch_pool(KERN<__nv_bfloat16, 32>, ARGS, N, SMEM, st);                               \
    }                                                                                               \
    return MEANT_EUNSUPPORTED;                                                                      \
  } while (0)

extern "C" int meant_softmax_pool_fwd(const void* x, const void* h, const float* ln_w, const float* ln_b, float ln_eps,
                                      const float* w2, float b2, void* out, float* prob, int64_t N, int S, int d,
                                      int degenerate, int dtype, meant_stream_t stream) {
  PoolArgs a{x, h, ln_w, ln_b, w2, b2, ln_eps, out, prob, N, S, d, degenerate};
  if (!pool_ok(a) || N > 0x7fffffff) return MEANT_EARG;
  const size_t smem = sizeof(float) * (((size_t)S + 3) / 4 * 4 + (size_t)PW * d);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  POOL_DISPATCH(softmax_pool_fwd_kernel, a, d, smem);
}

extern "C" int meant_softmax_pool_bwd(const void* dout, const void* x, const void* h, const float* ln_w, const float* ln_b,
                                      float ln_eps, const float* w2, const float* prob, void* dx, void* dh, float* dln_w,
                                      float* dln_b, float* dw2, int64_t N, int S, int d, int degenerate, int dtype,
                                      meant_stream_t stream) {
  PoolBwdArgs b{{x, h, ln_w, ln_b, w2, 0.f, ln_eps, dx /* stands in for `out`: unused */, const_cast<float*>(prob), N, S, d,
                 degenerate},
                dout, dx, dh, dln_w, dln_b, dw2};
  if (!pool_ok(b.f) || !dout || !dx || !dh || N > 0x7fffffff) return MEANT_EARG;
  if (!degenerate && (!prob || !dln_w || !dln_b || !dw2)) return MEANT_EARG;
  const size_t smem = sizeof(float) * (((size_t)S + 3) / 4 * 4 + (size_t)d + (size_t)3 * PW * d);
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  POOL_DISPATCH(softmax_pool_bwd_kernel, b, d, smem);
}

extern "C" int meant_token_ce_fwd(const void* logits, const int64_t* target, float* lse, float* loss_sum, float* count,
                                  int64_t rows, int64_t V, int64_t ld, int64_t ignore_index, int dtype, meant_stream_t stream) {
  if (!logits || !target || !lse || !loss_sum || !count || rows <= 0 || V <= 0 || ld < V || rows > 0x7fffffff) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MEANT_DT_F32)
    token_ce_fwd_kernel<float><<<(unsigned)rows, 256, 0, st>>>((const float*)logits, target, lse, loss_sum, count, V, ld, ignore_index);
  else if (dtype == MEANT_DT_BF16)
    token_ce_fwd_kernel<__nv_bfloat16><<<(unsigned)rows, 256, 0, st>>>((const __nv_bfloat16*)logits, target, lse, loss_sum, count, V, ld,
                                                                      ignore_index);
  else
    return MEANT_EUNSUPPORTED;
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_token_ce_bwd(const void* logits, const int64_t* target, const float* lse, const float* count,
                                  const float* dloss, void* dlogits, int64_t rows, int64_t V, int64_t ld, int64_t ignore_index,
                                  int dtype, meant_stream_t stream) {
  if (!logits || !target || !lse || !count || !dloss || !dlogits || rows <= 0 || V <= 0 || ld < V || rows > 0x7fffffff) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MEANT_DT_F32)
    token_ce_bwd_kernel<float><<<(unsigned)rows, 256, 0, st>>>((const float*)logits, target, lse, count, dloss, (float*)dlogits, V, ld,
                                                               ignore_index);
  else if (dtype == MEANT_DT_BF16)
    token_ce_bwd_kernel<__nv_bfloat16><<<(unsigned)rows, 256, 0, st>>>((const __nv_bfloat16*)logits, target, lse, count, dloss,
                                                                      (__nv_bfloat16*)dlogits, V, ld, ignore_index);
  else
    return MEANT_EUNSUPPORTED;
  MEANT_LAUNCH_CHECK();
  return 0;
}
