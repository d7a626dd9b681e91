// Memory-bound helpers around the encoder blocks: mean-pool over the token axis
// (meant/meant.py:231 `torch.mean(words, dim=2)`), GEGLU (src/meant/timesformer_pytorch.py:59-62)
// and the TimeSformer time token-shift (:29-55).  128-bit vectorised, coalesced, fp32 math.
#include "common.cuh"

namespace meant {

// y[n, :] = mean_s x[n, s, :]   — one thread per 16-byte column group, S split over blockIdx.y
template <typename T>
__global__ void __launch_bounds__(256) meanpool_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t N,
                                                           int S, int D) {
  constexpr int V = Vec16<T>::N;
  const int vec_per_row = D / V;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= N * vec_per_row) return;
  const int64_t n = gid / vec_per_row;
  const int c = (int)(gid - n * vec_per_row) * V;
  float acc[V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] = 0.f;
  const T* p = x + (n * S) * D + c;
#pragma unroll 4
  for (int s = 0; s < S; ++s) {
    Vec16<T> t;
    t.load(p + (int64_t)s * D);
    float f[V];
    t.unpack(f);
#pragma unroll
    for (int i = 0; i < V; ++i) acc[i] += f[i];
  }
  const float inv = 1.0f / (float)S;
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] *= inv;
  Vec16<T> o;
  o.pack(acc);
  o.store(y + n * D + c);
}

// dx[n, s, :] = dy[n, :] / S
template <typename T>
__global__ void __launch_bounds__(256) meanpool_bwd_kernel(const T* __restrict__ dy, T* __restrict__ dx, int64_t N,
                                                           int S, int D) {
  constexpr int V = Vec16<T>::N;
  const int vec_per_row = D / V;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= N * S * vec_per_row) return;
  const int64_t row = gid / vec_per_row;  // n * S + s
  const int c = (int)(gid - row * vec_per_row) * V;
  const int64_t n = row / S;
  Vec16<T> t;
  t.load(dy + n * D + c);
  float f[V];
  t.unpack(f);
  const float inv = 1.0f / (float)S;
#pragma unroll
  for (int i = 0; i < V; ++i) f[i] *= inv;
  t.pack(f);
  t.store(dx + row * D + c);
}

// y[t, :F] = x[t, :F] * gelu(x[t, F:2F])
template <typename T>
__global__ void __launch_bounds__(256) geglu_fwd_kernel(const T* __restrict__ x, T* __restrict__ y, int64_t rows,
                                                        int F) {
  constexpr int V = Vec16<T>::N;
  const int vpr = F / V;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows * vpr) return;
  const int64_t r = gid / vpr;
  const int c = (int)(gid - r * vpr) * V;
  Vec16<T> a, g;
  a.load(x + r * 2 * F + c);
  g.load(x + r * 2 * F + F + c);
  float fa[V], fg[V];
  a.unpack(fa);
  g.unpack(fg);
#pragma unroll
  for (int i = 0; i < V; ++i) fa[i] *= gelu_f(fg[i]);
  a.pack(fa);
  a.store(y + r * F + c);
}

template <typename T>
__global__ void __launch_bounds__(256) geglu_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ x,
                                                        T* __restrict__ dx, int64_t rows, int F) {
  constexpr int V = Vec16<T>::N;
  const int vpr = F / V;
  const int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= rows * vpr) return;
  const int64_t r = gid / vpr;
  const int c = (int)(gid - r * vpr) * V;
  Vec16<T> a, g, d;
  a.load(x + r * 2 * F + c);
  g.load(x + r * 2 * F + F + c);
  d.load(dy + r * F + c);
  float fa[V], fg[V], fd[V], da[V], dg[V];
  a.unpack(fa);
  g.unpack(fg);
  d.unpack(fd);
#pragma unroll
  for (int i = 0; i < V; ++i) {
    da[i] = fd[i] * gelu_f(fg[i]);
    dg[i] = fd[i] * fa[i] * gelu_grad_f(fg[i]);
  }
  a.pack(da);
  g.pack(dg);
  a.store(dx + r * 2 * F + c);
  g.store(dx + r * 2 * F + F + c);
}

// Time token-shift on x[b, 1 + f*n, d]: the cls row is copied; in the body, feature chunk 0
// ([0, d/3)) takes frame t+1, chunk 2 ([2d/3', 3d/3')) takes frame t-1, zero padded; the rest is
// copied.  dir = +1 forward op, dir = -1 its transpose (the backward).
template <typename T>
__global__ void __launch_bounds__(256) token_shift_kernel(const T* __restrict__ x, T* __restrict__ y, int B, int f,
                                                          int n, int d, int dir) {
  const int64_t total = (int64_t)B * (1 + f * n) * d;
  const int chunk = d / 3;
  for (int64_t gid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; gid < total;
       gid += (int64_t)gridDim.x * blockDim.x) {
    const int col = (int)(gid % d);
    const int64_t tok = gid / d;
    const int t = (int)(tok % (1 + f * n));
    T v = x[gid];
    if (t > 0 && col < 3 * chunk) {
      const int which = col / chunk;  // 0: from t+1, 1: unchanged, 2: from t-1
      int shift = which == 0 ? 1 : (which == 2 ? -1 : 0);
      shift *= dir;
      if (shift != 0) {
        const int fr = (t - 1) / n + shift;
        v = (fr >= 0 && fr < f) ? x[gid + (int64_t)shift * n * d] : from_f<T>(0.f);
      }
    }
    y[gid] = v;
  }
}


// Bias gradient of a Linear layer: db[c] = sum over rows of dy[r, c] (what autograd's sum-to-size does for
// the q/k/v, multi_mad and MLP biases: meant/xPosAttention.py:24-28, meant/meant.py:61-66).  A pure
// streaming read: 32 x 8 threads, each keeps 8 independent 16-byte loads in flight over its row lane of
// the CTA's row slab, partial sums meet in shared memory and one atomicAdd per column leaves the CTA.
template <typename T>
__global__ void __launch_bounds__(256) bias_grad_kernel(const T* __restrict__ dy, float* __restrict__ db, int64_t rows,
                                                        int cols, int64_t ld, int rows_per_cta) {
  constexpr int V = Vec16<T>::N;
  __shared__ float red[8][32 * V + 1];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col = (blockIdx.x * 32 + tx) * V;
  const int64_t r0 = (int64_t)blockIdx.y * rows_per_cta;
  const int64_t r1 = r0 + rows_per_cta < rows ? r0 + rows_per_cta : rows;
  float acc[V];
#pragma unroll
  for (int i = 0; i < V; ++i) acc[i] = 0.f;
  if (col < cols) {
    const T* p = dy + col;
    int64_t r = r0 + ty;
    for (; r + 56 < r1; r += 64) {
      Vec16<T> v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u].load(p + (r + 8 * u) * ld);
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        float f[V];
        v[u].unpack(f);
#pragma unroll
        for (int i = 0; i < V; ++i) acc[i] += f[i];
      }
    }
    for (; r < r1; r += 8) {
      Vec16<T> v;
      v.load(p + r * ld);
      float f[V];
      v.unpack(f);
#pragma unroll
      for (int i = 0; i < V; ++i) acc[i] += f[i];
    }
  }
#pragma unroll
  for (int i = 0; i < V; ++i) red[ty][tx * V + i] = acc[i];
  __syncthreads();
  const int c = threadIdx.x;
  if (c < 32 * V && blockIdx.x * 32 * V + c < cols) {
    float t = 0.f;
#pragma unroll
    for (int y = 0; y < 8; ++y) t += red[y][c];
    atomicAdd(db + blockIdx.x * 32 * V + c, t);
  }
}

// Token embedding lookup (the `embedding` module handed to the model constructors: meant/meant_tweet.py:152,
// nn.Embedding in bench / tests): out[t, :] = T(W[ids[t], :]) -- the gather writes the compute dtype directly, so no
// fp32 [tokens, dim] intermediate exists.  One warp per token, 16-byte loads.  Ids outside [0, vocab) give a zero row.
template <typename T>
__global__ void __launch_bounds__(256) embedding_fwd_kernel(const int64_t* __restrict__ ids, const float* __restrict__ W,
                                                            T* __restrict__ out, int64_t tokens, int D, int64_t vocab) {
  const int lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); t < tokens; t += (int64_t)gridDim.x * 8) {
    const int64_t id = ids[t];
    const bool ok = id >= 0 && id < vocab;
    const float4* src = reinterpret_cast<const float4*>(W + (ok ? id : 0) * D);
    for (int c = lane; c < D / 4; c += 32) {
      float4 v = ok ? __ldg(src + c) : make_float4(0.f, 0.f, 0.f, 0.f);
      if constexpr (sizeof(T) == 4) {
        reinterpret_cast<float4*>(out + t * D)[c] = v;
      } else {
        const __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&lo);
        pk.y = *reinterpret_cast<const uint32_t*>(&hi);
        reinterpret_cast<uint2*>(out + t * D)[c] = pk;
      }
    }
  }
}

// dW[ids[t], :] += dy[t, :]  (dW fp32, zero-filled by the entry point): one warp per token, one 16-byte vector
// reduction (red.global.add.v4.f32) per 4 columns -- no sort, no segmented scan.
template <typename T>
__global__ void __launch_bounds__(256) embedding_bwd_kernel(const int64_t* __restrict__ ids, const T* __restrict__ dy,
                                                            float* __restrict__ dW, int64_t tokens, int D, int64_t vocab) {
  const int lane = threadIdx.x & 31;
  for (int64_t t = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); t < tokens; t += (int64_t)gridDim.x * 8) {
    const int64_t id = ids[t];
    if (id < 0 || id >= vocab) continue;
    float4* dst = reinterpret_cast<float4*>(dW + id * D);
    for (int c = lane; c < D / 4; c += 32) {
      float4 v;
      if constexpr (sizeof(T) == 4) {
        v = reinterpret_cast<const float4*>(dy + t * D)[c];
      } else {
        const uint2 pk = reinterpret_cast<const uint2*>(dy + t * D)[c];
        const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
        const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
        v = make_float4(lo.x, lo.y, hi.x, hi.y);
      }
      atomicAdd(dst + c, v);
    }
  }
}

// Patchify: 'b c (h p1) (w p2) -> b (h w) (p1 p2 c)' (meant/meant.py:193-195; per frame the same pattern in
// src/meant/timesformer_pytorch.py:219-220) fused with the cast to the compute dtype: the im2col operand of the
// patch-embedding GEMM is written once, in bf16 under autocast, instead of a permuted fp32 copy plus a cast pass.
// A thread owns one pixel position (p1, p2) of one patch: it reads the C channel values (each read is part of a
// 16-pixel, 64-byte run of an image row) and writes them as C consecutive outputs, so a warp writes one contiguous
// 32 * C element run.  TRANSPOSE: the adjoint (patch gradients back to the image layout).
template <typename TI, typename TO, bool TRANSPOSE>
__global__ void __launch_bounds__(256) patchify_kernel(const TI* __restrict__ in, TO* __restrict__ out, int64_t B, int C,
                                                       int H, int W, int P) {
  const int hp = H / P, wp = W / P;
  const int64_t total = B * hp * wp * P * P;
  for (int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x; idx < total; idx += (int64_t)gridDim.x * 256) {
    const int p2 = (int)(idx % P);
    int64_t r = idx / P;
    const int p1 = (int)(r % P);
    r /= P;
    const int w = (int)(r % wp);
    r /= wp;
    const int h = (int)(r % hp);
    const int64_t b = r / hp;
    const int64_t img = (b * C * H + (int64_t)(h * P + p1)) * W + (w * P + p2);     // channel 0; + c * H * W
    const int64_t pat = idx * C;                                                      // (b, h, w, p1, p2, 0)
    for (int c = 0; c < C; ++c) {
      if (!TRANSPOSE) out[pat + c] = from_f<TO>(to_f(in[img + (int64_t)c * H * W]));
      else out[img + (int64_t)c * H * W] = from_f<TO>(to_f(in[pat + c]));
    }
  }
}

}  // namespace meant

using namespace meant;

extern "C" int meant_patchify(const void* in, void* out, int64_t B, int C, int H, int W, int P, int in_dtype, int out_dtype,
                              int transpose, cudaStream_t st) {
  if (!in || !out || B <= 0 || C <= 0 || H <= 0 || W <= 0 || P <= 0 || H % P || W % P) return MEANT_EARG;
  const int64_t total = B * (H / P) * (W / P) * P * P;
  unsigned blocks = (unsigned)((total + 255) / 256 < 148 * 32 ? (total + 255) / 256 : 148 * 32);
#define PATCHIFY_GO(TI, TO)                                                                                     \
  do {                                                                                                          \
    if (transpose) patchify_kernel<TI, TO, true><<<blocks, 256, 0, st>>>((const TI*)in, (TO*)out, B, C, H, W, P); \
    else patchify_kernel<TI, TO, false><<<blocks, 256, 0, st>>>((const TI*)in, (TO*)out, B, C, H, W, P);        \
  } while (0)
  if (in_dtype == MEANT_DT_F32 && out_dtype == MEANT_DT_F32) PATCHIFY_GO(float, float);
  else if (in_dtype == MEANT_DT_F32 && out_dtype == MEANT_DT_BF16) PATCHIFY_GO(float, __nv_bfloat16);
  else if (in_dtype == MEANT_DT_BF16 && out_dtype == MEANT_DT_BF16) PATCHIFY_GO(__nv_bfloat16, __nv_bfloat16);
  else if (in_dtype == MEANT_DT_BF16 && out_dtype == MEANT_DT_F32) PATCHIFY_GO(__nv_bfloat16, float);
  else return MEANT_EUNSUPPORTED;
#undef PATCHIFY_GO
  MEANT_LAUNCH_CHECK();
  return 0;
}

#define DISPATCH_T(dtype, CALL)                                   \
  if (dtype == MEANT_DT_F32) { using T = float; CALL; }           \
  else if (dtype == MEANT_DT_BF16) { using T = __nv_bfloat16; CALL; } \
  else return MEANT_EUNSUPPORTED;

static inline unsigned blocks_for(int64_t work, int per) { return (unsigned)((work + per - 1) / per); }

extern "C" int meant_meanpool_fwd(const void* x, void* y, int64_t N, int S, int D, int dtype, cudaStream_t st) {
  if (!x || !y || N < 0 || S <= 0 || D <= 0) return MEANT_EARG;
  if (D % (dtype == MEANT_DT_F32 ? 4 : 8)) return MEANT_EUNSUPPORTED;
  if (N == 0) return 0;
  DISPATCH_T(dtype, (meanpool_fwd_kernel<T><<<blocks_for(N * (D / Vec16<T>::N), 256), 256, 0, st>>>(
                        (const T*)x, (T*)y, N, S, D)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_meanpool_bwd(const void* dy, void* dx, int64_t N, int S, int D, int dtype, cudaStream_t st) {
  if (!dy || !dx || N < 0 || S <= 0 || D <= 0) return MEANT_EARG;
  if (D % (dtype == MEANT_DT_F32 ? 4 : 8)) return MEANT_EUNSUPPORTED;
  if (N == 0) return 0;
  DISPATCH_T(dtype, (meanpool_bwd_kernel<T><<<blocks_for(N * S * (D / Vec16<T>::N), 256), 256, 0, st>>>(
                        (const T*)dy, (T*)dx, N, S, D)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_geglu_fwd(const void* x, void* y, int64_t rows, int F, int dtype, cudaStream_t st) {
  if (!x || !y || rows < 0 || F <= 0) return MEANT_EARG;
  if (F % (dtype == MEANT_DT_F32 ? 4 : 8)) return MEANT_EUNSUPPORTED;
  if (rows == 0) return 0;
  DISPATCH_T(dtype, (geglu_fwd_kernel<T><<<blocks_for(rows * (F / Vec16<T>::N), 256), 256, 0, st>>>(
                        (const T*)x, (T*)y, rows, F)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_geglu_bwd(const void* dy, const void* x, void* dx, int64_t rows, int F, int dtype,
                               cudaStream_t st) {
  if (!dy || !x || !dx || rows < 0 || F <= 0) return MEANT_EARG;
  if (F % (dtype == MEANT_DT_F32 ? 4 : 8)) return MEANT_EUNSUPPORTED;
  if (rows == 0) return 0;
  DISPATCH_T(dtype, (geglu_bwd_kernel<T><<<blocks_for(rows * (F / Vec16<T>::N), 256), 256, 0, st>>>(
                        (const T*)dy, (const T*)x, (T*)dx, rows, F)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_token_shift(const void* x, void* y, int B, int f, int n, int d, int transpose, int dtype,
                                 cudaStream_t st) {
  if (!x || !y || B <= 0 || f <= 0 || n <= 0 || d <= 0 || x == y) return MEANT_EARG;
  const int64_t total = (int64_t)B * (1 + f * n) * d;
  unsigned blocks = blocks_for(total, 256);
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  DISPATCH_T(dtype, (token_shift_kernel<T><<<blocks, 256, 0, st>>>((const T*)x, (T*)y, B, f, n, d,
                                                                    transpose ? -1 : 1)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_bias_grad(const void* dy, float* db, int64_t rows, int cols, int64_t ld, int dtype,
                               cudaStream_t st) {
  if ((!dy && rows != 0) || !db || rows < 0 || cols <= 0 || ld < cols) return MEANT_EARG;
  const int v = dtype == MEANT_DT_F32 ? 4 : 8;
  if (cols % v || ld % v || ((uintptr_t)dy & 15)) return MEANT_EUNSUPPORTED;
  cudaError_t e = cudaMemsetAsync(db, 0, sizeof(float) * cols, st);
  if (e != cudaSuccess) return (int)e;
  if (rows == 0) return 0;
  const unsigned gx = blocks_for(cols, 32 * v);
  // ~4 CTAs per SM in flight, slabs of at least 64 rows
  int64_t slabs = (148 * 4 + gx - 1) / gx;
  if (slabs > (rows + 63) / 64) slabs = (rows + 63) / 64;
  const int rows_per_cta = (int)(((rows + slabs - 1) / slabs + 7) / 8 * 8);
  dim3 grid(gx, blocks_for(rows, rows_per_cta));
  DISPATCH_T(dtype, (bias_grad_kernel<T><<<grid, 256, 0, st>>>((const T*)dy, db, rows, cols, ld, rows_per_cta)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_embedding_fwd(const int64_t* ids, const float* W, void* out, int64_t tokens, int D, int64_t vocab,
                                   int dtype, cudaStream_t st) {
  if (!ids || !W || !out || tokens < 0 || D <= 0 || vocab <= 0) return MEANT_EARG;
  if (D % 4 || ((uintptr_t)W & 15) || ((uintptr_t)out & 15)) return MEANT_EUNSUPPORTED;
  if (tokens == 0) return 0;
  unsigned blocks = blocks_for(tokens, 8);
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  DISPATCH_T(dtype, (embedding_fwd_kernel<T><<<blocks, 256, 0, st>>>(ids, W, (T*)out, tokens, D, vocab)));
  MEANT_LAUNCH_CHECK();
  return 0;
}

extern "C" int meant_embedding_bwd(const int64_t* ids, const void* dy, float* dW, int64_t tokens, int D, int64_t vocab,
                                   int dtype, cudaStream_t st) {
  if (!ids || !dy || !dW || tokens < 0 || D <= 0 || vocab <= 0) return MEANT_EARG;
  if (D % 4 || ((uintptr_t)dW & 15) || ((uintptr_t)dy & 15)) return MEANT_EUNSUPPORTED;
  cudaError_t e = cudaMemsetAsync(dW, 0, sizeof(float) * (size_t)vocab * D, st);
  if (e != cudaSuccess) return (int)e;
  if (tokens == 0) return 0;
  unsigned blocks = blocks_for(tokens, 8);
  if (blocks > 148u * 16u) blocks = 148u * 16u;
  DISPATCH_T(dtype, (embedding_bwd_kernel<T><<<blocks, 256, 0, st>>>(ids, (const T*)dy, dW, tokens, D, vocab)));
  MEANT_LAUNCH_CHECK();
  return 0;
}
