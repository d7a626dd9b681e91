// The BERTweet (RoBERTa) embedding block the reference passes to every text model as `embedding`
// (src/in_loop_train.py:440,486-496: `embedding=bertweet.embeddings`; meant/meant.py:211-212 `words = mod(words)`):
//     pos    = cumsum(ids != pad) * (ids != pad) + pad                (create_position_ids_from_input_ids)
//     out[t] = Dropout( LayerNorm( word[ids[t]] + position[pos[t]] + type[0] ) )
// One kernel per direction instead of three gathers, two adds, a LayerNorm and a dropout pass:
//   forward   a CTA per sequence scans the padding mask once, then a warp per token gathers the three fp32 rows with
//             16-byte loads, normalises in registers and writes the compute dtype; pos / mean / rstd are kept
//   backward  LayerNorm backward per token from the re-gathered rows, then vector reductions (red.global.add.v4.f32)
//             into the word / position tables' gradients; the type row's gradient and the LayerNorm parameter
//             gradients are reduced per CTA first (every token hits the same addresses), one atomic per column per CTA
// HBM-bound on the output write (forward) and on the scatter (backward).  d = 32 E, E in {8, 16, 24, 32}.
#include "common.cuh"

namespace meant {

constexpr int EW = 8;      // warps per CTA

struct EmbArgs {
  const int64_t* ids;      // [N, S]
  const float* word;       // [vocab, d]
  const float* pos;        // [max_pos, d]
  const float* type;       // [types, d] (row 0 is used: token_type_ids default to zeros)
  const float* ln_w;
  const float* ln_b;
  void* out;               // [N, S, d]
  int* pos_ids;            // [N, S]
  float* mean;             // [N, S]
  float* rstd;
  int64_t N, vocab;
  int S, d, max_pos, pad;
  float eps;
  uint32_t drop_thresh;
  float drop_inv_keep;
  unsigned long long seed;
  const unsigned long long* seed_dev;
};

template <int E>
__device__ __forceinline__ void gather3(const EmbArgs& a, int64_t id, int p, int lane, float* f) {
  const float4* w = reinterpret_cast<const float4*>(a.word + id * a.d);
  const float4* q = reinterpret_cast<const float4*>(a.pos + (int64_t)p * a.d);
  const float4* ty = reinterpret_cast<const float4*>(a.type);
#pragma unroll
  for (int v = 0; v < E / 4; ++v) {
    const float4 x = __ldg(w + lane + 32 * v), y = __ldg(q + lane + 32 * v), z = __ldg(ty + lane + 32 * v);
    f[4 * v] = x.x + y.x + z.x;
    f[4 * v + 1] = x.y + y.y + z.y;
    f[4 * v + 2] = x.z + y.z + z.z;
    f[4 * v + 3] = x.w + y.w + z.w;
  }
}

// position ids of one sequence into shared memory (block-wide inclusive scan of the not-padding mask)
__device__ __forceinline__ void scan_positions(const int64_t* __restrict__ ids, int S, int pad, int* spos, int* swarp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int carry = 0;
  for (int base = 0; base < S; base += EW * 32) {
    const int t = base + threadIdx.x;
    const int m = t < S && ids[t] != pad ? 1 : 0;
    int x = m;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) swarp[warp] = x;
    __syncthreads();
    int off = carry;
    for (int w = 0; w < warp; ++w) off += swarp[w];
    if (t < S) spos[t] = m ? x + off + pad : pad;
    int tot = 0;
    for (int w = 0; w < EW; ++w) tot += swarp[w];
    carry += tot;
    __syncthreads();
  }
}

template <typename T, int E>
__global__ void __launch_bounds__(EW * 32) embed_ln_fwd_kernel(EmbArgs a) {
  extern __shared__ int spos[];
  __shared__ int swarp[EW];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t n = blockIdx.x;
  const int64_t* ids = a.ids + n * a.S;
  scan_positions(ids, a.S, a.pad, spos, swarp);
  unsigned long long seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  constexpr int VN = Vec16<T>::N;
  for (int t = warp; t < a.S; t += EW) {
    const int64_t id = ids[t], tok = n * a.S + t;
    const int p = min(spos[t], a.max_pos - 1);
    float f[E];
    if (id >= 0 && id < a.vocab) gather3<E>(a, id, p, lane, f);
    else {
#pragma unroll
      for (int i = 0; i < E; ++i) f[i] = 0.f;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < E; ++i) s += f[i];
    const float mean = warp_sum(s) / a.d;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < E; ++i) { const float c = f[i] - mean; q += c * c; }
    const float rstd = rsqrtf(warp_sum(q) / a.d + a.eps);
    if (lane == 0) {
      a.pos_ids[tok] = p;
      a.mean[tok] = mean;
      a.rstd[tok] = rstd;
    }
    // element (v, i) of this lane is column (lane + 32 v) * 4 + i; outputs are written as 16-byte vectors of T
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
      const int c0 = (lane + 32 * v) * 4;
      const float4 lw = __ldg(reinterpret_cast<const float4*>(a.ln_w + c0)), lb = __ldg(reinterpret_cast<const float4*>(a.ln_b + c0));
      float y[4] = {(f[4 * v] - mean) * rstd * lw.x + lb.x, (f[4 * v + 1] - mean) * rstd * lw.y + lb.y,
                    (f[4 * v + 2] - mean) * rstd * lw.z + lb.z, (f[4 * v + 3] - mean) * rstd * lw.w + lb.w};
      if (a.drop_thresh) {
        const uint32_t m4 = drop_mask4(seed, (uint64_t)(tok * a.d + c0) >> 2, a.drop_thresh);
#pragma unroll
        for (int j = 0; j < 4; ++j) y[j] = drop_sel(y[j], m4, j) * a.drop_inv_keep;
      }
      T* dst = static_cast<T*>(a.out) + tok * a.d + c0;
      if (VN == 4) {
        *reinterpret_cast<float4*>(dst) = make_float4(y[0], y[1], y[2], y[3]);
      } else {
        const __nv_bfloat162 lo = __floats2bfloat162_rn(y[0], y[1]), hi = __floats2bfloat162_rn(y[2], y[3]);
        uint2 pk;
        pk.x = *reinterpret_cast<const uint32_t*>(&lo);
        pk.y = *reinterpret_cast<const uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(dst) = pk;
      }
    }
  }
}

struct EmbBwdArgs {
  EmbArgs f;
  const void* dy;          // [N, S, d]
  float* dword;            // [vocab, d] zero on entry, scatter-added
  float* dpos;             // [max_pos, d]
  float* dtype_row;        // [d] gradient of type row 0
  float* dln_w;
  float* dln_b;
};

template <typename T, int E>
__global__ void __launch_bounds__(EW * 32) embed_ln_bwd_kernel(EmbBwdArgs b) {
  const EmbArgs& a = b.f;
  extern __shared__ float part[];                 // [3][EW][d]: type row, ln_w, ln_b partials
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t n = blockIdx.x;
  const int64_t* ids = a.ids + n * a.S;
  unsigned long long seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  constexpr int VN = Vec16<T>::N;
  float aty[E], alw[E], alb[E];
#pragma unroll
  for (int i = 0; i < E; ++i) aty[i] = alw[i] = alb[i] = 0.f;
  for (int t = warp; t < a.S; t += EW) {
    const int64_t id = ids[t], tok = n * a.S + t;
    if (id < 0 || id >= a.vocab) continue;          // warp-uniform
    const int p = a.pos_ids[tok];
    const float mean = a.mean[tok], rstd = a.rstd[tok];
    float f[E], g[E];
    gather3<E>(a, id, p, lane, f);
    float m1 = 0.f, m2 = 0.f;
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
      const int c0 = (lane + 32 * v) * 4;
      float dy[4];
      const T* src = static_cast<const T*>(b.dy) + tok * a.d + c0;
      if (VN == 4) {
        const float4 d4 = *reinterpret_cast<const float4*>(src);
        dy[0] = d4.x; dy[1] = d4.y; dy[2] = d4.z; dy[3] = d4.w;
      } else {
        const uint2 pk = *reinterpret_cast<const uint2*>(src);
        const float2 lo = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.x));
        const float2 hi = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&pk.y));
        dy[0] = lo.x; dy[1] = lo.y; dy[2] = hi.x; dy[3] = hi.y;
      }
      if (a.drop_thresh) {
        const uint32_t m4 = drop_mask4(seed, (uint64_t)(tok * a.d + c0) >> 2, a.drop_thresh);
#pragma unroll
        for (int j = 0; j < 4; ++j) dy[j] = drop_sel(dy[j], m4, j) * a.drop_inv_keep;
      }
      const float4 lw = __ldg(reinterpret_cast<const float4*>(a.ln_w + c0));
      const float lwv[4] = {lw.x, lw.y, lw.z, lw.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int e = 4 * v + i;
        const float xhat = (f[e] - mean) * rstd;
        alw[e] = fmaf(dy[i], xhat, alw[e]);
        alb[e] += dy[i];
        const float dxh = dy[i] * lwv[i];
        m1 += dxh;
        m2 = fmaf(dxh, xhat, m2);
        f[e] = xhat;
        g[e] = dxh;
      }
    }
    m1 = warp_sum(m1) / a.d;
    m2 = warp_sum(m2) / a.d;
    float* dw = b.dword + id * a.d;
    float* dp = b.dpos + (int64_t)p * a.d;
#pragma unroll
    for (int v = 0; v < E / 4; ++v) {
      const int c0 = (lane + 32 * v) * 4;
      float4 r;
      r.x = rstd * (g[4 * v] - m1 - f[4 * v] * m2);
      r.y = rstd * (g[4 * v + 1] - m1 - f[4 * v + 1] * m2);
      r.z = rstd * (g[4 * v + 2] - m1 - f[4 * v + 2] * m2);
      r.w = rstd * (g[4 * v + 3] - m1 - f[4 * v + 3] * m2);
      atomicAdd(reinterpret_cast<float4*>(dw + c0), r);
      atomicAdd(reinterpret_cast<float4*>(dp + c0), r);
      aty[4 * v] += r.x;
      aty[4 * v + 1] += r.y;
      aty[4 * v + 2] += r.z;
      aty[4 * v + 3] += r.w;
    }
  }
#pragma unroll
  for (int v = 0; v < E / 4; ++v)
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = (lane + 32 * v) * 4 + i, e = 4 * v + i;
      part[(0 * EW + warp) * a.d + c] = aty[e];
      part[(1 * EW + warp) * a.d + c] = alw[e];
      part[(2 * EW + warp) * a.d + c] = alb[e];
    }
  __syncthreads();
  for (int c = threadIdx.x; c < a.d; c += EW * 32) {
    float s0 = 0.f, s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int w = 0; w < EW; ++w) {
      s0 += part[(0 * EW + w) * a.d + c];
      s1 += part[(1 * EW + w) * a.d + c];
      s2 += part[(2 * EW + w) * a.d + c];
    }
    atomicAdd(b.dtype_row + c, s0);
    atomicAdd(b.dln_w + c, s1);
    atomicAdd(b.dln_b + c, s2);
  }
}

static uint32_t emb_thr8(float p) {   // round(256 p): common.cuh drop_mask4
  if (!(p > 0.f)) return 0u;
  const uint32_t t = (uint32_t)(p * 256.0f + 0.5f);
  return t < 1u ? 1u : (t > 255u ? 255u : t);
}
static uint32_t emb_thr4(float p) { return emb_thr8(p) * 0x01010101u; }
static float emb_scale(float p) { return 256.0f / (256.0f - (float)emb_thr8(p)); }

template <typename Kern, typename Args>
static int launch_emb(Kern kern, const Args& args, int64_t N, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024 && cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess)
    return MEANT_EUNSUPPORTED;
  kern<<<(unsigned)N, EW * 32, smem, st>>>(args);
  MEANT_LAUNCH_CHECK();
  return 0;
}

#define EMB_DISPATCH(KERN, ARGS, D, SMEM)                                                    \
  do {                                                                                       \
    if (dtype == MEANT_DT_F32) {                                                             \
      if (D == 256) return launch_emb(KERN<float, 8>, ARGS, N, SMEM, st);                    \
      if (D == 512) return launch_emb(KERN<float, 16>, ARGS, N, SMEM, st);                   \
      if (D == 768) return launch_emb(KERN<float, 24>, ARGS, N, SMEM, st);                   \
      return launch_emb(KERN<float, 32>, ARGS, N, SMEM, st);                                 \
    }                                                                                        \
    if (dtype == MEANT_DT_BF16) {                                                            \
      if (D == 256) return launch_emb(KERN<__nv_bfloat16, 8>, ARGS, N, SMEM, st);            \
      if (D == 512) return launch_emb(KERN<__nv_bfloat16, 16>, ARGS, N, SMEM, st);           \
      if (D == 768) return launch_emb(KERN<__nv_bfloat16, 24>, ARGS, N, SMEM, st);           \
      return launch_emb(KERN<__nv_bfloat16, 32>, ARGS, N, SMEM, st);                         \
    }                                                                                        \
    return MEANT_EUNSUPPORTED;                                                               \
  } while (0)

static bool emb_ok(const EmbArgs& a) {
  return a.ids && a.word && a.pos && a.type && a.ln_w && a.ln_b && a.pos_ids && a.mean && a.rstd && a.N > 0 && a.S > 0 &&
         a.vocab > 0 && a.max_pos > 0 && (a.d == 256 || a.d == 512 || a.d == 768 || a.d == 1024) && a.N <= 0x7fffffff;
}

}  // namespace meant

using namespace meant;

extern "C" int meant_embed_ln_fwd(const int64_t* ids, const float* word, const float* pos, const float* type, const float* ln_w,
                                  const float* ln_b, float eps, void* out, int* pos_ids, float* mean, float* rstd, int64_t N,
                                  int S, int d, int64_t vocab, int max_pos, int pad_idx, float drop_p, uint64_t seed,
                                  const uint64_t* seed_dev, int dtype, meant_stream_t stream) {
  EmbArgs a{ids, word, pos, type, ln_w, ln_b, out, pos_ids, mean, rstd, N, vocab, S, d, max_pos, pad_idx, eps,
            emb_thr4(drop_p), emb_scale(drop_p), seed, reinterpret_cast<const unsigned long long*>(seed_dev)};
  if (!emb_ok(a) || !out || drop_p < 0.f || drop_p >= 1.f) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t smem = sizeof(int) * (size_t)S;
  EMB_DISPATCH(embed_ln_fwd_kernel, a, d, smem);
}

extern "C" int meant_embed_ln_bwd(const void* dy, const int64_t* ids, const int* pos_ids, const float* mean, const float* rstd,
                                  const float* word, const float* pos, const float* type, const float* ln_w, float* dword,
                                  float* dpos, float* dtype_row, float* dln_w, float* dln_b, int64_t N, int S, int d,
                                  int64_t vocab, int max_pos, float drop_p, uint64_t seed, const uint64_t* seed_dev, int dtype,
                                  meant_stream_t stream) {
  EmbBwdArgs b{{ids, word, pos, type, ln_w, ln_w /* ln_b is not needed */, nullptr, const_cast<int*>(pos_ids),
                const_cast<float*>(mean), const_cast<float*>(rstd), N, vocab, S, d, max_pos, 0, 0.f, emb_thr4(drop_p),
                emb_scale(drop_p), seed, reinterpret_cast<const unsigned long long*>(seed_dev)},
               dy, dword, dpos, dtype_row, dln_w, dln_b};
  if (!emb_ok(b.f) || !dy || !dword || !dpos || !dtype_row || !dln_w || !dln_b || drop_p < 0.f || drop_p >= 1.f) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  const size_t smem = sizeof(float) * (size_t)3 * EW * d;
  EMB_DISPATCH(embed_ln_bwd_kernel, b, d, smem);
}
