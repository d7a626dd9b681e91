// Single-query attention (Sq = 1), CUDA cores, fp32 math: the TimeSformer cls query, which attends to every token
// (src/meant/timesformer_pytorch.py:116-119: `cls_out = attn(cls_q, k, v)`, one row against 1 + f*n keys), and the
// same shape as temporal.forward (meant/temporal.py:43-58).  One query row per (sequence, head) leaves nothing for
// the tensor cores (a 128-row MMA tile would carry one live row); the work is a stream over K and V, so the keys
// are split into chunks of 64 across CTAs (flash-decoding style): grid = (chunks, sequences), a warp per head.
//   forward   partial pass: per chunk (max, sum, sum p v) -> scratch; merge pass: one warp per (sequence, head)
//   backward  one pass: per key p, dS from the saved LSE; dK = dS q and dV = p dO are whole rows (written, or added
//             to what the buffer holds: the keys also get gradients from the group attention of the same module);
//             dq = sum dS k is reduced per CTA and added atomically to an fp32 row.
// HBM-bound: K and V are read once (forward), once more plus one read-modify-write of dK, dV (backward).
#include "attn_common.cuh"
#include <math_constants.h>
#include <atomic>
#include <cstdio>

namespace meant {
namespace q1 {

constexpr int KC = 64;        // keys per CTA
constexpr int MAXP = 4;       // feature pairs per lane: Dh <= 256
constexpr int NW = 8;         // warps per CTA (heads are dealt round-robin to warps)

template <typename T> struct Pair;
template <> struct Pair<float> {
  static __device__ __forceinline__ float2 ld(const float* p) { return *reinterpret_cast<const float2*>(p); }
  static __device__ __forceinline__ void st(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
};
template <> struct Pair<__nv_bfloat16> {
  static __device__ __forceinline__ float2 ld(const __nv_bfloat16* p) {
    return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float2 v) {
    *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(v.x, v.y);
  }
};

// dot of a shared-memory fp32 vector with one row of T (16-byte loads; Dh % 8 == 0)
template <typename T>
__device__ __forceinline__ float dot_row(const float* __restrict__ vec, const T* __restrict__ row, int Dh) {
  float acc = 0.f;
  constexpr int N = Vec16<T>::N;
  for (int d0 = 0; d0 < Dh; d0 += N) {
    Vec16<T> v;
    v.load(row + d0);
    float f[N];
    v.unpack(f);
#pragma unroll
    for (int i = 0; i < N; ++i) acc = fmaf(vec[d0 + i], f[i], acc);
  }
  return acc;
}

// ------------------------------------------------------------------------------------------ forward, partial
template <typename T>
__global__ void __launch_bounds__(NW * 32) q1_partial_kernel(AttnArgs a, float* __restrict__ partial, int nchunk) {
  extern __shared__ float qs[];                       // [H][Dh] the query rows of this sequence, pre-scaled
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x, n = blockIdx.y, Dh = a.Dh;
  for (int i = threadIdx.x; i < a.H * Dh; i += blockDim.x) {
    const int h = i / Dh, d = i - h * Dh;
    qs[i] = to_f(view_ptr<T>(a.q, n, a.n2_count, 0, h)[d]) * a.scale;
  }
  __syncthreads();
  const int k0 = c * KC;
  for (int h = warp; h < a.H; h += NW) {
    const float* q = qs + h * Dh;
    float s[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int kj = k0 + lane + 32 * u;
      s[u] = kj < a.Sk ? dot_row<T>(q, view_ptr<T>(a.k, n, a.n2_count, kj, h), Dh) : -CUDART_INF_F;
    }
    const float m = warp_max(fmaxf(s[0], s[1]));      // every chunk holds at least one key
    const float p0 = expf(s[0] - m), p1 = expf(s[1] - m);
    const float l = warp_sum(p0 + p1);
    float2 o[MAXP];
#pragma unroll
    for (int i = 0; i < MAXP; ++i) o[i] = make_float2(0.f, 0.f);
    const int kend = min(KC, a.Sk - k0);
#pragma unroll 8                                      // eight rows' loads in flight: the loop was one L2 latency per key
    for (int j = 0; j < kend; ++j) {
      const float pj = __shfl_sync(0xffffffffu, j < 32 ? p0 : p1, j & 31);
      const T* vrow = view_ptr<T>(a.v, n, a.n2_count, k0 + j, h);
#pragma unroll
      for (int i = 0; i < MAXP; ++i) {
        const int d = 2 * (lane + 32 * i);
        if (d < Dh) {
          const float2 v = Pair<T>::ld(vrow + d);
          o[i].x = fmaf(pj, v.x, o[i].x);
          o[i].y = fmaf(pj, v.y, o[i].y);
        }
      }
    }
    float* dst = partial + (((int64_t)n * a.H + h) * nchunk + c) * (Dh + 2);
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
      const int d = 2 * (lane + 32 * i);
      if (d < Dh) *reinterpret_cast<float2*>(dst + d) = o[i];
    }
    if (lane == 0) {
      dst[Dh] = m;
      dst[Dh + 1] = l;
    }
  }
}

// ------------------------------------------------------------------------------------------ forward, merge
template <typename T>
__global__ void __launch_bounds__(128) q1_merge_kernel(AttnArgs a, const float* __restrict__ partial, int nchunk) {
  const int lane = threadIdx.x & 31;
  const int64_t item = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
  if (item >= (int64_t)a.N * a.H) return;
  const int n = (int)(item / a.H), h = (int)(item - (int64_t)n * a.H), Dh = a.Dh;
  const float* src = partial + item * nchunk * (Dh + 2);
  float m = -CUDART_INF_F;
  for (int c = lane; c < nchunk; c += 32) m = fmaxf(m, src[c * (Dh + 2) + Dh]);
  m = warp_max(m);
  float l = 0.f;
  float2 o[MAXP];
#pragma unroll
  for (int i = 0; i < MAXP; ++i) o[i] = make_float2(0.f, 0.f);
  for (int c = 0; c < nchunk; ++c) {
    const float* pc = src + c * (Dh + 2);
    const float w = expf(pc[Dh] - m);
    l = fmaf(pc[Dh + 1], w, l);
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
      const int d = 2 * (lane + 32 * i);
      if (d < Dh) {
        const float2 v = *reinterpret_cast<const float2*>(pc + d);
        o[i].x = fmaf(v.x, w, o[i].x);
        o[i].y = fmaf(v.y, w, o[i].y);
      }
    }
  }
  const float inv = 1.0f / l;
  T* op = view_ptr<T>(a.o, n, a.n2_count, 0, h);
#pragma unroll
  for (int i = 0; i < MAXP; ++i) {
    const int d = 2 * (lane + 32 * i);
    if (d < Dh) Pair<T>::st(op + d, make_float2(o[i].x * inv, o[i].y * inv));
  }
  if (lane == 0 && a.lse) a.lse[item] = m + logf(l);
}

// ------------------------------------------------------------------------------------------ backward
template <typename T, bool ACC>
__global__ void __launch_bounds__(NW * 32) q1_bwd_kernel(AttnArgs a, float* __restrict__ dq32) {
  extern __shared__ float sm[];                       // [H][Dh] q * scale | [H][Dh] dO | [H] delta
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x, n = blockIdx.y, Dh = a.Dh, HD = a.H * Dh;
  float* qs = sm;
  float* gs = sm + HD;
  float* ds_ = sm + 2 * HD;
  for (int i = threadIdx.x; i < HD; i += blockDim.x) {
    const int h = i / Dh, d = i - h * Dh;
    qs[i] = to_f(view_ptr<T>(a.q, n, a.n2_count, 0, h)[d]) * a.scale;
    gs[i] = to_f(view_ptr<T>(a.o, n, a.n2_count, 0, h)[d]);
  }
  __syncthreads();
  for (int h = warp; h < a.H; h += NW) {              // delta = <dO, O>
    const T* op = view_ptr<T>(a.out_saved, n, a.n2_count, 0, h);
    float acc = 0.f;
    for (int d = lane; d < Dh; d += 32) acc = fmaf(gs[h * Dh + d], to_f(op[d]), acc);
    acc = warp_sum(acc);
    if (lane == 0) ds_[h] = acc;
  }
  __syncthreads();
  const int k0 = c * KC;
  constexpr int N = Vec16<T>::N;
  for (int h = warp; h < a.H; h += NW) {
    const float* q = qs + h * Dh;                     // already times scale
    const float* g = gs + h * Dh;
    const float lse = a.lse[(int64_t)n * a.H + h], delta = ds_[h];
    float dsv[2];
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      const int kj = k0 + lane + 32 * u;
      dsv[u] = 0.f;
      if (kj < a.Sk) {
        const T* krow = view_ptr<T>(a.k, n, a.n2_count, kj, h);
        const T* vrow = view_ptr<T>(a.v, n, a.n2_count, kj, h);
        const float p = expf(dot_row<T>(q, krow, Dh) - lse);
        const float dsc = p * (dot_row<T>(g, vrow, Dh) - delta);      // d loss / d score
        dsv[u] = dsc;
        T* gk = view_ptr<T>(a.dk, n, a.n2_count, kj, h);
        T* gv = view_ptr<T>(a.dv, n, a.n2_count, kj, h);
#pragma unroll 4
        for (int d0 = 0; d0 < Dh; d0 += N) {          // dK = dS (q * scale), dV = p dO: this lane's own rows
          float fk[N], fv[N];
          if (ACC) {
            Vec16<T> ok, ov;
            ok.load(gk + d0);
            ov.load(gv + d0);
            ok.unpack(fk);
            ov.unpack(fv);
          } else {
#pragma unroll
            for (int i = 0; i < N; ++i) fk[i] = fv[i] = 0.f;
          }
#pragma unroll
          for (int i = 0; i < N; ++i) {
            fk[i] = fmaf(dsc, q[d0 + i], fk[i]);
            fv[i] = fmaf(p, g[d0 + i], fv[i]);
          }
          Vec16<T> wk, wv;
          wk.pack(fk);
          wv.pack(fv);
          wk.store(gk + d0);
          wv.store(gv + d0);
        }
      }
    }
    // dq = scale * sum_j dS_j k_j over this chunk: lanes over feature pairs, keys walked in order
    float2 acc[MAXP];
#pragma unroll
    for (int i = 0; i < MAXP; ++i) acc[i] = make_float2(0.f, 0.f);
    const int kend = min(KC, a.Sk - k0);
#pragma unroll 8
    for (int j = 0; j < kend; ++j) {
      const float dj = __shfl_sync(0xffffffffu, j < 32 ? dsv[0] : dsv[1], j & 31);
      const T* krow = view_ptr<T>(a.k, n, a.n2_count, k0 + j, h);
#pragma unroll
      for (int i = 0; i < MAXP; ++i) {
        const int d = 2 * (lane + 32 * i);
        if (d < Dh) {
          const float2 kv = Pair<T>::ld(krow + d);
          acc[i].x = fmaf(dj, kv.x, acc[i].x);
          acc[i].y = fmaf(dj, kv.y, acc[i].y);
        }
      }
    }
    float* dst = dq32 + ((int64_t)n * a.H + h) * Dh;
#pragma unroll
    for (int i = 0; i < MAXP; ++i) {
      const int d = 2 * (lane + 32 * i);
      if (d < Dh) {
        atomicAdd(dst + d, acc[i].x * a.scale);
        atomicAdd(dst + d + 1, acc[i].y * a.scale);
      }
    }
  }
}

static bool q1_ok(const AttnArgs& a, bool backward) {
  if (a.N <= 0 || a.H <= 0 || a.Sq != 1 || a.Sk <= 0 || a.n2_count <= 0 || a.Dh <= 0 || a.Dh % 8 || a.Dh > 64 * MAXP) return false;
  if (a.qcos || a.key_mask || a.causal || a.kx.p || a.N > 65535) return false;
  const HeadView* vs[4] = {&a.q, &a.k, &a.v, &a.o};
  for (const HeadView* v : vs)
    if (!v->p || (reinterpret_cast<uintptr_t>(v->p) & 15) || v->ss % 8 || v->sh % 8 || v->sn1 % 8 || v->sn2 % 8) return false;
  if (backward) {
    const HeadView* gs[3] = {&a.dk, &a.dv, &a.out_saved};
    for (const HeadView* v : gs)
      if (!v->p || (reinterpret_cast<uintptr_t>(v->p) & 15) || v->ss % 8 || v->sh % 8 || v->sn1 % 8 || v->sn2 % 8) return false;
    if (!a.lse) return false;
  }
  return true;
}

template <typename T>
static int fwd(const AttnArgs& a, float* partial, cudaStream_t st) {
  const int nchunk = (a.Sk + KC - 1) / KC;
  const size_t smem = sizeof(float) * (size_t)a.H * a.Dh;
  if (smem > 48 * 1024) return MEANT_EUNSUPPORTED;
  q1_partial_kernel<T><<<dim3(nchunk, a.N), NW * 32, smem, st>>>(a, partial, nchunk);
  MEANT_LAUNCH_CHECK();
  const int64_t items = (int64_t)a.N * a.H;
  q1_merge_kernel<T><<<(unsigned)((items + 3) / 4), 128, 0, st>>>(a, partial, nchunk);
  MEANT_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int bwd(const AttnArgs& a, float* dq32, int accumulate, cudaStream_t st) {
  const int nchunk = (a.Sk + KC - 1) / KC;
  const size_t smem = sizeof(float) * ((size_t)2 * a.H * a.Dh + a.H);
  if (smem > 48 * 1024) return MEANT_EUNSUPPORTED;
  if (accumulate) q1_bwd_kernel<T, true><<<dim3(nchunk, a.N), NW * 32, smem, st>>>(a, dq32);
  else q1_bwd_kernel<T, false><<<dim3(nchunk, a.N), NW * 32, smem, st>>>(a, dq32);
  MEANT_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------ no-allocation forward
// Scratch for meant_attn_q1_fwd_static: a file-scope device array, so the call needs no caller buffer and no
// allocation (CUDA-graph capture, serving loops that may not call cudaMalloc).  Every GPU that loads the library holds
// these MEANT_Q1_STATIC_FLOATS * 4 bytes (4 MiB) until exit.  The forward's partial pass writes every entry its merge
// pass reads, so nothing has to start at zero; calls that overlap on one GPU share the array and race.
static __device__ __align__(256) float q1_static_scratch[MEANT_Q1_STATIC_FLOATS];

constexpr int Q1_MAX_DEV = 64;
static std::atomic<float*> q1_static_ptr[Q1_MAX_DEV];

// the array's DEVICE address on the current device (in host code its name is only a host-side stand-in): looked up
// once per device with that device current, then cached
static int q1_static_addr(float** out) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= Q1_MAX_DEV) return MEANT_EUNSUPPORTED;
  float* p = q1_static_ptr[dev].load(std::memory_order_acquire);
  if (!p) {
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), q1_static_scratch);
    if (e != cudaSuccess) return (int)e;
    if (!p) return MEANT_EUNSUPPORTED;
    q1_static_ptr[dev].store(p, std::memory_order_release);
  }
  *out = p;
  return 0;
}

}  // namespace q1
}  // namespace meant

using namespace meant;

extern "C" int meant_attn_q1_fwd(const meant_attn_args* args, float* partial, int dtype, meant_stream_t stream) {
  if (!args || !partial || !q1::q1_ok(*args, false)) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MEANT_DT_F32) return q1::fwd<float>(*args, partial, st);
  if (dtype == MEANT_DT_BF16) return q1::fwd<__nv_bfloat16>(*args, partial, st);
  return MEANT_EUNSUPPORTED;
}

extern "C" int meant_attn_q1_bwd(const meant_attn_args* args, float* partial, int accumulate, int dtype,
                                 meant_stream_t stream) {
  if (!args || !partial || !q1::q1_ok(*args, true)) return MEANT_EARG;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MEANT_DT_F32) return q1::bwd<float>(*args, partial, accumulate, st);
  if (dtype == MEANT_DT_BF16) return q1::bwd<__nv_bfloat16>(*args, partial, accumulate, st);
  return MEANT_EUNSUPPORTED;
}

extern "C" int meant_attn_q1_fwd_static(const meant_attn_args* args, int dtype, meant_stream_t stream) {
  if (!args || !q1::q1_ok(*args, false)) return MEANT_EARG;
  const int64_t nchunk = (args->Sk + q1::KC - 1) / q1::KC;
  const int64_t need = (int64_t)args->N * args->H * nchunk * (args->Dh + 2);
  if (need > (int64_t)MEANT_Q1_STATIC_FLOATS) {
    fprintf(stderr, "meant_attn_q1_fwd_static: N * H * ceil(Sk/64) * (Dh + 2) = %lld floats exceeds the static scratch of %d "
            "floats; use meant_attn_q1_fwd with a caller buffer\n", (long long)need, (int)MEANT_Q1_STATIC_FLOATS);
    return MEANT_ETOOBIG;
  }
  float* scratch = nullptr;
  const int rc = q1::q1_static_addr(&scratch);
  if (rc) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (dtype == MEANT_DT_F32) return q1::fwd<float>(*args, scratch, st);
  if (dtype == MEANT_DT_BF16) return q1::fwd<__nv_bfloat16>(*args, scratch, st);
  return MEANT_EUNSUPPORTED;
}
