// CUDA-core attention (fp32 math, fp32 or bf16 storage): the exact path.
// It serves (1) fp32 parity (logits max-abs <= 1e-4 needs true fp32 products, which the tensor
// cores do not offer), (2) shapes the tcgen05 kernels do not take (single-query temporal
// attention meant/temporal.py:34-60, odd head sizes), (3) an on-GPU cross-check of the
// tcgen05 kernels.  Flash-style: K/V streamed through shared memory in 32-key tiles, online
// softmax, nothing of size S^2 is materialised; backward recomputes the scores from the LSE.
#include "attn_common.cuh"
#include <math_constants.h>

namespace meant {

constexpr int QT = 16;   // query rows per CTA (2 per warp, 8 warps)
constexpr int BK = 32;   // keys per shared-memory tile (one per lane)
constexpr int MAXDL = 8; // Dh <= 256

// xv (optional, p != NULL): the shared extra row (meant_attn_args.kx / vx) that stands for row S-1, un-rotated
template <typename T>
__device__ __forceinline__ const T* xrow_ptr(const HeadView& xv, int n, int n2c, int h) {
  return static_cast<const T*>(xv.p) + (int64_t)(n / n2c) * xv.sn1 + (int64_t)h * xv.sh;
}

template <typename T>
__device__ __forceinline__ void load_tile(float* dst, int DhP, const HeadView& v, int n, int n2c, int h, int s0,
                                          int R, int S, int Dh, const float* c, const float* s, int rot_dim,
                                          int pos_is_head, float mul, const HeadView* xv = nullptr) {
  const int half = Dh >> 1;
  const bool has_x = xv != nullptr && xv->p != nullptr;
  for (int idx = threadIdx.x; idx < R * half; idx += blockDim.x) {
    const int r = idx / half, d0 = (idx - r * half) * 2;
    const int srow = s0 + r;
    float x0 = 0.f, x1 = 0.f;
    if (has_x && srow == S - 1) {
      const T* p = xrow_ptr<T>(*xv, n, n2c, h) + d0;
      x0 = to_f(p[0]) * mul;
      x1 = to_f(p[1]) * mul;
    } else if (srow < S) {
      const T* p = view_ptr<T>(v, n, n2c, srow, h) + d0;
      x0 = to_f(p[0]);
      x1 = to_f(p[1]);
      if (c && d0 < rot_dim) rot_pair(x0, x1, c, s, pos_is_head ? h : srow, rot_dim, d0);
      x0 *= mul;
      x1 *= mul;
    }
    dst[r * DhP + d0] = x0;
    dst[r * DhP + d0 + 1] = x1;
  }
}

__device__ __forceinline__ float dot_bcast(const float* a_bcast, const float* b_row, int Dh) {
  float acc = 0.f;
#pragma unroll 8
  for (int d = 0; d < Dh; ++d) acc = fmaf(a_bcast[d], b_row[d], acc);
  return acc;
}

// ------------------------------------------------------------------------------- forward
template <typename T>
__global__ void __launch_bounds__(256) attn_exact_fwd_kernel(AttnArgs a) {
  extern __shared__ float sm[];
  const int Dh = a.Dh, DhP = Dh + 1;
  float* Qs = sm;
  float* Ks = Qs + QT * DhP;
  float* Vs = Ks + BK * DhP;
  float* Ps = Vs + BK * DhP;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q0 = blockIdx.x * QT, h = blockIdx.y, n = blockIdx.z;
  load_tile<T>(Qs, DhP, a.q, n, a.n2_count, h, q0, QT, a.Sq, Dh, a.qcos, a.qsin, a.rot_dim, a.pos_is_head, a.scale);
  float m[2] = {-CUDART_INF_F, -CUDART_INF_F}, l[2] = {0.f, 0.f}, o[2][MAXDL];
#pragma unroll
  for (int rr = 0; rr < 2; ++rr)
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) o[rr][i] = 0.f;
  const int last_q = min(q0 + QT, a.Sq) - 1;
  const int n_tiles = (a.Sk + BK - 1) / BK;
  for (int kt = 0; kt < n_tiles; ++kt) {
    if (a.causal && kt * BK > last_q) break;
    __syncthreads();
    load_tile<T>(Ks, DhP, a.k, n, a.n2_count, h, kt * BK, BK, a.Sk, Dh, a.kcos, a.ksin, a.rot_dim, a.pos_is_head, 1.f, &a.kx);
    load_tile<T>(Vs, DhP, a.v, n, a.n2_count, h, kt * BK, BK, a.Sk, Dh, nullptr, nullptr, 0, 0, 1.f, &a.vx);
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const int r = warp * 2 + rr, qi = q0 + r;
      if (qi >= a.Sq) continue;
      const int kj = kt * BK + lane;
      float s = dot_bcast(Qs + r * DhP, Ks + lane * DhP, Dh);
      const bool valid = kj < a.Sk && (!a.causal || kj <= qi);
      if (a.key_mask && kj < a.Sk) s += (1.0f - a.key_mask[(int64_t)n * a.Sk + kj]) * -1e9f;
      if (!valid) s = -CUDART_INF_F;
      const float m_new = fmaxf(m[rr], warp_max(s));
      float p = 0.f, corr = 1.f;
      if (m_new != -CUDART_INF_F) {
        p = valid ? expf(s - m_new) : 0.f;
        corr = expf(m[rr] - m_new);
      }
      l[rr] = l[rr] * corr + warp_sum(p);
      m[rr] = m_new;
      Ps[r * BK + lane] = p;
      __syncwarp();
#pragma unroll
      for (int i = 0; i < MAXDL; ++i) {
        const int d = lane + 32 * i;
        if (d < Dh) {
          float acc = o[rr][i] * corr;
#pragma unroll 8
          for (int j = 0; j < BK; ++j) acc = fmaf(Ps[r * BK + j], Vs[j * DhP + d], acc);
          o[rr][i] = acc;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = warp * 2 + rr, qi = q0 + r;
    if (qi >= a.Sq) continue;
    const float inv = l[rr] > 0.f ? 1.0f / l[rr] : 0.f;
    T* op = view_ptr<T>(a.o, n, a.n2_count, qi, h);
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) {
      const int d = lane + 32 * i;
      if (d < Dh) op[d] = from_f<T>(o[rr][i] * inv);
    }
    if (lane == 0 && a.lse) a.lse[((int64_t)n * a.H + h) * a.Sq + qi] = m[rr] + logf(l[rr]);
  }
}

// ------------------------------------------------------------------------------- backward: dQ (+ delta)
template <typename T>
__global__ void __launch_bounds__(256) attn_exact_dq_kernel(AttnArgs a) {
  extern __shared__ float sm[];
  const int Dh = a.Dh, DhP = Dh + 1;
  float* Qs = sm;
  float* dOs = Qs + QT * DhP;
  float* Ks = dOs + QT * DhP;
  float* Vs = Ks + BK * DhP;
  float* Ps = Vs + BK * DhP;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int q0 = blockIdx.x * QT, h = blockIdx.y, n = blockIdx.z;
  load_tile<T>(Qs, DhP, a.q, n, a.n2_count, h, q0, QT, a.Sq, Dh, a.qcos, a.qsin, a.rot_dim, a.pos_is_head, a.scale);
  load_tile<T>(dOs, DhP, a.o, n, a.n2_count, h, q0, QT, a.Sq, Dh, nullptr, nullptr, 0, 0, 1.f);
  float lse[2], delta[2], dq[2][MAXDL];
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = warp * 2 + rr, qi = q0 + r;
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) dq[rr][i] = 0.f;
    lse[rr] = 0.f;
    delta[rr] = 0.f;
    if (qi < a.Sq) {
      const T* op = view_ptr<T>(a.out_saved, n, a.n2_count, qi, h);
      const T* gp = view_ptr<T>(a.o, n, a.n2_count, qi, h);
      float acc = 0.f;
      for (int d = lane; d < Dh; d += 32) acc += to_f(op[d]) * to_f(gp[d]);
      delta[rr] = warp_sum(acc);
      const int64_t li = ((int64_t)n * a.H + h) * a.Sq + qi;
      lse[rr] = a.lse[li];
      if (lane == 0) a.delta[li] = delta[rr];
    }
  }
  const int last_q = min(q0 + QT, a.Sq) - 1;
  const int n_tiles = (a.Sk + BK - 1) / BK;
  for (int kt = 0; kt < n_tiles; ++kt) {
    if (a.causal && kt * BK > last_q) break;
    __syncthreads();
    load_tile<T>(Ks, DhP, a.k, n, a.n2_count, h, kt * BK, BK, a.Sk, Dh, a.kcos, a.ksin, a.rot_dim, a.pos_is_head, 1.f, &a.kx);
    load_tile<T>(Vs, DhP, a.v, n, a.n2_count, h, kt * BK, BK, a.Sk, Dh, nullptr, nullptr, 0, 0, 1.f, &a.vx);
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const int r = warp * 2 + rr, qi = q0 + r;
      if (qi >= a.Sq) continue;
      const int kj = kt * BK + lane;
      float s = dot_bcast(Qs + r * DhP, Ks + lane * DhP, Dh);
      const bool valid = kj < a.Sk && (!a.causal || kj <= qi);
      if (a.key_mask && kj < a.Sk) s += (1.0f - a.key_mask[(int64_t)n * a.Sk + kj]) * -1e9f;
      const float p = valid ? expf(s - lse[rr]) : 0.f;
      const float dp = dot_bcast(dOs + r * DhP, Vs + lane * DhP, Dh);
      Ps[r * BK + lane] = p * (dp - delta[rr]);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < MAXDL; ++i) {
        const int d = lane + 32 * i;
        if (d < Dh) {
          float acc = dq[rr][i];
#pragma unroll 8
          for (int j = 0; j < BK; ++j) acc = fmaf(Ps[r * BK + j], Ks[j * DhP + d], acc);
          dq[rr][i] = acc;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = warp * 2 + rr, qi = q0 + r;
    if (qi >= a.Sq) continue;   // warp-uniform
    T* gp = view_ptr<T>(a.dq, n, a.n2_count, qi, h);
    const int pos = a.pos_is_head ? h : qi;
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) {
      const int d = lane + 32 * i;          // pair partner (d ^ 1) lives in lane ^ 1
      float g = dq[rr][i] * a.scale;
      const float partner = __shfl_xor_sync(0xffffffffu, g, 1);
      if (d < Dh) {
        if (a.qcos && d < a.rot_dim) {
          float g0 = (d & 1) ? partner : g, g1 = (d & 1) ? g : partner;
          unrot_pair(g0, g1, a.qcos, a.qsin, pos, a.rot_dim, d & ~1);
          g = (d & 1) ? g1 : g0;
        }
        gp[d] = from_f<T>(g);
      }
    }
  }
}

// ------------------------------------------------------------------------------- backward: dK, dV
template <typename T>
__global__ void __launch_bounds__(256) attn_exact_dkv_kernel(AttnArgs a) {
  extern __shared__ float sm[];
  const int Dh = a.Dh, DhP = Dh + 1;
  float* Kr = sm;
  float* Vr = Kr + QT * DhP;
  float* Qs = Vr + QT * DhP;
  float* dOs = Qs + BK * DhP;
  float* Ps = dOs + BK * DhP;
  float* dSs = Ps + QT * BK;
  float* lse_s = dSs + QT * BK;
  float* delta_s = lse_s + BK;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k0 = blockIdx.x * QT, h = blockIdx.y, n = blockIdx.z;
  load_tile<T>(Kr, DhP, a.k, n, a.n2_count, h, k0, QT, a.Sk, Dh, a.kcos, a.ksin, a.rot_dim, a.pos_is_head, 1.f, &a.kx);
  load_tile<T>(Vr, DhP, a.v, n, a.n2_count, h, k0, QT, a.Sk, Dh, nullptr, nullptr, 0, 0, 1.f, &a.vx);
  float dk[2][MAXDL], dv[2][MAXDL];
#pragma unroll
  for (int rr = 0; rr < 2; ++rr)
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) { dk[rr][i] = 0.f; dv[rr][i] = 0.f; }
  const int n_tiles = (a.Sq + BK - 1) / BK;
  for (int qt = a.causal ? k0 / BK : 0; qt < n_tiles; ++qt) {
    __syncthreads();
    load_tile<T>(Qs, DhP, a.q, n, a.n2_count, h, qt * BK, BK, a.Sq, Dh, a.qcos, a.qsin, a.rot_dim, a.pos_is_head, a.scale);
    load_tile<T>(dOs, DhP, a.o, n, a.n2_count, h, qt * BK, BK, a.Sq, Dh, nullptr, nullptr, 0, 0, 1.f);
    if (threadIdx.x < BK) {
      const int qi = qt * BK + threadIdx.x;
      const int64_t li = ((int64_t)n * a.H + h) * a.Sq + qi;
      lse_s[threadIdx.x] = qi < a.Sq ? a.lse[li] : 0.f;
      delta_s[threadIdx.x] = qi < a.Sq ? a.delta[li] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < 2; ++rr) {
      const int r = warp * 2 + rr, kj = k0 + r;
      if (kj >= a.Sk) continue;
      const int qi = qt * BK + lane;
      float s = dot_bcast(Kr + r * DhP, Qs + lane * DhP, Dh);
      if (a.key_mask) s += (1.0f - a.key_mask[(int64_t)n * a.Sk + kj]) * -1e9f;
      const bool valid = qi < a.Sq && (!a.causal || kj <= qi);
      const float p = valid ? expf(s - lse_s[lane]) : 0.f;
      const float dp = dot_bcast(Vr + r * DhP, dOs + lane * DhP, Dh);
      Ps[r * BK + lane] = p;
      dSs[r * BK + lane] = p * (dp - delta_s[lane]);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < MAXDL; ++i) {
        const int d = lane + 32 * i;
        if (d < Dh) {
          float av = dv[rr][i], ak = dk[rr][i];
#pragma unroll 8
          for (int j = 0; j < BK; ++j) {
            av = fmaf(Ps[r * BK + j], dOs[j * DhP + d], av);
            ak = fmaf(dSs[r * BK + j], Qs[j * DhP + d], ak);
          }
          dv[rr][i] = av;
          dk[rr][i] = ak;
        }
      }
      __syncwarp();
    }
  }
#pragma unroll
  for (int rr = 0; rr < 2; ++rr) {
    const int r = warp * 2 + rr, kj = k0 + r;
    if (kj >= a.Sk) continue;
    if (a.kx.p != nullptr && kj == a.Sk - 1) {
      // the shared extra row: its gradient sums over the n2_count sequences that read it (fp32 atomics, no rotation)
      float* xk = a.dkx + ((int64_t)(n / a.n2_count) * a.H + h) * Dh;
      float* xv = a.dvx + ((int64_t)(n / a.n2_count) * a.H + h) * Dh;
#pragma unroll
      for (int i = 0; i < MAXDL; ++i) {
        const int d = lane + 32 * i;
        if (d < Dh) {
          atomicAdd(xk + d, dk[rr][i]);
          atomicAdd(xv + d, dv[rr][i]);
        }
      }
      continue;
    }
    T* gk = view_ptr<T>(a.dk, n, a.n2_count, kj, h);
    T* gv = view_ptr<T>(a.dv, n, a.n2_count, kj, h);
    const int pos = a.pos_is_head ? h : kj;
#pragma unroll
    for (int i = 0; i < MAXDL; ++i) {
      const int d = lane + 32 * i;
      float g = dk[rr][i];
      const float partner = __shfl_xor_sync(0xffffffffu, g, 1);
      if (d < Dh) {
        if (a.kcos && d < a.rot_dim) {
          float g0 = (d & 1) ? partner : g, g1 = (d & 1) ? g : partner;
          unrot_pair(g0, g1, a.kcos, a.ksin, pos, a.rot_dim, d & ~1);
          g = (d & 1) ? g1 : g0;
        }
        gk[d] = from_f<T>(g);
        gv[d] = from_f<T>(dv[rr][i]);
      }
    }
  }
}

static bool args_ok(const AttnArgs& a) {
  if (a.kx.p != nullptr && (a.vx.p == nullptr || a.causal || a.key_mask || a.pos_is_head || a.Sk < 1)) return false;
  return a.N > 0 && a.H > 0 && a.Sq > 0 && a.Sk > 0 && a.Dh > 0 && (a.Dh % 2 == 0) && a.Dh <= 32 * MAXDL &&
         a.n2_count > 0 && (a.rot_dim % 2 == 0) && a.rot_dim <= a.Dh && a.H <= 65535 && a.N <= 65535;
}

template <typename K>
static int set_smem(K kern, size_t bytes) {
  if (bytes > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return (int)e;
  }
  return 0;
}

template <typename T>
static int run_fwd(const AttnArgs& a, cudaStream_t st) {
  const int DhP = a.Dh + 1;
  const size_t smem = sizeof(float) * ((size_t)(QT + 2 * BK) * DhP + QT * BK);
  if (int e = set_smem(attn_exact_fwd_kernel<T>, smem)) return e;
  dim3 grid((a.Sq + QT - 1) / QT, a.H, a.N);
  attn_exact_fwd_kernel<T><<<grid, 256, smem, st>>>(a);
  MEANT_LAUNCH_CHECK();
  return 0;
}

template <typename T>
static int run_bwd(const AttnArgs& a, cudaStream_t st) {
  const int DhP = a.Dh + 1;
  const size_t smem_q = sizeof(float) * ((size_t)(2 * QT + 2 * BK) * DhP + QT * BK);
  const size_t smem_kv = sizeof(float) * ((size_t)(2 * QT + 2 * BK) * DhP + 2 * QT * BK + 2 * BK);
  if (int e = set_smem(attn_exact_dq_kernel<T>, smem_q)) return e;
  if (int e = set_smem(attn_exact_dkv_kernel<T>, smem_kv)) return e;
  attn_exact_dq_kernel<T><<<dim3((a.Sq + QT - 1) / QT, a.H, a.N), 256, smem_q, st>>>(a);
  MEANT_LAUNCH_CHECK();
  attn_exact_dkv_kernel<T><<<dim3((a.Sk + QT - 1) / QT, a.H, a.N), 256, smem_kv, st>>>(a);
  MEANT_LAUNCH_CHECK();
  return 0;
}

int attn_exact_fwd(const AttnArgs& a, int dtype, cudaStream_t st) {
  if (!args_ok(a) || !a.q.p || !a.k.p || !a.v.p || !a.o.p) return MEANT_EARG;
  if (dtype == MEANT_DT_F32) return run_fwd<float>(a, st);
  if (dtype == MEANT_DT_BF16) return run_fwd<__nv_bfloat16>(a, st);
  return MEANT_EUNSUPPORTED;
}

int attn_exact_bwd(const AttnArgs& a, int dtype, cudaStream_t st) {
  if (!args_ok(a) || !a.q.p || !a.k.p || !a.v.p || !a.o.p || !a.out_saved.p || !a.dq.p || !a.dk.p || !a.dv.p ||
      !a.lse || !a.delta || (a.kx.p != nullptr && (!a.dkx || !a.dvx)))
    return MEANT_EARG;
  if (dtype == MEANT_DT_F32) return run_bwd<float>(a, st);
  if (dtype == MEANT_DT_BF16) return run_bwd<__nv_bfloat16>(a, st);
  return MEANT_EUNSUPPORTED;
}

}  // namespace meant
