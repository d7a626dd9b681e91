// RMSNorm (reference semantics: utils/rms_norm.py:40-57, eps added to the RMS) and LayerNorm,
// forward + backward, with the memory-bound neighbours of the MEANT encoder block fused in:
//   * residual add before the norm          (meant/meant.py:116 `inter + input` -> encode2[0])
//   * exact GELU before the norm            (meant/meant.py:107 GELU -> RMSNorm)
//   * dropout after the norm                (meant/meant.py:105,107 Dropout(p), Dropout(0.5))
// One warp per row, 128-bit coalesced loads, the row kept in registers (single HBM read),
// warp-shuffle reductions, fp32 math.  HBM-bound: algorithmic bytes are listed in DESIGN.md.
#include "common.cuh"
#include <algorithm>

namespace meant {

struct NormArgs {
  const void* x;      // [rows, d] input (pre-GELU input when pre_gelu)
  const void* res;    // optional residual, added before the norm
  const float* gain;  // [d] fp32
  const float* bias;  // [d] fp32 (LayerNorm only)
  void* y;            // [rows, d]
  void* sum_out;      // optional: x + res (kept for the next residual / backward)
  float* stat0;       // rms: 1/(rms+eps)   ln: mean
  float* stat1;       // ln: rstd
  int64_t rows;
  int d;
  float eps;
  int pre_gelu;
  uint32_t drop_thresh;  // round(256 p) in each byte (drop_mask4); 0 = no dropout
  float drop_inv_keep;
  uint64_t seed;
  const unsigned long long* seed_dev;   // optional device-resident addend (CUDA-graph replays advance it)
};

// ------------------------------------------------------------------ forward, row in registers
// RES: a residual is added (its two rows of prefetched vectors cost 24 registers: without them a fifth CTA fits an SM)
template <typename T, int CHUNKS, bool LN, bool RES>
__global__ void __launch_bounds__(128, (RES || CHUNKS > 3) ? 4 : 5) norm_fwd_kernel(NormArgs a) {
  uint64_t seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  constexpr int V = Vec16<T>::N;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* __restrict__ x = static_cast<const T*>(a.x);
  const T* __restrict__ res = static_cast<const T*>(a.res);
  T* __restrict__ y = static_cast<T*>(a.y);
  T* __restrict__ so = static_cast<T*>(a.sum_out);
  float g[CHUNKS][V], bb[CHUNKS][V];
#pragma unroll
  for (int c = 0; c < CHUNKS; ++c) {
    const int col = (lane + 32 * c) * V;
#pragma unroll
    for (int i = 0; i < V; ++i) {
      g[c][i] = a.gain[col + i];
      bb[c][i] = LN ? a.bias[col + i] : 0.f;
    }
  }
  const float inv_d = 1.0f / (float)a.d;
  // The next row's loads are issued before this row is touched: a warp always has two rows of requests in flight, which
  // is what keeps HBM busy while the GELU / dropout arithmetic of the current row runs (the kernel was latency-bound
  // with one row in flight per warp: 20 warps x 1.5 KB per SM against the ~27 KB Little's law asks for).
  const int64_t stride = (int64_t)gridDim.x * 4;
  Vec16<T> t[CHUNKS], r[CHUNKS];
  auto fetch = [&](int64_t row_, Vec16<T>* tt, Vec16<T>* rr) {
    const int64_t b_ = row_ * a.d;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) tt[c].load(x + b_ + (lane + 32 * c) * V);
    if (RES) {
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) rr[c].load(res + b_ + (lane + 32 * c) * V);
    }
  };
  const int64_t row0 = (int64_t)blockIdx.x * 4 + warp;
  if (row0 < a.rows) fetch(row0, t, r);
  for (int64_t row = row0; row < a.rows; row += stride) {
    const int64_t base = row * a.d;
    float v[CHUNKS][V];
    Vec16<T> tn[CHUNKS], rn[CHUNKS];
    if (row + stride < a.rows) fetch(row + stride, tn, rn);
    float ss = 0.f, sm = 0.f;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      t[c].unpack(v[c]);
      if (a.pre_gelu) {
#pragma unroll
        for (int i = 0; i < V; ++i) v[c][i] = gelu_f(v[c][i]);
      }
      if (RES) {
        float rv[V];
        r[c].unpack(rv);
#pragma unroll
        for (int i = 0; i < V; ++i) v[c][i] += rv[i];
        if (so) {
          // the sum is rounded to T once; the norm then uses the rounded value so that
          // backward (which re-reads sum_out) sees exactly the forward's input
          Vec16<T> o;
          o.pack(v[c]);
          o.store(so + base + (lane + 32 * c) * V);
          o.unpack(v[c]);
        }
      }
#pragma unroll
      for (int i = 0; i < V; ++i) { ss += v[c][i] * v[c][i]; sm += v[c][i]; }
    }
    float mean = 0.f, rinv;
    if (LN) {
      mean = warp_sum(sm) * inv_d;
      float var = 0.f;
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
        for (int i = 0; i < V; ++i) { const float dl = v[c][i] - mean; var += dl * dl; }
      rinv = rsqrtf(warp_sum(var) * inv_d + a.eps);
      if (lane == 0) { a.stat0[row] = mean; a.stat1[row] = rinv; }
    } else {
      ss = warp_sum(ss);
      rinv = 1.0f / (sqrtf(ss) * sqrtf(inv_d) + a.eps);   // ||x|| * d^-1/2 + eps
      if (lane == 0) a.stat0[row] = rinv;
    }
    // inverted dropout: the 1/(1-p) rides in the row factor, the mask is a select (LayerNorm is never dropped)
    const float rinv_o = (!LN && a.drop_thresh) ? rinv * a.drop_inv_keep : rinv;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      float o[V];
      const int col = (lane + 32 * c) * V;
#pragma unroll
      for (int i = 0; i < V; ++i) o[i] = LN ? (v[c][i] - mean) * rinv * g[c][i] + bb[c][i] : g[c][i] * (v[c][i] * rinv_o);
      if (!LN && a.drop_thresh) {
#pragma unroll
        for (int i = 0; i < V; i += 4) {
          const uint32_t m4 = drop_mask4(seed, (uint64_t)(base + col + i) >> 2, a.drop_thresh);
#pragma unroll
          for (int j = 0; j < 4; ++j) o[i + j] = drop_sel(o[i + j], m4, j);
        }
      }
      Vec16<T> ov;
      ov.pack(o);
      ov.store(y + base + col);
    }
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      t[c] = tn[c];
      if (RES) r[c] = rn[c];
    }
  }
}

// ------------------------------------------------------------------ forward, any d (two passes)
template <typename T, bool LN>
__global__ void __launch_bounds__(128) norm_fwd_generic_kernel(NormArgs a) {
  uint64_t seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* __restrict__ x = static_cast<const T*>(a.x);
  const T* __restrict__ res = static_cast<const T*>(a.res);
  T* __restrict__ y = static_cast<T*>(a.y);
  T* __restrict__ so = static_cast<T*>(a.sum_out);
  const float inv_d = 1.0f / (float)a.d;
  for (int64_t row = (int64_t)blockIdx.x * 4 + warp; row < a.rows; row += (int64_t)gridDim.x * 4) {
    const int64_t base = row * a.d;
    auto val = [&](int c) -> float {
      float v = to_f(x[base + c]);
      if (a.pre_gelu) v = gelu_f(v);
      if (res) { v += to_f(res[base + c]); if (so) v = to_f(from_f<T>(v)); }
      return v;
    };
    float ss = 0.f, sm = 0.f;
    for (int c = lane; c < a.d; c += 32) { const float v = val(c); ss += v * v; sm += v; }
    float mean = 0.f, rinv;
    if (LN) {
      mean = warp_sum(sm) * inv_d;
      float var = 0.f;
      for (int c = lane; c < a.d; c += 32) { const float dl = val(c) - mean; var += dl * dl; }
      rinv = rsqrtf(warp_sum(var) * inv_d + a.eps);
      if (lane == 0) { a.stat0[row] = mean; a.stat1[row] = rinv; }
    } else {
      rinv = 1.0f / (sqrtf(warp_sum(ss)) * sqrtf(inv_d) + a.eps);
      if (lane == 0) a.stat0[row] = rinv;
    }
    for (int c = lane; c < a.d; c += 32) {
      const float v = val(c);
      if (res && so) so[base + c] = from_f<T>(v);
      float o = LN ? (v - mean) * rinv * a.gain[c] + a.bias[c] : a.gain[c] * (v * rinv);
      if (a.drop_thresh) o *= drop_keep(seed, (uint64_t)(base + c), a.drop_thresh, a.drop_inv_keep);
      y[base + c] = from_f<T>(o);
    }
  }
}

struct NormBwdArgs {
  const void* dy;       // [rows, d]
  const void* src;      // forward input of the norm: x, or pre-GELU u, or sum_out
  const float* gain;
  const float* stat0;
  const float* stat1;
  const void* dsum_in;  // optional gradient arriving at the same tensor from the residual path
  void* dx;             // [rows, d]
  float* dgain;         // [d] fp32, atomically accumulated (caller zeroes)
  float* dbias;         // [d] fp32 (LayerNorm)
  int64_t rows;
  int d;
  float eps;
  int pre_gelu;
  uint32_t drop_thresh;
  float drop_inv_keep;
  uint64_t seed;
  const unsigned long long* seed_dev;   // optional device-resident addend (CUDA-graph replays advance it)
};

// ------------------------------------------------------------------ backward, row in registers
// The row stays PACKED (bf16/fp32 vectors) between the two sweeps, so only the gain, the per-column
// gradient accumulators and (GELU form) the activation / its derivative live in fp32 registers:
// 3-4 CTAs of 4 warps per SM keep >= 36 KB of loads in flight.
// COLSUM (RMSNorm only): also accumulate sum_rows dx into a.dbias -- the bias gradient of the Linear that produced x.
template <typename T, int CHUNKS, bool LN, bool GELU, bool COLSUM = false>
__global__ void __launch_bounds__(128, GELU ? 2 : ((LN || COLSUM) ? 3 : 4)) norm_bwd_kernel(NormBwdArgs a) {
  static_assert(!(LN && COLSUM), "the dbias accumulators serve LayerNorm's bias OR the column sum of dx");
  uint64_t seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  constexpr int V = Vec16<T>::N;
  __shared__ float red[4][32 * CHUNKS * V + 1];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* __restrict__ dy = static_cast<const T*>(a.dy);
  const T* __restrict__ src = static_cast<const T*>(a.src);
  const T* __restrict__ dsi = static_cast<const T*>(a.dsum_in);
  T* __restrict__ dx = static_cast<T*>(a.dx);
  float g[CHUNKS][V], dg[CHUNKS][V], db[CHUNKS][V];
#pragma unroll
  for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
    for (int i = 0; i < V; ++i) { g[c][i] = a.gain[(lane + 32 * c) * V + i]; dg[c][i] = 0.f; db[c][i] = 0.f; }
  const float inv_d = 1.0f / (float)a.d;
  // GELU forms (arithmetic-heavy: erf polynomial, its derivative, the dropout hash): as in the forward kernel, the next
  // row's loads (src, dy, the row statistics) are in flight while this row computes -- measured 195 -> 166 us per launch.
  // The other forms are faster WITHOUT it (the extra registers cost a resident CTA): PF is off for them and the "next"
  // fetch below degenerates to loading the current row at the top of each iteration.
  constexpr bool PF = GELU;
  const int64_t stride = (int64_t)gridDim.x * 4;
  Vec16<T> tu[CHUNKS], tg[CHUNKS], ts[CHUNKS];
  float mean_n = 0.f, rinv_n = 0.f;
  auto fetch = [&](int64_t row_, Vec16<T>* uu, Vec16<T>* gg, float& mn, float& ri) {
    const int64_t b_ = row_ * a.d;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      uu[c].load(src + b_ + (lane + 32 * c) * V);
      gg[c].load(dy + b_ + (lane + 32 * c) * V);
    }
    mn = LN ? a.stat0[row_] : 0.f;
    ri = LN ? a.stat1[row_] : a.stat0[row_];
  };
  const int64_t row0 = (int64_t)blockIdx.x * 4 + warp;
  if (PF && row0 < a.rows) fetch(row0, tu, tg, mean_n, rinv_n);
  for (int64_t row = row0; row < a.rows; row += stride) {
    const int64_t base = row * a.d;
    if (!PF) fetch(row, tu, tg, mean_n, rinv_n);
    const float mean = mean_n, rinv = rinv_n;
    if (dsi) {
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) ts[c].load(dsi + base + (lane + 32 * c) * V);
    }
    Vec16<T> un[PF ? CHUNKS : 1], gn[PF ? CHUNKS : 1];
    if (PF && row + stride < a.rows) fetch(row + stride, un, gn, mean_n, rinv_n);
    float sk[GELU ? CHUNKS : 1][V], gk[GELU ? CHUNKS : 1][V];   // GELU form: activation and its derivative
    float t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      float u[V], gy[V];
      tu[c].unpack(u);
      tg[c].unpack(gy);
      if (a.drop_thresh) {
#pragma unroll
        for (int i = 0; i < V; i += 4) {
          const uint32_t m4 = drop_mask4(seed, (uint64_t)(base + (lane + 32 * c) * V + i) >> 2, a.drop_thresh);
#pragma unroll
          for (int j = 0; j < 4; ++j) gy[i + j] = drop_sel(gy[i + j], m4, j) * a.drop_inv_keep;
        }
        tg[c].pack(gy);     // the second sweep re-reads the masked gradient (exact: 0 or a power-of-two scale for p = 0.5;
                            // otherwise one extra rounding of dy * inv_keep to T, as if dropout's backward had run in T)
      }
#pragma unroll
      for (int i = 0; i < V; ++i) {
        float sv = u[i];
        if (GELU) {
          gelu_val_grad(u[i], sv, gk[c][i]);
          sk[c][i] = sv;
        }
        if (LN) sv = (sv - mean) * rinv;  // xhat
        const float gd = g[c][i] * gy[i];
        t1 = fmaf(gd, sv, t1);
        t2 += gd;
        dg[c][i] = fmaf(gy[i], sv * (LN ? 1.f : rinv), dg[c][i]);
        if (LN) db[c][i] += gy[i];
      }
    }
    t1 = warp_sum(t1);
    float coef, m2 = 0.f;
    if (LN) {
      m2 = warp_sum(t2) * inv_d;
      coef = t1 * inv_d;
    } else {
      const float rms = 1.0f / rinv - a.eps;
      coef = rms > 0.f ? t1 * rinv * rinv * inv_d / rms : 0.f;
    }
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c) {
      float u[V], gy[V], o[V], sv[V];
      tu[c].unpack(u);
      tg[c].unpack(gy);
      if (dsi) ts[c].unpack(sv);
#pragma unroll
      for (int i = 0; i < V; ++i) {
        float s1 = GELU ? sk[c][i] : u[i];
        if (LN) s1 = (s1 - mean) * rinv;
        const float gd = g[c][i] * gy[i];
        float ds = LN ? rinv * (gd - m2 - s1 * coef) : gd * rinv - s1 * coef;
        if (GELU) ds *= gk[c][i];
        if (dsi) ds += sv[i];
        o[i] = ds;
        if (COLSUM) db[c][i] += ds;
      }
      Vec16<T> ov;
      ov.pack(o);
      ov.store(dx + base + (lane + 32 * c) * V);
    }
    if (PF) {
#pragma unroll
      for (int c = 0; c < CHUNKS; ++c) { tu[c] = un[c]; tg[c] = gn[c]; }
    }
  }
  // block-level reduction of the per-warp column sums, then one atomic per column per block
#pragma unroll 1
  for (int pass = 0; pass < ((LN || COLSUM) ? 2 : 1); ++pass) {
    __syncthreads();
#pragma unroll
    for (int c = 0; c < CHUNKS; ++c)
#pragma unroll
      for (int i = 0; i < V; ++i) red[warp][(lane + 32 * c) * V + i] = pass ? db[c][i] : dg[c][i];
    __syncthreads();
    float* dst = pass ? a.dbias : a.dgain;
    for (int col = threadIdx.x; col < a.d; col += 128)
      atomicAdd(dst + col, red[0][col] + red[1][col] + red[2][col] + red[3][col]);
  }
}

// ------------------------------------------------------------------ backward, any d
template <typename T, bool LN>
__global__ void __launch_bounds__(128) norm_bwd_generic_kernel(NormBwdArgs a) {
  uint64_t seed = a.seed;
  if (a.drop_thresh && a.seed_dev) seed += __ldg(a.seed_dev);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const T* __restrict__ dy = static_cast<const T*>(a.dy);
  const T* __restrict__ src = static_cast<const T*>(a.src);
  const T* __restrict__ dsi = static_cast<const T*>(a.dsum_in);
  T* __restrict__ dx = static_cast<T*>(a.dx);
  const float inv_d = 1.0f / (float)a.d;
  for (int64_t row = (int64_t)blockIdx.x * 4 + warp; row < a.rows; row += (int64_t)gridDim.x * 4) {
    const int64_t base = row * a.d;
    const float mean = LN ? a.stat0[row] : 0.f;
    const float rinv = LN ? a.stat1[row] : a.stat0[row];
    auto sval = [&](int c) -> float {
      float v = to_f(src[base + c]);
      if (a.pre_gelu) v = gelu_f(v);
      return LN ? (v - mean) * rinv : v;
    };
    auto gval = [&](int c) -> float {
      float v = to_f(dy[base + c]);
      if (a.drop_thresh) v *= drop_keep(seed, (uint64_t)(base + c), a.drop_thresh, a.drop_inv_keep);
      return v;
    };
    float t1 = 0.f, t2 = 0.f;
    for (int c = lane; c < a.d; c += 32) {
      const float gd = a.gain[c] * gval(c);
      t1 += gd * sval(c);
      t2 += gd;
    }
    t1 = warp_sum(t1);
    float coef, m2 = 0.f;
    if (LN) { m2 = warp_sum(t2) * inv_d; coef = t1 * inv_d; }
    else { const float rms = 1.0f / rinv - a.eps; coef = rms > 0.f ? t1 * rinv * rinv * inv_d / rms : 0.f; }
    for (int c = lane; c < a.d; c += 32) {
      const float s = sval(c), gy = gval(c), gd = a.gain[c] * gy;
      float ds = LN ? rinv * (gd - m2 - s * coef) : gd * rinv - s * coef;
      if (a.pre_gelu) ds *= gelu_grad_f(to_f(src[base + c]));
      if (dsi) ds += to_f(dsi[base + c]);
      dx[base + c] = from_f<T>(ds);
      atomicAdd(a.dgain + c, gy * s * (LN ? 1.f : rinv));
      if (LN) atomicAdd(a.dbias + c, gy);
    }
  }
}

static uint32_t drop_thr8(float p) {   // 8-bit drop threshold round(256 p) (common.cuh: drop_mask4); 0 = dropout off
  if (!(p > 0.f)) return 0u;
  const uint32_t t = (uint32_t)(p * 256.0f + 0.5f);
  return t < 1u ? 1u : (t > 255u ? 255u : t);
}
static uint32_t drop_thr4(float p) { return drop_thr8(p) * 0x01010101u; }
static float drop_scale(float p) { return 256.0f / (256.0f - (float)drop_thr8(p)); }   // 1 / realised keep probability

static int grid_for(int64_t rows) {
  int64_t blocks = (rows + 3) / 4;
  const int64_t cap = 148 * 4;  // persistent: one resident wave of 128-thread CTAs (4 per SM at ~125 registers)
  return (int)(blocks < cap ? (blocks > 0 ? blocks : 1) : cap);
}

template <typename T, bool LN>
static int launch_fwd(const NormArgs& a, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  const int g = grid_for(a.rows);
  const bool vec_ok = (a.d % (32 * V) == 0) &&
                      ((reinterpret_cast<uintptr_t>(a.x) | reinterpret_cast<uintptr_t>(a.y) |
                        reinterpret_cast<uintptr_t>(a.res) | reinterpret_cast<uintptr_t>(a.sum_out)) % 16 == 0);
  const int chunks = vec_ok ? a.d / (32 * V) : 0;
  const bool res = a.res != nullptr;
  const int g5 = (chunks == 3 && !res) ? (int)std::min<int64_t>((a.rows + 3) / 4 > 0 ? (a.rows + 3) / 4 : 1, 148 * 5) : g;   // five CTAs per SM
  switch (chunks) {
    case 3:
      if (res) norm_fwd_kernel<T, 3, LN, true><<<g, 128, 0, st>>>(a);
      else norm_fwd_kernel<T, 3, LN, false><<<g5, 128, 0, st>>>(a);
      break;
    case 6:
      if (res) norm_fwd_kernel<T, 6, LN, true><<<g, 128, 0, st>>>(a);
      else norm_fwd_kernel<T, 6, LN, false><<<g, 128, 0, st>>>(a);
      break;
    case 12:
      if (res) norm_fwd_kernel<T, 12, LN, true><<<g, 128, 0, st>>>(a);
      else norm_fwd_kernel<T, 12, LN, false><<<g, 128, 0, st>>>(a);
      break;
    default: norm_fwd_generic_kernel<T, LN><<<g, 128, 0, st>>>(a); break;
  }
  MEANT_LAUNCH_CHECK();
  return 0;
}

template <typename T, bool LN>
static int launch_bwd(const NormBwdArgs& a, cudaStream_t st) {
  constexpr int V = Vec16<T>::N;
  int g = grid_for(a.rows);
  if (g > 148 * 2) g = 148 * (a.pre_gelu ? 2 : ((LN || a.dbias) ? 3 : 4));  // one resident wave: fewer dgain atomics
  const bool vec_ok = (a.d % (32 * V) == 0) &&
                      ((reinterpret_cast<uintptr_t>(a.dy) | reinterpret_cast<uintptr_t>(a.src) |
                        reinterpret_cast<uintptr_t>(a.dx) | reinterpret_cast<uintptr_t>(a.dsum_in)) % 16 == 0);
  const int chunks = vec_ok ? a.d / (32 * V) : 0;
  // CHUNKS * V <= 24 values per lane keeps the per-lane arrays in registers
  const bool gelu = a.pre_gelu != 0;
  bool done = false;
  if constexpr (!LN) {
    const bool colsum = a.dbias != nullptr;
    if (chunks == 3 && colsum) {
      if (g > 148 * 3) g = 148 * 3;
      if (gelu) norm_bwd_kernel<T, 3, false, true, true><<<g, 128, 0, st>>>(a);
      else      norm_bwd_kernel<T, 3, false, false, true><<<g, 128, 0, st>>>(a);
      done = true;
    } else if (colsum) {
      return MEANT_EUNSUPPORTED;    // the entry point falls back to a separate column-sum launch
    } else if (chunks == 3 && gelu) { norm_bwd_kernel<T, 3, false, true><<<g, 128, 0, st>>>(a); done = true; }
  }
  if (done) {}
  else if (chunks == 3 && !gelu) norm_bwd_kernel<T, 3, LN, false><<<g, 128, 0, st>>>(a);
  else if (chunks == 6 && V == 4 && !gelu) norm_bwd_kernel<T, (V == 4 ? 6 : 3), LN, false><<<g, 128, 0, st>>>(a);
  else norm_bwd_generic_kernel<T, LN><<<g, 128, 0, st>>>(a);
  MEANT_LAUNCH_CHECK();
  return 0;
}

}  // namespace meant

using namespace meant;

extern "C" int meant_rmsnorm_fwd(const void* x, const void* res, const float* gain, void* y, void* sum_out,
                                 float* rinv, int64_t rows, int d, float eps, int pre_gelu, float drop_p,
                                 uint64_t seed, const uint64_t* seed_dev, int dtype, cudaStream_t stream) {
  if (!x || !gain || !y || !rinv || rows < 0 || d <= 0 || drop_p < 0.f || drop_p >= 1.f) return MEANT_EARG;
  // gelu(x) + res: the backward keeps only the sum and could not separate the two gradients (no caller needs it)
  if (pre_gelu && res) return MEANT_EARG;
  if (rows == 0) return 0;
  NormArgs a{x, res, gain, nullptr, y, sum_out, rinv, nullptr, rows, d, eps, pre_gelu,
             drop_thr4(drop_p), drop_scale(drop_p), seed,
             reinterpret_cast<const unsigned long long*>(seed_dev)};
  if (dtype == MEANT_DT_F32) return launch_fwd<float, false>(a, stream);
  if (dtype == MEANT_DT_BF16) return launch_fwd<__nv_bfloat16, false>(a, stream);
  return MEANT_EUNSUPPORTED;
}

extern "C" int meant_rmsnorm_bwd(const void* dy, const void* src, const float* gain, const float* rinv,
                                 const void* dsum_in, void* dx, float* dgain, int64_t rows, int d, float eps,
                                 int pre_gelu, float drop_p, uint64_t seed, const uint64_t* seed_dev, float* dxsum,
                                 int dtype, cudaStream_t stream) {
  if (!dy || !src || !gain || !rinv || !dx || !dgain || rows < 0 || d <= 0) return MEANT_EARG;
  if (pre_gelu && dsum_in) return MEANT_EARG;   // see meant_rmsnorm_fwd
  if (rows == 0) return 0;
  NormBwdArgs a{dy, src, gain, rinv, nullptr, dsum_in, dx, dgain, dxsum, rows, d, eps, pre_gelu,
                drop_thr4(drop_p), drop_scale(drop_p), seed,
                reinterpret_cast<const unsigned long long*>(seed_dev)};
  if (dtype != MEANT_DT_F32 && dtype != MEANT_DT_BF16) return MEANT_EUNSUPPORTED;
  int rc = dtype == MEANT_DT_F32 ? launch_bwd<float, false>(a, stream) : launch_bwd<__nv_bfloat16, false>(a, stream);
  if (rc == MEANT_EUNSUPPORTED && dxsum) {   // row shapes without a fused column sum: one extra streaming pass over dx
    a.dbias = nullptr;
    rc = dtype == MEANT_DT_F32 ? launch_bwd<float, false>(a, stream) : launch_bwd<__nv_bfloat16, false>(a, stream);
    if (rc == 0) rc = meant_bias_grad(dx, dxsum, rows, d, d, dtype, stream);
  }
  return rc;
}

extern "C" int meant_layernorm_fwd(const void* x, const float* gain, const float* bias, void* y, float* mean,
                                   float* rstd, int64_t rows, int d, float eps, int dtype, cudaStream_t stream) {
  if (!x || !gain || !bias || !y || !mean || !rstd || rows < 0 || d <= 0) return MEANT_EARG;
  if (rows == 0) return 0;
  NormArgs a{x, nullptr, gain, bias, y, nullptr, mean, rstd, rows, d, eps, 0, 0u, 1.0f, 0};
  if (dtype == MEANT_DT_F32) return launch_fwd<float, true>(a, stream);
  if (dtype == MEANT_DT_BF16) return launch_fwd<__nv_bfloat16, true>(a, stream);
  return MEANT_EUNSUPPORTED;
}

extern "C" int meant_layernorm_bwd(const void* dy, const void* x, const float* gain, const float* mean,
                                   const float* rstd, void* dx, float* dgain, float* dbias, int64_t rows, int d,
                                   int dtype, cudaStream_t stream) {
  if (!dy || !x || !gain || !mean || !rstd || !dx || !dgain || !dbias || rows < 0 || d <= 0) return MEANT_EARG;
  if (rows == 0) return 0;
  NormBwdArgs a{dy, x, gain, mean, rstd, nullptr, dx, dgain, dbias, rows, d, 0.f, 0, 0u, 1.0f, 0};
  if (dtype == MEANT_DT_F32) return launch_bwd<float, true>(a, stream);
  if (dtype == MEANT_DT_BF16) return launch_bwd<__nv_bfloat16, true>(a, stream);
  return MEANT_EUNSUPPORTED;
}
