/* meant_b200 -- C ABI of the B200 (sm_100a) kernels behind the MEANT encoder hot path.
 *
 * The reference (biirving/meant) is pure PyTorch: it has no FFI for this path.  What a
 * maintainer would bind is therefore the *operator* boundary -- the bodies of the nn.Module
 * forwards listed beside each entry point below (file:line relative to the reference tree).
 * INTEGRATION.md shows the ctypes binding that replaces each reference forward.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer into caller-owned memory; nothing is allocated here;
 *   - `stream` is the CUDA stream to launch on (the caller's current stream);
 *   - return value: 0 = ok, > 0 = cudaError_t of the failed launch, < 0 = argument error
 *     (MEANT_EARG -1 bad argument, MEANT_EUNSUPPORTED -2 unsupported dtype/shape, MEANT_ETOOBIG -3 problem
 *     larger than a fixed-size internal buffer: meant_attn_q1_fwd_static only);
 *     no exception crosses the boundary;
 *   - re-entrant and thread-safe (no global state besides the launch counter and the
 *     one-time cudaFuncSetAttribute calls);
 *   - dtype: MEANT_F32 = 0, MEANT_BF16 = 1 -- the storage type of activations; math is fp32
 *     (tensor-core products bf16 x bf16 -> fp32).  Norm gains/biases and every statistic /
 *     LSE buffer are always fp32.
 */
#ifndef MEANT_B200_H_
#define MEANT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct CUstream_st* meant_stream_t; /* == cudaStream_t */

enum { MEANT_F32 = 0, MEANT_BF16 = 1 };
enum { MEANT_IMPL_AUTO = 0, MEANT_IMPL_EXACT = 1, MEANT_IMPL_TCGEN05 = 2 };

/* ---- library ------------------------------------------------------------------------- */
int meant_version(void);                 /* ABI version (5), bumped on any signature change */
int64_t meant_launch_count(void);        /* kernels launched by this library so far       */
void meant_reset_launch_count(void);

/* ---- RMSNorm: utils/rms_norm.py:40-57  y = gain * x / (||x||_2 d^-1/2 + eps) ----------
 * Fused neighbours (all optional):
 *   res/sum_out : s = x + res is normalised and also written to sum_out
 *                 (meant/meant.py:116,120 residual adds feeding encode2[0] / the next block)
 *   pre_gelu    : s = gelu(x), exact erf form (meant/meant.py:107 GELU -> RMSNorm); not together with
 *                 res (MEANT_EARG): the backward keeps one tensor and could not split the two gradients
 *   drop_p,seed : inverted dropout on y with a counter-based mask (meant/meant.py:105,107); the probability is
 *                 realised as round(256 p) / 256 (one hashed byte per element: exact for the block's p = 0.5) and the
 *                 survivors are scaled by the reciprocal of the realised keep probability;
 *                 seed_dev (optional, device memory) is added to seed when the kernel runs, so a
 *                 captured CUDA graph draws a fresh mask on every replay
 * rinv[rows] receives 1/(rms+eps) for the backward pass.                                  */
int meant_rmsnorm_fwd(const void* x, const void* res, const float* gain, void* y, void* sum_out, float* rinv,
                      int64_t rows, int d, float eps, int pre_gelu, float drop_p, uint64_t seed,
                      const uint64_t* seed_dev, int dtype, meant_stream_t stream);
/* src = the forward's normalised-input source: x (plain), x (pre_gelu: pre-activation) or
 * sum_out (residual form).  dsum_in (optional) is added to dx: the gradient that reaches the
 * same tensor through the residual path.  dgain[d] (fp32) is ACCUMULATED atomically.
 * dxsum[d] (optional, fp32, zero on entry) receives sum_rows dx: when x came out of an nn.Linear this is that
 * layer's bias gradient, produced here instead of by a separate reduction over dx.           */
int meant_rmsnorm_bwd(const void* dy, const void* src, const float* gain, const float* rinv, const void* dsum_in,
                      void* dx, float* dgain, int64_t rows, int d, float eps, int pre_gelu, float drop_p,
                      uint64_t seed, const uint64_t* seed_dev, float* dxsum, int dtype, meant_stream_t stream);

/* ---- LayerNorm: nn.LayerNorm heads (meant/meant_tweet.py:143, meant/meant_vision.py:143) and
 *      TimeSformer PreNorm (src/meant/timesformer_pytorch.py:17-25) ------------------------ */
int meant_layernorm_fwd(const void* x, const float* gain, const float* bias, void* y, float* mean, float* rstd,
                        int64_t rows, int d, float eps, int dtype, meant_stream_t stream);
int meant_layernorm_bwd(const void* dy, const void* x, const float* gain, const float* mean, const float* rstd,
                        void* dx, float* dgain, float* dbias, int64_t rows, int d, int dtype,
                        meant_stream_t stream);

/* ---- attention core --------------------------------------------------------------------
 * Replaces the body of xPosAttention.forward between the projections and multi_mad
 * (meant/xPosAttention.py:39-65), attention.forward (meant/attention.py:39-59), the flash
 * variants (meant/xPosAttention_flash.py:40-43, meant/flash_attention.py:40-45: pos_is_head=1),
 * temporal.forward (meant/temporal.py:43-58: Sq=1) and TimeSformer's grouped attn
 * (src/meant/timesformer_pytorch.py:78-87,119-139).
 *
 * A head view addresses element (n, s, h, d) at  p + n1*sn1 + n2*sn2 + s*ss + h*sh + d  with
 * n = n1*n2_count + n2 and strides in ELEMENTS, so `b s (h d)` projections, a fused
 * [tokens, 3*dim] QKV buffer and the TimeSformer '(b n) f d' / '(b f) n d' groups
 * (src/meant/timesformer_pytorch.py:122: n1 = b, n2 = patch or frame) are all read and written in
 * place: meant_b200/timesformer_pytorch.py passes such views, no rearranged copy is made.   */
typedef struct {
  void* p;
  int64_t sn1, sn2, ss, sh;
} meant_head_view;

typedef struct {
  meant_head_view q, k, v, o;   /* o: forward output; backward: dO                         */
  meant_head_view dq, dk, dv;   /* backward outputs                                        */
  meant_head_view out_saved;    /* backward: the forward output                            */
  float* lse;                   /* [N, H, Sq] log-sum-exp of the scaled, masked scores     */
  float* delta;                 /* [N, H, Sq] backward scratch: rowsum(dO * O)             */
  const float* qcos;            /* [P, rot_dim] cos*xpos_scale; NULL = no rotation         */
  const float* qsin;            /* [P, rot_dim] sin*xpos_scale                             */
  const float* kcos;            /* [P, rot_dim] cos/xpos_scale                             */
  const float* ksin;            /* [P, rot_dim] sin/xpos_scale                             */
  const float* key_mask;        /* [N, Sk] 1 = keep, adds (1-m)*-1e9; NULL = none          */
  int N, n2_count, H, Sq, Sk, Dh, rot_dim;
  int pos_is_head;              /* rotary position = head index (reference flash variants) */
  int causal;
  float scale;                  /* softmax scale applied to q k^T                          */
  meant_head_view q_rot, k_rot; /* optional (p != NULL): Rq q and Rk k written by meant_attn_prerotate; the
                                 * multi-tile kernels (Sq or Sk > 128) then read these instead of rotating every
                                 * tile on the way in (a K tile is needed by every query tile)              */
  /* Shared extra key/value row -- the TimeSformer cls token, which every time / space group attends to
   * un-rotated (src/meant/timesformer_pytorch.py:125-133; the reference `repeat`s and `cat`s it).  With
   * kx.p != NULL the LAST key (index Sk-1) of every sequence n = n1*n2_count + n2 is the row kx / vx (n1, h, :)
   * at p + n1*sn1 + h*sh (sn2, ss ignored) and k / v hold the other Sk-1 rows; the rotary tables still have Sk
   * rows (row Sk-1 is not applied).  Backward: the gradient of that row, summed over the n2_count sequences
   * sharing it, is ACCUMULATED atomically into dkx / dvx: fp32 [N / n2_count, H, Dh], zero on entry.
   * Not combined with causal, key_mask, pos_is_head or q_rot / k_rot.                                       */
  meant_head_view kx, vx;
  float* dkx;
  float* dvx;
} meant_attn_args;

/* Rotates q -> q_rot and k -> k_rot (caller-owned buffers, any head-view strides) with the qcos.. tables of `args`:
 * the per-position pair rotation of meant/rotary_embedding_torch.py:37-44,96-110 applied once per element.
 * bf16 only.  Optional: without it (q_rot.p == NULL) the attention kernels rotate in flight.  IN PLACE when q_rot and
 * k_rot are the very views q and k (same pointer and strides): only the rotated columns are read and written, and the
 * caller then passes q_rot = q, k_rot = k to meant_attn_fwd / meant_attn_bwd (which un-rotate dq / dk as always).   */
int meant_attn_prerotate(const meant_attn_args* args, int dtype, meant_stream_t stream);

int meant_attn_fwd(const meant_attn_args* args, int dtype, int impl, meant_stream_t stream);
int meant_attn_bwd(const meant_attn_args* args, int dtype, int impl, meant_stream_t stream);
/* 1 if the tcgen05 kernel takes this problem (bf16, Dh in {64, 96}, ...), else 0            */
int meant_attn_tcgen05_supported(const meant_attn_args* args, int dtype, int backward);
/* diagnostics: single-tile (Sq, Sk <= 128) calls whose views the TMA tensor maps could not describe and that
 * therefore ran on the multi-tile kernels (0 on every path the models use)                   */
int64_t meant_attn_tma_fallbacks(void);

/* ---- single-query attention (Sq = 1): temporal.forward (meant/temporal.py:43-58: the target day over the lag
 *      window) and the TimeSformer cls query, which sees every token (src/meant/timesformer_pytorch.py:116-119).
 *      No rotation, no masks.  The keys are split into chunks of 64 across CTAs: `partial` is caller-owned scratch,
 *      fp32 [N, H, ceil(Sk/64), Dh + 2]; forward = partial pass + merge (2 launches), writes o (1 row) and lse.
 *      Backward reads q, k, v, dO (args->o), out_saved, lse; writes dq (1 row; via `partial`, fp32 [N, H, Dh], zero
 *      on entry) and dk, dv for every key -- ADDED to what dk / dv already hold when accumulate != 0 (the keys
 *      also receive gradients from the group attention of the same module).                                   */
int meant_attn_q1_fwd(const meant_attn_args* args, float* partial, int dtype, meant_stream_t stream);
int meant_attn_q1_bwd(const meant_attn_args* args, float* partial, int accumulate, int dtype, meant_stream_t stream);
/* meant_attn_q1_fwd without the caller-owned scratch and without any allocation.  NO COUNTERPART IN THE REFERENCE:
 * added for callers that cannot call cudaMalloc (code under CUDA-graph
 * capture, tight serving loops, harnesses that forbid it).  Same inputs, outputs and results as meant_attn_q1_fwd;
 * the partial (max, sum, sum p v) records live in a file-scope `static __device__` array of
 * MEANT_Q1_STATIC_FLOATS fp32 (4 MiB, held by every GPU that loads the library until exit).  Problems with
 * N * H * ceil(Sk/64) * (Dh + 2) > MEANT_Q1_STATIC_FLOATS are rejected with MEANT_ETOOBIG (-3) and a message on
 * stderr; nothing is launched.  The call itself calls no cudaMalloc / cudaFree / allocating routine (the array's
 * device address is looked up once per device with cudaGetSymbolAddress and cached).  Every call on a GPU shares the
 * array: calls that overlap on that GPU (different streams, or graphs replayed concurrently) RACE -- serialise them.
 * Nothing has to start at zero (the partial pass writes every record the merge pass reads).                  */
#define MEANT_Q1_STATIC_FLOATS (1 << 20)
int meant_attn_q1_fwd_static(const meant_attn_args* args, int dtype, meant_stream_t stream);

/* ---- mean over the token axis: meant/meant.py:231, meant/meant_tweet.py:157 ------------- */
int meant_meanpool_fwd(const void* x, void* y, int64_t N, int S, int D, int dtype, meant_stream_t stream);
int meant_meanpool_bwd(const void* dy, void* dx, int64_t N, int S, int D, int dtype, meant_stream_t stream);

/* ---- GEGLU: src/meant/timesformer_pytorch.py:59-62   y = x[:, :F] * gelu(x[:, F:]) ------ */
int meant_geglu_fwd(const void* x, void* y, int64_t rows, int F, int dtype, meant_stream_t stream);
int meant_geglu_bwd(const void* dy, const void* x, void* dx, int64_t rows, int F, int dtype,
                    meant_stream_t stream);

/* ---- time token-shift: src/meant/timesformer_pytorch.py:29-55 on x[B, 1+f*n, d];
 *      transpose != 0 applies the adjoint (backward) ---------------------------------------- */
int meant_token_shift(const void* x, void* y, int B, int f, int n, int d, int transpose, int dtype,
                      meant_stream_t stream);

/* ---- bias gradient of nn.Linear (the q/k/v, multi_mad and MLP Linears of meant/xPosAttention.py:24-28,
 *      meant/meant.py:61-66):  db[c] (fp32, overwritten) = sum_r dy[r*ld + c] ------------------------- */
int meant_bias_grad(const void* dy, float* db, int64_t rows, int cols, int64_t ld, int dtype,
                    meant_stream_t stream);

/* ---- token embedding (the `embedding` module of the model constructors, meant/meant_tweet.py:152 `mod(words)`;
 *      nn.Embedding semantics without padding_idx / max_norm): out[t, :] = W[ids[t], :] cast to `dtype`;
 *      backward: dW (fp32 [vocab, D], overwritten) = scatter-add of dy rows.  Ids outside [0, vocab) read as a
 *      zero row and contribute nothing (torch raises a device assert instead). ------------------------------ */
int meant_embedding_fwd(const int64_t* ids, const float* W, void* out, int64_t tokens, int D, int64_t vocab,
                        int dtype, meant_stream_t stream);
int meant_embedding_bwd(const int64_t* ids, const void* dy, float* dW, int64_t tokens, int D, int64_t vocab,
                        int dtype, meant_stream_t stream);

/* ---- patchify, the im2col operand of the patch-embedding GEMM: 'b c (h p1) (w p2) -> b (h w) (p1 p2 c)'
 *      (meant/meant.py:193-195; src/meant/timesformer_pytorch.py:219-220 per frame) fused with the cast to out_dtype.
 *      in [B, C, H, W] -> out [B, (H/P)(W/P), P*P*C]; transpose != 0: the adjoint, in = patches, out = image layout. -- */
int meant_patchify(const void* in, void* out, int64_t B, int C, int H, int W, int P, int in_dtype, int out_dtype,
                   int transpose, meant_stream_t stream);

/* ---- the BERTweet / RoBERTa embedding block the reference hands to its text models as `embedding`
 *      (src/in_loop_train.py:440,486-496 `embedding=bertweet.embeddings`, applied at meant/meant.py:211-212):
 *        pos = cumsum(ids != pad) * (ids != pad) + pad;  out = Dropout(LayerNorm(word[ids] + position[pos] + type[0]))
 *      ids [N, S]; word / pos / type tables and LayerNorm parameters fp32; out [N, S, d] in `dtype`; pos_ids, mean, rstd
 *      [N, S] are kept for the backward pass, which scatter-adds into dword [vocab, d], dpos [max_pos, d] and accumulates
 *      dtype_row, dln_w, dln_b [d] (all fp32, zero on entry).  d = 256, 512, 768 or 1024.  drop_p / seed / seed_dev as in
 *      meant_rmsnorm_fwd.  Ids outside [0, vocab) produce LayerNorm(0) rows and receive no gradient. --------------- */
int meant_embed_ln_fwd(const int64_t* ids, const float* word, const float* pos, const float* type, const float* ln_w,
                       const float* ln_b, float eps, void* out, int* pos_ids, float* mean, float* rstd, int64_t N, int S,
                       int d, int64_t vocab, int max_pos, int pad_idx, float drop_p, uint64_t seed,
                       const uint64_t* seed_dev, int dtype, meant_stream_t stream);
int meant_embed_ln_bwd(const void* dy, const int64_t* ids, const int* pos_ids, const float* mean, const float* rstd,
                       const float* word, const float* pos, const float* type, const float* ln_w, float* dword,
                       float* dpos, float* dtype_row, float* dln_w, float* dln_b, int64_t N, int S, int d, int64_t vocab,
                       int max_pos, float drop_p, uint64_t seed, const uint64_t* seed_dev, int dtype,
                       meant_stream_t stream);

/* ---- nn.Linear forward as a hand-written tcgen05 GEMM (meant/meant.py:61-66, meant/xPosAttention.py:24-28):
 *      C[M, N] (bf16) = A[M, K] (bf16) * W[N, K]^T (bf16, nn.Linear's weight layout) + bias[N] (fp32, optional).
 *      lda / ldw / ldc: row pitches in elements.  N and K multiples of 8; pointers 16-byte aligned. ---------- */
int meant_gemm_bf16_tn(const void* A, const void* W, const float* bias, void* C, int64_t M, int N, int K,
                       int64_t lda, int64_t ldw, int64_t ldc, meant_stream_t stream);
/* The other forms of the same Linear layers (backward of meant/meant.py:101-107, meant/xPosAttention.py:24-32), same
 * pipeline.  C[M, N] (bf16) = op(A) op(B) (+ bias[N], fp32) (+ residual[M, N], bf16; row pitch ldr):
 *   a_mn = 0: A is [M, K] row-major (lda >= K);   a_mn = 1: A is stored [K, M] row-major (lda >= M; M % 8 == 0)
 *   b_mn = 0: B is [N, K] row-major (C = A B^T: nn.Linear forward);  b_mn = 1: B is stored [K, N] (C = A B: dX = dY W)
 * Nothing is transposed in memory: MN-major operands are read with 64 x 64 TMA boxes and MN-major MMA descriptors. */
int meant_gemm_bf16(const void* A, const void* B, void* C, int64_t M, int N, int K, int64_t lda, int64_t ldb,
                    int64_t ldc, int a_mn, int b_mn, const float* bias, const void* residual, int64_t ldr,
                    meant_stream_t stream);
/* Weight and bias gradient of a Linear: dW[N, K] (fp32) += dY[T, N]^T X[T, K] and, when db != NULL, db[N] (fp32) +=
 * column sums of dY (from an extra MMA column against a tile of ones: no pass over dY).  ACCUMULATES: zero dW / db
 * first for a plain gradient.  Split over the T tokens with TMA reduce-add of fp32 partial tiles: the summation
 * order is not fixed (results can differ in the last bits between runs).                                        */
int meant_gemm_bf16_wgrad(const void* dY, const void* X, float* dW, float* db, int64_t T, int N, int K, int64_t lddy,
                          int64_t ldx, int64_t lddw, meant_stream_t stream);

/* ---- the step after backward: src/in_loop_train.py:237-244 (CrossEntropyLoss on the sigmoid outputs, clip_grad_norm_(1.0),
 *      AdamW: :568), as multi-tensor kernels.  A launch takes a BY-VALUE table of up to MEANT_OPT_MAX_TENSORS fp32 tensors
 *      cut into MEANT_OPT_CHUNK-element chunks: block i of the launch works on chunk block_chunk[i] of tensor
 *      block_tensor[i] (n_blocks <= MEANT_OPT_MAX_BLOCKS); the caller issues as many launches as the parameter list needs. */
#define MEANT_OPT_CHUNK 65536
#define MEANT_OPT_MAX_TENSORS 48
#define MEANT_OPT_MAX_BLOCKS 320
typedef struct {
  void* p[MEANT_OPT_MAX_TENSORS];       /* parameters (fp32, updated in place); unused by meant_multi_sqnorm   */
  void* g[MEANT_OPT_MAX_TENSORS];       /* gradients: fp32, or bf16 when g_bf16 != 0 (bf16 data-parallel buckets) */
  void* m[MEANT_OPT_MAX_TENSORS];       /* exp_avg (fp32)                                                      */
  void* v[MEANT_OPT_MAX_TENSORS];       /* exp_avg_sq (fp32)                                                   */
  void* shadow[MEANT_OPT_MAX_TENSORS];  /* optional bf16 copy of the updated parameter (NULL: none)            */
  int64_t numel[MEANT_OPT_MAX_TENSORS];
  int32_t block_chunk[MEANT_OPT_MAX_BLOCKS];
  uint8_t block_tensor[MEANT_OPT_MAX_BLOCKS];
  int32_t n_tensors, n_blocks;
  int32_t g_bf16, reserved_;
} meant_tensor_table;

/* *sumsq (device, fp32) += sum over the table's gradients of g^2: the squared total norm of clip_grad_norm_.          */
int meant_multi_sqnorm(const meant_tensor_table* table, float* sumsq, meant_stream_t stream);
/* torch.optim.AdamW (decoupled weight decay, bias correction) on every tensor of the table.  *step (device) is the 1-based
 * step count t; sumsq (optional, device) is the finished squared gradient norm: gradients are used as
 * g * min(1, max_norm / (sqrt(*sumsq) + 1e-6)), i.e. clip_grad_norm_(max_norm) without a pass that rescales them.     */
int meant_adamw_step(const meant_tensor_table* table, const float* step, const float* sumsq, float lr, float beta1,
                     float beta2, float eps, float weight_decay, float max_norm, meant_stream_t stream);
/* *loss = mean_i CrossEntropy(probs[i, :], target[i]) over B rows of C values (the reference feeds the model's sigmoid
 * outputs to nn.CrossEntropyLoss, src/in_loop_train.py:237); dprobs (optional) = grad_scale * d loss / d probs.        */
int meant_prob_ce(const float* probs, const int64_t* target, float* loss, float* dprobs, int B, int C, float grad_scale,
                  meant_stream_t stream);

/* ---- softmax-weighted pooling head of the `src` tree (src/meant/meant_tweet.py:173,259-261 lang_prep;
 *      src/meant/meant_timesformer.py:274-279,335-345 lang_prep / image_prep):
 *        w = Softmax(dim=2)(Linear(d,1)(GELU(LayerNorm(h)))),  out[n, :] = sum_t w[n, t] x[n, t, :],  h = Linear(d,d)(x)
 *      x, h: [N, S, d] (the first Linear is the caller's GEMM); prob [N, S] receives w for the backward pass.
 *      degenerate != 0: Softmax over an axis of length one (what the reference's image_prep does on its 3-D input):
 *      every weight is 1, out = sum_t x[n, t, :], h and the parameters are not read.
 *      Backward: dx is the gradient through the weighted sum only (the caller adds the Linear(d,d) path);
 *      dln_w, dln_b, dw2 [d] fp32 are accumulated atomically (zero on entry).  d = 256, 512, 768 or 1024. --------- */
int meant_softmax_pool_fwd(const void* x, const void* h, const float* ln_w, const float* ln_b, float ln_eps,
                           const float* w2, float b2, void* out, float* prob, int64_t N, int S, int d,
                           int degenerate, int dtype, meant_stream_t stream);
int meant_softmax_pool_bwd(const void* dout, const void* x, const void* h, const float* ln_w, const float* ln_b,
                           float ln_eps, const float* w2, const float* prob, void* dx, void* dh, float* dln_w,
                           float* dln_b, float* dw2, int64_t N, int S, int d, int degenerate, int dtype,
                           meant_stream_t stream);

/* ---- token cross entropy of the MLM head (pretrain_mlm.py:154,182: nn.CrossEntropyLoss()(out.view(-1, vocab),
 *      target.view(-1)), ignore_index): *loss_sum += sum over counted rows of (logsumexp - logit[target]), *count +=
 *      counted rows (both fp32 device scalars, zero on entry; loss = loss_sum / count), lse[rows] kept for the backward
 *      pass, which writes dlogits = *dloss / *count * (softmax - onehot) (0 for ignored rows; may alias logits).
 *      ld: row pitch in elements. ------------------------------------------------------------------------------ */
int meant_token_ce_fwd(const void* logits, const int64_t* target, float* lse, float* loss_sum, float* count,
                       int64_t rows, int64_t V, int64_t ld, int64_t ignore_index, int dtype, meant_stream_t stream);
int meant_token_ce_bwd(const void* logits, const int64_t* target, const float* lse, const float* count,
                       const float* dloss, void* dlogits, int64_t rows, int64_t V, int64_t ld, int64_t ignore_index,
                       int dtype, meant_stream_t stream);

/* ---- tcgen05 self-test: D[128,N] (fp32) = op(A) op(B)^T for one bf16 tile; a_mn / b_mn select
 *      MN-major operands (A given as [K,128], B as [K,N]) instead of K-major ([128,K], [N,K]).
 *      Exists so the tests can pin the descriptor conventions the attention kernels rely on. */
int meant_tc_probe(const void* A, const void* B, float* D, int N, int K, int a_mn, int b_mn,
                   meant_stream_t stream);

/* Same tile with SWIZZLED operand sub-tiles (64-column SWIZZLE_128B, or one 32-column SWIZZLE_64B), the
 * layout the attention kernels fill with cp.async; an MN-major B wider than a sub-tile is issued per sub-tile. */
int meant_tc_probe_sw(const void* A, const void* B, float* D, int N, int K, int a_mn, int b_mn,
                      meant_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* MEANT_B200_H_ */
