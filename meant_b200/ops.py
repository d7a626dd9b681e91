"""torch.autograd wrappers over the C ABI (include/meant_b200.h).

torch is the plumbing here (device memory, streams, autograd graph); every op below runs a
hand-written sm_100a kernel from libmeant_b200.so and raises if that is impossible.
"""
import contextlib
import math
import threading
import weakref

import torch

from . import _native as nv
from ._native import AttnArgs, HeadView


def compute_dtype(x: torch.Tensor) -> torch.dtype:
    """bf16 under torch.autocast('cuda', dtype=torch.bfloat16), else the tensor's dtype."""
    if torch.is_autocast_enabled("cuda"):
        dt = torch.get_autocast_dtype("cuda")
        if dt != torch.bfloat16:
            raise RuntimeError(
                f"meant_b200 computes in bfloat16 or float32; autocast dtype {dt} is not supported "
                "(the reference's fp16 autocast + GradScaler maps to torch.autocast('cuda', torch.bfloat16))")
        return dt
    if x.dtype not in (torch.float32, torch.bfloat16):
        raise RuntimeError(f"meant_b200 computes in bfloat16 or float32, got {x.dtype}")
    return x.dtype


def _prep(x, dtype):
    nv.require_cuda(x)
    return x.to(dtype).contiguous()


def _prep_rows(x, dtype):
    """Like _prep but keeps strided [N, S, W] views whose rows are dense (sliced projections)."""
    if x is None:
        return None
    nv.require_cuda(x)
    x = x.to(dtype)
    return x if x.dim() == 3 and x.stride(-1) == 1 and x.data_ptr() % 16 == 0 else x.contiguous()


def _f32(t):
    return t.detach().float().contiguous()


# ---- operator context ------------------------------------------------------------------------------------------
class OpsContext:
    """State the operator wrappers consult, owned by whoever drives a step (train.TrainStep builds one per instance and
    activates it around its forward + backward with `use`).  Two models or two TrainSteps therefore never share it, and
    a backward pass running on an autograd worker thread sees the context of ITS forward: every autograd Function
    stores `current()` at forward time and reads only that object in backward.

      shadow        {id(param): (param, bf16 copy)} kept current by the optimizer update (or one multi-tensor cast per
                    step); ("packed", id(wq), id(wv), id(wk)) -> (packed weight, packed bias, weights) for the fused
                    q | v | k projection
      arena         zeroed fp32 buffer handed out in slices for the small per-column accumulators of the norm backward
                    kernels (one fill per step instead of ~100); arena_off is the bump pointer, reset by begin_step()
      graph_seed    device int64 added to every dropout seed by the kernels (advanced inside a captured step)
      fused_linear  route ops.linear / ops.linear3 through _LinearFn / _Linear3Fn
      native_gemm   forward + dgrad + wgrad of those through the hand-written tcgen05 GEMMs
      native_wgrad  wgrad + bias gradient in one tcgen05 kernel wherever the bias gradient would need its own pass
      attn_impl, prerotate   None = the module-level defaults DEFAULT_ATTN_IMPL / PREROTATE
    """

    def __init__(self, shadow=None, arena=None, fused_linear=False, native_gemm=None, native_wgrad=None, attn_impl=None,
                 prerotate=None):
        self.shadow = shadow
        self.arena = arena
        self.arena_off = 0
        self.graph_seed = None
        self.fused_linear = bool(fused_linear)
        self._native_gemm = native_gemm
        self._native_wgrad = native_wgrad
        self.attn_impl = attn_impl
        self.prerotate = prerotate
        self._colsum = {}

    # a context without its own choice follows the process-wide default context (use_native_gemm / use_native_wgrad)
    @property
    def native_gemm(self):
        return _default.native_gemm if (self._native_gemm is None and self is not _default) else bool(self._native_gemm)

    @native_gemm.setter
    def native_gemm(self, v):
        self._native_gemm = v

    @property
    def native_wgrad(self):
        if self._native_wgrad is None:
            return True if self is _default else _default.native_wgrad
        return bool(self._native_wgrad)

    @native_wgrad.setter
    def native_wgrad(self, v):
        self._native_wgrad = v

    def begin_step(self):
        self.arena_off = 0
        self._colsum.clear()

    def shadow_of(self, p, dtype):
        if self.shadow is None or p is None:
            return None
        e = self.shadow.get(id(p))
        return e[1] if e is not None and e[0] is p and e[1].dtype == dtype else None

    def zeros(self, n, device):
        if self.arena is not None and self.arena.device == device:
            start = (self.arena_off + 63) // 64 * 64
            if start + n <= self.arena.numel():
                self.arena_off = start + n
                return self.arena[start:start + n]
        return torch.zeros(n, dtype=torch.float32, device=device)

    # The norm backward also sums dx over rows; when that dx is the dy of a Linear, the sum IS the layer's bias gradient.
    # The hand-off is keyed by the tensor object autograd passes on (identity + version), and lives here, not on the tensor.
    def put_colsum(self, dx, colsum):
        self._colsum[id(dx)] = (weakref.ref(dx), dx._version, colsum)

    def take_colsum(self, dy, n_out, dtype):
        e = self._colsum.pop(id(dy), None)
        if e is None or e[0]() is not dy or e[1] != dy._version or e[2].numel() != n_out or dy.dtype != dtype:
            return None
        return e[2]


_default = OpsContext()
_tls = threading.local()


def current():
    """The active context of this thread (the process-wide default one outside any `use`)."""
    return getattr(_tls, "ctx", None) or _default


@contextlib.contextmanager
def use(ctx):
    prev = getattr(_tls, "ctx", None)
    _tls.ctx = ctx
    try:
        yield ctx
    finally:
        _tls.ctx = prev


# process-wide defaults (benchmarks, tests, eager use without a TrainStep) act on the default context
def set_weight_shadows(d):
    _default.shadow = d


def set_zero_arena(buf):
    """buf: fp32 tensor that the caller zeroes at the start of every step (None switches the arena off)."""
    _default.arena, _default.arena_off = buf, 0


def set_graph_seed(t):
    _default.graph_seed = t


# ------------------------------------------------------------------------------ RMSNorm
class _RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, res, gain, eps, pre_gelu, drop_p, seed):
        dtype = compute_dtype(x)
        x2 = _prep(x, dtype)
        d = x2.shape[-1]
        rows = x2.numel() // d
        r2 = _prep(res, dtype) if res is not None else None
        g32 = _f32(gain)
        y = torch.empty_like(x2)
        s = torch.empty_like(x2) if r2 is not None else None
        rinv = torch.empty(rows, dtype=torch.float32, device=x2.device)
        tag = f"rmsnorm_fwd/d{d}/res{int(r2 is not None)}/gelu{int(pre_gelu)}"
        octx = ctx.octx = current()
        sdev = octx.graph_seed if drop_p > 0.0 else None
        nv.check(nv.timed(tag, nv.lib().meant_rmsnorm_fwd, nv.ptr(x2), nv.ptr(r2), nv.ptr(g32), nv.ptr(y), nv.ptr(s),
                          nv.ptr(rinv), rows, d, eps, int(pre_gelu), drop_p, seed, nv.ptr(sdev), nv.dt(x2),
                          nv.stream()), "meant_rmsnorm_fwd")
        ctx.sdev = sdev
        ctx.save_for_backward(s if r2 is not None else x2, g32, rinv)
        ctx.cfg = (eps, int(pre_gelu), drop_p, seed, res is not None, gain.dtype, x.dtype,
                   res.dtype if res is not None else None)
        if r2 is None:
            return y
        return y, s

    @staticmethod
    def backward(ctx, dy, dsum=None):
        src, g32, rinv = ctx.saved_tensors
        eps, pre_gelu, drop_p, seed, has_res, gdt, xdt, rdt = ctx.cfg
        d = src.shape[-1]
        rows = src.numel() // d
        dy = torch.zeros_like(src) if dy is None else dy.to(src.dtype).contiguous()
        ds = dsum.to(src.dtype).contiguous() if dsum is not None else None
        dx = torch.empty_like(src)
        octx = ctx.octx
        dgain = octx.zeros(d, src.device)
        # with the Linear wrapper active the kernel also sums dx over rows: if x came out of a Linear, that is the
        # layer's bias gradient (picked up by _LinearFn.backward through the context's colsum hand-off)
        dxsum = octx.zeros(d, src.device) if octx.fused_linear and xdt == src.dtype else None
        tag = f"rmsnorm_bwd/d{d}/res{int(ds is not None)}/gelu{pre_gelu}" + ("/colsum" if dxsum is not None else "")
        nv.check(nv.timed(tag, nv.lib().meant_rmsnorm_bwd, nv.ptr(dy), nv.ptr(src), nv.ptr(g32), nv.ptr(rinv),
                          nv.ptr(ds), nv.ptr(dx), nv.ptr(dgain), rows, d, eps, pre_gelu, drop_p, seed,
                          nv.ptr(ctx.sdev), nv.ptr(dxsum), nv.dt(src), nv.stream()), "meant_rmsnorm_bwd")
        dxo = dx.to(xdt)
        if dxsum is not None:
            octx.put_colsum(dxo, dxsum)                 # valid only while dx is not modified (version check)
        return dxo, (dx.to(rdt) if has_res else None), dgain.to(gdt), None, None, None, None


def _seed():
    # drawn from torch's CPU generator so torch.manual_seed() makes dropout reproducible
    return int(torch.randint(0, 2 ** 62, (1,)).item())


def rms_norm(x, gain, eps=1e-8, residual=None, pre_gelu=False, drop_p=0.0):
    """y = gain * s / (rms(s) + eps), s = gelu?(x) (+ residual); returns y, or (y, s) when a
    residual is given.  drop_p > 0 applies inverted dropout to y (training only: the caller
    passes 0 in eval mode)."""
    if pre_gelu and residual is not None:
        raise ValueError("rms_norm: pre_gelu together with a residual is not supported (the backward keeps only the "
                         "sum, so the gradients of x and of the residual could not be separated)")
    seed = _seed() if drop_p > 0.0 else 0
    return _RMSNormFn.apply(x, residual, gain, float(eps), bool(pre_gelu), float(drop_p), seed)


# ------------------------------------------------------------------------------- Linear
def _mm_f32(a, b):
    """a @ b with an fp32 result straight from the GEMM epilogue (weight gradients: saves the bf16 -> fp32 pass)."""
    if a.dtype == torch.float32:
        return torch.mm(a, b)
    try:
        return torch.mm(a, b, out_dtype=torch.float32)
    except (TypeError, RuntimeError):
        return torch.mm(a, b).float()


def use_native_gemm(on=True):
    """Forward, dgrad and wgrad GEMMs of the Linear wrappers through the hand-written tcgen05 kernels (meant_gemm_bf16_tn,
    meant_gemm_bf16, meant_gemm_bf16_wgrad) instead of cuBLASLt, for every context that has no choice of its own.
    Off by default: the library GEMM is still 7-10 % faster on the forward / dgrad shapes."""
    _default.native_gemm = bool(on)


def use_native_wgrad(on=True):
    """Weight + bias gradient of the Linear wrappers through meant_gemm_bf16_wgrad whenever the bias gradient would
    otherwise need its own pass over dY (on by default: 112 vs 103 + 41 us at 98304 x 768 x 768, 328 vs 290 + 97 us at
    N = 2304).  With use_native_gemm(True) every weight gradient takes it."""
    _default.native_wgrad = bool(on)


def _gemm_ok(t2):
    return t2.dtype == torch.bfloat16 and t2.dim() == 2 and t2.stride(1) == 1 and t2.stride(0) % 8 == 0 and t2.data_ptr() % 16 == 0


def _wgrad_native(dy2, x2, want_db):
    """(dW fp32 [n_out, n_in], db fp32 [n_out] | None) from ONE kernel: dW = dy2^T x2, db = column sums of dy2."""
    n_out, n_in = dy2.shape[1], x2.shape[1]
    slab = torch.zeros(n_out * n_in + (n_out if want_db else 0), dtype=torch.float32, device=dy2.device)
    dw = slab[:n_out * n_in].view(n_out, n_in)
    db = slab[n_out * n_in:] if want_db else None
    nv.check(nv.timed(f"gemm_wgrad/t{dy2.shape[0]}/n{n_out}/k{n_in}/db{int(want_db)}", nv.lib().meant_gemm_bf16_wgrad, nv.ptr(dy2),
                      nv.ptr(x2), nv.ptr(dw), nv.ptr(db), dy2.shape[0], n_out, n_in, dy2.stride(0), x2.stride(0), n_in,
                      nv.stream()), "meant_gemm_bf16_wgrad")
    return dw, db


def _dgrad_native(dy2, wc):
    """dX = dy2 @ wc (wc is the [n_out, n_in] weight, read in place as an MN-major operand)."""
    n_out, n_in = wc.shape
    dx = torch.empty(dy2.shape[0], n_in, dtype=dy2.dtype, device=dy2.device)
    nv.check(nv.timed(f"gemm_dgrad/m{dy2.shape[0]}/n{n_in}/k{n_out}", nv.lib().meant_gemm_bf16, nv.ptr(dy2), nv.ptr(wc), nv.ptr(dx),
                      dy2.shape[0], n_in, n_out, dy2.stride(0), wc.stride(0), n_in, 0, 1, None, None, 0, nv.stream()),
             "meant_gemm_bf16")
    return dx


def _linear_bwd(octx, ctx_needs, dy2, x2, wc, colsum, want_db):
    """The three gradients of a Linear from 2-D bf16/fp32 operands: (dx | None, dw fp32 | None, db fp32 | None)."""
    n_out, n_in = wc.shape
    native_ok = _gemm_ok(dy2) and _gemm_ok(x2) and _gemm_ok(wc) and n_out % 8 == 0 and n_in % 8 == 0 and dy2.shape[0] >= 1024
    dx = dw = db = None
    if ctx_needs[0]:
        dx = _dgrad_native(dy2, wc) if (octx.native_gemm and native_ok) else torch.mm(dy2, wc)
    need_db = want_db and colsum is None
    if ctx_needs[1] and native_ok and (octx.native_gemm or (octx.native_wgrad and need_db)):
        dw, db = _wgrad_native(dy2, x2, need_db)
        need_db = False
    elif ctx_needs[1]:
        dw = _mm_f32(dy2.t(), x2)
    if want_db and colsum is not None:
        db = colsum
    elif need_db:
        db = torch.empty(n_out, dtype=torch.float32, device=dy2.device)
        nv.check(nv.timed(f"bias_grad/r{dy2.shape[0]}/c{n_out}", nv.lib().meant_bias_grad, nv.ptr(dy2), nv.ptr(db),
                          dy2.shape[0], n_out, dy2.stride(0), nv.dt(dy2), nv.stream()), "meant_bias_grad")
    return dx, dw, db


def _gemm_fwd(octx, xc, wc, bc, bias_f32):
    """y = xc @ wc^T + bias for bf16 operands: the tcgen05 kernel when selected and applicable, else F.linear."""
    n_out, n_in = wc.shape
    if octx.native_gemm and xc.dtype == torch.bfloat16 and n_out % 8 == 0 and n_in % 8 == 0 and xc.is_contiguous() \
            and wc.stride(1) == 1 and wc.stride(0) % 8 == 0 and (bias_f32 is None or bias_f32.dtype == torch.float32):
        x2 = xc.reshape(-1, n_in)
        y = torch.empty(x2.shape[0], n_out, dtype=xc.dtype, device=xc.device)
        nv.check(nv.timed(f"gemm_tn/m{x2.shape[0]}/n{n_out}/k{n_in}", nv.lib().meant_gemm_bf16_tn, nv.ptr(x2), nv.ptr(wc),
                          nv.ptr(bias_f32), nv.ptr(y), x2.shape[0], n_out, n_in, n_in, wc.stride(0), n_out, nv.stream()),
                 "meant_gemm_bf16_tn")
        return y.view(*xc.shape[:-1], n_out)
    return torch.nn.functional.linear(xc, wc, bc)


class _LinearFn(torch.autograd.Function):
    """nn.Linear (meant/xPosAttention.py:24-28, meant/meant.py:61-66).  The GEMMs are plain library GEMMs
    (cuBLASLt through torch: SURVEY 8f); what is ours is the backward's bias gradient -- one streaming
    column-sum kernel instead of autograd's generic reduction -- and the fp32 weight gradient written by the
    GEMM itself.  Worth it only when launches are not host-bound (TrainStep(graph=True))."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        dtype = compute_dtype(x)
        nv.require_cuda(x, weight, bias)
        with torch.autocast("cuda", enabled=False):
            xc = x.to(dtype)
            octx = ctx.octx = current()
            wc = octx.shadow_of(weight, dtype)
            wc = weight.to(dtype) if wc is None else wc
            bc = octx.shadow_of(bias, dtype)
            bc = (bias.to(dtype) if bias is not None else None) if bc is None else bc
            y = _gemm_fwd(octx, xc, wc, bc, bias.detach() if bias is not None else None)
        ctx.save_for_backward(xc, wc)
        ctx.cfg = (x.dtype, weight.dtype, bias.dtype if bias is not None else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, wc = ctx.saved_tensors
        xdt, wdt, bdt = ctx.cfg
        n_out, n_in = wc.shape
        octx = ctx.octx
        colsum = octx.take_colsum(dy, n_out, wc.dtype)    # left by _RMSNormFn.backward: sum over rows of this very tensor
        with torch.autocast("cuda", enabled=False):
            dy2 = dy.to(wc.dtype).reshape(-1, n_out)
            if dy2.stride(-1) != 1 or dy2.stride(0) % 8 or dy2.data_ptr() % 16:
                dy2 = dy2.contiguous()
            x2 = xc.reshape(-1, n_in)
            want_db = bdt is not None and ctx.needs_input_grad[2]
            dx, dw, db = _linear_bwd(octx, ctx.needs_input_grad, dy2, x2, wc, colsum, want_db)
            dx = dx.view(xc.shape).to(xdt) if dx is not None else None
            dw = dw.to(wdt) if dw is not None else None
            db = db.to(bdt) if db is not None else None
        return dx, dw, db


class _Linear3Fn(torch.autograd.Function):
    """Three nn.Linear layers over the same input as ONE GEMM (the q | v | k projection: the parameters stay three
    separate modules, meant/xPosAttention.py:24-27).  With TrainStep's weight shadows the packed bf16 weight already
    exists (the three shadows are row blocks of one buffer), so no concatenation runs at all."""

    @staticmethod
    def forward(ctx, x, w0, w1, w2, b0, b1, b2):
        dtype = compute_dtype(x)
        nv.require_cuda(x, w0)
        octx = ctx.octx = current()
        packed = octx.shadow.get(("packed", id(w0), id(w1), id(w2))) if octx.shadow is not None else None
        with torch.autocast("cuda", enabled=False):
            if packed is not None and packed[0].dtype == dtype and packed[2][0] is w0:
                wc, bc = packed[0], packed[1]
            else:
                wc, bc = torch.cat((w0, w1, w2)).to(dtype), torch.cat((b0, b1, b2)).to(dtype)
            xc = x.to(dtype)
            y = torch.nn.functional.linear(xc, wc, bc)
        ctx.save_for_backward(xc, wc)
        ctx.cfg = (x.dtype, w0.dtype, b0.dtype, (w0.shape[0], w1.shape[0], w2.shape[0]))
        return y

    @staticmethod
    def backward(ctx, dy):
        xc, wc = ctx.saved_tensors
        xdt, wdt, bdt, sizes = ctx.cfg
        n_out, n_in = wc.shape
        with torch.autocast("cuda", enabled=False):
            dy2 = dy.to(wc.dtype).reshape(-1, n_out)
            if dy2.stride(-1) != 1 or dy2.stride(0) % 8 or dy2.data_ptr() % 16:
                dy2 = dy2.contiguous()
            dx, dw, db = _linear_bwd(ctx.octx, (ctx.needs_input_grad[0], True), dy2, xc.reshape(-1, n_in), wc, None, True)
            dx = dx.view(xc.shape).to(xdt) if dx is not None else None
            dw = dw.to(wdt).split(sizes, 0)
            db = db.to(bdt).split(sizes)
        return (dx,) + tuple(dw) + tuple(db)


def linear3(x, m0, m1, m2):
    """cat(m0(x), m1(x), m2(x)) along the feature axis for three nn.Linear modules, as one GEMM."""
    if current().fused_linear and x.is_cuda and torch.is_grad_enabled() and m0.bias is not None and m1.bias is not None \
            and m2.bias is not None and x.numel() // x.shape[-1] >= 1024 and m0.weight.shape[0] % 8 == 0 \
            and m1.weight.shape[0] % 8 == 0:
        return _Linear3Fn.apply(x, m0.weight, m1.weight, m2.weight, m0.bias, m1.bias, m2.bias)
    w = torch.cat((m0.weight, m1.weight, m2.weight), dim=0)
    b = torch.cat((m0.bias, m1.bias, m2.bias), dim=0)
    return linear(x, w, b)


def use_fused_linear(on=True):
    """Route ops.linear through _LinearFn in the default context (TrainStep(graph=True) sets it in its own context)."""
    _default.fused_linear = bool(on)


def linear(x, weight, bias=None):
    """nn.Linear forward (meant/meant.py:102,106,107)."""
    if current().fused_linear and bias is not None and x.is_cuda and torch.is_grad_enabled() and \
            (weight.shape[0] % 8 == 0) and x.numel() // x.shape[-1] >= 1024:
        return _LinearFn.apply(x, weight, bias)
    return torch.nn.functional.linear(x, weight, bias)


# ---------------------------------------------------------------------------- Embedding
class _EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids, weight, dtype):
        ids2 = ids.reshape(-1).contiguous()
        out = torch.empty(ids2.numel(), weight.shape[1], dtype=dtype, device=weight.device)
        nv.check(nv.lib().meant_embedding_fwd(nv.ptr(ids2), nv.ptr(weight), nv.ptr(out), ids2.numel(), weight.shape[1],
                                              weight.shape[0], nv.dt(out), nv.stream()), "meant_embedding_fwd")
        ctx.save_for_backward(ids2)
        ctx.shape = weight.shape
        return out.view(*ids.shape, weight.shape[1])

    @staticmethod
    def backward(ctx, dy):
        (ids2,) = ctx.saved_tensors
        dy2 = dy.reshape(-1, ctx.shape[1]).contiguous()
        dw = torch.empty(ctx.shape, dtype=torch.float32, device=dy.device)
        nv.check(nv.timed(f"embedding_bwd/t{ids2.numel()}", nv.lib().meant_embedding_bwd, nv.ptr(ids2), nv.ptr(dy2),
                          nv.ptr(dw), ids2.numel(), ctx.shape[1], ctx.shape[0], nv.dt(dy2), nv.stream()),
                 "meant_embedding_bwd")
        return None, dw, None


class _EmbedBlockFn(torch.autograd.Function):
    """word + position + type gather, position ids from the padding mask, LayerNorm and dropout in one kernel
    (csrc/embed.cu); backward scatter-adds into the three tables' gradients."""

    @staticmethod
    def forward(ctx, ids, word, pos, typ, ln_w, ln_b, eps, pad, drop_p, seed, dtype, frozen_rows):
        ids2 = ids.reshape(-1, ids.shape[-1]).contiguous()
        n, s = ids2.shape
        d = word.shape[1]
        dev = word.device
        lw, lb = _f32(ln_w), _f32(ln_b)
        out = torch.empty(n, s, d, dtype=dtype, device=dev)
        pos_ids = torch.empty(n, s, dtype=torch.int32, device=dev)
        mean = torch.empty(n, s, dtype=torch.float32, device=dev)
        rstd = torch.empty_like(mean)
        sdev = current().graph_seed if drop_p > 0.0 else None
        nv.check(nv.timed(f"embed_ln_fwd/S{s}/d{d}", nv.lib().meant_embed_ln_fwd, nv.ptr(ids2), nv.ptr(word), nv.ptr(pos), nv.ptr(typ),
                          nv.ptr(lw), nv.ptr(lb), float(eps), nv.ptr(out), nv.ptr(pos_ids), nv.ptr(mean), nv.ptr(rstd), n, s, d,
                          word.shape[0], pos.shape[0], int(pad), float(drop_p), seed, nv.ptr(sdev), nv.dt(out), nv.stream()),
                 "meant_embed_ln_fwd")
        ctx.save_for_backward(ids2, pos_ids, mean, rstd, word, pos, typ, lw)
        ctx.sdev = sdev
        ctx.cfg = (float(drop_p), seed, ln_w.dtype, typ.shape, frozen_rows)
        return out.view(*ids.shape, d)

    @staticmethod
    def backward(ctx, dy):
        ids2, pos_ids, mean, rstd, word, pos, typ, lw = ctx.saved_tensors
        drop_p, seed, lndt, tshape, frozen_rows = ctx.cfg
        n, s = ids2.shape
        d = word.shape[1]
        dy2 = dy.reshape(n, s, d).contiguous()
        dword = torch.zeros_like(word)
        dpos = torch.zeros_like(pos)
        acc = torch.zeros(3, d, dtype=torch.float32, device=word.device)
        nv.check(nv.timed(f"embed_ln_bwd/S{s}/d{d}", nv.lib().meant_embed_ln_bwd, nv.ptr(dy2), nv.ptr(ids2), nv.ptr(pos_ids), nv.ptr(mean),
                          nv.ptr(rstd), nv.ptr(word), nv.ptr(pos), nv.ptr(typ), nv.ptr(lw), nv.ptr(dword), nv.ptr(dpos), nv.ptr(acc[0]),
                          nv.ptr(acc[1]), nv.ptr(acc[2]), n, s, d, word.shape[0], pos.shape[0], drop_p, seed, nv.ptr(ctx.sdev),
                          nv.dt(dy2), nv.stream()), "meant_embed_ln_bwd")
        dtyp = torch.zeros(tshape, dtype=torch.float32, device=word.device)
        dtyp[0].copy_(acc[0])
        # nn.Embedding(padding_idx=...) keeps that row's gradient at zero (Roberta: the pad row of word and position tables)
        if frozen_rows[0] is not None:
            dword[frozen_rows[0]].zero_()
        if frozen_rows[1] is not None:
            dpos[frozen_rows[1]].zero_()
        if frozen_rows[2] is not None:
            dtyp[frozen_rows[2]].zero_()
        return None, dword, dpos, dtyp, acc[1].to(lndt), acc[2].to(lndt), None, None, None, None, None, None


def _is_bert_block(mod):
    """A Hugging Face Bert/Roberta-style embedding block (what `bertweet.embeddings` is): duck-typed, no import."""
    names = ("word_embeddings", "position_embeddings", "token_type_embeddings", "LayerNorm", "dropout")
    if not all(hasattr(mod, k) for k in names):
        return False
    we, pe, te, ln = mod.word_embeddings, mod.position_embeddings, mod.token_type_embeddings, mod.LayerNorm
    return all(isinstance(e, torch.nn.Embedding) and e.weight.is_cuda and e.weight.dtype == torch.float32 and
               e.weight.is_contiguous() and e.max_norm is None for e in (we, pe, te)) and \
        isinstance(ln, torch.nn.LayerNorm) and ln.elementwise_affine and ln.bias is not None and \
        we.weight.shape[1] in (256, 512, 768, 1024) and getattr(mod, "position_embedding_type", "absolute") == "absolute"


def embed(mod, ids):
    """The model's `embedding` module applied to token ids (meant/meant_tweet.py:152).  A plain nn.Embedding on the
    GPU runs the gather kernel, which writes the compute dtype directly, and a scatter-add backward; any other module
    (e.g. a full BERT embedding block) is called as is."""
    w = getattr(mod, "weight", None)
    if type(mod) is torch.nn.Embedding and mod.padding_idx is None and mod.max_norm is None and not mod.sparse \
            and not mod.scale_grad_by_freq and w.is_cuda and w.dtype == torch.float32 and w.shape[1] % 8 == 0 \
            and w.is_contiguous() and ids.dtype == torch.int64:
        if ids.device != w.device:
            # nn.Embedding raises for this too; a host pointer must never reach the gather kernel
            raise RuntimeError(f"embed: token ids are on {ids.device} but the embedding table is on {w.device}")
        return _EmbeddingFn.apply(ids, w, compute_dtype(w))
    if ids.dtype == torch.int64 and ids.dim() >= 2 and _is_bert_block(mod):
        we = mod.word_embeddings
        if ids.device != we.weight.device:
            raise RuntimeError(f"embed: token ids are on {ids.device} but the embedding tables are on {we.weight.device}")
        # RoBERTa derives position ids from the padding mask (padding_idx on the block); BERT counts 0..S-1, which is
        # the same formula with a pad id no token carries and an offset of -1: not built, falls through to the module
        pad = getattr(mod, "padding_idx", None)
        if pad is not None:
            p = float(mod.dropout.p) if mod.training else 0.0
            return _EmbedBlockFn.apply(ids, we.weight, mod.position_embeddings.weight, mod.token_type_embeddings.weight,
                                       mod.LayerNorm.weight, mod.LayerNorm.bias, mod.LayerNorm.eps, int(pad), p,
                                       _seed() if p > 0.0 else 0, compute_dtype(we.weight),
                                       (we.padding_idx, mod.position_embeddings.padding_idx, mod.token_type_embeddings.padding_idx))
    return mod(ids)


# ----------------------------------------------------------------------------- patchify
class _PatchifyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, images, patch, dtype):
        nv.require_cuda(images)
        if images.dtype not in (torch.float32, torch.bfloat16):
            images = images.float()
        x = images.contiguous()
        b, c, hh, ww = x.shape
        if hh % patch or ww % patch:
            raise ValueError(f"image size {hh}x{ww} is not divisible by the patch size {patch}")
        out = torch.empty(b, (hh // patch) * (ww // patch), patch * patch * c, dtype=dtype, device=x.device)
        nv.check(nv.timed(f"patchify/c{c}/{hh}x{ww}", nv.lib().meant_patchify, nv.ptr(x), nv.ptr(out), b, c, hh, ww, patch, nv.dt(x),
                          nv.dt(out), 0, nv.stream()), "meant_patchify")
        ctx.cfg = (b, c, hh, ww, patch, images.dtype)
        return out

    @staticmethod
    def backward(ctx, dy):
        b, c, hh, ww, patch, idt = ctx.cfg
        dy = dy.contiguous()
        if dy.dtype not in (torch.float32, torch.bfloat16):
            dy = dy.float()
        dx = torch.empty(b, c, hh, ww, dtype=idt if idt in (torch.float32, torch.bfloat16) else torch.float32, device=dy.device)
        nv.check(nv.lib().meant_patchify(nv.ptr(dy), nv.ptr(dx), b, c, hh, ww, patch, nv.dt(dy), nv.dt(dx), 1, nv.stream()),
                 "meant_patchify")
        return dx, None, None


def patchify(images, patch):
    """'b c (h p1) (w p2) -> b (h w) (p1 p2 c)' (meant/meant.py:193-195) written once in the compute dtype."""
    return _PatchifyFn.apply(images, patch, compute_dtype(images))


# ---------------------------------------------------------------------------- LayerNorm
class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        dtype = compute_dtype(x)
        x2 = _prep(x, dtype)
        d = x2.shape[-1]
        rows = x2.numel() // d
        w32, b32 = _f32(weight), _f32(bias)
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x2.device)
        rstd = torch.empty_like(mean)
        nv.check(nv.lib().meant_layernorm_fwd(nv.ptr(x2), nv.ptr(w32), nv.ptr(b32), nv.ptr(y), nv.ptr(mean),
                                              nv.ptr(rstd), rows, d, eps, nv.dt(x2), nv.stream()),
                 "meant_layernorm_fwd")
        ctx.save_for_backward(x2, w32, mean, rstd)
        ctx.cfg = (weight.dtype, x.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, w32, mean, rstd = ctx.saved_tensors
        wdt, xdt = ctx.cfg
        d = x2.shape[-1]
        rows = x2.numel() // d
        dy = dy.to(x2.dtype).contiguous()
        dx = torch.empty_like(x2)
        dw = torch.zeros(d, dtype=torch.float32, device=x2.device)
        db = torch.zeros_like(dw)
        nv.check(nv.lib().meant_layernorm_bwd(nv.ptr(dy), nv.ptr(x2), nv.ptr(w32), nv.ptr(mean), nv.ptr(rstd),
                                              nv.ptr(dx), nv.ptr(dw), nv.ptr(db), rows, d, nv.dt(x2),
                                              nv.stream()), "meant_layernorm_bwd")
        return dx.to(xdt), dw.to(wdt), db.to(wdt), None


def layer_norm(x, weight, bias, eps=1e-5):
    return _LayerNormFn.apply(x, weight, bias, float(eps))


# ---------------------------------------------------------------------------- attention
def _view(t, col0, heads, dh, n2_count=1, group=None):
    """HeadView of columns [col0, col0 + heads*dh) of a [N, S, W] tensor (last stride 1)."""
    assert t.stride(-1) == 1
    es = t.element_size()
    if group is None:
        return HeadView(t.data_ptr() + col0 * es, t.stride(0), 0, t.stride(1), dh)
    sn1, sn2, ss = group
    return HeadView(t.data_ptr() + col0 * es, sn1, sn2, ss, dh)


class _AttnFn(torch.autograd.Function):
    """mode 'qkv'  : t0 = [N, S, 3*dim] packed (q | k | v), t1 = t2 = None
       mode 'q_kv' : t0 = [N, Sq, dim], t1 = [N, Sk, 2*dim] packed (k | v)
       mode 'q_k_v': three [N, S*, dim] tensors."""

    @staticmethod
    def forward(ctx, t0, t1, t2, mode, heads, scale, causal, rot, pos_is_head, key_mask, impl, consume):
        dtype = compute_dtype(t0)
        given = (t0, t1, t2)
        t0, t1, t2 = (_prep_rows(t, dtype) for t in (t0, t1, t2))
        # Q and K may be rotated in place when the caller gives the buffers up (`consume`: a projection output nobody
        # else reads) or when _prep_rows made private copies of them anyway
        own_qk = consume or (t0 is not given[0] and (mode == "qkv" or t1 is not given[1]))
        if mode == "qkv":
            dim = t0.shape[-1] // 3
            srcs = ((t0, 0), (t0, dim), (t0, 2 * dim))
        elif mode == "q_kv":
            dim = t0.shape[-1]
            srcs = ((t0, 0), (t1, 0), (t1, dim))
        else:
            dim = t0.shape[-1]
            srcs = ((t0, 0), (t1, 0), (t2, 0))
        if dim % heads:
            raise ValueError(f"dim {dim} is not divisible by heads {heads}")
        dh = dim // heads
        n, sq = t0.shape[0], t0.shape[1]
        sk = srcs[1][0].shape[1]
        out = torch.empty(n, sq, dim, dtype=dtype, device=t0.device)
        lse = torch.empty(n, heads, sq, dtype=torch.float32, device=t0.device)
        a = AttnArgs()
        a.q, a.k, a.v = (_view(t, c, heads, dh) for t, c in srcs)
        a.o = _view(out, 0, heads, dh)
        a.lse = nv.ptr(lse)
        tabs = None
        if rot is not None:
            tabs = tuple(r.contiguous() for r in rot)
            a.qcos, a.qsin, a.kcos, a.ksin = (nv.ptr(r) for r in tabs)
            a.rot_dim = tabs[0].shape[-1]
            need_q, need_k = (heads, heads) if pos_is_head else (sq, sk)
            if tabs[0].shape[0] < need_q or tabs[2].shape[0] < need_k or a.rot_dim > dh:
                raise ValueError("rotary tables do not cover the positions / head size")
        km = None
        if key_mask is not None:
            km = key_mask.to(device=t0.device, dtype=torch.float32).contiguous()
            if km.shape != (n, sk):
                raise ValueError(f"attention_mask must be [{n}, {sk}], got {tuple(km.shape)}")
            a.key_mask = nv.ptr(km)
        a.N, a.n2_count, a.H, a.Sq, a.Sk, a.Dh = n, 1, heads, sq, sk, dh
        a.pos_is_head, a.causal, a.scale = int(pos_is_head), int(causal), float(scale)
        # Multi-tile problems on the tensor-core path: rotate Q and K ONCE into scratch (every query tile needs every
        # key tile; the kernels then stage plain tiles), kept for the backward pass.
        qrot = krot = None
        inplace = False
        octx = current()
        if ((PREROTATE if octx.prerotate is None else octx.prerotate) and tabs is not None and (sq > 128 or sk > 128) and impl != nv.IMPL_EXACT
                and nv.lib().meant_attn_tcgen05_supported(a, nv.dt(t0), 0)):
            inplace = bool(own_qk)
            if inplace:                                   # no q_rot / k_rot buffers: the projection output itself is rotated
                a.q_rot, a.k_rot = a.q, a.k
                impl = nv.IMPL_TCGEN05                    # the exact kernels would rotate again: fail loudly instead
            else:
                qrot = torch.empty(n, sq, dim, dtype=dtype, device=t0.device)
                krot = torch.empty(n, sk, dim, dtype=dtype, device=t0.device)
                a.q_rot, a.k_rot = _view(qrot, 0, heads, dh), _view(krot, 0, heads, dh)
            nv.check(nv.timed(f"attn_prerotate/S{sq}x{sk}/Dh{dh}", nv.lib().meant_attn_prerotate, a, nv.dt(t0), nv.stream()),
                     "meant_attn_prerotate")
        nv.check(nv.timed(f"attn_fwd/S{sq}x{sk}/Dh{dh}/c{int(causal)}", nv.lib().meant_attn_fwd, a, nv.dt(t0), impl, nv.stream()),
                 "meant_attn_fwd")
        ctx.save_for_backward(t0, t1, t2, out, lse, km, qrot, krot, *(tabs or ()))
        ctx.cfg = (mode, heads, float(scale), int(causal), int(pos_is_head), impl, rot is not None, inplace)
        return out

    @staticmethod
    def backward(ctx, dout):
        t0, t1, t2, out, lse, km, qrot, krot, *tabs = ctx.saved_tensors
        mode, heads, scale, causal, pos_is_head, impl, has_rot, inplace = ctx.cfg
        dout = dout.to(out.dtype).contiguous()
        g0 = torch.empty_like(t0)
        g1 = torch.empty_like(t1) if t1 is not None else None
        g2 = torch.empty_like(t2) if t2 is not None else None
        if mode == "qkv":
            dim = t0.shape[-1] // 3
            srcs = ((t0, 0), (t0, dim), (t0, 2 * dim))
            dsts = ((g0, 0), (g0, dim), (g0, 2 * dim))
        elif mode == "q_kv":
            dim = t0.shape[-1]
            srcs = ((t0, 0), (t1, 0), (t1, dim))
            dsts = ((g0, 0), (g1, 0), (g1, dim))
        else:
            dim = t0.shape[-1]
            srcs = ((t0, 0), (t1, 0), (t2, 0))
            dsts = ((g0, 0), (g1, 0), (g2, 0))
        dh = dim // heads
        n, sq, sk = t0.shape[0], t0.shape[1], srcs[1][0].shape[1]
        delta = torch.empty_like(lse)
        a = AttnArgs()
        a.q, a.k, a.v = (_view(t, c, heads, dh) for t, c in srcs)
        a.dq, a.dk, a.dv = (_view(t, c, heads, dh) for t, c in dsts)
        a.o = _view(dout, 0, heads, dh)
        a.out_saved = _view(out, 0, heads, dh)
        a.lse, a.delta = nv.ptr(lse), nv.ptr(delta)
        if has_rot:
            a.qcos, a.qsin, a.kcos, a.ksin = (nv.ptr(r) for r in tabs)
            a.rot_dim = tabs[0].shape[-1]
        if km is not None:
            a.key_mask = nv.ptr(km)
        if qrot is not None:
            a.q_rot, a.k_rot = _view(qrot, 0, heads, dh), _view(krot, 0, heads, dh)
        elif inplace:                                     # the saved projections ARE the rotated Q and K
            a.q_rot, a.k_rot = a.q, a.k
        a.N, a.n2_count, a.H, a.Sq, a.Sk, a.Dh = n, 1, heads, sq, sk, dh
        a.pos_is_head, a.causal, a.scale = pos_is_head, causal, scale
        nv.check(nv.timed(f"attn_bwd/S{sq}x{sk}/Dh{dh}/c{causal}", nv.lib().meant_attn_bwd, a, nv.dt(t0), impl, nv.stream()),
                 "meant_attn_bwd")
        return g0, g1, g2, None, None, None, None, None, None, None, None, None


DEFAULT_ATTN_IMPL = nv.IMPL_AUTO   # bench/tests may force IMPL_EXACT or IMPL_TCGEN05
PREROTATE = True                   # multi-tile tensor-core attention: rotate Q, K once (meant_attn_prerotate)


def attention(qkv=None, q=None, kv=None, k=None, v=None, *, heads, scale, causal=False, rot=None,
              pos_is_head=False, key_mask=None, impl=None, consume=False):
    """softmax((Rq q)(Rk k)^T * scale + masks) v on packed `b s (h d)` projections.
    Returns [N, Sq, dim] in `b s (h d)` layout (no rearrange is materialised).
    consume=True: the caller gives the projection buffers up (nobody else reads them, and the op that produced them does
    not need its own output in backward -- a Linear): the multi-tile path then rotates Q and K in place instead of
    writing rotated copies (half the traffic of meant_attn_prerotate, two activations less per layer)."""
    if impl is None:
        impl = DEFAULT_ATTN_IMPL if current().attn_impl is None else current().attn_impl
    if qkv is not None:
        return _AttnFn.apply(qkv, None, None, "qkv", heads, scale, causal, rot, pos_is_head, key_mask, impl, bool(consume))
    if kv is not None:
        return _AttnFn.apply(q, kv, None, "q_kv", heads, scale, causal, rot, pos_is_head, key_mask, impl, bool(consume))
    return _AttnFn.apply(q, k, v, "q_k_v", heads, scale, causal, rot, pos_is_head, key_mask, impl, bool(consume))


# ------------------------------------------------- TimeSformer divided space-time attention
class _DividedAttnFn(torch.autograd.Function):
    """The attention core of src/meant/timesformer_pytorch.py:108-145 on the packed [b, 1 + f*n, 3*inner] projection
    (q | k | v), IN PLACE: the '(b n) f d' / '(b f) n d' groups of :122 are head views over that buffer
    (n1 = b, n2 = patch or frame), the cls key / value every group attends to (:125-133, `repeat` + `cat` in the
    reference) is the kernels' shared extra key read from the un-replicated row 0, outputs and gradients are written
    straight into `b (f n) (h d)` buffers, and the cls query (:116-119) runs on the single-query kernels.  No
    rearranged, repeated or concatenated copy of q, k, v or of their gradients exists."""

    @staticmethod
    def forward(ctx, qkv, heads, scale, f, n, time_mode, rot, impl):
        ctx.octx = current()
        dtype = compute_dtype(qkv)
        nv.require_cuda(qkv)
        qkv = qkv.to(dtype)
        if qkv.dim() != 3 or qkv.stride(-1) != 1 or qkv.stride(1) != qkv.shape[-1] or qkv.data_ptr() % 16:
            qkv = qkv.contiguous()
        b, tokens, w = qkv.shape
        inner = w // 3
        dh = inner // heads
        if tokens != 1 + f * n or inner * 3 != w or dh * heads != inner:
            raise ValueError(f"divided attention: {tuple(qkv.shape)} is not [b, 1 + {f}*{n}, 3 * heads * dim_head]")
        out = torch.empty(b, tokens, inner, dtype=dtype, device=qkv.device)
        tabs = None
        if rot is not None:
            tabs = tuple(r.contiguous() for r in rot)
        ctx.cfg = (heads, float(scale), f, n, bool(time_mode), impl)
        lse_c = torch.empty(b, heads, dtype=torch.float32, device=qkv.device)
        nchunk = (tokens + 63) // 64
        partial = torch.empty(b, heads, nchunk, dh + 2, dtype=torch.float32, device=qkv.device)
        a = _DividedAttnFn._cls_args(qkv, out, lse_c, heads, scale)
        nv.check(nv.timed(f"attn_q1_fwd/Sk{tokens}/Dh{dh}", nv.lib().meant_attn_q1_fwd, a, nv.ptr(partial), nv.dt(qkv), nv.stream()),
                 "meant_attn_q1_fwd")
        g, lse_g = _DividedAttnFn._group_args(qkv, out, None, heads, scale, f, n, time_mode, tabs)
        nv.check(nv.timed(f"attn_grp_fwd/{'time' if time_mode else 'space'}/S{g.Sq}/Dh{dh}", nv.lib().meant_attn_fwd, g,
                          nv.dt(qkv), impl, nv.stream()), "meant_attn_fwd")
        ctx.save_for_backward(qkv, out, lse_c, lse_g, *(tabs or ()))
        return out

    @staticmethod
    def _cls_args(qkv, out, lse_c, heads, scale):
        """Single-query problem: row 0 of q against every row of k, v."""
        b, tokens, w = qkv.shape
        inner = w // 3
        dh = inner // heads
        a = AttnArgs()
        a.q, a.k, a.v = (_view(qkv, c, heads, dh) for c in (0, inner, 2 * inner))
        a.o = _view(out, 0, heads, dh)
        a.lse = nv.ptr(lse_c)
        a.N, a.n2_count, a.H, a.Sq, a.Sk, a.Dh = b, 1, heads, 1, tokens, dh
        a.scale = float(scale)
        return a

    @staticmethod
    def _group_args(qkv, out, lse_g, heads, scale, f, n, time_mode, tabs):
        """Grouped problem over the body tokens 1..: time groups (b, patch) of f rows, or space groups (b, frame) of
        n rows; the cls row is the shared extra key (Sk = rows + 1)."""
        b, tokens, w = qkv.shape
        inner = w // 3
        dh = inner // heads
        es = qkv.element_size()
        rows, n2c = (f, n) if time_mode else (n, f)

        def body(t, col):        # head view of the body tokens of t[b, tokens, width] at column `col`, grouped
            width = t.stride(1)
            sn2, ss = (width, n * width) if time_mode else (n * width, width)
            return HeadView(t.data_ptr() + (width + col) * es, t.stride(0), sn2, ss, dh)

        def xrow(t, col):        # row 0 of every batch element
            return HeadView(t.data_ptr() + col * es, t.stride(0), 0, 0, dh)

        g = AttnArgs()
        g.q, g.k, g.v = (body(qkv, c) for c in (0, inner, 2 * inner))
        g.kx, g.vx = xrow(qkv, inner), xrow(qkv, 2 * inner)
        g.o = body(out, 0)
        if lse_g is None:
            lse_g = torch.empty(b * n2c, heads, rows, dtype=torch.float32, device=qkv.device)
        g.lse = nv.ptr(lse_g)
        if tabs is not None:
            g.qcos, g.qsin, g.kcos, g.ksin = (nv.ptr(r) for r in tabs)
            g.rot_dim = tabs[0].shape[-1]
            if tabs[0].shape[0] < rows or tabs[2].shape[0] < rows + 1 or g.rot_dim > dh:
                raise ValueError("rotary tables do not cover the group length / head size")
        g.N, g.n2_count, g.H, g.Sq, g.Sk, g.Dh = b * n2c, n2c, heads, rows, rows + 1, dh
        g.scale = float(scale)
        return g, lse_g

    @staticmethod
    def backward(ctx, dout):
        qkv, out, lse_c, lse_g, *tabs = ctx.saved_tensors
        heads, scale, f, n, time_mode, impl = ctx.cfg
        tabs = tuple(tabs) or None
        b, tokens, w = qkv.shape
        inner = w // 3
        dh = inner // heads
        dout = dout.to(out.dtype).contiguous()
        dqkv = torch.empty_like(qkv)
        acc = ctx.octx.zeros(3 * b * inner, qkv.device).view(3, b, inner)          # dkx | dvx | dq of the cls row, fp32
        # groups: dq, dk, dv of the body rows (written), gradient of the shared cls key / value (accumulated)
        g, _ = _DividedAttnFn._group_args(qkv, out, lse_g, heads, scale, f, n, time_mode, tabs)
        es = qkv.element_size()
        rows, n2c = (f, n) if time_mode else (n, f)

        def body(t, col):
            width = t.stride(1)
            sn2, ss = (width, n * width) if time_mode else (n * width, width)
            return HeadView(t.data_ptr() + (width + col) * es, t.stride(0), sn2, ss, dh)

        g.dq, g.dk, g.dv = (body(dqkv, c) for c in (0, inner, 2 * inner))
        g.o = body(dout, 0)
        g.out_saved = body(out, 0)
        delta = torch.empty_like(lse_g)
        g.delta = nv.ptr(delta)
        g.dkx, g.dvx = nv.ptr(acc[0]), nv.ptr(acc[1])
        nv.check(nv.timed(f"attn_grp_bwd/{'time' if time_mode else 'space'}/S{rows}/Dh{dh}", nv.lib().meant_attn_bwd, g,
                          nv.dt(qkv), impl, nv.stream()), "meant_attn_bwd")
        dqkv[:, 0, inner:2 * inner].copy_(acc[0])
        dqkv[:, 0, 2 * inner:].copy_(acc[1])
        # cls query: dq of row 0; its dk, dv are ADDED to what the groups left in every row
        a = _DividedAttnFn._cls_args(qkv, dout, lse_c, heads, scale)      # a.o = dO
        a.out_saved = _view(out, 0, heads, dh)
        a.dk, a.dv = _view(dqkv, inner, heads, dh), _view(dqkv, 2 * inner, heads, dh)
        nv.check(nv.timed(f"attn_q1_bwd/Sk{tokens}/Dh{dh}", nv.lib().meant_attn_q1_bwd, a, nv.ptr(acc[2]), 1, nv.dt(qkv),
                          nv.stream()), "meant_attn_q1_bwd")
        dqkv[:, 0, :inner].copy_(acc[2])
        return dqkv, None, None, None, None, None, None, None


def divided_attention(qkv, *, heads, scale, frames, patches, time_mode, rot=None, impl=None):
    """TimeSformer time ('(b n) f d' groups) or space ('(b f) n d' groups) attention plus the cls query, on the packed
    q | k | v projection [b, 1 + frames*patches, 3*heads*dim_head]; returns [b, 1 + frames*patches, heads*dim_head].
    rot = (qcos, qsin, kcos, ksin): query tables with one row per group position, key tables with one more row (the
    cls key, not rotated)."""
    if impl is None:
        impl = DEFAULT_ATTN_IMPL if current().attn_impl is None else current().attn_impl
    return _DividedAttnFn.apply(qkv, heads, scale, frames, patches, time_mode, rot, impl)


# ------------------------------------------------------------------- pooling, GEGLU, shift
class _MeanPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        dtype = compute_dtype(x)
        x2 = _prep(x, dtype)
        n, s, d = x2.shape
        y = torch.empty(n, d, dtype=dtype, device=x2.device)
        nv.check(nv.lib().meant_meanpool_fwd(nv.ptr(x2), nv.ptr(y), n, s, d, nv.dt(x2), nv.stream()),
                 "meant_meanpool_fwd")
        ctx.cfg = (n, s, d, x.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        n, s, d, xdt = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty(n, s, d, dtype=dy.dtype, device=dy.device)
        nv.check(nv.lib().meant_meanpool_bwd(nv.ptr(dy), nv.ptr(dx), n, s, d, nv.dt(dy), nv.stream()),
                 "meant_meanpool_bwd")
        return dx.to(xdt)


def mean_pool(x):
    """mean over axis 1 of [N, S, D] (meant/meant.py:231)."""
    return _MeanPoolFn.apply(x)


class _GegluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        dtype = compute_dtype(x)
        x2 = _prep(x, dtype)
        f = x2.shape[-1] // 2
        rows = x2.numel() // (2 * f)
        y = torch.empty(*x2.shape[:-1], f, dtype=dtype, device=x2.device)
        nv.check(nv.lib().meant_geglu_fwd(nv.ptr(x2), nv.ptr(y), rows, f, nv.dt(x2), nv.stream()),
                 "meant_geglu_fwd")
        ctx.save_for_backward(x2)
        ctx.xdt = x.dtype
        return y

    @staticmethod
    def backward(ctx, dy):
        (x2,) = ctx.saved_tensors
        f = x2.shape[-1] // 2
        rows = x2.numel() // (2 * f)
        dy = dy.to(x2.dtype).contiguous()
        dx = torch.empty_like(x2)
        nv.check(nv.lib().meant_geglu_bwd(nv.ptr(dy), nv.ptr(x2), nv.ptr(dx), rows, f, nv.dt(x2), nv.stream()),
                 "meant_geglu_bwd")
        return dx.to(ctx.xdt)


def geglu(x):
    return _GegluFn.apply(x)


class _TokenShiftFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, frames):
        dtype = compute_dtype(x)
        x2 = _prep(x, dtype)
        b, ntok, d = x2.shape
        n = (ntok - 1) // frames
        y = torch.empty_like(x2)
        nv.check(nv.lib().meant_token_shift(nv.ptr(x2), nv.ptr(y), b, frames, n, d, 0, nv.dt(x2), nv.stream()),
                 "meant_token_shift")
        ctx.cfg = (b, frames, n, d, x.dtype)
        return y

    @staticmethod
    def backward(ctx, dy):
        b, frames, n, d, xdt = ctx.cfg
        dy = dy.contiguous()
        dx = torch.empty_like(dy)
        nv.check(nv.lib().meant_token_shift(nv.ptr(dy), nv.ptr(dx), b, frames, n, d, 1, nv.dt(dy), nv.stream()),
                 "meant_token_shift")
        return dx.to(xdt), None


def token_shift(x, frames):
    return _TokenShiftFn.apply(x, frames)
