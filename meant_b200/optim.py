"""The step after backward on the sm_100a multi-tensor kernels (csrc/optim.cu): gradient norm, clip + AdamW, and the
cross entropy on the model's sigmoid outputs.  Reference: src/in_loop_train.py:237-244 (CrossEntropyLoss(out, target);
clip_grad_norm_(model.parameters(), 1.0); optimizer.step()) and :568 (torch.optim.AdamW)."""
import ctypes as C

import torch

from . import _native as nv


def _tables(entries, g_bf16=False):
    """entries: [(p, g, m, v, shadow or None)] -> list of TensorTable, each <= 48 tensors and <= 320 chunks.  A tensor
    larger than one launch's chunks (the 64001 x 768 embedding: 751 chunks) is cut across launches by offsetting its
    pointers (chunk boundaries are multiples of 65536 elements, so the 16-byte alignment of the vector path survives)."""
    tabs, cur = [], None

    def fresh():
        t = nv.TensorTable()
        t.n_tensors = t.n_blocks = 0
        t.g_bf16 = 1 if g_bf16 else 0
        return t

    for p, g, m, v, sh in entries:
        numel, done = p.numel(), 0
        while done < numel:
            if cur is None or cur.n_tensors == nv.OPT_MAX_TENSORS or cur.n_blocks == nv.OPT_MAX_BLOCKS:
                if cur is not None:
                    tabs.append(cur)
                cur = fresh()
            room = nv.OPT_MAX_BLOCKS - cur.n_blocks
            part = min(numel - done, room * nv.OPT_CHUNK)
            i = cur.n_tensors
            cur.p[i] = p.data_ptr() + 4 * done
            cur.g[i] = g.data_ptr() + (2 if g_bf16 else 4) * done
            cur.m[i] = m.data_ptr() + 4 * done
            cur.v[i] = v.data_ptr() + 4 * done
            cur.shadow[i] = (sh.data_ptr() + 2 * done) if sh is not None else None
            cur.numel[i] = part
            for c in range((part + nv.OPT_CHUNK - 1) // nv.OPT_CHUNK):
                cur.block_tensor[cur.n_blocks] = i
                cur.block_chunk[cur.n_blocks] = c
                cur.n_blocks += 1
            cur.n_tensors += 1
            done += part
    if cur is not None and cur.n_tensors:
        tabs.append(cur)
    return tabs


class FusedAdamW:
    """torch.optim.AdamW semantics (decoupled weight decay, bias correction, eps outside the sqrt) with gradient-norm
    clipping folded in: step() = clip_grad_norm_(params, max_norm) + AdamW.step() in two multi-tensor passes
    (sum of squares, update).  State (exp_avg, exp_avg_sq) lives in two flat fp32 buffers; the step counter and the
    squared norm stay on the device, so a captured CUDA graph replays it.  `shadows` maps id(param) -> a bf16 tensor that
    receives a copy of the updated parameter (the GEMM operands of the next forward: no separate cast pass).
    Parameters must be fp32 CUDA tensors; parameters without a gradient are skipped, like torch does."""

    def __init__(self, params, lr=5e-5, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, max_norm=1.0, shadows=None):
        self.params = [p for p in params if p.requires_grad]
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32 or not p.is_contiguous():
                raise ValueError("FusedAdamW takes contiguous fp32 CUDA parameters")
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), (float(betas[0]), float(betas[1])), float(eps), float(weight_decay)
        self.max_norm = float(max_norm or 0.0)
        dev = self.params[0].device
        total = sum((p.numel() + 3) // 4 * 4 for p in self.params)
        self._m = torch.zeros(total, dtype=torch.float32, device=dev)
        self._v = torch.zeros(total, dtype=torch.float32, device=dev)
        self.state, off = {}, 0
        for p in self.params:
            n = p.numel()
            self.state[p] = (self._m[off:off + n].view_as(p), self._v[off:off + n].view_as(p))
            off += (n + 3) // 4 * 4
        self._step = torch.zeros(1, dtype=torch.float32, device=dev)
        self._sumsq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.shadows = shadows or {}
        self._cache_key, self._cache = None, None
        self.launches = 0

    def _build(self, grads=None):
        """grads: optional {param: gradient tensor} (the data-parallel reducer's reduced bucket slices, fp32 or bf16);
        default: every parameter's .grad."""
        if grads is None:
            live = [(p, p.grad) for p in self.params if p.grad is not None]
        else:
            live = [(p, grads[p]) for p in self.params if grads.get(p) is not None]
        key = tuple((p.data_ptr(), g.data_ptr(), g.dtype) for p, g in live)
        if key != self._cache_key:
            by_dtype = {torch.float32: [], torch.bfloat16: []}
            for p, g in live:
                if g.dtype not in by_dtype or not g.is_contiguous() or g.device != p.device or g.numel() != p.numel():
                    raise ValueError("FusedAdamW takes contiguous fp32 or bf16 gradients on the parameter's device")
                m, v = self.state[p]
                sh = self.shadows.get(id(p))
                by_dtype[g.dtype].append((p.data, g, m, v, sh[1] if sh is not None else None))
            self._cache_key = key
            self._cache = _tables(by_dtype[torch.float32]) + _tables(by_dtype[torch.bfloat16], g_bf16=True)
        return self._cache

    @torch.no_grad()
    def step(self, grads=None):
        tabs = self._build(grads)
        st = nv.stream()
        self._step.add_(1.0)
        sumsq = None
        if self.max_norm > 0.0:
            self._sumsq.zero_()
            for t in tabs:
                nv.check(nv.timed("multi_sqnorm", nv.lib().meant_multi_sqnorm, C.byref(t), nv.ptr(self._sumsq), st), "meant_multi_sqnorm")
            sumsq = self._sumsq
        for t in tabs:
            nv.check(nv.timed("adamw_step", nv.lib().meant_adamw_step, C.byref(t), nv.ptr(self._step), nv.ptr(sumsq), self.lr,
                              self.betas[0], self.betas[1], self.eps, self.weight_decay, self.max_norm, st), "meant_adamw_step")
        self.launches = len(tabs) * (2 if sumsq is not None else 1)

    def grad_norm(self):
        """Total gradient norm of the last step() (before clipping), as a device tensor."""
        return self._sumsq.sqrt()

    def zero_grad(self, set_to_none=True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()


class _ProbCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, probs, target):
        nv.require_cuda(probs, target)
        p32 = probs.detach().float().contiguous()
        if p32.dim() != 2 or target.shape != (p32.shape[0],) or target.dtype != torch.int64:
            raise ValueError("prob_cross_entropy takes probs [B, C] and int64 targets [B]")
        loss = torch.empty((), dtype=torch.float32, device=p32.device)
        dp = torch.empty_like(p32)
        nv.check(nv.lib().meant_prob_ce(nv.ptr(p32), nv.ptr(target.contiguous()), nv.ptr(loss), nv.ptr(dp), p32.shape[0],
                                        p32.shape[1], 1.0, nv.stream()), "meant_prob_ce")
        ctx.save_for_backward(dp)
        ctx.dtype = probs.dtype
        return loss

    @staticmethod
    def backward(ctx, dloss):
        (dp,) = ctx.saved_tensors
        return (dp * dloss).to(ctx.dtype), None


def prob_cross_entropy(probs, target):
    """nn.CrossEntropyLoss()(probs, target) with mean reduction, forward value and gradient from one kernel launch."""
    return _ProbCE.apply(probs, target)
