"""Heads around the encoder stacks: the `src` tree's softmax-weighted pooling (lang_prep / image_prep), the MLM head's
token cross entropy and the MIM pre-trainer shell.  References: src/meant/meant_tweet.py:173,259-261,
src/meant/meant_timesformer.py:274-279,335-345, pretrain_mlm.py:74-88,154,182, pretrain_mim.py:77-100."""
import math

import torch
from torch import nn

from . import _native as nv
from . import ops


def make_prep(dim):
    """The reference's pooling scorer, module for module (state_dict keys 0.*, 1.*, 3.*):
    nn.Sequential(Linear(d, d), LayerNorm(d), GELU(), Linear(d, 1), Softmax(dim=2))  -- src/meant/meant_tweet.py:173."""
    return nn.Sequential(nn.Linear(dim, dim), nn.LayerNorm(dim), nn.GELU(), nn.Linear(dim, 1), nn.Softmax(dim=2))


class _SoftmaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, h, ln_w, ln_b, eps, w2, b2, degenerate):
        dtype = ops.compute_dtype(x)
        nv.require_cuda(x, h, ln_w, w2)
        x2 = x.to(dtype).contiguous()
        n, s, d = x2.shape
        h2 = h.to(dtype).contiguous() if not degenerate else None
        lw, lb, w = (t.detach().float().contiguous() for t in (ln_w, ln_b, w2.reshape(-1)))
        out = torch.empty(n, d, dtype=dtype, device=x2.device)
        prob = torch.empty(n, s, dtype=torch.float32, device=x2.device) if not degenerate else None
        nv.check(nv.timed(f"softmax_pool_fwd/S{s}/d{d}", nv.lib().meant_softmax_pool_fwd, nv.ptr(x2), nv.ptr(h2), nv.ptr(lw),
                          nv.ptr(lb), float(eps), nv.ptr(w), float(b2), nv.ptr(out), nv.ptr(prob), n, s, d, int(degenerate),
                          nv.dt(x2), nv.stream()), "meant_softmax_pool_fwd")
        ctx.save_for_backward(x2, h2, lw, lb, w, prob)
        ctx.cfg = (float(eps), bool(degenerate), x.dtype, h.dtype, ln_w.dtype, w2.dtype, tuple(w2.shape))
        return out

    @staticmethod
    def backward(ctx, dout):
        x2, h2, lw, lb, w, prob = ctx.saved_tensors
        eps, degenerate, xdt, hdt, lndt, wdt, wshape = ctx.cfg
        n, s, d = x2.shape
        dout = dout.to(x2.dtype).contiguous()
        dx = torch.empty_like(x2)
        dh = torch.empty_like(x2)
        acc = torch.zeros(3, d, dtype=torch.float32, device=x2.device)
        nv.check(nv.timed(f"softmax_pool_bwd/S{s}/d{d}", nv.lib().meant_softmax_pool_bwd, nv.ptr(dout), nv.ptr(x2), nv.ptr(h2),
                          nv.ptr(lw), nv.ptr(lb), eps, nv.ptr(w), nv.ptr(prob), nv.ptr(dx), nv.ptr(dh), nv.ptr(acc[0]),
                          nv.ptr(acc[1]), nv.ptr(acc[2]), n, s, d, int(degenerate), nv.dt(x2), nv.stream()),
                 "meant_softmax_pool_bwd")
        # the Linear(d, 1) bias shifts every score of a sequence equally: its gradient is exactly zero
        return dx.to(xdt), dh.to(hdt), acc[0].to(lndt), acc[1].to(lndt), None, acc[2].to(wdt).view(wshape), None, None


def softmax_pool(x, prep):
    """torch.matmul(x.transpose(-2, -1), prep(x)).squeeze(-1) for the reference's `prep` Sequential (make_prep):
    x [B, lag, S, d] -> [B, lag, d] with the softmax over the S tokens (src/meant/meant_tweet.py:259-261), or
    x [B, T, d] -> [B, d], where Softmax(dim=2) runs over the scorer's length-1 output axis, every weight is 1 and the
    result is the plain token sum (src/meant/meant_timesformer.py:342-344: reproduced as is)."""
    lin1, ln, _, lin2, sm = prep
    if not isinstance(sm, nn.Softmax) or sm.dim != 2 or lin2.out_features != 1:
        raise ValueError("softmax_pool expects Sequential(Linear(d,d), LayerNorm(d), GELU(), Linear(d,1), Softmax(dim=2))")
    lead = x.shape[:-2]
    s, d = x.shape[-2:]
    degenerate = x.dim() == 3
    x3 = x.reshape(-1, s, d)
    h = ops.linear(x3, lin1.weight, lin1.bias) if not degenerate else x3
    out = _SoftmaxPoolFn.apply(x3, h, ln.weight, ln.bias, ln.eps, lin2.weight, float(lin2.bias.detach()) if lin2.bias is not None
                               and not lin2.bias.requires_grad else 0.0, degenerate)
    # (b2 only shifts the scores; it is passed as 0 when it is a trainable parameter so that no host sync is needed)
    return out.reshape(*lead, d)


class _TokenCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target, ignore_index):
        nv.require_cuda(logits, target)
        if logits.dtype not in (torch.float32, torch.bfloat16):
            logits = logits.float()
        lg = logits.reshape(-1, logits.shape[-1]).contiguous()       # dense rows: the gradient shares the pitch
        rows, v = lg.shape
        tgt = target.reshape(-1).contiguous()
        if tgt.numel() != rows or tgt.dtype != torch.int64:
            raise ValueError("token_cross_entropy takes logits [rows, V] and int64 targets [rows]")
        lse = torch.empty(rows, dtype=torch.float32, device=lg.device)
        acc = torch.zeros(2, dtype=torch.float32, device=lg.device)
        nv.check(nv.timed(f"token_ce_fwd/V{v}", nv.lib().meant_token_ce_fwd, nv.ptr(lg), nv.ptr(tgt), nv.ptr(lse), nv.ptr(acc[0:1]),
                          nv.ptr(acc[1:2]), rows, v, lg.stride(0), int(ignore_index), nv.dt(lg), nv.stream()), "meant_token_ce_fwd")
        ctx.save_for_backward(lg, tgt, lse, acc)
        ctx.cfg = (int(ignore_index), logits.shape)
        return acc[0] / acc[1]

    @staticmethod
    def backward(ctx, dloss):
        lg, tgt, lse, acc = ctx.saved_tensors
        ignore_index, shape = ctx.cfg
        rows, v = lg.shape
        dl = torch.empty(rows, v, dtype=lg.dtype, device=lg.device)
        d32 = dloss.detach().float().reshape(1).contiguous()
        nv.check(nv.timed(f"token_ce_bwd/V{v}", nv.lib().meant_token_ce_bwd, nv.ptr(lg), nv.ptr(tgt), nv.ptr(lse), nv.ptr(acc[1:2]),
                          nv.ptr(d32), nv.ptr(dl), rows, v, lg.stride(0), ignore_index, nv.dt(lg), nv.stream()), "meant_token_ce_bwd")
        return dl.view(shape), None, None


def token_cross_entropy(logits, target, ignore_index=-100):
    """nn.CrossEntropyLoss(ignore_index=...)(logits.view(-1, V), target.view(-1)) (pretrain_mlm.py:154,182): one pass for
    the loss, one for the gradient, no [rows, V] softmax in between."""
    return _TokenCEFn.apply(logits, target, ignore_index)


class meant_vision_pretrainer(nn.Module):
    """pretrain_mim.py:77-100: patchEmbed -> visionEncoder(flash=True) -> `b n d -> b d h w` -> the caller's decoder
    (the masked-image-modeling head: 1x1 convolution + PixelShuffle in the reference's training script)."""

    def __init__(self, num_encoders, decoder, mlm_input_dim, patch_res=16, channels=4, height=224, width=224, image_dim=768,
                 num_heads=8):
        super().__init__()
        from .meant import visionEncoder, _PatchRearrange
        self.channels = channels
        self.patch_dim = channels * patch_res * patch_res
        self.n = int((height * width) / (patch_res ** 2))
        self.patchEmbed = nn.Sequential(_PatchRearrange(patch_res), nn.Linear(self.patch_dim, image_dim))
        self.visionEncoders = nn.ModuleList([visionEncoder(image_dim, num_heads, flash=True)])   # one block, as in :86
        self.decoder = decoder

    def forward(self, images):
        x = self.patchEmbed(images)
        for enc in self.visionEncoders:
            x = enc(x)
        b, s, c = x.shape
        hw = math.floor(s ** 0.5)
        return self.decoder(x.permute(0, 2, 1).reshape(b, c, hw, hw))
