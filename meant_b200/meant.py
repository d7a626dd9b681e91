"""MEANT encoder blocks and the full-modality model with the reference's constructors,
attribute names and state_dict layout, running the sm_100a kernels.
Reference: meant/meant.py:35-237."""
import math

import torch
from torch import nn

from . import ops
from .attention import attention, flash_attention
from .rms_norm import RMSNorm
from .rotary_embedding_torch import RotaryEmbedding
from .temporal import temporal
from .xPosAttention import xPosAttention, xPosAttention_flash

MAX_SEQ_LENGTH = 3333


def _drop_p(mod: nn.Module) -> float:
    return float(mod.p) if mod.training else 0.0


def _lin(mod, x):
    return ops.linear(x, mod.weight, mod.bias)


def _fused_block(x, res_in, norm_a, lin_a, attn, norm_b, drop_b, lin_b, norm_c, lin_c, norm_d, drop_d, lin_d):
    """One MEANT encoder block (meant/meant.py:66-75 / :109-120) as 7 GEMMs + 5 fused kernels:
         [residual add +] RMSNorm | Linear | attention (QKV GEMM, fused rotary flash kernel, out GEMM)
         RMSNorm(+dropout) | Linear | [residual add + RMSNorm] | Linear
         [GELU + RMSNorm + dropout] | Linear
    The block input is x (+ res_in: the previous block's un-added (branch, residual) pair, so that its
    final add runs inside this block's first norm kernel).  Returns (branch, residual) un-added."""
    if res_in is None:
        h, xin = ops.rms_norm(x, norm_a.scale, norm_a.eps), x
    else:
        h, xin = ops.rms_norm(x, norm_a.scale, norm_a.eps, residual=res_in)
    h = attn(_lin(lin_a, h))
    h = ops.rms_norm(h, norm_b.scale, norm_b.eps, drop_p=drop_b)
    h = _lin(lin_b, h)
    g, r = ops.rms_norm(h, norm_c.scale, norm_c.eps, residual=xin)     # r = h + block input
    g = _lin(lin_c, g)
    g = ops.rms_norm(g, norm_d.scale, norm_d.eps, pre_gelu=True, drop_p=drop_d)
    return _lin(lin_d, g), r


class visionEncoder(nn.Module):
    def __init__(self, dim, num_heads, flash=False):
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.posEmbed = RotaryEmbedding(dim=math.floor(dim / num_heads / 2), freqs_for='pixel')
        # flash=True selects the reference's flash semantics (head-index rotary, scale 1/dim);
        # there is no device gate here: both variants run the same sm_100a kernel
        atten = (flash_attention if flash else attention)(num_heads, dim, self.posEmbed)
        self.encode = nn.ModuleList([RMSNorm(dim), nn.Linear(dim, dim), atten, RMSNorm(dim), nn.Linear(dim, dim)])
        self.encode2 = nn.ModuleList([RMSNorm(dim), nn.Linear(dim, dim), nn.GELU(), RMSNorm(dim),
                                      nn.Linear(dim, dim)])

    def forward_fused(self, x, res_in=None):
        e, e2 = self.encode, self.encode2
        return _fused_block(x, res_in, e[0], e[1], e[2], e[3], 0.0, e[4], e2[0], e2[1], e2[3], 0.0, e2[4])

    def forward(self, input):
        branch, resid = self.forward_fused(input)
        return branch + resid


class languageEncoder(nn.Module):
    def __init__(self, dim, num_heads, dropout=0.0, flash=False):
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.xPos = RotaryEmbedding(dim=48, use_xpos=True)
        att = (xPosAttention_flash if flash else xPosAttention)(num_heads, dim, self.xPos)
        self.encode = nn.ModuleList([RMSNorm(dim), nn.Linear(dim, dim), att, RMSNorm(dim), nn.Dropout(dropout),
                                     nn.Linear(dim, dim)])
        # nn.Dropout() -> p = 0.5 in training mode, as in the reference (meant/meant.py:107)
        self.encode2 = nn.ModuleList([RMSNorm(dim), nn.Linear(dim, dim), nn.GELU(), RMSNorm(dim), nn.Dropout(),
                                      nn.Linear(dim, dim)])

    def forward_fused(self, x, res_in=None, attention_mask=None):
        e, e2 = self.encode, self.encode2
        att = e[2]
        # the reference forwards the padding mask only to the eager class (meant/meant.py:112-115)
        call = (lambda t: att(t, attention_mask)) if type(att).__name__ == 'xPosAttention' else att
        return _fused_block(x, res_in, e[0], e[1], call, e[3], _drop_p(e[4]), e[5],
                            e2[0], e2[1], e2[3], _drop_p(e2[4]), e2[5])

    def forward(self, input, attention_mask=None):
        branch, resid = self.forward_fused(input, None, attention_mask)
        return branch + resid


class temporalEncoder(nn.Module):
    """meant/meant.py:124-145: temp_embedding add, RMSNorm, Linear, temporal, RMSNorm, Linear."""

    def __init__(self, dim, num_heads, lag):
        super().__init__()
        self.dim, self.num_heads = dim, num_heads
        self.temp_embedding = nn.Parameter(torch.randn(1, lag, dim))
        self.temp_encode = nn.ModuleList([RMSNorm(dim), nn.Linear(dim, dim), temporal(num_heads, dim),
                                          RMSNorm(dim), nn.Linear(dim, dim)])

    def forward(self, x):
        x = x + self.temp_embedding.to(x.dtype)  # the reference writes in place (`x +=`, :142)
        for mod in self.temp_encode:
            x = mod(x)
        return x


def patchify(images, patch):
    """'b c (h p1) (w p2) -> b (h w) (p1 p2 c)' (meant/meant.py:193-195): the library's patchify kernel, which also casts
    to the compute dtype (the GEMM operand is written once)."""
    return ops.patchify(images, patch)


class _PatchRearrange(nn.Module):
    """Parameter-free stand-in for einops' Rearrange at patchEmbed[0] (keeps `patchEmbed.1.*` keys)."""

    def __init__(self, patch):
        super().__init__()
        self.patch = patch

    def forward(self, images):
        return patchify(images, self.patch)


def _run_stack(encoders, x, *extra):
    """Run a ModuleList of encoder blocks; the add that closes block i is fused into block i+1's first norm."""
    x = x.to(ops.compute_dtype(x))   # one cast of the residual stream (bf16 under autocast)
    res = None
    for enc in encoders:
        if hasattr(enc, "forward_fused"):
            x, res = enc.forward_fused(x, res, *extra)
        else:
            if res is not None:
                x, res = x + res, None
            x = enc(x, *extra)
    return x if res is None else x + res


class meant(nn.Module):
    def __init__(self, text_dim, image_dim, price_dim, height, width, patch_res, lag, num_classes, embedding,
                 flash=False, num_heads=8, num_encoders=1, channels=4):
        super().__init__()
        self.lag, self.text_dim, self.image_dim = lag, text_dim, image_dim
        self.dim = text_dim + image_dim
        self.num_heads = num_heads
        self.channels = channels
        self.patch_dim = channels * patch_res * patch_res
        self.n = int((height * width) / (patch_res ** 2))
        self.embedding = nn.ModuleList([embedding])
        self.patchEmbed = nn.Sequential(_PatchRearrange(patch_res), nn.Linear(self.patch_dim, image_dim))
        self.visionEncoders = nn.ModuleList([visionEncoder(image_dim, num_heads, flash=flash)
                                             for _ in range(num_encoders)])
        self.languageEncoders = nn.ModuleList([languageEncoder(text_dim, num_heads, flash=flash)
                                               for _ in range(num_encoders)])
        self.temporal_encoding = nn.ModuleList([temporalEncoder(self.dim, num_heads, lag)])
        self.mlpHead = nn.ModuleList([RMSNorm(self.dim), nn.Linear(self.dim, num_classes), nn.Sigmoid()])

    def forward(self, tweets=None, images=None, prices=None, attention_mask=None, **kw):
        """forward(tweets, images, attention_mask=None) as meant/meant.py:206, and
        forward(tweets, images, prices) as the north star / meant/meantPrice.py:178 (prices are
        not consumed by the top-level model: its temporal width is text_dim + image_dim).
        Keyword aliases of the `src` tree (input_ids, pixels) are accepted."""
        tweets = kw.get("input_ids", tweets)
        images = kw.get("pixels", images)
        if prices is not None and attention_mask is None and prices.shape == tweets.shape:
            attention_mask, prices = prices, None          # positional (tweets, images, mask)
        b = images.shape[0]
        words = tweets.reshape(b * self.lag, tweets.shape[2])
        for mod in self.embedding:
            words = ops.embed(mod, words)
        if attention_mask is not None:
            attention_mask = attention_mask.reshape(b * self.lag, attention_mask.shape[2])
        words = _run_stack(self.languageEncoders, words, attention_mask)
        img = self.patchEmbed(images.reshape(b * self.lag, *images.shape[2:]))
        img = _run_stack(self.visionEncoders, img)
        t = torch.cat((ops.mean_pool(words).view(b, self.lag, -1), ops.mean_pool(img).view(b, self.lag, -1)), dim=2)
        for enc in self.temporal_encoding:
            t = enc(t)
        for mod in self.mlpHead:
            t = mod(t)
        return t.squeeze(dim=1).float()
