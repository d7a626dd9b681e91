"""TimeSformer divided space-time attention path with the reference's constructor and
state_dict, on the sm_100a kernels.  Reference: src/meant/timesformer_pytorch.py:17-263,
src/utils/rotary.py:7-62."""
from math import log, pi

import torch
from torch import nn

from . import ops


class _FrameRotary(nn.Module):
    """src/utils/rotary.py:49-61 (`RotaryEmbedding`): angles half-split cat(freqs, freqs), applied
    by the interleaved rotate_every_two -- reproduced as is."""

    def __init__(self, dim):
        super().__init__()
        self.register_buffer('inv_freqs', 1. / (10000 ** (torch.arange(0, dim, 2).float() / dim)))

    @torch.no_grad()
    def tables(self, n, device):
        ang = torch.arange(n, device=device, dtype=torch.float32)[:, None] * self.inv_freqs.float().to(device)[None, :]
        ang = torch.cat((ang, ang), dim=-1)
        return ang.cos().contiguous(), ang.sin().contiguous()


class _AxialRotary(nn.Module):
    """src/utils/rotary.py:21-47 (`AxialRotaryEmbedding`)."""

    def __init__(self, dim, max_freq=10):
        super().__init__()
        self.dim = dim
        self.register_buffer('scales', torch.logspace(0., log(max_freq / 2) / log(2), dim // 4, base=2))

    @torch.no_grad()
    def tables(self, h, w, device):
        sc = self.scales.float().to(device)[None, :]
        hs = torch.linspace(-1., 1., steps=h, device=device)[:, None] * sc * pi
        ws = torch.linspace(-1., 1., steps=w, device=device)[:, None] * sc * pi
        ang = torch.cat((hs[:, None, :].expand(h, w, -1), ws[None, :, :].expand(h, w, -1)), dim=-1)
        ang = torch.repeat_interleave(ang.reshape(h * w, -1), 2, dim=-1)
        return ang.cos().contiguous(), ang.sin().contiguous()


def _with_cls_row(cos, sin):
    """Key-side tables for [tokens..., cls]: the last row is the identity rotation (the cls key joins every group
    un-rotated, src/meant/timesformer_pytorch.py:125-133; softmax attention does not depend on the key order, and the
    kernels take the shared cls row as the LAST key)."""
    one = torch.ones_like(cos[:1])
    return torch.cat((cos, one), 0).contiguous(), torch.cat((sin, torch.zeros_like(sin[:1])), 0).contiguous()


class PreNorm(nn.Module):
    def __init__(self, dim, fn):
        super().__init__()
        self.fn = fn
        self.norm = nn.LayerNorm(dim)

    def forward(self, x, *args, **kwargs):
        x = ops.layer_norm(x, self.norm.weight, self.norm.bias, self.norm.eps)
        return self.fn(x, *args, **kwargs)


class PreTokenShift(nn.Module):
    def __init__(self, frames, fn):
        super().__init__()
        self.frames = frames
        self.fn = fn

    def forward(self, x, *args, **kwargs):
        return self.fn(ops.token_shift(x, self.frames), *args, **kwargs)


class GEGLU(nn.Module):
    def forward(self, x):
        return ops.geglu(x)


class FeedForward(nn.Module):
    def __init__(self, dim, mult=4, dropout=0.):
        super().__init__()
        self.net = nn.Sequential(nn.Linear(dim, dim * mult * 2), GEGLU(), nn.Dropout(dropout),
                                 nn.Linear(dim * mult, dim))

    def forward(self, x):
        # the two Linears through the fused wrapper (bias gradient from the weight-gradient kernel instead of autograd's
        # generic reduction); the parameters stay in `net` (state_dict contract)
        x = ops.linear(x, self.net[0].weight, self.net[0].bias)
        x = self.net[2](self.net[1](x))
        return ops.linear(x, self.net[3].weight, self.net[3].bias)


class Attention(nn.Module):
    """src/meant/timesformer_pytorch.py:89-148.  `mode` is 'time' ('(b n) f d' groups) or 'space'
    ('(b f) n d' groups); the einops patterns of the reference signature are accepted and mapped."""

    def __init__(self, dim, dim_head=64, heads=8, dropout=0.):
        super().__init__()
        self.heads = heads
        self.dim_head = dim_head
        self.scale = dim_head ** -0.5
        inner_dim = dim_head * heads
        self.to_qkv = nn.Linear(dim, inner_dim * 3, bias=False)
        self.to_out = nn.Sequential(nn.Linear(inner_dim, dim), nn.Dropout(dropout))

    def forward(self, x, einops_from=None, einops_to=None, mask=None, cls_mask=None, rot_emb=None, **dims):
        if mask is not None or cls_mask is not None:
            raise NotImplementedError("frame masks are not used on the MEANT path (meant_forward passes none)")
        ntok = x.shape[1]
        time_mode = einops_to.replace(" ", "").startswith("(bn)")
        if time_mode:
            n = dims['n']
            f = (ntok - 1) // n
        else:
            f = dims['f']
            n = (ntok - 1) // f
        qkv = self.to_qkv(x)                                     # [b, 1 + f*n, 3*inner], packed q | k | v
        rot = None
        if rot_emb is not None:
            cos, sin = rot_emb
            kc, ks = _with_cls_row(cos, sin)
            rot = (cos, sin, kc, ks)
        # cls query over every key (:119) + the grouped attention (:122-139), read and written in place
        out = ops.divided_attention(qkv, heads=self.heads, scale=self.scale, frames=f, patches=n, time_mode=time_mode,
                                    rot=rot)
        return self.to_out[1](ops.linear(out, self.to_out[0].weight, self.to_out[0].bias))


class TimeSformer(nn.Module):
    def __init__(self, *, dim, num_frames, num_classes, image_size=224, patch_size=16, channels=3, depth=12,
                 heads=8, dim_head=64, attn_dropout=0., ff_dropout=0., rotary_emb=True, shift_tokens=False):
        super().__init__()
        assert image_size % patch_size == 0, 'Image dimensions must be divisible by the patch size.'
        num_patches = (image_size // patch_size) ** 2
        patch_dim = channels * patch_size ** 2
        self.heads = heads
        self.patch_size = patch_size
        self.to_patch_embedding = nn.Linear(patch_dim, dim)
        self.cls_token = nn.Parameter(torch.randn(1, dim))
        self.use_rotary_emb = rotary_emb
        if rotary_emb:
            self.frame_rot_emb = _FrameRotary(dim_head)
            self.image_rot_emb = _AxialRotary(dim_head)
        else:
            self.pos_emb = nn.Embedding(num_frames * num_patches + 1, dim)
        self.layers = nn.ModuleList([])
        for _ in range(depth):
            parts = [Attention(dim, dim_head=dim_head, heads=heads, dropout=attn_dropout),
                     Attention(dim, dim_head=dim_head, heads=heads, dropout=attn_dropout),
                     FeedForward(dim, dropout=ff_dropout)]
            if shift_tokens:
                parts = [PreTokenShift(num_frames, t) for t in parts]
            self.layers.append(nn.ModuleList([PreNorm(dim, parts[0]), PreNorm(dim, parts[1]), PreNorm(dim, parts[2])]))
        self.to_out = nn.Sequential(nn.LayerNorm(dim), nn.Linear(dim, num_classes))

    def meant_forward(self, video, mask=None):
        if mask is not None:
            raise NotImplementedError("frame masks are not used on the MEANT path")
        b, f, c, hh, ww = video.shape
        p = self.patch_size
        assert hh % p == 0 and ww % p == 0, f'height {hh} and width {ww} of video must be divisible by the patch size {p}'
        hp, wp = hh // p, ww // p
        n = hp * wp
        v = ops.patchify(video.reshape(b * f, c, hh, ww), p).view(b, f * n, p * p * c)
        tokens = ops.linear(v, self.to_patch_embedding.weight, self.to_patch_embedding.bias)
        x = torch.cat((self.cls_token.to(tokens.dtype)[None].expand(b, -1, -1), tokens), dim=1)
        x = x.to(ops.compute_dtype(x))
        frot = arot = None
        if self.use_rotary_emb:
            frot = self.frame_rot_emb.tables(f, video.device)
            arot = self.image_rot_emb.tables(hp, wp, video.device)
        else:
            x = x + self.pos_emb(torch.arange(x.shape[1], device=video.device)).to(x.dtype)
        for time_attn, spatial_attn, ff in self.layers:
            x = time_attn(x, 'b (f n) d', '(b n) f d', n=n, rot_emb=frot) + x
            x = spatial_attn(x, 'b (f n) d', '(b f) n d', f=f, rot_emb=arot) + x
            x = ff(x) + x
        return x

    def forward(self, video, mask=None):
        x = self.meant_forward(video, mask=mask)
        x = ops.layer_norm(x[:, 0], self.to_out[0].weight, self.to_out[0].bias, self.to_out[0].eps)
        return self.to_out[1](x)


class dividedSpaceTimeAttention(nn.Module):
    """The reference file of this name does not parse (duplicate parameters,
    meant/dividedSpaceTimeAttention.py:14) and has no oracle.  The name is kept as an alias
    with the intended arguments, running the TimeSformer divided attention kernels: one time
    pass then one space pass over the lag x patch grid."""

    def __init__(self, num_heads, dim, n, num_frames, pos_emb=None, mask=False, droput=0.):
        super().__init__()
        self.n, self.num_frames = n, num_frames
        self.time = Attention(dim, dim_head=dim // num_heads, heads=num_heads, dropout=droput)
        self.space = Attention(dim, dim_head=dim // num_heads, heads=num_heads, dropout=droput)

    def forward(self, x):
        x = self.time(x, 'b (f n) d', '(b n) f d', n=self.n) + x
        return self.space(x, 'b (f n) d', '(b f) n d', f=self.num_frames) + x
