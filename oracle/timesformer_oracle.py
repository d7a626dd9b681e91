"""Oracle for the TimeSformer divided space-time path and the `src` temporal variant.
TEST INFRASTRUCTURE ONLY.  Citations relative to /root/reference.
"""
import math

import torch
import torch.nn.functional as F

from .meant_oracle import lin, rotate_pairs, split_heads, merge_heads, softmax_attention


def frame_rotary(inv_freqs, f):
    """RotaryEmbedding.forward, src/utils/rotary.py:49-61: angles are HALF-SPLIT
    (cat(freqs, freqs)) but applied with the interleaved rotate_every_two (:7-11)."""
    ang = torch.arange(f, dtype=inv_freqs.dtype)[:, None] * inv_freqs[None, :]
    ang = torch.cat((ang, ang), dim=-1)
    return ang.sin(), ang.cos()           # (f, dim_head)


def axial_rotary(scales, hp, wp):
    """AxialRotaryEmbedding.forward, src/utils/rotary.py:21-47."""
    hs = torch.linspace(-1.0, 1.0, hp)[:, None] * scales[None, :] * math.pi   # (hp, d/4)
    ws = torch.linspace(-1.0, 1.0, wp)[:, None] * scales[None, :] * math.pi
    x = hs[:, None, :].expand(hp, wp, -1)
    y = ws[None, :, :].expand(hp, wp, -1)
    ang = torch.cat((x, y), dim=-1).reshape(hp * wp, -1)                      # (n, d/2)
    ang = torch.repeat_interleave(ang, 2, dim=-1)                             # (n, d)
    return ang.sin(), ang.cos()


def _rot(t, sin, cos):
    """apply_rot_emb, src/utils/rotary.py:13-19 (rot_dim == dim_head for both tables)."""
    rd = sin.shape[-1]
    head, tail = t[..., :rd], t[..., rd:]
    return torch.cat((head * cos + rotate_pairs(head) * sin, tail), dim=-1)


def divided_attention(p, pre, x, heads, mode, f, n, rot):
    """Attention.forward, src/meant/timesformer_pytorch.py:108-148.
    x: (b, 1+f*n, dim).  mode 'time': groups '(b n) f d'; 'space': groups '(b f) n d'.
    The cls query attends to every key; every group also sees the cls key/value."""
    b, ntok, _ = x.shape
    q, k, v = lin(p, pre + ".to_qkv", x).chunk(3, dim=-1)
    q, k, v = (split_heads(t, heads) for t in (q, k, v))          # b h N d
    dh = q.shape[-1]
    q = q * dh ** -0.5
    cls_out = softmax_attention(q[:, :, :1], k, v, 1.0)           # b h 1 d
    qg, kg, vg = (t[:, :, 1:].reshape(b, heads, f, n, dh) for t in (q, k, v))
    if mode == "time":
        qg, kg, vg = (t.transpose(2, 3) for t in (qg, kg, vg))    # b h n f d
    sin, cos = rot                                                 # (group_len, dh)
    qg, kg = _rot(qg, sin, cos), _rot(kg, sin, cos)
    g = qg.shape[2]
    ck = k[:, :, None, :1].expand(b, heads, g, 1, dh)
    cv = v[:, :, None, :1].expand(b, heads, g, 1, dh)
    kg = torch.cat((ck, kg), dim=3)
    vg = torch.cat((cv, vg), dim=3)
    att = torch.softmax(torch.matmul(qg, kg.transpose(-1, -2)), dim=-1)
    og = torch.matmul(att, vg)                                     # b h g len d
    if mode == "time":
        og = og.transpose(2, 3)
    out = torch.cat((cls_out, og.reshape(b, heads, f * n, dh)), dim=2)
    return lin(p, pre + ".to_out.0", merge_heads(out))


def token_shift(x, f):
    """PreTokenShift.forward, src/meant/timesformer_pytorch.py:35-55: split the feature
    axis in thirds (dim//3 each, remainder untouched); chunk 0 is shifted one frame
    back (t <- t+1), chunk 2 one frame forward, zero padded; the cls token is not shifted."""
    cls_x, body = x[:, :1], x[:, 1:]
    b, fn, d = body.shape
    body = body.view(b, f, fn // f, d)
    c = d // 3
    back = F.pad(body[..., :c], (0, 0, 0, 0, -1, 1))       # shift(t, -1)
    fwd = F.pad(body[..., 2 * c:3 * c], (0, 0, 0, 0, 1, -1))  # shift(t, +1)
    body = torch.cat((back, body[..., c:2 * c], fwd, body[..., 3 * c:]), dim=-1)
    return torch.cat((cls_x, body.reshape(b, fn, d)), dim=1)


def timesformer_forward(p, video, heads=8, patch=16, shift_tokens=False):
    """TimeSformer.meant_forward, src/meant/timesformer_pytorch.py:208-259 (rotary_emb=True,
    no frame mask).  Returns (b, 1+f*n, dim)."""
    b, f, c, hh, ww = video.shape
    hp, wp = hh // patch, ww // patch
    n = hp * wp
    v = video.view(b, f, c, hp, patch, wp, patch).permute(0, 1, 3, 5, 4, 6, 2)
    tokens = lin(p, "to_patch_embedding", v.reshape(b, f * n, patch * patch * c))
    x = torch.cat((p["cls_token"][None].expand(b, -1, -1), tokens), dim=1)
    frot = frame_rotary(p["frame_rot_emb.inv_freqs"], f)
    arot = axial_rotary(p["image_rot_emb.scales"], hp, wp)
    inner = ".fn.fn" if shift_tokens else ".fn"
    depth = 1 + max(int(k.split(".")[1]) for k in p if k.startswith("layers."))
    for i in range(depth):
        def pre_norm(j, t):
            t = F.layer_norm(t, t.shape[-1:], p[f"layers.{i}.{j}.norm.weight"],
                             p[f"layers.{i}.{j}.norm.bias"])
            return token_shift(t, f) if shift_tokens else t
        x = divided_attention(p, f"layers.{i}.0{inner}", pre_norm(0, x), heads, "time", f, n, frot) + x
        x = divided_attention(p, f"layers.{i}.1{inner}", pre_norm(1, x), heads, "space", f, n, arot) + x
        h = lin(p, f"layers.{i}.2{inner}.net.0", pre_norm(2, x))
        a, gates = h.chunk(2, dim=-1)                    # GEGLU, :59-62
        x = lin(p, f"layers.{i}.2{inner}.net.3", a * F.gelu(gates)) + x
    return x


def temporal_attention_src(p, pre, x, heads):
    """src/meant/temporal.py:38-74 with rot_embed=None: K from `k`, V from `v` (NOT
    cross-wired), scale 1/sqrt(Dh), output flattened 'b (l h d)' -> (B, dim).  With a
    rot_embed the xPos table has one row at position 0 (angle 0, scale**0): the
    identity, so it is not restated."""
    dim = x.shape[-1]
    dh = max(dim // heads, 1)
    q = split_heads(lin(p, pre + ".q", x[:, -1:, :]), heads)
    k = split_heads(lin(p, pre + ".k", x), heads)
    v = split_heads(lin(p, pre + ".v", x), heads)
    o = softmax_attention(q, k, v, 1.0 / math.sqrt(dh))
    return lin(p, pre + ".multi_mad", merge_heads(o).reshape(x.shape[0], -1))
