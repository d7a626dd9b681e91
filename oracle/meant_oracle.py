"""Functional fp32 CPU restatement of the MEANT hot path.  TEST INFRASTRUCTURE ONLY.

Every function takes `p`, a flat {state_dict key: tensor} mapping with the reference's
key names, plus a key prefix, so the same hashed state dict drives the reference
modules (golden generation), this oracle and the CUDA build.  All `file:line`
citations are relative to /root/reference.
"""
import math

import torch
import torch.nn.functional as F

ROT_DIM = 48          # meant/meant.py:88-90 (language), :46-48 (vision: floor(768/8/2))
XPOS_SCALE_BASE = 512  # meant/rotary_embedding_torch.py:71


# ----------------------------------------------------------------------------- helpers
def lin(p, prefix, x):
    return F.linear(x, p[prefix + ".weight"], p.get(prefix + ".bias"))


def rmsnorm(x, scale, eps=1e-8):
    """utils/rms_norm.py:40-57: scale * x / (||x||_2 * d^-1/2 + eps); eps is added to
    the RMS, not under the square root."""
    d = x.shape[-1]
    rms = x.norm(2, dim=-1, keepdim=True) * d ** -0.5
    return scale * (x / (rms + eps))


def split_heads(t, h):
    b, s, hd = t.shape
    return t.view(b, s, h, hd // h).permute(0, 2, 1, 3)  # b h s d


def merge_heads(t):
    b, h, s, d = t.shape
    return t.permute(0, 2, 1, 3).reshape(b, s, h * d)


# ----------------------------------------------------------------------------- rotary
def rotary_angles(freqs, n):
    """meant/rotary_embedding_torch.py:127-147: angle[p, 2i] = angle[p, 2i+1] = p * freqs[i]."""
    pos = torch.arange(n, dtype=freqs.dtype)
    return torch.repeat_interleave(pos[:, None] * freqs[None, :], 2, dim=-1)


def xpos_scale(scale_buf, n):
    """meant/rotary_embedding_torch.py:112-125: scale_buf ** ((p - n//2)/512), laid out
    HALF-SPLIT (cat(scale, scale)) although the angles are interleaved."""
    power = (torch.arange(n, dtype=scale_buf.dtype) - n // 2) / XPOS_SCALE_BASE
    s = scale_buf[None, :] ** power[:, None]
    return torch.cat((s, s), dim=-1)


def rotate_pairs(x):
    """meant/rotary_embedding_torch.py:31-35: (x0, x1) -> (-x1, x0) on interleaved pairs."""
    x2 = x.reshape(*x.shape[:-1], -1, 2)
    return torch.stack((-x2[..., 1], x2[..., 0]), dim=-1).reshape(x.shape)


def apply_rotary(t, angles, scale=None):
    """meant/rotary_embedding_torch.py:37-44: first rot_dim features are rotated, the rest
    pass through.  `angles`/`scale` broadcast against t[..., :rot_dim] on dim -2."""
    rd = angles.shape[-1]
    head, tail = t[..., :rd], t[..., rd:]
    s = 1.0 if scale is None else scale
    head = head * angles.cos() * s + rotate_pairs(head) * angles.sin() * s
    return torch.cat((head, tail), dim=-1)


def xpos_rotate_qk(p, prefix, q, k):
    """rotate_queries_and_keys, meant/rotary_embedding_torch.py:101-110.  The table length
    is q.shape[-2] -- the token axis for `b h s d`, the HEAD axis for `b s h d`."""
    n = q.shape[-2]
    ang = rotary_angles(p[prefix + ".freqs"], n)
    sc = xpos_scale(p[prefix + ".scale"], n)
    return apply_rotary(q, ang, sc), apply_rotary(k, ang, sc ** -1)


def plain_rotate(p, prefix, t):
    """rotate_queries_or_keys, meant/rotary_embedding_torch.py:96-99."""
    return apply_rotary(t, rotary_angles(p[prefix + ".freqs"], t.shape[-2]))


# -------------------------------------------------------------------------- attention
def softmax_attention(q, k, v, scale, causal=False, key_mask=None):
    """softmax(q k^T * scale + causal(-inf) + (1-mask) * -1e9) v on `b h s d` tensors
    (meant/xPosAttention.py:41-63)."""
    scores = torch.matmul(q, k.transpose(-1, -2)) * scale
    if causal:
        sq, sk = scores.shape[-2:]
        keep = torch.ones(sq, sk, dtype=torch.bool).tril()
        scores = scores.masked_fill(~keep, float("-inf"))
    if key_mask is not None:
        scores = scores + (1 - key_mask[:, None, None, :].to(scores.dtype)) * -1e9
    return torch.matmul(torch.softmax(scores, dim=-1), v)


def xpos_attention(p, prefix, x, heads, attention_mask=None, causal=True):
    """meant/xPosAttention.py:35-67.  NOTE the cross-wiring at :37-38: keys come from the
    `v` Linear and values from the `k` Linear; scale is 1/sqrt(dim), not 1/sqrt(Dh)."""
    dim = x.shape[-1]
    q = split_heads(lin(p, prefix + ".q", x), heads)
    k = split_heads(lin(p, prefix + ".v", x), heads)
    v = split_heads(lin(p, prefix + ".k", x), heads)
    q, k = xpos_rotate_qk(p, prefix + ".xPos", q, k)
    o = softmax_attention(q, k, v, 1.0 / math.sqrt(dim), causal, attention_mask)
    return lin(p, prefix + ".multi_mad", merge_heads(o))


def pixel_attention(p, prefix, x, heads):
    """meant/attention.py:35-62 (same cross-wiring, pixel rotary on q and k, no mask)."""
    dim = x.shape[-1]
    q = split_heads(lin(p, prefix + ".q", x), heads)
    k = split_heads(lin(p, prefix + ".v", x), heads)
    v = split_heads(lin(p, prefix + ".k", x), heads)
    q = plain_rotate(p, prefix + ".pos_emb", q)
    k = plain_rotate(p, prefix + ".pos_emb", k)
    o = softmax_attention(q, k, v, 1.0 / math.sqrt(dim))
    return lin(p, prefix + ".multi_mad", merge_heads(o))


def xpos_attention_flash(p, prefix, x, heads, causal=True):
    """meant/xPosAttention_flash.py:35-45.  Tensors are `b s h d`, so the rotary position is the
    head index; the softmax scale is Dh^-1/2 (softmax_scale=None); no padding mask.  Pinned to the
    unmodified reference module with flash_attn_func replaced by its published definition
    (oracle/gen_golden.py); the flash-attn 2.5.2 CUDA kernel itself stays PARITY UNPINNED."""
    b, s, dim = x.shape
    dh = dim // heads
    q = lin(p, prefix + ".q", x).view(b, s, heads, dh)
    k = lin(p, prefix + ".v", x).view(b, s, heads, dh)
    v = lin(p, prefix + ".k", x).view(b, s, heads, dh)
    q, k = xpos_rotate_qk(p, prefix + ".xPos", q, k)
    o = softmax_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2),
                          dh ** -0.5, causal)
    return lin(p, prefix + ".multi_mad", merge_heads(o))


def flash_attention(p, prefix, x, heads, half_inputs=True):
    """meant/flash_attention.py:35-47.  Head-index pixel rotary, scale 1/dim, non-causal; q, k, v are
    rounded to fp16 before the attention core (`.half()`, :42) when half_inputs is set (the pinned
    form; the flash-attn kernel itself stays PARITY UNPINNED, see xpos_attention_flash)."""
    b, s, dim = x.shape
    dh = dim // heads
    rnd = (lambda t: t.half().float()) if half_inputs else (lambda t: t)
    q = rnd(plain_rotate(p, prefix + ".pos_emb", lin(p, prefix + ".q", x).view(b, s, heads, dh)))
    k = rnd(plain_rotate(p, prefix + ".pos_emb", lin(p, prefix + ".v", x).view(b, s, heads, dh)))
    v = rnd(lin(p, prefix + ".k", x).view(b, s, heads, dh))
    o = softmax_attention(q.transpose(1, 2), k.transpose(1, 2), v.transpose(1, 2), 1.0 / dim)
    return lin(p, prefix + ".multi_mad", merge_heads(o))


def temporal_attention(p, prefix, x, heads):
    """meant/temporal.py:34-60: one query (the last lag day) over all lag days; keys from
    the `v` Linear, values from the `k` Linear; scale 1/sqrt(dim).  (B,lag,D) -> (B,1,D)."""
    dim = x.shape[-1]
    q = split_heads(lin(p, prefix + ".q", x[:, -1:, :]), heads)
    k = split_heads(lin(p, prefix + ".v", x), heads)
    v = split_heads(lin(p, prefix + ".k", x), heads)
    o = softmax_attention(q, k, v, 1.0 / math.sqrt(dim))
    return lin(p, prefix + ".multi_mad", merge_heads(o))
