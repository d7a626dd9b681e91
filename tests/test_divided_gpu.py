"""TimeSformer divided space-time attention in place (ops.divided_attention: packed-group tcgen05 kernels, the
shared cls key, the single-query kernels) against a plain torch fp32 restatement of
src/meant/timesformer_pytorch.py:108-145, which rearranges, repeats and concatenates like the reference does;
and the `dividedSpaceTimeAttention` alias (meant/dividedSpaceTimeAttention.py does not parse: no oracle)."""
import pytest
import torch

import meant_b200 as mb
from meant_b200 import ops, _native as nv
from meant_b200.timesformer_pytorch import Attention, _with_cls_row

pytestmark = pytest.mark.gpu


def rotate(t, cos, sin):
    """interleaved-pair rotation, tables [len, d] broadcast over the leading axis"""
    x = t.reshape(*t.shape[:-1], -1, 2)
    c, s = cos.view(*cos.shape[:-1], -1, 2), sin.view(*sin.shape[:-1], -1, 2)
    y0 = x[..., 0] * c[..., 0] - x[..., 1] * s[..., 0]
    y1 = x[..., 1] * c[..., 1] + x[..., 0] * s[..., 1]
    return torch.stack((y0, y1), -1).reshape(t.shape)


def ref_divided(qkv, heads, scale, f, n, time_mode, rot):
    b, tokens, w = qkv.shape
    inner = w // 3
    dh = inner // heads
    q, k, v = (t.reshape(b, tokens, heads, dh).permute(0, 2, 1, 3).reshape(b * heads, tokens, dh)
               for t in qkv.float().split(inner, dim=-1))
    q = q * scale
    cls_out = torch.softmax(q[:, :1] @ k.transpose(1, 2), -1) @ v                      # :119
    q_, k_, v_ = q[:, 1:], k[:, 1:], v[:, 1:]

    def group(t):                                                                       # :122
        t = t.reshape(b * heads, f, n, dh)
        return t.transpose(1, 2).reshape(b * heads * n, f, dh) if time_mode else t.reshape(b * heads * f, n, dh)

    q_, k_, v_ = group(q_), group(k_), group(v_)
    if rot is not None:                                                                 # :125-126
        q_, k_ = rotate(q_, rot[0], rot[1]), rotate(k_, rot[0], rot[1])
    r = q_.shape[0] // (b * heads)
    cls_k, cls_v = (t[:, :1].repeat_interleave(r, dim=0) for t in (k, v))               # :129-130
    k_, v_ = torch.cat((cls_k, k_), 1), torch.cat((cls_v, v_), 1)                      # :132-133
    o = torch.softmax(q_ @ k_.transpose(1, 2), -1) @ v_                                 # :136
    if time_mode:                                                                       # :139
        o = o.reshape(b * heads, n, f, dh).transpose(1, 2)
    o = torch.cat((cls_out, o.reshape(b * heads, f * n, dh)), 1)                        # :142
    return o.reshape(b, heads, tokens, dh).permute(0, 2, 1, 3).reshape(b, tokens, inner)   # :145


def tables(length, dh, device):
    g = torch.Generator().manual_seed(length)
    ang = (torch.rand(length, dh // 2, generator=g) * 6.0).repeat_interleave(2, dim=-1).to(device)
    return ang.cos().contiguous(), ang.sin().contiguous()


CASES = [  # b, f, n, heads, dh, time_mode, rotary
    (2, 12, 196, 8, 64, True, True),      # BASELINE configs[2]: 10 time groups of 12 rows per tile, partial last tile
    (1, 12, 196, 8, 64, False, True),     # space groups of 196 (+ cls = 197 keys): multi-tile kernels
    (2, 5, 16, 8, 64, False, True),       # space groups of 16: 5 groups per tile
    (2, 3, 4, 8, 64, True, False),
    (1, 24, 9, 8, 64, True, True),        # lag 24
    (1, 2, 100, 8, 64, False, False),     # one group of 100 rows per tile + the cls key
    (3, 12, 49, 4, 96, True, True),       # Dh = 96 operand tiles
    (1, 4, 127, 8, 64, False, True),      # 127 rows + cls: a full tile
    (1, 3, 128, 8, 64, False, True),      # 128 rows + cls = 129 keys: the cls key alone in the second key tile
]


@pytest.mark.parametrize("b,f,n,heads,dh,time_mode,rotary", CASES)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_divided_attention_in_place(b, f, n, heads, dh, time_mode, rotary, dtype):
    torch.manual_seed(b * 1000 + f * 10 + n)
    dev = "cuda"
    inner = heads * dh
    tokens = 1 + f * n
    base = (torch.randn(b, tokens, 3 * inner, device=dev) * 0.7).to(dtype)
    rot = ref_rot = None
    if rotary:
        rd = dh if dh == 64 else 48
        cos, sin = tables(f if time_mode else n, rd, dev)
        kc, ks = _with_cls_row(cos, sin)
        rot = (cos, sin, kc, ks)
    scale = dh ** -0.5
    probe = torch.randn(b, tokens, inner, device=dev)
    x = base.clone().requires_grad_(True)
    out = ops.divided_attention(x, heads=heads, scale=scale, frames=f, patches=n, time_mode=time_mode, rot=rot)
    (out.float() * probe).sum().backward()
    xr = base.float().clone().requires_grad_(True)

    def ref_fn(t):
        if not rotary:
            return ref_divided(t, heads, scale, f, n, time_mode, None)
        # rotate the first rd features only (rd == dh for the TimeSformer tables)
        rd = rot[0].shape[-1]
        if rd == dh:
            return ref_divided(t, heads, scale, f, n, time_mode, (rot[0], rot[1]))
        cpad = torch.cat((rot[0], torch.ones(rot[0].shape[0], dh - rd, device=dev)), -1)
        spad = torch.cat((rot[1], torch.zeros(rot[1].shape[0], dh - rd, device=dev)), -1)
        return ref_divided(t, heads, scale, f, n, time_mode, (cpad, spad))

    ref = ref_fn(xr)
    (ref * probe).sum().backward()
    torch.cuda.synchronize()
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    err = float((out.detach().float() - ref.detach()).abs().max())
    assert err <= tol * max(1.0, float(ref.detach().abs().max())), f"forward max-abs {err}"
    g, gr = x.grad.float(), xr.grad
    # per projection (q | k | v) and separately for the cls row, whose k / v gradients sum over every group
    for name, sl in (("q", slice(0, inner)), ("k", slice(inner, 2 * inner)), ("v", slice(2 * inner, 3 * inner))):
        for part, rows in (("cls", slice(0, 1)), ("body", slice(1, None))):
            a_, b_ = g[:, rows, sl].flatten().double(), gr[:, rows, sl].flatten().double()
            cos_ = float(torch.dot(a_, b_) / (a_.norm() * b_.norm()).clamp_min(1e-300))
            ratio = float(a_.norm() / b_.norm().clamp_min(1e-300))
            lim = 0.99999 if dtype == torch.float32 else 0.999
            assert cos_ >= lim, f"d{name} {part}: cosine {cos_}"
            assert abs(ratio - 1.0) <= (1e-4 if dtype == torch.float32 else 3e-2), f"d{name} {part}: norm ratio {ratio}"


def test_divided_attention_makes_no_copies():
    """The step must not launch rearrange / repeat / cat copies: only this library's kernels and the two tiny casts of
    the cls-row gradients run inside ops.divided_attention."""
    from torch.profiler import profile, ProfilerActivity
    b, f, n, heads, dh = 2, 12, 196, 8, 64
    x = torch.randn(b, 1 + f * n, 3 * heads * dh, device="cuda").bfloat16().requires_grad_(True)
    cos, sin = tables(f, dh, "cuda")
    rot = (cos, sin, *_with_cls_row(cos, sin))
    probe = torch.randn(b, 1 + f * n, heads * dh, device="cuda").bfloat16()
    for _ in range(2):
        out = ops.divided_attention(x, heads=heads, scale=0.125, frames=f, patches=n, time_mode=True, rot=rot)
        out.backward(probe)
        x.grad = None
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        out = ops.divided_attention(x, heads=heads, scale=0.125, frames=f, patches=n, time_mode=True, rot=rot)
        out.backward(probe)
        torch.cuda.synchronize()
    names = [ev.name for ev in prof.events() if ev.device_type == torch.autograd.DeviceType.CUDA]
    big = b * f * n * heads * dh           # elements of one projection: anything that size would be a rearranged copy
    assert not any("CatArray" in k or "repeat" in k.lower() for k in names), names
    copies = [k for k in names if "copy" in k.lower() or "elementwise" in k.lower()]
    assert len(copies) <= 4, copies        # the cls-row gradient casts (3) + the zero fill of their fp32 accumulators
    assert any("attn2_fwd_kernel" in k for k in names) and any("attn2_bwd_kernel" in k for k in names), names
    assert any("q1_" in k for k in names), names
    del big


def test_divided_space_time_attention_alias():
    """meant/dividedSpaceTimeAttention.py:14 (intended ctor: num_heads, dim, n, num_frames, pos_emb, mask, droput): one
    time pass then one space pass, each with its residual, built from the TimeSformer attention kernels."""
    torch.manual_seed(3)
    heads, dim, n, frames, b = 8, 512, 16, 4, 2
    m = mb.dividedSpaceTimeAttention(heads, dim, n, frames, pos_emb=None, mask=False, droput=0.).cuda()
    assert isinstance(m.time, Attention) and isinstance(m.space, Attention)
    assert m.time.heads == heads and m.time.dim_head == dim // heads
    x = torch.randn(b, 1 + frames * n, dim, device="cuda", requires_grad=True)
    out = m(x)
    assert out.shape == x.shape and torch.isfinite(out).all()
    # equals the two chained Attention calls with their residuals
    t = m.time(x, 'b (f n) d', '(b n) f d', n=n) + x
    want = m.space(t, 'b (f n) d', '(b f) n d', f=frames) + t
    assert torch.allclose(out, want, atol=1e-6, rtol=1e-6)
    # and the plain torch restatement of the same two passes
    def ref_attn(att, y, time_mode):
        qkv = torch.nn.functional.linear(y, att.to_qkv.weight)
        o = ref_divided(qkv, heads, att.scale, frames, n, time_mode, None)
        return torch.nn.functional.linear(o, att.to_out[0].weight, att.to_out[0].bias)
    rt = ref_attn(m.time, x, True) + x
    rw = ref_attn(m.space, rt, False) + rt
    assert float((out - rw).abs().max()) <= 1e-4 * max(1.0, float(rw.abs().max()))
    out.square().mean().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all() and float(x.grad.abs().max()) > 0
    for name, p in m.named_parameters():
        assert p.grad is not None and torch.isfinite(p.grad).all() and float(p.grad.abs().max()) > 0, name
    with torch.autocast("cuda", dtype=torch.bfloat16):
        ob = m(x)
    assert float((ob.float() - rw).abs().max()) <= 2e-2 * max(1.0, float(rw.abs().max()))
