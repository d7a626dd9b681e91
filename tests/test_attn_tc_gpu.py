"""tcgen05 attention kernels against the CUDA-core exact kernels (same C ABI, same inputs) and a
torch fp32 reference of the same op: forward and backward for every tile configuration (single-tile
Sq, Sk <= 128: the persistent TMA kernels; larger: the multi-tile kernels of attn_tc.cu)."""
import math

import pytest
import torch

from meant_b200 import ops, RotaryEmbedding, _native as nv

pytestmark = pytest.mark.gpu


def torch_ref(qkv, heads, scale, causal, rot, mask):
    n, s, w = qkv.shape
    dim = w // 3
    dh = dim // heads
    q, k, v = (t.float().view(n, s, heads, dh).transpose(1, 2) for t in qkv.split(dim, dim=-1))

    def rotate(t, c, sn):
        rd = c.shape[-1]
        h_, tail = t[..., :rd], t[..., rd:]
        x = h_.reshape(*h_.shape[:-1], rd // 2, 2)
        c2, s2 = c[:s].view(s, rd // 2, 2), sn[:s].view(s, rd // 2, 2)
        y0 = x[..., 0] * c2[..., 0] - x[..., 1] * s2[..., 0]
        y1 = x[..., 1] * c2[..., 1] + x[..., 0] * s2[..., 1]
        return torch.cat((torch.stack((y0, y1), -1).reshape(h_.shape), tail), -1)

    if rot is not None:
        q, k = rotate(q, rot[0], rot[1]), rotate(k, rot[2], rot[3])
    sc = q @ k.transpose(-1, -2) * scale
    if causal:
        sc = sc.masked_fill(~torch.ones(s, s, dtype=torch.bool, device=sc.device).tril(), float("-inf"))
    if mask is not None:
        sc = sc + (1 - mask[:, None, None, :]) * -1e9
    return (sc.softmax(-1) @ v).transpose(1, 2).reshape(n, s, dim)


CASES = [  # n, s, heads, dh, causal, xpos, masked, backward
    (3, 128, 8, 96, True, True, True, True),
    (2, 128, 8, 96, False, False, False, True),
    (2, 77, 8, 96, True, True, True, True),
    (1, 1, 8, 96, True, True, False, True),
    (2, 196, 8, 96, False, True, False, True),
    (1, 512, 8, 96, True, True, True, True),
    (5, 13, 8, 64, False, False, False, True),
    (2, 197, 8, 64, False, False, False, True),
    (2, 300, 8, 96, True, True, True, True),      # three ragged tiles, causal diagonal skipping
    (1, 129, 8, 64, True, False, False, True),    # one row in the second tile
    (2, 256, 8, 96, False, False, True, True),
    # more work items than CTAs: the persistent multi-tile backward walks 4-5 items per CTA
    (40, 196, 8, 96, False, True, False, True),   # 640 items per pass
    (26, 300, 8, 96, True, True, True, True),     # 624 items, causal tiles of unequal length
    # multi-step kernels (attn_tc2.cu): Dh = 64 with a 48-wide rotation, five key tiles, three full causal tiles
    (3, 260, 8, 64, True, True, True, True),
    (2, 640, 8, 96, False, True, True, True),
    (21, 384, 8, 96, True, True, False, True),    # 504 items in the dQ pass: several per CTA, steps 1..3
    # wide-key forward (attn2_fwd_wk_kernel: 128 < Sk <= 208, not causal): padding masks, the 208-key limit, S a
    # multiple of 16 / of 32, more items than CTAs
    (3, 160, 8, 96, False, True, True, True),
    (2, 208, 8, 64, False, False, True, True),
    (2, 192, 8, 96, False, False, False, True),
    (2, 209, 8, 96, False, True, True, True),     # one key too many: the multi-step kernel
    (30, 201, 8, 64, False, True, True, True),    # 480 items
]


@pytest.mark.parametrize("n,s,heads,dh,causal,xpos,masked,backward", CASES)
def test_tcgen05_matches_exact_and_torch(n, s, heads, dh, causal, xpos, masked, backward):
    torch.manual_seed(s * 7 + n)
    dim = heads * dh
    base = torch.randn(n, s, 3 * dim, device="cuda").bfloat16()
    rot = None
    if xpos:
        rot = RotaryEmbedding(dim=48, use_xpos=True).tables(s, "cuda")
    mask = None
    if masked:
        cut = torch.randint(0, s // 4 + 1, (n, 1), device="cuda")
        mask = (torch.arange(s, device="cuda")[None] < (s - cut)).float()
    scale = 1.0 / math.sqrt(dim)
    probe = torch.randn(n, s, dim, device="cuda")
    outs, grads = {}, {}
    for name, impl in (("tc", nv.IMPL_TCGEN05), ("exact", nv.IMPL_EXACT)):
        x = base.clone().requires_grad_(True)
        o = ops.attention(x, heads=heads, scale=scale, causal=causal, rot=rot, key_mask=mask, impl=impl)
        outs[name] = o.float()
        if backward:
            (o.float() * probe).sum().backward()
            grads[name] = x.grad.float()
    torch.cuda.synchronize()
    ref = torch_ref(base, heads, scale, causal, rot, mask)
    assert torch.isfinite(outs["tc"]).all()
    assert float((outs["exact"] - ref).abs().max()) < 2e-2
    assert float((outs["tc"] - ref).abs().max()) < 2e-2, float((outs["tc"] - ref).abs().max())
    if backward:
        xr = base.float().requires_grad_(True)
        (torch_ref(xr, heads, scale, causal, rot, mask) * probe).sum().backward()
        for name in ("exact", "tc"):
            g = grads[name]
            assert torch.isfinite(g).all()
            cos = torch.nn.functional.cosine_similarity(g.flatten(), xr.grad.flatten(), dim=0)
            assert float(cos) > 0.999, (name, float(cos))
            if n > 8:   # per sequence: a gradient written to the wrong work item cannot hide in the global cosine
                per = torch.nn.functional.cosine_similarity(g.reshape(n, -1), xr.grad.reshape(n, -1), dim=1)
                assert float(per.min()) > 0.999, (name, int(per.argmin()), float(per.min()))
            for part, sl in (("dq", slice(0, dim)), ("dk", slice(dim, 2 * dim)), ("dv", slice(2 * dim, 3 * dim))):
                want = xr.grad[..., sl].flatten()
                if float(want.norm()) < 1e-6 * float(xr.grad.norm()):   # e.g. one key: dq = dk = 0 exactly
                    assert float(g[..., sl].norm()) < 1e-2 * float(xr.grad.norm())
                    continue
                c = torch.nn.functional.cosine_similarity(g[..., sl].flatten(), want, dim=0)
                assert float(c) > 0.998, (name, part, float(c))
                # magnitude too: a uniform scale error (softmax scale / log2e folded wrongly) keeps the cosine at 1
                ratio = float(g[..., sl].norm() / want.norm())
                assert abs(ratio - 1.0) <= 3e-2, (name, part, ratio)


@pytest.mark.parametrize("n,sq,sk", [(2, 200, 333), (3, 100, 300), (3, 300, 100), (20, 129, 257),
                                     (3, 300, 150), (20, 129, 208), (2, 64, 193),     # these three: wide-key forward
                                     (3, 150, 300), (2, 208, 100)])   # wide-query dK | dV pass beside multi-step / single-key-tile kernels
def test_multitile_backward_is_reproducible_and_cross_shaped(n, sq, sk):
    """Multi-tile backward: bit-identical gradients run to run (no atomics), and Sq != Sk (q | kv inputs): one query tile
    against several key tiles, several against one, and more work items than CTAs with one-row tail tiles."""
    torch.manual_seed(5)
    heads, dh = 8, 96
    dim = heads * dh
    q0 = torch.randn(n, sq, dim, device="cuda").bfloat16()
    kv0 = torch.randn(n, sk, 2 * dim, device="cuda").bfloat16()
    probe = torch.randn(n, sq, dim, device="cuda")
    scale = dh ** -0.5
    got = []
    for impl in (nv.IMPL_TCGEN05, nv.IMPL_TCGEN05, nv.IMPL_EXACT):
        q, kv = q0.clone().requires_grad_(True), kv0.clone().requires_grad_(True)
        o = ops.attention(q=q, kv=kv, heads=heads, scale=scale, impl=impl)
        (o.float() * probe).sum().backward()
        got.append((o.float(), q.grad.float(), kv.grad.float()))
    torch.cuda.synchronize()
    for a, b in zip(got[0], got[1]):
        assert torch.equal(a, b)
    qf, kvf = q0.float().requires_grad_(True), kv0.float().requires_grad_(True)
    k, v = kvf.split(dim, dim=-1)
    qh, kh, vh = (t.view(n, -1, heads, dh).transpose(1, 2) for t in (qf, k, v))
    ref = ((qh @ kh.transpose(-1, -2) * scale).softmax(-1) @ vh).transpose(1, 2).reshape(n, sq, dim)
    (ref * probe).sum().backward()
    assert float((got[0][0] - ref).abs().max()) < 2e-2
    for name, g, want in (("dq", got[0][1], qf.grad), ("dkv", got[0][2], kvf.grad)):
        c = torch.nn.functional.cosine_similarity(g.flatten(), want.flatten(), dim=0)
        assert float(c) > 0.999, (name, float(c))
        assert float((g - want).abs().max()) < 0.05 * float(want.abs().max()) + 1e-3, name


@pytest.mark.parametrize("s,causal", [(196, False), (300, True)])
def test_prerotated_and_in_flight_rotation_agree(s, causal):
    """Multi-tile kernels: Q, K rotated once by meant_attn_prerotate (default) vs rotated per tile in the kernel."""
    torch.manual_seed(s)
    heads, dh, n = 8, 96, 2
    dim = heads * dh
    base = torch.randn(n, s, 3 * dim, device="cuda").bfloat16()
    rot = RotaryEmbedding(dim=48, use_xpos=True).tables(s, "cuda")
    probe = torch.randn(n, s, dim, device="cuda")
    got = []
    try:
        for pre in (True, False):
            ops.PREROTATE = pre
            x = base.clone().requires_grad_(True)
            o = ops.attention(x, heads=heads, scale=dim ** -0.5, causal=causal, rot=rot, impl=nv.IMPL_TCGEN05)
            (o.float() * probe).sum().backward()
            got.append((o.float(), x.grad.float()))
    finally:
        ops.PREROTATE = True
    torch.cuda.synchronize()
    # outputs agree to one bf16 ulp of the value (rounding of the accumulation order; the in-flight path of the
    # multi-step forward keeps its rotation tables in fp16): 2^-7 relative, 1e-2 absolute below 1
    diff, mag = (got[0][0] - got[1][0]).abs(), got[0][0].abs().clamp(min=1.0)
    assert bool((diff <= 1e-2 * mag).all()), float((diff / mag).max())
    c = torch.nn.functional.cosine_similarity(got[0][1].flatten(), got[1][1].flatten(), dim=0)
    assert float(c) > 0.9999, float(c)


@pytest.mark.parametrize("s,causal", [(196, False), (512, True)])
def test_in_place_rotation_matches_rotated_copies(s, causal):
    """consume=True: Q and K of the projection buffer are rotated in place (no q_rot / k_rot copies).  Same outputs and
    gradients as the copying path, V untouched, Q / K overwritten with exactly what meant_attn_prerotate writes."""
    torch.manual_seed(s + 1)
    heads, dh, n = 8, 96, 3
    dim = heads * dh
    base = torch.randn(n, s, 3 * dim, device="cuda").bfloat16()
    rot = RotaryEmbedding(dim=48, use_xpos=True).tables(s, "cuda")
    probe = torch.randn(n, s, dim, device="cuda")
    got = []
    for consume in (False, True):
        x = base.clone().requires_grad_(True)
        o = ops.attention(x, heads=heads, scale=dim ** -0.5, causal=causal, rot=rot, impl=nv.IMPL_TCGEN05, consume=consume)
        (o.float() * probe).sum().backward()
        got.append((o.float(), x.grad.float(), x.detach().clone()))
    torch.cuda.synchronize()
    assert torch.equal(got[0][0], got[1][0])                      # the kernels read bit-identical rotated operands
    assert torch.equal(got[0][1], got[1][1])
    assert torch.equal(got[0][2], base)                           # copying path: input untouched
    after = got[1][2]
    assert torch.equal(after[..., 2 * dim:], base[..., 2 * dim:])                      # V
    qk = after[..., :2 * dim].view(n, s, 2 * heads, dh)
    assert torch.equal(qk[..., 48:], base[..., :2 * dim].view(n, s, 2 * heads, dh)[..., 48:])   # un-rotated tail of every head
    assert not torch.equal(qk[..., :48], base[..., :2 * dim].view(n, s, 2 * heads, dh)[..., :48])
