"""Size-independent properties of the hot-path kernels at the FULL bench size (BASELINE configs[1]: N = 64 * 12 = 768
sequences, S = 128, 8 heads of 96), where the oracle would take minutes: batch independence, causality, the padding
mask, softmax rows summing to one, linearity in V / dO, and the norm kernels' invariances."""
import math

import pytest
import torch

from meant_b200 import ops, RotaryEmbedding, _native as nv

pytestmark = pytest.mark.gpu
N, S, H, DH = 768, 128, 8, 96
DIM = H * DH
SCALE = 1.0 / math.sqrt(DIM)


def _inputs(seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    qkv = torch.randn(N, S, 3 * DIM, device="cuda", generator=g).bfloat16()
    cut = torch.randint(0, S // 4 + 1, (N, 1), device="cuda", generator=g)
    mask = (torch.arange(S, device="cuda")[None] < (S - cut)).float()
    rot = RotaryEmbedding(dim=48, use_xpos=True).tables(S, "cuda")
    return qkv, mask, rot


def _attn(qkv, mask, rot, causal=True):
    return ops.attention(qkv, heads=H, scale=SCALE, causal=causal, rot=rot, key_mask=mask, impl=nv.IMPL_TCGEN05)


def test_attention_batch_independence_and_determinism():
    qkv, mask, rot = _inputs()
    full = _attn(qkv, mask, rot)
    assert torch.equal(full, _attn(qkv, mask, rot))                      # no atomics, no race: bit-identical reruns
    idx = torch.tensor([0, 1, 147, 148, 400, 767], device="cuda")        # first / last work items of several CTAs
    part = _attn(qkv[idx].contiguous(), mask[idx].contiguous(), rot)
    assert torch.equal(full[idx], part)                                  # a sequence does not see its neighbours


def test_attention_causality_and_padding():
    qkv, mask, rot = _inputs(1)
    out = _attn(qkv, mask, rot)
    # perturb the keys/values of the second half: the first 64 query rows must not move (causal mask)
    q2 = qkv.clone()
    q2[:, 64:, DIM:] += 1.0
    out2 = _attn(q2, mask, rot)
    assert torch.equal(out[:, :64], out2[:, :64])
    assert not torch.equal(out[:, 64:], out2[:, 64:])
    # perturb masked key positions only: with the mask (-1e9 bias) nothing changes
    q3 = qkv.clone()
    pad = (mask == 0)[:, :, None].expand(-1, -1, 2 * DIM)
    q3[:, :, DIM:][pad] = 3.0
    assert torch.equal(out, _attn(q3, mask, rot))


def test_attention_rows_are_convex_combinations_and_linear_in_v():
    qkv, mask, rot = _inputs(2)
    ones = qkv.clone()
    ones[:, :, 2 * DIM:] = 1.0
    out = _attn(ones, mask, rot).float()
    assert float((out - 1.0).abs().max()) <= 2 ** -7                     # softmax rows sum to one (bf16 P, fp32 sum)
    a = _attn(qkv, mask, rot).float()
    dbl = qkv.clone()
    dbl[:, :, 2 * DIM:] *= 2                                             # exact in bf16
    assert torch.equal(_attn(dbl, mask, rot).float(), 2 * a)             # linear in V, bit for bit


def test_attention_backward_properties_full_size():
    qkv, mask, rot = _inputs(3)
    qkv.requires_grad_(True)
    out = _attn(qkv, mask, rot)
    g = torch.Generator(device="cuda").manual_seed(9)
    dout = torch.randn(out.shape, device="cuda", generator=g).bfloat16()
    (grad,) = torch.autograd.grad(out, qkv, dout, retain_graph=True)
    (grad2,) = torch.autograd.grad(out, qkv, dout * 2, retain_graph=True)
    # linear in dO: exact in fp32 up to subnormal intermediate products (P can be ~1e-38), i.e. at most a stray
    # 1-ulp flip of a bf16 rounding among the 226 M gradient elements
    d = (grad2.float() - 2 * grad.float()).abs()
    print("tma fallbacks", nv.lib().meant_attn_tma_fallbacks(), "mismatch fraction", float((d > 0).float().mean()))
    assert float((d > 0).float().mean()) < 1e-6
    assert float((d / (2 * grad.float()).abs().clamp_min(1e-20)).max()) <= 2 ** -7
    (grad3,) = torch.autograd.grad(out, qkv, dout, retain_graph=True)
    assert torch.equal(grad3, grad)                                      # deterministic: no atomics, no races
    assert nv.lib().meant_attn_tma_fallbacks() == 0                      # everything ran on the TMA kernels
    dq, dk, dv = grad.float().split(DIM, dim=-1)
    pad = mask == 0
    assert float(dk[pad].abs().max()) == 0.0 and float(dv[pad].abs().max()) == 0.0    # masked keys get no gradient
    # sum_k dV[k] = sum_q dO[q] per (sequence, head) because every softmax row sums to one
    lhs, rhs = dv.sum(1), dout.float().sum(1)
    assert float((lhs - rhs).abs().max()) <= 2e-2 * max(1.0, float(rhs.abs().max()))


def test_rmsnorm_invariances_full_size():
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(N * S, DIM, device="cuda", generator=g).bfloat16()
    gain = torch.rand(DIM, device="cuda", generator=g) + 0.5
    y = ops.rms_norm(x, gain)
    assert torch.equal(y, ops.rms_norm(x, gain))
    y4 = ops.rms_norm(x * 4, gain)                                       # scale invariance up to eps = 1e-8
    diff = (y4.float() - y.float()).abs()
    assert float((diff > 0).float().mean()) < 1e-4 and float((diff / y.float().abs().clamp_min(1e-6)).max()) <= 2 ** -7
    rms = (y.float() / gain).pow(2).mean(-1).sqrt()
    assert float((rms - 1).abs().max()) < 2e-2                           # unit RMS before the gain
    idx = torch.randperm(N * S, device="cuda", generator=g)[:4096]
    assert torch.equal(ops.rms_norm(x[idx].contiguous(), gain), y[idx])  # rows are independent
    # fused residual form == separate add (the sum is rounded to bf16 once, as stored)
    r = torch.randn(N * S, DIM, device="cuda", generator=g).bfloat16()
    y2, s2 = ops.rms_norm(x, gain, residual=r)
    assert torch.equal(s2, (x.float() + r.float()).bfloat16())
    assert torch.equal(y2, ops.rms_norm(s2, gain))


def test_dropout_mask_statistics_and_backward_consistency():
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(N * S, DIM, device="cuda", generator=g).bfloat16().requires_grad_(True)
    gain = torch.ones(DIM, device="cuda", requires_grad=True)
    torch.manual_seed(11)
    y = ops.rms_norm(x, gain, pre_gelu=True, drop_p=0.5)
    keep = y != 0
    frac = float(keep.float().mean())
    assert abs(frac - 0.5) < 2e-3
    assert abs(float(keep.float().mean(0).std()) - math.sqrt(0.25 / (N * S))) < 2e-3   # columns are unbiased
    # backward regenerates the same mask: dropped outputs pass no gradient to the gain
    (dg,) = torch.autograd.grad(y, gain, torch.ones_like(y), retain_graph=True)
    yhat = torch.where(keep, y.float(), torch.zeros((), device="cuda"))
    assert float((dg - yhat.sum(0)).abs().max()) <= 2e-2 * float(yhat.sum(0).abs().max())
