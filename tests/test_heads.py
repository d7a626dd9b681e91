"""The oracle's head restatements against the reference's own building blocks on CPU (the nn.Sequential of
src/meant/meant_tweet.py:173 executed as at :259-261; nn.CrossEntropyLoss as at pretrain_mlm.py:154,182), and -- on the
GPU -- the library's kernels against the oracle."""
import pytest
import torch
from torch import nn

from oracle import heads_oracle as ho


def _prep(d, seed=0):
    torch.manual_seed(seed)
    return nn.Sequential(nn.Linear(d, d), nn.LayerNorm(d), nn.GELU(), nn.Linear(d, 1), nn.Softmax(dim=2))


def _params(prep, prefix="p"):
    return {f"{prefix}.{k}": v.detach().clone() for k, v in prep.state_dict().items()}


def test_oracle_pool_is_the_reference_sequential():
    prep = _prep(64)
    words = torch.randn(2, 3, 10, 64)                           # [B, lag, S, d]
    want = torch.matmul(words.transpose(2, 3), prep(words)).squeeze(dim=-1)
    assert torch.allclose(ho.softmax_pool(_params(prep), "p", words), want, atol=1e-6)
    images = torch.randn(2, 11, 64)                             # [B, T, d]: Softmax(dim=2) over a length-1 axis
    want = torch.matmul(images.transpose(1, 2), prep(images)).squeeze(dim=-1)
    got = ho.softmax_pool(_params(prep), "p", images)
    assert torch.allclose(got, want, atol=1e-6) and torch.allclose(got, images.sum(1), atol=1e-5)


def test_oracle_token_ce_is_cross_entropy_loss():
    torch.manual_seed(1)
    logits, target = torch.randn(4, 7, 50), torch.randint(0, 50, (4, 7))
    target[0, :3] = -100
    want = nn.CrossEntropyLoss()(logits.view(-1, 50), target.view(-1))
    assert abs(float(ho.token_cross_entropy(logits, target)) - float(want)) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("shape,dtype", [((2, 5, 128, 768), torch.float32), ((2, 5, 128, 768), torch.bfloat16),
                                         ((3, 1, 77, 256), torch.float32), ((2, 981, 768), torch.float32),
                                         ((2, 981, 768), torch.bfloat16), ((1, 2, 512, 1024), torch.bfloat16)])
def test_softmax_pool_kernel(shape, dtype):
    import meant_b200 as mb
    d = shape[-1]
    prep = mb.make_prep(d).cuda()
    torch.manual_seed(3)
    with torch.no_grad():
        for prm in prep.parameters():
            prm.copy_(torch.randn_like(prm) * (0.3 if prm.dim() == 1 else 1.0 / d ** 0.5))
        prep[1].weight.add_(1.0)
    x = torch.randn(*shape, device="cuda")
    probe = torch.randn(*shape[:-2], d, device="cuda")
    xr = x.cpu().clone().requires_grad_(True)
    p = {k: v.detach().cpu().clone().requires_grad_(True) for k, v in _params(prep).items()}
    ref = ho.softmax_pool(p, "p", xr)
    (ref * probe.cpu()).sum().backward()
    xg = x.clone().requires_grad_(True)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=dtype == torch.bfloat16):
        out = mb.softmax_pool(xg, prep)
    (out.float() * probe).sum().backward()
    tol = 2e-5 if dtype == torch.float32 else 2e-2
    scale = max(1.0, float(ref.detach().abs().max()))
    assert float((out.detach().float().cpu() - ref.detach()).abs().max()) <= tol * scale

    def close(a, b, what):
        a, b = a.double().flatten().cpu(), b.double().flatten()
        if float(b.norm()) < 1e-9:
            assert float(a.norm()) < 1e-6, what
            return
        cos = float(torch.dot(a, b) / (a.norm() * b.norm()))
        assert cos >= (0.99999 if dtype == torch.float32 else 0.999), (what, cos)
        assert abs(float(a.norm() / b.norm()) - 1) <= (1e-4 if dtype == torch.float32 else 3e-2), what

    close(xg.grad, xr.grad, "dx")
    if len(shape) == 4:                      # 3-D input: the weights are constant, no gradient reaches the scorer
        for k, prm in prep.named_parameters():
            if k == "3.bias":
                continue                     # shifts every score equally: exactly zero
            close(prm.grad, p[f"p.{k}"].grad, k)


@pytest.mark.gpu
@pytest.mark.parametrize("rows,vocab,dtype", [(64, 64001, torch.bfloat16), (33, 512, torch.float32), (7, 1000, torch.float32)])
def test_token_cross_entropy_kernel(rows, vocab, dtype):
    import meant_b200 as mb
    torch.manual_seed(rows)
    logits = (torch.randn(rows, vocab, device="cuda") * 3).to(dtype)
    target = torch.randint(0, vocab, (rows,), device="cuda")
    target[::3] = -100
    lr = logits.float().cpu().clone().requires_grad_(True)
    ref = ho.token_cross_entropy(lr, target.cpu())
    (ref * 2.0).backward()
    lg = logits.clone().requires_grad_(True)
    loss = mb.token_cross_entropy(lg, target)
    (loss * 2.0).backward()
    assert abs(float(loss) - float(ref)) <= (1e-5 if dtype == torch.float32 else 2e-3) * max(1.0, abs(float(ref)))
    g, gr = lg.grad.float().cpu(), lr.grad
    assert float((g - gr).abs().max()) <= (1e-6 if dtype == torch.float32 else 2e-3) * max(1e-3, float(gr.abs().max()))
    assert float(g[::3].abs().max()) == 0.0            # ignored rows get no gradient


@pytest.mark.gpu
def test_vision_pretrainer_shell():
    """pretrain_mim.py:77-100: runs, keeps the reference's parameter names, differentiates."""
    import meant_b200 as mb
    dec = nn.Sequential(nn.Conv2d(768, 16 * 16 * 3, kernel_size=1), nn.PixelShuffle(16))
    m = mb.meant_vision_pretrainer(1, dec, 768, channels=4, height=64, width=64).cuda()
    keys = set(m.state_dict())
    assert {"patchEmbed.1.weight", "visionEncoders.0.encode.2.q.weight", "decoder.0.weight"} <= keys
    out = m(torch.randn(2, 4, 64, 64, device="cuda"))
    assert out.shape == (2, 3, 64, 64) and torch.isfinite(out).all()
    out.abs().mean().backward()
    assert all(p.grad is not None for p in m.parameters() if p.requires_grad)
