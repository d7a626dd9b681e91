"""The library's step-after-backward kernels (csrc/optim.cu, meant_b200/optim.py) against torch:
clip_grad_norm_(1.0) + torch.optim.AdamW and nn.CrossEntropyLoss on probabilities (src/in_loop_train.py:237-244,568)."""
import pytest
import torch
import torch.nn.functional as F

from meant_b200.optim import FusedAdamW, prob_cross_entropy

pytestmark = pytest.mark.gpu


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    shapes = [(768, 768), (768,), (3, 5), (1,), (70000, 33), (2304, 768), (7,), (640001,)] + [(96, 96)] * 60
    return [torch.randn(*s, generator=g).cuda().requires_grad_(True) for s in shapes]


@pytest.mark.parametrize("wd,max_norm,gscale", [(0.0, 1.0, 1.0), (0.01, 1.0, 40.0), (0.0, 0.0, 1.0), (0.05, 2.5, 1e-3)])
def test_fused_adamw_matches_torch(wd, max_norm, gscale):
    mine, ref = _params(1), _params(1)
    shadows = {id(mine[0]): (mine[0], torch.empty(768, 768, dtype=torch.bfloat16, device="cuda")),
               id(mine[4]): (mine[4], torch.empty(70000, 33, dtype=torch.bfloat16, device="cuda"))}
    opt = FusedAdamW(mine, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd, max_norm=max_norm, shadows=shadows)
    topt = torch.optim.AdamW(ref, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=wd)
    g = torch.Generator().manual_seed(5)
    for it in range(4):
        for a, b in zip(mine, ref):
            if it == 3 and a.shape == (7,):            # a parameter without a gradient this step is skipped
                a.grad = b.grad = None
                continue
            gr = (torch.randn(*a.shape, generator=g) * gscale).cuda()
            a.grad, b.grad = gr.clone(), gr.clone()
        total = None
        if max_norm > 0:
            total = torch.nn.utils.clip_grad_norm_([p for p in ref if p.grad is not None], max_norm)
        topt.step()
        opt.step()
        if total is not None:
            assert abs(float(opt.grad_norm()) / float(total) - 1.0) < 1e-5
    for a, b in zip(mine, ref):
            assert torch.allclose(a.detach(), b.detach(), rtol=2e-5, atol=2e-6), (a.shape, float((a.detach() - b.detach()).abs().max()))
    for p, sh in shadows.values():
        assert torch.equal(sh, p.detach().bfloat16())      # the shadow is the bf16 rounding of the updated parameter
    assert opt.launches >= 2


def test_fused_adamw_takes_bf16_gradient_slices():
    """The data-parallel reducer's compressed buckets: gradients arrive as bf16 slices of flat buffers through grads=."""
    mine, ref = _params(2), _params(2)
    opt = FusedAdamW(mine, lr=1e-3, weight_decay=0.0, max_norm=1.0)
    topt = torch.optim.AdamW(ref, lr=1e-3, weight_decay=0.0)
    g = torch.Generator().manual_seed(9)
    flat = torch.empty(sum(p.numel() for p in mine) + 8, dtype=torch.bfloat16, device="cuda")
    for _ in range(3):
        grads, off = {}, 1                        # odd offset: slices are only 2-byte aligned, the scalar path must cope
        for a, b in zip(mine, ref):
            gr = torch.randn(*a.shape, generator=g).cuda().bfloat16()
            view = flat[off:off + a.numel()].view_as(a)
            view.copy_(gr)
            grads[a] = view
            b.grad = gr.float()
            off += a.numel()
        torch.nn.utils.clip_grad_norm_(ref, 1.0)
        topt.step()
        opt.step(grads)
    for a, b in zip(mine, ref):
        assert torch.allclose(a.detach(), b.detach(), rtol=2e-5, atol=2e-6), a.shape


def test_prob_cross_entropy_matches_torch():
    torch.manual_seed(0)
    for b, c in ((64, 2), (1, 2), (300, 5)):
        probs = torch.sigmoid(torch.randn(b, c, device="cuda")).requires_grad_(True)
        ref_p = probs.detach().clone().requires_grad_(True)
        y = torch.randint(0, c, (b,), device="cuda")
        loss = prob_cross_entropy(probs, y)
        ref = F.cross_entropy(ref_p, y)
        (loss * 3.0).backward()
        (ref * 3.0).backward()
        assert abs(float(loss) - float(ref)) < 1e-6
        assert torch.allclose(probs.grad, ref_p.grad, rtol=1e-5, atol=1e-7)


def test_train_step_own_optimizer_follows_torch_trajectory():
    """TrainStep with the library's optimizer / loss against the same step on torch's fused AdamW + F.cross_entropy."""
    import torch.nn as nn
    import meant_b200 as mb
    from meant_b200.train import TrainStep
    losses = {}
    for own in (True, False):
        torch.manual_seed(0)
        model = mb.meant_tweet(768, 4, 3, 2, nn.Embedding(512, 768), num_encoders=1).cuda().eval()   # eval: no dropout noise
        step = TrainStep(model, lr=1e-3, bf16=True, graph=False, own_optimizer=own)
        g = torch.Generator().manual_seed(3)
        tweets = torch.randint(0, 512, (4, 3, 32), generator=g).cuda()
        mask = torch.ones(4, 3, 32).cuda()
        labels = torch.randint(0, 2, (4,), generator=g).cuda()
        losses[own] = [float(step(labels, tweets, mask)) for _ in range(5)]
    for a, b in zip(losses[True], losses[False]):
        assert abs(a - b) < 2e-3, (losses[True], losses[False])
    assert all(l == l and l < 10 for l in losses[True])
