"""GPU tests of the training-step plumbing around the hot path: the bias-gradient kernel (meant_bias_grad),
the Linear autograd wrapper that uses it, device-side dropout seeds and the CUDA-graph TrainStep."""
import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _lib():
    from meant_b200 import _native as nv
    return nv


@pytest.mark.parametrize("rows,cols,dtype", [(4097, 768, torch.bfloat16), (98304, 768, torch.bfloat16),
                                             (1000, 2304, torch.bfloat16), (63, 8, torch.bfloat16),
                                             (513, 772, torch.float32), (0, 768, torch.bfloat16)])
def test_bias_grad_matches_column_sum(rows, cols, dtype):
    nv = _lib()
    torch.manual_seed(rows + cols)
    dy = torch.randn(rows, cols, device="cuda").to(dtype)
    db = torch.full((cols,), 7.0, device="cuda")
    rc = nv.lib().meant_bias_grad(nv.ptr(dy), nv.ptr(db), rows, cols, cols, nv.dt(dy), nv.stream())
    assert rc == 0
    ref = dy.double().sum(0)
    tol = 1e-3 * max(1.0, rows ** 0.5)
    assert float((db.double() - ref).abs().max()) <= tol


def test_bias_grad_strided_rows_and_argument_errors():
    nv = _lib()
    wide = torch.randn(2000, 3 * 768, device="cuda").bfloat16()
    dy = wide[:, 768:1536]                                   # the k slice of a packed qkv gradient
    db = torch.empty(768, device="cuda")
    assert nv.lib().meant_bias_grad(nv.ptr(dy), nv.ptr(db), 2000, 768, wide.stride(0), nv.dt(dy), nv.stream()) == 0
    assert float((db.double() - dy.double().sum(0)).abs().max()) < 0.1
    assert nv.lib().meant_bias_grad(None, nv.ptr(db), 4, 768, 768, nv.BF16, None) == -1
    assert nv.lib().meant_bias_grad(nv.ptr(dy), nv.ptr(db), 4, 766, 768, nv.BF16, None) == -2     # cols % 8


@pytest.mark.parametrize("autocast", [True, False])
def test_fused_linear_matches_autograd(autocast):
    from meant_b200 import ops
    torch.manual_seed(0)
    lin = nn.Linear(768, 1536).cuda()
    x = torch.randn(16, 128, 768, device="cuda", requires_grad=True)
    dy = torch.randn(16, 128, 1536, device="cuda")

    def run(fused):
        ops.use_fused_linear(fused)
        try:
            for t in (x, lin.weight, lin.bias):
                t.grad = None
            with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
                y = ops.linear(x, lin.weight, lin.bias)
            y.backward(dy.to(y.dtype))
            return y.detach().float(), x.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone()
        finally:
            ops.use_fused_linear(False)

    got, ref = run(True), run(False)
    tol = 2e-2 if autocast else 2e-3          # fp32 GEMMs may use TF32 on either side
    for g, r in zip(got, ref):
        assert g.dtype == r.dtype and g.shape == r.shape
        assert float((g - r).abs().max()) <= tol * max(1.0, float(r.abs().max()))


def test_device_seed_advances_dropout_masks():
    from meant_b200 import ops
    x = torch.randn(256, 768, device="cuda").bfloat16()
    g = torch.ones(768, device="cuda")
    seed_t = torch.zeros(1, dtype=torch.int64, device="cuda")
    ops.set_graph_seed(seed_t)
    try:
        torch.manual_seed(3)
        a = ops.rms_norm(x, g, drop_p=0.5)
        seed_t.add_(12345)
        torch.manual_seed(3)                  # same host seed: only the device addend differs
        b = ops.rms_norm(x, g, drop_p=0.5)
        seed_t.sub_(12345)
        torch.manual_seed(3)
        c = ops.rms_norm(x, g, drop_p=0.5)
    finally:
        ops.set_graph_seed(None)
    assert torch.equal(a, c)
    assert not torch.equal(a == 0, b == 0)
    assert 0.45 < float((b == 0).float().mean()) < 0.55


def _tiny_model():
    import meant_b200 as mb
    torch.manual_seed(0)
    return mb.meant_tweet(768, 4, 3, 2, nn.Embedding(1000, 768), num_encoders=2).cuda()


def _batch(seed):
    g = torch.Generator().manual_seed(seed)
    tweets = torch.randint(0, 1000, (8, 3, 128), generator=g).cuda()
    mask = (torch.arange(128).view(1, 1, 128) < 100).float().expand(8, 3, 128).contiguous().cuda()
    labels = torch.randint(0, 2, (8,), generator=g).cuda()
    return labels, tweets, mask


def test_graph_step_follows_eager_trajectory():
    """eval() switches the dropouts off, so the captured step must reproduce the eager one step for step
    (fresh inputs every step: the static buffers really are reloaded)."""
    from meant_b200 import ops
    from meant_b200.train import TrainStep
    losses = {}
    try:
        for mode in (False, True):
            model = _tiny_model().eval()
            step = TrainStep(model, lr=1e-4, bf16=True, graph=mode, graph_warmup=2)
            losses[mode] = [float(step(*_batch(k))) for k in range(6)]
            if mode:
                assert step._g is not None and step.launches_per_replay > 20
    finally:
        ops.use_fused_linear(False)
    for a, b in zip(losses[False], losses[True]):
        assert abs(a - b) <= 5e-3 * max(1.0, abs(a)), (losses[False], losses[True])


def test_graph_step_train_mode_runs_and_learns():
    from meant_b200 import ops
    from meant_b200.train import TrainStep
    try:
        model = _tiny_model().train()
        step = TrainStep(model, lr=2e-4, bf16=True, graph=True, graph_warmup=2)
        batch = _batch(0)
        losses = [float(step(*batch)) for _ in range(12)]
    finally:
        ops.use_fused_linear(False)
    assert all(l == l for l in losses)                       # finite
    assert len(set(round(l, 6) for l in losses[3:])) > 1     # replays are not frozen
    assert min(losses[6:]) < losses[0]


@pytest.mark.parametrize("autocast", [True, False])
def test_embedding_matches_torch(autocast):
    from meant_b200 import ops
    torch.manual_seed(1)
    emb = nn.Embedding(5000, 768).cuda()
    ids = torch.randint(0, 5000, (7, 3, 128), device="cuda")
    ids[0, 0, :64] = 17                                      # heavy collisions on one row
    dy = torch.randn(7, 3, 128, 768, device="cuda")
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        got = ops.embed(emb, ids)
    assert got.dtype == (torch.bfloat16 if autocast else torch.float32)
    ref = emb(ids)
    assert torch.equal(got.float(), ref.to(got.dtype).float())
    got.backward(dy.to(got.dtype))
    g = emb.weight.grad.clone()
    emb.weight.grad = None
    ref.backward(dy.to(got.dtype).float())
    assert float((g - emb.weight.grad).abs().max()) <= 1e-3 * max(1.0, float(emb.weight.grad.abs().max()))


def test_embedding_other_modules_pass_through():
    from meant_b200 import ops
    emb = nn.Sequential(nn.Embedding(100, 64), nn.LayerNorm(64)).cuda()
    ids = torch.randint(0, 100, (2, 5), device="cuda")
    assert torch.equal(ops.embed(emb, ids), emb(ids))


def test_norm_backward_supplies_linear_bias_gradient():
    """Linear -> RMSNorm(+GELU/dropout/residual): with the Linear wrapper active the norm's backward kernel also
    sums dx over rows and the Linear's backward uses it instead of launching the column-sum kernel."""
    from meant_b200 import ops
    nv = _lib()
    torch.manual_seed(0)
    lin = nn.Linear(768, 768).cuda()
    gain = torch.rand(768, device="cuda", requires_grad=True)
    x = torch.randn(4096, 768, device="cuda", requires_grad=True)
    res = torch.randn(4096, 768, device="cuda").bfloat16()
    dz = torch.randn(4096, 768, device="cuda").bfloat16()

    def run(fused, **kw):
        ops.use_fused_linear(fused)
        try:
            for t in (x, lin.weight, lin.bias, gain):
                t.grad = None
            nv.reset_launch_count()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                z = ops.rms_norm(ops.linear(x, lin.weight, lin.bias), gain, **kw)
            z = z[0] if isinstance(z, tuple) else z
            z.backward(dz)
            return lin.bias.grad.clone(), x.grad.clone(), nv.launch_count()
        finally:
            ops.use_fused_linear(False)

    for kw in ({}, {"pre_gelu": True}, {"residual": res}):
        b1, x1, n1 = run(True, **kw)
        b0, x0, n0 = run(False, **kw)
        assert n1 == 2                                   # norm forward + norm backward: no bias_grad launch
        assert float((b1 - b0).abs().max()) <= 2e-2 * max(1.0, float(b0.abs().max()))
        assert float((x1 - x0).abs().max()) <= 2e-2 * max(1.0, float(x0.abs().max()))


@pytest.mark.parametrize("own", [True, False])
def test_clip_folded_into_adamw_matches_clip_grad_norm(own):
    from meant_b200.train import TrainStep
    torch.manual_seed(0)
    ref = nn.Sequential(nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, 3)).cuda()
    mine = nn.Sequential(nn.Linear(64, 64), nn.Tanh(), nn.Linear(64, 3)).cuda()
    mine.load_state_dict(ref.state_dict())
    x = torch.randn(32, 64, device="cuda") * 30.0            # large gradients: the clip is active
    y = torch.randint(0, 3, (32,), device="cuda")
    opt = torch.optim.AdamW(ref.parameters(), lr=1e-2, weight_decay=0.0, fused=True)
    step = TrainStep(mine, lr=1e-2, bf16=False, max_grad_norm=1.0, own_optimizer=own)
    for _ in range(3):
        opt.zero_grad()
        F.cross_entropy(ref(x), y).backward()
        norm = torch.nn.utils.clip_grad_norm_(ref.parameters(), 1.0)
        opt.step()
        step(y, x)
    assert float(norm) > 1.0
    for a, b in zip(ref.parameters(), mine.parameters()):
        assert torch.allclose(a, b, atol=1e-6, rtol=1e-5)
        if not own:   # torch's fused AdamW writes the clipped gradient back; the library's update leaves gradients alone
            assert torch.allclose(a.grad, b.grad, atol=1e-7, rtol=1e-4)


def test_native_gemm_in_the_linear_wrapper():
    """The Linear wrapper's three GEMM routes: library everywhere (+ meant_bias_grad), the default (library forward and
    dgrad, meant_gemm_bf16_wgrad for dW and db in one kernel), and use_native_gemm (forward, dgrad and wgrad through
    the hand-written tcgen05 kernels): same outputs and gradients (the native forward adds the fp32 bias before
    rounding, the library path adds a bf16 bias: <= 1 bf16 ulp apart)."""
    from meant_b200 import ops
    torch.manual_seed(0)
    lin = nn.Linear(768, 2304).cuda()
    x = torch.randn(8, 128, 768, device="cuda", requires_grad=True)
    dy = torch.randn(8, 128, 2304, device="cuda").bfloat16()
    res = {}
    try:
        ops.use_fused_linear(True)
        for mode in ("library", "default", "native"):
            ops.use_native_gemm(mode == "native")
            ops.use_native_wgrad(mode != "library")
            for t in (x, lin.weight, lin.bias):
                t.grad = None
            nv = _lib()
            nv.reset_launch_count()
            with torch.autocast("cuda", dtype=torch.bfloat16):
                y = ops.linear(x, lin.weight, lin.bias)
            y.backward(dy)
            res[mode] = (y.detach().float(), x.grad.clone(), lin.weight.grad.clone(), lin.bias.grad.clone(), nv.launch_count())
    finally:
        ops.use_native_gemm(False)
        ops.use_native_wgrad(True)
        ops.use_fused_linear(False)
    assert res["library"][4] == 1 and res["default"][4] == 1 and res["native"][4] == 3     # bias_grad | wgrad | fwd, dgrad, wgrad
    for mode in ("default", "native"):
        for a, b in zip(res[mode][:4], res["library"][:4]):
            assert float((a - b).abs().max()) <= 2e-2 * max(1.0, float(b.abs().max())), mode
