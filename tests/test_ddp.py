"""World-size-2 gloo tests (CPU) of the data-parallel plumbing around the hot path: meant_b200/ddp.py
(bucketed gradient all-reduce, parameter broadcast) and meant_b200/train.py (TrainStep).  The path shards by
sample with one gradient exchange per step, so two ranks on half batches must reproduce the single-process
step on the whole batch."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
import torch.nn as nn

from meant_b200.ddp import BucketedGradAllReduce
from meant_b200.train import TrainStep


class Net(nn.Module):
    def __init__(self):
        super().__init__()
        self.a = nn.Linear(12, 32)
        self.b = nn.Linear(32, 32)
        self.unused = nn.Linear(4, 4)            # never gets a gradient: exercises finish()
        self.head = nn.Linear(32, 3)

    def forward(self, x):
        return self.head(torch.tanh(self.b(torch.tanh(self.a(x)))))


def _data():
    g = torch.Generator().manual_seed(7)
    return torch.randn(16, 12, generator=g), torch.randint(0, 3, (16,), generator=g)


def _single(steps):
    torch.manual_seed(0)
    model = Net()
    step = TrainStep(model, lr=1e-2, bf16=False, bucket_mb=0.002)
    x, y = _data()
    losses = [float(step(y, x)) for _ in range(steps)]
    return [p.detach().clone() for p in model.parameters()], losses


def _worker(rank, world, port, steps, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(100 + rank)            # ranks start different: broadcast_parameters must fix it
        model = Net()
        if rank == 0:
            torch.manual_seed(0)
            model = Net()
        step = TrainStep(model, lr=1e-2, bf16=False, bucket_mb=0.002)
        assert len(step.reducer.buckets) > 1     # small cap: several buckets in flight
        x, y = _data()
        n = x.shape[0] // world
        xs, ys = x[rank * n:(rank + 1) * n], y[rank * n:(rank + 1) * n]
        losses = [float(step(ys, xs)) for _ in range(steps)]
        # a parameter no rank used still gets the (all-zero) reduced slice: every rank hands the optimizer the same set
        assert model.unused.weight.grad is not None and float(model.unused.weight.grad.abs().max()) == 0.0
        out[rank] = ([p.detach().clone() for p in model.parameters()], losses)
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.timeout(180)
def test_two_ranks_match_single_process():
    steps = 3
    ref_params, ref_losses = _single(steps)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, _free_port(), steps, out), nprocs=2, join=True)
    (p0, l0), (p1, l1) = out[0], out[1]
    for a, b, r in zip(p0, p1, ref_params):
        assert torch.equal(a, b)                                   # replicas stay identical
        assert torch.allclose(a, r, atol=1e-6, rtol=1e-5)          # and follow the single-process trajectory
    for k in range(steps):
        assert abs(0.5 * (l0[k] + l1[k]) - ref_losses[k]) < 1e-5


class BranchNet(nn.Module):
    """`extra` is used by rank 0 only (data-dependent control flow): its gradient exists on one rank."""

    def __init__(self):
        super().__init__()
        self.a = nn.Linear(12, 16)
        self.extra = nn.Linear(12, 16)
        self.head = nn.Linear(16, 3)

    def forward(self, x, use_extra):
        h = torch.tanh(self.a(x))
        if use_extra:
            h = h + torch.tanh(self.extra(x))
        return self.head(h)


def _branch_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        model = BranchNet()
        red = BucketedGradAllReduce(list(model.parameters()), bucket_mb=0.0005)
        opt = torch.optim.AdamW(model.parameters(), lr=1e-2, weight_decay=0.0)
        x, y = _data()
        n = x.shape[0] // world
        for _ in range(2):
            red.zero_grad()
            nn.functional.cross_entropy(model(x[rank * n:(rank + 1) * n], rank == 0), y[rank * n:(rank + 1) * n]).backward()
            red.finish()
            assert model.extra.weight.grad is not None            # also on the rank that never ran `extra`
            opt.step()
        # a second backward without zero_grad must not silently skip the exchange
        raised = False
        try:
            nn.functional.cross_entropy(model(x[:2], True), y[:2]).backward()
        except RuntimeError as e:
            raised = "zero_grad" in str(e)
        out[rank] = ([p.detach().clone() for p in model.parameters()], raised)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_parameter_used_on_one_rank_only_stays_in_sync():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_branch_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    (p0, r0), (p1, r1) = out[0], out[1]
    assert r0 and r1
    torch.manual_seed(0)
    init = [p.detach().clone() for p in BranchNet().parameters()]
    for a, b in zip(p0, p1):
        assert torch.equal(a, b)                                   # replicas identical although one rank skipped `extra`
    assert not torch.equal(p0[2], init[2])                         # and `extra.weight` did move on both


def _bf16_worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        model = Net()
        red = BucketedGradAllReduce(list(model.parameters()), bucket_mb=0.002, grad_dtype=torch.bfloat16)
        x, y = _data()
        n = x.shape[0] // world
        red.zero_grad()
        nn.functional.cross_entropy(model(x[rank * n:(rank + 1) * n]), y[rank * n:(rank + 1) * n]).backward()
        local = {k: p.grad.clone() for k, p in model.named_parameters() if p.grad is not None}
        red.finish()
        g = red.grads()
        assert all(p.grad is None for p in model.parameters())          # compressed mode hands gradients out via grads()
        assert all(t.dtype == torch.bfloat16 for t in g.values())
        out[rank] = ({k: g[p].float().clone() for k, p in model.named_parameters()}, local)
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_bf16_buckets_average_within_bf16_rounding():
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_bf16_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    (g0, l0), (g1, l1) = out[0], out[1]
    for k in l0:
        want = 0.5 * (l0[k] + l1[k])
        assert torch.equal(g0[k], g1[k])                                    # every rank ends with the same bits
        assert float((g0[k] - want).abs().max()) <= 2e-2 * float(want.abs().max()) + 1e-6
    assert float(g0["unused.weight"].abs().max()) == 0.0


def test_single_rank_keeps_autograd_gradients():
    """With one rank nothing is packed: .grad is whatever autograd produced, None at the start of a step."""
    torch.manual_seed(0)
    model = Net()
    red = BucketedGradAllReduce(list(model.parameters()), bucket_mb=0.002)
    assert all(flat is None for flat, _ in red.buckets)
    x, y = _data()
    nn.functional.cross_entropy(model(x), y).backward()
    red.finish()
    g = model.a.weight.grad.clone()
    red.zero_grad()
    assert model.a.weight.grad is None
    nn.functional.cross_entropy(model(x), y).backward()
    assert torch.equal(model.a.weight.grad, g)
    assert sum(red.bucket_bytes()) == sum(p.numel() * 4 for p in model.parameters())
