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
        assert model.unused.weight.grad is None
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
