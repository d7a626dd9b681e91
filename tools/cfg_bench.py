#!/usr/bin/env python
"""Training-step timing of the BASELINE.json configs that are NOT the bench.py line (configs[2..4]: parity-test
cases there), through the same TrainStep, on one GPU, with the per-kernel time of one step (torch.profiler).

    python tools/cfg_bench.py vision   [--batch 16] [--encoders 12]    # configs[2]: meant_vision lag 12, 224x224 / 16x16
    python tools/cfg_bench.py timesformer [--batch 8] [--depth 1]      # configs[2]: TimeSformer.meant_forward, 12 frames
    python tools/cfg_bench.py full     [--batch 16] [--encoders 12]    # configs[3]: meant lag 24, tweets + images
    python tools/cfg_bench.py long     [--batch 32] [--encoders 12]    # configs[4]: S = 512 causal xPos text stack

Synthetic seeded inputs, random-init weights, bf16 autocast, train mode, CrossEntropy + clip + fused AdamW
(SURVEY.md 8d).  CUDA events around `--steps` eager steps after `--warmup`; not a bench.py line.
"""
import argparse
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch import nn  # noqa: E402

import meant_b200 as mb  # noqa: E402
from meant_b200 import _native as nv  # noqa: E402
from meant_b200.train import TrainStep  # noqa: E402

sys.path.insert(0, os.path.join(ROOT, "tools"))
import cfg_models  # noqa: E402


def build(a, dev):
    model, labels, args, desc, _ = cfg_models.build(a.what, a.batch, dev, encoders=a.encoders, depth=a.depth)
    return model, labels, args, {}, desc


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["vision", "timesformer", "full", "full12", "long"])
    ap.add_argument("--batch", type=int, default=None)
    ap.add_argument("--encoders", type=int, default=12)
    ap.add_argument("--depth", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--top", type=int, default=25)
    ap.add_argument("--attn-impl", default="auto", choices=["auto", "exact", "tcgen05"])
    a = ap.parse_args()
    if a.batch is None:
        a.batch = cfg_models.DEFAULT_BATCH[a.what]
    from meant_b200 import ops
    ops.DEFAULT_ATTN_IMPL = {"auto": nv.IMPL_AUTO, "exact": nv.IMPL_EXACT, "tcgen05": nv.IMPL_TCGEN05}[a.attn_impl]
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model, labels, args, kwargs, desc = build(a, dev)
    model = model.to(dev).train()
    step = TrainStep(model, bf16=True, graph=True, graph_warmup=1 << 30)   # eager launches, fused-linear path on
    for _ in range(a.warmup):
        loss = step(labels, *args, **kwargs)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = step(labels, *args, **kwargs)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    print(f"# {desc}: {ms:.2f} ms/step (eager launches), {a.batch / ms * 1e3:.1f} samples/s, loss {float(loss):.4f}, "
          f"attn {a.attn_impl}, peak memory {torch.cuda.max_memory_allocated() / 2**30:.1f} GiB")
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step(labels, *args, **kwargs)
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            k = agg[ev.name[:110]]
            k[0] += ev.device_time
            k[1] += 1
    total = sum(v[0] for v in agg.values())
    print(f"# one step: {total / 1e3:.2f} ms of kernel time, {sum(v[1] for v in agg.values())} launches")
    for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:a.top]:
        print(f"{100 * us / total:6.2f}%  {us / 1e3:8.3f} ms  {n:5d}x  {us / n:8.1f} us  {name}")


if __name__ == "__main__":
    main()
