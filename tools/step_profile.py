#!/usr/bin/env python
"""Per-kernel time of one bench.py training step (torch.profiler / CUPTI, eager launches of the same
kernels the graph replays).  python tools/step_profile.py [--batch 64] [--encoders 12] [--top 40]"""
import argparse
import collections
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch import nn  # noqa: E402

import bench  # noqa: E402
import meant_b200 as mb  # noqa: E402
from meant_b200 import ops  # noqa: E402
from meant_b200.train import TrainStep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--encoders", type=int, default=12)
    ap.add_argument("--top", type=int, default=40)
    ap.add_argument("--fused-linear", type=int, default=1)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = mb.meant_tweet(bench.DIM, 4, bench.LAG, bench.CLASSES, nn.Embedding(bench.VOCAB, bench.DIM),
                           num_encoders=a.encoders).to(dev).train()
    # graph=True installs everything the captured step uses (Linear wrapper, weight shadows, zero arena);
    # a huge graph_warmup keeps the steps eager so that the profiler sees individual kernels
    step = TrainStep(model, bf16=True, graph=bool(a.fused_linear), graph_warmup=1 << 30)
    tweets, mask, labels = bench.synth_batch(a.batch, 0)
    tweets, mask, labels = tweets.to(dev), mask.to(dev), labels.to(dev)
    for _ in range(3):
        step(labels, tweets, mask)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        step(labels, tweets, mask)
        torch.cuda.synchronize()
    agg = collections.defaultdict(lambda: [0.0, 0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            k = agg[ev.name[:110]]
            k[0] += ev.device_time
            k[1] += 1
    total = sum(v[0] for v in agg.values())
    print(f"# one step, batch {a.batch}, {a.encoders} encoders: {total / 1e3:.2f} ms of kernel time, "
          f"{sum(v[1] for v in agg.values())} launches")
    for name, (us, n) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:a.top]:
        print(f"{100 * us / total:6.2f}%  {us / 1e3:8.3f} ms  {n:5d}x  {us / n:8.1f} us  {name}")


if __name__ == "__main__":
    main()
