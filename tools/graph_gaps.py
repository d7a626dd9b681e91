"""Where does a replayed step spend its time between kernels?  Replays the headline TrainStep graph under
torch.profiler (CUDA activities only) and prints: span of one replay, sum of kernel durations, idle time, and the idle
time grouped by the kernel that FOLLOWS each gap (a kernel that starts late is usually waiting for launch latency).
  python tools/graph_gaps.py [--batch 64]"""
import argparse
import collections
import os
import sys

import torch
from torch import nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import meant_b200 as mb  # noqa: E402
from meant_b200.train import TrainStep  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--top", type=int, default=25)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = mb.meant_tweet(bench.DIM, 4, bench.LAG, bench.CLASSES, nn.Embedding(bench.VOCAB, bench.DIM), num_encoders=12).to(dev)
    model.train()
    step = TrainStep(model, bf16=True, graph=True, graph_warmup=3)
    tweets, mask, labels = (t.to(dev) for t in bench.synth_batch(args.batch, 0))
    for _ in range(6):
        step(labels, tweets, mask)
    torch.cuda.synchronize()
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as pr:
        for _ in range(3):
            step(labels, tweets, mask)
        torch.cuda.synchronize()
    ev = sorted((e for e in pr.events() if e.device_type.name == "CUDA" and e.time_range.end > e.time_range.start),
                key=lambda e: e.time_range.start)
    # the last replay: events after the last long gap
    n = len(ev) // 3
    ev = ev[-n:]
    span = (ev[-1].time_range.end - ev[0].time_range.start) / 1e3
    busy = sum(e.time_range.end - e.time_range.start for e in ev) / 1e3
    gaps = collections.defaultdict(lambda: [0, 0.0])
    idle = 0.0
    for a, b in zip(ev, ev[1:]):
        g = (b.time_range.start - a.time_range.end) / 1e3
        if g > 0:
            idle += g
            k = gaps[b.name[:70]]
            k[0] += 1
            k[1] += g
    print(f"# one replay: {len(ev)} kernels, span {span:.2f} ms, kernels {busy:.2f} ms, idle {idle:.2f} ms "
          f"({idle / len(ev) * 1e3:.2f} us per launch)")
    for name, (c, t) in sorted(gaps.items(), key=lambda kv: -kv[1][1])[:5]:
        print(f"  {t:7.3f} ms idle before {c:4d}x ({t / c * 1e3:6.2f} us each)  {name}")
    by = collections.defaultdict(lambda: [0, 0.0])
    for e in ev:
        k = by[e.name[:90]]
        k[0] += 1
        k[1] += (e.time_range.end - e.time_range.start) / 1e3
    print("# kernel time inside the replay")
    for name, (c, t) in sorted(by.items(), key=lambda kv: -kv[1][1])[:args.top]:
        print(f"  {t:7.3f} ms {100 * t / busy:5.1f} %  {c:4d}x {t / c * 1e3:8.1f} us  {name}")
    step.close()


if __name__ == "__main__":
    main()
