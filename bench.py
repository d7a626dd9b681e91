#!/usr/bin/env python
"""MEANT lag-12 training-step benchmark (BASELINE.json metric: train samples/sec).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): meant_tweet, lag 12, S=128 Stocknet-shaped synthetic
text, bf16 autocast, forward + CrossEntropy + backward + grad-clip + AdamW, one process per GPU,
the batch sharded per rank, gradients all-reduced in buckets over NCCL (weak scaling).
Prints ONE JSON line on rank 0 (contract in the task statement; extra keys: roofline,
roofline_extra, cpu_baseline, e2e, clocks, gpu_launches).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))

import torch  # noqa: E402

LAG, SEQ, VOCAB, DIM, HEADS, CLASSES = 12, 128, 64001, 768, 8, 2
METRIC, UNIT = "MEANT lag-12 train samples/sec", "samples/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=64, help="samples per GPU per step")
    ap.add_argument("--encoders", type=int, default=12, help="languageEncoder depth (reference default 12)")
    ap.add_argument("--cpu-batch", type=int, default=2, help="samples per CPU-baseline step (bounded sample)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--attn-impl", default="auto", choices=["auto", "exact", "tcgen05"])
    ap.add_argument("--graph", type=int, default=1, help="capture the step into a CUDA graph")
    ap.add_argument("--gemm", default="library", choices=["library", "native"],
                    help="forward GEMM of the Linear layers: cuBLASLt (default) or the hand-written tcgen05 kernel")
    ap.add_argument("--configs", default="auto",
                    help="extra BASELINE configs measured in the same process and reported in the `configs` array: "
                         "auto (N=1: vision,timesformer,full12,full,long; N>1: full), none, or a comma list")
    ap.add_argument("--config-steps", type=int, default=5)
    ap.add_argument("--grad-dtype", default="auto", choices=["auto", "fp32", "bf16"],
                    help="dtype of the data-parallel gradient buckets (auto: bf16 when more than one rank)")
    return ap.parse_args()


def workload_config(args, n):
    return {"workload": "meant_tweet lag=12 S=128 text+price-shaped synthetic batch (BASELINE configs[1])",
            "lag": LAG, "seq_len": SEQ, "dim": DIM, "heads": HEADS, "vocab": VOCAB,
            "num_encoders": args.encoders, "batch_per_gpu": args.batch, "global_batch": args.batch * n,
            "parallelism": f"dp{n}", "optimizer": "AdamW + clip_grad_norm 1.0 (meant_b200.optim multi-tensor kernels)",
            "grad_buckets": "bf16" if args.grad_dtype == "bf16" or (args.grad_dtype == "auto" and n > 1) else "fp32",
            "mode": "train (dropout on)",
            "launch": "whole step (incl. the NCCL gradient all-reduces) replayed from one CUDA graph" if args.graph else "eager launches",
            "linear_forward_gemm": getattr(args, "gemm", "library"),
            "linear_weight_gradient": "meant_gemm_bf16_wgrad (hand-written tcgen05, bias gradient from an extra MMA column) where the bias "
                                      "gradient would need its own pass (q|v|k and norm-less Linears); library GEMM elsewhere"
                                      if getattr(args, "gemm", "library") == "library" else "native",
            "l2": "no flush needed: each step streams >10 GB of activations/weights, far above the 126 MB L2"}


def synth_batch(batch, rank, device="cpu", pin=False):
    g = torch.Generator().manual_seed(1234 + rank)
    tweets = torch.randint(0, VOCAB, (batch, LAG, SEQ), generator=g)
    cut = torch.randint(0, SEQ // 4 + 1, (batch, LAG, 1), generator=g)
    mask = (torch.arange(SEQ).view(1, 1, SEQ) < (SEQ - cut)).float()
    labels = torch.randint(0, CLASSES, (batch,), generator=g)
    out = (tweets, mask, labels)
    if pin:
        out = tuple(t.pin_memory() for t in out)
    return tuple(t.to(device) for t in out)


# --------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(args, steps, warmup):
    """The reference algorithm on the host cores: the oracle port (oracle/, a restatement of the
    reference's PyTorch code pinned to reference-generated goldens; /root/reference itself is
    Python and cannot travel to the GPU box).  Same model config, bounded batch."""
    from oracle import meant_models as mm
    import torch.nn.functional as F
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    from oracle.params import make_params
    p = make_params("meant_tweet_l12", encoders=args.encoders, vocab=VOCAB)
    leaves = [v for v in p.values() if v.requires_grad]
    opt = torch.optim.AdamW(leaves, lr=5e-5)
    tweets, mask, labels = synth_batch(args.cpu_batch, 0)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        out = mm.meant_tweet_forward(p, tweets, mask, HEADS)
        loss = F.cross_entropy(out, labels)
        loss.backward()
        torch.nn.utils.clip_grad_norm_(leaves, 1.0)
        opt.step()
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return {"value": args.cpu_batch / sec, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{steps} steps of batch {args.cpu_batch} (the GPU arm steps batch {args.batch} per GPU), "
                      f"fwd+bwd+clip+AdamW in fp32, {args.encoders} encoders, {sec:.2f} s/step.  Two things understate "
                      f"the CPU rate slightly: the oracle runs eval-mode dropout (identity) while the GPU arm trains "
                      f"with dropout on, and the per-step cost of clip + AdamW over all {sum(v.numel() for v in leaves) / 1e6:.0f} M "
                      f"parameters is spread over {args.cpu_batch} samples instead of {args.batch}"}, sec


# ------------------------------------------------------------------------------ clock sampling
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown," \
        "clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm = [float(r[0]) for r in self.rows if len(r) >= 7 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 7 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 7 and r[3 + i] == "Active" for r in self.rows)]
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------------------------ GPU arm
def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return d["hbm_gbs"], d["bf16_tflops"], d["bf16_tflops_sustained"], "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1590.0, 1400.0, "fallback (B200_PROFILING.md)"


def kernel_stats(timers):
    out = {}
    for name, evs in timers.items():
        ms = [a.elapsed_time(b) for a, b in evs]
        out[name] = (sum(ms) / len(ms), len(ms))
    return out


# ncu `dram__bytes_read.sum + dram__bytes_write.sum` per launch of the dominant kernels at THIS workload, read from a
# committed `ncu --set full` capture (N=768, H=8, S=128, Dh=96): equal to the algorithmic bytes, nothing is re-read.
# It is a profile-derived constant, NOT a measurement of the run that prints it: `traffic_source` says so in the line.
NCU_TRAFFIC = {"attn_fwd": 584.9e6, "attn_bwd": 1019.7e6}    # dram__bytes_read.sum + dram__bytes_write.sum per launch at 768 sequences
NCU_TRAFFIC_SOURCE = "profiles/r2s_ncu/r2s_attn128_summary.txt (ncu --set full capture of this kernel at this shape, scaled by " \
                     "sequences/768; not re-measured by this run)"


def rooflines(args, stats, steps):
    """Roofline objects from the live CUDA-event timings of this run.
    Attention at S=128 is HBM-bound on B200 (65 FLOP/B against a ridge of ~209): algorithmic bytes per
    (sequence, head) = q,k,v read + o written + LSE = 4*S*Dh*2 + 4*S forward, q,k,v,dO read + dq,dk,dv written
    + LSE = 7*S*Dh*2 + 4*S backward.  Its tensor-core fraction (the north star's other figure; causal FLOPs
    2*S^2*Dh*H per sequence, backward 2x) is listed in roofline_extra.  RMSNorm fwd 2*T*d*2 B (+1 tensor read
    for the residual form, +1 write for the kept sum), bwd 3*T*d*2 B (+1 read for the residual gradient)."""
    hbm, tf_burst, tf_sus, src = peaks()
    nseq, tokens = args.batch * LAG, args.batch * LAG * SEQ
    dh = DIM // HEADS
    flops_fwd = 2.0 * SEQ * SEQ * dh * HEADS * nseq
    main, extra = None, []
    for name, (ms, count) in sorted(stats.items()):
        if name.startswith("attn_fwd/S128x128") or name.startswith("attn_bwd/S128x128"):
            bwd = "bwd" in name
            byts = nseq * HEADS * ((7 if bwd else 4) * SEQ * dh * 2 + 4 * SEQ)
            ach = byts / (ms * 1e-3) / 1e9
            scale_n = nseq / 768.0
            r = {"kernel": name, "bound": "hbm", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s",
                 "frac": round(ach / hbm, 4), "traffic": NCU_TRAFFIC["attn_bwd" if bwd else "attn_fwd"] * scale_n,
                 "traffic_source": NCU_TRAFFIC_SOURCE, "avg_ms": round(ms, 4), "launches_per_step": count / steps, "peak_source": src,
                 "algorithmic_bytes_per_launch": byts}
            fl = flops_fwd * (2.0 if bwd else 1.0)
            tf = fl / (ms * 1e-3) / 1e12
            extra.append({"kernel": name, "bound": "tensor", "achieved": round(tf, 2), "peak": tf_sus,
                          "unit": "TFLOP/s", "frac": round(tf / tf_sus, 4), "traffic": None, "avg_ms": round(ms, 4),
                          "peak_source": src + ", sustained",
                          "note": "S=128 attention is HBM-bound (65 FLOP/B); tensor fraction listed for the north star"})
            if not bwd:
                main = r
            else:
                extra.append(r)
        elif name.startswith("rmsnorm_") and "/d768/" in name:
            res, bwd = "/res1" in name, "bwd" in name
            tensors = (3 + res) if bwd else (2 + 2 * res)
            byts = tensors * tokens * DIM * 2
            ach = byts / (ms * 1e-3) / 1e9
            extra.append({"kernel": name, "bound": "hbm", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s",
                          "frac": round(ach / hbm, 4), "traffic": None, "avg_ms": round(ms, 4),
                          "launches_per_step": count / steps, "peak_source": src})
        elif name.startswith("gemm_wgrad/") or name.startswith("gemm_dgrad/"):
            dims = [int(f[1:]) for f in name.split("/")[1:4]]
            fl = 2.0 * dims[0] * dims[1] * dims[2]
            tf = fl / (ms * 1e-3) / 1e12
            extra.append({"kernel": name + " (hand-written tcgen05, gemm_tc2.cu)", "bound": "tensor", "achieved": round(tf, 1),
                          "peak": tf_sus, "unit": "TFLOP/s", "frac": round(tf / tf_sus, 4), "traffic": None,
                          "avg_ms": round(ms, 4), "launches_per_step": count / steps, "peak_source": src + ", sustained"})
        elif name.startswith("bias_grad/"):
            r_, c_ = (int(f[1:]) for f in name.split("/")[1:3])
            if r_ < tokens:
                continue
            ach = r_ * c_ * 2 / (ms * 1e-3) / 1e9
            extra.append({"kernel": name, "bound": "hbm", "achieved": round(ach, 1), "peak": hbm, "unit": "GB/s",
                          "frac": round(ach / hbm, 4), "traffic": None, "avg_ms": round(ms, 4),
                          "launches_per_step": count / steps, "peak_source": src})
    return main, extra


def gemm_probe(args, dev):
    """The hand-written tcgen05 GEMM on the block's two Linear shapes, timed with CUDA events next
    to the library GEMM on the same operands (151 MB of activations per call: no L2 reuse between iterations)."""
    import torch.nn.functional as F
    from meant_b200 import _native as nv
    _, _, tf_sus, src = peaks()
    rows, out = args.batch * LAG * SEQ, []
    for n_out in (DIM, 3 * DIM):
        x = torch.randn(rows, DIM, device=dev).bfloat16()
        w = (torch.randn(n_out, DIM, device=dev) * 0.02).bfloat16()
        b = torch.randn(n_out, device=dev)
        bh = b.bfloat16()
        y = torch.empty(rows, n_out, device=dev, dtype=torch.bfloat16)

        def native():
            nv.check(nv.lib().meant_gemm_bf16_tn(nv.ptr(x), nv.ptr(w), nv.ptr(b), nv.ptr(y), rows, n_out, DIM, DIM, DIM, n_out,
                                                 nv.stream()), "meant_gemm_bf16_tn")

        times = {}
        for name, fn in (("native", native), ("library", lambda: F.linear(x, w, bh))):
            for _ in range(3):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                fn()
            e1.record()
            torch.cuda.synchronize()
            times[name] = e0.elapsed_time(e1) / 10
        fl = 2.0 * rows * DIM * n_out
        ach = fl / (times["native"] * 1e-3) / 1e12
        out.append({"kernel": f"gemm_bf16_tn/m{rows}/n{n_out}/k{DIM} (hand-written tcgen05; opt-in: --gemm native)",
                    "bound": "tensor", "achieved": round(ach, 1), "peak": tf_sus, "unit": "TFLOP/s",
                    "frac": round(ach / tf_sus, 4), "traffic": None, "avg_ms": round(times["native"], 4),
                    "library_same_problem_tflops": round(fl / (times["library"] * 1e-3) / 1e12, 1),
                    "peak_source": src + ", sustained"})
    return out


def extra_configs(args, dev, rank, world, barrier, max_over_ranks):
    """configs[2..4] of BASELINE.json (and the north star's full-modality lag-12 step) through the same TrainStep, a
    few steps each, in this process: ms/step, samples/s, the step's kernel-time split and a tensor-core roofline object
    for every attention kernel of the config (FLOPs per SURVEY.md 8d), the dominant one first."""
    import gc
    import cfg_models
    from meant_b200 import _native as nv
    from meant_b200.train import TrainStep
    names = args.configs
    if names == "auto":
        names = "vision,timesformer,full12,full,long" if world == 1 else "full"
    if names in ("none", ""):
        return []
    _, tf_burst, tf_sus, src = peaks()
    out = []
    for what in names.split(","):
        batch = cfg_models.DEFAULT_BATCH[what]
        torch.manual_seed(0)
        model, labels, inputs, desc, shapes = cfg_models.build(what, batch, dev, encoders=args.encoders, seed=1234 + rank)
        model = model.to(dev).train()
        gdt = torch.bfloat16 if args.grad_dtype == "bf16" or (args.grad_dtype == "auto" and world > 1) else None
        step = TrainStep(model, bf16=True, graph=bool(args.graph), graph_warmup=2, grad_dtype=gdt)
        for _ in range(4 if step.graph else 3):       # eager warm-up, capture, one replay
            step(labels, *inputs)
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(args.config_steps):
            step(labels, *inputs)
        e1.record()
        barrier()
        ms = max_over_ranks(e0.elapsed_time(e1)) / args.config_steps
        row = {"config": desc, "batch_per_gpu": batch, "n_gpus": world, "steps": args.config_steps,
               "ms_per_step": round(ms, 3), "value": round(batch * world / (ms * 1e-3), 1), "unit": UNIT,
               "launch": "CUDA graph replay" if step.graph else "eager"}
        # one eager pass of the same step for the per-kernel CUDA events and one under torch.profiler for the split:
        # EVERY rank runs them (the step holds collectives), rank 0 keeps the numbers
        nv.enable_timers(rank == 0)
        step._step(labels, inputs, {})
        torch.cuda.synchronize()
        stats = kernel_stats(nv.timers()) if rank == 0 else {}
        nv.enable_timers(False)
        split_c = cfg_models.step_split_ms(lambda: step._step(labels, inputs, {}))
        if rank == 0:
            att = []
            for tag, sq, sk, dh, seqs, causal in shapes:
                for name, (kms, count) in sorted(stats.items()):
                    if not name.startswith("attn_") or tag not in name or "q1" in name or "prerotate" in name:
                        continue
                    bwd = "_bwd" in name
                    fl = cfg_models.attention_flops(sq, sk, dh, HEADS, seqs, causal, bwd)
                    tf = fl / (kms * 1e-3) / 1e12
                    att.append({"kernel": name, "bound": "tensor", "achieved": round(tf, 2), "peak": tf_sus, "unit": "TFLOP/s",
                                "frac": round(tf / tf_sus, 4), "traffic": None, "avg_ms": round(kms, 4),
                                "launches_per_step": count, "ms_per_step": round(kms * count, 3),
                                "algorithmic_flops_per_launch": fl, "peak_source": src + ", sustained"})
            att.sort(key=lambda r: -r["ms_per_step"])
            row["roofline"] = att[0] if att else None
            row["attention_kernels"] = att
            row["aux_kernels_ms"] = {k: round(v[0] * v[1], 3) for k, v in stats.items()
                                     if k.startswith("attn_") and ("q1" in k or "prerotate" in k)}
            row["step_split_ms"] = split_c
        out.append(row)
        step.close()
        del step, model, labels, inputs
        gc.collect()
        torch.cuda.empty_cache()
    return out


def gpu_run(args):
    import torch.distributed as dist
    from torch import nn
    import meant_b200 as mb
    from meant_b200 import _native as nv
    from meant_b200.train import TrainStep
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (meant_b200 has no CPU path); use --impl reference for the CPU arm")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.attn_impl != "auto":
        import meant_b200.ops as ops
        ops.DEFAULT_ATTN_IMPL = {"exact": nv.IMPL_EXACT, "tcgen05": nv.IMPL_TCGEN05}[args.attn_impl]
    if args.gemm == "native":
        import meant_b200.ops as ops_
        ops_.use_native_gemm(True)
    torch.manual_seed(0)
    model = mb.meant_tweet(DIM, 4, LAG, CLASSES, nn.Embedding(VOCAB, DIM), num_encoders=args.encoders).to(dev)
    model.train()
    gdt = torch.bfloat16 if args.grad_dtype == "bf16" or (args.grad_dtype == "auto" and world > 1) else None
    step = TrainStep(model, bf16=True, graph=bool(args.graph), graph_warmup=min(3, args.warmup), grad_dtype=gdt)
    h_tweets, h_mask, h_labels = synth_batch(args.batch, rank, pin=True)
    d_tweets, d_mask, d_labels = (t.to(dev) for t in (h_tweets, h_mask, h_labels))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world == 1:
            return ms
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    for _ in range(args.warmup):
        step(d_labels, d_tweets, d_mask)
    if step.graph:      # the capture itself plus one replay stay outside the timed regions
        for _ in range(2):
            step(d_labels, d_tweets, d_mask)
        assert step._g is not None
    # ---- timed region 1: device-resident inputs ("value")
    sampler = ClockSampler(local) if rank == 0 else None
    barrier()
    if sampler:
        sampler.start()
    nv.reset_launch_count()
    nv.enable_timers(rank == 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        loss = step(d_labels, d_tweets, d_mask)
    e1.record()
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1))
    launches = step.launches_per_replay * args.steps if step.graph else nv.launch_count()
    stats = kernel_stats(nv.timers()) if rank == 0 and not step.graph else {}
    nv.enable_timers(False)
    # ---- timed region 2: host buffers in, loss out, every step ("e2e")
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # The loss of EVERY step is copied to pinned host memory (4 bytes, asynchronous, stream-ordered before the next
    # replay overwrites the static loss tensor) and read by the host one step later, the way a training loop logs its
    # loss: the host never stalls the GPU between a step's read-back and the next step's copies and launch.
    pin_loss = torch.empty(2, dtype=torch.float32).pin_memory()
    ev_loss = [torch.cuda.Event(), torch.cuda.Event()]
    pending, loss_host = None, float("nan")
    e2.record()
    for i in range(args.steps):
        if step.graph:      # pinned host buffers are copied straight into the graph's static inputs
            tw, mk, lb = h_tweets, h_mask, h_labels
        else:
            tw, mk, lb = (t.to(dev, non_blocking=True) for t in (h_tweets, h_mask, h_labels))
        pin_loss[i & 1].copy_(step(lb, tw, mk).detach().reshape(()), non_blocking=True)
        ev_loss[i & 1].record()
        if pending is not None:
            ev_loss[pending].synchronize()
            loss_host = float(pin_loss[pending])
        pending = i & 1
    ev_loss[pending].synchronize()
    loss_host = float(pin_loss[pending])
    e3.record()
    barrier()
    ms_e2e = max_over_ranks(e2.elapsed_time(e3))
    clocks = sampler.stop() if sampler else None
    timer_steps = args.steps
    if step.graph:
        # graph replays carry no per-kernel events: the per-launch durations behind `roofline` come from an
        # eager pass of the very same step (same kernels, shapes and stream), timed live with CUDA events
        nv.enable_timers(True)
        timer_steps = min(args.steps, 3)
        for _ in range(timer_steps):
            step._step(d_labels, (d_tweets, d_mask), {})
        torch.cuda.synchronize()
        stats = kernel_stats(nv.timers())
        nv.enable_timers(False)
    gemm_rows = gemm_probe(args, dev) if rank == 0 else []
    import cfg_models
    split = cfg_models.step_split_ms(lambda: step._step(d_labels, (d_tweets, d_mask), {}))   # every rank: collectives inside
    cfg_rows = extra_configs(args, dev, rank, world, barrier, max_over_ranks)
    if rank == 0:
        n = world
        total = args.batch * n * args.steps
        main, extra = rooflines(args, stats, timer_steps)
        extra += gemm_rows
        line = {"metric": METRIC, "value": round(total / (ms * 1e-3), 2), "unit": UNIT, "n_gpus": n,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3),
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16",
                "data": "synthetic (seeded token ids + padding masks, random-init weights)",
                "config": workload_config(args, n),
                "e2e": {"value": round(total / (ms_e2e * 1e-3), 2), "unit": UNIT,
                        "h2d_bytes_per_step": sum(t.numel() * t.element_size() for t in (h_tweets, h_mask, h_labels)),
                        "d2h_bytes_per_step": 4, "ms_per_step": round(ms_e2e / args.steps, 3),
                        "loss_read": "every step: 4-byte asynchronous copy to pinned memory, read by the host one step later"},
                "gpu_launches": int(launches), "clocks": clocks, "roofline": main, "roofline_extra": extra,
                "step_split_ms": split, "configs": cfg_rows,
                "final_loss": loss_host, "native_library": os.path.relpath(nv.lib_path(), ROOT),
                "attn_tma_fallbacks": int(nv.lib().meant_attn_tma_fallbacks())}
        if n == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"], _ = cpu_reference_run(args, steps=2, warmup=1)
        print(json.dumps(line), flush=True)
    step.close()           # the graph holds NCCL kernels: it has to go before the communicator does
    if world > 1:
        barrier()
        sys.stdout.flush()
        # TrainStep.close() released the captured graph AND the work handles created under capture (they never complete
        # outside the graph and kept the communicator's teardown waiting in round 1): a plain teardown now returns
        if os.environ.get("MEANT_BENCH_HARD_EXIT") == "1":
            os._exit(0)
        dist.destroy_process_group()


def reference_run(args):
    """--impl reference: the reference algorithm's CPU path (oracle port) on the host cores.
    Under torchrun only rank 0 works."""
    if int(os.environ.get("RANK", 0)) != 0:
        return
    base, sec = cpu_reference_run(args, steps=args.steps, warmup=args.warmup)
    cfg = workload_config(args, args.gpus)
    print(json.dumps({"impl": "reference", "note": "CPU arm: same model/config, bounded batch per step (cpu_baseline.sample)",
                      "metric": METRIC, "value": round(base["value"], 4), "unit": UNIT,
                      "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
                      "ms_per_step": round(sec * 1e3, 2), "higher_is_better": True, "scaling": "weak",
                      "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
                      "cpu_baseline": base,
                      "e2e": {"value": round(base["value"], 4), "unit": UNIT, "h2d_bytes_per_step": 0,
                              "d2h_bytes_per_step": 0}}), flush=True)


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        reference_run(a)
    else:
        gpu_run(a)
