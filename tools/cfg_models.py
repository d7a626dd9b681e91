"""Model + synthetic-batch builders for the BASELINE.json configs that are not the headline bench line
(configs[2..4]), shared by bench.py (the `configs` array of its JSON line) and tools/cfg_bench.py.
Synthetic seeded inputs, random-init weights."""
import torch
from torch import nn

import meant_b200 as mb

DIM, HEADS, VOCAB, CLASSES = 768, 8, 64001, 2
DEFAULT_BATCH = {"vision": 16, "timesformer": 8, "full12": 16, "full": 16, "long": 32}


def build(what, batch, dev, encoders=12, depth=1, seed=1234):
    """-> (model, labels, args, description, attention shapes [(tag, Sq, Sk, Dh, sequences, causal)])"""
    g = torch.Generator().manual_seed(seed)
    b = batch
    labels = torch.randint(0, CLASSES, (b,), generator=g).to(dev)
    if what == "vision":
        lag = 12
        model = mb.meant_vision(DIM, 4, 224, 224, 16, lag, CLASSES, num_encoders=encoders)
        images = torch.randn(b, lag, 4, 224, 224, generator=g).to(dev)
        return model, labels, (images,), f"meant_vision lag=12 224x224/16 C=4 B={b} encoders={encoders} (BASELINE configs[2])", \
            [("S196x196", 196, 196, 96, b * lag, False)]
    if what == "timesformer":
        from meant_b200.timesformer_pytorch import TimeSformer
        frames = 12

        class Wrap(nn.Module):   # meant_forward returns tokens; pool + classify so that a training step exists
            def __init__(self):
                super().__init__()
                self.body = TimeSformer(dim=DIM, num_frames=frames, num_classes=CLASSES, image_size=224, patch_size=16,
                                        depth=depth, heads=8, dim_head=64)
                self.head = nn.Linear(DIM, CLASSES)

            def forward(self, video):
                return torch.sigmoid(self.head(self.body.meant_forward(video).mean(dim=1).float()))

        video = torch.randn(b, frames, 3, 224, 224, generator=g).to(dev)
        return Wrap(), labels, (video,), \
            f"TimeSformer.meant_forward 12 frames 224x224/16 B={b} depth={depth} (BASELINE configs[2])", \
            [("time/S12", 12, 13, 64, b * 196, False), ("space/S196", 196, 197, 64, b * frames, False)]
    if what in ("full", "full12"):
        lag, s = (24, 128) if what == "full" else (12, 128)
        model = mb.meant(DIM, DIM, 4, 224, 224, 16, lag, CLASSES, nn.Embedding(VOCAB, DIM), num_encoders=encoders)
        tweets = torch.randint(0, VOCAB, (b, lag, s), generator=g).to(dev)
        cut = torch.randint(0, s // 4 + 1, (b, lag, 1), generator=g)
        mask = (torch.arange(s).view(1, 1, s) < (s - cut)).float().to(dev)
        images = torch.randn(b, lag, 4, 224, 224, generator=g).to(dev)
        tag = "BASELINE configs[3]" if what == "full" else "the north star's full-modality lag-12 step"
        return model, labels, (tweets, images, mask), \
            f"meant lag={lag} S=128 text + 224x224/16 C=4 images B={b} encoders={encoders} ({tag})", \
            [("S196x196", 196, 196, 96, b * lag, False), ("S128x128", 128, 128, 96, b * lag, True)]
    if what == "long":   # the S = 512 causal xPos language stack of meant_vqa / the MLM pre-trainer (configs[4])
        s = 512

        class Stack(nn.Module):
            def __init__(self):
                super().__init__()
                self.embedding = nn.Embedding(VOCAB, DIM)
                self.languageEncoders = nn.ModuleList([mb.languageEncoder(DIM, HEADS) for _ in range(encoders)])
                self.head = nn.Linear(DIM, CLASSES)

            def forward(self, ids, mask):
                from meant_b200 import ops
                from meant_b200.meant import _run_stack      # how meant_vqa / meant_language_pretrainer run this stack
                x = _run_stack(self.languageEncoders, ops.embed(self.embedding, ids), mask)
                return torch.sigmoid(self.head(x.mean(dim=1).float()))

        ids = torch.randint(0, VOCAB, (b, s), generator=g).to(dev)
        cut = torch.randint(0, s // 4 + 1, (b, 1), generator=g)
        mask = (torch.arange(s).view(1, s) < (s - cut)).float().to(dev)
        return Stack(), labels, (ids, mask), f"languageEncoder x{encoders} S=512 causal xPos B={b} (BASELINE configs[4])", \
            [("S512x512", 512, 512, 96, b, True)]
    raise KeyError(what)


def attention_flops(sq, sk, dh, heads, seqs, causal, backward):
    """forward 4 Sq Sk Dh H per sequence (2 S^2 Dh H causal: only the lower triangle counts);
    backward 2x (4 GEMMs, the recomputed QK^T is not counted)."""
    f = (2.0 if causal else 4.0) * sq * sk * dh * heads * seqs
    return f * (2.0 if backward else 1.0)


CATEGORIES = (("library_gemm", ("nvjet", "cutlass", "gemm", "cublas", "sm100_xmma", "sm90_xmma")),
              ("attention", ("attn", "q1_")),
              ("norm", ("norm_fwd", "norm_bwd", "layernorm")),
              ("optimizer", ("adamw", "sqnorm", "multi_tensor_apply", "fusedoptimizer")),
              ("nccl", ("nccl",)))


def step_split_ms(run_step):
    """Kernel time of ONE eager step by family (torch.profiler; a profiled pass, used for the SHARES only)."""
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        run_step()
        torch.cuda.synchronize()
    split = {k: 0.0 for k, _ in CATEGORIES}
    split["other"] = 0.0
    launches = 0
    for ev in prof.events():
        if ev.device_type != torch.autograd.DeviceType.CUDA:
            continue
        launches += 1
        name = ev.name.lower()
        if "meant::g1::gemm" in name or "meant::g2::gemm" in name:
            split.setdefault("own_gemm", 0.0)
            split["own_gemm"] += ev.device_time / 1e3
            continue
        for cat, keys in CATEGORIES:
            if any(k in name for k in keys):
                split[cat] += ev.device_time / 1e3
                break
        else:
            split["other"] += ev.device_time / 1e3
    out = {k: round(v, 3) for k, v in split.items()}
    out["kernel_total"] = round(sum(split.values()), 3)
    out["launches"] = launches
    out["source"] = "torch.profiler pass over one eager step (shares; not the timed region)"
    return out
