"""Parity cases shared by the golden generator (reference side), the oracle tests and
the GPU parity tests.  TEST INFRASTRUCTURE ONLY.

A case is (name, kind, cfg).  Inputs and weights are hashed from names (detfill), so
the three sides agree without exchanging tensors.  The scalar that is differentiated
is always  loss = sum(out * probe)  with a hashed probe of the output's shape.
"""
import torch

from . import detfill

VOCAB = 512  # stand-in for BERTweet's 64001-row table (weights cannot be downloaded)

CASES = [
    # ---- operators (meant/xPosAttention.py, attention.py, temporal.py, utils/rms_norm.py)
    ("rmsnorm_768", "rmsnorm", dict(shape=(3, 50, 768))),
    ("rmsnorm_1536", "rmsnorm", dict(shape=(2, 5, 1536))),
    ("xpos_s128_mask", "xpos", dict(b=2, s=128, dim=768, heads=8, mask=True)),
    ("xpos_s128_nomask", "xpos", dict(b=2, s=128, dim=768, heads=8, mask=False)),
    ("xpos_s77_mask", "xpos", dict(b=3, s=77, dim=768, heads=8, mask=True)),
    ("xpos_s512", "xpos", dict(b=1, s=512, dim=768, heads=8, mask=False)),
    ("pixel_n196", "pixel", dict(b=2, s=196, dim=768, heads=8)),
    ("pixel_n50", "pixel", dict(b=1, s=50, dim=768, heads=8)),
    ("temporal_1536_l5", "temporal", dict(b=4, lag=5, dim=1536, heads=8)),
    ("temporal_768_l12", "temporal", dict(b=3, lag=12, dim=768, heads=8)),
    ("temporal_768_l24", "temporal", dict(b=2, lag=24, dim=768, heads=8)),
    ("temporal_src_768_l5", "temporal_src", dict(b=3, lag=5, dim=768, heads=8)),
    ("temporal_src_772_l5", "temporal_src", dict(b=2, lag=5, dim=772, heads=8)),
    # ---- flash variants (meant/xPosAttention_flash.py, meant/flash_attention.py): head-index rotary.  Goldens come from
    #      the unmodified reference MODULES with the third-party flash_attn_func replaced by its published definition
    #      (oracle/gen_golden.py:flash_attn stub): the kernel of flash-attn 2.5.2 itself stays parity-unpinned.
    ("xpos_flash_s128", "xpos_flash", dict(b=2, s=128, dim=768, heads=8)),
    ("xpos_flash_s200", "xpos_flash", dict(b=1, s=200, dim=768, heads=8)),
    ("pixel_flash_n196", "pixel_flash", dict(b=2, s=196, dim=768, heads=8)),
    ("pixel_flash_n50", "pixel_flash", dict(b=1, s=50, dim=768, heads=8)),
    # ---- encoder blocks (meant/meant.py:35-145)
    ("lang_enc_s128", "lang_enc", dict(b=2, s=128, dim=768, heads=8, mask=True)),
    ("vis_enc_n196", "vis_enc", dict(b=1, s=196, dim=768, heads=8)),
    ("lang_enc_flash_s128", "lang_enc", dict(b=2, s=128, dim=768, heads=8, mask=False, flash=True)),
    ("vis_enc_flash_n196", "vis_enc", dict(b=1, s=196, dim=768, heads=8, flash=True)),
    ("temp_enc_1536_l5", "temp_enc", dict(b=3, lag=5, dim=1536, heads=8, normed=True)),
    ("temp_enc_768_l12", "temp_enc", dict(b=2, lag=12, dim=768, heads=8, normed=False)),
    # ---- model assemblies
    ("meant_cfg1", "meant", dict(b=2, lag=5, s=128, c=4, hw=224, enc=1, mask=True)),
    ("meant_small_2enc", "meant", dict(b=1, lag=2, s=32, c=4, hw=64, enc=2, mask=False)),
    ("meant_tweet_l12", "meant_tweet", dict(b=2, lag=12, s=128, enc=1)),
    ("meant_vision_l12", "meant_vision", dict(b=1, lag=12, c=4, hw=224, enc=1)),
    ("meant_vqa_s40", "meant_vqa", dict(b=2, s=40, c=4, hw=224, enc=1, mask=True)),
    ("lang_pretrainer_s512", "lang_pretrainer", dict(b=1, s=512, enc=2, mask=True)),
    # ---- the depth and sizes the bench runs (BASELINE.json configs[1..4]) at b=1
    ("meant_tweet_l12_12enc", "meant_tweet", dict(b=1, lag=12, s=128, enc=12)),
    ("meant_l24", "meant", dict(b=1, lag=24, s=128, c=4, hw=224, enc=1, mask=True)),
    ("lang_pretrainer_s512_12enc", "lang_pretrainer", dict(b=1, s=512, enc=12, mask=True)),
    # ---- TimeSformer divided space-time path (src/meant/timesformer_pytorch.py)
    ("timesformer_f5", "timesformer", dict(b=1, f=5, hw=224, depth=1, shift=False)),
    ("timesformer_f12_small", "timesformer", dict(b=1, f=12, hw=64, depth=1, shift=False)),
    ("timesformer_shift_small", "timesformer", dict(b=2, f=3, hw=32, depth=2, shift=True)),
    ("timesformer_f12_224", "timesformer", dict(b=1, f=12, hw=224, depth=1, shift=False)),
]

CASE_BY_NAME = {c[0]: c for c in CASES}


def make_inputs(name, kind, cfg):
    """Hashed inputs for a case: dict of tensors (floats fp32, ids int64)."""
    u = detfill.uniform
    if kind == "rmsnorm":
        return dict(x=u(cfg["shape"], name + "/x", 2.0))
    if kind in ("xpos", "lang_enc"):
        d = dict(x=u((cfg["b"], cfg["s"], cfg["dim"]), name + "/x", 1.0))
        if cfg["mask"]:
            d["attention_mask"] = detfill.padding_mask(cfg["b"], cfg["s"], name + "/mask")
        return d
    if kind in ("pixel", "vis_enc", "xpos_flash", "pixel_flash"):
        return dict(x=u((cfg["b"], cfg["s"], cfg["dim"]), name + "/x", 1.0))
    if kind in ("temporal", "temporal_src", "temp_enc"):
        return dict(x=u((cfg["b"], cfg["lag"], cfg["dim"]), name + "/x", 1.0))
    if kind == "meant":
        d = dict(tweets=detfill.integers((cfg["b"], cfg["lag"], cfg["s"]), name + "/tweets", VOCAB),
                 images=u((cfg["b"], cfg["lag"], cfg["c"], cfg["hw"], cfg["hw"]), name + "/images", 1.0))
        if cfg["mask"]:
            d["attention_mask"] = detfill.padding_mask(cfg["b"] * cfg["lag"], cfg["s"], name + "/mask") \
                .view(cfg["b"], cfg["lag"], cfg["s"])
        return d
    if kind == "meant_tweet":
        return dict(tweets=detfill.integers((cfg["b"], cfg["lag"], cfg["s"]), name + "/tweets", VOCAB),
                    attention_mask=detfill.padding_mask(cfg["b"] * cfg["lag"], cfg["s"], name + "/mask")
                    .view(cfg["b"], cfg["lag"], cfg["s"]))
    if kind == "meant_vision":
        return dict(images=u((cfg["b"], cfg["lag"], cfg["c"], cfg["hw"], cfg["hw"]), name + "/images", 1.0))
    if kind == "meant_vqa":
        d = dict(tweets=detfill.integers((cfg["b"], cfg["s"]), name + "/tweets", VOCAB),
                 images=u((cfg["b"], cfg["c"], cfg["hw"], cfg["hw"]), name + "/images", 1.0))
        if cfg["mask"]:
            d["attention_mask"] = detfill.padding_mask(cfg["b"], cfg["s"], name + "/mask")
        return d
    if kind == "lang_pretrainer":
        return dict(words=detfill.integers((cfg["b"], cfg["s"]), name + "/words", VOCAB),
                    attention_mask=detfill.padding_mask(cfg["b"], cfg["s"], name + "/mask"))
    if kind == "timesformer":
        return dict(video=u((cfg["b"], cfg["f"], 3, cfg["hw"], cfg["hw"]), name + "/video", 1.0))
    raise KeyError(kind)


def probe_for(name, out):
    return detfill.uniform(tuple(out.shape), name + "/probe", 1.0)


SAMPLE = 2048


def digest_output(out):
    """Strided sample (<= SAMPLE values) + [l2 norm, sum, numel] of an output tensor."""
    flat = out.detach().double().flatten()
    stride = max(1, flat.numel() // SAMPLE)
    return flat[::stride][:SAMPLE].float().numpy(), \
        torch.tensor([flat.norm().item(), flat.sum().item(), float(flat.numel())]).numpy()


def digest_grad(key, g):
    """[l2 norm, <g, hashed direction>] of one gradient tensor."""
    flat = g.detach().double().flatten()
    direction = detfill.uniform((flat.numel(),), key + "/dir", 1.0).double()
    return torch.tensor([flat.norm().item(), torch.dot(flat, direction).item()]).numpy()
