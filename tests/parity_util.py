"""Build the meant_b200 module for a parity case (the mirror of oracle/gen_golden.build_reference)."""
import torch
from torch import nn

import meant_b200 as mb
from oracle import cases, detfill
from oracle.cases import VOCAB


def build_module(kind, cfg):
    """-> (module, call(module, inputs))"""
    if kind == "rmsnorm":
        return mb.RMSNorm(cfg["shape"][-1]), lambda m, i: m(i["x"])
    if kind == "xpos":
        m = mb.xPosAttention(cfg["heads"], cfg["dim"], mb.RotaryEmbedding(dim=48, use_xpos=True))
        return m, lambda m, i: m(i["x"], i.get("attention_mask"))
    if kind == "pixel":
        pe = mb.RotaryEmbedding(dim=cfg["dim"] // cfg["heads"] // 2, freqs_for="pixel")
        return mb.attention(cfg["heads"], cfg["dim"], pe), lambda m, i: m(i["x"])
    if kind == "xpos_flash":
        m = mb.xPosAttention_flash(cfg["heads"], cfg["dim"], mb.RotaryEmbedding(dim=48, use_xpos=True))
        return m, lambda m, i: m(i["x"])
    if kind == "pixel_flash":
        pe = mb.RotaryEmbedding(dim=cfg["dim"] // cfg["heads"] // 2, freqs_for="pixel")
        return mb.flash_attention(cfg["heads"], cfg["dim"], pe), lambda m, i: m(i["x"])
    if kind == "temporal":
        return mb.temporal(cfg["heads"], cfg["dim"]), lambda m, i: m(i["x"])
    if kind == "temporal_src":
        return mb.temporal_src(cfg["heads"], cfg["dim"]), lambda m, i: m(i["x"])
    if kind == "lang_enc":
        return mb.languageEncoder(cfg["dim"], cfg["heads"], flash=bool(cfg.get("flash"))), \
            lambda m, i: m(i["x"], i.get("attention_mask"))
    if kind == "vis_enc":
        return mb.visionEncoder(cfg["dim"], cfg["heads"], flash=bool(cfg.get("flash"))), lambda m, i: m(i["x"])
    if kind == "temp_enc":
        if cfg["normed"]:
            return mb.temporalEncoder(cfg["dim"], cfg["heads"], cfg["lag"]), lambda m, i: m(i["x"])
        from meant_b200.meant_tweet import temporalEncoder as plain
        return plain(cfg["dim"], cfg["heads"], cfg["lag"]), lambda m, i: m(i["x"])
    emb = nn.Embedding(VOCAB, 768)
    if kind == "meant":
        m = mb.meant(768, 768, 4, cfg["hw"], cfg["hw"], 16, cfg["lag"], 2, emb, flash=False,
                     num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["tweets"], i["images"], attention_mask=i.get("attention_mask"))
    if kind == "meant_tweet":
        m = mb.meant_tweet(768, 4, cfg["lag"], 2, emb, flash=False, num_encoders=cfg["enc"])
        return m, lambda m, i: m(i["tweets"], i["attention_mask"])
    if kind == "meant_vision":
        m = mb.meant_vision(768, 4, cfg["hw"], cfg["hw"], 16, cfg["lag"], 2, flash=False,
                            num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["images"])
    if kind == "meant_vqa":
        m = mb.meant_vqa(768, 768, 4, cfg["hw"], cfg["hw"], 16, 1, 2, emb, flash=False,
                         num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["tweets"], i["images"], i.get("attention_mask"))
    if kind == "lang_pretrainer":
        m = mb.meant_language_pretrainer(cfg["enc"], 768, emb, nn.Linear(768, VOCAB))
        return m, lambda m, i: m(i["words"], i["attention_mask"])
    if kind == "timesformer":
        m = mb.TimeSformer(dim=768, num_frames=cfg["f"], num_classes=2, image_size=cfg["hw"], patch_size=16,
                           depth=cfg["depth"], shift_tokens=cfg["shift"])
        return m, lambda m, i: m.meant_forward(i["video"])
    raise KeyError(kind)


def run_build(name, device="cuda", autocast=False, want_grads=True):
    """Run a case through meant_b200 -> (out fp32 cpu, {param: grad}, {input: grad})."""
    _, kind, cfg = cases.CASE_BY_NAME[name]
    module, call = build_module(kind, cfg)
    module.load_state_dict(detfill.fill_state_dict(module.state_dict()), strict=True)
    module = module.to(device).eval()
    inp = {}
    for k, v in cases.make_inputs(name, kind, cfg).items():
        v = v.to(device)
        if torch.is_floating_point(v) and k != "attention_mask":
            v.requires_grad_(want_grads)
        inp[k] = v
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=autocast):
        out = call(module, inp)
    pg, ig = {}, {}
    if want_grads:
        probe = cases.probe_for(name, out).to(device)
        (out.float() * probe).sum().backward()
        pg = {k: p.grad.detach().float().cpu() for k, p in module.named_parameters() if p.grad is not None}
        ig = {k: v.grad.detach().float().cpu() for k, v in inp.items() if v.requires_grad and v.grad is not None}
    return out.detach().float().cpu(), pg, ig


def cosine(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    na, nb = a.norm(), b.norm()
    if na == 0 and nb == 0:
        return 1.0
    return float(torch.dot(a, b) / (na * nb).clamp_min(1e-300))


def compare_grads(mine, ref, min_cos=0.999, what="", norm_tol=None):
    """Per-tensor gradient check: cosine >= min_cos AND (norm_tol given) | |mine| / |ref| - 1 | <= norm_tol, so a
    uniform scale error (a wrong dropout keep-scale, a missing 1/world, a mis-folded softmax scale) cannot pass.
    Tensors whose true gradient is (numerically) zero -- e.g. the key bias of single-query softmax attention,
    which shifts every score equally -- are noise in both implementations: for those only the magnitude is
    bounded."""
    top = max(float(g.norm()) for g in ref.values()) if ref else 0.0
    worst = 1.0
    for k, g in ref.items():
        assert k in mine, f"{what}: no gradient for {k}"
        gn = float(g.norm())
        if gn <= 1e-5 * top:
            assert float(mine[k].norm()) <= 2e-2 * top + 1e-12, f"{what}: {k} should be ~0"
            continue
        c = cosine(mine[k], g)
        worst = min(worst, c)
        assert c >= min_cos, f"{what}: grad cosine {k}: {c}"
        if norm_tol is not None:
            ratio = float(mine[k].double().norm()) / gn
            assert abs(ratio - 1.0) <= norm_tol, f"{what}: grad norm ratio {k}: {ratio}"
    return worst
