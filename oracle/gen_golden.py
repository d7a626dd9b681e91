"""Generate tests/golden/{manifest.json,golden.npz} by running the UNMODIFIED reference
(/root/reference) on hashed weights and inputs.  TEST INFRASTRUCTURE ONLY; runs only in
the build container (the GPU box has no /root/reference).

    python -m oracle.gen_golden            # regenerate every case
    python -m oracle.gen_golden xpos_s128_mask meant_cfg1

Shims (nothing under /root/reference is edited):
  * /root/reference/meant on sys.path so `import rotary_embedding_torch` resolves to the
    vendored copy (pip rotary-embedding-torch==0.5.3 is not installed);
  * a stub `utils` package exposing RMSNorm loaded from utils/rms_norm.py by path
    (utils/__init__.py imports torchmetrics, which is missing);
  * stub torchmetrics / seaborn / matplotlib for the `src` tree;
  * meant_tweet's languageEncoder is replaced by meant.meant.languageEncoder
    (meant/meant_tweet.py:81 returns the undefined `final_resi`);
  * `flash_attn` (third party, CUDA only: flash-attn==2.5.2) is replaced by a stub whose
    flash_attn_func is the published definition of that call (softmax(scale q k^T [+ causal]) v on
    `b s h d` tensors, scale default d^-1/2, fp32 accumulation).  The reference MODULES that call it
    (meant/xPosAttention_flash.py, meant/flash_attention.py) run unmodified, so their projections,
    head-index rotary quirk and scale choices are pinned; the flash-attn kernel itself is not.
    The stub returns fp32 even for the fp16 inputs meant/flash_attention.py:42 hands it, because the
    fp32 `multi_mad` that follows only accepts fp16 under CUDA autocast;
  * `meant.meant.ampere = True` while a flash=True encoder is constructed (meant/meant.py:22-25 gates the
    flash classes on the CUDA device name; there is no device here).
"""
import importlib.util
import json
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"


def _flash_attn_func(q, k, v, dropout_p=0.0, softmax_scale=None, causal=False, window_size=(-1, -1),
                     alibi_slopes=None, deterministic=False, **kw):
    """flash_attn.flash_attn_func as published (flash_attn_interface.py docstring): q, k, v are
    (batch, seqlen, nheads, headdim); softmax_scale defaults to headdim^-1/2; causal masks key j > query i
    (equal lengths); accumulation in fp32."""
    assert dropout_p == 0.0 and tuple(window_size) == (-1, -1) and alibi_slopes is None
    qf, kf, vf = (t.float().transpose(1, 2) for t in (q, k, v))
    scale = qf.shape[-1] ** -0.5 if softmax_scale is None else softmax_scale
    scores = torch.matmul(qf, kf.transpose(-1, -2)) * scale
    if causal:
        sq, sk = scores.shape[-2:]
        scores = scores.masked_fill(~torch.ones(sq, sk, dtype=torch.bool).tril(), float("-inf"))
    return torch.matmul(torch.softmax(scores, dim=-1), vf).transpose(1, 2)


def install_shims():
    fa = types.ModuleType("flash_attn")
    fa.__path__ = []
    fa.flash_attn_func = _flash_attn_func
    fa.flash_attn_qkvpacked_func = None
    sys.modules["flash_attn"] = fa
    for m in ("torchmetrics", "seaborn", "matplotlib", "matplotlib.pyplot"):
        if m not in sys.modules:
            stub = types.ModuleType(m)
            stub.__path__ = []
            sys.modules[m] = stub
    sys.modules["torchmetrics"].F1Score = sys.modules["torchmetrics"].Precision = \
        sys.modules["torchmetrics"].Recall = sys.modules["torchmetrics"].Accuracy = object
    spec = importlib.util.spec_from_file_location("_ref_rms_norm", f"{REF}/utils/rms_norm.py")
    rms = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(rms)
    utils = types.ModuleType("utils")
    utils.RMSNorm = rms.RMSNorm
    utils.__path__ = []
    sys.modules["utils"] = utils
    sys.path.insert(0, f"{REF}/meant")
    sys.path.insert(0, REF)
    return rms.RMSNorm


def build_reference(kind, cfg, RMSNorm):
    """-> (module, call) where call(module, inputs) runs the reference forward."""
    from torch import nn
    from oracle.cases import VOCAB
    if kind == "rmsnorm":
        return RMSNorm(cfg["shape"][-1]), lambda m, i: m(i["x"])
    import meant as ref  # noqa: the reference package
    from rotary_embedding_torch import RotaryEmbedding
    if kind == "xpos":
        m = ref.xPosAttention(cfg["heads"], cfg["dim"], RotaryEmbedding(dim=48, use_xpos=True))
        return m, lambda m, i: m(i["x"], i.get("attention_mask"))
    if kind == "pixel":
        pe = RotaryEmbedding(dim=cfg["dim"] // cfg["heads"] // 2, freqs_for="pixel")
        return ref.attention(cfg["heads"], cfg["dim"], pe), lambda m, i: m(i["x"])
    if kind == "xpos_flash":
        m = ref.xPosAttention_flash(cfg["heads"], cfg["dim"], RotaryEmbedding(dim=48, use_xpos=True))
        return m, lambda m, i: m(i["x"])
    if kind == "pixel_flash":
        pe = RotaryEmbedding(dim=cfg["dim"] // cfg["heads"] // 2, freqs_for="pixel")
        return ref.flash_attention(cfg["heads"], cfg["dim"], pe), lambda m, i: m(i["x"])
    if kind == "temporal":
        return ref.temporal(cfg["heads"], cfg["dim"]), lambda m, i: m(i["x"])
    if kind == "temporal_src":
        from src.meant.temporal import temporal as t_src
        return t_src(cfg["heads"], cfg["dim"]), lambda m, i: m(i["x"])
    if kind in ("lang_enc", "vis_enc"):
        flash = bool(cfg.get("flash"))
        sys.modules["meant.meant"].ampere = flash       # device-name gate of meant/meant.py:22-25
        try:
            if kind == "lang_enc":
                m = ref.languageEncoder(cfg["dim"], cfg["heads"], flash=flash)
                assert type(m.encode[2]).__name__ == ("xPosAttention_flash" if flash else "xPosAttention")
                return m, lambda m, i: m(i["x"], i.get("attention_mask"))
            m = ref.visionEncoder(cfg["dim"], cfg["heads"], flash=flash)
            assert type(m.encode[2]).__name__ == ("flash_attention" if flash else "attention")
            return m, lambda m, i: m(i["x"])
        finally:
            sys.modules["meant.meant"].ampere = False
    if kind == "temp_enc":
        mm, mv = sys.modules["meant.meant"], sys.modules["meant.meant_vision"]
        cls = mm.temporalEncoder if cfg["normed"] else mv.temporalEncoder
        return cls(cfg["dim"], cfg["heads"], cfg["lag"]), lambda m, i: m(i["x"].clone())
    emb = nn.Embedding(VOCAB, 768)
    if kind == "meant":
        m = ref.meant(768, 768, 4, cfg["hw"], cfg["hw"], 16, cfg["lag"], 2, emb, flash=False,
                      num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["tweets"], i["images"], i.get("attention_mask"))
    if kind == "meant_tweet":
        mm, mt = sys.modules["meant.meant"], sys.modules["meant.meant_tweet"]
        mt.languageEncoder = mm.languageEncoder
        m = mt.meant_tweet(768, 4, cfg["lag"], 2, emb, flash=False, num_encoders=cfg["enc"])
        return m, lambda m, i: m(i["tweets"], i["attention_mask"])
    if kind == "meant_vision":
        m = ref.meant_vision(768, 4, cfg["hw"], cfg["hw"], 16, cfg["lag"], 2, flash=False,
                             num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["images"])
    if kind == "meant_vqa":
        m = ref.meant_vqa(768, 768, 4, cfg["hw"], cfg["hw"], 16, 1, 2, emb, flash=False,
                          num_encoders=cfg["enc"], channels=cfg["c"])
        return m, lambda m, i: m(i["tweets"], i["images"], i.get("attention_mask"))
    if kind == "lang_pretrainer":
        # pretrain_mlm.py runs argparse/wandb at import; execute only the class statement
        src = open(f"{REF}/pretrain_mlm.py").read()
        start = src.index("class meant_language_pretrainer")
        end = src.index("class mlm_pretrainer")
        ns = {"nn": nn, "languageEncoder": ref.languageEncoder}
        exec(compile(src[start:end], f"{REF}/pretrain_mlm.py", "exec"), ns)  # class body only
        m = ns["meant_language_pretrainer"](cfg["enc"], 768, emb, nn.Linear(768, VOCAB))
        return m, lambda m, i: m(i["words"], i["attention_mask"])
    if kind == "timesformer":
        from src.meant.timesformer_pytorch import TimeSformer
        m = TimeSformer(dim=768, num_frames=cfg["f"], num_classes=2, image_size=cfg["hw"],
                        patch_size=16, depth=cfg["depth"], shift_tokens=cfg["shift"])
        return m, lambda m, i: m.meant_forward(i["video"])
    raise KeyError(kind)


def main(argv):
    torch.set_num_threads(os.cpu_count())
    RMSNorm = install_shims()
    from oracle import cases, detfill
    gdir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
    os.makedirs(gdir, exist_ok=True)
    mpath, npath = os.path.join(gdir, "manifest.json"), os.path.join(gdir, "golden.npz")
    manifest, arrays = {}, {}
    if argv and os.path.exists(mpath):
        manifest = json.load(open(mpath))
        arrays = dict(np.load(npath))
    for name, kind, cfg in cases.CASES:
        if argv and name not in argv:
            continue
        torch.manual_seed(0)
        module, call = build_reference(kind, cfg, RMSNorm)
        sd0 = module.state_dict()
        module.load_state_dict(detfill.fill_state_dict(sd0), strict=True)
        module.eval()
        inp = cases.make_inputs(name, kind, cfg)
        for k, v in inp.items():
            if torch.is_floating_point(v) and k != "attention_mask":
                v.requires_grad_(True)
        out = call(module, inp)
        (out * cases.probe_for(name, out)).sum().backward()
        for k in [k for k in arrays if k.startswith(name + "::")]:
            del arrays[k]
        arrays[f"{name}::out_sample"], arrays[f"{name}::out_stats"] = cases.digest_output(out)
        for k, prm in module.named_parameters():
            if prm.grad is not None:
                arrays[f"{name}::grad::{k}"] = cases.digest_grad(k, prm.grad)
        for k, v in inp.items():
            if v.requires_grad and v.grad is not None:
                arrays[f"{name}::igrad::{k}"] = cases.digest_grad(name + "/" + k, v.grad)
        for k, v in sd0.items():
            if detfill.is_rotary_key(k):
                arrays[f"{name}::rot::{k}"] = v.detach().numpy().copy()
        manifest[name] = dict(kind=kind, cfg=cfg, out_shape=list(out.shape),
                              params=[[k, list(v.shape)] for k, v in sd0.items()])
        print(f"{name:28s} out{tuple(out.shape)} |out|={out.norm().item():.6g} "
              f"params={len(sd0)}", flush=True)
    json.dump(manifest, open(mpath, "w"), indent=1, sort_keys=True)
    np.savez_compressed(npath, **arrays)
    print("wrote", mpath, npath, os.path.getsize(npath), "bytes")


if __name__ == "__main__":
    main(sys.argv[1:])
