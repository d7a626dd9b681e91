"""Run one parity case through the oracle.  TEST INFRASTRUCTURE ONLY."""
import json
import os

import numpy as np
import torch

from . import cases, detfill, meant_oracle as mo, meant_models as mm, timesformer_oracle as to

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def load_golden():
    with open(os.path.join(GOLDEN_DIR, "manifest.json")) as f:
        manifest = json.load(f)
    data = np.load(os.path.join(GOLDEN_DIR, "golden.npz"))
    return manifest, data


def state_from_manifest(name, manifest, data):
    """Rebuild the hashed state dict of a case from the key/shape list the reference
    constructor produced (stored at golden-generation time)."""
    sd = {}
    for key, shape in manifest[name]["params"]:
        rk = f"{name}::rot::{key}"
        if rk in data.files:
            sd[key] = torch.from_numpy(data[rk]).clone()
        else:
            sd[key] = torch.zeros(tuple(shape), dtype=torch.float32)
    return detfill.fill_state_dict(sd)


def oracle_forward(kind, cfg, p, inp):
    h = cfg.get("heads", 8)
    if kind == "rmsnorm":
        return mo.rmsnorm(inp["x"], p["scale"])
    if kind == "xpos":
        return mo.xpos_attention(p, "", inp["x"], h, inp.get("attention_mask"))
    if kind == "pixel":
        return mo.pixel_attention(p, "", inp["x"], h)
    if kind == "xpos_flash":
        return mo.xpos_attention_flash(p, "", inp["x"], h)
    if kind == "pixel_flash":
        return mo.flash_attention(p, "", inp["x"], h, half_inputs=cfg.get("half_inputs", True))
    if kind == "temporal":
        return mo.temporal_attention(p, "", inp["x"], h)
    if kind == "temporal_src":
        return to.temporal_attention_src(p, "", inp["x"], h)
    if kind == "lang_enc":
        return mm.language_encoder(p, "", inp["x"], h, inp.get("attention_mask"), flash=bool(cfg.get("flash")))
    if kind == "vis_enc":
        return mm.vision_encoder(p, "", inp["x"], h, flash=bool(cfg.get("flash")))
    if kind == "temp_enc":
        return mm.temporal_encoder(p, "", inp["x"], h, cfg["normed"])
    if kind == "meant":
        return mm.meant_forward(p, inp["tweets"], inp["images"], inp.get("attention_mask"))
    if kind == "meant_tweet":
        return mm.meant_tweet_forward(p, inp["tweets"], inp["attention_mask"])
    if kind == "meant_vision":
        return mm.meant_vision_forward(p, inp["images"])
    if kind == "meant_vqa":
        return mm.meant_vqa_forward(p, inp["tweets"], inp["images"], inp.get("attention_mask"))
    if kind == "lang_pretrainer":
        return mm.language_pretrainer_forward(p, inp["words"], inp["attention_mask"])
    if kind == "timesformer":
        return to.timesformer_forward(p, inp["video"], 8, 16, cfg["shift"])
    raise KeyError(kind)


def _strip(p):
    """Operator-level cases use un-prefixed keys ('q.weight'); the oracle functions build
    keys as prefix + '.name', so store them under a leading dot for prefix ''."""
    return {"." + k: v for k, v in p.items()} | dict(p)


def run_case(name, manifest=None, data=None, want_grads=True, cfg_override=None):
    """-> (out, {param key: grad}, {input name: grad}, state dict, inputs)."""
    if manifest is None:
        manifest, data = load_golden()
    _, kind, cfg = cases.CASE_BY_NAME[name]
    if cfg_override:
        cfg = dict(cfg, **cfg_override)
    sd = state_from_manifest(name, manifest, data)
    inp = cases.make_inputs(name, kind, cfg)
    leaves = {}
    for k, v in sd.items():
        if torch.is_floating_point(v) and not detfill.is_rotary_key(k):
            v = v.clone().requires_grad_(want_grads)
        leaves[k] = v
    fin = {}
    for k, v in inp.items():
        if torch.is_floating_point(v) and k != "attention_mask":
            v = v.clone().requires_grad_(want_grads)
        fin[k] = v
    out = oracle_forward(kind, cfg, _strip(leaves), fin)
    pg, ig = {}, {}
    if want_grads:
        (out * cases.probe_for(name, out)).sum().backward()
        pg = {k: v.grad for k, v in leaves.items() if v.requires_grad and v.grad is not None}
        ig = {k: v.grad for k, v in fin.items() if v.requires_grad and v.grad is not None}
    return out.detach(), pg, ig, sd, inp
