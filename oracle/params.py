"""Random-init parameter dicts with the reference's state_dict layout, built WITHOUT the product
package (used by the CPU baseline arm of bench.py).  TEST INFRASTRUCTURE ONLY.
Layout and rotary constants come from the golden manifest (written from the reference's own
constructors by oracle/gen_golden.py); values follow nn.Linear / nn.Embedding default init."""
import re

import torch

from .run_oracle import load_golden


def make_params(case="meant_tweet_l12", encoders=1, vocab=None, seed=0):
    manifest, data = load_golden()
    g = torch.Generator().manual_seed(seed)
    p = {}
    for key, shape in manifest[case]["params"]:
        m = re.match(r"(languageEncoders|visionEncoders)\.0\.(.*)", key)
        keys = [f"{m.group(1)}.{i}.{m.group(2)}" for i in range(encoders)] if m else [key]
        for k in keys:
            rk = f"{case}::rot::{key}"
            if rk in data.files:
                p[k] = torch.from_numpy(data[rk]).clone()
                continue
            shp = list(shape)
            if k == "embedding.0.weight":
                if vocab:
                    shp[0] = vocab
                t = torch.randn(shp, generator=g)
            elif k.endswith(".scale") or (k.endswith(".weight") and len(shp) == 1):
                t = torch.ones(shp)
            elif len(shp) == 2:
                t = (torch.rand(shp, generator=g) * 2 - 1) / shp[1] ** 0.5
            elif k.endswith(".bias"):
                t = (torch.rand(shp, generator=g) * 2 - 1) * 0.03
            else:
                t = torch.randn(shp, generator=g)
            p[k] = t.requires_grad_(True)
    return p
