"""Encoder blocks and model assemblies of the oracle.  TEST INFRASTRUCTURE ONLY.

Dropout layers of the reference (`encode.4` p=dropout, `encode2.4` p=0.5,
meant/meant.py:105,107) are identities here: the oracle restates `.eval()` behaviour,
which is the mode every parity test runs in.
"""
import torch
import torch.nn.functional as F

from .meant_oracle import (lin, rmsnorm, xpos_attention, xpos_attention_flash,
                           pixel_attention, flash_attention, temporal_attention)


def language_encoder(p, pre, x, heads, attention_mask=None, flash=False):
    """languageEncoder.forward, meant/meant.py:109-120 (module list at :101-107)."""
    h = rmsnorm(x, p[pre + ".encode.0.scale"])
    h = lin(p, pre + ".encode.1", h)
    if flash:
        h = xpos_attention_flash(p, pre + ".encode.2", h, heads)
    else:
        h = xpos_attention(p, pre + ".encode.2", h, heads, attention_mask)
    h = rmsnorm(h, p[pre + ".encode.3.scale"])
    h = lin(p, pre + ".encode.5", h)
    r = h + x
    g = rmsnorm(r, p[pre + ".encode2.0.scale"])
    g = F.gelu(lin(p, pre + ".encode2.1", g))
    g = rmsnorm(g, p[pre + ".encode2.3.scale"])
    g = lin(p, pre + ".encode2.5", g)
    return g + r


def vision_encoder(p, pre, x, heads, flash=False):
    """visionEncoder.forward, meant/meant.py:66-75 (module list at :59-64)."""
    h = rmsnorm(x, p[pre + ".encode.0.scale"])
    h = lin(p, pre + ".encode.1", h)
    h = (flash_attention if flash else pixel_attention)(p, pre + ".encode.2", h, heads)
    h = rmsnorm(h, p[pre + ".encode.3.scale"])
    h = lin(p, pre + ".encode.4", h)
    r = h + x
    g = rmsnorm(r, p[pre + ".encode2.0.scale"])
    g = F.gelu(lin(p, pre + ".encode2.1", g))
    g = rmsnorm(g, p[pre + ".encode2.3.scale"])
    g = lin(p, pre + ".encode2.4", g)
    return g + r


def temporal_encoder(p, pre, x, heads, normed=True):
    """temporalEncoder.forward.  normed=True: meant/meant.py:138-145 (RMSNorm, Linear,
    temporal, RMSNorm, Linear).  normed=False: meant/meant_vision.py:89-105 and
    meant/meant_tweet.py:84-110 (Linear, temporal, Linear).  The reference adds the
    embedding in place on its input (`x += ...`, meant/meant.py:142); the oracle is
    functional, the build mirrors the in-place write."""
    x = x + p[pre + ".temp_embedding"]
    if normed:
        x = rmsnorm(x, p[pre + ".temp_encode.0.scale"])
        x = lin(p, pre + ".temp_encode.1", x)
        x = temporal_attention(p, pre + ".temp_encode.2", x, heads)
        x = rmsnorm(x, p[pre + ".temp_encode.3.scale"])
        return lin(p, pre + ".temp_encode.4", x)
    x = lin(p, pre + ".temp_encode.0", x)
    x = temporal_attention(p, pre + ".temp_encode.1", x, heads)
    return lin(p, pre + ".temp_encode.2", x)


def patch_embed(p, pre, images, patch):
    """Rearrange 'b c (h p1) (w p2) -> b (h w) (p1 p2 c)' + Linear, meant/meant.py:193-195."""
    b, c, hh, ww = images.shape
    x = images.view(b, c, hh // patch, patch, ww // patch, patch)
    x = x.permute(0, 2, 4, 3, 5, 1).reshape(b, (hh // patch) * (ww // patch), patch * patch * c)
    return lin(p, pre + ".1", x)


def _n_layers(p, stack):
    return 1 + max(int(k.split(".")[1]) for k in p if k.startswith(stack + "."))


def meant_forward(p, tweets, images, attention_mask=None, heads=8, patch=16, flash=False):
    """meant.forward(tweets, images, attention_mask), meant/meant.py:206-237."""
    b, lag, s = tweets.shape
    words = F.embedding(tweets.reshape(b * lag, s), p["embedding.0.weight"])
    if attention_mask is not None:
        attention_mask = attention_mask.reshape(b * lag, s)
    for i in range(_n_layers(p, "languageEncoders")):
        words = language_encoder(p, f"languageEncoders.{i}", words, heads, attention_mask, flash)
    img = images.reshape(b * lag, *images.shape[2:])
    img = patch_embed(p, "patchEmbed", img, patch)
    for i in range(_n_layers(p, "visionEncoders")):
        img = vision_encoder(p, f"visionEncoders.{i}", img, heads, flash)
    words = words.view(b, lag, s, -1).mean(dim=2)
    img = img.view(b, lag, img.shape[1], -1).mean(dim=2)
    t = temporal_encoder(p, "temporal_encoding.0", torch.cat((words, img), dim=2), heads)
    t = rmsnorm(t, p["mlpHead.0.scale"])
    return torch.sigmoid(lin(p, "mlpHead.1", t)).squeeze(1)


def meant_tweet_forward(p, tweets, attention_mask, heads=8, flash=False):
    """meant_tweet.forward, meant/meant_tweet.py:147-167 (its languageEncoder returns the
    misspelt `final_resi`, :81 -- NameError; the oracle uses the textually identical block
    from meant/meant.py).  Head is LayerNorm/Linear/Sigmoid (:143)."""
    b, lag, s = tweets.shape
    words = F.embedding(tweets.reshape(b * lag, s), p["embedding.0.weight"])
    attention_mask = attention_mask.reshape(b * lag, s)
    for i in range(_n_layers(p, "languageEncoders")):
        words = language_encoder(p, f"languageEncoders.{i}", words, heads, attention_mask, flash)
    t = words.view(b, lag, s, -1).mean(dim=2)
    t = temporal_encoder(p, "temporal_encoding.0", t, heads, normed=False)
    t = F.layer_norm(t, t.shape[-1:], p["mlpHead.0.weight"], p["mlpHead.0.bias"])
    return torch.sigmoid(lin(p, "mlpHead.1", t)).squeeze(1)


def meant_vision_forward(p, images, heads=8, patch=16, flash=False):
    """meant_vision.forward, meant/meant_vision.py:152-165."""
    b, lag = images.shape[:2]
    img = patch_embed(p, "patchEmbed", images.reshape(b * lag, *images.shape[2:]), patch)
    for i in range(_n_layers(p, "visionEncoders")):
        img = vision_encoder(p, f"visionEncoders.{i}", img, heads, flash)
    t = img.view(b, lag, img.shape[1], -1).mean(dim=2)
    t = temporal_encoder(p, "temporal_encoding.0", t, heads, normed=False)
    t = F.layer_norm(t, t.shape[-1:], p["mlpHead.0.weight"], p["mlpHead.0.bias"])
    return torch.sigmoid(lin(p, "mlpHead.1", t)).squeeze(1)


def meant_vqa_forward(p, tweets, images, attention_mask=None, heads=8, patch=16, flash=False):
    """meant_vqa.forward, meant/meant_vqa.py:205-233 (no lag axis, no temporal encoder)."""
    words = F.embedding(tweets, p["embedding.0.weight"])
    for i in range(_n_layers(p, "languageEncoders")):
        words = language_encoder(p, f"languageEncoders.{i}", words, heads, attention_mask, flash)
    img = patch_embed(p, "patchEmbed", images, patch)
    for i in range(_n_layers(p, "visionEncoders")):
        img = vision_encoder(p, f"visionEncoders.{i}", img, heads, flash)
    t = torch.cat((words.mean(dim=1), img.mean(dim=1)), dim=1)
    t = rmsnorm(t, p["mlpHead.0.scale"])
    return torch.sigmoid(lin(p, "mlpHead.1", t))


def language_pretrainer_forward(p, words_ids, attention_mask=None, heads=8):
    """meant_language_pretrainer.forward, pretrain_mlm.py:83-88 /
    meant/hf_wrapper.py:111-149: embedding -> languageEncoder x N -> `mlm_head` (an nn.Linear here)."""
    x = F.embedding(words_ids, p["embedding.0.weight"])
    for i in range(_n_layers(p, "languageEncoders")):
        x = language_encoder(p, f"languageEncoders.{i}", x, heads, attention_mask)
    return lin(p, "mlm_head", x)
