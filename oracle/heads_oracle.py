"""fp32 CPU restatement of the heads around the encoder stacks.  TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
Citations relative to /root/reference."""
import torch
import torch.nn.functional as F


def softmax_pool(p, prefix, x):
    """src/meant/meant_tweet.py:173 (lang_prep = Sequential(Linear(d,d), LayerNorm(d), GELU(), Linear(d,1),
    Softmax(dim=2))) applied as at :259-261:  w = lang_prep(x);  pooled = matmul(x.transpose(-2,-1), w).squeeze(-1).
    For the 4-D text input [B, lag, S, d] dim 2 is the token axis.  For the 3-D image input [B, T, d]
    (src/meant/meant_timesformer.py:342-344, image_prep) dim 2 is the scorer's output axis of length ONE, so every
    weight is exp(s - s) = 1 and the result is the plain sum over the T tokens -- restated as is."""
    h = F.linear(x, p[prefix + ".0.weight"], p[prefix + ".0.bias"])
    h = F.layer_norm(h, (h.shape[-1],), p[prefix + ".1.weight"], p[prefix + ".1.bias"], 1e-5)
    s = F.linear(F.gelu(h), p[prefix + ".3.weight"], p[prefix + ".3.bias"])
    w = torch.softmax(s, dim=2)
    return torch.matmul(x.transpose(-2, -1), w).squeeze(-1)


def token_cross_entropy(logits, target, ignore_index=-100):
    """pretrain_mlm.py:154,182: nn.CrossEntropyLoss()(out.view(-1, vocab), target.view(-1)); the data collator marks
    un-masked positions with -100, CrossEntropyLoss's default ignore_index."""
    v = logits.shape[-1]
    lg, tg = logits.reshape(-1, v).float(), target.reshape(-1)
    keep = tg != ignore_index
    lse = torch.logsumexp(lg, dim=-1)
    picked = lg.gather(1, tg.clamp_min(0)[:, None]).squeeze(1)
    return ((lse - picked) * keep).sum() / keep.sum()
