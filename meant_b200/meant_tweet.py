"""meant_tweet: language-only MEANT (text + lag).  Reference: meant/meant_tweet.py:39-167."""
import torch
from torch import nn

from . import ops
from .meant import languageEncoder, _run_stack
from .temporal import temporal


class temporalEncoder(nn.Module):
    """Un-normed variant used by meant_tweet / meant_vision: Linear, temporal, Linear
    (meant/meant_tweet.py:84-110, meant/meant_vision.py:75-105)."""

    def __init__(self, dim, num_heads, lag):
        super().__init__()
        self.dim, self.num_heads, self.n, self.lag = dim, num_heads, 196, lag
        self.temp_embedding = nn.Parameter(torch.randn(1, lag, dim))
        self.temp_encode = nn.ModuleList([nn.Linear(dim, dim), temporal(num_heads, dim), nn.Linear(dim, dim)])

    def forward(self, x):
        x = x + self.temp_embedding.to(x.dtype)
        for mod in self.temp_encode:
            x = mod(x)
        return x


class _LayerNorm(nn.LayerNorm):
    """nn.LayerNorm parameters/state_dict, sm_100a kernel."""

    def forward(self, x):
        return ops.layer_norm(x, self.weight, self.bias, self.eps)


class meant_tweet(nn.Module):
    def __init__(self, text_dim, price_dim, lag, num_classes, embedding, flash=False, num_heads=8, num_encoders=1,
                 channels=4):
        super().__init__()
        self.dim, self.num_heads, self.lag = text_dim, num_heads, lag
        self.embedding = nn.ModuleList([embedding])
        self.languageEncoders = nn.ModuleList([languageEncoder(text_dim, num_heads, flash=flash)
                                               for _ in range(num_encoders)])
        self.temporal_encoding = nn.ModuleList([temporalEncoder(self.dim, num_heads, lag)])
        self.mlpHead = nn.ModuleList([_LayerNorm(self.dim), nn.Linear(self.dim, num_classes), nn.Sigmoid()])

    def forward(self, tweets=None, attention_mask=None, prices=None, **kw):
        tweets = kw.get("input_ids", tweets)
        b = tweets.shape[0]
        words = tweets.reshape(b * self.lag, tweets.shape[2])
        if attention_mask is not None:   # the reference requires it (meant/meant_tweet.py:150)
            attention_mask = attention_mask.reshape(b * self.lag, attention_mask.shape[2])
        for mod in self.embedding:
            words = ops.embed(mod, words)
        words = _run_stack(self.languageEncoders, words, attention_mask)
        t = ops.mean_pool(words).view(b, self.lag, -1)
        for enc in self.temporal_encoding:
            t = enc(t)
        for mod in self.mlpHead:
            t = mod(t)
        return t.squeeze(dim=1).float()
