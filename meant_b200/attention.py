"""Pixel-rotary attention (vision encoder) and its flash variant.
Reference: meant/attention.py:11-62, meant/flash_attention.py:11-47."""
import math

from torch import nn

from . import ops
from .rotary_embedding_torch import RotaryEmbedding
from .xPosAttention import _packed_qkv


class attention(nn.Module):
    def __init__(self, num_heads, dim, pos_emb: RotaryEmbedding, mask=False, droput=0.):
        super().__init__()
        self.num_heads, self.dim, self.Dh = num_heads, dim, int(dim / num_heads)
        self.dropout = nn.Dropout(droput)
        self.pos_emb = pos_emb
        self.mask = mask
        self.softmax = nn.Softmax(dim=-1)
        self.multi_mad = nn.Linear(self.num_heads * self.Dh, self.dim)
        self.q = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.v = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.k = nn.Linear(self.dim, self.Dh * self.num_heads)

    def forward(self, input):
        if self.mask:
            # mask=True raises UnboundLocalError in the reference (meant/attention.py:46-52)
            raise NotImplementedError("attention(mask=True) is unusable in the reference; only mask=False exists")
        n = input.shape[1]
        qkv = _packed_qkv(self, input)
        o = ops.attention(qkv, consume=True, heads=self.num_heads, scale=1.0 / math.sqrt(self.Dh * self.num_heads),
                          rot=self.pos_emb.tables(n, input.device))
        return self.dropout(ops.linear(o, self.multi_mad.weight, self.multi_mad.bias))


class flash_attention(nn.Module):
    """Head-index pixel rotary, scale 1/dim, non-causal (meant/flash_attention.py:37-43)."""

    def __init__(self, num_heads, dim, pos_emb: RotaryEmbedding, mask=False, droput=0.):
        super().__init__()
        self.num_heads, self.dim, self.Dh = num_heads, dim, int(dim / num_heads)
        self.dropout = nn.Dropout(droput)
        self.pos_emb = pos_emb
        self.mask = mask
        self.softmax = nn.Softmax(dim=-1)
        self.multi_mad = nn.Linear(self.num_heads * self.Dh, self.dim)
        self.q = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.v = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.k = nn.Linear(self.dim, self.Dh * self.num_heads)

    def forward(self, input):
        qkv = _packed_qkv(self, input)
        o = ops.attention(qkv, consume=True, heads=self.num_heads, scale=1.0 / self.dim, causal=bool(self.mask),
                          rot=self.pos_emb.tables(self.num_heads, input.device), pos_is_head=True)
        return ops.linear(o, self.multi_mad.weight, self.multi_mad.bias)
