"""xPosAttention / xPosAttention_flash with the reference's constructor and state_dict
(`q`, `k`, `v`, `multi_mad`, shared `xPos`), running the fused sm_100a attention kernel.
Reference: meant/xPosAttention.py:10-67, meant/xPosAttention_flash.py:12-45."""
import math

import torch
import torch.nn.functional as F
from torch import nn

from . import ops
from .rotary_embedding_torch import RotaryEmbedding


def _packed_qkv(mod, x):
    """One N=3*dim projection.  The reference cross-wires the projections (keys come from the
    `v` Linear, values from the `k` Linear: meant/xPosAttention.py:37-38), so the packed weight
    is (q | v | k).  Parameters stay three separate nn.Linear modules (state_dict contract)."""
    return ops.linear3(x, mod.q, mod.v, mod.k)


class xPosAttention(nn.Module):
    def __init__(self, num_heads, dim, xPos: RotaryEmbedding, mask=True, droput=0.):
        super().__init__()
        self.num_heads, self.dim, self.Dh = num_heads, dim, int(dim / num_heads)
        self.dropout = nn.Dropout(droput)
        self.xPos = xPos
        self.mask = mask
        self.softmax = nn.Softmax(dim=-1)
        self.multi_mad = nn.Linear(self.num_heads * self.Dh, self.dim)
        self.q = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.v = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.k = nn.Linear(self.dim, self.Dh * self.num_heads)

    def forward(self, input, attention_mask=None):
        if self.training and self.dropout.p > 0:
            # the reference drops pre-softmax scores (meant/xPosAttention.py:59); no model sets it
            raise NotImplementedError("dropout on attention scores is not on the MEANT hot path (droput=0 everywhere)")
        s = input.shape[1]
        qkv = _packed_qkv(self, input)
        o = ops.attention(qkv, consume=True, heads=self.num_heads, scale=1.0 / math.sqrt(self.Dh * self.num_heads),
                          causal=bool(self.mask), rot=self.xPos.tables(s, input.device), key_mask=attention_mask)
        return ops.linear(o, self.multi_mad.weight, self.multi_mad.bias)


class xPosAttention_flash(nn.Module):
    """The reference's flash variant keeps tensors `b s h d`, which makes the rotary position the
    HEAD index, uses softmax scale Dh^-1/2 and takes no padding mask
    (meant/xPosAttention_flash.py:37-41).  Those semantics are reproduced, not 'fixed'."""

    def __init__(self, num_heads, dim, xPos: RotaryEmbedding, mask=True, droput=0.):
        super().__init__()
        self.num_heads, self.dim, self.Dh = num_heads, dim, int(dim / num_heads)
        self.dropout = nn.Dropout(droput)
        self.xPos = xPos
        self.mask = mask
        self.softmax = nn.Softmax(dim=-1)
        self.multi_mad = nn.Linear(self.num_heads * self.Dh, self.dim)
        self.q = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.k = nn.Linear(self.dim, self.Dh * self.num_heads)
        self.v = nn.Linear(self.dim, self.Dh * self.num_heads)

    def forward(self, input):
        qkv = _packed_qkv(self, input)
        o = ops.attention(qkv, consume=True, heads=self.num_heads, scale=self.Dh ** -0.5, causal=bool(self.mask),
                          rot=self.xPos.tables(self.num_heads, input.device), pos_is_head=True)
        return ops.linear(o, self.multi_mad.weight, self.multi_mad.bias)
