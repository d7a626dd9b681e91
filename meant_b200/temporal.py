"""Temporal lag attention: the target day (last row) queries the whole antecedent window.
Reference: meant/temporal.py:12-60 (top-level, cross-wired k/v, scale 1/sqrt(dim)) and
src/meant/temporal.py:12-74 (`temporal_src`: un-swapped, scale 1/sqrt(Dh), flattened output)."""
import math

import torch
import torch.nn.functional as F
from torch import nn

from . import ops


class temporal(nn.Module):
    def __init__(self, num_heads, dim, mask=False, droput=0.):
        super().__init__()
        self.num_heads, self.dim, self.Dh = num_heads, dim, int(dim / num_heads)
        self.dropout = nn.Dropout(droput)
        self.mask = mask
        self.softmax = nn.Softmax(dim=-1)
        self.multi_mad = nn.Linear(self.num_heads * self.Dh, self.dim)
        self.atten_size = self.Dh * self.num_heads
        self.q = nn.Linear(self.dim, self.atten_size)
        self.v = nn.Linear(self.dim, self.atten_size)
        self.k = nn.Linear(self.dim, self.atten_size)

    def _kv(self, x):
        # keys from `v`, values from `k` (meant/temporal.py:38-39)
        return F.linear(x, torch.cat((self.v.weight, self.k.weight), 0), torch.cat((self.v.bias, self.k.bias), 0))

    def _scale(self):
        return 1.0 / math.sqrt(self.Dh * self.num_heads)

    def forward(self, input):
        if self.mask:
            raise NotImplementedError("temporal(mask=True) raises in the reference (meant/temporal.py:46-51)")
        q = self.q(input[:, -1:, :])
        o = ops.attention(q=q, kv=self._kv(input), heads=self.num_heads, scale=self._scale())
        return self.multi_mad(o)


class temporal_src(temporal):
    """src/meant/temporal.py: same parameters; K from `k`, V from `v`, scale 1/sqrt(Dh), output
    (B, dim).  `rot_embed` is accepted and ignored: with one query row the xPos table is the
    identity (position 0: angle 0, scale**0)."""

    def __init__(self, num_heads, dim, sequence_length=512, mask=False, droput=0., rot_embed=None):
        super().__init__(num_heads, dim, mask, droput)
        self.Dh = max(self.Dh, 1)
        self.xPos = rot_embed

    def _kv(self, x):
        return F.linear(x, torch.cat((self.k.weight, self.v.weight), 0), torch.cat((self.k.bias, self.v.bias), 0))

    def _scale(self):
        return 1.0 / math.sqrt(self.Dh)

    def forward(self, input, attention_mask=None):
        return super().forward(input).reshape(input.shape[0], -1)
