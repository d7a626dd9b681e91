"""RMSNorm module with the reference's constructor and state_dict (`scale`), backed by the
fused sm_100a kernel.  Reference: utils/rms_norm.py:16-57."""
import torch
from torch import nn

from . import ops


class RMSNorm(nn.Module):
    def __init__(self, d, p=-1., eps=1e-8, bias=False):
        super().__init__()
        if 0. <= p <= 1. or bias:
            # every reference call site is RMSNorm(dim)
            raise NotImplementedError("partial RMSNorm (p in [0,1]) and bias=True are not on the MEANT hot path")
        self.eps, self.d, self.p, self.bias = eps, d, p, bias
        self.scale = nn.Parameter(torch.ones(d))

    def forward(self, x):
        return ops.rms_norm(x, self.scale, self.eps)
