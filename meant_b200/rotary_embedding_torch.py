"""Rotary / xPos tables.  Keeps the reference module's constructor, `freqs` parameter and
`scale` buffer (both are in the reference state_dict) but never rotates anything itself: it
only builds the per-element cos/sin tables the attention kernels apply in registers.
Reference: meant/rotary_embedding_torch.py:56-147."""
from math import pi

import torch
from torch import nn


class RotaryEmbedding(nn.Module):
    def __init__(self, dim, custom_freqs=None, freqs_for='lang', theta=10000, max_freq=10, num_freqs=1,
                 learned_freq=False, use_xpos=False, xpos_scale_base=512):
        super().__init__()
        if custom_freqs is not None:
            freqs = custom_freqs
        elif freqs_for == 'lang':
            freqs = 1. / (theta ** (torch.arange(0, dim, 2)[:(dim // 2)].float() / dim))
        elif freqs_for == 'pixel':
            freqs = torch.linspace(1., max_freq / 2, dim // 2) * pi
        elif freqs_for == 'constant':
            freqs = torch.ones(num_freqs).float()
        else:
            raise ValueError(f'unknown modality {freqs_for}')
        if learned_freq:
            raise NotImplementedError("learned rotary frequencies are not used by MEANT")
        self.freqs = nn.Parameter(freqs, requires_grad=False)
        self.use_xpos = use_xpos
        self.scale_base = xpos_scale_base
        self.register_buffer('scale', (torch.arange(0, dim, 2) + 0.4 * dim) / (1.4 * dim) if use_xpos else None)
        self._tables = {}

    @torch.no_grad()
    def tables(self, positions: int, device):
        """(qcos, qsin, kcos, ksin), each fp32 [positions, rot_dim].
        angle[p, j] = p * freqs[j // 2] (interleaved, :142); xPos scale[p, j] =
        scale[j mod half] ** ((p - positions//2) / scale_base) (half-split, :123-125);
        q tables carry *scale, k tables /scale (:108-109)."""
        key = (positions, str(device), self.freqs._version, None if self.scale is None else self.scale._version,
               self.freqs.data_ptr())
        hit = self._tables.get(key)
        if hit is not None:
            return hit
        f = self.freqs.detach().to(device=device, dtype=torch.float32)
        pos = torch.arange(positions, device=device, dtype=torch.float32)
        ang = torch.repeat_interleave(pos[:, None] * f[None, :], 2, dim=-1)
        cos, sin = ang.cos(), ang.sin()
        if self.use_xpos:
            base = self.scale.detach().to(device=device, dtype=torch.float32)
            power = (pos - positions // 2) / self.scale_base
            sc = base[None, :] ** power[:, None]
            sc = torch.cat((sc, sc), dim=-1)
            out = (cos * sc, sin * sc, cos / sc, sin / sc)
        else:
            out = (cos, sin, cos, sin)
        out = tuple(t.contiguous() for t in out)
        self._tables = {key: out} if len(self._tables) > 8 else {**self._tables, key: out}
        return out
