"""Deterministic, RNG-free tensor filling shared by the golden generator, the oracle
tests and the GPU parity tests.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  A splitmix64 hash of the element
index and a per-tensor tag gives values that are bit-identical on every machine and
torch version, so fixtures only need to store *outputs*: weights and inputs are
re-created from their names.
"""
import zlib

import numpy as np
import torch

_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def _splitmix64(z):
    z = (z + np.uint64(0x9E3779B97F4A7C15)) & _M64
    z = ((z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)) & _M64
    z = ((z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)) & _M64
    return z ^ (z >> np.uint64(31))


def tag_of(name: str) -> int:
    return zlib.crc32(name.encode()) & 0xFFFFFFFF


def uniform(shape, name: str, scale: float = 1.0, shift: float = 0.0) -> torch.Tensor:
    """U(-scale, scale) + shift, fp32, hashed from (name, flat index)."""
    n = int(np.prod(shape)) if len(shape) else 1
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64) + (np.uint64(tag_of(name)) << np.uint64(32))
        h = _splitmix64(idx)
    u = (h >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))  # [0,1)
    v = ((u * 2.0 - 1.0) * scale + shift).astype(np.float32)
    return torch.from_numpy(v).reshape(tuple(shape))


def integers(shape, name: str, high: int) -> torch.Tensor:
    n = int(np.prod(shape))
    with np.errstate(over="ignore"):
        idx = np.arange(n, dtype=np.uint64) + (np.uint64(tag_of(name)) << np.uint64(32))
        h = _splitmix64(idx)
    return torch.from_numpy((h % np.uint64(high)).astype(np.int64)).reshape(tuple(shape))


_ROTARY_MARKERS = ("xPos.", "posEmbed.", "pos_emb.", "rot_emb.", "inv_freqs", ".scales")


def is_rotary_key(key: str) -> bool:
    """Rotary tables (`freqs` parameter / `scale` buffer) keep their constructor values."""
    return any(m in key for m in _ROTARY_MARKERS)


def fill_state_dict(sd: dict) -> dict:
    """Return a new state dict with every learnable tensor replaced by hashed values.

    Rules (by key name and rank): 2-D Linear weights U(+-1/sqrt(fan_in)); embedding
    tables (keys containing 'embedding.0') U(+-1); norm gains 1 + U(+-0.1);
    biases U(+-0.05); everything else (tokens, temporal embeddings) U(+-0.5).
    """
    out = {}
    for k, v in sd.items():
        if is_rotary_key(k) or not torch.is_floating_point(v):
            out[k] = v.clone()
            continue
        shape = tuple(v.shape)
        if k.endswith(".scale") or (k.endswith(".weight") and v.dim() == 1):
            t = uniform(shape, k, 0.1, 1.0)
        elif "embedding.0" in k and v.dim() == 2:
            t = uniform(shape, k, 1.0)
        elif v.dim() == 2:
            t = uniform(shape, k, 1.0 / float(shape[1]) ** 0.5)
        elif k.endswith(".bias"):
            t = uniform(shape, k, 0.05)
        else:
            t = uniform(shape, k, 0.5)
        out[k] = t.to(v.dtype)
    return out


def padding_mask(n_rows: int, seq: int, name: str) -> torch.Tensor:
    """Ones with a per-row zero suffix of hashed length in [0, seq/4]."""
    cut = integers((n_rows,), name, max(seq // 4, 1) + 1)
    ar = torch.arange(seq).unsqueeze(0)
    return (ar < (seq - cut).unsqueeze(1)).to(torch.float32)
