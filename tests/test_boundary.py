"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/meant_b200.h declares, argument errors come back as codes (no compute, no GPU), the
module API reproduces the reference's state_dict layout, and there is no CPU fallback."""
import ctypes
import json
import os
import re

import pytest
import torch

import meant_b200
from meant_b200 import _native as nv
from oracle import cases
from parity_util import build_module

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "meant_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(meant_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = nv.lib()
    names = _declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/meant_b200.h but not exported"
    assert set(names) == set(nv.SIGNATURES), set(names) ^ set(nv.SIGNATURES)
    assert lib.meant_version() == nv.ABI_VERSION == 5


def test_argument_errors_are_codes_not_crashes():
    lib = nv.lib()
    assert lib.meant_rmsnorm_fwd(None, None, None, None, None, None, 4, 768, 1e-8, 0, 0.0, 0, None, nv.F32, None) == -1
    assert lib.meant_layernorm_fwd(None, None, None, None, None, None, 4, 768, 1e-5, nv.F32, None) == -1
    assert lib.meant_meanpool_fwd(None, None, 1, 1, 8, nv.F32, None) == -1
    assert lib.meant_attn_fwd(None, nv.F32, 0, None) == -1
    a = nv.AttnArgs()
    assert lib.meant_attn_fwd(ctypes.byref(a), nv.F32, nv.IMPL_EXACT, None) == -1
    assert lib.meant_geglu_fwd(ctypes.c_void_p(16), ctypes.c_void_p(16), 0, 6, nv.BF16, None) == -2  # F % 8


def test_attn_args_struct_layout_matches_header():
    # 8 head views of 40 bytes, 7 pointers, 9 ints, 1 float (+ tail padding to 8)
    assert ctypes.sizeof(nv.HeadView) == 40
    assert ctypes.sizeof(nv.AttnArgs) == 8 * 40 + 7 * 8 + 10 * 4 + 2 * 40 + 2 * 40 + 2 * 8


@pytest.mark.parametrize("name", [c[0] for c in cases.CASES])
def test_state_dict_layout_matches_reference(name, golden):
    manifest, _ = golden
    _, kind, cfg = cases.CASE_BY_NAME[name]
    module, _ = build_module(kind, cfg)
    mine = [(k, list(v.shape)) for k, v in module.state_dict().items()]
    ref = [(k, list(s)) for k, s in manifest[name]["params"]]
    assert mine == ref


def test_reference_package_exports_are_present():
    for n in ("attention", "meant", "languageEncoder", "visionEncoder", "meant_vision", "meant_tweet", "meant_vqa",
              "xPosAttention", "xPosAttention_flash", "flash_attention", "temporal"):   # meant/__init__.py:1-11
        assert hasattr(meant_b200, n)


def test_no_cpu_fallback():
    m = meant_b200.RMSNorm(768)
    with pytest.raises(RuntimeError, match="no CPU path"):
        m(torch.randn(2, 768))
    att = meant_b200.xPosAttention(8, 768, meant_b200.RotaryEmbedding(dim=48, use_xpos=True))
    with pytest.raises(RuntimeError, match="no CPU path"):
        att(torch.randn(1, 16, 768))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "meant_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("no oracle", ""), f"{f} mentions the oracle"


def test_rotary_tables_match_oracle():
    """Host-side table construction (the only rotary math outside the kernels)."""
    from oracle import meant_oracle as mo
    re_ = meant_b200.RotaryEmbedding(dim=48, use_xpos=True)
    qc, qs, kc, ks = re_.tables(128, "cpu")
    ang = mo.rotary_angles(re_.freqs.detach(), 128)
    sc = mo.xpos_scale(re_.scale, 128)
    assert torch.allclose(qc, ang.cos() * sc, atol=1e-6) and torch.allclose(ks, ang.sin() / sc, atol=1e-6)
    x = torch.randn(2, 8, 128, 96)
    want_q, want_k = mo.apply_rotary(x, ang, sc), mo.apply_rotary(x, ang, sc ** -1)
    h = x[..., :48].reshape(2, 8, 128, 24, 2)
    got = torch.stack((h[..., 0] * qc[:, 0::2] - h[..., 1] * qs[:, 0::2],
                       h[..., 1] * qc[:, 1::2] + h[..., 0] * qs[:, 1::2]), -1).reshape(2, 8, 128, 48)
    assert torch.allclose(got, want_q[..., :48], atol=1e-5)
    got_k = torch.stack((h[..., 0] * kc[:, 0::2] - h[..., 1] * ks[:, 0::2],
                         h[..., 1] * kc[:, 1::2] + h[..., 0] * ks[:, 1::2]), -1).reshape(2, 8, 128, 48)
    assert torch.allclose(got_k, want_k[..., :48], atol=1e-5)
