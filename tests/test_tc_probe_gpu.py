"""Pins the tcgen05 descriptor conventions (K-major / MN-major SWIZZLE_NONE operands from the
shared chunk layout, instruction descriptor, TMEM lane/column mapping) against torch."""
import pytest
import torch

from meant_b200 import _native as nv

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("a_mn,b_mn", [(0, 0), (0, 1), (1, 0), (1, 1)])
@pytest.mark.parametrize("n,k", [(128, 96), (96, 128), (64, 64), (128, 16), (208, 64), (16, 32)])
def test_probe_gemm(a_mn, b_mn, n, k):
    g = torch.Generator(device="cuda").manual_seed(n * 1000 + k + a_mn * 2 + b_mn)
    a = torch.randn(128, k, device="cuda", generator=g).bfloat16()
    b = torch.randn(n, k, device="cuda", generator=g).bfloat16()
    a_in = a.t().contiguous() if a_mn else a
    b_in = b.t().contiguous() if b_mn else b
    d = torch.full((128, n), float("nan"), device="cuda")
    nv.check(nv.lib().meant_tc_probe(a_in.data_ptr(), b_in.data_ptr(), d.data_ptr(), n, k, a_mn, b_mn, nv.stream()),
             "meant_tc_probe")
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    assert torch.allclose(d, ref, atol=1e-3, rtol=1e-4), float((d - ref).abs().max())


# (n, k, a_mn, b_mn): every operand form the attention kernels issue with swizzled sub-tiles
SW_CASES = [
    (128, 64, 0, 0),    # S = Q K^T, K steps inside one SWIZZLE_128B sub-tile
    (128, 32, 0, 0),    # ... inside the SWIZZLE_64B sub-tile (columns 64..95 of Dh = 96)
    (128, 128, 0, 0),   # two K sub-tiles (P / dS as a K-major A operand)
    (64, 128, 0, 1),    # P V: B = V MN-major, one 64-column sub-tile
    (32, 128, 0, 1),    # ... the 32-column SWIZZLE_64B sub-tile
    (64, 128, 1, 1),    # dV = P^T dO: A MN-major over two sub-tiles (M = 128), B MN-major
    (32, 64, 1, 1),
    (128, 64, 1, 0),
]


@pytest.mark.parametrize("n,k,a_mn,b_mn", SW_CASES)
def test_probe_gemm_swizzled(n, k, a_mn, b_mn):
    g = torch.Generator(device="cuda").manual_seed(n * 1000 + k + a_mn * 2 + b_mn)
    a = torch.randn(128, k, device="cuda", generator=g).bfloat16()
    b = torch.randn(n, k, device="cuda", generator=g).bfloat16()
    a_in = a.t().contiguous() if a_mn else a
    b_in = b.t().contiguous() if b_mn else b
    d = torch.full((128, n), float("nan"), device="cuda")
    nv.check(nv.lib().meant_tc_probe_sw(a_in.data_ptr(), b_in.data_ptr(), d.data_ptr(), n, k, a_mn, b_mn, nv.stream()),
             "meant_tc_probe_sw")
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t()
    assert torch.allclose(d, ref, atol=1e-3, rtol=1e-4), float((d - ref).abs().max())
