"""The hand-written tcgen05 GEMM (meant_gemm_bf16_tn) against torch on the same bf16 inputs: full tiles, ragged M / N /
K (TMA zero-fill and clipping), strided rows, with and without bias."""
import pytest
import torch

from meant_b200 import _native as nv

pytestmark = pytest.mark.gpu

CASES = [  # M, N, K, bias
    (128, 256, 64, False), (256, 512, 128, True), (98304, 768, 768, True), (4096, 2304, 768, True),
    (1000, 768, 768, True), (130, 264, 72, True), (1, 8, 8, False), (77, 1000, 200, True),
]


@pytest.mark.parametrize("m,n,k,bias", CASES)
def test_gemm_matches_torch(m, n, k, bias):
    torch.manual_seed(m + n + k)
    a = torch.randn(m, k, device="cuda").bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    b = torch.randn(n, device="cuda") if bias else None
    c = torch.full((m, n), 7.0, device="cuda").bfloat16()
    rc = nv.lib().meant_gemm_bf16_tn(nv.ptr(a), nv.ptr(w), nv.ptr(b), nv.ptr(c), m, n, k, k, k, n, nv.stream())
    assert rc == 0
    ref = a.float() @ w.float().t() + (b if bias else 0.0)
    err = float((c.float() - ref).abs().max())
    assert err <= 1e-2 * max(1.0, float(ref.abs().max())), err


def test_gemm_strided_rows_leave_neighbours_untouched():
    torch.manual_seed(0)
    wide_a = torch.randn(300, 1024, device="cuda").bfloat16()
    wide_c = torch.full((300, 1024), 3.0, device="cuda").bfloat16()
    a, c = wide_a[:, 128:128 + 256], wide_c[:, 256:256 + 384]
    w = (torch.randn(384, 256, device="cuda") * 0.05).bfloat16()
    rc = nv.lib().meant_gemm_bf16_tn(nv.ptr(a), nv.ptr(w), None, nv.ptr(c), 300, 384, 256, 1024, 256, 1024, nv.stream())
    assert rc == 0
    ref = a.float() @ w.float().t()
    assert float((c.float() - ref).abs().max()) <= 1e-2 * max(1.0, float(ref.abs().max()))
    assert float((wide_c[:, :256].float() - 3.0).abs().max()) == 0.0 and float((wide_c[:, 640:].float() - 3.0).abs().max()) == 0.0


def test_gemm_argument_errors():
    x = torch.zeros(16, 16, device="cuda").bfloat16()
    lib = nv.lib()
    assert lib.meant_gemm_bf16_tn(None, nv.ptr(x), None, nv.ptr(x), 16, 16, 16, 16, 16, 16, None) == -1
    assert lib.meant_gemm_bf16_tn(nv.ptr(x), nv.ptr(x), None, nv.ptr(x), 16, 12, 16, 16, 16, 16, None) == -2   # N % 8
