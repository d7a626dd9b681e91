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


# ---- the general forms (gemm_tc2.cu): every operand major-ness, bias / residual epilogue, weight gradient ----------
GEN_CASES = [  # M, N, K, a_mn, b_mn, bias, residual
    (256, 512, 128, 0, 0, True, False), (256, 512, 128, 0, 1, False, False), (256, 512, 128, 1, 0, False, False),
    (256, 512, 128, 1, 1, False, False), (98304, 768, 768, 0, 1, False, False), (4096, 768, 2304, 0, 1, False, False),
    (4096, 768, 768, 0, 0, True, True), (1000, 320, 200, 0, 1, True, True), (136, 264, 72, 1, 1, True, False),
    (8, 8, 8, 0, 1, False, False), (77, 1000, 200, 0, 0, True, True),
]


def _gen(m, n, k, a_mn, b_mn, bias, res, env=None):
    torch.manual_seed(m + n + k + 2 * a_mn + b_mn)
    a = torch.randn(m, k, device="cuda").bfloat16()
    b = (torch.randn(n, k, device="cuda") * 0.05).bfloat16()
    bi = torch.randn(n, device="cuda") if bias else None
    r = torch.randn(m, n, device="cuda").bfloat16() if res else None
    a_st = a.t().contiguous() if a_mn else a          # stored [K, M] when MN-major
    b_st = b.t().contiguous() if b_mn else b          # stored [K, N] when MN-major
    c = torch.full((m, n), 7.0, device="cuda").bfloat16()
    rc = nv.lib().meant_gemm_bf16(nv.ptr(a_st), nv.ptr(b_st), nv.ptr(c), m, n, k, a_st.stride(0), b_st.stride(0), n, a_mn, b_mn,
                                  nv.ptr(bi), nv.ptr(r), n if res else 0, nv.stream())
    assert rc == 0, rc
    ref = a.float() @ b.float().t() + (bi if bias else 0.0) + (r.float() if res else 0.0)
    err = float((c.float() - ref).abs().max())
    assert err <= 1e-2 * max(1.0, float(ref.abs().max())), err


@pytest.mark.parametrize("m,n,k,a_mn,b_mn,bias,res", GEN_CASES)
def test_gemm_general_forms_match_torch(m, n, k, a_mn, b_mn, bias, res):
    _gen(m, n, k, a_mn, b_mn, bias, res)


WGRAD_CASES = [  # T, N, K, bias gradient
    (98304, 768, 768, True), (98304, 2304, 768, True), (4096, 768, 2304, False), (1000, 320, 200, True), (64, 128, 256, True),
    (72, 8, 8, True), (30000, 1000, 136, True),
]


@pytest.mark.parametrize("t,n,k,want_db", WGRAD_CASES)
def test_gemm_wgrad_matches_torch_and_accumulates(t, n, k, want_db):
    torch.manual_seed(t + n + k)
    dy = (torch.randn(t, n, device="cuda") * 0.1).bfloat16()
    x = torch.randn(t, k, device="cuda").bfloat16()
    dw = torch.full((n, k), 0.5, device="cuda")            # accumulates on top of what the buffers hold
    db = torch.full((n,), -2.0, device="cuda") if want_db else None
    rc = nv.lib().meant_gemm_bf16_wgrad(nv.ptr(dy), nv.ptr(x), nv.ptr(dw), nv.ptr(db), t, n, k, n, k, k, nv.stream())
    assert rc == 0, rc
    ref = dy.double().t() @ x.double()
    scale = max(1.0, float(ref.abs().max()))
    assert float((dw.double() - 0.5 - ref).abs().max()) <= 2e-3 * scale
    if want_db:
        rb = dy.double().sum(0)
        assert float((db.double() + 2.0 - rb).abs().max()) <= 2e-3 * max(1.0, float(rb.abs().max()))


def test_gemm_general_argument_errors():
    x = torch.zeros(16, 16, device="cuda").bfloat16()
    f = torch.zeros(16, 16, device="cuda")
    lib = nv.lib()
    assert lib.meant_gemm_bf16(None, nv.ptr(x), nv.ptr(x), 16, 16, 16, 16, 16, 16, 0, 0, None, None, 0, None) == -1
    assert lib.meant_gemm_bf16(nv.ptr(x), nv.ptr(x), nv.ptr(x), 16, 12, 16, 16, 16, 16, 0, 1, None, None, 0, None) == -2
    assert lib.meant_gemm_bf16_wgrad(nv.ptr(x), nv.ptr(x), None, None, 16, 16, 16, 16, 16, 16, None) == -1
    assert lib.meant_gemm_bf16_wgrad(nv.ptr(x), nv.ptr(x), nv.ptr(f), None, 16, 16, 16, 8, 16, 16, None) == -1   # lddy < N


def test_gemm_cta_pair_kernel_in_a_subprocess():
    """The cta_group::2 kernel (two CTAs per 256 x 256 tile, opt-in through MEANT_GEMM_PAIR, which is read once per
    process): forward and dgrad forms, full and ragged shapes, against torch (tools/gemm_pair_check.py)."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, MEANT_GEMM_PAIR="1", MEANT_DEBUG="1")
    r = subprocess.run([sys.executable, os.path.join(root, "tools", "gemm_pair_check.py")], env=env, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.count(" OK") == 6 and "MISMATCH" not in r.stdout, r.stdout
    assert "resident CTA pairs" in r.stderr, r.stderr[-500:]          # the pair kernel really ran
