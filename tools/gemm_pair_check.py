import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from meant_b200 import _native as nv
torch.manual_seed(0)
L = nv.lib()
def run(M, N, K, b_mn, bias):
    a = torch.randn(M, K, device="cuda").bfloat16()
    b = (torch.randn(N, K, device="cuda") * 0.05).bfloat16()
    bi = torch.randn(N, device="cuda") if bias else None
    b_st = b.t().contiguous() if b_mn else b
    c = torch.full((M, N), 7.0, device="cuda").bfloat16()
    rc = L.meant_gemm_bf16(nv.ptr(a), nv.ptr(b_st), nv.ptr(c), M, N, K, K, b_st.stride(0), N, 0, b_mn, nv.ptr(bi), None, 0, nv.stream())
    torch.cuda.synchronize()
    ref = a.float() @ b.float().t() + (bi if bias else 0.0)
    err = float((c.float() - ref).abs().max()); sc = float(ref.abs().max())
    print(f"M={M} N={N} K={K} b_mn={b_mn} bias={bias}: rc={rc} err={err:.4f} scale={sc:.2f}", "OK" if err <= 1e-2 * max(1, sc) else "MISMATCH", flush=True)
for cfg in ((98304, 768, 768, 0, True), (98304, 768, 768, 1, False), (98304, 2304, 768, 0, True), (98304, 768, 2304, 1, False), (20000, 1000, 200, 0, True), (19000, 776, 136, 1, False)):
    run(*cfg)
