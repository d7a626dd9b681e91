"""meant_attn_q1_fwd_static: the single-query attention forward with its scratch in a file-scope device array
(include/meant_b200.h; no counterpart in the reference -- added for callers that cannot call cudaMalloc).

Collected last on purpose.  Called through the C ABI and compared with the existing
entry point `meant_attn_q1_fwd` (caller-owned scratch) on the same inputs, twice in a row: the array keeps the previous
call's contents, so the second call -- a different, smaller problem, then the first one again -- must not see them.
Both entry points run the same two kernels on the same records, so the comparison is bit-exact."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_FLOATS = 1 << 20          # MEANT_Q1_STATIC_FLOATS
ETOOBIG = -3                     # MEANT_ETOOBIG


def _problem(b, tokens, heads, dh, dtype, seed):
    from meant_b200 import _native as nv
    from meant_b200._native import AttnArgs, HeadView
    gen = torch.Generator(device="cuda").manual_seed(seed)
    inner = heads * dh
    qkv = torch.randn(b, tokens, 3 * inner, device="cuda", generator=gen).to(dtype)
    es = qkv.element_size()

    def args(out, lse):
        a = AttnArgs()
        for name, col in (("q", 0), ("k", inner), ("v", 2 * inner)):
            setattr(a, name, HeadView(qkv.data_ptr() + col * es, qkv.stride(0), 0, qkv.stride(1), dh))
        a.o = HeadView(out.data_ptr(), out.stride(0), 0, out.stride(1), dh)
        a.lse = nv.ptr(lse)
        a.N, a.n2_count, a.H, a.Sq, a.Sk, a.Dh = b, 1, heads, 1, tokens, dh
        a.scale = dh ** -0.5
        return a

    return qkv, args


def _run_pair(b, tokens, heads, dh, dtype, seed):
    from meant_b200 import _native as nv
    qkv, args = _problem(b, tokens, heads, dh, dtype, seed)
    inner = heads * dh
    nchunk = (tokens + 63) // 64
    outs = []
    for static in (False, True):
        out = torch.zeros(b, 1, inner, dtype=dtype, device="cuda")
        lse = torch.zeros(b, heads, dtype=torch.float32, device="cuda")
        a = args(out, lse)
        if static:
            rc = nv.lib().meant_attn_q1_fwd_static(C.byref(a), nv.dt(qkv), nv.stream())
        else:
            partial = torch.empty(b, heads, nchunk, dh + 2, dtype=torch.float32, device="cuda")
            rc = nv.lib().meant_attn_q1_fwd(C.byref(a), nv.ptr(partial), nv.dt(qkv), nv.stream())
        assert rc == 0, rc
        torch.cuda.synchronize()
        outs.append((out, lse))
    (o0, l0), (o1, l1) = outs
    assert torch.isfinite(o0.float()).all() and o0.float().abs().sum() > 0
    assert torch.equal(o0, o1), (o0.float() - o1.float()).abs().max()
    assert torch.equal(l0, l1), (l0 - l1).abs().max()
    # and against plain torch (the tolerance tests/test_divided_gpu.py uses for the cls query)
    q, k, v = (t.float().view(b, tokens, heads, dh).transpose(1, 2) for t in qkv.split(inner, dim=-1))
    ref = torch.softmax(q[:, :, :1] @ k.transpose(-1, -2) * dh ** -0.5, dim=-1) @ v
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-4
    assert torch.allclose(o1.float().view(b, heads, dh), ref[:, :, 0], atol=tol, rtol=tol)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_static_scratch_matches_caller_scratch_twice(dtype):
    # the TimeSformer cls query of the bench config (1 + 8*196 keys, 12 heads of 64), then a smaller ragged problem
    # that leaves stale records of the first call behind it in the array, then the first problem again
    _run_pair(2, 1 + 8 * 196, 12, 64, dtype, seed=1)
    _run_pair(3, 77, 4, 32, dtype, seed=2)
    _run_pair(2, 1 + 8 * 196, 12, 64, dtype, seed=1)


def test_static_scratch_rejects_oversize():
    from meant_b200 import _native as nv
    b, tokens, heads, dh = 32, 4096, 8, 64                    # 32 * 8 * 64 * 66 floats > 2^20
    assert b * heads * ((tokens + 63) // 64) * (dh + 2) > STATIC_FLOATS
    qkv, args = _problem(b, tokens, heads, dh, torch.bfloat16, seed=3)
    out = torch.full((b, 1, heads * dh), 7.0, dtype=torch.bfloat16, device="cuda")
    lse = torch.zeros(b, heads, dtype=torch.float32, device="cuda")
    a = args(out, lse)
    assert nv.lib().meant_attn_q1_fwd_static(C.byref(a), nv.dt(qkv), nv.stream()) == ETOOBIG
    torch.cuda.synchronize()
    assert (out == 7.0).all()                                 # nothing was launched
