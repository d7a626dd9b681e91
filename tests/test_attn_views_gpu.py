"""Multi-tile attention through the C ABI on GROUPED head views (n2_count > 1, strided rows), without the shared extra
key: the '(b n) f d' grouping of src/meant/timesformer_pytorch.py:122 addressed in place in a [b, f, n, 3*dim] buffer
-- group (b, patch) owns the rows over f, f * n * W apart per b, W apart per patch, n * W apart per row.  Checked
against the same problem on a rearranged, contiguous [(b n), f, 3*dim] copy (n2_count = 1), which the other attention
tests pin to the oracle.  Covers the (n1, n2) decomposition of the sequence index in the wide-key / multi-step kernels'
tensor maps, forward and backward."""
import pytest
import torch

from meant_b200 import ops, _native as nv
from meant_b200._native import AttnArgs, HeadView

pytestmark = pytest.mark.gpu


def _grouped(t, col, f, n, dh):
    """view of columns [col, col + heads*dh) of t[b, f, n, W]: sequences (b, patch), rows over f"""
    w = t.stride(2)
    return HeadView(t.data_ptr() + col * t.element_size(), t.stride(0), w, n * w, dh)


@pytest.mark.parametrize("f,n", [(150, 3), (196, 2), (300, 2)])      # wide-key kernels, their 196-row case, multi-step
def test_grouped_views_match_contiguous(f, n):
    torch.manual_seed(f)
    b, heads, dh = 2, 8, 96
    dim = heads * dh
    scale = dh ** -0.5
    qkv = torch.randn(b, f, n, 3 * dim, device="cuda").bfloat16()
    dout = torch.randn(b, f, n, dim, device="cuda").bfloat16()
    # reference: contiguous copy, one sequence per (b, patch)
    ref_in = qkv.permute(0, 2, 1, 3).reshape(b * n, f, 3 * dim).contiguous().requires_grad_(True)
    ref_out = ops.attention(ref_in, heads=heads, scale=scale, impl=nv.IMPL_TCGEN05)
    ref_out.backward(dout.permute(0, 2, 1, 3).reshape(b * n, f, dim))
    # in place through the C ABI
    out = torch.zeros(b, f, n, dim, device="cuda", dtype=torch.bfloat16)
    dqkv = torch.zeros_like(qkv)
    lse = torch.empty(b * n, heads, f, dtype=torch.float32, device="cuda")
    delta = torch.empty_like(lse)
    a = AttnArgs()
    a.q, a.k, a.v = (_grouped(qkv, c, f, n, dh) for c in (0, dim, 2 * dim))
    a.o = _grouped(out, 0, f, n, dh)
    a.lse = nv.ptr(lse)
    a.N, a.n2_count, a.H, a.Sq, a.Sk, a.Dh = b * n, n, heads, f, f, dh
    a.scale = scale
    nv.check(nv.lib().meant_attn_fwd(a, nv.dt(qkv), nv.IMPL_TCGEN05, nv.stream()), "meant_attn_fwd")
    a.dq, a.dk, a.dv = (_grouped(dqkv, c, f, n, dh) for c in (0, dim, 2 * dim))
    a.o = _grouped(dout, 0, f, n, dh)
    a.out_saved = _grouped(out, 0, f, n, dh)
    a.delta = nv.ptr(delta)
    nv.check(nv.lib().meant_attn_bwd(a, nv.dt(qkv), nv.IMPL_TCGEN05, nv.stream()), "meant_attn_bwd")
    torch.cuda.synchronize()
    got_out = out.permute(0, 2, 1, 3).reshape(b * n, f, dim)
    got_grad = dqkv.permute(0, 2, 1, 3).reshape(b * n, f, 3 * dim)
    assert torch.equal(got_out, ref_out.detach())                  # same kernels, same operands: bit-identical
    assert torch.equal(got_grad, ref_in.grad)
