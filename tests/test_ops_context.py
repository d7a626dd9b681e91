"""ops.OpsContext: the per-TrainStep state of the operator wrappers (no process-global mutable state)."""
import threading

import torch

from meant_b200 import ops


def test_context_is_thread_local_and_nests():
    a, b = ops.OpsContext(fused_linear=True), ops.OpsContext()
    assert ops.current() is ops._default
    with ops.use(a):
        assert ops.current() is a and ops.current().fused_linear
        with ops.use(b):
            assert ops.current() is b and not ops.current().fused_linear
        assert ops.current() is a
        seen = []
        t = threading.Thread(target=lambda: seen.append(ops.current()))
        t.start()
        t.join()
        assert seen[0] is ops._default            # another thread does not inherit it: Functions capture it at forward
    assert ops.current() is ops._default


def test_contexts_do_not_share_flags_or_buffers():
    a, b = ops.OpsContext(arena=torch.zeros(256)), ops.OpsContext(arena=torch.zeros(256))
    za = a.zeros(10, torch.device("cpu"))
    assert za.data_ptr() == a.arena.data_ptr() and a.arena_off == 10 and b.arena_off == 0
    zb = b.zeros(10, torch.device("cpu"))
    assert zb.data_ptr() == b.arena.data_ptr()
    a.begin_step()
    assert a.arena_off == 0 and b.arena_off == 10
    # process-wide shims act on the default context only; a context without its own choice follows it
    try:
        ops.use_native_gemm(True)
        assert a.native_gemm and ops._default.native_gemm
        c = ops.OpsContext(native_gemm=False)
        assert not c.native_gemm
        ops.use_fused_linear(True)
        assert ops._default.fused_linear and not a.fused_linear
    finally:
        ops.use_native_gemm(False)
        ops.use_fused_linear(False)
    assert not a.native_gemm and a.native_wgrad


def test_colsum_handoff_is_keyed_by_identity_and_version():
    c = ops.OpsContext()
    dx, s = torch.zeros(4, 8), torch.ones(8)
    c.put_colsum(dx, s)
    assert c.take_colsum(dx, 8, torch.float32) is s
    assert c.take_colsum(dx, 8, torch.float32) is None            # consumed
    c.put_colsum(dx, s)
    dx.add_(1)                                                    # modified after the norm backward: stale
    assert c.take_colsum(dx, 8, torch.float32) is None
    c.put_colsum(dx, s)
    assert c.take_colsum(dx.clone(), 8, torch.float32) is None    # another tensor
    c.put_colsum(dx, s)
    c.begin_step()
    assert c.take_colsum(dx, 8, torch.float32) is None            # cleared at the start of a step


def test_trainstep_owns_its_context():
    from meant_b200.train import TrainStep
    m1, m2 = torch.nn.Linear(8, 2), torch.nn.Linear(8, 2)
    s1, s2 = TrainStep(m1, bf16=False), TrainStep(m2, bf16=False)
    assert s1.ctx is not s2.ctx and s1.ctx is not ops._default
    assert ops.current() is ops._default and not ops._default.fused_linear
