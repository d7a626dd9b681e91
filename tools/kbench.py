#!/usr/bin/env python
"""Kernel micro-benchmarks through the C ABI (CUDA events on the launching stream, inputs larger
than L2 or an L2 flush between iterations).  Used for the ncu captures under profiles/.

    python tools/kbench.py attn  [--n 768] [--s 128] [--dh 96] [--iters 20]
    python tools/kbench.py norm  [--rows 98304] [--d 768]
    python tools/kbench.py gemm  [--rows 98304] [--d 768] [--nout 2304]   # hand-written tcgen05 GEMM vs the library
"""
import argparse
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from meant_b200 import ops, RotaryEmbedding, _native as nv  # noqa: E402


def timeit(fn, iters, flush=None):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        if flush is not None:
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms.sort()
    return ms[len(ms) // 2], ms[0]


def prof(tag, items, launches):
    """Phase timers of CTA 0 (library built with MEANT_PROF=1): cycles per item per phase."""
    import ctypes
    fn = getattr(nv.lib(), "meant_debug_prof", None)
    if fn is None:
        return
    buf = (ctypes.c_ulonglong * 64)()
    fn.restype = ctypes.c_int
    if fn(buf, 1) != 1:
        return
    per = launches * ((items + 147) // 148)
    print(f"  prof {tag}: compute", [round(buf[i] / per) for i in range(16)], "loader", [round(buf[16 + i] / per) for i in range(16)],
          "mma", [round(buf[32 + i] / per) for i in range(8)])
    fn2 = getattr(nv.lib(), "meant_debug_prof_mt", None)   # multi-tile forward: cycles per phase, CTA (0,0,0), per launch
    if fn2 is not None:
        b2 = (ctypes.c_ulonglong * 32)()
        fn2.restype = ctypes.c_int
        if fn2(b2, 1) == 1 and any(b2):
            print(f"  prof_mt {tag}: cycles/launch fwd", [round(b2[i] / launches) for i in range(12)],
                  "bwd dK|dV pass", [round(b2[i] / launches) for i in range(12, 22)],
                  "bwd dQ pass", [round(b2[i] / launches) for i in range(22, 32)])


def attn(args):
    heads, dh, n, s = args.heads, args.dh, args.n, args.s
    dim = heads * dh
    torch.manual_seed(0)
    qkv = torch.randn(n, s, 3 * dim, device="cuda").bfloat16().requires_grad_(True)
    rot = RotaryEmbedding(dim=48 if dh == 96 else dh, use_xpos=dh == 96).tables(s, "cuda") if not args.norot else None
    cut = torch.randint(0, s // 4 + 1, (n, 1), device="cuda")
    mask = (torch.arange(s, device="cuda")[None] < (s - cut)).float() if args.causal else None
    scale = 1.0 / math.sqrt(dim)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    impl = {"auto": nv.IMPL_AUTO, "exact": nv.IMPL_EXACT, "tcgen05": nv.IMPL_TCGEN05}[args.impl]

    def fwd():
        return ops.attention(qkv, heads=heads, scale=scale, causal=args.causal, rot=rot, key_mask=mask, impl=impl)

    out = fwd()
    dout = torch.randn_like(out)
    prof("warm", 1, 1)
    f_med, f_min = timeit(lambda: ops.attention(qkv.detach(), heads=heads, scale=scale, causal=args.causal, rot=rot,
                                                key_mask=mask, impl=impl), args.iters, flush)
    prof("fwd", n * heads, 3 + args.iters)

    def bwd():
        qkv.grad = None
        out.backward(dout, retain_graph=True)

    b_med, b_min = timeit(bwd, args.iters, flush)
    prof("bwd", n * heads, 3 + args.iters)
    if args.kernels:                       # per-kernel device times of one forward + backward (torch.profiler)
        from torch.profiler import profile, ProfilerActivity
        with profile(activities=[ProfilerActivity.CUDA]) as pr:
            for _ in range(5):
                flush.zero_()
                out2 = fwd()
                out2.backward(dout)
            torch.cuda.synchronize()
        agg = {}
        for ev in pr.events():
            if ev.device_type == torch.autograd.DeviceType.CUDA:
                k = ev.name[:90]
                a_ = agg.setdefault(k, [0, 0.0])
                a_[0] += 1
                a_[1] += ev.device_time
        for k, (c, tot) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:12]:
            print(f"    {tot / c:9.1f} us x{c // 5}/iter  {k}")
    # algorithmic work: fwd reads q,k,v + writes o; bwd reads q,k,v,dO (+lse) + writes dq,dk,dv
    tile = n * s * dim * 2
    fl = (2.0 if args.causal else 4.0) * s * s * dh * heads * n
    print(f"attn n={n} s={s} dh={dh} causal={args.causal}: fwd {f_med*1e3:.1f} us (min {f_min*1e3:.1f}) "
          f"{4*tile/f_med/1e6:.0f} GB/s {fl/f_med/1e9:.1f} TFLOP/s | bwd(autograd, incl. torch glue) {b_med*1e3:.1f} us "
          f"(min {b_min*1e3:.1f}) {7*tile/b_med/1e6:.0f} GB/s {2*fl/b_med/1e9:.1f} TFLOP/s")


def norm(args):
    rows, d = args.rows, args.d
    if args.colsum:      # the backward also sums dx over rows (the form a TrainStep(graph=True) step runs)
        ops.current().fused_linear = True
    x = torch.randn(rows, d, device="cuda").bfloat16().requires_grad_(True)
    res = torch.randn(rows, d, device="cuda").bfloat16()
    g = torch.ones(d, device="cuda", requires_grad=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    byts = rows * d * 2
    for name, kw, tf, tb in (("plain", {}, 2, 3), ("gelu", {"pre_gelu": True}, 2, 3), ("gelu+drop", {"pre_gelu": True, "drop_p": 0.5}, 2, 3),
                             ("plain+drop", {"drop_p": 0.5}, 2, 3), ("res", {"residual": res}, 4, 4)):
        if args.only and name != args.only:
            continue
        f_med, f_min = timeit(lambda: ops.rms_norm(x.detach(), g.detach(), **kw), args.iters, flush)
        y = ops.rms_norm(x, g, **kw)
        outs = y if isinstance(y, tuple) else (y,)
        douts = tuple(torch.randn_like(o) for o in outs)

        def bwd():
            x.grad = None
            g.grad = None
            torch.autograd.backward(outs, douts, retain_graph=True)

        b_med, b_min = timeit(bwd, args.iters, flush)
        print(f"rmsnorm {name} rows={rows} d={d}: fwd {f_med*1e3:.1f} us {tf*byts/f_med/1e6:.0f} GB/s | "
              f"bwd(autograd) {b_med*1e3:.1f} us {tb*byts/b_med/1e6:.0f} GB/s")


def gemm(args):
    import torch.nn.functional as F
    m, k, n = args.rows, args.d, args.nout
    x = torch.randn(m, k, device="cuda").bfloat16()
    w = (torch.randn(n, k, device="cuda") * 0.02).bfloat16()
    b = torch.randn(n, device="cuda")
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def native():
        rc = nv.lib().meant_gemm_bf16_tn(nv.ptr(x), nv.ptr(w), nv.ptr(b), nv.ptr(c), m, n, k, k, k, n, nv.stream())
        assert rc == 0, rc

    bb = b.bfloat16()
    t_nat, _ = timeit(native, args.iters, flush)
    t_lib, _ = timeit(lambda: F.linear(x, w, bb), args.iters, flush)
    err = (c.float() - F.linear(x, w, bb).float()).abs().max().item()
    fl = 2.0 * m * n * k
    print(f"gemm M={m} N={n} K={k} bf16: tcgen05 {t_nat*1e3:.1f} us {fl/t_nat/1e9:.0f} TFLOP/s | library {t_lib*1e3:.1f} us "
          f"{fl/t_lib/1e9:.0f} TFLOP/s | max|diff| {err:.3g}")
    # the backward forms (gemm_tc2.cu): launch order after the forward pair is [dgrad x (3 + iters)], [wgrad x (3 + iters)]
    dy = (torch.randn(m, n, device="cuda") * 0.1).bfloat16()
    dx = torch.empty(m, k, device="cuda", dtype=torch.bfloat16)
    dw, db = torch.zeros(n, k, device="cuda"), torch.zeros(n, device="cuda")

    def dgrad():
        rc = nv.lib().meant_gemm_bf16(nv.ptr(dy), nv.ptr(w), nv.ptr(dx), m, k, n, n, k, k, 0, 1, None, None, 0, nv.stream())
        assert rc == 0, rc

    def wgrad():
        rc = nv.lib().meant_gemm_bf16_wgrad(nv.ptr(dy), nv.ptr(x), nv.ptr(dw), nv.ptr(db), m, n, k, n, k, k, nv.stream())
        assert rc == 0, rc

    t_dg, _ = timeit(dgrad, args.iters, flush)
    t_wg, _ = timeit(wgrad, args.iters, flush)
    t_ldg, _ = timeit(lambda: torch.mm(dy, w), args.iters, flush)
    t_lwg, _ = timeit(lambda: torch.mm(dy.t(), x, out_dtype=torch.float32), args.iters, flush)
    print(f"  dgrad: tcgen05 {t_dg*1e3:.1f} us {fl/t_dg/1e9:.0f} TFLOP/s | library {t_ldg*1e3:.1f} us;  wgrad + bias grad: tcgen05 "
          f"{t_wg*1e3:.1f} us {fl/t_wg/1e9:.0f} TFLOP/s | library wgrad alone {t_lwg*1e3:.1f} us")


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["attn", "norm", "gemm"])
    ap.add_argument("--n", type=int, default=768)
    ap.add_argument("--s", type=int, default=128)
    ap.add_argument("--dh", type=int, default=96)
    ap.add_argument("--heads", type=int, default=8)
    ap.add_argument("--rows", type=int, default=98304)
    ap.add_argument("--d", type=int, default=768)
    ap.add_argument("--nout", type=int, default=2304)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--causal", type=int, default=1)
    ap.add_argument("--norot", action="store_true")
    ap.add_argument("--colsum", action="store_true", help="norm: backward with the fused column sum of dx")
    ap.add_argument("--impl", default="auto")
    ap.add_argument("--only", default="", help="norm: run one variant only (plain, gelu, gelu+drop, plain+drop, res)")
    ap.add_argument("--kernels", action="store_true", help="attn: print per-kernel device times (torch.profiler)")
    ap.add_argument("--prerotate", type=int, default=1, help="multi-tile attention: rotate Q, K once into scratch (0: in the kernels)")
    a = ap.parse_args()
    a.causal = bool(a.causal)
    ops.PREROTATE = bool(a.prerotate)
    {"attn": attn, "norm": norm, "gemm": gemm}[a.what](a)
