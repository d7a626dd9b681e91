import torch, torch.nn.functional as F
torch.manual_seed(0)
M=98304
flush = torch.empty(256<<20, dtype=torch.uint8, device="cuda")
def t(fn, it=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ms=[]
    for _ in range(it):
        flush.zero_()
        a,b=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ms.append(a.elapsed_time(b))
    ms.sort(); return ms[len(ms)//2]*1e3
for n_in,n_out in ((768,768),(768,2304)):
    x=torch.randn(M,n_in,device="cuda").bfloat16(); w=torch.randn(n_out,n_in,device="cuda").bfloat16()*0.02; b=torch.randn(n_out,device="cuda").bfloat16()
    wt=w.t().contiguous(); dy=torch.randn(M,n_out,device="cuda").bfloat16()
    fl=2*M*n_in*n_out
    r={}
    r["linear TN+bias"]=t(lambda: F.linear(x,w,b))
    r["addmm NN+bias"]=t(lambda: torch.addmm(b,x,wt))
    r["mm NN nobias"]=t(lambda: torch.mm(x,wt))
    r["linear nobias"]=t(lambda: F.linear(x,w))
    r["dx = dy@w (NN)"]=t(lambda: torch.mm(dy,w))
    r["dx = dy@wt.t() (NT)"]=t(lambda: torch.mm(dy,wt.t()))
    r["dw = dy.t()@x bf16"]=t(lambda: torch.mm(dy.t(),x))
    r["dw f32 out"]=t(lambda: torch.mm(dy.t(),x,out_dtype=torch.float32))
    r["dwT = x.t()@dy bf16"]=t(lambda: torch.mm(x.t(),dy))
    for k,v in r.items(): print(f"{n_in}->{n_out} {k:24s} {v:8.1f} us  {fl/v/1e6:7.1f} TF/s")

# ---- the hand-written tcgen05 GEMM against the library on the same problem
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from meant_b200 import _native as nv
for n_in, n_out in ((768, 768), (768, 2304)):
    x = torch.randn(M, n_in, device="cuda").bfloat16(); w = (torch.randn(n_out, n_in, device="cuda") * 0.02).bfloat16()
    b = torch.randn(n_out, device="cuda"); c = torch.empty(M, n_out, device="cuda").bfloat16()
    fl = 2 * M * n_in * n_out
    def mine():
        rc = nv.lib().meant_gemm_bf16_tn(nv.ptr(x), nv.ptr(w), nv.ptr(b), nv.ptr(c), M, n_out, n_in, n_in, n_in, n_out, nv.stream())
        assert rc == 0
    v = t(mine)
    ref = F.linear(x, w, b.bfloat16())
    print(f"{n_in}->{n_out} meant_gemm_bf16_tn       {v:8.1f} us  {fl/v/1e6:7.1f} TF/s   max|diff| vs F.linear {float((c.float()-ref.float()).abs().max()):.3e}")

# ---- the general forms (gemm_tc2.cu): forward through the general kernel, dgrad (NN), wgrad (stream-K, fp32 + bias grad)
for n_in, n_out in ((768, 768), (768, 2304)):
    x = torch.randn(M, n_in, device="cuda").bfloat16(); w = (torch.randn(n_out, n_in, device="cuda") * 0.02).bfloat16()
    dy = (torch.randn(M, n_out, device="cuda") * 0.1).bfloat16(); b = torch.randn(n_out, device="cuda")
    res = torch.randn(M, n_out, device="cuda").bfloat16()
    c = torch.empty(M, n_out, device="cuda").bfloat16(); dx = torch.empty(M, n_in, device="cuda").bfloat16()
    dw = torch.zeros(n_out, n_in, device="cuda"); db = torch.zeros(n_out, device="cuda")
    fl = 2 * M * n_in * n_out
    L = nv.lib()
    def fwd(): assert L.meant_gemm_bf16(nv.ptr(x), nv.ptr(w), nv.ptr(c), M, n_out, n_in, n_in, n_in, n_out, 0, 0, nv.ptr(b), None, 0, nv.stream()) == 0
    def fwd_res(): assert L.meant_gemm_bf16(nv.ptr(x), nv.ptr(w), nv.ptr(c), M, n_out, n_in, n_in, n_in, n_out, 0, 0, nv.ptr(b), nv.ptr(res), n_out, nv.stream()) == 0
    def dgrad(): assert L.meant_gemm_bf16(nv.ptr(dy), nv.ptr(w), nv.ptr(dx), M, n_in, n_out, n_out, n_in, n_in, 0, 1, None, None, 0, nv.stream()) == 0
    def wgrad(): assert L.meant_gemm_bf16_wgrad(nv.ptr(dy), nv.ptr(x), nv.ptr(dw), nv.ptr(db), M, n_out, n_in, n_out, n_in, n_in, nv.stream()) == 0
    def wgrad_nodb(): assert L.meant_gemm_bf16_wgrad(nv.ptr(dy), nv.ptr(x), nv.ptr(dw), None, M, n_out, n_in, n_out, n_in, n_in, nv.stream()) == 0
    def lib_wgrad_db():
        torch.mm(dy.t(), x, out_dtype=torch.float32); dy.float().sum(0)
    for name, fn in (("general fwd + bias", fwd), ("general fwd + bias + res", fwd_res), ("dgrad dX = dY W", dgrad),
                     ("wgrad fp32 + db", wgrad), ("wgrad fp32", wgrad_nodb)):
        v = t(fn)
        print(f"{n_in}->{n_out} {name:26s} {v:8.1f} us  {fl/v/1e6:7.1f} TF/s")
    v = t(lambda: torch.mm(dy.t(), x, out_dtype=torch.float32)); print(f"{n_in}->{n_out} {'library wgrad fp32':26s} {v:8.1f} us  {fl/v/1e6:7.1f} TF/s")
    v = t(lambda: torch.mm(dy, w)); print(f"{n_in}->{n_out} {'library dgrad':26s} {v:8.1f} us  {fl/v/1e6:7.1f} TF/s")
    dbk = torch.empty(n_out, device="cuda")
    v = t(lambda: L.meant_bias_grad(nv.ptr(dy), nv.ptr(dbk), M, n_out, n_out, 1, nv.stream())); print(f"{n_in}->{n_out} {'meant_bias_grad pass':26s} {v:8.1f} us")
