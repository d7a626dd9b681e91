"""Debug helper: is the attention backward independent of what earlier kernels left in shared memory / TMEM?"""
import math, torch, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from meant_b200 import ops, RotaryEmbedding, _native as nv
N,S,H,DH=768,128,8,96; DIM=H*DH
def inputs(seed):
    g=torch.Generator(device="cuda").manual_seed(seed)
    qkv=torch.randn(N,S,3*DIM,device="cuda",generator=g).bfloat16()
    cut=torch.randint(0,S//4+1,(N,1),device="cuda",generator=g)
    mask=(torch.arange(S,device="cuda")[None]<(S-cut)).float()
    return qkv,mask
rot=RotaryEmbedding(dim=48,use_xpos=True).tables(S,"cuda")
def attn(qkv,mask): return ops.attention(qkv,heads=H,scale=1/math.sqrt(DIM),causal=True,rot=rot,key_mask=mask,impl=nv.IMPL_TCGEN05)
qkv,mask=inputs(3); qkv.requires_grad_(True)
out=attn(qkv,mask)
dout=torch.randn(out.shape,device="cuda",generator=torch.Generator(device="cuda").manual_seed(9)).bfloat16()
clean=torch.autograd.grad(out,qkv,dout,retain_graph=True)[0].float()
# pollute: forward-only launches on other data (V = ones), as in tests/test_properties_gpu.py
q2,m2=inputs(2); q2[:,:,2*DIM:]=1.0
_=attn(q2,m2); _=attn(q2*2,m2)
dirty=torch.autograd.grad(out,qkv,dout,retain_graph=True)[0].float()
d=(dirty!=clean)
print("after fwd pollution: mismatches",int(d.sum()))
if d.any():
    idx=d.nonzero()
    part=idx[:,2]//DIM
    print("by part q/k/v", [(int((part==i).sum())) for i in range(3)])
    item=idx[:,0]*H+(idx[:,2]%DIM)//DH
    u=torch.unique(item)
    print("distinct items",int(u.numel()),"first",u[:12].tolist(),"max",int(u.max()))
    print("items mod 148 (CTA) first:", (u[:12]%148).tolist(), " j = item//148:", (u[:12]//148).tolist())
    print("j histogram:", torch.bincount(u//148, minlength=42).tolist())
    print("rows hist/16", torch.bincount(idx[:,1]//16, minlength=8).tolist())
    rel=((dirty-clean).abs()/clean.abs().clamp_min(1e-20))[d]
    print("max rel", float(rel.max()), "median rel", float(rel.median()))
again=torch.autograd.grad(out,qkv,dout,retain_graph=True)[0].float()
print("third run vs clean mismatches", int((again!=clean).sum()), " vs dirty", int((again!=dirty).sum()))
