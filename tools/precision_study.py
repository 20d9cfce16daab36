import torch, math, sys
sys.path.insert(0,'/root/repo')
from oracle import flows_oracle as O
torch.manual_seed(0)
def split_bf16(a, n):
    parts=[]; r=a.clone()
    for _ in range(n):
        h=r.bfloat16().float(); parts.append(h); r=r-h
    return parts
def trunc_tf32(a):
    return (a.view(torch.int32) & ~0x1fff).view(torch.float32)
def split_tf32(a,n):
    parts=[]; r=a.clone()
    for _ in range(n):
        h=trunc_tf32(r); parts.append(h); r=r-h
    return parts
def mm(a,W,mode):
    a=a.float(); W=W.float()
    if mode=='fp32': return a@W
    if mode=='bf16': return (a.bfloat16().float().double()@W.bfloat16().float().double()).float()
    if mode=='tf32': return (trunc_tf32(a).double()@trunc_tf32(W).double()).float()
    if mode=='bf16x3':
        ah,al=split_bf16(a,2); wh,wl=split_bf16(W,2)
        return (ah.double()@wh.double()+ah.double()@wl.double()+al.double()@wh.double()).float()
    if mode=='tf32x3':
        ah,al=split_tf32(a,2); wh,wl=split_tf32(W,2)
        return (ah.double()@wh.double()+ah.double()@wl.double()+al.double()@wh.double()).float()
    if mode=='bf16x6':
        a0,a1,a2=split_bf16(a,3); w0,w1,w2=split_bf16(W,3)
        d=lambda x,y:x.double()@y.double()
        return (d(a0,w0)+d(a0,w1)+d(a1,w0)+d(a0,w2)+d(a2,w0)+d(a1,w1)).float()
def cond(layer_params,x,mode,first_fp32=True):
    act=x
    for i,(W,b) in enumerate(layer_params[:-1]):
        m = 'fp32' if (first_fp32 and i==0) else mode
        act=torch.relu(mm(act,W,m)+b.float())
    W,b=layer_params[-1]
    return mm(act,W,mode)+b.float()
def logprob(spec,params,v,mode,first_fp32):
    v=v.float().reshape(v.shape[0],-1); tot=torch.zeros(v.shape[0])
    for l in reversed(range(spec.num_layers)):
        mask=spec.layer_mask(l)
        mv=torch.where(mask,v,torch.zeros_like(v))
        c=cond(params[l],mv,mode,first_fp32)
        out0,ld=O._inner_bijector(spec,c,v,False)
        v=torch.where(mask,v,out0); tot=tot+torch.where(mask,torch.zeros_like(ld),ld).sum(-1)
    return O.base_log_prob(spec,v)+tot
which=sys.argv[1]
if which=='toy':
    for scale in (0.1,0.3,0.5):
        spec=O.RQS_TOY; p=O.init_coupling_params(spec,seed=2,scale=scale)
        x=torch.randn(4096,2,dtype=torch.float64)*2
        ref=O.coupling_flow_log_prob(spec,p,x)
        for mode in ('fp32','tf32','bf16x3','tf32x3','bf16x6'):
            lp=logprob(spec,p,x,mode,True).double()
            e=((lp-ref).abs()/ref.abs().clamp(min=1)).max().item()
            print(scale,mode,'max scaled err %.2e'%e)
else:
    spec=O.RQS_MNIST; p=O.init_coupling_params(spec,seed=1234,scale=0.1)
    g=torch.Generator().manual_seed(1)
    x=((torch.randint(0,256,(64,784),generator=g).float()+torch.rand(64,784,generator=g))/256).double()
    ref=O.coupling_flow_log_prob(spec,p,x)
    print('ref',ref[:4])
    for mode in ('fp32','bf16','tf32','bf16x3'):
        lp=logprob(spec,p,x,mode,False).double()
        print(mode,'max abs err %.2e'%(lp-ref).abs().max().item(), 'mean abs %.2e'%(lp-ref).abs().mean().item())
