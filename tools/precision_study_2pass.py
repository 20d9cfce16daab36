import torch, sys
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tools')
from oracle import flows_oracle as O
torch.manual_seed(0)
def split(a,n,dt):
    parts=[]; r=a.clone()
    for _ in range(n):
        h=r.to(dt).float(); parts.append(h); r=r-h
    return parts
d=lambda x,y:(x.double()@y.double())
def mm(a,W,mode):
    a=a.float(); W=W.float()
    if mode=='fp32': return a@W
    dt = torch.float16 if mode.startswith('fp16') else torch.bfloat16
    if mode in('fp16','bf16'): return d(a.to(dt).float(),W.to(dt).float()).float()
    ah,al=split(a,2,dt); wh,wl=split(W,2,dt)
    if mode.endswith('x3'): return (d(ah,wh)+d(ah,wl)+d(al,wh)).float()
    if mode.endswith('x2a'): return (d(ah,wh)+d(al,wh)).float()   # A split, W hi only
    if mode.endswith('x2w'): return (d(ah,wh)+d(ah,wl)).float()   # W split, A hi only
def cond(layer_params,x,modes):
    act=x
    n=len(layer_params)
    for i,(W,b) in enumerate(layer_params[:-1]):
        act=torch.relu(mm(act,W,modes[0])+b.float())
    W,b=layer_params[-1]
    return mm(act,W,modes[1])+b.float()
def logprob(spec,params,v,modes):
    v=v.float().reshape(v.shape[0],-1); tot=torch.zeros(v.shape[0])
    for l in reversed(range(spec.num_layers)):
        mask=spec.layer_mask(l)
        mv=torch.where(mask,v,torch.zeros_like(v))
        c=cond(params[l],mv,modes)
        out0,ld=O._inner_bijector(spec,c,v,False)
        v=torch.where(mask,v,out0); tot=tot+torch.where(mask,torch.zeros_like(ld),ld).sum(-1)
    return O.base_log_prob(spec,v)+tot
spec=O.RQS_MNIST
for scale in (0.1,):
  p=O.init_coupling_params(spec,seed=1234,scale=scale)
  g=torch.Generator().manual_seed(1)
  n=int(sys.argv[1]) if len(sys.argv)>1 else 64
  x=((torch.randint(0,256,(n,784),generator=g).float()+torch.rand(n,784,generator=g))/256).double()
  ref=O.coupling_flow_log_prob(spec,p,x)
  print('ref',ref[:4])
  for modes in (('fp32','fp32'),('bf16x3','bf16x3'),('fp16','fp16'),('fp16x2a','fp16x2a'),('fp16x2w','fp16x2w'),('bf16x2a','bf16x2a'),('bf16x2w','bf16x2w'),('bf16x3','fp16x2a'),('bf16x3','fp16x2w'),('bf16x3','fp16'),('fp16x3','fp16x3'),('fp16x3','fp16x2w'), ('fp16x3','fp16x2a')):
    lp=logprob(spec,p,x,modes).double()
    e=(lp-ref).abs()
    print(scale,modes,'max abs err %.2e'%e.max().item(), 'mean abs %.2e'%e.mean().item())
