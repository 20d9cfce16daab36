"""Smallest cases of the hand-synchronised kernels (mbarrier / TMEM / TMA), for compute-sanitizer:
coupling_toy_tc_kernel, dense_tc_gemm_kernel (hidden / RQS / affine epilogues, n-half and tail-split items),
dense_tc2_gemm_kernel (CTA pairs and 4-CTA multicast clusters), mgtc_gemm_kernel, made_sample_kernel.
    compute-sanitizer --tool {memcheck|racecheck|synccheck} python tools/sanitize_case.py
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import _lib, ops
from oracle import flows_oracle as O

dev = "cuda"
g = torch.Generator().manual_seed(0)
# D=2 tensor-core kernel
p = O.init_coupling_params(O.RQS_TOY, seed=2, scale=0.3)
spec = ops.make_spec(2, 8, [64, 64], 2, 12, -8.0, 8.0, 0, _lib.PRECISION_3XTF32)
x = torch.randn(300, 2, generator=g) * 2
lp, _, _ = ops.coupling_log_prob(spec, O.flatten_layers(p).float().to(dev), x.to(dev))
print("toy_tc", float(lp.sum()))
# D=40 RQS coupling on the bf16x3 GEMMs, all three kernels
so = O.CouplingSpec(40, 2, [32, 32, 32], "rqs", 8, 0.0, 1.0, "uniform")
dp = O.flatten_layers(O.init_coupling_params(so, seed=11, scale=0.3)).float().to(dev)
ds = ops.make_spec(40, 2, [32, 32, 32], 2, 8, 0.0, 1.0, 1, _lib.PRECISION_BF16X3)
xd = torch.rand(300, 40, generator=g).to(dev)
for mode in ("", "pair", "quad"):
    if mode:
        os.environ["FLOWCUDA_DENSE_TC"] = mode
    else:
        os.environ.pop("FLOWCUDA_DENSE_TC", None)
    lp, _, _ = ops.coupling_log_prob(ds, dp, xd)
    print("dense rqs", mode or "1cta", float(lp[torch.isfinite(lp)].sum()))
os.environ.pop("FLOWCUDA_DENSE_TC", None)
# affine epilogue + MADE (masked K ranges), sampler, reverse-mode GEMMs
mo = O.MadeSpec(32, 3, [40, 40])
mp = O.flatten_layers(O.init_made_params(mo, seed=3, scale=0.4)).float().to(dev)
ms = ops.make_made_spec(32, 3, [40, 40])
xm = torch.randn(300, 32, generator=g).to(dev)
print("made log_prob", float(ops.made_log_prob(ms, mp, xm)[0].sum()))
print("made sample", float(ops.made_sample(ms, mp, xm[:16]).sum()))
ts = ops.make_made_spec(32, 3, [40, 40], precision=_lib.PRECISION_BF16X3)
loss, grad = ops.made_nll_grad(ts, mp, xm)
print("made grad (mgtc)", float(loss), float(grad.abs().sum()))
torch.cuda.synchronize()
print("done")
