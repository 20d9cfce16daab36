"""One coupling-flow training step (value_and_grad + Adam) at the maf_mnist shape, for `ncu` launch lists."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import ops, _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
prec = int(sys.argv[2]) if len(sys.argv) > 2 else 3
spec = ops.make_spec(784, 8, [512, 512, 512], 2, 8, 0.0, 1.0, _lib.BASE_UNIFORM01, prec)
cnt = ops.coupling_param_count(spec) if hasattr(ops, "coupling_param_count") else ops.param_count(spec)
p = torch.randn(cnt, device="cuda") * 0.02
m, v = torch.zeros_like(p), torch.zeros_like(p)
x = torch.rand(n, 784, device="cuda")
for i in range(2):
    loss, g = ops.coupling_nll_grad(spec, p, x)
    ops.adam_step(p, g, m, v, i + 1, 1e-4)
torch.cuda.synchronize()
print(float(loss))
