"""One MADE training step (value_and_grad) at the bench shape, for `ncu` launch lists."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
prec = int(sys.argv[2]) if len(sys.argv) > 2 else 0
spec = ops.make_made_spec(784, 8, [512, 512], precision=prec)
p = torch.randn(ops.made_param_count(spec), device="cuda") * 0.02
x = torch.rand(n, 784, device="cuda")
for _ in range(2):
    loss, g = ops.made_nll_grad(spec, p, x)
torch.cuda.synchronize()
print(float(loss))
