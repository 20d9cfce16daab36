"""A few toy-flow training steps (fused FP32 reverse mode), for `ncu` captures."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import ops, _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
rqs = (sys.argv[2] if len(sys.argv) > 2 else "rqs") == "rqs"
spec = ops.make_spec(2, 8, [64, 64], 2 if rqs else 0, 12, -8.0, 8.0, _lib.BASE_NORMAL, 0)
p = torch.randn(ops.coupling_param_count(spec), device="cuda") * 0.05
x = torch.randn(n, 2, device="cuda") * 2
for _ in range(2):
    loss, g = ops.coupling_nll_grad(spec, p, x)
torch.cuda.synchronize()
print(float(loss))
