"""A few MADE log_prob calls at the bench shape, for `ncu` captures of the density-path GEMMs."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
spec = ops.make_made_spec(784, 8, [512, 512])
p = torch.randn(ops.made_param_count(spec), device="cuda") * 0.02
x = torch.rand(n, 784, device="cuda")
for _ in range(2):
    lp = ops.made_log_prob(spec, p, x)[0]
torch.cuda.synchronize()
print(float(lp.mean()))
