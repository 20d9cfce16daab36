"""Development check: toy RQS flow training step, fused FP32 kernel vs the stored-activation tensor-core engine."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from normalizing_flows_b200 import ops, _lib
from oracle import flows_oracle as O

sp = O.RQS_TOY
params = O.init_coupling_params(sp, seed=1234, dtype=torch.float32, scale=0.1)
flat = O.flatten_layers(params).float().cuda()
for n in (4096, 1 << 18):
    x = (torch.randn(n, 2, generator=torch.Generator().manual_seed(1)) * 2).cuda()
    out = {}
    for prec in (0, 3):
        spec = ops.make_spec(2, 8, [64, 64], 2, 12, -8.0, 8.0, _lib.BASE_NORMAL, prec)
        for _ in range(2):
            loss, grad = ops.coupling_nll_grad(spec, flat, x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            loss, grad = ops.coupling_nll_grad(spec, flat, x)
        e1.record(); torch.cuda.synchronize()
        out[prec] = (e0.elapsed_time(e1) / 3, loss.item(), grad)
        print(f"n={n} prec={prec}: {out[prec][0]:.3f} ms, loss {out[prec][1]:.6f}", flush=True)
    d = (out[0][2] - out[3][2]).abs().max().item() / out[0][2].abs().max().item()
    print(f"  tc vs fp32 grad max diff / max |g| = {d:.2e}")
