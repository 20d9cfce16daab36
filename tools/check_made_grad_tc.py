"""Development check: made_nll_grad on the tensor-core engine vs float64 autograd (errors) and vs the FP32 engine (time)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import torch
from normalizing_flows_b200 import ops, _lib
from oracle import flows_oracle as O

CASES = {
    "d6": (O.MadeSpec(6, 2, [16, 16]), 0.5, 33),
    "d50_ragged": (O.MadeSpec(50, 3, [40, 70]), 0.4, 100),
    "toy_d2": (O.MadeSpec(2, 8, [64, 64]), 0.3, 257),
    "d784": (O.MadeSpec(784, 2, [512, 512]), 0.3, 48),
    "d50_splitk": (O.MadeSpec(50, 2, [40, 70]), 0.4, 1500),
    "d300_splitk": (O.MadeSpec(300, 2, [200, 260]), 0.3, 1100),
}
for name, (ospec, scale, n) in CASES.items():
    params = O.init_made_params(ospec, seed=8, scale=scale)
    leaves = [t for layer in params for (W, b) in layer for t in (W, b)]
    x = torch.randn(n, ospec.event_dim, generator=torch.Generator().manual_seed(3), dtype=torch.float64).float().double()
    def lp_fn(lv):
        it = iter(lv)
        p = [[(next(it), next(it)) for _ in layer] for layer in params]
        return O.maf_log_prob(ospec, p, x)
    loss_ref, grads_ref = O.nll_and_grads(lp_fn, leaves)
    flat = O.flatten_layers(params).float().cuda()
    for prec in (0, 3):
        spec = ops.make_made_spec(ospec.event_dim, ospec.num_layers, ospec.hidden, precision=prec)
        loss, grad = ops.made_nll_grad(spec, flat, x.float().cuda())
        loss2, grad2 = ops.made_nll_grad(spec, flat, x.float().cuda())
        g = grad.double().cpu(); off = 0; worst = 0.0
        for ref in grads_ref:
            got = g[off:off + ref.numel()].reshape(ref.shape); off += ref.numel()
            worst = max(worst, (got - ref).abs().max().item() / max(ref.abs().max().item(), 1e-6))
        print(f"{name:12s} prec={prec} loss_rel={abs(loss.item()-loss_ref.item())/max(1,abs(loss_ref.item())):.2e} grad_rel_max={worst:.2e} "
              f"repro={torch.equal(grad, grad2)}", flush=True)

ospec = O.MadeSpec(784, 8, [512, 512])
params = O.init_made_params(ospec, seed=1, dtype=torch.float32, scale=0.1)
flat = O.flatten_layers(params).float().cuda()
for n in (1024, 8192):
    x = torch.rand(n, 784, device="cuda")
    out = {}
    for prec in (0, 3):
        spec = ops.make_made_spec(784, 8, [512, 512], precision=prec)
        for _ in range(2):
            loss, grad = ops.made_nll_grad(spec, flat, x)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            loss, grad = ops.made_nll_grad(spec, flat, x)
        e1.record(); torch.cuda.synchronize()
        out[prec] = (e0.elapsed_time(e1) / 3, loss.item(), grad)
        print(f"n={n} prec={prec}: {out[prec][0]:.2f} ms, loss {out[prec][1]:.6f}", flush=True)
    d = (out[0][2] - out[3][2]).abs().max().item() / out[0][2].abs().max().item()
    print(f"  tc vs fp32 grad max diff / max |g| = {d:.2e}")
