"""Error of the FP32 and 3xTF32 toy kernels and of the float32 oracle against the float64 oracle."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from oracle import flows_oracle as O
from normalizing_flows_b200 import ops, _lib
from test_gpu_parity import COUPLING_CASES, _spec_pair, G

def sc(got, ref):
    got = got.double().cpu().numpy(); ref = ref.numpy()
    e = np.abs(got - ref) / np.maximum(1, np.abs(ref))
    return e.max(), np.quantile(e, 0.999)

for case in ["realnvp_toy", "rqs_toy", "rqs_toy_k8", "affine_toy"]:
    ospec, scale, xs = COUPLING_CASES[case]
    for n in (1027, 70001, 300000):
        params = O.init_coupling_params(ospec, seed=11, scale=scale)
        flat = O.flatten_layers(params)
        p32 = [[(W.float(), b.float()) for (W, b) in layer] for layer in params]
        x = (torch.randn(n, 2, generator=torch.Generator().manual_seed(n), dtype=torch.float64) * xs).float()
        zref, ldref = O.coupling_flow_to_base(ospec, params, x.double())
        z32, ld32 = O.coupling_flow_to_base(ospec, p32, x)
        row = [case, n, "oracle32 z %.1e/%.1e ld %.1e/%.1e" % (*sc(z32, zref), *sc(ld32, ldref))]
        for name, prec in (("fp32", 0), ("tf32x3", 1)):
            _, z, ld = ops.coupling_log_prob(_spec_pair(ops, ospec, precision=prec), G(flat), x.cuda(), want_z=True, want_log_det=True, want_log_prob=False)
            row.append("%s z %.1e/%.1e ld %.1e/%.1e" % (name, *sc(z, zref), *sc(ld, ldref)))
        print(" | ".join(map(str, row)), flush=True)
