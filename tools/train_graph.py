"""Does the training step gain from CUDA-graph replay?  value_and_grad (flowcuda_coupling_nll_grad, its internal side
stream joined by events) captured once per distinct batch, Adam launched eagerly after each replay (its step count is a
host scalar).  Prints eager and replayed ms per step and whether the replayed gradient equals the eager one bit for bit.
    python tools/train_graph.py [workload] [K]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from normalizing_flows_b200 import _lib, ops

name = sys.argv[1] if len(sys.argv) > 1 else "maf_mnist"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
w = bench.WORKLOADS[name]()
w.setup(dev)
sp = w._ospec()
batch = 4096 if sp.event_dim != 2 else 1 << 18
tr = {"affine_scalar": 0, "affine": 1, "rqs": 2}[sp.transform]
spec = ops.make_spec(sp.event_dim, sp.num_layers, sp.hidden, tr, sp.num_bins, sp.range_min, sp.range_max,
                     _lib.BASE_UNIFORM01 if sp.base == "uniform" else _lib.BASE_NORMAL,
                     _lib.PRECISION_BF16X3 if sp.event_dim != 2 else 0)
p0 = w.model.params.clone()
gen = torch.Generator().manual_seed(77)
xs = [w.make_inputs(gen, batch).to(dev) for _ in range(4)]


def timed(step):
    for i in range(3):
        step(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        step(3 + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / K


out = {"workload": name, "batch": batch, "steps": K}

p = p0.clone()
m, v = torch.zeros_like(p), torch.zeros_like(p)


def eager(i):
    loss, grad = ops.coupling_nll_grad(spec, p, xs[i % len(xs)])
    ops.adam_step(p, grad, m, v, i + 1, 1e-4)


out["eager_ms"] = timed(eager)
p_eager = p.clone()

p = p0.clone()
m, v = torch.zeros_like(p), torch.zeros_like(p)
caps = []
side = torch.cuda.Stream()
try:
    for x in xs:
        g = torch.cuda.CUDAGraph()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            ops.coupling_nll_grad(spec, p, x)  # warm-up on the capture stream
        with torch.cuda.graph(g, stream=side):
            loss, grad = ops.coupling_nll_grad(spec, p, x)
        caps.append((g, loss, grad))
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()

    def replay(i):
        g, loss, grad = caps[i % len(caps)]
        g.replay()
        ops.adam_step(p, grad, m, v, i + 1, 1e-4)

    out["graph_ms"] = timed(replay)
    out["params_equal_after_training"] = bool(torch.equal(p, p_eager))
    out["max_abs_param_diff"] = float((p - p_eager).abs().max().item())
except Exception as e:  # noqa: BLE001
    out["graph_error"] = repr(e)[:400]
print(json.dumps(out))
