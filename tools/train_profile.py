"""The maf_mnist training step (value_and_grad + Adam, src/maf_mnist.py:155-166) on its own.
    python tools/train_profile.py [workload] [K]          -> bench.time_train's figure (eager launches, K timed steps)
    ncu --metrics gpu__time_duration.sum,... --csv --log-file gpurun_out/launches_train.csv \
        python tools/train_profile.py maf_mnist 0         -> two bare steps for a launch list (tools/launch_table.py ... 2)"""
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
if K > 0:
    print(json.dumps(bench.time_train(w, dev, 0, 1, None, ops, K=K)))
    sys.exit(0)
sp = w._ospec()
batch = 4096 if sp.event_dim != 2 else 1 << 18
tr = {"affine_scalar": 0, "affine": 1, "rqs": 2}[sp.transform]
spec = ops.make_spec(sp.event_dim, sp.num_layers, sp.hidden, tr, sp.num_bins, sp.range_min, sp.range_max,
                     _lib.BASE_UNIFORM01 if sp.base == "uniform" else _lib.BASE_NORMAL,
                     _lib.PRECISION_BF16X3 if sp.event_dim != 2 else 0)
p = w.model.params.clone()
m, v = torch.zeros_like(p), torch.zeros_like(p)
x = w.make_inputs(torch.Generator().manual_seed(77), batch).to(dev)
for i in range(2):
    loss, grad = ops.coupling_nll_grad(spec, p, x)
    ops.adam_step(p, grad, m, v, i + 1, 1e-4)
torch.cuda.synchronize()
