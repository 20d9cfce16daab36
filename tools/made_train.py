import os, sys, json
sys.path.insert(0, "/root/repo")
import torch, bench
from normalizing_flows_b200 import ops
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
for name in sys.argv[1:]:
    w = bench.WORKLOADS[name](); w.setup(dev)
    print(name, json.dumps(w.train(dev, 0, 1, None, ops, K=10))[:160])
