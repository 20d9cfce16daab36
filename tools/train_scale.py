"""Training step of one workload under torchrun (N ranks): bench.time_train's figure only.
    NCCL_MAX_CTAS=8 python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29540 tools/train_scale.py maf_mnist 10"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import bench
from normalizing_flows_b200 import ops

name = sys.argv[1] if len(sys.argv) > 1 else "maf_mnist"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 10
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dev = torch.device("cuda", local)
torch.cuda.set_device(dev)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
w = bench.WORKLOADS[name]()
w.setup(dev)
r = bench.time_train(w, dev, rank, world, dist if world > 1 else None, ops, K=K)
if rank == 0:
    print(json.dumps({"world": world, "nccl_max_ctas": os.environ.get("NCCL_MAX_CTAS"), "ms_per_step": r["ms_per_step"], "value": r["value"]}))
if world > 1:
    dist.destroy_process_group()
