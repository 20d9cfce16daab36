"""CPU, world_size 2 over gloo: the host-side multi-GPU logic (batch sharding, scalar mean
log-likelihood all-reduce, gradient all-reduce + 1/N folded into Adam) gives the same result as a
single process on the concatenated batch.  Gradients come from the oracle here (no GPU)."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import flows_oracle as O

SPEC = O.CouplingSpec(2, 3, [16, 16], "rqs", 8, -4.0, 4.0)


def _grad_fn(params, x_local):
    shapes = [(m, n) for _ in range(SPEC.num_layers) for (m, n) in SPEC.layer_shapes()]

    def lp_fn(leaves):
        flat, o, ps = leaves[0], 0, []
        for l in range(SPEC.num_layers):
            layer = []
            for (m, n) in SPEC.layer_shapes():
                W = flat[o:o + m * n].reshape(m, n); o += m * n
                b = flat[o:o + n]; o += n
                layer.append((W, b))
            ps.append(layer)
        return O.coupling_flow_log_prob(SPEC, ps, x_local)

    p = params.detach().clone().requires_grad_(True)
    loss = -lp_fn([p]).sum()
    (g,) = torch.autograd.grad(loss, [p])
    return loss.detach().reshape(1), g


def _adam_fn(params, grad, m, v, step, lr, grad_scale):
    pn, mn, vn = O.adam_step(params, grad * grad_scale, m, v, step, lr)
    params.copy_(pn); m.copy_(mn); v.copy_(vn)


def _data():
    g = torch.Generator().manual_seed(5)
    x = torch.randn(37, 2, generator=g, dtype=torch.float64) * 1.5
    p0 = O.flatten_layers(O.init_coupling_params(SPEC, seed=9, scale=0.3))
    return x, p0


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from normalizing_flows_b200 import data_parallel as dp

    x, p0 = _data()
    xl = dp.shard_batch(x)
    tr = dp.DataParallelTrainer(p0.clone(), _grad_fn, _adam_fn, lr=1e-2)
    losses = [tr.step(xl).item() for _ in range(3)]
    lp = O.coupling_flow_log_prob(SPEC, _unflat(tr.params), xl)
    mean_lp = dp.mean_log_prob(lp).item()
    if rank == 0:
        torch.save({"params": tr.params, "losses": losses, "mean_lp": mean_lp}, out)
    dist.destroy_process_group()


def _unflat(flat):
    o, ps = 0, []
    for l in range(SPEC.num_layers):
        layer = []
        for (m, n) in SPEC.layer_shapes():
            W = flat[o:o + m * n].reshape(m, n); o += m * n
            b = flat[o:o + n]; o += n
            layer.append((W, b))
        ps.append(layer)
    return ps


def test_two_ranks_equal_single_process(tmp_path):
    from normalizing_flows_b200 import data_parallel as dp

    assert [dp.shard_bounds(37, r, 2) for r in range(2)] == [(0, 19), (19, 37)]
    assert [dp.shard_bounds(8, r, 4) for r in range(4)] == [(0, 2), (2, 4), (4, 6), (6, 8)]
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = torch.load(out)
    x, p0 = _data()
    tr = dp.DataParallelTrainer(p0.clone(), _grad_fn, _adam_fn, lr=1e-2)  # single process, whole batch
    losses = [tr.step(x).item() for _ in range(3)]
    np.testing.assert_allclose(got["losses"], losses, rtol=1e-12)
    np.testing.assert_allclose(got["params"].numpy(), tr.params.numpy(), rtol=1e-10, atol=1e-12)
    ref_mean = O.coupling_flow_log_prob(SPEC, _unflat(tr.params), x).mean().item()
    np.testing.assert_allclose(got["mean_lp"], ref_mean, rtol=1e-12)
