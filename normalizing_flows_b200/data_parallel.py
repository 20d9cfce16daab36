"""Batch-sharded evaluation and training across the GPUs of one box: one process per
GPU, parameters replicated, samples split evenly over ranks, NO collective on the data path.
The only exchanges are (i) one scalar all-reduce for the mean log-likelihood and (ii) one sum
all-reduce of the flat fp32 gradient buffer per training step; the 1/world factor is folded
into the optimiser kernel's `grad_scale`.  `torch.distributed` (NCCL on GPUs, gloo in the CPU
tests) is plumbing only; gradients and updates come from libflowcuda kernels unless the caller
injects other callables (the CPU tests inject CPU ones)."""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced split: the first n % world ranks get one extra sample."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_batch(x: torch.Tensor, rank: Optional[int] = None, world: Optional[int] = None) -> torch.Tensor:
    rank = dist.get_rank() if rank is None else rank
    world = dist.get_world_size() if world is None else world
    lo, hi = shard_bounds(x.shape[0], rank, world)
    return x[lo:hi]


def mean_log_prob(local_log_prob: torch.Tensor) -> torch.Tensor:
    """Global mean of per-sample log_prob over all ranks: all-reduce of (sum, count)."""
    acc = torch.stack([local_log_prob.double().sum(), torch.tensor(float(local_log_prob.numel()), dtype=torch.float64,
                                                                   device=local_log_prob.device)])
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(acc)
    return acc[0] / acc[1]


class DataParallelTrainer:
    """update() of src/real_nvp_toy_densities.py:164-173 across ranks.

    grad_fn(params, x_local) -> (sum over local samples of -log_prob as a 1-element tensor,
                                 gradient of that SUM w.r.t. params)
    adam_fn(params, grad, m, v, step, lr, grad_scale) updates params/m/v in place."""

    def __init__(self, params: torch.Tensor, grad_fn: Callable, adam_fn: Callable, lr: float = 1e-4):
        self.params = params
        self.m = torch.zeros_like(params)
        self.v = torch.zeros_like(params)
        self.grad_fn, self.adam_fn, self.lr, self.t = grad_fn, adam_fn, lr, 0

    def step(self, x_local: torch.Tensor) -> torch.Tensor:
        loss_sum, grad_sum = self.grad_fn(self.params, x_local)
        count = torch.tensor([float(x_local.shape[0])], dtype=grad_sum.dtype, device=grad_sum.device)
        if dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(grad_sum)            # the one data-dependent exchange of the step
            tail = torch.cat([loss_sum.reshape(1).to(grad_sum.dtype), count])
            dist.all_reduce(tail)
            loss_sum, count = tail[:1], tail[1:]
        self.t += 1
        self.adam_fn(self.params, grad_sum, self.m, self.v, self.t, self.lr, 1.0 / float(count.item()))
        return loss_sum / count


class LayerwiseAllReduce:
    """Sum all-reduce of the flat gradient in the buckets the reverse mode reports (per flow layer: its last Linear as
    soon as that is done, then its hidden Linears), issued while the reverse mode is still working
    on the remaining layers (351 MB per step at the maf_mnist shape) — and, optionally, that layer's
    optimiser step right behind it.  Use as the `on_layer` hook of ops.coupling_nll_grad, then `wait()`:

        upd = LayerwiseAllReduce(after=lambda layer, lo, hi, g: ops.adam_step(p[lo:hi], g, m[lo:hi], v[lo:hi], t, lr,
                                                                              grad_scale=1.0 / world))
        loss, grad = ops.coupling_nll_grad(spec, p, x, on_layer=upd)
        upd.wait()                                 # the current stream now sees the reduced gradient / updated params

    The hook makes a communication stream wait for the event the layer completes on and enqueues the collective (and
    `after`) there; nothing blocks the host.  The C call guarantees that the layer's parameters are no longer read
    at that point, so `after` may update them in place.  Single process: no collective, `after` still runs there."""

    def __init__(self, group=None, after=None):
        self.group = group
        self.after = after
        self.works = []
        self.comm = None

    def __call__(self, layer, grad_slice, stream):
        multi = dist.is_initialized() and dist.get_world_size(self.group) > 1
        if not multi and self.after is None:
            return
        if self.comm is None:
            self.comm = torch.cuda.Stream(device=grad_slice.device)
        ev = torch.cuda.Event()
        ev.record(stream)
        self.comm.wait_event(ev)
        with torch.cuda.stream(self.comm):
            if multi:
                w = dist.all_reduce(grad_slice, group=self.group, async_op=True)
                if self.after is None:
                    self.works.append(w)
                else:
                    w.wait()  # the communication stream (not the host) waits for the collective
            if self.after is not None:
                lo = grad_slice.storage_offset()  # the slice is a view of the flat gradient
                self.after(layer, lo, lo + grad_slice.numel(), grad_slice)

    def wait(self):
        for w in self.works:
            w.wait()  # the CURRENT stream waits for the collective; the host does not
        self.works = []
        if self.comm is not None:
            torch.cuda.current_stream().wait_stream(self.comm)


def cuda_grad_fn(spec):
    """libflowcuda gradient of the SUM of -log_prob over the local shard (coupling or MADE spec)."""
    from . import _lib, ops

    nll_grad = ops.made_nll_grad if isinstance(spec, _lib.MadeSpec) else ops.coupling_nll_grad

    def fn(params, x_local):
        loss, grad = nll_grad(spec, params, x_local)
        n = float(x_local.shape[0])
        return loss * n, grad * n

    return fn


def cuda_adam_fn(params, grad, m, v, step, lr, grad_scale):
    from . import ops

    ops.adam_step(params, grad, m, v, step, lr, grad_scale=grad_scale)
