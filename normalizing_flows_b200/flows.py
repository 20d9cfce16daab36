"""Model builders mirroring the reference scripts' `make_conditioner` / `make_flow_model`
(real_nvp_toy_densities.py:56-110, maf_toy_densities.py:55-111, maf_mnist.py:34-89): same
arguments, same masks, same base distributions; parameters live in ONE flat CUDA buffer so
that `Transformed.log_prob` / `.sample` are single fused kernel calls."""
from __future__ import annotations

import math
from typing import Optional, Sequence

import torch

from . import _lib
from . import distrax_api as distrax


def make_conditioner(event_shape: Sequence[int], hidden_sizes: Sequence[int], num_bijector_params: int,
                     params: Optional[torch.Tensor] = None, per_dim: bool = False, seed: int = 0) -> distrax.MLPConditioner:
    """hk.Sequential([Flatten, MLP(hidden, activate_final=True), Linear(zeros), Reshape]).
    per_dim=True reshapes to event_shape + (num_bijector_params,) (maf_toy_densities.py:63-68)."""
    D = int(math.prod(event_shape))
    out = D * num_bijector_params if per_dim else num_bijector_params
    return distrax.MLPConditioner(event_shape, hidden_sizes, num_bijector_params, params=params, out_features=out, seed=seed)


def make_flow_model(event_shape: Sequence[int], num_layers: int, hidden_sizes: Sequence[int], *,
                    transform: str = "scalar_affine", num_bins: int = 12, range_min: float = -8.0,
                    range_max: float = 8.0, base: str = "normal", params: Optional[torch.Tensor] = None,
                    precision: int = _lib.PRECISION_FP32, seed: int = 42) -> distrax.Transformed:
    """transform="scalar_affine": real_nvp_toy_densities.py:74-110 (ScalarAffine, 2 outputs);
    "rqs": maf_toy_densities.py:73-111 / maf_mnist.py:52-89 (3*num_bins+1 outputs per dim);
    "affine": per-dimension affine.  base "normal" | "uniform" (maf_mnist.py:85-88)."""
    event_shape = tuple(int(s) for s in event_shape)
    D = int(math.prod(event_shape))
    mask = (torch.arange(D) % 2).reshape(event_shape).bool()  # every second element is masked
    if transform == "scalar_affine":
        fn, nbp, per_dim = distrax.scalar_affine_fn, 2, False
    elif transform == "affine":
        fn, nbp, per_dim = distrax.affine_fn, 2, True
    elif transform == "rqs":
        fn, nbp, per_dim = distrax.rational_quadratic_spline_fn(range_min, range_max, num_bins), 3 * num_bins + 1, True
    else:
        raise ValueError(transform)
    probe = distrax.MLPConditioner(event_shape, hidden_sizes, nbp, params=torch.empty(0) if False else None,
                                   out_features=D * nbp if per_dim else nbp, device="cpu", seed=seed)
    per_layer = probe.count
    if params is None:
        gen = torch.Generator().manual_seed(seed)
        params = torch.cat([probe.init(gen) for _ in range(num_layers)]).to("cuda")
    params = params.to("cuda", torch.float32).contiguous()
    if params.numel() != per_layer * num_layers:
        raise ValueError(f"expected {per_layer * num_layers} parameters, got {params.numel()}")
    layers = []
    for l in range(num_layers):
        cond = make_conditioner(event_shape, hidden_sizes, nbp, params=params[l * per_layer:(l + 1) * per_layer], per_dim=per_dim)
        layers.append(distrax.MaskedCoupling(mask=mask, bijector=fn, conditioner=cond))
        mask = torch.logical_not(mask)  # flip mask after each layer
    flow = distrax.Inverse(distrax.Chain(layers))
    if base == "normal":
        base_dist = distrax.Independent(distrax.Normal(torch.zeros(event_shape), torch.ones(event_shape)), len(event_shape))
    elif base == "uniform":
        base_dist = distrax.Independent(distrax.Uniform(torch.zeros(event_shape), torch.ones(event_shape)), len(event_shape))
    else:
        raise ValueError(base)
    model = distrax.Transformed(base_dist, flow, precision=precision)
    model.params = params
    return model


def make_maf_model(event_shape: Sequence[int], num_layers: int, hidden_sizes: Sequence[int], *, base: str = "normal",
                   params: Optional[torch.Tensor] = None, seed: int = 42) -> distrax.Transformed:
    """MADE-conditioned masked autoregressive flow (the `tfb.MaskedAutoregressiveFlow` stack north_star
    names for the D=784 configs) over a flattened event, with the reference scripts'
    base distributions (maf_mnist.py:84-89)."""
    event_shape = tuple(int(s) for s in event_shape)
    D = int(math.prod(event_shape))
    flow = distrax.MaskedAutoregressiveFlow(D, num_layers, hidden_sizes, params=params, seed=seed)
    if base == "normal":
        base_dist = distrax.Independent(distrax.Normal(torch.zeros(D), torch.ones(D)), 1)
    elif base == "uniform":
        base_dist = distrax.Independent(distrax.Uniform(torch.zeros(D), torch.ones(D)), 1)
    else:
        raise ValueError(base)
    model = distrax.Transformed(base_dist, flow)
    model.params = flow.params
    return model
