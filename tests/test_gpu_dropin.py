"""The reference's own model-building code running against the host mirror, IMPORT LINES CHANGED ONLY.

The string constants below are verbatim excerpts of the reference (cited line ranges; kept to the functions under test):
`make_conditioner` / `make_flow_model` / the hk.transform-ed `log_prob` and `sample` of src/real_nvp_toy_densities.py and
src/maf_toy_densities.py, and the bijector chain of src/bflow.py.  They are exec'd in a namespace whose `distrax`, `hk`,
`jnp`, `jax`, `tfb`, `tfd`, `BernsteinBijector` come from normalizing_flows_b200 instead of the JAX libraries; the results
are compared with the golden vectors that the SAME code produced under the numpy stand-ins (tests/golden/
make_golden_flows.py) and with the float64 oracle.
"""
import os
from typing import Any, Mapping, Sequence, Tuple

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")

# ---- src/real_nvp_toy_densities.py:56-110 -------------------------------------------------------------------------
REAL_NVP_TOY_56_110 = r'''
def make_conditioner(
    event_shape: Sequence[int], hidden_sizes: Sequence[int], num_bijector_params: int
) -> hk.Sequential:

    return hk.Sequential(
        [
            hk.Flatten(preserve_dims=-1),
            hk.nets.MLP(hidden_sizes, activate_final=True),
            hk.Linear(
                num_bijector_params,
                w_init=jnp.zeros,
                b_init=jnp.zeros,
            ),
            hk.Reshape((num_bijector_params,), preserve_dims=-1),
        ]
    )


def make_flow_model(
    event_shape: Sequence[int],
    num_layers: int,
    hidden_sizes: Sequence[int],
) -> distrax.Transformed:

    mask = jnp.arange(0, np.prod(event_shape)) % 2  # every second element is masked
    mask = jnp.reshape(mask, event_shape)
    mask = mask.astype(bool)

    def bijector_fn(params: Array):
        shift, log_scale = jnp.split(params, 2, axis=-1)
        return distrax.ScalarAffine(shift=shift, log_scale=log_scale)

    layers = []
    for _ in range(num_layers):
        layer = distrax.MaskedCoupling(
            mask=mask,
            bijector=bijector_fn,
            conditioner=make_conditioner(
                event_shape=event_shape,
                hidden_sizes=hidden_sizes,
                num_bijector_params=2,
            ),
        )

        layers.append(layer)
        # Flip mask after each layer
        mask = jnp.logical_not(mask)

    flow = distrax.Inverse(distrax.Chain(layers))
    base_distribution = distrax.Independent(
        distrax.Normal(loc=jnp.zeros(event_shape), scale=jnp.ones(event_shape)),
        reinterpreted_batch_ndims=len(event_shape),
    )

    return distrax.Transformed(base_distribution, flow)
'''

# ---- src/real_nvp_toy_densities.py:112-136 ------------------------------------------------------------------------
REAL_NVP_TOY_112_136 = r'''

@hk.without_apply_rng
@hk.transform
def log_prob(data: Array) -> Array:

    model = make_flow_model(
        event_shape=data.shape[-1:],
        num_layers=FLOW_NUM_LAYERS,
        hidden_sizes=[HIDDEN_SIZE] * MLP_NUM_LAYERS,
    )

    return model.log_prob(data)


@hk.without_apply_rng
@hk.transform
def sample(event_shape: Tuple, prng_key: PRNGKey, num_samples: int):  # type: ignore

    model = make_flow_model(
        event_shape=event_shape,
        num_layers=FLOW_NUM_LAYERS,
        hidden_sizes=[HIDDEN_SIZE] * MLP_NUM_LAYERS,
    )

    return model.sample(seed=prng_key, sample_shape=(num_samples,))
'''

# ---- src/maf_toy_densities.py:55-111 ------------------------------------------------------------------------------
MAF_TOY_55_111 = r'''
def make_conditioner(
    event_shape: Sequence[int], hidden_sizes: Sequence[int], num_bijector_params: int
) -> hk.Sequential:

    return hk.Sequential(
        [
            hk.Flatten(preserve_dims=-1),
            hk.nets.MLP(hidden_sizes, activate_final=True),
            hk.Linear(
                np.prod(event_shape) * num_bijector_params,
                w_init=jnp.zeros,
                b_init=jnp.zeros,
            ),
            hk.Reshape(tuple(event_shape) + (num_bijector_params,), preserve_dims=-1),
        ]
    )


def make_flow_model(
    event_shape: Sequence[int],
    num_layers: int,
    hidden_sizes: Sequence[int],
    flow_num_params: int,
) -> distrax.Transformed:

    mask = jnp.arange(0, np.prod(event_shape)) % 2  # every second element is masked
    mask = jnp.reshape(mask, event_shape)
    mask = mask.astype(bool)

    flow_num_params = 3 * flow_num_params + 1

    def bijector_fn(params: Array):
        return distrax.RationalQuadraticSpline(params, range_min=-8.0, range_max=8.0)

    layers = []
    for _ in range(num_layers):
        layer = distrax.MaskedCoupling(
            mask=mask,
            bijector=bijector_fn,
            conditioner=make_conditioner(
                event_shape=event_shape,
                hidden_sizes=hidden_sizes,
                num_bijector_params=flow_num_params,
            ),
        )

        layers.append(layer)
        # Flip mask after each layer
        mask = jnp.logical_not(mask)

    flow = distrax.Inverse(distrax.Chain(layers))
    base_distribution = distrax.Independent(
        distrax.Normal(loc=jnp.zeros(event_shape), scale=jnp.ones(event_shape)),
        reinterpreted_batch_ndims=len(event_shape),
    )

    return distrax.Transformed(base_distribution, flow)
'''

# ---- src/bflow.py:22-34 -------------------------------------------------------------------------------------------
BFLOW_22_34 = r'''
a = jnp.array([2.0], dtype=np.float32)
b = jnp.array([-0.1], dtype=np.float32)
theta = jnp.array([-2.5, 2, 3, -7, -7, -7, -7, -7, 7])
alpha = jnp.array([0.2], dtype=np.float32)
beta = jnp.array([-2.50], dtype=np.float32)

T1 = distrax.Chain([tfb.Scale(softplus(a)), tfb.Shift(b)])
T2 = tfb.SoftClip(low=0, high=1, hinge_softness=1.5)
T3 = BernsteinBijector(thetas=constrain_thetas(theta))
T4 = distrax.Chain([tfb.Scale(softplus(alpha)), tfb.Shift(beta)])
chained_bijectors = [T4, T3, T2, T1]

bij = distrax.Inverse(distrax.Chain(chained_bijectors))
'''


def mirror_namespace():
    """What the changed import lines bind."""
    from normalizing_flows_b200 import distrax_api as distrax
    from normalizing_flows_b200.BernsteinBijector import BernsteinBijector, constrain_thetas
    from normalizing_flows_b200.compat import hk, jax, jnp, softplus, tfp

    return dict(distrax=distrax, hk=hk, jax=jax, jnp=jnp, np=np, tfp=tfp, tfd=tfp.distributions, tfb=tfp.bijectors,
                softplus=softplus, BernsteinBijector=BernsteinBijector, constrain_thetas=constrain_thetas,
                Sequence=Sequence, Tuple=Tuple, Any=Any, Mapping=Mapping, Array=torch.Tensor, PRNGKey=int,
                FLOW_NUM_LAYERS=8, HIDDEN_SIZE=64, MLP_NUM_LAYERS=2, FLOW_NUM_PARAMS=12)


def test_mirror_builds_the_reference_parameter_tree():
    """CPU: hk.transform(...).init creates exactly the haiku tree of the reference (names, shapes, zero-initialised final
    Linear) — the same assertion the golden generator makes about the reference itself."""
    ns = mirror_namespace()
    exec(REAL_NVP_TOY_56_110, ns)
    exec(REAL_NVP_TOY_112_136, ns)
    params = ns["log_prob"].init(ns["jax"].random.PRNGKey(0), np.zeros((1, 2), dtype=np.float32))
    g = np.load(os.path.join(GOLDEN, "coupling_toy_golden.npz"))
    assert sorted(params) == sorted(str(s) for s in g["rnvp_names"])
    assert tuple(params["mlp_3/~/linear_1"]["w"].shape) == (64, 64) and tuple(params["linear_7"]["w"].shape) == (64, 2)
    assert not params["linear"]["w"].any() and params["mlp/~/linear_0"]["w"].abs().sum() > 0


def tree_from_flat(flat, out):
    from normalizing_flows_b200 import params_io

    return {m: {k: torch.from_numpy(v).cuda() for k, v in d.items()}
            for m, d in params_io.flat_to_haiku(torch.from_numpy(np.asarray(flat)), 8, 2, [64, 64], out).items()}


@pytest.mark.gpu
@pytest.mark.parametrize("script,key,out", [("rnvp", "rnvp", 2), ("maf", "maf", 74)])
def test_reference_make_flow_model_runs_on_the_mirror(script, key, out):
    from parity_util import hold

    ns = mirror_namespace()
    exec(REAL_NVP_TOY_56_110 if script == "rnvp" else MAF_TOY_55_111, ns)
    if script == "maf":  # maf_toy_densities.py:114-139 differ from the Real NVP wrappers only by flow_num_params
        wrappers = REAL_NVP_TOY_112_136.replace("hidden_sizes=[HIDDEN_SIZE] * MLP_NUM_LAYERS,",
                                                "hidden_sizes=[HIDDEN_SIZE] * MLP_NUM_LAYERS,\n        flow_num_params=FLOW_NUM_PARAMS,")
    else:
        wrappers = REAL_NVP_TOY_112_136
    exec(wrappers, ns)
    g = np.load(os.path.join(GOLDEN, "coupling_toy_golden.npz"))
    params = tree_from_flat(g[f"{key}_flat"], out)
    x = torch.from_numpy(g[f"{key}_x"]).cuda()
    lp = ns["log_prob"].apply(params, x)
    hold(f"dropin_{key}_log_prob", lp, g[f"{key}_log_prob"], 1e-5 if key == "rnvp" else 1.5e-4,
         kind="toy" if key == "rnvp" else "stress")
    # the model the reference's code built is the fused one: a Chain of MaskedCouplings collapsed into one kernel
    model = None

    def grab(data):
        nonlocal model
        model = ns["make_flow_model"](event_shape=data.shape[-1:], num_layers=8, hidden_sizes=[64, 64],
                                      **({"flow_num_params": 12} if script == "maf" else {}))
        return model.log_prob(data)

    ns["hk"].without_apply_rng(ns["hk"].transform(grab)).apply(params, x)
    assert model._stack is not None and len(model._stack.layers) == 8
    # in-place update of the ONE parameter buffer is seen by the next call (single source of truth)
    before = model.log_prob(x).clone()
    model.params.mul_(0.5)
    assert (model.log_prob(x) - before).abs().max().item() > 1e-3
    # sampling: Transformed.sample(seed=, sample_shape=) through the reference's wrapper
    s = ns["sample"].apply(params, (2,), 3, 16)
    assert tuple(s.shape) == (16, 2) and torch.isfinite(s).all()


@pytest.mark.gpu
def test_reference_bflow_chain_runs_on_the_mirror():
    from parity_util import hold

    ns = mirror_namespace()
    exec(BFLOW_22_34, ns)
    distrax = ns["distrax"]
    g = np.load(os.path.join(GOLDEN, "bflow_golden.npz"))
    # bflow.py:82-88
    mu, sigma = 0.0, 1.0
    dist_base = distrax.Normal(loc=mu, scale=sigma)
    bij = distrax.Inverse(distrax.Chain(ns["chained_bijectors"]))
    dist_transformed = distrax.Transformed(dist_base, bij)
    x1 = np.linspace(-3, 3, 200)
    Y = dist_base.log_prob(x1)
    Z = dist_transformed.log_prob(x1)
    assert dist_transformed._bflow is not None, "the bflow chain should dispatch the fused kernel"
    hold("dropin_bflow_log_prob_fused", Z, g["bflow_log_prob"], 3e-5, kind="toy")
    hold("dropin_normal_log_prob", Y, -0.5 * x1 ** 2 - 0.5 * np.log(2 * np.pi), 1e-6, kind="toy")
    # the same through the bijector-by-bijector composition (elementwise kernels)
    dist_transformed._bflow = None
    hold("dropin_bflow_log_prob_composed", dist_transformed.log_prob(x1), g["bflow_log_prob"], 3e-5, kind="toy")
    # bflow.py:133-142 on the golden grid: MultivariateNormalDiag base, Block(..., ndims=1)
    jnp = ns["jnp"]
    dist_base2 = distrax.MultivariateNormalDiag(loc=jnp.zeros(2), scale_diag=jnp.ones(2))
    bij2 = distrax.Block(distrax.Inverse(distrax.Chain(ns["chained_bijectors"])), ndims=1)
    Z2 = distrax.Transformed(dist_base2, bij2).log_prob(g["bflow_X"])
    hold("dropin_bflow_mvn_log_prob", Z2, g["bflow_mvn_log_prob"], 5e-5, kind="toy")
    # each bijector of the chain on its own
    for nm, bj in zip(("T4", "T3", "T2", "T1"), ns["chained_bijectors"]):
        y, ld = bj.forward_and_log_det(g[f"bflow_{nm}_x"])
        hold(f"dropin_bflow_{nm}_y", y, g[f"bflow_{nm}_y"], 2e-5, kind="toy")
        hold(f"dropin_bflow_{nm}_ld", ld, g[f"bflow_{nm}_ld"], 2e-5, kind="toy")
    xr, _ = ns["T2"].inverse_and_log_det(ns["T2"].forward(np.linspace(-4, 4, 101)))
    hold("dropin_softclip_round_trip", xr, np.linspace(-4, 4, 101), 2e-4, kind="toy")


@pytest.mark.gpu
def test_elementwise_bijector_classes_vs_oracle():
    """distrax.ScalarAffine / RationalQuadraticSpline as classes, called the way a bijector_fn's result is called."""
    from normalizing_flows_b200 import distrax_api as distrax
    from oracle import flows_oracle as O
    from parity_util import hold

    gen = torch.Generator().manual_seed(0)
    x = torch.randn(300, 2, generator=gen, dtype=torch.float64) * 2
    p = torch.randn(300, 2, generator=gen, dtype=torch.float64) * 0.5
    b = distrax.scalar_affine_fn(p.float())  # shift / log_scale [N,1] broadcast over D
    y, ld = b.forward_and_log_det(x.float())
    hold("class_scalar_affine_y", y, x * torch.exp(p[:, 1:2]) + p[:, :1], 1e-5)
    hold("class_scalar_affine_ld", ld, p[:, 1:2].expand(300, 2), 1e-6)
    xi, _ = b.inverse_and_log_det(y)
    hold("class_scalar_affine_round_trip", xi, x, 1e-5)
    q = torch.randn(300, 2, 37, generator=gen, dtype=torch.float64) * 0.5
    r = distrax.RationalQuadraticSpline(q.float(), range_min=-8.0, range_max=8.0)
    y, ld = r.forward_and_log_det(x.float())
    yr, lr = O.rqs_forward(q.float().double(), x.float().double(), -8.0, 8.0)
    hold("class_rqs_y", y, yr, 2e-5)
    hold("class_rqs_ld", ld, lr, 5e-5)
    xr, _ = r.inverse_and_log_det(y)
    hold("class_rqs_round_trip", xr, x.float().double(), 5e-5)
    with pytest.raises(ValueError):
        distrax.RationalQuadraticSpline(torch.zeros(4, 36), -1.0, 1.0)
    with pytest.raises(ValueError):
        distrax.RationalQuadraticSpline(torch.zeros(4, 37), 1.0, -1.0)
    with pytest.raises(ValueError):
        distrax.ScalarAffine(shift=0.0, scale=1.0, log_scale=0.0)
    u = distrax.Uniform(low=torch.zeros(3), high=torch.ones(3) * 2)
    lp = distrax.Independent(u, 1).log_prob(torch.tensor([[0.5, 1.5, 0.1], [0.5, 2.5, 0.1]]))
    assert abs(lp[0].item() + 3 * np.log(2.0)) < 1e-6 and lp[1].item() == -np.inf
