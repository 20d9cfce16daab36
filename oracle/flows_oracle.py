"""CPU restatement of the reference's flow hot path (torch-CPU, any float dtype).

TEST INFRASTRUCTURE — never imported by ``normalizing_flows_b200`` (the product).
float64 = parity truth; float32 with all host threads = the timed CPU baseline
("CPU restatement of the reference; JAX is not installable on either box").

Pinning status:
  * rows #1 (hand-rolled Real NVP) and #5 (Bernstein): PINNED against
    ``tests/golden/*.npz`` — outputs of the reference's own source files executed
    in this container with numpy standing in for ``jax.numpy``
    (``tests/golden/make_golden.py``; tests/test_oracle_golden.py).
  * rows #2-#4 and the bflow chain (§8f-1): WIRING PINNED, PRIMITIVES RESTATED.
    ``tests/golden/make_golden_flows.py`` imports the reference's own scripts
    (``src/real_nvp_toy_densities.py``, ``src/maf_toy_densities.py``, ``src/maf_mnist.py``) and
    executes ``src/bflow.py`` lines 1-34 UNMODIFIED, with the numpy stand-ins of
    ``tests/golden/ref_shim.py`` in place of jax / haiku / distrax / TFP, and commits the
    outputs (``coupling_toy_golden.npz``, ``coupling_mnist_golden.npz``, ``bflow_golden.npz``).
    Everything the reference wrote is therefore pinned — ``make_conditioner`` /
    ``make_flow_model``, the ``arange % 2`` masks and the per-layer ``logical_not``, ``3*n+1``,
    Flatten / Reshape, the ``bijector_fn`` closures, ``Inverse(Chain(layers))``, the base
    distributions, the haiku module names and creation order — and this oracle is held to those
    vectors at 1e-9 ... 1e-12 (tests/test_oracle_golden.py).  What remains restated from the
    libraries' published definitions, here AND in the shim, are the
    library primitives: the rational-quadratic spline, ScalarAffine, Linear / MLP, SoftClip,
    tfd.Beta, the Normal / Uniform densities.  They are additionally held by the
    known-answer properties K1-K4 of SURVEY.md §4.
  * row #9 (optax.adam): restated from Kingma & Ba / optax's documented update; SGD is
    ``real_nvp.py:298`` verbatim.  UNPINNED (optax is absent).
  * row #8 (MADE-affine MAF): PARITY UNPINNED — no reference code exists.

All ``file:line`` citations are into /root/reference/.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Sequence, Tuple

import torch

Tensor = torch.Tensor
LOG_2PI = math.log(2.0 * math.pi)

# ----------------------------------------------------------------------------
# Row #1 — hand-rolled Real NVP  (src/real_nvp.py)
# ----------------------------------------------------------------------------


def leaky_relu(x: Tensor) -> Tensor:
    """jax.nn.leaky_relu, negative_slope=0.01 (src/real_nvp.py:167)."""
    return torch.where(x >= 0, x, 0.01 * x)


def shift_and_log_scale_fn(net_params, x: Tensor) -> Tuple[Tensor, Tensor]:
    """src/real_nvp.py:161-171."""
    act = x
    for W, b in net_params[:-1]:
        act = leaky_relu(act @ W + b)
    W, b = net_params[-1]
    out = act @ W + b
    half = out.shape[1] // 2
    return out[:, :half], out[:, half:]


def forward_nvp(net_params, x: Tensor, flip: bool = False) -> Tensor:
    """src/real_nvp.py:71-101."""
    d = x.shape[-1] // 2
    x1, x2 = x[:, :d], x[:, d:]
    if flip:
        x2, x1 = x1, x2
    shift, log_scale = shift_and_log_scale_fn(net_params, x1)
    y1 = x1
    y2 = x2 * torch.exp(log_scale) + shift
    if flip:
        y1, y2 = y2, y1
    return torch.cat([y1, y2], dim=-1)


def inverse_nvp(net_params, y: Tensor, flip: bool = False) -> Tuple[Tensor, Tensor]:
    """src/real_nvp.py:104-134.  Returns (x, log_scale)."""
    d = y.shape[-1] // 2
    y1, y2 = y[:, :d], y[:, d:]
    if flip:
        y1, y2 = y2, y1
    shift, log_scale = shift_and_log_scale_fn(net_params, y1)
    x1 = y1
    x2 = (y2 - shift) * torch.exp(-log_scale)
    if flip:
        x1, x2 = x2, x1
    return torch.cat([x1, x2], dim=-1), log_scale


def log_prob_N01(x: Tensor) -> Tensor:
    """tfd.MultivariateNormalDiag(0, I).log_prob (src/real_nvp.py:57-68)."""
    return -0.5 * torch.sum(x * x, dim=-1) - 0.5 * x.shape[-1] * LOG_2PI


def log_prob_nvp(net_params, base_log_prob_fn, y: Tensor, flip: bool = False) -> Tensor:
    """src/real_nvp.py:137-158."""
    x, log_scale = inverse_nvp(net_params, y, flip)
    return base_log_prob_fn(x) - torch.sum(log_scale, dim=-1)


def log_prob_nvp_chains(params, flips: Sequence[bool], y: Tensor, base_log_prob_fn=log_prob_N01) -> Tensor:
    """src/real_nvp.py:251-283: closures are folded so the LAST layer's inverse
    is applied to the data first."""
    fn = base_log_prob_fn
    for net_params, flip in zip(params, flips):
        fn = (lambda p, f, inner: (lambda v: log_prob_nvp(p, inner, v, flip=f)))(net_params, flip, fn)
    return fn(y)


def sample_nvp_chain(params, flips: Sequence[bool], z: Tensor) -> Tensor:
    """src/real_nvp.py:228-248 with the base noise ``z`` given (JAX threefry
    streams cannot be reproduced here; parity is same-z-in → same-x-out)."""
    for net_params, flip in zip(params, flips):
        z = forward_nvp(net_params, z, flip=flip)
    return z


def init_nvp_chain_like(n: int, hidden: int = 4, seed: int = 534, dtype=torch.float64, scale: float = 0.1):
    """Shape-compatible stand-in for init_nvp_chain (src/real_nvp.py:174-220):
    layer MLP [1, hidden, 2], W ~ 0.1*N(0,1), b = 0, flips alternate from False.
    (The reference re-seeds PRNGKey(534) per layer so all layers start equal; we
    draw distinct weights so every layer is exercised.)"""
    g = torch.Generator().manual_seed(seed)
    params, flips, flip = [], [], False
    for _ in range(n):
        layer = []
        for m, k in ((1, hidden), (hidden, 2)):
            W = (torch.randn(m, k, generator=g, dtype=torch.float64) * scale).to(dtype)
            b = (torch.randn(k, generator=g, dtype=torch.float64) * scale).to(dtype)
            layer.append([W, b])
        params.append(layer)
        flips.append(flip)
        flip = not flip
    return params, flips


# ----------------------------------------------------------------------------
# Row #5 — Bernstein bijector  (src/BernsteinBijector.py)
# ----------------------------------------------------------------------------


def softplus(x: Tensor) -> Tensor:
    return torch.nn.functional.softplus(x, beta=1.0, threshold=1e9)


def constrain_thetas(theta_unconstrained: Tensor) -> Tensor:
    """src/BernsteinBijector.py:60-71."""
    d = torch.cat(
        (
            torch.zeros_like(theta_unconstrained[..., :1]),
            theta_unconstrained[..., :1],
            softplus(theta_unconstrained[..., 1:]) + 1e-4,
        ),
        dim=-1,
    )
    return torch.cumsum(d[..., 1:], dim=-1)


def _beta_prob(alpha: Tensor, beta: Tensor, x: Tensor) -> Tensor:
    """tfd.Beta(alpha, beta).prob(x) = exp(xlogy(a-1,x) + xlog1py(b-1,-x) - lbeta)."""
    lbeta = torch.lgamma(alpha) + torch.lgamma(beta) - torch.lgamma(alpha + beta)
    return torch.exp(torch.xlogy(alpha - 1, x) + torch.special.xlog1py(beta - 1, -x) - lbeta)


def bernstein_poly(theta: Tensor, x: Tensor) -> Tensor:
    """get_bernstein_poly (src/BernsteinBijector.py:17-36): mean_k Beta(k, M-k+1).pdf(x)*theta_k."""
    order = theta.shape[-1]
    alpha = torch.arange(1, order + 1, dtype=theta.dtype)
    beta = torch.flip(alpha, dims=(0,))
    bx = _beta_prob(alpha, beta, x[..., None])
    return torch.mean(bx * theta, dim=-1)


def bernstein_poly_jac(theta: Tensor, x: Tensor) -> Tensor:
    """get_bernstein_poly_jac (src/BernsteinBijector.py:39-57)."""
    order = theta.shape[-1]
    alpha = torch.arange(1, order, dtype=theta.dtype)
    beta = torch.flip(alpha, dims=(0,))
    by = _beta_prob(alpha, beta, x[..., None])
    dtheta = theta[..., 1:] - theta[..., :-1]
    return torch.sum(by * dtheta, dim=-1)


def _bern_clip(x: Tensor) -> Tensor:
    """jnp.clip(x, 1e-7, 1-1e-7) in x's dtype (src/BernsteinBijector.py:84-85,91-92)."""
    lo = torch.tensor(1e-7, dtype=x.dtype)
    hi = torch.tensor(1.0, dtype=x.dtype) - torch.tensor(1e-7, dtype=x.dtype)
    return torch.clamp(x, lo, hi)


def bernstein_forward(thetas_constrained: Tensor, x: Tensor) -> Tensor:
    """BernsteinBijector.forward (src/BernsteinBijector.py:81-86)."""
    return bernstein_poly(thetas_constrained, _bern_clip(x))


def bernstein_forward_log_det_jacobian(thetas_constrained: Tensor, x: Tensor) -> Tensor:
    """src/BernsteinBijector.py:88-93."""
    return torch.log(bernstein_poly_jac(thetas_constrained, _bern_clip(x)))


def bernstein_forward_and_log_det(thetas_constrained: Tensor, x: Tensor):
    """src/BernsteinBijector.py:113-117."""
    return bernstein_forward(thetas_constrained, x), bernstein_forward_log_det_jacobian(thetas_constrained, x)


def _interp(xq: Tensor, xp: Tensor, fp: Tensor) -> Tensor:
    """jnp.interp / np.interp for increasing xp: linear inside, clamped outside."""
    n = xp.shape[0]
    idx = torch.searchsorted(xp.contiguous(), xq.contiguous(), right=True)
    i1 = idx.clamp(1, n - 1)
    i0 = i1 - 1
    x0, x1, f0, f1 = xp[i0], xp[i1], fp[i0], fp[i1]
    dx = x1 - x0
    t = torch.where(dx > 0, (xq - x0) / torch.where(dx > 0, dx, torch.ones_like(dx)), torch.zeros_like(dx))
    out = f0 + t * (f1 - f0)
    out = torch.where(xq <= xp[0], fp[0].expand_as(out), out)
    out = torch.where(xq >= xp[-1], fp[-1].expand_as(out), out)
    return out


def bernstein_inverse(thetas_constrained: Tensor, y: Tensor, n_points: int = 200) -> Tensor:
    """BernsteinBijector.inverse (src/BernsteinBijector.py:95-111).

    thetas [E, M] (or [M] == E=1 with the event axis implied); y [N, E].
    jax.vmap(inp, in_axes=-1) maps the LAST axis and stacks on axis 0, so the
    result is [E, N] (transposed w.r.t. y) — reference behaviour, preserved."""
    th = thetas_constrained if thetas_constrained.dim() == 2 else thetas_constrained[None, :]
    E = th.shape[0]
    assert y.dim() == 2 and y.shape[1] == E
    clip = 1e-7
    x_fit = torch.linspace(clip, 1 - clip, n_points, dtype=torch.float64).to(y.dtype)[:, None] * torch.ones(
        (1, E), dtype=y.dtype
    )
    y_fit = bernstein_forward(th, x_fit)  # [n_points, E]
    cols = [_interp(y[:, e], y_fit[:, e], x_fit[:, e]) for e in range(E)]
    return torch.stack(cols, dim=0)


# ---- bflow.py chain: Scale∘Shift → SoftClip → Bernstein → Scale∘Shift ----


def softclip_forward_and_log_det(x: Tensor, low: float = 0.0, high: float = 1.0, c: float = 1.5):
    """tfb.SoftClip(low, high, hinge_softness=c) (src/bflow.py:29) [unpinned]:
    y = high - k*sp_c(high - low - sp_c(x - low)),  k = (high-low)/sp_c(high-low),
    sp_c(u) = c*softplus(u/c)."""
    def sp(u):
        return c * softplus(u / c)

    r = torch.tensor(high - low, dtype=x.dtype)
    k = r / sp(r)
    inner = sp(x - low)
    u = r - inner
    y = high - k * sp(u)
    # dy/dx = k * sigmoid(u/c) * sigmoid((x-low)/c)
    ld = torch.log(k) + torch.nn.functional.logsigmoid(u / c) + torch.nn.functional.logsigmoid((x - low) / c)
    return y, ld


def bernstein_flow_log_prob(a, b, theta_raw, alpha, beta, x: Tensor, double_constrain: bool = True) -> Tensor:
    """Transformed(Normal(0,1), Inverse(Chain([T4,T3,T2,T1]))).log_prob(x) of
    src/bflow.py:28-34,82-88.  Chain.forward applies T1 first … T4 last;
    distrax.Chain([Scale(s), Shift(c)]) applies Shift then Scale: y = s*(x + c).
    bflow.py:30 constrains theta before BernsteinBijector.__init__ constrains it
    again (``double_constrain``; reference quirk)."""
    dt = x.dtype
    a, b, alpha, beta = (torch.as_tensor(v, dtype=dt) for v in (a, b, alpha, beta))
    th = constrain_thetas(torch.as_tensor(theta_raw, dtype=dt))
    if double_constrain:
        th = constrain_thetas(th)
    s1 = softplus(a)
    v = s1 * (x + b)
    ld = torch.log(s1).expand_as(x).clone()
    v, l2 = softclip_forward_and_log_det(v)
    ld = ld + l2
    l3 = bernstein_forward_log_det_jacobian(th, v)
    v = bernstein_forward(th, v)
    ld = ld + l3
    s4 = softplus(alpha)
    v = s4 * (v + beta)
    ld = ld + torch.log(s4)
    return -0.5 * v * v - 0.5 * LOG_2PI + ld


# ----------------------------------------------------------------------------
# Rows #2-#4, #6 — distrax MaskedCoupling flows  [semantics from memory: UNPINNED]
# ----------------------------------------------------------------------------


@dataclass
class CouplingSpec:
    """Workload definition of the three distrax scripts.

    transform: 'affine_scalar' = real_nvp_toy_densities.py:84-98 (2 conditioner outputs,
               shift/log_scale [N,1] broadcast over D); 'affine' = per-dimension
               (shift, log_scale); 'rqs' = maf_toy_densities.py:84-98 /
               maf_mnist.py:63-77 (3K+1 params per dimension).
    base: 'normal' (Independent Normal(0,1)) or 'uniform' (Independent Uniform(0,1))."""

    event_dim: int = 2
    num_layers: int = 8
    hidden: List[int] = field(default_factory=lambda: [64, 64])
    transform: str = "affine_scalar"
    num_bins: int = 12
    range_min: float = -8.0
    range_max: float = 8.0
    base: str = "normal"

    @property
    def params_per_dim(self) -> int:
        return {"affine_scalar": 2, "affine": 2, "rqs": 3 * self.num_bins + 1}[self.transform]

    @property
    def cond_out(self) -> int:
        if self.transform == "affine_scalar":
            return 2
        return self.event_dim * self.params_per_dim

    def layer_shapes(self):
        sizes = [self.event_dim] + list(self.hidden) + [self.cond_out]
        return [(m, n) for m, n in zip(sizes[:-1], sizes[1:])]

    def layer_mask(self, layer: int) -> Tensor:
        """mask = arange(D) % 2 (True = pass-through & visible to the conditioner),
        logical_not after each layer (real_nvp_toy_densities.py:80-82,102)."""
        m = (torch.arange(self.event_dim) % 2).bool()
        return ~m if (layer % 2) else m


REAL_NVP_TOY = CouplingSpec(2, 8, [64, 64], "affine_scalar")  # src/real_nvp_toy_densities.py:34-36
RQS_TOY = CouplingSpec(2, 8, [64, 64], "rqs", 12, -8.0, 8.0, "normal")  # src/maf_toy_densities.py:32-35,87
RQS_MNIST = CouplingSpec(784, 8, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform")  # src/maf_mnist.py:28-31,66


def init_coupling_params(spec: CouplingSpec, seed: int = 0, dtype=torch.float64, scale: float = 0.1, final_scale=None):
    """Random params ~ N(0, scale^2) in haiku order: per flow layer,
    [(W[in,out], b[out]) for each Linear].  ``final_scale=0`` reproduces the
    reference's zero-initialised last Linear (identity flow, K1/K2)."""
    g = torch.Generator().manual_seed(seed)
    out = []
    shapes = spec.layer_shapes()
    for _ in range(spec.num_layers):
        layer = []
        for i, (m, n) in enumerate(shapes):
            s = scale / math.sqrt(max(m, 1) / 2.0) if m > 8 else scale
            if final_scale is not None and i == len(shapes) - 1:
                s = final_scale
            W = (torch.randn(m, n, generator=g, dtype=torch.float64) * s).to(dtype)
            b = (torch.randn(n, generator=g, dtype=torch.float64) * (s if s == 0 else scale)).to(dtype)
            layer.append((W, b))
        out.append(layer)
    return out


def mlp_conditioner(layer_params, x: Tensor) -> Tensor:
    """hk.Sequential[Flatten, nets.MLP(hidden, activate_final=True) (relu), Linear]
    (real_nvp_toy_densities.py:56-71)."""
    act = x
    for W, b in layer_params[:-1]:
        act = torch.relu(act @ W + b)
    W, b = layer_params[-1]
    return act @ W + b


RQS_MIN_BIN = 1e-4
RQS_MIN_SLOPE = 1e-4


def rqs_knots(params: Tensor, range_min: float, range_max: float):
    """distrax.RationalQuadraticSpline.__init__: params [..., 3K+1] →
    x_pos, y_pos, knot_slopes each [..., K+1]."""
    K = (params.shape[-1] - 1) // 3
    uw, uh, us = params[..., :K], params[..., K : 2 * K], params[..., 2 * K :]
    R = range_max - range_min

    def norm_bins(u):
        return torch.softmax(u, dim=-1) * (R - K * RQS_MIN_BIN) + RQS_MIN_BIN

    w, h = norm_bins(uw), norm_bins(uh)
    lo = torch.full(params.shape[:-1] + (1,), range_min, dtype=params.dtype)
    hi = torch.full(params.shape[:-1] + (1,), range_max, dtype=params.dtype)
    x_pos = torch.cat([lo, range_min + torch.cumsum(w[..., :-1], -1), hi], -1)
    y_pos = torch.cat([lo, range_min + torch.cumsum(h[..., :-1], -1), hi], -1)
    offset = math.log(math.exp(1.0 - RQS_MIN_SLOPE) - 1.0)
    slopes = softplus(us + offset) + RQS_MIN_SLOPE
    return x_pos, y_pos, slopes


def _select_bin(v: Tensor, pos: Tensor):
    """correct_bin = (v >= pos[:-1]) & (v < pos[1:]); first bin if none."""
    inb = (v[..., None] >= pos[..., :-1]) & (v[..., None] < pos[..., 1:])
    anyb = inb.any(-1, keepdim=True)
    first = torch.zeros_like(inb)
    first[..., 0] = True
    return torch.where(anyb, inb, first)


def _gather_bin(sel: Tensor, arr: Tensor):
    s = sel.to(arr.dtype)
    return (s * arr[..., :-1]).sum(-1), (s * arr[..., 1:]).sum(-1)


def rqs_forward(params: Tensor, x: Tensor, range_min: float, range_max: float):
    """_rational_quadratic_spline_fwd."""
    x_pos, y_pos, sl = rqs_knots(params, range_min, range_max)
    below, above = x <= x_pos[..., 0], x >= x_pos[..., -1]
    sel = _select_bin(x, x_pos)
    xl, xr = _gather_bin(sel, x_pos)
    yl, yr = _gather_bin(sel, y_pos)
    sl_l, sl_r = _gather_bin(sel, sl)
    w, h = xr - xl, yr - yl
    s = h / w
    z = torch.clamp((x - xl) / w, 0.0, 1.0)
    sq_z = z * z
    z1mz = z - sq_z
    sq_1mz = (1.0 - z) ** 2
    t = sl_r + sl_l - 2.0 * s
    num = h * (s * sq_z + sl_l * z1mz)
    den = s + t * z1mz
    y = yl + num / den
    ld = 2.0 * torch.log(s) + torch.log(sl_r * sq_z + 2.0 * s * z1mz + sl_l * sq_1mz) - 2.0 * torch.log(den)
    y = torch.where(below, (x - x_pos[..., 0]) * sl[..., 0] + y_pos[..., 0], y)
    y = torch.where(above, (x - x_pos[..., -1]) * sl[..., -1] + y_pos[..., -1], y)
    ld = torch.where(below, torch.log(sl[..., 0]), ld)
    ld = torch.where(above, torch.log(sl[..., -1]), ld)
    return y, ld


def rqs_inverse(params: Tensor, y: Tensor, range_min: float, range_max: float):
    """_rational_quadratic_spline_inv."""
    x_pos, y_pos, sl = rqs_knots(params, range_min, range_max)
    below, above = y <= y_pos[..., 0], y >= y_pos[..., -1]
    sel = _select_bin(y, y_pos)
    xl, xr = _gather_bin(sel, x_pos)
    yl, yr = _gather_bin(sel, y_pos)
    sl_l, sl_r = _gather_bin(sel, sl)
    w, h = xr - xl, yr - yl
    s = h / w
    om = torch.clamp((y - yl) / h, 0.0, 1.0)
    t = sl_r + sl_l - 2.0 * s
    c = -s * om
    b = sl_l - t * om
    a = s - b
    z = torch.clamp(-2.0 * c / (b + torch.sqrt(b * b - 4.0 * a * c)), 0.0, 1.0)
    x = w * z + xl
    sq_z = z * z
    z1mz = z - sq_z
    sq_1mz = (1.0 - z) ** 2
    den = s + t * z1mz
    ld = -2.0 * torch.log(s) - torch.log(sl_r * sq_z + 2.0 * s * z1mz + sl_l * sq_1mz) + 2.0 * torch.log(den)
    x = torch.where(below, (y - y_pos[..., 0]) / sl[..., 0] + x_pos[..., 0], x)
    x = torch.where(above, (y - y_pos[..., -1]) / sl[..., -1] + x_pos[..., -1], x)
    ld = torch.where(below, -torch.log(sl[..., 0]), ld)
    ld = torch.where(above, -torch.log(sl[..., -1]), ld)
    return x, ld


def _inner_bijector(spec: CouplingSpec, cond_out: Tensor, v: Tensor, inverse: bool):
    """bijector_fn(params).forward_and_log_det / inverse_and_log_det on the FULL x
    (MaskedCoupling applies it everywhere, then masks)."""
    N, D = v.shape
    if spec.transform == "affine_scalar":
        shift, ls = cond_out[:, :1], cond_out[:, 1:2]  # jnp.split(params, 2, -1), broadcast over D
    elif spec.transform == "affine":
        p = cond_out.reshape(N, D, 2)
        shift, ls = p[..., 0], p[..., 1]
    else:
        p = cond_out.reshape(N, D, spec.params_per_dim)
        return (rqs_inverse if inverse else rqs_forward)(p, v, spec.range_min, spec.range_max)
    if inverse:  # ScalarAffine
        return (v - shift) * torch.exp(-ls), (-ls).expand(N, D)
    return v * torch.exp(ls) + shift, ls.expand(N, D)


def masked_coupling(spec: CouplingSpec, layer_params, layer: int, v: Tensor, inverse: bool):
    """distrax.MaskedCoupling.{forward,inverse}_and_log_det."""
    mask = spec.layer_mask(layer)
    mv = torch.where(mask, v, torch.zeros_like(v))
    cond = mlp_conditioner(layer_params, mv)
    out0, ld = _inner_bijector(spec, cond, v, inverse)
    out = torch.where(mask, v, out0)
    logdet = torch.where(mask, torch.zeros_like(ld), ld).sum(-1)
    return out, logdet


def base_log_prob(spec: CouplingSpec, x: Tensor) -> Tensor:
    """distrax.Independent(Normal(0,1)) / Independent(Uniform(0,1))."""
    if spec.base == "normal":
        return (-0.5 * x * x - 0.5 * LOG_2PI).sum(-1)
    inside = ((x >= 0.0) & (x <= 1.0)).all(-1)
    return torch.where(inside, torch.zeros(x.shape[0], dtype=x.dtype), torch.full((x.shape[0],), -math.inf, dtype=x.dtype))


def coupling_flow_to_base(spec: CouplingSpec, params, v: Tensor):
    """Inverse(Chain(layers)).inverse_and_log_det = Chain.forward_and_log_det:
    layers[-1] first … layers[0] last, each layer's FORWARD."""
    v = v.reshape(v.shape[0], -1)
    total = torch.zeros(v.shape[0], dtype=v.dtype)
    for l in reversed(range(spec.num_layers)):
        v, ld = masked_coupling(spec, params[l], l, v, inverse=False)
        total = total + ld
    return v, total


def coupling_flow_log_prob(spec: CouplingSpec, params, v: Tensor) -> Tensor:
    """distrax.Transformed.log_prob."""
    x, ld = coupling_flow_to_base(spec, params, v)
    return base_log_prob(spec, x) + ld


def coupling_flow_sample(spec: CouplingSpec, params, z: Tensor) -> Tensor:
    """Transformed.sample with base noise z given: Inverse.forward = Chain.inverse:
    layers[0].inverse first … layers[-1].inverse last."""
    v = z.reshape(z.shape[0], -1)
    for l in range(spec.num_layers):
        v, _ = masked_coupling(spec, params[l], l, v, inverse=True)
    return v


# ----------------------------------------------------------------------------
# Row #8 — MADE-affine MAF (north_star; no reference code — SURVEY A.7) [UNPINNED]
# ----------------------------------------------------------------------------


@dataclass
class MadeSpec:
    event_dim: int = 784
    num_layers: int = 8
    hidden: List[int] = field(default_factory=lambda: [512, 512])
    base: str = "normal"

    def degrees(self):
        D = self.event_dim
        deg_in = torch.arange(1, D + 1)
        degs = [deg_in]
        for H in self.hidden:
            k = torch.arange(1, H + 1, dtype=torch.float64)
            degs.append(torch.clamp(torch.ceil(k * (D - 1) / (H + 1)), min=1).long())
        return degs

    def masks(self):
        """[in,out] 0/1 masks: hidden deg_in <= deg_out; output deg_in < deg_out;
        output columns ordered [shift_0..shift_{D-1}, log_scale_0..log_scale_{D-1}]."""
        degs = self.degrees()
        ms = []
        for a, b in zip(degs[:-1], degs[1:]):
            ms.append((a[:, None] <= b[None, :]).to(torch.float64))
        out_deg = torch.cat([degs[0], degs[0]])
        ms.append((degs[-1][:, None] < out_deg[None, :]).to(torch.float64))
        return ms

    def layer_shapes(self):
        sizes = [self.event_dim] + list(self.hidden) + [2 * self.event_dim]
        return [(m, n) for m, n in zip(sizes[:-1], sizes[1:])]


def init_made_params(spec: MadeSpec, seed: int = 0, dtype=torch.float64, scale: float = 0.1):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(spec.num_layers):
        layer = []
        for m, n in spec.layer_shapes():
            s = scale / math.sqrt(max(m, 1) / 2.0) if m > 8 else scale
            W = (torch.randn(m, n, generator=g, dtype=torch.float64) * s).to(dtype)
            b = (torch.randn(n, generator=g, dtype=torch.float64) * scale * 0.1).to(dtype)
            layer.append((W, b))
        out.append(layer)
    return out


def made_conditioner(spec: MadeSpec, layer_params, x: Tensor):
    masks = [m.to(x.dtype) for m in spec.masks()]
    act = x
    for (W, b), m in zip(layer_params[:-1], masks[:-1]):
        act = torch.relu(act @ (W * m) + b)
    W, b = layer_params[-1]
    out = act @ (W * masks[-1]) + b
    D = spec.event_dim
    return out[:, :D], out[:, D:]


def _maf_perm(layer: int, v: Tensor) -> Tensor:
    """Reversal permutation between layers: odd layers see the reversed ordering."""
    return torch.flip(v, dims=(-1,)) if (layer % 2) else v


def maf_log_prob(spec: MadeSpec, params, x: Tensor) -> Tensor:
    """Density pass, one MADE evaluation per layer: u = (x - mu)*exp(-alpha),
    ildj = -sum(alpha).  Data → base applies layer L-1 first … layer 0 last."""
    v = x.reshape(x.shape[0], -1)
    total = torch.zeros(v.shape[0], dtype=v.dtype)
    for l in reversed(range(spec.num_layers)):
        vp = _maf_perm(l, v)
        mu, alpha = made_conditioner(spec, params[l], vp)
        u = (vp - mu) * torch.exp(-alpha)
        v = _maf_perm(l, u)
        total = total - alpha.sum(-1)
    if spec.base == "normal":
        return (-0.5 * v * v - 0.5 * LOG_2PI).sum(-1) + total
    raise ValueError(spec.base)


def maf_sample(spec: MadeSpec, params, z: Tensor) -> Tensor:
    """Sequential sampling: layer 0 first … layer L-1 last; within a layer D steps."""
    v = z.reshape(z.shape[0], -1)
    D = spec.event_dim
    for l in range(spec.num_layers):
        u = _maf_perm(l, v)
        x = torch.zeros_like(u)
        for i in range(D):
            mu, alpha = made_conditioner(spec, params[l], x)
            x = x.clone()
            x[:, i] = u[:, i] * torch.exp(alpha[:, i]) + mu[:, i]
        v = _maf_perm(l, x)
    return v


# ----------------------------------------------------------------------------
# Rows #7, #9 — loss, gradient, optimiser steps
# ----------------------------------------------------------------------------


def nll_and_grads(log_prob_fn, leaves: List[Tensor]):
    """-mean(log_prob) and d/d(leaves) by autograd (jax.value_and_grad analogue;
    src/real_nvp.py:292-297, src/real_nvp_toy_densities.py:139-151,170)."""
    leaves = [p.detach().clone().requires_grad_(True) for p in leaves]
    loss = -torch.mean(log_prob_fn(leaves))
    grads = torch.autograd.grad(loss, leaves, allow_unused=True)
    grads = [torch.zeros_like(p) if g is None else g for p, g in zip(leaves, grads)]
    return loss.detach(), grads


def sgd_step(p: Tensor, g: Tensor, lr: float) -> Tensor:
    """p - lr*g (src/real_nvp.py:298)."""
    return p - lr * g


def adam_step(p, g, m, v, step: int, lr: float, b1=0.9, b2=0.999, eps=1e-8, eps_root=0.0):
    """optax.adam (scale_by_adam then scale(-lr)) [from memory, UNPINNED]:
    m = b1 m + (1-b1) g; v = b2 v + (1-b2) g²; m̂ = m/(1-b1^t); v̂ = v/(1-b2^t);
    p ← p - lr · m̂ / (sqrt(v̂ + eps_root) + eps); t = step (1-based)."""
    m = b1 * m + (1 - b1) * g
    v = b2 * v + (1 - b2) * g * g
    mh = m / (1 - b1**step)
    vh = v / (1 - b2**step)
    return p - lr * mh / (torch.sqrt(vh + eps_root) + eps), m, v


# ----------------------------------------------------------------------------
# Synthetic inputs restated from src/densities/*.py
# ----------------------------------------------------------------------------


def flatten_layers(params) -> Tensor:
    """[(W,b), ...] per layer → one flat fp vector in (W row-major, b) order per Linear."""
    return torch.cat([t.reshape(-1) for layer in params for (W, b) in layer for t in (W, b)])
