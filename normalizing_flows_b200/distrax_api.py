"""Host-side mirror of the distrax / TFP surface the reference scripts use.

Bijectors: `Bijector`, `Chain`, `Inverse`, `Block`, `ScalarAffine`, `RationalQuadraticSpline`, `MaskedCoupling`,
`MaskedAutoregressiveFlow`, and (namespace `tfb`) `Scale`, `Shift`, `SoftClip`; distributions: `Normal`, `Uniform`,
`Independent`, `MultivariateNormalDiag`, `Transformed` — same names, argument meaning, composition rules
(`Chain.forward` applies the LAST bijector first; `Transformed.log_prob(v) = base.log_prob(inverse(v)) + ildj`) and error
behaviour as the libraries, every numeric step a libflowcuda kernel:

  * `Transformed(base, Inverse(Chain([MaskedCoupling...])))` — what every `make_flow_model` of the reference builds —
    collapses into ONE fused call (`flowcuda_coupling_log_prob` / `flowcuda_coupling_sample`).  `MaskedCoupling` takes the
    reference's own `bijector_fn` callables (real_nvp_toy_densities.py:84-86, maf_toy_densities.py:86-87): the callable
    is invoked once on probe parameters to learn which bijector it builds (and its range / bin count);
  * the bflow.py chain `Inverse(Chain([Scale∘Shift, Bernstein, SoftClip, Scale∘Shift]))` under a Normal base dispatches
    the fused `flowcuda_bernstein_flow_log_prob`;
  * anything else composes bijector by bijector through the elementwise kernels (csrc/elementwise.cu).

Inputs may be torch tensors, numpy arrays or Python numbers; results are float32 CUDA tensors.
"""
from __future__ import annotations

import math
import types as _types
from typing import Callable, Optional, Sequence, Tuple

import torch

from . import _lib, ops

Array = torch.Tensor


def _t(x) -> Array:
    """numpy array / number / tensor -> float32 torch tensor (device unchanged for tensors)."""
    if isinstance(x, torch.Tensor):
        return x.detach().to(torch.float32)
    return torch.as_tensor(x, dtype=torch.float32)


def _cuda(x) -> Array:
    return _t(x).to("cuda").contiguous()


def _add_ld(a, b):
    """Sum of two log-det terms on the device; plain numbers (the reference's `return x, 0`) stay numbers."""
    if not isinstance(a, torch.Tensor):
        a, b = b, a
    if not isinstance(a, torch.Tensor):
        return a + b
    if not isinstance(b, torch.Tensor):
        return a if b == 0 else ops.affine_elementwise(a.contiguous(), shift=torch.tensor(float(b)), want_ld=False)[0]
    if b.numel() > a.numel():
        a, b = b, a
    return ops.add_elementwise(a.contiguous(), b)


# ------------------------------------------------------------------------------- bijectors
class Bijector:
    """distrax.Bijector protocol (BernsteinBijector.py:74-79, rotation_flow.py:13-37)."""

    def __init__(self, event_ndims_in: int = 0, event_ndims_out: Optional[int] = None,
                 is_constant_jacobian: bool = False, is_constant_log_det: Optional[bool] = None):
        self.event_ndims_in = event_ndims_in
        self.event_ndims_out = event_ndims_in if event_ndims_out is None else event_ndims_out
        self.is_constant_jacobian = is_constant_jacobian
        self.is_constant_log_det = is_constant_jacobian if is_constant_log_det is None else is_constant_log_det

    def forward(self, x: Array) -> Array:
        return self.forward_and_log_det(x)[0]

    def inverse(self, y: Array) -> Array:
        return self.inverse_and_log_det(y)[0]

    def forward_log_det_jacobian(self, x: Array) -> Array:
        return self.forward_and_log_det(x)[1]

    def inverse_log_det_jacobian(self, y: Array) -> Array:
        return self.inverse_and_log_det(y)[1]

    def forward_and_log_det(self, x: Array) -> Tuple[Array, Array]:
        raise NotImplementedError(f"{type(self).__name__} does not implement forward_and_log_det")

    def inverse_and_log_det(self, y: Array) -> Tuple[Array, Array]:
        raise NotImplementedError(f"{type(self).__name__} does not implement inverse_and_log_det")


class Inverse(Bijector):
    def __init__(self, bijector: Bijector):
        super().__init__(bijector.event_ndims_out, bijector.event_ndims_in)
        self.bijector = bijector

    def forward_and_log_det(self, x):
        return self.bijector.inverse_and_log_det(x)

    def inverse_and_log_det(self, y):
        return self.bijector.forward_and_log_det(y)


class Chain(Bijector):
    """forward applies bijectors[-1] first ... bijectors[0] last (distrax semantics)."""

    def __init__(self, bijectors: Sequence[Bijector]):
        if not bijectors:
            raise ValueError("The sequence of bijectors cannot be empty.")
        self.bijectors = list(bijectors)
        super().__init__(self.bijectors[-1].event_ndims_in, self.bijectors[0].event_ndims_out)
        self._fused = _try_fuse(self.bijectors)

    def forward_and_log_det(self, x):
        if self._fused is not None:
            return self._fused.to_base(x)
        total = None
        for b in reversed(self.bijectors):
            x, ld = b.forward_and_log_det(x)
            total = ld if total is None else _add_ld(total, ld)
        return x, total

    def inverse_and_log_det(self, y):
        if self._fused is not None:
            return self._fused.to_data(y)
        total = None
        for b in self.bijectors:
            y, ld = b.inverse_and_log_det(y)
            total = ld if total is None else _add_ld(total, ld)
        return y, total


class Block(Bijector):
    """distrax.Block(b, ndims): sums b's log-det over the last `ndims` dims (bflow.py:134)."""

    def __init__(self, bijector: Bijector, ndims: int):
        super().__init__(bijector.event_ndims_in + ndims, bijector.event_ndims_out + ndims)
        self.bijector, self.ndims = bijector, ndims

    def _sum(self, ld):
        return ld if (self.ndims == 0 or not isinstance(ld, torch.Tensor)) else ops.sum_last(ld.contiguous(), self.ndims)

    def forward_and_log_det(self, x):
        y, ld = self.bijector.forward_and_log_det(x)
        return y, self._sum(ld)

    def inverse_and_log_det(self, y):
        x, ld = self.bijector.inverse_and_log_det(y)
        return x, self._sum(ld)


class ScalarAffine(Bijector):
    """distrax.ScalarAffine(shift, scale=None, log_scale=None): y = scale * x + shift elementwise, parameters broadcast
    against x (real_nvp_toy_densities.py:84-86 passes the [N,1] halves of the conditioner output)."""

    def __init__(self, shift, scale=None, log_scale=None):
        super().__init__(event_ndims_in=0, is_constant_jacobian=True)
        if scale is not None and log_scale is not None:
            raise ValueError("Only one of `scale` and `log_scale` can be specified, not both.")
        self._shift = _t(shift)
        self._scale = None if scale is None else _t(scale)
        self._log_scale = None if log_scale is None else _t(log_scale)

    @property
    def shift(self):
        return self._shift

    @property
    def log_scale(self):
        if self._log_scale is not None:
            return self._log_scale
        return torch.zeros_like(self._shift) if self._scale is None else torch.log(torch.abs(self._scale))

    @property
    def scale(self):
        if self._scale is not None:
            return self._scale
        return torch.ones_like(self._shift) if self._log_scale is None else torch.exp(self._log_scale)

    def _run(self, v, inverse):
        return ops.affine_elementwise(_cuda(v), shift=self._shift, scale=self._scale, log_scale=self._log_scale, inverse=inverse)

    def forward_and_log_det(self, x):
        return self._run(x, False)

    def inverse_and_log_det(self, y):
        return self._run(y, True)


class RationalQuadraticSpline(Bijector):
    """distrax.RationalQuadraticSpline(params[..., 3*num_bins+1], range_min, range_max) with the library's defaults
    (unconstrained boundary slopes, min_bin_size = min_knot_slope = 1e-4) — maf_toy_densities.py:86-87, maf_mnist.py:65-66."""

    def __init__(self, params, range_min: float, range_max: float, boundary_slopes: str = "unconstrained",
                 min_bin_size: float = 1e-4, min_knot_slope: float = 1e-4):
        super().__init__(event_ndims_in=0)
        params = _t(params)
        if params.shape[-1] % 3 != 1 or params.shape[-1] < 4:
            raise ValueError("The last dimension of `params` must have size `3 * num_bins + 1` and `num_bins` must be at "
                             f"least 1. Got size {params.shape[-1]}.")
        if range_min >= range_max:
            raise ValueError(f"`range_min` must be less than `range_max`. Got `range_min={range_min}` and "
                             f"`range_max={range_max}`.")
        if boundary_slopes != "unconstrained":
            raise NotImplementedError("the kernels implement boundary_slopes='unconstrained' (the library default, used by "
                                      "every reference script)")
        if min_bin_size != 1e-4 or min_knot_slope != 1e-4:
            raise NotImplementedError("the kernels implement min_bin_size = min_knot_slope = 1e-4 (the library defaults)")
        self.params = params
        self.num_bins = (params.shape[-1] - 1) // 3
        self.range_min, self.range_max = float(range_min), float(range_max)

    def forward_and_log_det(self, x):
        return ops.rqs_elementwise(_cuda(x), self.params, self.range_min, self.range_max, inverse=False)

    def inverse_and_log_det(self, y):
        return ops.rqs_elementwise(_cuda(y), self.params, self.range_min, self.range_max, inverse=True)


class Scale(ScalarAffine):
    """tfb.Scale(scale) (bflow.py:28,31)."""

    def __init__(self, scale):
        s = _t(scale)
        super().__init__(shift=torch.zeros_like(s), scale=s)


class Shift(ScalarAffine):
    """tfb.Shift(shift) (bflow.py:28,31)."""

    def __init__(self, shift):
        super().__init__(shift=shift)


class SoftClip(Bijector):
    """tfb.SoftClip(low, high, hinge_softness) (bflow.py:29)."""

    def __init__(self, low=None, high=None, hinge_softness=None):
        super().__init__(event_ndims_in=0)
        if low is None or high is None:
            raise NotImplementedError("SoftClip kernels implement the two-sided clip (bflow.py:29 passes low and high)")
        self.low, self.high = float(low), float(high)
        self.hinge_softness = 1.0 if hinge_softness is None else float(hinge_softness)

    def forward_and_log_det(self, x):
        return ops.softclip_elementwise(_cuda(x), self.low, self.high, self.hinge_softness, inverse=False)

    def inverse_and_log_det(self, y):
        return ops.softclip_elementwise(_cuda(y), self.low, self.high, self.hinge_softness, inverse=True)


tfb = _types.SimpleNamespace(Scale=Scale, Shift=Shift, SoftClip=SoftClip)  # `tfb = tfp.bijectors` of the reference scripts


# ---- the reference's `bijector_fn` closures, ready-made (and what MaskedCoupling's probe recognises) --------------------
def scalar_affine_fn(params):
    """real_nvp_toy_densities.py:84-86: `shift, log_scale = jnp.split(params, 2, axis=-1)` -> ScalarAffine."""
    shift, log_scale = torch.split(_t(params), 1, dim=-1)
    return ScalarAffine(shift=shift, log_scale=log_scale)


def affine_fn(params):
    """Per-dimension (shift, log_scale): conditioner output reshaped [.., D, 2]."""
    p = _t(params)
    return ScalarAffine(shift=p[..., 0], log_scale=p[..., 1])


def rational_quadratic_spline_fn(range_min: float, range_max: float, num_bins: Optional[int] = None) -> Callable:
    """`lambda params: distrax.RationalQuadraticSpline(params, range_min, range_max)`
    (maf_toy_densities.py:86-87, maf_mnist.py:65-66); params[..., 3*num_bins+1]."""

    def fn(params):
        return RationalQuadraticSpline(params, range_min=range_min, range_max=range_max)

    return fn


def _probe_bijector_fn(fn: Callable, event_dim: int, out_features: int):
    """Learn what a `bijector_fn(params) -> bijector` builds by calling it ONCE on probe parameters.  Returns
    (transform code, kwargs) for the fused kernels, or None when it builds something they do not fuse."""
    gen = torch.Generator().manual_seed(7)
    if out_features == 2:  # hk.Reshape((num_bijector_params,)): one parameter vector per sample
        probe = torch.randn(3, 2, generator=gen)
    elif event_dim > 0 and out_features % event_dim == 0:
        probe = torch.randn(3, event_dim, out_features // event_dim, generator=gen)
    else:
        return None
    try:
        b = fn(probe)
    except Exception:
        return None
    if isinstance(b, RationalQuadraticSpline) and b.params.shape == probe.shape and torch.equal(b.params, probe):
        return _lib.TRANSFORM_RQS, dict(range_min=b.range_min, range_max=b.range_max, num_bins=b.num_bins)
    if isinstance(b, ScalarAffine) and b._log_scale is not None and b._scale is None:
        sh, ls = b._shift, b._log_scale
        if probe.dim() == 2 and sh.shape == (3, 1) and torch.equal(sh, probe[:, :1]) and torch.equal(ls, probe[:, 1:2]):
            return _lib.TRANSFORM_AFFINE_SCALAR, {}  # jnp.split(params, 2, -1): one (shift, log_scale) per sample
        if probe.dim() == 3 and probe.shape[-1] == 2 and sh.shape == probe.shape[:-1] and torch.equal(sh, probe[..., 0]) \
                and torch.equal(ls, probe[..., 1]):
            return _lib.TRANSFORM_AFFINE, {}
    return None


class MLPConditioner:
    """hk.Sequential([Flatten, nets.MLP(hidden, activate_final=True), Linear(n_out), Reshape])
    (real_nvp_toy_densities.py:56-71).  `params` is a flat float32 CUDA tensor holding, per
    Linear, W[in,out] row-major then b[out] (haiku `w`, `b`)."""

    def __init__(self, event_shape: Sequence[int], hidden_sizes: Sequence[int], num_bijector_params: int,
                 params: Optional[Array] = None, out_features: Optional[int] = None, device="cuda", seed: int = 0):
        self.event_shape = tuple(int(s) for s in event_shape)
        self.event_dim = int(math.prod(self.event_shape))
        self.hidden_sizes = [int(h) for h in hidden_sizes]
        self.num_bijector_params = int(num_bijector_params)
        self.out_features = int(out_features if out_features is not None else num_bijector_params)
        sizes = [self.event_dim] + self.hidden_sizes + [self.out_features]
        self.shapes = list(zip(sizes[:-1], sizes[1:]))
        self.count = sum(m * n + n for m, n in self.shapes)
        if params is None:
            params = self.init(torch.Generator().manual_seed(seed)).to(device)
        if params.numel() != self.count:
            raise ValueError(f"MLPConditioner needs {self.count} parameters, got {params.numel()}")
        self.params = params

    def init(self, gen: torch.Generator) -> Array:
        """haiku defaults: hidden Linear w ~ TruncatedNormal(1/sqrt(fan_in)), b = 0; the final
        Linear is zero-initialised (real_nvp_toy_densities.py:64-68) => identity flow at init."""
        chunks = []
        for i, (m, n) in enumerate(self.shapes):
            if i == len(self.shapes) - 1:
                w = torch.zeros(m, n)
            else:
                w = torch.empty(m, n)
                torch.nn.init.trunc_normal_(w, std=1.0 / math.sqrt(m), a=-2.0 / math.sqrt(m), b=2.0 / math.sqrt(m), generator=gen)
            chunks += [w.reshape(-1), torch.zeros(n)]
        return torch.cat(chunks).float()

    def __call__(self, x):
        raise TypeError("MLPConditioner is evaluated inside the fused coupling kernels, not called from Python")


class MaskedCoupling(Bijector):
    """distrax.MaskedCoupling(mask, conditioner, bijector): mask True = passed through and visible to the
    conditioner; `bijector` is the reference's `bijector_fn(params) -> bijector` callable.

    Fused form (one kernel per call, and one for a whole Chain of them): an `MLPConditioner`, one of the alternating
    masks `arange(D) % 2` / its negation every reference script builds (real_nvp_toy_densities.py:80-82,102) and a
    `bijector_fn` that builds ScalarAffine from the split parameters or RationalQuadraticSpline.  Any other combination
    (a Python conditioner callable, another inner bijector, another mask) composes step by step: conditioner call,
    `bijector_fn(params)`, the inner bijector's own kernels, masking."""

    def __init__(self, mask: Array, conditioner, bijector: Callable, event_ndims: Optional[int] = None,
                 inner_event_ndims: int = 0):
        mask_t = torch.as_tensor(mask)
        if mask_t.dtype != torch.bool:
            raise ValueError(f"`mask` must have values of type `bool`; got values of type `{mask_t.dtype}`.")
        nd = mask_t.dim() if event_ndims is None else event_ndims
        if nd < mask_t.dim():
            raise ValueError(f"`event_ndims={nd}` should be at least as large as `mask.ndim={mask_t.dim()}`.")
        if inner_event_ndims > nd:
            raise ValueError(f"`inner_event_ndims={inner_event_ndims}` should be at most `event_ndims={nd}`.")
        super().__init__(nd, nd)
        if not callable(bijector):
            raise TypeError("`bijector` must be a callable params -> bijector")
        if hasattr(conditioner, "resolve_for_mask"):  # compat.hk.Sequential([...]): the reference's make_conditioner
            conditioner = conditioner.resolve_for_mask(mask_t)
        self.mask, self.conditioner, self.bijector = mask_t, conditioner, bijector
        self.inner_event_ndims = inner_event_ndims
        self.parity = None   # alternating-mask parity when the layer can run fused
        self.transform, self.targs = None, {}
        flat = mask_t.reshape(-1).cpu()
        D = flat.numel()
        if isinstance(conditioner, MLPConditioner):
            if conditioner.event_dim != D:
                raise ValueError("mask and conditioner disagree on the event size")
            odd = (torch.arange(D) % 2).bool()
            found = _probe_bijector_fn(bijector, D, conditioner.out_features)
            if found is not None and inner_event_ndims == 0 and (torch.equal(flat, odd) or torch.equal(flat, ~odd)):
                self.parity = 0 if torch.equal(flat, odd) else 1
                self.transform, self.targs = found
                ppd = 3 * self.targs["num_bins"] + 1 if self.transform == _lib.TRANSFORM_RQS else 2
                want = 2 if self.transform == _lib.TRANSFORM_AFFINE_SCALAR else D * ppd
                if conditioner.out_features != want:
                    raise ValueError(f"conditioner emits {conditioner.out_features} values; this bijector needs {want}")
            else:
                raise NotImplementedError(
                    "an MLPConditioner is evaluated inside the fused kernels, which need the alternating masks "
                    "arange(D) % 2 / its negation and a bijector_fn building ScalarAffine(split params) or "
                    "RationalQuadraticSpline(params, ...); pass a Python callable as conditioner for anything else")

    # ---- fused path -----------------------------------------------------------------------------------------------
    def _spec(self, num_layers: int, base: int = _lib.BASE_NORMAL, precision: int = _lib.PRECISION_FP32):
        c = self.conditioner
        return ops.make_spec(c.event_dim, num_layers, c.hidden_sizes, self.transform, self.targs.get("num_bins", 0),
                             self.targs.get("range_min", 0.0), self.targs.get("range_max", 1.0), base, precision)

    def _run_fused(self, v: Array, inverse: bool):
        # a one-layer stack of a spec whose layer index carries this mask's parity: for the odd parity the layer sits at
        # index 1 of a two-layer parameter buffer whose layer 0 is never touched (built once, refreshed when the
        # conditioner's parameters move or are rewritten in place)
        v = _cuda(v)
        batch = v.shape[: v.dim() - self.event_ndims_in]
        flat = v.reshape(-1, self.conditioner.event_dim) if batch else v.reshape(1, -1)
        spec = self._spec(2)
        p = self.conditioner.params
        if self.parity == 1:
            key = (p.data_ptr(), p._version)
            if getattr(self, "_pad_key", None) != key:
                buf = getattr(self, "_pad_buf", None)
                if buf is None or buf.numel() != 2 * p.numel() or buf.device != p.device:
                    buf = torch.zeros(2 * p.numel(), device=p.device, dtype=torch.float32)
                buf[p.numel():].copy_(p.reshape(-1))
                self._pad_buf, self._pad_key = buf, key
            p = self._pad_buf
        rng = (self.parity, self.parity + 1)
        if inverse:
            out, ld = ops.coupling_sample(spec, p, flat, want_log_det=True, layers=rng)
        else:
            _, out, ld = ops.coupling_log_prob(spec, p, flat, want_z=True, want_log_det=True, want_log_prob=False,
                                               layers=rng, with_base=False)
        return out.reshape(v.shape), ld.reshape(batch)

    # ---- generic path ---------------------------------------------------------------------------------------------
    def _run_generic(self, v: Array, inverse: bool):
        v = _cuda(v)
        mask = self.mask.to(v.device)
        masked = torch.where(mask, v, torch.zeros_like(v))
        inner = self.bijector(self.conditioner(masked))
        out0, ld = inner.inverse_and_log_det(v) if inverse else inner.forward_and_log_det(v)
        out = torch.where(mask, v, out0)
        ld = torch.where(mask, torch.zeros_like(ld), ld) if isinstance(ld, torch.Tensor) else ld
        nsum = self.event_ndims_in - self.inner_event_ndims
        if isinstance(ld, torch.Tensor) and nsum > 0:
            ld = ops.sum_last(ld.contiguous(), nsum)
        return out, ld

    def forward_and_log_det(self, x):
        return self._run_fused(x, False) if self.parity is not None else self._run_generic(x, False)

    def inverse_and_log_det(self, y):
        return self._run_fused(y, True) if self.parity is not None else self._run_generic(y, True)


class _FusedCouplingStack:
    """A Chain of MaskedCouplings with alternating masks, identical conditioner shapes and one
    bijector kind, whose parameters are consecutive slices of one flat buffer."""

    def __init__(self, layers: Sequence[MaskedCoupling], params: Array):
        self.layers, self.params = list(layers), params
        self.first = layers[0]

    def spec(self, base=_lib.BASE_NORMAL, precision=_lib.PRECISION_FP32):
        return self.first._spec(len(self.layers), base, precision)

    def _shape(self, v):
        nd = self.first.event_ndims_in
        return v.shape[: v.dim() - nd]

    def to_base(self, x, precision=_lib.PRECISION_FP32):
        x = _cuda(x)
        batch = self._shape(x)
        _, z, ld = ops.coupling_log_prob(self.spec(precision=precision), self.params, x.reshape(-1, self.first.conditioner.event_dim),
                                         want_z=True, want_log_det=True, want_log_prob=False, with_base=False)
        return z.reshape(x.shape), ld.reshape(batch)

    def to_data(self, z, precision=_lib.PRECISION_FP32):
        z = _cuda(z)
        batch = self._shape(z)
        x, ld = ops.coupling_sample(self.spec(precision=precision), self.params, z.reshape(-1, self.first.conditioner.event_dim),
                                    want_log_det=True)
        return x.reshape(z.shape), ld.reshape(batch)


def _try_fuse(bijectors) -> Optional[_FusedCouplingStack]:
    if not all(isinstance(b, MaskedCoupling) and b.parity is not None for b in bijectors):
        return None
    f = bijectors[0]
    if f.parity != 0:
        return None
    for i, b in enumerate(bijectors):
        c, c0 = b.conditioner, f.conditioner
        if b.parity != i % 2 or b.transform != f.transform or b.targs != f.targs:
            return None
        if c.hidden_sizes != c0.hidden_sizes or c.event_dim != c0.event_dim or c.out_features != c0.out_features:
            return None
    ps = [b.conditioner.params for b in bijectors]
    base = ps[0]
    contiguous = all(
        p.is_cuda and p.is_contiguous() and p.untyped_storage().data_ptr() == base.untyped_storage().data_ptr()
        and p.storage_offset() == base.storage_offset() + i * base.numel()
        for i, p in enumerate(ps)
    )
    if contiguous:
        flat = torch.as_strided(base, (base.numel() * len(ps),), (1,), base.storage_offset())
    else:
        # the reference's construction pattern (one make_conditioner(...) per layer, separate parameters): gather them
        # ONCE into one flat CUDA buffer and rebind every conditioner to its slice, so that there is a single source of
        # truth — later in-place updates (an optimiser step on `model.params` or on a conditioner's view) are seen by
        # the fused kernels, and `Transformed.params` is trainable
        flat = torch.cat([p.reshape(-1).to("cuda" if torch.cuda.is_available() else p.device, torch.float32) for p in ps]).contiguous()
        n = base.numel()
        for i, b in enumerate(bijectors):
            b.conditioner.params = flat[i * n:(i + 1) * n]
    return _FusedCouplingStack(bijectors, flat)


class MaskedAutoregressiveFlow(Bijector):
    """A stack of MADE-conditioned masked autoregressive affine layers (tfb.MaskedAutoregressiveFlow
    over tfb.AutoregressiveNetwork(params=2, hidden_units=hidden, activation="relu"), with a reversal
    between layers — the `tfb` component north_star names; the reference has no code for it,
    SURVEY App. A.7).  As in TFP, `forward` is the sequential sampling direction (base -> data) and
    `inverse` the one-pass density direction.  `params`: flat fp32 buffer, per layer per Linear
    W[in,out] then b[out], unmasked (csrc/made.cu folds the masks in)."""

    def __init__(self, event_dim: int, num_layers: int, hidden_sizes: Sequence[int], params: Optional[Array] = None,
                 seed: int = 0):
        super().__init__(event_ndims_in=1)
        self.event_dim, self.num_layers, self.hidden_sizes = int(event_dim), int(num_layers), [int(h) for h in hidden_sizes]
        self._mspec = ops.make_made_spec(self.event_dim, self.num_layers, self.hidden_sizes)
        count = ops.made_param_count(self._mspec)
        if params is None:
            gen = torch.Generator().manual_seed(seed)
            sizes = [self.event_dim] + self.hidden_sizes + [2 * self.event_dim]
            chunks = []
            for _ in range(self.num_layers):
                for i, (m, n) in enumerate(zip(sizes[:-1], sizes[1:])):
                    last = i == len(sizes) - 2  # zeros-initialised output Linear: the flow starts as the identity
                    w = torch.zeros(m, n) if last else torch.nn.init.trunc_normal_(torch.empty(m, n), std=1.0 / math.sqrt(m), generator=gen)
                    chunks += [w.reshape(-1), torch.zeros(n)]
            params = torch.cat(chunks)
        params = params.to("cuda", torch.float32).contiguous()
        if params.numel() != count:
            raise ValueError(f"expected {count} parameters, got {params.numel()}")
        self.params = params

    def spec(self, base: int = _lib.BASE_NORMAL):
        return ops.make_made_spec(self.event_dim, self.num_layers, self.hidden_sizes, base)

    def inverse_and_log_det(self, y):
        y = _cuda(y)
        batch = y.shape[:-1]
        _, u, ld = ops.made_log_prob(self._mspec, self.params, y.reshape(-1, self.event_dim), want_u=True,
                                     want_log_det=True, want_log_prob=False)
        return u.reshape(y.shape), ld.reshape(batch)

    def forward_and_log_det(self, x):
        x = _cuda(x)
        batch = x.shape[:-1]
        y = ops.made_sample(self._mspec, self.params, x.reshape(-1, self.event_dim))
        _, _, ld = ops.made_log_prob(self._mspec, self.params, y, want_log_det=True, want_log_prob=False)
        return y.reshape(x.shape), ops.affine_elementwise(ld, scale=torch.tensor(-1.0), want_ld=False)[0].reshape(batch)


# ------------------------------------------------------------------------------ distributions
class Distribution:
    event_shape: Tuple[int, ...] = ()

    def log_prob(self, value: Array) -> Array:
        raise NotImplementedError

    def prob(self, value: Array) -> Array:
        return torch.exp(self.log_prob(value))

    def _sample_n(self, gen: Optional[torch.Generator], n: int) -> Array:
        raise NotImplementedError

    def sample(self, *, seed, sample_shape=()) -> Array:
        """`seed` is an int (base noise from the library's own Philox kernel, csrc/densities.cu) or a
        torch.Generator (JAX threefry keys cannot be reproduced: parity is defined as
        same-base-noise-in -> same-sample-out)."""
        shape = (sample_shape,) if isinstance(sample_shape, int) else tuple(sample_shape)
        n = int(math.prod(shape)) if shape else 1
        gen = seed if isinstance(seed, torch.Generator) else _PhiloxSeed(int(seed))
        out = self._sample_n(gen, n)
        return out.reshape(*shape, *out.shape[1:])


class _PhiloxSeed:
    """An integer seed for ops.base_noise; `offset` advances so that successive draws do not overlap."""

    def __init__(self, seed: int):
        self.seed, self.offset = seed, 0


def _draw(gen, shape, uniform: bool) -> Array:
    if isinstance(gen, _PhiloxSeed):
        out = ops.base_noise(shape, gen.seed, uniform=uniform, offset=gen.offset)
        gen.offset += (out.numel() + 3) // 4 * 4
        return out
    return (torch.rand if uniform else torch.randn)(*shape, device="cuda", generator=gen)


class Normal(Distribution):
    """distrax.Normal(loc, scale): log_prob = -((x-loc)/scale)^2/2 - log(2 pi)/2 - log(scale) (bflow.py:44,82-87)."""

    def __init__(self, loc, scale):
        self.loc, self.scale = _t(loc), _t(scale)
        self.batch_shape = tuple(torch.broadcast_shapes(self.loc.shape, self.scale.shape))
        self.standard = bool((self.loc == 0).all() and (self.scale == 1).all())

    def log_prob(self, value):
        v = _cuda(value)
        if self.standard:
            return ops.base_log_prob(v)
        return ops.base_log_prob(v, self.loc, self.scale)

    def _sample_n(self, gen, n):
        z = _draw(gen, (n, *self.batch_shape), uniform=False)
        return z if self.standard else z * self.scale.to(z.device) + self.loc.to(z.device)


class Uniform(Distribution):
    """distrax.Uniform(low, high): log_prob = -log(high-low) inside [low, high], -inf outside (maf_mnist.py:85-88)."""

    def __init__(self, low=0.0, high=1.0):
        self.low, self.high = _t(low), _t(high)
        self.batch_shape = tuple(torch.broadcast_shapes(self.low.shape, self.high.shape))
        self.standard = bool((self.low == 0).all() and (self.high == 1).all())

    def log_prob(self, value):
        v = _cuda(value)
        if self.standard:
            return ops.base_log_prob(v, uniform=True)
        return ops.base_log_prob(v, self.low, self.high, uniform=True)

    def _sample_n(self, gen, n):
        u = _draw(gen, (n, *self.batch_shape), uniform=True)
        return u if self.standard else u * (self.high - self.low).to(u.device) + self.low.to(u.device)


class Independent(Distribution):
    """distrax.Independent(distribution, reinterpreted_batch_ndims): log_prob summed over the last k batch dims."""

    def __init__(self, distribution: Distribution, reinterpreted_batch_ndims: Optional[int] = None):
        self.distribution = distribution
        nd = len(distribution.batch_shape) if reinterpreted_batch_ndims is None else reinterpreted_batch_ndims
        self.reinterpreted_batch_ndims = nd
        self.event_shape = tuple(distribution.batch_shape[len(distribution.batch_shape) - nd:])
        self.batch_shape = tuple(distribution.batch_shape[: len(distribution.batch_shape) - nd])

    def log_prob(self, value):
        return ops.sum_last(self.distribution.log_prob(value), self.reinterpreted_batch_ndims)

    def _sample_n(self, gen, n):
        return self.distribution._sample_n(gen, n)

    def _base_code(self) -> int:
        d = self.distribution
        if isinstance(d, Normal) and d.standard:
            return _lib.BASE_NORMAL
        if isinstance(d, Uniform) and d.standard:
            return _lib.BASE_UNIFORM01
        raise NotImplementedError("the kernels fuse the standard Normal(0,1) and Uniform(0,1) bases the reference uses")


def MultivariateNormalDiag(loc=None, scale_diag=None) -> Independent:
    """tfd / distrax MultivariateNormalDiag = Independent Normal over the last dim (real_nvp.py:66, bflow.py:133)."""
    if loc is None and scale_diag is None:
        raise ValueError("at least one of `loc` and `scale_diag` must be given")
    loc = _t(loc) if loc is not None else torch.zeros_like(_t(scale_diag))
    scale_diag = _t(scale_diag) if scale_diag is not None else torch.ones_like(loc)
    return Independent(Normal(loc, scale_diag), 1)


def _inv_softplus(s: float) -> float:
    """a with softplus(a) = s, in float64."""
    return s + math.log(-math.expm1(-s)) if s < 30 else s


def _match_bflow(dist, bijector):
    """Transformed(Normal(0,1), Inverse(Chain([Chain([Scale, Shift]), BernsteinBijector, SoftClip(0,1,1.5),
    Chain([Scale, Shift])]))) with scalar parameters (bflow.py:28-34,82-84) -> (theta_constrained, scal[4]) for the fused
    kernel, else None."""
    from .BernsteinBijector import BernsteinBijector  # local: that module imports this one

    if not (isinstance(dist, Normal) and dist.standard and dist.loc.numel() == 1 and isinstance(bijector, Inverse)):
        return None
    ch = bijector.bijector
    if not (isinstance(ch, Chain) and len(ch.bijectors) == 4):
        return None
    t4, t3, t2, t1 = ch.bijectors

    def scale_shift(c):
        if not (isinstance(c, Chain) and len(c.bijectors) == 2):
            return None
        sc, sh = c.bijectors
        if type(sc) is not Scale or type(sh) is not Shift or sc._scale.numel() != 1 or sh._shift.numel() != 1:
            return None
        s = float(sc._scale.reshape(-1)[0])
        return (s, float(sh._shift.reshape(-1)[0])) if s > 0 else None

    p1, p4 = scale_shift(t1), scale_shift(t4)
    if p1 is None or p4 is None or not isinstance(t3, BernsteinBijector) or t3.thetas.dim() != 1:
        return None
    if not (isinstance(t2, SoftClip) and (t2.low, t2.high, t2.hinge_softness) == (0.0, 1.0, 1.5)):
        return None
    scal = torch.tensor([_inv_softplus(p1[0]), p1[1], _inv_softplus(p4[0]), p4[1]], dtype=torch.float32, device="cuda")
    return t3.thetas, scal


class Transformed(Distribution):
    """distrax.Transformed(distribution, bijector) (real_nvp_toy_densities.py:110, bflow.py:84)."""

    def __init__(self, distribution: Distribution, bijector: Bijector, precision: int = _lib.PRECISION_FP32):
        self.distribution, self.bijector = distribution, bijector
        self.event_shape = distribution.event_shape
        self.precision = precision
        self._stack = None
        inner = bijector.bijector if isinstance(bijector, Inverse) else None
        self._maf = bijector if isinstance(bijector, MaskedAutoregressiveFlow) and isinstance(distribution, Independent) else None
        if self._maf is not None:
            self._base = distribution._base_code()
            self.params = self._maf.params
        if isinstance(inner, Chain) and inner._fused is not None and isinstance(distribution, Independent):
            try:
                self._base = distribution._base_code()
                self._stack = inner._fused
                self.params = self._stack.params  # ONE flat buffer: what an optimiser updates (in place)
            except NotImplementedError:
                self._stack = None
        self._bflow = _match_bflow(distribution, bijector) if self._stack is None and self._maf is None else None

    # ---- serving: keep the weights packed as tensor-core tile images between calls -----------------------
    def prepare(self, max_batch: int) -> "Transformed":
        """Pack the flow's weights once for batches up to `max_batch`; later log_prob / sample calls skip the
        per-call packing pass.  Valid while the parameters stay as they are: call `prepare` again (or
        `invalidate`) after any update — the library cannot observe writes to the parameter buffer.
        No-op for flows whose path has no packed form (FP32 / 3xTF32 kernels stage their weights themselves)."""
        self._packed = None
        if self._maf is not None:
            self._packed = ops.made_pack_weights(self._maf.spec(self._base), self._maf.params, max_batch)
        elif self._stack is not None and self.precision == _lib.PRECISION_BF16X3:
            self._packed = ops.coupling_pack_weights(self._stack.spec(self._base, self.precision), self._stack.params, max_batch)
        return self

    def invalidate(self) -> None:
        self._packed = None

    def _packed_for(self, n: int):
        pk = getattr(self, "_packed", None)
        return pk if pk is not None and n <= pk.max_batch else None

    def log_prob(self, value: Array) -> Array:
        value = _cuda(value)
        if self._maf is not None:
            D = self._maf.event_dim
            batch = value.shape[:-1] if value.shape[-1] == D else (value.numel() // D,)
            flat = value.reshape(-1, self._maf.event_dim)
            lp, _, _ = ops.made_log_prob(self._maf.spec(self._base), self._maf.params, flat, packed=self._packed_for(flat.shape[0]))
            return lp.reshape(batch)
        if self._stack is not None:
            D = self._stack.first.conditioner.event_dim
            nd = len(self.event_shape)
            batch = value.shape[: value.dim() - nd]
            flat = value.reshape(-1, D)
            lp, _, _ = ops.coupling_log_prob(self._stack.spec(self._base, self.precision), self._stack.params, flat,
                                             packed=self._packed_for(flat.shape[0]))
            return lp.reshape(batch)
        if self._bflow is not None:
            theta, scal = self._bflow
            return ops.bernstein_flow_log_prob(theta, scal, value.reshape(-1)).reshape(value.shape)
        # general composition: x, ildj = bijector.inverse_and_log_det(value); base.log_prob(x) + ildj
        x, ildj = self.bijector.inverse_and_log_det(value)
        return _add_ld(self.distribution.log_prob(x), ildj)

    def sample_from_noise(self, z: Array) -> Array:
        """Transformed.sample with the base noise given: bijector.forward(z)."""
        z = _cuda(z)
        if self._maf is not None:
            return ops.made_sample(self._maf._mspec, self._maf.params, z.reshape(-1, self._maf.event_dim)).reshape(z.shape)
        if self._stack is None:
            return self.bijector.forward(z)
        D = self._stack.first.conditioner.event_dim
        flat = z.reshape(-1, D)
        x, _ = ops.coupling_sample(self._stack.spec(self._base, self.precision), self._stack.params, flat,
                                   packed=self._packed_for(flat.shape[0]))
        return x.reshape(z.shape)

    def _sample_n(self, gen, n):
        return self.sample_from_noise(self.distribution._sample_n(gen, n))

    def sample_and_log_prob(self, *, seed, sample_shape=()):
        x = self.sample(seed=seed, sample_shape=sample_shape)
        return x, self.log_prob(x)
