"""Host-side mirror of the distrax surface the reference scripts use: Bijector,
Chain, Inverse, Block, ScalarAffine, MaskedCoupling, Transformed, Independent, Normal, Uniform,
MultivariateNormalDiag.  Same names, argument meaning and composition rules
(`Chain.forward` applies the LAST bijector first; `Transformed.log_prob(v) =
base.log_prob(inverse(v)) + ildj`), but every numeric step is a libflowcuda kernel:

  * `Transformed(base, Inverse(Chain([MaskedCoupling...])))` built from `MLPConditioner`s that
    live in one flat parameter buffer collapses into ONE fused kernel call
    (`flowcuda_coupling_log_prob` / `flowcuda_coupling_sample`);
  * a single `MaskedCoupling` runs the same kernels on a one-layer range;
  * `BernsteinBijector` (BernsteinBijector.py) and the bflow chain have their own kernels.

Arbitrary Python conditioners cannot be fused into a CUDA kernel; `MaskedCoupling` therefore
requires an `MLPConditioner` (the only conditioner the reference builds:
real_nvp_toy_densities.py:56-71) and one of the tagged bijector factories below.
"""
from __future__ import annotations

import math
from typing import Callable, Optional, Sequence, Tuple

import torch

from . import _lib, ops

Array = torch.Tensor


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
            total = ld if total is None else total + ld
        return x, total

    def inverse_and_log_det(self, y):
        if self._fused is not None:
            return self._fused.to_data(y)
        total = None
        for b in self.bijectors:
            y, ld = b.inverse_and_log_det(y)
            total = ld if total is None else total + ld
        return y, total


class Block(Bijector):
    """distrax.Block(b, ndims): sums b's log-det over the last `ndims` dims (bflow.py:134)."""

    def __init__(self, bijector: Bijector, ndims: int):
        super().__init__(bijector.event_ndims_in + ndims, bijector.event_ndims_out + ndims)
        self.bijector, self.ndims = bijector, ndims

    def _sum(self, ld):
        return ld if self.ndims == 0 else ld.sum(dim=tuple(range(-self.ndims, 0)))

    def forward_and_log_det(self, x):
        y, ld = self.bijector.forward_and_log_det(x)
        return y, self._sum(ld)

    def inverse_and_log_det(self, y):
        x, ld = self.bijector.inverse_and_log_det(y)
        return x, self._sum(ld)


def scalar_affine_fn(params=None):
    """The reference's `bijector_fn` of real_nvp_toy_densities.py:84-86: split the
    conditioner output into (shift, log_scale) [.., 1] and broadcast over the event."""
    raise TypeError("scalar_affine_fn is a tag consumed by MaskedCoupling; it is never called")


scalar_affine_fn._flowcuda_transform = _lib.TRANSFORM_AFFINE_SCALAR  # type: ignore[attr-defined]
scalar_affine_fn._flowcuda_args = {}  # type: ignore[attr-defined]


def affine_fn(params=None):
    """Per-dimension (shift, log_scale): conditioner output reshaped [.., D, 2]."""
    raise TypeError("affine_fn is a tag consumed by MaskedCoupling; it is never called")


affine_fn._flowcuda_transform = _lib.TRANSFORM_AFFINE  # type: ignore[attr-defined]
affine_fn._flowcuda_args = {}  # type: ignore[attr-defined]


def rational_quadratic_spline_fn(range_min: float, range_max: float, num_bins: int) -> Callable:
    """`lambda params: distrax.RationalQuadraticSpline(params, range_min, range_max)`
    (maf_toy_densities.py:86-87, maf_mnist.py:65-66); params[..., 3*num_bins+1]."""

    def fn(params=None):
        raise TypeError("rational_quadratic_spline_fn(...) is a tag consumed by MaskedCoupling")

    fn._flowcuda_transform = _lib.TRANSFORM_RQS  # type: ignore[attr-defined]
    fn._flowcuda_args = dict(range_min=float(range_min), range_max=float(range_max), num_bins=int(num_bins))  # type: ignore[attr-defined]
    return fn


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
    """distrax.MaskedCoupling(mask, conditioner, bijector): mask True = passed
    through and visible to the conditioner.  The kernels implement the alternating
    `arange(D) % 2` masks every reference script builds (real_nvp_toy_densities.py:80-82,102)."""

    def __init__(self, mask: Array, conditioner: MLPConditioner, bijector: Callable, event_ndims: Optional[int] = None,
                 inner_event_ndims: int = 0):
        mask = torch.as_tensor(mask).bool()
        nd = mask.dim() if event_ndims is None else event_ndims
        super().__init__(nd, nd)
        if not isinstance(conditioner, MLPConditioner):
            raise TypeError("MaskedCoupling needs an MLPConditioner: arbitrary Python conditioners cannot run inside the CUDA kernel")
        if not hasattr(bijector, "_flowcuda_transform"):
            raise TypeError("MaskedCoupling needs one of scalar_affine_fn / affine_fn / rational_quadratic_spline_fn(...)")
        flat = mask.reshape(-1).cpu()
        D = flat.numel()
        odd = (torch.arange(D) % 2).bool()
        if torch.equal(flat, odd):
            self.parity = 0  # behaves like an even layer of the alternating stack
        elif torch.equal(flat, ~odd):
            self.parity = 1
        else:
            raise ValueError("MaskedCoupling kernels support the alternating masks arange(D)%2 and its negation")
        if conditioner.event_dim != D:
            raise ValueError("mask and conditioner disagree on the event size")
        self.mask, self.conditioner, self.bijector = mask, conditioner, bijector
        self.transform = bijector._flowcuda_transform
        self.targs = dict(bijector._flowcuda_args)
        ppd = 3 * self.targs["num_bins"] + 1 if self.transform == _lib.TRANSFORM_RQS else 2
        want = 2 if self.transform == _lib.TRANSFORM_AFFINE_SCALAR else D * ppd
        if conditioner.out_features != want:
            raise ValueError(f"conditioner emits {conditioner.out_features} values; this bijector needs {want}")

    def _spec(self, num_layers: int, base: int = _lib.BASE_NORMAL, precision: int = _lib.PRECISION_FP32):
        c = self.conditioner
        return ops.make_spec(c.event_dim, num_layers, c.hidden_sizes, self.transform, self.targs.get("num_bins", 0),
                             self.targs.get("range_min", 0.0), self.targs.get("range_max", 1.0), base, precision)

    def _run(self, v: Array, inverse: bool):
        # one-layer range of a 2-layer spec whose layer index carries this mask's parity
        batch = v.shape[: v.dim() - self.event_ndims_in]
        flat = v.reshape(-1, self.conditioner.event_dim) if batch else v.reshape(1, -1)
        spec = self._spec(2)
        p = self.conditioner.params
        if self.parity == 1:  # pad a dummy layer 0 in front so that this layer sits at index 1
            p = torch.cat([torch.zeros_like(p), p])
        rng = (self.parity, self.parity + 1)
        if inverse:
            out, ld = ops.coupling_sample(spec, p, flat, want_log_det=True, layers=rng)
        else:
            _, out, ld = ops.coupling_log_prob(spec, p, flat, want_z=True, want_log_det=True, want_log_prob=False,
                                               layers=rng, with_base=False)
        return out.reshape(v.shape), ld.reshape(batch)

    def forward_and_log_det(self, x):
        return self._run(x, inverse=False)

    def inverse_and_log_det(self, y):
        return self._run(y, inverse=True)


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
        batch = self._shape(x)
        _, z, ld = ops.coupling_log_prob(self.spec(precision=precision), self.params, x.reshape(-1, self.first.conditioner.event_dim),
                                         want_z=True, want_log_det=True, want_log_prob=False, with_base=False)
        return z.reshape(x.shape), ld.reshape(batch)

    def to_data(self, z, precision=_lib.PRECISION_FP32):
        batch = self._shape(z)
        x, ld = ops.coupling_sample(self.spec(precision=precision), self.params, z.reshape(-1, self.first.conditioner.event_dim),
                                    want_log_det=True)
        return x.reshape(z.shape), ld.reshape(batch)


def _try_fuse(bijectors) -> Optional[_FusedCouplingStack]:
    if not all(isinstance(b, MaskedCoupling) for b in bijectors):
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
        flat = torch.cat([p.reshape(-1) for p in ps])
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
        batch = y.shape[:-1]
        _, u, ld = ops.made_log_prob(self._mspec, self.params, y.reshape(-1, self.event_dim), want_u=True,
                                     want_log_det=True, want_log_prob=False)
        return u.reshape(y.shape), ld.reshape(batch)

    def forward_and_log_det(self, x):
        batch = x.shape[:-1]
        y = ops.made_sample(self._mspec, self.params, x.reshape(-1, self.event_dim))
        _, _, ld = ops.made_log_prob(self._mspec, self.params, y, want_log_det=True, want_log_prob=False)
        return y.reshape(x.shape), (-ld).reshape(batch)


# ------------------------------------------------------------------------------ distributions
class Distribution:
    event_shape: Tuple[int, ...] = ()

    def log_prob(self, value: Array) -> Array:
        raise NotImplementedError

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
    def __init__(self, loc, scale):
        self.loc = torch.as_tensor(loc, dtype=torch.float32)
        self.scale = torch.as_tensor(scale, dtype=torch.float32)
        self.batch_shape = tuple(torch.broadcast_shapes(self.loc.shape, self.scale.shape))
        self.standard = bool((self.loc == 0).all() and (self.scale == 1).all())

    def _sample_n(self, gen, n):
        z = _draw(gen, (n, *self.batch_shape), uniform=False)
        return z if self.standard else z * self.scale.to(z.device) + self.loc.to(z.device)


class Uniform(Distribution):
    def __init__(self, low=0.0, high=1.0):
        self.low = torch.as_tensor(low, dtype=torch.float32)
        self.high = torch.as_tensor(high, dtype=torch.float32)
        self.batch_shape = tuple(torch.broadcast_shapes(self.low.shape, self.high.shape))
        self.standard = bool((self.low == 0).all() and (self.high == 1).all())

    def _sample_n(self, gen, n):
        u = _draw(gen, (n, *self.batch_shape), uniform=True)
        return u if self.standard else u * (self.high - self.low).to(u.device) + self.low.to(u.device)


class Independent(Distribution):
    def __init__(self, distribution: Distribution, reinterpreted_batch_ndims: Optional[int] = None):
        self.distribution = distribution
        nd = len(distribution.batch_shape) if reinterpreted_batch_ndims is None else reinterpreted_batch_ndims
        self.reinterpreted_batch_ndims = nd
        self.event_shape = tuple(distribution.batch_shape[len(distribution.batch_shape) - nd:])

    def _sample_n(self, gen, n):
        return self.distribution._sample_n(gen, n)

    def _base_code(self) -> int:
        d = self.distribution
        if isinstance(d, Normal) and d.standard:
            return _lib.BASE_NORMAL
        if isinstance(d, Uniform) and d.standard:
            return _lib.BASE_UNIFORM01
        raise NotImplementedError("the kernels fuse the standard Normal(0,1) and Uniform(0,1) bases the reference uses")


def MultivariateNormalDiag(loc, scale_diag) -> Independent:
    """tfd / distrax MultivariateNormalDiag = Independent Normal over the last dim (real_nvp.py:66)."""
    return Independent(Normal(loc, scale_diag), 1)


class Transformed(Distribution):
    """distrax.Transformed(distribution, bijector) (real_nvp_toy_densities.py:110)."""

    def __init__(self, distribution: Distribution, bijector: Bijector, precision: int = _lib.PRECISION_FP32):
        self.distribution, self.bijector = distribution, bijector
        self.event_shape = distribution.event_shape
        self.precision = precision
        self._stack = None
        inner = bijector.bijector if isinstance(bijector, Inverse) else None
        self._maf = bijector if isinstance(bijector, MaskedAutoregressiveFlow) and isinstance(distribution, Independent) else None
        if self._maf is not None:
            self._base = distribution._base_code()
        if isinstance(inner, Chain) and inner._fused is not None and isinstance(distribution, Independent):
            self._stack = inner._fused
            self._base = distribution._base_code()

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
        if self._maf is not None:
            D = self._maf.event_dim
            batch = value.shape[:-1] if value.shape[-1] == D else (value.numel() // D,)
            flat = value.reshape(-1, self._maf.event_dim)
            lp, _, _ = ops.made_log_prob(self._maf.spec(self._base), self._maf.params, flat, packed=self._packed_for(flat.shape[0]))
            return lp.reshape(batch)
        if self._stack is None:
            raise NotImplementedError("Transformed.log_prob is implemented for Inverse(Chain([MaskedCoupling...])) flows; "
                                      "see BernsteinBijector.bernstein_flow_log_prob for the bflow chain")
        D = self._stack.first.conditioner.event_dim
        nd = len(self.event_shape)
        batch = value.shape[: value.dim() - nd]
        flat = value.reshape(-1, D)
        lp, _, _ = ops.coupling_log_prob(self._stack.spec(self._base, self.precision), self._stack.params, flat,
                                         packed=self._packed_for(flat.shape[0]))
        return lp.reshape(batch)

    def sample_from_noise(self, z: Array) -> Array:
        """Transformed.sample with the base noise given: bijector.forward(z)."""
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
