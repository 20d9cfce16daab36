"""numpy stand-ins for the third-party libraries the reference scripts import, so that
``/root/reference/src/{real_nvp_toy_densities,maf_toy_densities,maf_mnist,bflow}.py`` can be
EXECUTED UNMODIFIED in a container without JAX (tests/golden/make_golden_flows.py).

What this pins and what it does not
-----------------------------------
Everything the *reference* wrote runs as the reference wrote it: ``make_conditioner``,
``make_flow_model`` (mask = ``arange % 2`` reshaped, ``logical_not`` per layer, ``3*n+1``,
the ``hk.Flatten`` / ``hk.Reshape`` dims, the ``bijector_fn`` closures, ``Inverse(Chain(layers))``,
the base distributions), the ``hk.transform``-ed ``log_prob`` / ``sample`` wrappers and — through
haiku's creation-order module naming — the parameter tree keys ``params_io.py`` reads.

The library primitives below are a float64 numpy restatement of distrax / haiku / TFP written
from their published definitions; the libraries themselves are absent
from /root/reference and from this image, so THESE remain "restated, not executed".  They are
deliberately a second, independent implementation (numpy, per-element bin search) of what
``oracle/flows_oracle.py`` restates in torch — the golden test therefore also cross-checks
the oracle's primitives against a differently written copy.

TEST INFRASTRUCTURE: imported only by tests/golden/make_golden_flows.py (build container).
"""
from __future__ import annotations

import math
import sys
import types

import numpy as np

F = np.float64


# ------------------------------------------------------------------------------------------
# jax
# ------------------------------------------------------------------------------------------
def softplus(x):
    return np.logaddexp(np.asarray(x, dtype=F), 0.0)


def softmax(x, axis=-1):
    x = np.asarray(x, dtype=F)
    e = np.exp(x - np.max(x, axis=axis, keepdims=True))
    return e / e.sum(axis=axis, keepdims=True)


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-np.asarray(x, dtype=F)))


def _vmap(fn, in_axes=0, out_axes=0):
    def mapped(*args):
        n = args[0].shape[in_axes]
        outs = [fn(*[np.take(a, i, axis=in_axes) for a in args]) for i in range(n)]
        return np.stack(outs, axis=out_axes)

    return mapped


# ------------------------------------------------------------------------------------------
# haiku: modules are named at construction in creation order inside the current transform
# ------------------------------------------------------------------------------------------
class _Frame:
    def __init__(self, params, init_rng):
        self.params = params  # {module_name: {"w": , "b": }}
        self.init_rng = init_rng  # None in apply mode
        self.counts = {}
        self.scope = ""

    def unique(self, base):
        key = self.scope + base
        k = self.counts.get(key, 0)
        self.counts[key] = k + 1
        return key if k == 0 else f"{key}_{k}"


_frames = []


def _frame() -> _Frame:
    if not _frames:
        raise RuntimeError("haiku modules must be built inside hk.transform")
    return _frames[-1]


class Module:
    def __init__(self, name=None):
        fr = _frame()
        base = name if name is not None else _snake(type(self).__name__)
        # an explicit name is still uniquified by haiku; inside a parent scope it is prefixed "parent/~/"
        self.module_name = fr.unique(base)


def _snake(s):
    out = ""
    for i, ch in enumerate(s):
        if ch.isupper() and i and not s[i - 1].isupper():
            out += "_"
        out += ch.lower()
    return out


def get_parameter(module: Module, name, shape, init):
    fr = _frame()
    slot = fr.params.setdefault(module.module_name, {})
    if name not in slot:
        if fr.init_rng is None:
            raise KeyError(f"params have no '{module.module_name}/{name}'")
        slot[name] = np.asarray(init(tuple(shape), fr.init_rng), dtype=F)
    p = np.asarray(slot[name])
    if tuple(p.shape) != tuple(shape):
        raise ValueError(f"{module.module_name}/{name}: expected {tuple(shape)}, got {p.shape}")
    return p


def _trunc_normal(stddev):
    def init(shape, rng):
        out = rng.normal(size=shape)
        bad = np.abs(out) > 2.0
        while bad.any():
            out[bad] = rng.normal(size=int(bad.sum()))
            bad = np.abs(out) > 2.0
        return out * stddev

    return init


def _zeros(shape, rng=None, dtype=None):
    return np.zeros(shape, dtype=F)


class Linear(Module):
    def __init__(self, output_size, with_bias=True, w_init=None, b_init=None, name=None):
        super().__init__(name)
        self.output_size, self.with_bias = int(output_size), with_bias
        self.w_init, self.b_init = w_init, b_init

    @staticmethod
    def _wrap(init):
        # the reference passes jnp.zeros (shape, dtype) as an initialiser
        return lambda shape, rng: np.asarray(init(shape, F), dtype=F)

    def __call__(self, x):
        n_in = x.shape[-1]
        w_init = self._wrap(self.w_init) if self.w_init is not None else _trunc_normal(1.0 / math.sqrt(n_in))
        w = get_parameter(self, "w", (n_in, self.output_size), w_init)
        out = x @ w
        if self.with_bias:
            b_init = self._wrap(self.b_init) if self.b_init is not None else _zeros
            out = out + get_parameter(self, "b", (self.output_size,), b_init)
        return out


class MLP(Module):
    def __init__(self, output_sizes, w_init=None, b_init=None, with_bias=True, activation=None, activate_final=False,
                 name=None):
        super().__init__(name)
        fr = _frame()
        self.activation = (lambda x: np.maximum(x, 0.0)) if activation is None else activation  # jax.nn.relu default
        self.activate_final = activate_final
        outer, fr.scope = fr.scope, self.module_name + "/~/"
        try:
            self.layers = [Linear(n, with_bias=with_bias, w_init=w_init, b_init=b_init, name=f"linear_{i}")
                           for i, n in enumerate(output_sizes)]
        finally:
            fr.scope = outer

    def __call__(self, x):
        n = len(self.layers)
        for i, layer in enumerate(self.layers):
            x = layer(x)
            if i < n - 1 or self.activate_final:
                x = self.activation(x)
        return x


class Reshape(Module):
    """hk.Reshape(output_shape, preserve_dims): negative preserve_dims = number of TRAILING dims replaced."""

    def __init__(self, output_shape, preserve_dims=1, name=None):
        super().__init__(name)
        self.output_shape, self.preserve_dims = tuple(int(s) for s in output_shape), int(preserve_dims)

    def __call__(self, x):
        if self.preserve_dims >= 0:
            lead = x.shape[: self.preserve_dims]
        else:
            lead = x.shape[: x.ndim + self.preserve_dims]
        return np.reshape(x, tuple(lead) + self.output_shape)


class Flatten(Reshape):
    def __init__(self, preserve_dims=1, name=None):
        super().__init__((-1,), preserve_dims=preserve_dims, name=name)


class Sequential(Module):
    def __init__(self, layers, name=None):
        super().__init__(name)
        self.layers = tuple(layers)

    def __call__(self, x):
        for f in self.layers:
            x = f(x)
        return x


class Transformed:
    def __init__(self, fn, with_rng=True):
        self.fn, self.with_rng = fn, with_rng

    def init(self, rng, *args, **kw):
        fr = _Frame({}, np.random.default_rng(rng if isinstance(rng, (int, np.integer)) else 0))
        _frames.append(fr)
        try:
            self.fn(*args, **kw)
        finally:
            _frames.pop()
        return fr.params

    def apply(self, params, *args, **kw):
        if self.with_rng:
            args = args[1:]
        _frames.append(_Frame(params, None))
        try:
            return self.fn(*args, **kw)
        finally:
            _frames.pop()


def transform(fn):
    return Transformed(fn, with_rng=True)


def without_apply_rng(t: Transformed):
    return Transformed(t.fn, with_rng=False)


# ------------------------------------------------------------------------------------------
# distrax
# ------------------------------------------------------------------------------------------
class Bijector:
    def __init__(self, event_ndims_in=0, event_ndims_out=None, is_constant_jacobian=False, is_constant_log_det=None):
        self.event_ndims_in = event_ndims_in
        self.event_ndims_out = event_ndims_in if event_ndims_out is None else event_ndims_out

    def forward(self, x):
        return self.forward_and_log_det(x)[0]

    def inverse(self, y):
        return self.inverse_and_log_det(y)[0]

    def forward_log_det_jacobian(self, x):
        return self.forward_and_log_det(x)[1]

    def inverse_log_det_jacobian(self, y):
        return self.inverse_and_log_det(y)[1]


def _as_bijector(b):
    """distrax accepts TFP bijectors and wraps them; ours already share the protocol."""
    return b


class Chain(Bijector):
    def __init__(self, bijectors):
        self.bijectors = [_as_bijector(b) for b in bijectors]
        super().__init__(self.bijectors[-1].event_ndims_in, self.bijectors[0].event_ndims_out)

    def forward_and_log_det(self, x):
        x, ld = self.bijectors[-1].forward_and_log_det(x)
        for b in reversed(self.bijectors[:-1]):
            x, l = b.forward_and_log_det(x)
            ld = ld + l
        return x, ld

    def inverse_and_log_det(self, y):
        y, ld = self.bijectors[0].inverse_and_log_det(y)
        for b in self.bijectors[1:]:
            y, l = b.inverse_and_log_det(y)
            ld = ld + l
        return y, ld


class Inverse(Bijector):
    def __init__(self, bijector):
        self.bijector = _as_bijector(bijector)
        super().__init__(self.bijector.event_ndims_out, self.bijector.event_ndims_in)

    def forward_and_log_det(self, x):
        return self.bijector.inverse_and_log_det(x)

    def inverse_and_log_det(self, y):
        return self.bijector.forward_and_log_det(y)


def _sum_last(x, ndims):
    return x if ndims == 0 else np.sum(x, axis=tuple(range(-ndims, 0)))


class Block(Bijector):
    def __init__(self, bijector, ndims):
        self.bijector, self.ndims = _as_bijector(bijector), int(ndims)
        super().__init__(self.bijector.event_ndims_in + ndims, self.bijector.event_ndims_out + ndims)

    def forward_and_log_det(self, x):
        y, ld = self.bijector.forward_and_log_det(x)
        return y, _sum_last(ld, self.ndims)

    def inverse_and_log_det(self, y):
        x, ld = self.bijector.inverse_and_log_det(y)
        return x, _sum_last(ld, self.ndims)


class ScalarAffine(Bijector):
    def __init__(self, shift, scale=None, log_scale=None):
        super().__init__(0)
        self.shift = np.asarray(shift, dtype=F)
        if scale is not None and log_scale is not None:
            raise ValueError("Only one of `scale` and `log_scale` can be specified, not both.")
        if scale is None and log_scale is None:
            self.scale, self.log_scale = np.ones_like(self.shift), np.zeros_like(self.shift)
        elif log_scale is not None:
            self.log_scale = np.asarray(log_scale, dtype=F)
            self.scale = np.exp(self.log_scale)
        else:
            self.scale = np.asarray(scale, dtype=F)
            self.log_scale = np.log(np.abs(self.scale))

    def forward_and_log_det(self, x):
        y = self.scale * x + self.shift
        return y, np.broadcast_to(self.log_scale, y.shape)

    def inverse_and_log_det(self, y):
        x = (y - self.shift) / self.scale
        return x, np.broadcast_to(-self.log_scale, x.shape)


class RationalQuadraticSpline(Bijector):
    """Durkan et al. 2019 as distrax implements it: unconstrained boundary slopes, linear tails."""

    def __init__(self, params, range_min, range_max, boundary_slopes="unconstrained", min_bin_size=1e-4,
                 min_knot_slope=1e-4):
        super().__init__(0)
        params = np.asarray(params, dtype=F)
        if params.shape[-1] % 3 != 1 or params.shape[-1] < 4:
            raise ValueError(f"The last dimension of `params` must have size `3 * num_bins + 1` (got {params.shape[-1]}).")
        if range_min >= range_max:
            raise ValueError("`range_min` must be less than `range_max`.")
        if boundary_slopes != "unconstrained":
            raise NotImplementedError(boundary_slopes)
        K = (params.shape[-1] - 1) // 3
        self.num_bins = K
        total = float(range_max) - float(range_min)

        def sizes(u):
            return softmax(u, axis=-1) * (total - K * min_bin_size) + min_bin_size

        widths, heights = sizes(params[..., :K]), sizes(params[..., K:2 * K])
        pad_lo = np.full(params.shape[:-1] + (1,), float(range_min))
        pad_hi = np.full(params.shape[:-1] + (1,), float(range_max))
        self.x_pos = np.concatenate([pad_lo, range_min + np.cumsum(widths[..., :-1], axis=-1), pad_hi], axis=-1)
        self.y_pos = np.concatenate([pad_lo, range_min + np.cumsum(heights[..., :-1], axis=-1), pad_hi], axis=-1)
        offset = math.log(math.exp(1.0 - min_knot_slope) - 1.0)
        self.knot_slopes = softplus(params[..., 2 * K:] + offset) + min_knot_slope

    def _bin(self, v, pos):
        """index of the bin with pos[i] <= v < pos[i+1]; bin 0 when there is none (outside the range)."""
        inside = (v[..., None] >= pos[..., :-1]) & (v[..., None] < pos[..., 1:])
        idx = np.argmax(inside, axis=-1)
        return np.where(inside.any(axis=-1), idx, 0)

    @staticmethod
    def _take(arr, idx):
        return np.take_along_axis(arr, idx[..., None], axis=-1)[..., 0]

    def _bcast(self, v):
        v = np.asarray(v, dtype=F)
        shp = np.broadcast_shapes(v.shape, self.x_pos.shape[:-1])
        bc = lambda a: np.broadcast_to(a, shp + a.shape[-1:])
        return np.broadcast_to(v, shp), bc(self.x_pos), bc(self.y_pos), bc(self.knot_slopes)

    def forward_and_log_det(self, x):
        x, xp, yp, sl = self._bcast(x)
        i = self._bin(x, xp)
        x0, x1 = self._take(xp, i), self._take(xp, i + 1)
        y0, y1 = self._take(yp, i), self._take(yp, i + 1)
        d0, d1 = self._take(sl, i), self._take(sl, i + 1)
        w, h = x1 - x0, y1 - y0
        s = h / w
        z = np.clip((x - x0) / w, 0.0, 1.0)
        zz, z1 = z * z, z - z * z
        omz2 = (1.0 - z) ** 2
        t = d1 + d0 - 2.0 * s
        den = s + t * z1
        y = y0 + h * (s * zz + d0 * z1) / den
        ld = 2.0 * np.log(s) + np.log(d1 * zz + 2.0 * s * z1 + d0 * omz2) - 2.0 * np.log(den)
        below, above = x <= xp[..., 0], x >= xp[..., -1]
        y = np.where(below, (x - xp[..., 0]) * sl[..., 0] + yp[..., 0], y)
        y = np.where(above, (x - xp[..., -1]) * sl[..., -1] + yp[..., -1], y)
        ld = np.where(below, np.log(sl[..., 0]), ld)
        ld = np.where(above, np.log(sl[..., -1]), ld)
        return y, ld

    def inverse_and_log_det(self, y):
        y, xp, yp, sl = self._bcast(y)
        i = self._bin(y, yp)
        x0, x1 = self._take(xp, i), self._take(xp, i + 1)
        y0, y1 = self._take(yp, i), self._take(yp, i + 1)
        d0, d1 = self._take(sl, i), self._take(sl, i + 1)
        w, h = x1 - x0, y1 - y0
        s = h / w
        om = np.clip((y - y0) / h, 0.0, 1.0)
        t = d1 + d0 - 2.0 * s
        c = -s * om
        b = d0 - t * om
        a = s - b
        z = np.clip(-2.0 * c / (b + np.sqrt(b * b - 4.0 * a * c)), 0.0, 1.0)
        x = w * z + x0
        zz, z1 = z * z, z - z * z
        omz2 = (1.0 - z) ** 2
        den = s + t * z1
        ld = -2.0 * np.log(s) - np.log(d1 * zz + 2.0 * s * z1 + d0 * omz2) + 2.0 * np.log(den)
        below, above = y <= yp[..., 0], y >= yp[..., -1]
        x = np.where(below, (y - yp[..., 0]) / sl[..., 0] + xp[..., 0], x)
        x = np.where(above, (y - yp[..., -1]) / sl[..., -1] + xp[..., -1], x)
        ld = np.where(below, -np.log(sl[..., 0]), ld)
        ld = np.where(above, -np.log(sl[..., -1]), ld)
        return x, ld


class MaskedCoupling(Bijector):
    def __init__(self, mask, conditioner, bijector, event_ndims=None, inner_event_ndims=0):
        mask = np.asarray(mask)
        if mask.dtype != bool:
            raise ValueError(f"`mask` must have values of type `bool`; got values of type `{mask.dtype}`.")
        nd = mask.ndim if event_ndims is None else event_ndims
        super().__init__(nd)
        self.mask, self.conditioner, self.bijector_fn = mask, conditioner, bijector
        self.inner_event_ndims = inner_event_ndims

    def _run(self, v, inverse):
        masked = np.where(self.mask, v, 0.0)
        params = self.conditioner(masked)
        inner = self.bijector_fn(params)
        out0, ld = inner.inverse_and_log_det(v) if inverse else inner.forward_and_log_det(v)
        out = np.where(self.mask, v, out0)
        ld = np.where(self.mask, 0.0, ld)
        return out, _sum_last(ld, self.event_ndims_in - self.inner_event_ndims)

    def forward_and_log_det(self, x):
        return self._run(np.asarray(x, dtype=F), False)

    def inverse_and_log_det(self, y):
        return self._run(np.asarray(y, dtype=F), True)


class Distribution:
    pass


class Normal(Distribution):
    def __init__(self, loc, scale):
        self.loc, self.scale = np.asarray(loc, dtype=F), np.asarray(scale, dtype=F)
        self.batch_shape = np.broadcast_shapes(self.loc.shape, self.scale.shape)

    def log_prob(self, value):
        z = (np.asarray(value, dtype=F) - self.loc) / self.scale
        return -0.5 * z * z - 0.5 * math.log(2.0 * math.pi) - np.log(self.scale)

    def sample(self, seed, sample_shape=()):
        return _rng(seed).normal(size=tuple(sample_shape) + tuple(self.batch_shape)) * self.scale + self.loc


class Uniform(Distribution):
    def __init__(self, low=0.0, high=1.0):
        self.low, self.high = np.asarray(low, dtype=F), np.asarray(high, dtype=F)
        self.batch_shape = np.broadcast_shapes(self.low.shape, self.high.shape)

    def log_prob(self, value):
        value = np.asarray(value, dtype=F)
        outside = (value < self.low) | (value > self.high)
        with np.errstate(divide="ignore"):
            return np.log(np.where(outside, 0.0, 1.0 / (self.high - self.low)))

    def sample(self, seed, sample_shape=()):
        u = _rng(seed).uniform(size=tuple(sample_shape) + tuple(self.batch_shape))
        return self.low + (self.high - self.low) * u


class Independent(Distribution):
    def __init__(self, distribution, reinterpreted_batch_ndims=None):
        self.distribution = distribution
        nd = len(distribution.batch_shape) if reinterpreted_batch_ndims is None else reinterpreted_batch_ndims
        self.reinterpreted_batch_ndims = nd

    def log_prob(self, value):
        return _sum_last(self.distribution.log_prob(value), self.reinterpreted_batch_ndims)

    def sample(self, seed, sample_shape=()):
        return self.distribution.sample(seed, sample_shape)


class MultivariateNormalDiag(Independent):
    def __init__(self, loc, scale_diag):
        super().__init__(Normal(loc, scale_diag), 1)


class TransformedDist(Distribution):
    def __init__(self, distribution, bijector):
        self.distribution, self.bijector = distribution, _as_bijector(bijector)

    def log_prob(self, value):
        x, ildj = self.bijector.inverse_and_log_det(np.asarray(value, dtype=F))
        return self.distribution.log_prob(x) + ildj

    def sample(self, seed, sample_shape=()):
        z = self.distribution.sample(seed, sample_shape)
        self.last_base_sample = z  # exposed so that the golden file can record the base noise that was used
        LAST_BASE_SAMPLE[0] = z
        return self.bijector.forward(z)


LAST_BASE_SAMPLE = [None]


def _rng(seed):
    return seed if isinstance(seed, np.random.Generator) else np.random.default_rng(seed)


# ------------------------------------------------------------------------------------------
# TFP bijectors used by src/bflow.py:28-31
# ------------------------------------------------------------------------------------------
class Scale(Bijector):
    def __init__(self, scale):
        super().__init__(0)
        self.scale = np.asarray(scale, dtype=F)

    def forward_and_log_det(self, x):
        y = np.asarray(x, dtype=F) * self.scale
        return y, np.broadcast_to(np.log(np.abs(self.scale)), y.shape)

    def inverse_and_log_det(self, y):
        x = np.asarray(y, dtype=F) / self.scale
        return x, np.broadcast_to(-np.log(np.abs(self.scale)), x.shape)


class Shift(Bijector):
    def __init__(self, shift):
        super().__init__(0)
        self.shift = np.asarray(shift, dtype=F)

    def forward_and_log_det(self, x):
        y = np.asarray(x, dtype=F) + self.shift
        return y, np.zeros(y.shape)

    def inverse_and_log_det(self, y):
        x = np.asarray(y, dtype=F) - self.shift
        return x, np.zeros(x.shape)


class SoftClip(Bijector):
    """tfb.SoftClip(low, high, hinge_softness=c) = Chain([Shift(high), Scale(-k), Softplus(c), Shift(high-low) . Scale(-1),
    Softplus(c), Shift(-low)]) in TFP's own composition: y = high - k*sp(high - low - sp(x - low)),
    k = (high-low)/sp(high-low), sp(u) = c*softplus(u/c)."""

    def __init__(self, low=None, high=None, hinge_softness=None):
        super().__init__(0)
        self.low, self.high = float(low), float(high)
        self.c = 1.0 if hinge_softness is None else float(hinge_softness)

    def _sp(self, u):
        return self.c * softplus(np.asarray(u, dtype=F) / self.c)

    def forward_and_log_det(self, x):
        x = np.asarray(x, dtype=F)
        r = self.high - self.low
        k = r / self._sp(r)
        inner = self._sp(x - self.low)
        u = r - inner
        y = self.high - k * self._sp(u)
        ld = np.log(k) + np.log(sigmoid(u / self.c)) + np.log(sigmoid((x - self.low) / self.c))
        return y, ld

    def _sp_inv(self, v):
        # inverse of c*softplus(u/c): u = c*log(expm1(v/c))
        return self.c * np.log(np.expm1(np.asarray(v, dtype=F) / self.c))

    def inverse_and_log_det(self, y):
        y = np.asarray(y, dtype=F)
        r = self.high - self.low
        k = r / self._sp(r)
        u = self._sp_inv((self.high - y) / k)
        x = self.low + self._sp_inv(r - u)
        return x, -self.forward_and_log_det(x)[1]


class Beta:
    """tfd.Beta(concentration1, concentration0).prob (src/BernsteinBijector.py:17-20,39-42)."""

    def __init__(self, concentration1, concentration0):
        self.a, self.b = np.asarray(concentration1, dtype=F), np.asarray(concentration0, dtype=F)

    def prob(self, x):
        import scipy.stats

        return scipy.stats.beta.pdf(np.asarray(x, dtype=F), self.a, self.b)


# ------------------------------------------------------------------------------------------
def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class _Anything(types.ModuleType):
    """A module whose every attribute is a harmless callable/namespace (plotting, optimisers, datasets:
    imported at the top of the reference scripts, never touched by the functions we call)."""

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _Anything(f"{self.__name__}.{name}")

    def __call__(self, *a, **k):
        return _Anything(self.__name__ + "()")

    def __iter__(self):
        return iter(())


def install():
    """Put the stand-ins into sys.modules under the names the reference imports."""
    np.seterr(all="ignore")
    jnp = _mod("jax.numpy")
    jnp.__dict__.update({k: getattr(np, k) for k in dir(np) if not k.startswith("_")})
    jnp.ndarray = np.ndarray
    jnp.array = lambda a, dtype=None: np.asarray(a, dtype=F)
    jnp.zeros = lambda shape, dtype=None: np.zeros(shape, dtype=F)
    jnp.ones = lambda shape, dtype=None: np.ones(shape, dtype=F)
    jnn = _mod("jax.nn", softplus=softplus, softmax=softmax, sigmoid=sigmoid,
               leaky_relu=lambda x, negative_slope=0.01: np.where(x >= 0, x, negative_slope * x),
               relu=lambda x: np.maximum(x, 0.0))
    jrandom = _mod("jax.random", KeyArray=object, PRNGKey=lambda s: int(s),
                   uniform=lambda key, shape: _rng(key).uniform(size=shape))
    cfg = types.SimpleNamespace(update=lambda *a, **k: None)
    jnp.newaxis = None
    _mod("jax", numpy=jnp, nn=jnn, random=jrandom, config=cfg, jit=lambda f: f, vmap=_vmap,
         grad=_Anything("jax.grad"), value_and_grad=_Anything("jax.value_and_grad"))
    nets = _mod("haiku.nets", MLP=MLP)
    _mod("haiku", Module=Module, Linear=Linear, Flatten=Flatten, Reshape=Reshape, Sequential=Sequential, nets=nets,
         transform=transform, without_apply_rng=without_apply_rng, Params=dict, PRNGSequence=_Anything("hk.PRNGSequence"))
    _mod("distrax", Bijector=Bijector, Chain=Chain, Inverse=Inverse, Block=Block, ScalarAffine=ScalarAffine,
         RationalQuadraticSpline=RationalQuadraticSpline, MaskedCoupling=MaskedCoupling, Normal=Normal, Uniform=Uniform,
         Independent=Independent, MultivariateNormalDiag=MultivariateNormalDiag, Transformed=TransformedDist,
         Tanh=_Anything("distrax.Tanh"))
    _mod("distrax._src")
    _mod("distrax._src.bijectors")
    _mod("distrax._src.bijectors.bijector", Array=np.ndarray)
    _mod("distrax._src.distributions")
    _mod("distrax._src.distributions.distribution", PRNGKey=object)
    tfb = types.SimpleNamespace(Scale=Scale, Shift=Shift, SoftClip=SoftClip)
    tfd = types.SimpleNamespace(Beta=Beta, MultivariateNormalDiag=MultivariateNormalDiag)
    tfp_jax = _mod("tensorflow_probability.substrates.jax", bijectors=tfb, distributions=tfd)
    subs = _mod("tensorflow_probability.substrates", jax=tfp_jax)
    _mod("tensorflow_probability", substrates=subs)
    for name in ("optax", "tensorflow_datasets", "matplotlib", "matplotlib.pyplot", "seaborn", "imageio", "webbrowser",
                 "densities", "densities.banana", "densities.energies", "densities.gaussian_blobs", "scipy.interpolate"):
        if name == "scipy.interpolate":
            continue  # real scipy is installed
        sys.modules[name] = _Anything(name)
    sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]
