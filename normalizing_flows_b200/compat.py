"""Import-line compatibility for the reference scripts: the names they pull from jax / haiku / TFP, backed by
torch tensors and the libflowcuda mirror.  With

    from normalizing_flows_b200 import distrax_api as distrax
    from normalizing_flows_b200.compat import hk, jax, jnp, tfp, softplus
    tfd, tfb = tfp.distributions, tfp.bijectors

`make_conditioner` / `make_flow_model` / `log_prob` / `sample` of src/real_nvp_toy_densities.py:56-136,
src/maf_toy_densities.py:55-139, src/maf_mnist.py:34-131 and the chain of src/bflow.py:22-34,82-88,133-142 run as written
(tests/test_gpu_dropin.py holds them verbatim, import lines changed).  Only the surface those files use is provided:

  jnp   arange, reshape, zeros, ones, zeros_like, ones_like, logical_not, split, array, mean, exp, log, stack, linspace,
        meshgrid, ndarray, newaxis; arrays are torch tensors with `.astype`
  jax   nn.softplus/sigmoid/relu/leaky_relu, jit (identity), random.PRNGKey / uniform, config.update (no-op)
  hk    Sequential([Flatten, nets.MLP(..., activate_final=True), Linear, Reshape]) -> the conditioner the fused kernels
        evaluate; transform / without_apply_rng with haiku's creation-order parameter names
        (`mlp/~/linear_0`, ..., `linear`, `mlp_1/~/linear_0`, ...), so trees saved by the reference load as they are
  tfp   distributions.MultivariateNormalDiag, bijectors.Scale / Shift / SoftClip

Everything here is host-side glue (shapes, names, parameter trees); the arithmetic is in the kernels.
"""
from __future__ import annotations

import math
import types
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from . import distrax_api as _dx

_DT = {bool: torch.bool, np.bool_: torch.bool, np.float32: torch.float32, float: torch.float32, np.float64: torch.float32,
       int: torch.int64, np.int32: torch.int32, np.int64: torch.int64}


class Array(torch.Tensor):
    """A torch tensor that also answers to the few numpy-isms the reference uses (`mask.astype(bool)`)."""

    def astype(self, dtype):
        return self.to(_DT.get(dtype, dtype)).as_subclass(Array)


def _arr(t) -> Array:
    return (t if isinstance(t, torch.Tensor) else torch.as_tensor(t)).as_subclass(Array)


def _to_t(x, dtype=None):
    if isinstance(x, torch.Tensor):
        return x
    a = np.asarray(x)
    if dtype is None and a.dtype == np.float64:
        a = a.astype(np.float32)  # jax's default: float32
    return torch.as_tensor(a)


def _split(x, n, axis=-1):
    return tuple(_arr(p) for p in torch.tensor_split(_to_t(x), n, dim=axis))


jnp = types.SimpleNamespace(
    ndarray=torch.Tensor, newaxis=None, float32=np.float32, bool_=np.bool_,
    arange=lambda *a, **k: _arr(torch.arange(*[int(v) for v in a])),
    reshape=lambda x, shape: _arr(_to_t(x).reshape(tuple(int(s) for s in np.atleast_1d(shape)))),
    zeros=lambda shape, dtype=None: _arr(torch.zeros(tuple(int(s) for s in np.atleast_1d(shape)))),
    ones=lambda shape, dtype=None: _arr(torch.ones(tuple(int(s) for s in np.atleast_1d(shape)))),
    zeros_like=lambda x: _arr(torch.zeros_like(_to_t(x))),
    ones_like=lambda x: _arr(torch.ones_like(_to_t(x))),
    logical_not=lambda x: _arr(torch.logical_not(_to_t(x))),
    split=_split,
    array=lambda x, dtype=None: _arr(_to_t(x).to(_DT.get(dtype, torch.float32) if dtype is not None else
                                                (torch.float32 if _to_t(x).is_floating_point() else _to_t(x).dtype))),
    asarray=lambda x, dtype=None: _arr(_to_t(x)),
    mean=lambda x, axis=None: _arr(_to_t(x).float().mean() if axis is None else _to_t(x).float().mean(dim=axis)),
    exp=lambda x: _arr(torch.exp(_to_t(x))),
    log=lambda x: _arr(torch.log(_to_t(x))),
    stack=lambda xs, axis=0: _arr(torch.stack([_to_t(x) for x in xs], dim=axis)),
    linspace=lambda a, b, n: _arr(torch.linspace(float(a), float(b), int(n))),
    meshgrid=lambda *xs: tuple(_arr(g) for g in torch.meshgrid(*[_to_t(x) for x in xs], indexing="xy")),
)


def softplus(x):
    return _arr(torch.nn.functional.softplus(_to_t(x).float()))


_nn = types.SimpleNamespace(
    softplus=softplus,
    sigmoid=lambda x: _arr(torch.sigmoid(_to_t(x).float())),
    relu=lambda x: _arr(torch.relu(_to_t(x).float())),
    leaky_relu=lambda x, negative_slope=0.01: _arr(torch.nn.functional.leaky_relu(_to_t(x).float(), negative_slope)),
    softmax=lambda x, axis=-1: _arr(torch.softmax(_to_t(x).float(), dim=axis)),
)
_random = types.SimpleNamespace(
    PRNGKey=lambda seed: int(seed), KeyArray=int,
    uniform=lambda key, shape: _arr(torch.rand(tuple(shape), generator=torch.Generator().manual_seed(int(key)))),
    normal=lambda key, shape: _arr(torch.randn(tuple(shape), generator=torch.Generator().manual_seed(int(key)))),
)
jax = types.SimpleNamespace(nn=_nn, random=_random, numpy=jnp, jit=lambda f: f,
                            config=types.SimpleNamespace(update=lambda *a, **k: None))

tfp = types.SimpleNamespace(
    bijectors=_dx.tfb,
    distributions=types.SimpleNamespace(MultivariateNormalDiag=_dx.MultivariateNormalDiag, Normal=_dx.Normal,
                                        Uniform=_dx.Uniform, Independent=_dx.Independent),
)


# ------------------------------------------------------------------------------------------- haiku
class _Frame:
    """One hk.transform call: the parameter tree and haiku's creation-order module counters."""

    def __init__(self, params: Optional[Dict], init_seed: Optional[int]):
        self.params = params            # apply: {module: {"w", "b"}}; init: filled as conditioners are created
        self.init_seed = init_seed
        self.counts: Dict[str, int] = {}
        self.flat: Optional[torch.Tensor] = None   # apply: all leaves in haiku order, one buffer
        self.offset = 0

    def unique(self, base: str) -> str:
        k = self.counts.get(base, 0)
        self.counts[base] = k + 1
        return base if k == 0 else f"{base}_{k}"


_frames: List[_Frame] = []


class _Flatten:
    def __init__(self, preserve_dims=1, name=None):
        self.preserve_dims = int(preserve_dims)
        if _frames:
            _frames[-1].unique("flatten")


class _Reshape:
    def __init__(self, output_shape, preserve_dims=1, name=None):
        self.output_shape, self.preserve_dims = tuple(int(s) for s in output_shape), int(preserve_dims)
        if _frames:
            _frames[-1].unique("reshape")


class _MLP:
    def __init__(self, output_sizes, w_init=None, b_init=None, with_bias=True, activation=None, activate_final=False, name=None):
        if activation is not None or not with_bias:
            raise NotImplementedError("the fused conditioner is hk.nets.MLP with relu and biases (haiku's defaults)")
        self.output_sizes, self.activate_final = [int(s) for s in output_sizes], bool(activate_final)
        self.name = _frames[-1].unique("mlp") if _frames else "mlp"


class _Linear:
    def __init__(self, output_size, with_bias=True, w_init=None, b_init=None, name=None):
        if not with_bias:
            raise NotImplementedError("the fused conditioner's Linears have biases")
        self.output_size, self.w_init, self.b_init = int(output_size), w_init, b_init
        self.name = _frames[-1].unique("linear") if _frames else "linear"


def _is_zeros_init(init) -> bool:
    if init is None:
        return False
    try:
        return not bool(torch.as_tensor(np.asarray(init((2, 2), np.float32))).any())
    except Exception:
        return False


class _Sequential:
    """hk.Sequential([hk.Flatten, hk.nets.MLP(hidden, activate_final=True), hk.Linear(out), hk.Reshape]) — the conditioner
    of every reference script (real_nvp_toy_densities.py:60-71).  It is not evaluated in Python: MaskedCoupling resolves it
    into the MLPConditioner the fused kernels run, with parameters taken from the surrounding hk.transform frame."""

    def __init__(self, layers, name=None):
        layers = list(layers)
        ok = (len(layers) == 4 and isinstance(layers[0], _Flatten) and isinstance(layers[1], _MLP) and layers[1].activate_final
              and isinstance(layers[2], _Linear) and isinstance(layers[3], _Reshape))
        if not ok:
            raise NotImplementedError("hk.Sequential here is the reference's conditioner: [Flatten, nets.MLP(activate_final=True), "
                                      "Linear, Reshape]")
        self.flatten, self.mlp, self.linear, self.reshape = layers
        self.frame = _frames[-1] if _frames else None
        if self.frame is not None:
            self.frame.unique("sequential")

    def resolve_for_mask(self, mask: torch.Tensor) -> "_dx.MLPConditioner":
        event_shape = tuple(mask.shape)
        D = int(mask.numel())
        hidden, out = self.mlp.output_sizes, self.linear.output_size
        names = [f"{self.mlp.name}/~/linear_{i}" for i in range(len(hidden))] + [self.linear.name]
        sizes = [D] + hidden + [out]
        shapes = list(zip(sizes[:-1], sizes[1:]))
        count = sum(m * n + n for m, n in shapes)
        fr = self.frame
        if fr is None or (fr.params is None and fr.init_seed is None):
            return _dx.MLPConditioner(event_shape, hidden, out, out_features=out)  # eager use outside hk.transform
        if fr.params is not None and fr.init_seed is None:  # apply: a slice of the frame's flat buffer
            flat = fr.flat[fr.offset:fr.offset + count]
            fr.offset += count
            for n_, (m, k) in zip(names, shapes):
                if n_ not in fr.params:
                    raise KeyError(f"params have no module '{n_}'")
                w = fr.params[n_]["w"]
                if tuple(w.shape) != (m, k):
                    raise ValueError(f"{n_}: expected w{(m, k)}, got {tuple(w.shape)}")
            return _dx.MLPConditioner(event_shape, hidden, out, params=flat, out_features=out)
        # init: haiku defaults (TruncatedNormal(1/sqrt(fan_in)) hidden weights, zero biases) and the script's own
        # initialisers for the final Linear (jnp.zeros in every reference script)
        cond = _dx.MLPConditioner(event_shape, hidden, out, out_features=out, device="cpu", seed=fr.init_seed + len(fr.params))
        flat = cond.params
        if not _is_zeros_init(self.linear.w_init):
            m, k = shapes[-1]
            off = count - (m * k + k)
            torch.nn.init.trunc_normal_(flat[off:off + m * k], std=1.0 / math.sqrt(m), a=-2.0 / math.sqrt(m), b=2.0 / math.sqrt(m))
        off = 0
        for n_, (m, k) in zip(names, shapes):
            fr.params[n_] = {"w": flat[off:off + m * k].reshape(m, k), "b": flat[off + m * k:off + m * k + k]}
            off += m * k + k
        return cond


def _haiku_order(params: Dict) -> List[str]:
    """Module names in haiku's creation order for the reference's conditioners: mlp{_k}/~/linear_i ..., linear{_k}."""
    def layer_of(name):
        head = name.split("/")[0]
        return int(head.rsplit("_", 1)[1]) if "_" in head and head.rsplit("_", 1)[1].isdigit() else 0

    def key(name):
        is_final = not name.startswith("mlp")
        sub = int(name.rsplit("_", 1)[1]) if not is_final else 0
        return (layer_of(name), 1 if is_final else 0, sub)

    return sorted(params, key=key)


class _Transformed:
    """hk.transform(fn): `.init(rng, *args)` builds the parameter tree, `.apply(params, [rng,] *args)` runs fn with it."""

    def __init__(self, fn, with_rng=True):
        self.fn, self.with_rng = fn, with_rng
        self._cache_key, self._cache_flat = None, None

    def init(self, rng, *args, **kw):
        fr = _Frame({}, int(rng) if not isinstance(rng, torch.Tensor) else int(rng.reshape(-1)[0]))
        _frames.append(fr)
        try:
            self.fn(*args, **kw)
        except (_lib.FlowCudaError, RuntimeError, AssertionError):
            # the function body also EVALUATES the model (log_prob / sample); the parameters exist once the model is
            # built, and on a machine without a GPU (or with host-resident init parameters) the evaluation cannot run
            if not fr.params:
                raise
        finally:
            _frames.pop()
        dev = "cuda" if torch.cuda.is_available() else "cpu"
        return {m: {k: v.to(dev) for k, v in d.items()} for m, d in fr.params.items()}

    def _flat_of(self, params: Dict) -> torch.Tensor:
        order = _haiku_order(params)
        leaves = [torch.as_tensor(params[m][k]) for m in order for k in ("w", "b")]
        key = tuple((t.data_ptr(), t._version, tuple(t.shape)) for t in leaves)
        if key != self._cache_key:
            dev = "cuda" if torch.cuda.is_available() else "cpu"
            self._cache_flat = torch.cat([t.reshape(-1).to(dev, torch.float32) for t in leaves]).contiguous()
            self._cache_key = key
        return self._cache_flat

    def apply(self, params, *args, **kw):
        if self.with_rng:
            args = args[1:]
        fr = _Frame(params, None)
        fr.flat = self._flat_of(params)
        _frames.append(fr)
        try:
            return self.fn(*args, **kw)
        finally:
            _frames.pop()


hk = types.SimpleNamespace(
    Sequential=_Sequential, Flatten=_Flatten, Reshape=_Reshape, Linear=_Linear, nets=types.SimpleNamespace(MLP=_MLP),
    transform=lambda fn: _Transformed(fn, True), without_apply_rng=lambda t: _Transformed(t.fn, False), Params=dict,
)
