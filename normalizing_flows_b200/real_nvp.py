"""Mirror of the reference's src/real_nvp.py functional API on libflowcuda kernels.

`params` keeps the reference's structure params[layer] = [[W1[1,H], b1[H]], [W2[H,2], b2[2]]]
(real_nvp.py:188-195); `pack_params` flattens it into the device layout the kernels read.
`fns_config` is the reference's list of (fn, flip); `fn` must be `shift_and_log_scale_fn`
(the leaky-relu MLP of real_nvp.py:161-171 is what the fused kernel implements)."""
from __future__ import annotations

from typing import Callable, List, Sequence, Tuple

import torch

from . import ops


def shift_and_log_scale_fn(net_params, x):
    """real_nvp.py:161-171.  A tag: the conditioner runs inside the fused kernels."""
    raise TypeError("shift_and_log_scale_fn is evaluated inside the CUDA kernels; pass it as `fn` only")


def _check_fn(fn):
    if fn is not shift_and_log_scale_fn:
        raise TypeError("the fused Real NVP kernels implement real_nvp.shift_and_log_scale_fn only")


_PACK_CACHE = {}  # fingerprint of the leaves -> (the leaves themselves, flat CUDA tensor)


def pack_params(params) -> torch.Tensor:
    """[[W1,b1],[W2,b2]] per layer -> flat float32 CUDA tensor (W row-major, then b).

    The reference passes the nested list on every call (`forward_nvp(net_params, ...)`, real_nvp.py:71); packing and
    copying it each time would put a host round trip in front of a 40 us kernel.  The flat buffer is cached and reused
    while every leaf is the same tensor object at the same version (torch bumps `_version` on in-place writes; the
    cache keeps the leaves alive, so an id cannot be recycled under it); numpy leaves are fingerprinted by their bytes
    (18 floats per layer)."""
    if isinstance(params, torch.Tensor):
        return params.to("cuda", torch.float32).contiguous()
    leaves = [t for layer in params for (W, b) in layer for t in (W, b)]
    fp = tuple((id(t), t._version) if isinstance(t, torch.Tensor) else (t.tobytes() if hasattr(t, "tobytes") else repr(t))
               for t in leaves)
    hit = _PACK_CACHE.get(fp)
    if hit is not None:
        return hit[1]
    chunks = [torch.as_tensor(t, dtype=torch.float32).reshape(-1) for t in leaves]
    if all(c.is_cuda for c in chunks):
        flat = torch.cat(chunks).contiguous()
    else:
        flat = torch.cat([c.cpu() for c in chunks]).to("cuda").contiguous()
    if len(_PACK_CACHE) >= 64:
        _PACK_CACHE.clear()
    _PACK_CACHE[fp] = (leaves, flat)
    return flat


def unpack_params(flat: torch.Tensor, n_layers: int, hidden: int = 4):
    out, o = [], 0
    for _ in range(n_layers):
        W1 = flat[o:o + hidden].reshape(1, hidden); o += hidden
        b1 = flat[o:o + hidden]; o += hidden
        W2 = flat[o:o + 2 * hidden].reshape(hidden, 2); o += 2 * hidden
        b2 = flat[o:o + 2]; o += 2
        out.append([[W1, b1], [W2, b2]])
    return out


def _hidden_of(params) -> int:
    if isinstance(params, torch.Tensor):
        raise TypeError("pass hidden= explicitly with flat params")
    return int(params[0][0][0].shape[-1])


def init_mlp(hidden_units: List[int] = [4], seed: int = 534):
    """real_nvp.py:174-195: ignores its argument, [1,4,2], W ~ 0.1*N(0,1), b = 0."""
    g = torch.Generator().manual_seed(seed)
    params = []
    for m, n in ((1, 4), (4, 2)):
        params.append([torch.randn(m, n, generator=g) * 0.1, torch.zeros(n)])
    return params


def init_nvp_chain(n: int = 2):
    """real_nvp.py:198-220 (every layer starts from the same seed, as in the reference)."""
    flip, params, fns_config = False, [], []
    for _ in range(n):
        params.append(init_mlp())
        fns_config.append((shift_and_log_scale_fn, flip))
        flip = not flip
    return params, fns_config


def _flat_single(net_params) -> Tuple[torch.Tensor, int]:
    H = int(net_params[0][0].shape[-1])
    return pack_params([net_params]), H


def forward_nvp(net_params, shift_log_scale_fn: Callable, x, flip: bool = False):
    """real_nvp.py:71-101."""
    _check_fn(shift_log_scale_fn)
    p, H = _flat_single(net_params)
    return ops.realnvp_forward(p, H, flip, _cuda(x))


def inverse_nvp(net_params, shift_log_scale_fn: Callable, y, flip: bool = False):
    """real_nvp.py:104-134: returns (x, log_scale[N,1])."""
    _check_fn(shift_log_scale_fn)
    p, H = _flat_single(net_params)
    return ops.realnvp_inverse(p, H, flip, _cuda(y))


def log_prob_N01(x):
    """real_nvp.py:57-68.  A tag for the fused base density; also callable on its own."""
    x = _cuda(x)
    ident = torch.tensor([0.0] * 4 + [0.0] * 4 + [0.0] * 8 + [0.0, 0.0], device="cuda")  # identity layer
    return ops.realnvp_log_prob(ident, 1, 4, 0, x)


def sample_N01(N: int, seed):
    """real_nvp.py:40-53; `seed` is an int or torch.Generator (threefry keys are not reproducible)."""
    gen = seed if isinstance(seed, torch.Generator) else torch.Generator(device="cuda").manual_seed(int(seed))
    return torch.randn(N, 2, device="cuda", generator=gen)


def log_prob_nvp(net_params, shift_log_scale_fn, base_log_prob_fn, y, flip: bool = False):
    """real_nvp.py:137-158."""
    return log_prob_nvp_chains([net_params], [(shift_log_scale_fn, flip)], base_log_prob_fn, y)


def log_prob_nvp_chains(params, fns_config: Sequence[Tuple[Callable, bool]], base_log_prob_fn, y, hidden: int = None):
    """real_nvp.py:268-283: one fused kernel for the whole chain + N(0,I) log-density."""
    if base_log_prob_fn is not log_prob_N01:
        raise TypeError("the fused kernel implements the reference's base density log_prob_N01")
    for fn, _ in fns_config:
        _check_fn(fn)
    H = hidden if hidden is not None else _hidden_of(params)
    return ops.realnvp_log_prob(pack_params(params), len(fns_config), H, ops.flip_mask_of([f for _, f in fns_config]), _cuda(y))


def sample_nvp_chain(params, fns_config, base_sample_fn: Callable, N: int, key, hidden: int = None):
    """real_nvp.py:228-248."""
    for fn, _ in fns_config:
        _check_fn(fn)
    z = _cuda(base_sample_fn(N, key))
    H = hidden if hidden is not None else _hidden_of(params)
    return ops.realnvp_sample(pack_params(params), len(fns_config), H, ops.flip_mask_of([f for _, f in fns_config]), z)


def loss(params, fns_config, batch, hidden: int = None):
    """-mean(log_prob_nvp_chains(...)) (real_nvp.py:292-293)."""
    return ops.nll(log_prob_nvp_chains(params, fns_config, log_prob_N01, batch, hidden=hidden))


class SGDTrainer:
    """jit(sgd_update) of real_nvp.py:295-306 with parameters resident on the device: one call =
    value_and_grad + p <- p - lr*g in two kernels (stage 2 fuses the update)."""

    def __init__(self, params, fns_config, learning_rate: float = 1e-2, hidden: int = None):
        self.hidden = hidden if hidden is not None else _hidden_of(params)
        self.n_layers = len(fns_config)
        self.flip_mask = ops.flip_mask_of([f for _, f in fns_config])
        self.flat = pack_params(params).clone()
        self.lr = float(learning_rate)
        self._ws = None

    def sgd_update(self, batch):
        batch = _cuda(batch)
        from . import _lib
        need = _lib.load().flowcuda_realnvp_nll_grad_workspace_bytes(self.n_layers, self.hidden, batch.shape[0])
        if self._ws is None or self._ws.numel() * 4 < need:
            self._ws = torch.empty((need + 3) // 4, device="cuda", dtype=torch.float32)
        nll, grad = ops.realnvp_nll_grad(self.flat, self.n_layers, self.hidden, self.flip_mask, batch, lr=self.lr,
                                         params_out=self.flat, workspace=self._ws)
        return nll, grad

    @property
    def params(self):
        return unpack_params(self.flat, self.n_layers, self.hidden)


def _cuda(x):
    return torch.as_tensor(x, dtype=torch.float32).to("cuda").contiguous()
