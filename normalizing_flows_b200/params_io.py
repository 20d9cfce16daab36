"""Parameter interchange with the reference's formats (SURVEY §8b "params wire format", §8f-4).

The library keeps a flow's parameters in ONE flat float32 buffer: per flow layer, per Linear,
`W[in,out]` row-major followed by `b[out]`.  The reference holds them as

  * a haiku params tree (`hk.transform(...).init`, src/real_nvp_toy_densities.py:113-136):
    `{module: {"w": [in,out], "b": [out]}}`; inside `make_flow_model` the k-th conditioner creates
    `mlp{_k}/~/linear_{i}` for its hidden Linears and `linear{_k}` for the final one (suffix absent
    for k = 0) — haiku's creation-order naming;
  * or nested lists `params[layer][sublayer] = [W, b]` (src/real_nvp.py:188-195,210-220).

`.npz` files carry the tree flattened as `"<module>/<w|b>"` keys — what `np.savez(path,
**{f"{m}/{n}": a for m, d in params.items() for n, a in d.items()})` writes on the reference side.
Pure host-side layout code: no arithmetic happens here.
"""
from __future__ import annotations

from typing import Dict, List, Mapping, Sequence, Tuple

import numpy as np
import torch


def haiku_module_names(num_layers: int, num_hidden: int) -> List[List[str]]:
    """Per flow layer, the haiku module name of every Linear in evaluation order."""
    names = []
    for k in range(num_layers):
        suf = "" if k == 0 else f"_{k}"
        names.append([f"mlp{suf}/~/linear_{i}" for i in range(num_hidden)] + [f"linear{suf}"])
    return names


def _linear_shapes(in_features: int, hidden_sizes: Sequence[int], out_features: int) -> List[Tuple[int, int]]:
    sizes = [int(in_features)] + [int(h) for h in hidden_sizes] + [int(out_features)]
    return list(zip(sizes[:-1], sizes[1:]))


def haiku_to_flat(tree: Mapping[str, Mapping[str, np.ndarray]], num_layers: int, in_features: int,
                  hidden_sizes: Sequence[int], out_features: int) -> torch.Tensor:
    """haiku params tree -> flat float32 buffer (host).  Raises KeyError / ValueError on a missing
    module or a shape that does not fit the model."""
    shapes = _linear_shapes(in_features, hidden_sizes, out_features)
    parts = []
    for mods in haiku_module_names(num_layers, len(hidden_sizes)):
        for mod, (m, n) in zip(mods, shapes):
            if mod not in tree:
                raise KeyError(f"haiku params have no module '{mod}' (have: {sorted(tree)[:6]} ...)")
            w = np.asarray(tree[mod]["w"], dtype=np.float32)
            b = np.asarray(tree[mod]["b"], dtype=np.float32)
            if w.shape != (m, n) or b.shape != (n,):
                raise ValueError(f"{mod}: expected w{(m, n)} b{(n,)}, got w{w.shape} b{b.shape}")
            parts += [w.reshape(-1), b]
    return torch.from_numpy(np.concatenate(parts))


def flat_to_haiku(flat: torch.Tensor, num_layers: int, in_features: int, hidden_sizes: Sequence[int],
                  out_features: int) -> Dict[str, Dict[str, np.ndarray]]:
    """Inverse of haiku_to_flat."""
    shapes = _linear_shapes(in_features, hidden_sizes, out_features)
    need = num_layers * sum(m * n + n for m, n in shapes)
    a = flat.detach().to("cpu", torch.float32).reshape(-1).numpy()
    if a.size != need:
        raise ValueError(f"expected {need} parameters, got {a.size}")
    tree, off = {}, 0
    for mods in haiku_module_names(num_layers, len(hidden_sizes)):
        for mod, (m, n) in zip(mods, shapes):
            w = a[off:off + m * n].reshape(m, n).copy()
            off += m * n
            b = a[off:off + n].copy()
            off += n
            tree[mod] = {"w": w, "b": b}
    return tree


def save_haiku_npz(path: str, flat: torch.Tensor, num_layers: int, in_features: int, hidden_sizes: Sequence[int],
                   out_features: int) -> None:
    tree = flat_to_haiku(flat, num_layers, in_features, hidden_sizes, out_features)
    np.savez(path, **{f"{m}/{n}": arr for m, d in tree.items() for n, arr in d.items()})


def load_haiku_npz(path: str, num_layers: int, in_features: int, hidden_sizes: Sequence[int],
                   out_features: int) -> torch.Tensor:
    """`.npz` keyed by `<haiku module>/<w|b>` -> flat float32 buffer (host; `.cuda()` it or pass it as
    `params=` to flows.make_flow_model)."""
    with np.load(path) as z:
        tree: Dict[str, Dict[str, np.ndarray]] = {}
        for key in z.files:
            mod, _, name = key.rpartition("/")
            if name not in ("w", "b") or not mod:
                raise ValueError(f"{path}: key '{key}' is not '<module>/w' or '<module>/b'")
            tree.setdefault(mod, {})[name] = z[key]
    return haiku_to_flat(tree, num_layers, in_features, hidden_sizes, out_features)


def nested_to_flat(params) -> torch.Tensor:
    """`params[layer][sublayer] = [W[in,out], b[out]]` (src/real_nvp.py:188-195) -> flat buffer."""
    parts = []
    for layer in params:
        for W, b in layer:
            W = np.asarray(W, dtype=np.float32)
            b = np.asarray(b, dtype=np.float32)
            if W.ndim != 2 or b.shape != (W.shape[1],):
                raise ValueError(f"expected [W[in,out], b[out]], got W{W.shape} b{b.shape}")
            parts += [W.reshape(-1), b]
    return torch.from_numpy(np.concatenate(parts)) if parts else torch.zeros(0)


def flat_to_nested(flat: torch.Tensor, num_layers: int, sizes: Sequence[int]):
    """flat buffer -> `params[layer][sublayer] = [W, b]` with Linear sizes `sizes` (e.g. [1, 4, 2])."""
    a = flat.detach().to("cpu", torch.float32).reshape(-1).numpy()
    shapes = list(zip(sizes[:-1], sizes[1:]))
    if a.size != num_layers * sum(m * n + n for m, n in shapes):
        raise ValueError("parameter count does not match num_layers x sizes")
    out, off = [], 0
    for _ in range(num_layers):
        layer = []
        for m, n in shapes:
            W = a[off:off + m * n].reshape(m, n).copy()
            off += m * n
            layer.append([W, a[off:off + n].copy()])
            off += n
        out.append(layer)
    return out
