"""Deterministic parameter trees for the golden fixtures of the distrax-built flows.

The MNIST-scale flow of src/maf_mnist.py has 87.9 M parameters (351 MB): they cannot be committed,
so the golden generator (make_golden_flows.py, build container) and the tests (anywhere) both
rebuild them from a seed with this function and only the outputs are stored.  numpy's PCG64 +
``standard_normal(dtype=float32)`` stream is part of numpy's compatibility guarantee for
``default_rng``; the golden file also stores a checksum of the generated tree so that a changed
stream fails loudly instead of comparing against stale outputs.

Shapes and names are the haiku tree of the reference's ``make_conditioner`` (module names in
creation order: ``mlp/~/linear_i``, ``linear``, ``mlp_1/~/linear_i``, ``linear_1``, ...).
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import numpy as np


def module_shapes(num_layers: int, in_features: int, hidden: Sequence[int], out_features: int) -> List[Tuple[str, int, int]]:
    sizes = [int(in_features)] + [int(h) for h in hidden] + [int(out_features)]
    out = []
    for k in range(num_layers):
        suf = "" if k == 0 else f"_{k}"
        for i, (m, n) in enumerate(zip(sizes[:-1], sizes[1:])):
            name = f"mlp{suf}/~/linear_{i}" if i < len(hidden) else f"linear{suf}"
            out.append((name, m, n))
    return out


def make_tree(seed: int, num_layers: int, in_features: int, hidden: Sequence[int], out_features: int,
              scale: float = 0.1) -> Dict[str, Dict[str, np.ndarray]]:
    """float32 tree: w ~ N(0, s^2) with s = scale/sqrt(fan_in/2) for fan_in > 8 else scale; b ~ N(0, scale^2)."""
    rng = np.random.default_rng(seed)
    tree = {}
    for name, m, n in module_shapes(num_layers, in_features, hidden, out_features):
        s = scale / math.sqrt(m / 2.0) if m > 8 else scale
        w = rng.standard_normal((m, n), dtype=np.float32) * np.float32(s)
        b = rng.standard_normal((n,), dtype=np.float32) * np.float32(scale)
        tree[name] = {"w": w, "b": b}
    return tree


def tree_checksum(tree) -> np.ndarray:
    """[sum, sum of squares, strided sample sum] in float64 — cheap fingerprint of the generated parameters."""
    s = ss = st = 0.0
    for name in tree:
        for k in ("w", "b"):
            a = tree[name][k].astype(np.float64).reshape(-1)
            s += a.sum()
            ss += (a * a).sum()
            st += a[::997].sum()
    return np.array([s, ss, st])


def tree_to_flat(tree, num_layers, in_features, hidden, out_features) -> np.ndarray:
    parts = []
    for name, m, n in module_shapes(num_layers, in_features, hidden, out_features):
        parts += [tree[name]["w"].reshape(-1), tree[name]["b"].reshape(-1)]
    return np.concatenate(parts)
