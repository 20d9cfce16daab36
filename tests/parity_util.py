"""Shared helpers of the GPU parity tests: error statistics against the float64 oracle, FIXED numeric
ceilings (no tolerance derived from the data under test), and a machine-readable record of the bound
each case actually achieved (gpurun_out/parity_report.json, summarised in DESIGN.md "Parity achieved").

Contract (BASELINE.json north_star):
  * FP32 toy paths: per-sample |Δ| <= 1e-5 · max(1, |ref|)            ("toy" ceiling)
  * tensor-core D=784 path: per-sample |Δ log_prob| <= 1e-3 absolute   ("d784" ceiling)
Cases whose float32 *reference arithmetic itself* (the oracle evaluated in float32, i.e. what the JAX
reference computes) is further than 1e-5 from float64 truth are labelled "stress": they get their own
fixed ceiling, written next to the float32 oracle's measured error, and are listed as deviations in
DESIGN.md.  Nothing here widens a ceiling at run time.
"""
from __future__ import annotations

import json
import os

import numpy as np

_REPORT = {}


def stats(got, ref, absolute=False):
    """max and p99.9 of |got-ref| / max(1,|ref|) (or absolute) over the finite reference entries."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert got.shape == ref.shape, (got.shape, ref.shape)
    ok = np.isfinite(ref)
    assert np.array_equal(got[~ok], ref[~ok]) or not (~ok).any(), "non-finite entries differ"
    assert np.all(np.isfinite(got[ok])), "non-finite result where the reference is finite"
    e = np.abs(got[ok] - ref[ok])
    if not absolute:
        e = e / np.maximum(1.0, np.abs(ref[ok]))
    if e.size == 0:
        return 0.0, 0.0
    return float(e.max()), float(np.quantile(e, 0.999))


def hold(name, got, ref, ceiling_max, ceiling_p999=None, absolute=False, ref32=None, kind="toy"):
    """Assert the fixed ceilings and record what was measured (and, when given, the float32 oracle's own error)."""
    if hasattr(got, "detach"):
        got = got.detach().double().cpu().numpy()
    if hasattr(ref, "detach"):
        ref = ref.detach().double().cpu().numpy()
    mx, p = stats(got, ref, absolute)
    rec = {"max": mx, "p99.9": p, "ceiling_max": ceiling_max, "ceiling_p99.9": ceiling_p999, "n": int(np.size(ref)),
           "absolute": absolute, "kind": kind}
    if ref32 is not None:
        if hasattr(ref32, "detach"):
            ref32 = ref32.detach().double().cpu().numpy()
        m32, p32 = stats(ref32, ref, absolute)
        rec["oracle_fp32_max"], rec["oracle_fp32_p99.9"] = m32, p32
    _REPORT[name] = rec
    print(f"parity {name}: max {mx:.3e} (ceiling {ceiling_max:.1e})  p99.9 {p:.3e}"
          + (f" (ceiling {ceiling_p999:.1e})" if ceiling_p999 else "")
          + (f"  | float32 oracle: max {rec['oracle_fp32_max']:.3e} p99.9 {rec['oracle_fp32_p99.9']:.3e}" if ref32 is not None else ""))
    assert mx <= ceiling_max, f"{name}: max err {mx:.3e} > {ceiling_max:.1e}"
    if ceiling_p999 is not None:
        assert p <= ceiling_p999, f"{name}: p99.9 err {p:.3e} > {ceiling_p999:.1e}"
    return mx, p


def dump():
    if not _REPORT:
        return
    root = os.environ.get("GRAFT_REPO_ROOT") or os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = os.path.join(root, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        path = os.path.join(out, "parity_report.json")
        old = {}
        if os.path.exists(path):
            try:
                old = json.load(open(path))
            except Exception:
                old = {}
        old.update(_REPORT)
        json.dump(old, open(path, "w"), indent=1, sort_keys=True)
    except OSError:
        pass
