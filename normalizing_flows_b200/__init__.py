"""normalizing_flows_b200 — B200-native (sm_100a) flow density / sampling / training hot path
behind the reference's Bijector / functional API (kaijennissen/Normalizing_Flows).

Layout: csrc/ (CUDA kernels + the C ABI of include/flowcuda.h), _lib.py (ctypes binding),
ops.py (tensor-level calls), and the host-side mirrors of the reference interface
(real_nvp.py, distrax_api.py, BernsteinBijector.py, flows.py)."""
from . import _lib  # noqa: F401

__all__ = ["_lib"]
__version__ = "0.1.0"
