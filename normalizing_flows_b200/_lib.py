"""ctypes binding of libflowcuda.so (include/flowcuda.h).

The product path has NO fallback: if the shared library is missing or a call fails this
module raises.  (`oracle/` is test infrastructure and is never imported from here.)
"""
from __future__ import annotations

import ctypes as C
import os

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libflowcuda.so")

TRANSFORM_AFFINE_SCALAR, TRANSFORM_AFFINE, TRANSFORM_RQS = 0, 1, 2
BASE_NORMAL, BASE_UNIFORM01 = 0, 1
PRECISION_FP32, PRECISION_3XTF32, PRECISION_TF32, PRECISION_BF16X3 = 0, 1, 2, 3
MAX_HIDDEN_LAYERS = 4


class FlowCudaError(RuntimeError):
    pass


class CouplingSpec(C.Structure):
    """Mirror of `flowcuda_coupling_spec`."""

    _fields_ = [
        ("event_dim", C.c_int32),
        ("num_layers", C.c_int32),
        ("num_hidden", C.c_int32),
        ("hidden", C.c_int32 * MAX_HIDDEN_LAYERS),
        ("transform", C.c_int32),
        ("num_bins", C.c_int32),
        ("range_min", C.c_float),
        ("range_max", C.c_float),
        ("base", C.c_int32),
        ("precision", C.c_int32),
    ]


class MadeSpec(C.Structure):
    """Mirror of `flowcuda_made_spec`."""

    _fields_ = [
        ("event_dim", C.c_int32),
        ("num_layers", C.c_int32),
        ("num_hidden", C.c_int32),
        ("hidden", C.c_int32 * MAX_HIDDEN_LAYERS),
        ("base", C.c_int32),
        ("precision", C.c_int32),
    ]


_p = C.c_void_p
_i = C.c_int
_i64 = C.c_int64
_u64 = C.c_uint64
_f = C.c_float
_d = C.c_double
_sz = C.c_size_t
# flowcuda_layer_grad_fn: void (*)(void* user, int layer, int64_t offset, int64_t count, void* stream)
LAYER_GRAD_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int64, C.c_int64, C.c_void_p)
_spec = C.POINTER(CouplingSpec)
_mspec = C.POINTER(MadeSpec)

# name -> (restype, argtypes); every symbol include/flowcuda.h declares
SIGNATURES = {
    "flowcuda_version": (_i, []),
    "flowcuda_last_error": (C.c_char_p, []),
    "flowcuda_sm_count": (_i, []),
    "flowcuda_realnvp_log_prob": (_i, [_p, _i, _i, _u64, _p, _p, _p, _i64, _p]),
    "flowcuda_realnvp_sample": (_i, [_p, _i, _i, _u64, _p, _p, _i64, _p]),
    "flowcuda_realnvp_forward": (_i, [_p, _i, _i, _p, _p, _i64, _p]),
    "flowcuda_realnvp_inverse": (_i, [_p, _i, _i, _p, _p, _p, _i64, _p]),
    "flowcuda_realnvp_nll_grad_workspace_bytes": (_sz, [_i, _i, _i64]),
    "flowcuda_realnvp_nll_grad": (_i, [_p, _i, _i, _u64, _p, _i64, _p, _p, _f, _p, _p, _sz, _p]),
    "flowcuda_bernstein_constrain_thetas": (_i, [_p, _p, _i64, _i, _p]),
    "flowcuda_bernstein_forward_and_log_det": (_i, [_p, _i, _i64, _p, _p, _p, _i64, _p]),
    "flowcuda_bernstein_inverse": (_i, [_p, _i, _i64, _i, _p, _p, _i64, _p, _p]),
    "flowcuda_bernstein_flow_log_prob": (_i, [_p, _i, _p, _p, _p, _i64, _p]),
    "flowcuda_bernstein_vjp_workspace_bytes": (_sz, [_i, _i64, _i64]),
    "flowcuda_bernstein_forward_and_log_det_vjp": (_i, [_p, _i, _i64, _p, _p, _p, _p, _p, _i64, _p, _sz, _p]),
    "flowcuda_bernstein_constrain_thetas_vjp": (_i, [_p, _p, _p, _i64, _i, _p]),
    "flowcuda_coupling_param_count": (_i64, [_spec]),
    "flowcuda_coupling_workspace_bytes": (_sz, [_spec, _i64]),
    "flowcuda_coupling_log_prob": (_i, [_spec, _p, _p, _p, _p, _p, _i64, _i, _i, _i, _p, _sz, _p]),
    "flowcuda_coupling_sample": (_i, [_spec, _p, _p, _p, _p, _i64, _i, _i, _p, _sz, _p]),
    "flowcuda_coupling_nll_grad_workspace_bytes": (_sz, [_spec, _i64]),
    "flowcuda_coupling_nll_grad": (_i, [_spec, _p, _p, _i64, _p, _p, _p, _sz, _p]),
    "flowcuda_coupling_nll_grad_layers": (_i, [_spec, _p, _p, _i64, _p, _p, _p, _sz, _p, _p, _p]),
    "flowcuda_made_param_count": (_i64, [_mspec]),
    "flowcuda_made_workspace_bytes": (_sz, [_mspec, _i64]),
    "flowcuda_made_log_prob": (_i, [_mspec, _p, _p, _p, _p, _p, _i64, _i, _p, _sz, _p]),
    "flowcuda_made_sample": (_i, [_mspec, _p, _p, _p, _i64, _p]),
    "flowcuda_made_log_prob_fp32_workspace_bytes": (_sz, [_mspec, _i64]),
    "flowcuda_made_log_prob_fp32": (_i, [_mspec, _p, _p, _p, _p, _i64, _p, _sz, _p]),
    "flowcuda_made_nll_grad_workspace_bytes": (_sz, [_mspec, _i64]),
    "flowcuda_coupling_pack_weights": (_i, [_spec, _p, _p, _sz, _p]),
    "flowcuda_coupling_log_prob_packed": (_i, [_spec, _p, _p, _p, _p, _p, _i64, _i, _i, _i, _p, _sz, _p]),
    "flowcuda_coupling_sample_packed": (_i, [_spec, _p, _p, _p, _p, _i64, _i, _i, _p, _sz, _p]),
    "flowcuda_made_pack_weights": (_i, [_mspec, _p, _p, _sz, _p]),
    "flowcuda_made_log_prob_packed": (_i, [_mspec, _p, _p, _p, _p, _p, _i64, _i, _p, _sz, _p]),
    "flowcuda_made_nll_grad": (_i, [_mspec, _p, _p, _i64, _p, _p, _p, _sz, _p]),
    "flowcuda_density_sample": (_i, [_i, _u64, _u64, _i, _f, _p, _i64, _p]),
    "flowcuda_dequantize_u8": (_i, [_p, _u64, _u64, _p, _i64, _p]),
    "flowcuda_base_noise": (_i, [_i, _u64, _u64, _p, _i64, _p]),
    "flowcuda_rotate2d": (_i, [_p, _f, _f, _i, _p, _i64, _p]),
    "flowcuda_sum_scaled": (_i, [_p, _i64, _d, _p, _p, _p]),
    "flowcuda_sum_scaled_static": (_i, [_p, _i64, _d, _p, _p]),
    "flowcuda_sgd_step": (_i, [_p, _p, _i64, _f, _f, _p]),
    "flowcuda_adam_step": (_i, [_p, _p, _p, _p, _i64, _i64, _f, _f, _f, _f, _f, _p]),
    "flowcuda_affine_elementwise": (_i, [_p, _p, _p, _i, _i, _p, _p, _i64, _i64, _i64, _p]),
    "flowcuda_softclip_elementwise": (_i, [_p, _f, _f, _f, _i, _p, _p, _i64, _p]),
    "flowcuda_rqs_elementwise": (_i, [_p, _p, _i, _f, _f, _i, _p, _p, _i64, _i64, _i64, _p]),
    "flowcuda_base_log_prob": (_i, [_p, _p, _p, _i, _p, _i64, _i64, _i64, _p]),
    "flowcuda_sum_last": (_i, [_p, _p, _i64, _i64, _p]),
}

# include/flowcuda_selftest.h — diagnostics of the tcgen05 primitives, in a separate TEST library
SELFTEST_LIB_PATH = os.path.join(PKG, "libflowcuda_selftest.so")
SELFTEST_SIGNATURES = {
    "flowcuda_selftest_last_error": (C.c_char_p, []),
    "flowcuda_selftest_umma_workspace_bytes": (_sz, [_i, _i]),
    "flowcuda_selftest_mma_timing": (_i, [_p, _i, _p]),
    "flowcuda_selftest_umma": (_i, [_p, _p, _p, _p, _i, _i, _i, _p, _sz, _p]),
}

_lib = None
_selftest = None


def load_selftest() -> C.CDLL:
    """The diagnostics library (tests/ and tools/ only)."""
    global _selftest
    if _selftest is None:
        if not os.path.exists(SELFTEST_LIB_PATH):
            raise FlowCudaError(f"{SELFTEST_LIB_PATH} is missing: build it with `python -m normalizing_flows_b200.build`")
        lib = C.CDLL(SELFTEST_LIB_PATH)
        for name, (res, args) in SELFTEST_SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _selftest = lib
    return _selftest


def load() -> C.CDLL:
    """Load libflowcuda.so once; raise loudly when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FlowCudaError(
            f"{LIB_PATH} is missing: build it with `python -m normalizing_flows_b200.build` "
            "(there is no CPU or PyTorch fallback for the flow hot path)"
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so is stale
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().flowcuda_last_error().decode("utf-8", "replace")
        raise FlowCudaError(f"{what or 'flowcuda'} failed with code {rc}: {msg}")
