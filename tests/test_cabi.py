"""CPU-only: the C-ABI library builds, loads and exports every symbol include/flowcuda.h
declares; host-side helpers behave.  No compute call is made (no GPU here)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from normalizing_flows_b200 import _lib, build

    build.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    from normalizing_flows_b200 import _lib

    hdr = open(os.path.join(ROOT, "include", "flowcuda.h")).read()
    declared = set(re.findall(r"\b(flowcuda_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations found"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), name
    for dbg in ("flowcuda_debug_set_dt_trace", "flowcuda_debug_set_dt_trace_made"):
        assert not hasattr(raw, dbg), f"{dbg}: a FLOWCUDA_BUILD_TRACE diagnostics build is in place of the product library"
    assert lib.flowcuda_version() == 100
    # diagnostics live in their own test library with their own header; the product library does not export them
    hdr2 = open(os.path.join(ROOT, "include", "flowcuda_selftest.h")).read()
    declared2 = set(re.findall(r"\b(flowcuda_[a-z0-9_]+)\s*\(", hdr2))
    assert declared2 == set(_lib.SELFTEST_SIGNATURES), declared2 ^ set(_lib.SELFTEST_SIGNATURES)
    raw2 = ctypes.CDLL(_lib.SELFTEST_LIB_PATH)
    for name in declared2:
        assert hasattr(raw2, name), name
        assert not hasattr(raw, name) or name == "flowcuda_last_error", f"{name} must not ship in libflowcuda.so"


def test_spec_struct_matches_header():
    from normalizing_flows_b200 import _lib

    # 3 ints + 4 ints + 2 ints + 2 floats + 2 ints = 13 x 4 bytes
    assert ctypes.sizeof(_lib.CouplingSpec) == 13 * 4


def test_host_helpers_and_argument_errors(lib):
    from normalizing_flows_b200 import _lib, ops

    assert ops.flip_mask_of([False, True, False, True]) == 0b1010
    sp = ops.make_spec(2, 8, [64, 64], _lib.TRANSFORM_RQS, 12, -8.0, 8.0)
    assert lib.flowcuda_coupling_param_count(ctypes.byref(sp)) == 8 * (2 * 64 + 64 + 64 * 64 + 64 + 64 * 74 + 74)
    sp2 = ops.make_spec(784, 8, [512, 512, 512], _lib.TRANSFORM_RQS, 8, 0.0, 1.0, _lib.BASE_UNIFORM01)
    assert lib.flowcuda_coupling_param_count(ctypes.byref(sp2)) == 8 * (784 * 512 + 512 + 2 * (512 * 512 + 512) + 512 * 19600 + 19600)
    bad = ops.make_spec(1, 8, [64], 0)
    assert lib.flowcuda_coupling_param_count(ctypes.byref(bad)) == -1
    assert b"event_dim" in lib.flowcuda_last_error()
    # argument validation happens before any CUDA call
    assert lib.flowcuda_realnvp_log_prob(None, 8, 4, 0, None, None, None, 16, None) == -1
    assert lib.flowcuda_bernstein_forward_and_log_det(None, 9, 0, None, None, None, 4, None) == -1
    import torch

    with pytest.raises(_lib.FlowCudaError):
        ops.realnvp_log_prob(torch.zeros(18), 1, 4, 0, torch.zeros(4, 2))  # CPU tensors: no CPU path


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "normalizing_flows_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert "oracle" not in src.replace("(`oracle/`", "").replace("oracle/", ""), fn


def test_xla_ffi_adapter_is_inert_without_the_header(tmp_path):
    """integration/flowcuda_xla_ffi.cc (the jax.ffi handlers a maintainer builds on a machine that has JAX) must
    preprocess to nothing where xla/ffi/api/ffi.h does not exist — the only state that can be checked here — and the
    build must not try to compile it then."""
    import subprocess

    from normalizing_flows_b200 import build

    src = os.path.join(ROOT, "integration", "flowcuda_xla_ffi.cc")
    obj = str(tmp_path / "adapter.o")
    r = subprocess.run(["g++", "-std=c++17", "-c", "-I", os.path.join(ROOT, "include"), src, "-o", obj], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    syms = subprocess.run(["nm", obj], capture_output=True, text=True).stdout
    assert "_ffi" not in syms
    if build._xla_ffi_include() is None:
        assert build.build_xla_ffi_adapter() is None
