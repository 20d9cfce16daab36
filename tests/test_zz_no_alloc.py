"""flowcuda_sum_scaled_static — the allocation-free variant of flowcuda_sum_scaled whose stage-1 partials live in a
file-scope ``static __device__`` array of the library (include/flowcuda.h; no counterpart in the reference).
Called straight through the C ABI (ctypes), checked against the existing
entry point on the same inputs, TWICE IN A ROW per case: the array keeps the previous call's contents, so the
second call (different data, fewer blocks) is the one that would show a stale partial.

This file sorts last on purpose (``pytest --collect-only``): it must pass both alone in a fresh process and after
every other GPU test has used the library."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

LIMIT = 1 << 26  # include/flowcuda.h: n <= 2^26 elements
TOL = 1e-5  # the tests' FP32 tolerance: |Δ| <= 1e-5 · max(1, |ref|)


@pytest.fixture(scope="module")
def lib():
    from normalizing_flows_b200 import _lib

    return _lib.load()


def _p(t):
    return C.c_void_p(t.data_ptr())


def _both(lib, v, n, scale):
    """(static-scratch result, workspace result) for the first n elements of v, on the current stream."""
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    out_a = torch.full((1,), float("nan"), device="cuda")
    out_b = torch.full((1,), float("nan"), device="cuda")
    ws = torch.empty(1024, dtype=torch.float64, device="cuda")
    rc = lib.flowcuda_sum_scaled_static(_p(v), n, scale, _p(out_a), s)
    assert rc == 0, lib.flowcuda_last_error()
    rc = lib.flowcuda_sum_scaled(_p(v), n, scale, _p(out_b), _p(ws), s)
    assert rc == 0, lib.flowcuda_last_error()
    torch.cuda.synchronize()
    return float(out_a.item()), float(out_b.item())


def _check(lib, v, n, scale):
    a, b = _both(lib, v, n, scale)
    ref = float(v[:n].double().sum().item()) * scale
    assert abs(a - b) <= TOL * max(1.0, abs(b)), (n, a, b)
    assert a == b, (n, a, b)  # same kernels, same grid rule: the same bits
    assert abs(a - ref) <= TOL * max(1.0, abs(ref)), (n, a, ref)


@pytest.mark.parametrize("n_first,n_second", [(4 * 1024 * 1024 + 17, 4096), (1 << 20, 1000), (5000, 1), (LIMIT, 12345)])
def test_two_calls_in_a_row_match_the_workspace_entry_point(lib, n_first, n_second):
    g = torch.Generator(device="cuda").manual_seed(n_first % 9973)
    v1 = torch.randn(n_first, device="cuda", generator=g) * 3.0 + 0.25
    v2 = torch.randn(max(n_second, 1), device="cuda", generator=g) - 1.5
    _check(lib, v1, n_first, -1.0 / n_first)  # the loss reduction of the training step: -mean(log_prob)
    _check(lib, v2, n_second, -1.0 / n_second)  # fewer blocks than the call before: stale partials must not leak
    _check(lib, v1, n_first, 1.0)  # and back


def test_empty_input_and_back_to_back_on_one_stream(lib):
    v = torch.ones(8192, device="cuda")
    a, b = _both(lib, v, 0, 1.0)
    assert a == 0.0 and b == 0.0
    # many calls queued on one stream without a host sync in between: stream order is what serialises the scratch
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = torch.zeros(16, device="cuda")
    for i in range(16):
        n = 8192 - 500 * i
        assert lib.flowcuda_sum_scaled_static(_p(v), n, 1.0, C.c_void_p(outs.data_ptr() + 4 * i), s) == 0
    torch.cuda.synchronize()
    assert outs.cpu().tolist() == [float(8192 - 500 * i) for i in range(16)]


def test_limit_is_enforced_before_any_launch(lib):
    v = torch.ones(16, device="cuda")
    out = torch.full((1,), 7.0, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.flowcuda_sum_scaled_static(_p(v), LIMIT + 1, 1.0, _p(out), s)  # rejected on the size alone: v is not read
    assert rc == -1
    msg = lib.flowcuda_last_error().decode()
    assert "exceeds the limit" in msg and str(LIMIT) in msg, msg
    assert lib.flowcuda_sum_scaled_static(_p(v), -1, 1.0, _p(out), s) == -1
    assert lib.flowcuda_sum_scaled_static(_p(v), 16, 1.0, None, s) == -1
    torch.cuda.synchronize()
    assert out.item() == 7.0  # nothing was launched


def test_entry_point_allocates_nothing(lib):
    """Device memory in use (driver view: cudaMemGetInfo) is unchanged by a burst of calls; the inputs/outputs are
    the caller's and are created before the measurement."""
    v = torch.randn(1 << 22, device="cuda")
    out = torch.zeros(1, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert lib.flowcuda_sum_scaled_static(_p(v), v.numel(), 1.0, _p(out), s) == 0  # module + symbol already resolved
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(64):
        assert lib.flowcuda_sum_scaled_static(_p(v), v.numel(), 1.0, _p(out), s) == 0
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, (free0, free1)
    ref = float(v.double().sum().item())
    assert abs(out.item() - ref) <= TOL * max(1.0, abs(ref))
