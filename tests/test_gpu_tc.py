"""GPU: the hand-written tcgen05 primitives (csrc/tc.cuh) on one 128-row tile vs a float64
matmul — operand tile images, TMA bulk copy, kind::tf32 / kind::f16 MMAs in 1-pass and
error-compensated 3-pass form, TMEM loads at aligned and unaligned columns."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,tol", [(0, 2e-6), (1, 3e-5), (2, 2e-3), (3, 1e-2)])
@pytest.mark.parametrize("n,k", [(64, 64), (48, 64), (16, 32), (256, 64), (208, 128)])
def test_umma_tile_matches_float64(mode, tol, n, k):
    from normalizing_flows_b200 import ops

    if mode in (0, 2) and (n, k) == (208, 128):
        pytest.skip("tile image exceeds shared memory in fp32 words")
    g = torch.Generator().manual_seed(n * 1000 + k + mode)
    a = torch.randn(128, k, generator=g)
    w = torch.randn(k, n, generator=g)
    d, d2 = ops.selftest_umma(a.cuda(), w.cuda(), mode, want_unaligned=True)
    torch.cuda.synchronize()
    ref = a.double() @ w.double()
    scale = (a.double().abs() @ w.double().abs())  # per-entry magnitude of the summands
    err = ((d.double().cpu() - ref).abs() / scale).max().item()
    assert err <= tol, f"mode {mode} n={n} k={k}: scaled err {err:.3e} > {tol}"
    m = 25 * (n // 25)
    assert torch.equal(d2[:, :m], d[:, :m]), "unaligned-column tcgen05.ld differs from the aligned read"
