"""GPU: the hand-written tcgen05 primitives (csrc/tc.cuh) on one 128-row tile vs a float64
matmul — operand tile images, TMA bulk copy, kind::tf32 / kind::f16 MMAs in 1-pass and
error-compensated 3-pass form, TMEM loads at aligned and unaligned columns."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("mode,tol", [(0, 2e-6), (1, 3e-5), (2, 2e-3), (3, 1e-2), (4, 2e-6), (5, 3e-5)])
@pytest.mark.parametrize("n,k", [(64, 64), (48, 64), (16, 32), (256, 64), (208, 128)])
def test_umma_tile_matches_float64(mode, tol, n, k):
    from normalizing_flows_b200 import ops

    if mode in (0, 2, 4) and (n, k) == (208, 128):
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


# ------------------------------------------------------------------ 3xTF32 toy coupling stacks (rows #2/#3)
import numpy as np

from oracle import flows_oracle as O
from parity_util import hold
from test_gpu_parity import CEIL, COUPLING_CASES, G, _spec_pair, close, kind_of

TC_TOY = ["realnvp_toy", "rqs_toy", "rqs_toy_k8", "affine_toy"]


def close_tc(got, ref64, ref32, what, case, n):
    """Fixed per-case ceilings (test_gpu_parity.CEIL): max and p99.9 of |Δ| / max(1,|ref|); the float32 oracle's own
    error on the same inputs is recorded beside the measurement (gpurun_out/parity_report.json)."""
    cm, cp = CEIL[case]
    hold(f"tc_{case}_n{n}_{what}", got, ref64, cm, cp, ref32=ref32, kind=kind_of(case))


@pytest.mark.parametrize("case", TC_TOY)
@pytest.mark.parametrize("n", [1, 77, 1024 + 3, 70_001, 300_000])
def test_toy_coupling_3xtf32_parity(case, n):
    """Same fixed ceilings as the FP32 path (test_gpu_parity.CEIL)."""
    from normalizing_flows_b200 import _lib, ops

    ospec, scale, xs = COUPLING_CASES[case]
    params = O.init_coupling_params(ospec, seed=11, scale=scale)
    flat = O.flatten_layers(params)
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_3XTF32)
    gen = torch.Generator().manual_seed(n)
    x = (torch.randn(n, 2, generator=gen, dtype=torch.float64) * xs).float()
    x64 = x.double()
    p32 = [[(W.float(), b.float()) for (W, b) in layer] for layer in params]
    lp, z, ld = ops.coupling_log_prob(spec, G(flat), x.cuda(), want_z=True, want_log_det=True)
    zref, ldref = O.coupling_flow_to_base(ospec, params, x64)
    z32, ld32 = O.coupling_flow_to_base(ospec, p32, x)
    lpref = O.coupling_flow_log_prob(ospec, params, x64)
    close_tc(lp, lpref.numpy(), O.coupling_flow_log_prob(ospec, p32, x).numpy(), "log_prob", case, n)
    close_tc(z, zref.numpy(), z32.numpy(), "z", case, n)
    close_tc(ld, ldref.numpy(), ld32.numpy(), "log_det", case, n)
    xsamp, _ = ops.coupling_sample(spec, G(flat), x.cuda())
    sref = O.coupling_flow_sample(ospec, params, x64)
    close_tc(xsamp, sref.numpy(), O.coupling_flow_sample(ospec, p32, x).numpy(), "sample", case, n)
    if n > 1:
        _, z1, ld1 = ops.coupling_log_prob(spec, G(flat), x.cuda(), want_z=True, want_log_det=True, want_log_prob=False,
                                           layers=(1, 2), with_base=False)
        zr1, lr1 = O.masked_coupling(ospec, params[1], 1, x64, inverse=False)
        zs1, ls1 = O.masked_coupling(ospec, p32[1], 1, x, inverse=False)
        close_tc(z1, zr1.numpy(), zs1.numpy(), "layer1_z", case, n); close_tc(ld1, lr1.numpy(), ls1.numpy(), "layer1_log_det", case, n)


def test_toy_coupling_3xtf32_matches_fp32_kernel_at_scale():
    """2^22 samples (BASELINE configs[1] size): the tensor-core kernel against the FP32 CUDA-core
    kernel on the bench parameters; identical inputs, per-sample."""
    from normalizing_flows_b200 import _lib, ops

    ospec = O.RQS_TOY
    params = O.init_coupling_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    x = (torch.randn(1 << 22, 2, generator=torch.Generator().manual_seed(5)) * 2.0).cuda()
    lp32, _, _ = ops.coupling_log_prob(_spec_pair(ops, ospec), flat, x)
    lptc, _, _ = ops.coupling_log_prob(_spec_pair(ops, ospec, precision=_lib.PRECISION_3XTF32), flat, x)
    err = ((lptc - lp32).abs() / lp32.abs().clamp(min=1)).max().item()
    assert err <= 2e-5, err


# ------------------------------------------------------------------ bf16x3 dense coupling stacks (row #4, C5)
DENSE_TC = {
    "rqs_d40_uniform": (O.CouplingSpec(40, 2, [32, 32, 32], "rqs", 8, 0.0, 1.0, "uniform"), 0.3, None),
    "rqs_d7_k8": (O.CouplingSpec(7, 3, [24, 24], "rqs", 8, -3.0, 3.0), 0.3, 2.0),
    "affine_d6": (O.CouplingSpec(6, 3, [16], "affine"), 0.3, 1.5),
    "affine_d300": (O.CouplingSpec(300, 2, [320, 64], "affine"), 0.2, 1.0),
    "rqs_d50_wide": (O.CouplingSpec(50, 2, [300], "rqs", 8, -4.0, 4.0), 0.3, 2.0),
}


@pytest.mark.parametrize("case", sorted(DENSE_TC))
@pytest.mark.parametrize("n", [1, 77, 300])
def test_dense_coupling_bf16x3_parity(case, n):
    """Tensor-core path of the wide-conditioner flows; north_star tolerance for this path is 1e-3
    absolute on log_prob — held here at 2e-4·max(1,|ref|) on every output."""
    from normalizing_flows_b200 import _lib, ops

    ospec, scale, xs = DENSE_TC[case]
    params = O.init_coupling_params(ospec, seed=11, scale=scale)
    flat = O.flatten_layers(params)
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3)
    gen = torch.Generator().manual_seed(n)
    if xs is None:
        x = torch.rand(n, ospec.event_dim, generator=gen, dtype=torch.float64).float()
        if n > 3:
            x[3, 1] = 1.5  # outside the Uniform support -> -inf
    else:
        x = (torch.randn(n, ospec.event_dim, generator=gen, dtype=torch.float64) * xs).float()
    x64 = x.double()
    lp, z, ld = ops.coupling_log_prob(spec, G(flat), x.cuda(), want_z=True, want_log_det=True)
    zref, ldref = O.coupling_flow_to_base(ospec, params, x64)
    lpref = O.coupling_flow_log_prob(ospec, params, x64)
    close(lp, lpref.numpy(), tol=2e-4, what="log_prob", nonfinite="equal")
    close(z, zref.numpy(), tol=2e-4, what="z")
    close(ld, ldref.numpy(), tol=2e-4, what="log_det")
    xsamp, _ = ops.coupling_sample(spec, G(flat), x.cuda())
    close(xsamp, O.coupling_flow_sample(ospec, params, x64).numpy(), tol=2e-4, what="sample")


@pytest.mark.parametrize("mode", ["pair", "quad"])
@pytest.mark.parametrize("case", ["rqs_d40_uniform", "affine_d300", "rqs_d50_wide"])
def test_dense_coupling_cta_pair_kernels(case, mode, monkeypatch):
    """The same flows through the CTA-pair GEMM (tcgen05 cta_group::2, tensor-map TMA; `quad`: 4-CTA clusters with the A
    operand multicast) — csrc/dense_tc2.cuh; FLOWCUDA_DENSE_TC forces one kernel for every launch, unset the driver
    takes pairs for the RQS transform Linear and for hidden Linears of multi-wave batches."""
    from normalizing_flows_b200 import _lib, ops

    monkeypatch.setenv("FLOWCUDA_DENSE_TC", mode)
    ospec, scale, xs = DENSE_TC[case]
    params = O.init_coupling_params(ospec, seed=11, scale=scale)
    flat = O.flatten_layers(params)
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3)
    n = 700  # three pair tiles, the last one ragged
    gen = torch.Generator().manual_seed(n)
    x = (torch.rand(n, ospec.event_dim, generator=gen, dtype=torch.float64) if xs is None
         else torch.randn(n, ospec.event_dim, generator=gen, dtype=torch.float64) * xs).float()
    lp, z, ld = ops.coupling_log_prob(spec, G(flat), x.cuda(), want_z=True, want_log_det=True)
    zref, ldref = O.coupling_flow_to_base(ospec, params, x.double())
    close(lp, O.coupling_flow_log_prob(ospec, params, x.double()).numpy(), tol=2e-4, what="log_prob", nonfinite="equal")
    close(z, zref.numpy(), tol=2e-4, what="z")
    close(ld, ldref.numpy(), tol=2e-4, what="log_det")
    xsamp, _ = ops.coupling_sample(spec, G(flat), x.cuda())  # the inverse epilogues of the same kernels
    close(xsamp, O.coupling_flow_sample(ospec, params, x.double()).numpy(), tol=2e-4, what="sample")
    monkeypatch.setenv("FLOWCUDA_DENSE_TC", "1cta")
    lp1, _, _ = ops.coupling_log_prob(spec, G(flat), x.cuda())
    assert (lp1 - lp).abs().max().item() <= 1e-5 * max(1.0, lp.abs().max().item())  # same arithmetic, other tiling
    monkeypatch.delenv("FLOWCUDA_DENSE_TC")  # the default policy: pairs for the RQS transform Linear only at this size
    lp2, _, _ = ops.coupling_log_prob(spec, G(flat), x.cuda())
    assert (lp2 - lp).abs().max().item() <= 1e-5 * max(1.0, lp.abs().max().item())


def test_hidden_chain_cluster_kernel_matches_per_linear_kernels(monkeypatch):
    """The three 512-wide hidden Linears of a conditioner as one 4-CTA-cluster kernel (default at batches of one wave of
    clusters; FLOWCUDA_HIDDEN_CHAIN = 0 / 1 forces never / always) with the
    activation stage multicast and the weight images in 128-row tiles (csrc/dense_tc_chain.cuh) — same products in the
    same order as the per-Linear kernels, so log_prob, z and the gradient must be the same bits; ragged batch
    (three m tiles, the last one partial) and the prepared-weights path included."""
    from normalizing_flows_b200 import _lib, ops

    ospec = O.CouplingSpec(784, 3, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform")
    flat = O.flatten_layers(O.init_coupling_params(ospec, seed=9, scale=0.1))
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3)
    x = torch.rand(300, 784, generator=torch.Generator().manual_seed(4), dtype=torch.float64).float().cuda()
    p = G(flat)
    monkeypatch.setenv("FLOWCUDA_HIDDEN_CHAIN", "0")  # per-Linear kernels (the default above one wave of clusters)
    lp0, z0, _ = ops.coupling_log_prob(spec, p, x, want_z=True, want_log_det=True)
    loss0, grad0 = ops.coupling_nll_grad(spec, p, x)
    monkeypatch.setenv("FLOWCUDA_HIDDEN_CHAIN", "1")
    lp1, z1, _ = ops.coupling_log_prob(spec, p, x, want_z=True, want_log_det=True)
    loss1, grad1 = ops.coupling_nll_grad(spec, p, x)
    assert torch.equal(lp0, lp1) and torch.equal(z0, z1)
    assert torch.equal(loss0, loss1) and torch.equal(grad0, grad1)
    monkeypatch.delenv("FLOWCUDA_HIDDEN_CHAIN")  # default: by batch size — a workspace prepared once serves both kernels
    model_p = ops.coupling_pack_weights(spec, p, max_batch=20000) if hasattr(ops, "coupling_pack_weights") else None
    if model_p is not None:
        for n in (300, 20000):  # one wave of clusters -> chain kernel; 157 m tiles -> per-Linear kernels
            xb = torch.rand(n, 784, generator=torch.Generator().manual_seed(n), dtype=torch.float64).float().cuda()
            a, _, _ = ops.coupling_log_prob(spec, p, xb, packed=model_p)
            b, _, _ = ops.coupling_log_prob(spec, p, xb)
            assert torch.equal(a, b)


def test_dense_coupling_bf16x3_mnist_shape():
    """maf_mnist.py:28-31 conditioner 784->512->512->512->19600, 8 bins, Uniform base (2 flow layers to
    bound the CPU oracle's time): |log_prob - float64 oracle| <= 1e-3 absolute (north_star)."""
    from normalizing_flows_b200 import _lib, ops

    ospec = O.CouplingSpec(784, 2, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform")
    params = O.init_coupling_params(ospec, seed=5, scale=0.1)
    gen = torch.Generator().manual_seed(0)
    x = ((torch.randint(0, 256, (200, 784), generator=gen).double() + torch.rand(200, 784, generator=gen, dtype=torch.float64)) / 256).float()
    lp, _, _ = ops.coupling_log_prob(_spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3), G(O.flatten_layers(params)), x.cuda())
    ref = O.coupling_flow_log_prob(ospec, params, x.double()).numpy()
    err = np.abs(lp.double().cpu().numpy() - ref).max()
    assert err <= 1e-3, err


def test_deep16_affine_stack_d784():
    """configs[4] / SURVEY C5: 16 per-dimension affine couplings at D=784, conditioner 784->512->512->1568,
    all 16 layers: |log_prob - float64 oracle| <= 1e-3 absolute; sample inverts log_prob's z."""
    from normalizing_flows_b200 import _lib, ops

    ospec = O.CouplingSpec(784, 16, [512, 512], "affine")
    params = O.init_coupling_params(ospec, seed=7, scale=0.1)
    flat = G(O.flatten_layers(params))
    gen = torch.Generator().manual_seed(0)
    x = ((torch.randint(0, 256, (96, 784), generator=gen).double() + torch.rand(96, 784, generator=gen, dtype=torch.float64)) / 256).float()
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3)
    lp, z, _ = ops.coupling_log_prob(spec, flat, x.cuda(), want_z=True)
    ref = O.coupling_flow_log_prob(ospec, params, x.double()).numpy()
    err = np.abs(lp.double().cpu().numpy() - ref).max()
    assert err <= 1e-3, err
    xb, _ = ops.coupling_sample(spec, flat, z)
    assert (xb.cpu() - x).abs().max().item() <= 1e-3
