"""GPU parity at the BASELINE.json configurations' own sizes and parameters, and against the golden
vectors produced by executing the reference's own make_flow_model (tests/golden/make_golden_flows.py).

Fixed ceilings (tests/parity_util.py): 1e-5·max(1,|ref|) for the D=2 paths on the BASELINE parameter
scale (params ~ N(0, 0.1²)), 1e-3 absolute on log_prob for the D=784 tensor-core path —
the north_star tolerances, applied to the MAX over every sample.  What each case measured is recorded
in gpurun_out/parity_report.json.
"""
import os
import sys

import numpy as np
import pytest
import torch

from oracle import flows_oracle as O
from parity_util import hold

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
sys.path.insert(0, GOLDEN)
import mnist_params as MP  # noqa: E402

FP32, TOY_TC, BF16X3 = 0, 1, 3  # FLOWCUDA_PRECISION_FP32 / _3XTF32 (fp16x3 toy kernel) / _BF16X3


def G(a):
    return torch.as_tensor(np.asarray(a), dtype=torch.float32).cuda().contiguous()


def spec_of(ospec, precision=0):
    from normalizing_flows_b200 import ops

    tr = {"affine_scalar": 0, "affine": 1, "rqs": 2}[ospec.transform]
    base = {"normal": 0, "uniform": 1}[ospec.base]
    return ops.make_spec(ospec.event_dim, ospec.num_layers, ospec.hidden, tr, ospec.num_bins, ospec.range_min,
                         ospec.range_max, base, precision)


def oracle_chunks(fn, x, chunk):
    return torch.cat([fn(x[i:i + chunk]) for i in range(0, x.shape[0], chunk)])


# ----------------------------------------------------------------- golden: the reference's own model builders
@pytest.mark.parametrize("precision", [FP32, TOY_TC])
def test_toy_flows_match_reference_make_flow_model(precision):
    """real_nvp_toy_densities.py / maf_toy_densities.py executed unmodified (numpy stand-ins for the libraries):
    same parameters (haiku tree flattened by the names the reference created), same inputs."""
    from normalizing_flows_b200 import ops

    g = np.load(os.path.join(GOLDEN, "coupling_toy_golden.npz"))
    tag = "fp32" if precision == FP32 else "tc"
    for key, ospec, cmax in (("rnvp", O.REAL_NVP_TOY, 1e-5), ("maf", O.RQS_TOY, 1.5e-4)):
        # maf golden uses stress parameters (scale 0.3, |x| up to 18): the float32 oracle itself is > 1e-5 there
        spec = spec_of(ospec, precision)
        flat, x, z = G(g[f"{key}_flat"]), G(g[f"{key}_x"]), G(g[f"{key}_z"])
        lp, _, _ = ops.coupling_log_prob(spec, flat, x)
        kind = "toy" if key == "rnvp" else "stress"
        p32 = None
        if key == "maf":
            params = _unflatten(ospec, g[f"{key}_flat"], torch.float32)
            p32 = O.coupling_flow_log_prob(ospec, params, torch.from_numpy(g[f"{key}_x"]).float())
        hold(f"golden_{key}_toy_log_prob_{tag}", lp, g[f"{key}_log_prob"], cmax, ref32=p32, kind=kind)
        xs, _ = ops.coupling_sample(spec, flat, z)
        hold(f"golden_{key}_toy_sample_{tag}", xs, g[f"{key}_sample"], 1e-5 if key == "rnvp" else 1.5e-4, kind=kind)


def _unflatten(ospec, flat, dtype):
    flat = torch.as_tensor(np.asarray(flat)).to(dtype)
    out, o = [], 0
    for _ in range(ospec.num_layers):
        layer = []
        for m, n in ospec.layer_shapes():
            W = flat[o:o + m * n].reshape(m, n); o += m * n
            b = flat[o:o + n]; o += n
            layer.append((W, b))
        out.append(layer)
    return out


@pytest.mark.parametrize("precision", [FP32, BF16X3])
def test_mnist_flow_matches_reference_make_flow_model(precision):
    """maf_mnist.py executed unmodified on 24 dequantised 28x28x1 images, all 8 layers, Uniform base incl. one image
    outside the support (-inf); parameters rebuilt from the seed the golden generator used."""
    from normalizing_flows_b200 import ops

    g = np.load(os.path.join(GOLDEN, "coupling_mnist_golden.npz"))
    tree = MP.make_tree(int(g["mnist_seed"]), 8, 784, [512, 512, 512], 19600, scale=float(g["mnist_scale"]))
    np.testing.assert_allclose(MP.tree_checksum(tree), g["mnist_checksum"], rtol=1e-12)
    flat = torch.from_numpy(MP.tree_to_flat(tree, 8, 784, [512, 512, 512], 19600)).cuda()
    spec = spec_of(O.RQS_MNIST, precision)
    x = G(g["mnist_x"]).reshape(-1, 784)
    lp, _, _ = ops.coupling_log_prob(spec, flat, x)
    tag = "fp32" if precision == FP32 else "bf16x3"
    hold(f"golden_mnist_log_prob_{tag}", lp, g["mnist_log_prob"], 1e-3, absolute=True, kind="d784")
    xs, _ = ops.coupling_sample(spec, flat, G(g["mnist_z"]).reshape(-1, 784))
    hold(f"golden_mnist_sample_{tag}", xs, g["mnist_sample"].reshape(-1, 784), 1e-4, absolute=True, kind="d784")


def test_bflow_chain_matches_reference_script():
    """bflow.py:1-34 executed unmodified builds the chain; log_prob on its own 200-point grid (bflow.py:82-88)."""
    from normalizing_flows_b200 import ops

    g = np.load(os.path.join(GOLDEN, "bflow_golden.npz"))
    thc = ops.bernstein_constrain_thetas(ops.bernstein_constrain_thetas(G(g["bflow_theta"])))
    scal = G([g["bflow_a"][0], g["bflow_b"][0], g["bflow_alpha"][0], g["bflow_beta"][0]])
    lp = ops.bernstein_flow_log_prob(thc, scal, G(g["bflow_x"]))
    hold("golden_bflow_log_prob", lp, g["bflow_log_prob"], 3e-5, kind="toy")


# ----------------------------------------------------------------- BASELINE sizes, bench parameters
def test_maf_mnist_all_8_layers_vs_oracle():
    """configs[3] (src/maf_mnist.py:28-31): all 8 layers, 256 samples, bench parameters (seed 1234, scale 0.1),
    tensor-core path and FP32 path vs the float64 oracle, <= 1e-3 absolute on every sample."""
    from normalizing_flows_b200 import ops

    ospec = O.RQS_MNIST
    params = O.init_coupling_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    gen = torch.Generator().manual_seed(3)
    x = ((torch.randint(0, 256, (256, 784), generator=gen).float() + torch.rand(256, 784, generator=gen)) / 256.0)
    p64 = [[(W.double(), b.double()) for W, b in layer] for layer in params]
    zref, ldref = O.coupling_flow_to_base(ospec, p64, x.double())
    ref = O.base_log_prob(ospec, zref) + ldref
    del p64
    for prec, tag in ((BF16X3, "bf16x3"), (FP32, "fp32")):
        lp, z, _ = ops.coupling_log_prob(spec_of(ospec, prec), flat, x.cuda(), want_z=True)
        hold(f"maf_mnist_L8_log_prob_{tag}", lp, ref, 1e-3, absolute=True, kind="d784")
        hold(f"maf_mnist_L8_z_{tag}", z, zref, 1e-4, absolute=True, kind="d784")


def test_maf_mnist_bench_shape_tc_vs_fp32_kernel():
    """n = 4096 (1248 output tiles on 148 persistent CTAs — the multi-tile walk with the TMA ring carried
    across tiles and both TMEM accumulators reused), all 8 layers, bf16x3 vs the FP32 dense kernel per sample."""
    from normalizing_flows_b200 import ops

    ospec = O.RQS_MNIST
    params = O.init_coupling_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    gen = torch.Generator().manual_seed(4)
    x = ((torch.randint(0, 256, (4096, 784), generator=gen).float() + torch.rand(4096, 784, generator=gen)) / 256.0).cuda()
    lp32, z32, _ = ops.coupling_log_prob(spec_of(ospec, FP32), flat, x, want_z=True)
    lptc, ztc, _ = ops.coupling_log_prob(spec_of(ospec, BF16X3), flat, x, want_z=True)
    hold("maf_mnist_n4096_log_prob_tc_vs_fp32", lptc, lp32, 1e-3, absolute=True, kind="d784")
    hold("maf_mnist_n4096_z_tc_vs_fp32", ztc, z32, 1e-4, absolute=True, kind="d784")
    u = torch.rand(4096, 784, generator=gen).cuda()
    s32, _ = ops.coupling_sample(spec_of(ospec, FP32), flat, u)
    stc, _ = ops.coupling_sample(spec_of(ospec, BF16X3), flat, u)
    hold("maf_mnist_n4096_sample_tc_vs_fp32", stc, s32, 1e-4, absolute=True, kind="d784")
    # twice the same call: bit-identical (no atomics, fixed reduction orders)
    lptc2, _, _ = ops.coupling_log_prob(spec_of(ospec, BF16X3), flat, x)
    assert torch.equal(lptc, lptc2)


def test_maf_mnist_bench_shape_grad_tc_vs_fp32():
    """coupling_nll_grad at the maf_mnist shape, batch 1024, 2 flow layers: tensor-core reverse mode vs the FP32 engine."""
    from normalizing_flows_b200 import ops

    ospec = O.CouplingSpec(784, 2, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform")
    params = O.init_coupling_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    gen = torch.Generator().manual_seed(5)
    x = ((torch.randint(0, 256, (1024, 784), generator=gen).float() + torch.rand(1024, 784, generator=gen)) / 256.0).cuda()
    l32, g32 = ops.coupling_nll_grad(spec_of(ospec, FP32), flat, x)
    ltc, gtc = ops.coupling_nll_grad(spec_of(ospec, BF16X3), flat, x)
    assert abs(l32.item() - ltc.item()) <= 1e-3
    scale = g32.abs().max().item()
    err = (gtc - g32).abs().max().item() / scale
    print(f"parity maf_mnist grad tc vs fp32: max |dg| / max|g| = {err:.3e}")
    assert err <= 2e-4, err


def test_maf_made_all_8_layers_vs_oracle():
    """Row #8 at its BASELINE shape (D=784, H=512, 8 layers), 256 samples, bench parameters."""
    from normalizing_flows_b200 import ops

    ospec = O.MadeSpec(784, 8, [512, 512])
    params = O.init_made_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    x = torch.randn(256, 784, generator=torch.Generator().manual_seed(6))
    p64 = [[(W.double(), b.double()) for W, b in layer] for layer in params]
    ref = O.maf_log_prob(ospec, p64, x.double())
    lp, _, _ = ops.made_log_prob(ops.make_made_spec(784, 8, [512, 512]), flat, x.cuda())
    hold("maf_made_L8_log_prob_bf16x3", lp, ref, 1e-3, absolute=True, kind="d784")
    lp32, _ = ops.made_log_prob_fp32(ops.make_made_spec(784, 8, [512, 512]), flat, x.cuda())
    hold("maf_made_L8_log_prob_fp32", lp32, ref, 1e-3, absolute=True, kind="d784")


@pytest.mark.parametrize("name,ospec_name", [("maf_toy", "RQS_TOY"), ("realnvp_toy", "REAL_NVP_TOY")])
def test_toy_tc_kernel_vs_oracle_at_2e20(name, ospec_name):
    """configs[1] / configs[0] on 2^20 samples with the bench parameters: the TENSOR-CORE kernel against the
    float64 oracle (not against the repo's FP32 kernel), north_star's 1e-5 on the max."""
    from normalizing_flows_b200 import ops

    ospec = getattr(O, ospec_name)
    params = O.init_coupling_params(ospec, seed=1234, dtype=torch.float32, scale=0.1)
    flat = O.flatten_layers(params).float().cuda()
    n = 1 << 20
    gen = torch.Generator().manual_seed(9)
    k = torch.randint(0, 8, (n,), generator=gen)
    ang = k.double() * (2 * np.pi / 8)
    x = (torch.stack([2 * torch.cos(ang), 2 * torch.sin(ang)], -1) + 0.3 * torch.randn(n, 2, generator=gen, dtype=torch.float64)).float()
    p64 = [[(W.double(), b.double()) for W, b in layer] for layer in params]
    ref = oracle_chunks(lambda v: O.coupling_flow_log_prob(ospec, p64, v.double()), x, 1 << 16)
    ref32 = oracle_chunks(lambda v: O.coupling_flow_log_prob(ospec, params, v), x, 1 << 16)
    lp, _, _ = ops.coupling_log_prob(spec_of(ospec, TOY_TC), flat, x.cuda())
    hold(f"{name}_2e20_log_prob_tc", lp, ref, 1e-5, ref32=ref32, kind="toy")
    lp0, _, _ = ops.coupling_log_prob(spec_of(ospec, FP32), flat, x.cuda())
    hold(f"{name}_2e20_log_prob_fp32", lp0, ref, 1e-5, kind="toy")
