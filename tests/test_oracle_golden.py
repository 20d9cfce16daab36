"""Pin the oracle (rows #1, #5) against golden vectors produced by the reference's
own source (tests/golden/make_golden.py).  CPU only."""
import os

import numpy as np
import torch

from oracle import flows_oracle as O

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
T = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.float64)


def unflatten_nvp(flat, L, H):
    flat = T(flat)
    out, o = [], 0
    for _ in range(L):
        W1 = flat[o : o + H].reshape(1, H); o += H
        b1 = flat[o : o + H]; o += H
        W2 = flat[o : o + 2 * H].reshape(H, 2); o += 2 * H
        b2 = flat[o : o + 2]; o += 2
        out.append([[W1, b1], [W2, b2]])
    return out


def test_real_nvp_oracle_matches_reference_source():
    g = np.load(os.path.join(GOLDEN, "real_nvp_golden.npz"))
    L, H = int(g["n_layers"]), int(g["hidden"])
    params = unflatten_nvp(g["params_flat"], L, H)
    flips = [bool(f) for f in g["flips"]]
    y, z = T(g["y"]), T(g["z"])
    np.testing.assert_allclose(O.log_prob_nvp_chains(params, flips, y).numpy(), g["log_prob"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(O.sample_nvp_chain(params, flips, z).numpy(), g["sample"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(O.forward_nvp(params[0], z, False).numpy(), g["fwd0"], rtol=1e-13)
    np.testing.assert_allclose(O.forward_nvp(params[1], z, True).numpy(), g["fwd1"], rtol=1e-13)
    x0, ls0 = O.inverse_nvp(params[0], y, False)
    x1, ls1 = O.inverse_nvp(params[1], y, True)
    np.testing.assert_allclose(x0.numpy(), g["inv0"], rtol=1e-13)
    np.testing.assert_allclose(ls0.numpy(), g["ls0"], rtol=1e-13)
    np.testing.assert_allclose(x1.numpy(), g["inv1"], rtol=1e-13)
    np.testing.assert_allclose(ls1.numpy(), g["ls1"], rtol=1e-13)
    lp = O.log_prob_nvp(params[3], O.log_prob_N01, y, flip=True)
    np.testing.assert_allclose(lp.numpy(), g["log_prob_layer3"], rtol=1e-12)


def test_bernstein_oracle_matches_reference_source():
    g = np.load(os.path.join(GOLDEN, "bernstein_golden.npz"))
    x = T(g["x"])
    for name in ("demo", "tf", "m32", "m64", "m2", "m8"):
        th = O.constrain_thetas(T(g[f"{name}_theta_raw"]))
        np.testing.assert_allclose(th.numpy(), g[f"{name}_theta_constrained"], rtol=1e-13, atol=1e-13)
        yv, ld = O.bernstein_forward_and_log_det(th, x)
        np.testing.assert_allclose(yv.numpy(), g[f"{name}_y"], rtol=1e-9, atol=1e-10)
        np.testing.assert_allclose(ld.numpy(), g[f"{name}_ld"], rtol=1e-9, atol=1e-9)
    np.testing.assert_allclose(O.constrain_thetas(T(g["batch_theta_raw"])).numpy(), g["batch_theta_constrained"], rtol=1e-13)
    dd = O.constrain_thetas(O.constrain_thetas(T(g["demo_theta_raw"])))
    np.testing.assert_allclose(dd.numpy(), g["demo_theta_double"], rtol=1e-13)
    # per-sample thetas
    th = O.constrain_thetas(T(g["ps_theta_raw"]))
    yv, ld = O.bernstein_forward_and_log_det(th, T(g["ps_x"]))
    np.testing.assert_allclose(yv.numpy(), g["ps_y"], rtol=1e-10)
    np.testing.assert_allclose(ld.numpy(), g["ps_ld"], rtol=1e-10, atol=1e-12)


def test_bernstein_inverse_oracle_matches_reference_source():
    g = np.load(os.path.join(GOLDEN, "bernstein_golden.npz"))
    th = O.constrain_thetas(T(g["inv_theta_raw"]))
    xi = O.bernstein_inverse(th, T(g["inv_y"]))
    assert xi.shape == g["inv_x"].shape  # [E, N]: vmap(in_axes=-1) stacks on axis 0
    np.testing.assert_allclose(xi.numpy(), g["inv_x"], rtol=0, atol=1e-10)
    assert float(g["inv_logdet"]) == 0.0  # reference returns logdet = 0 (BernsteinBijector.py:119-123)
    th1 = O.constrain_thetas(T(g["tf_theta_raw"]))
    np.testing.assert_allclose(O.bernstein_inverse(th1, T(g["inv1_y"])).numpy(), g["inv1_x"], atol=1e-10)


def test_known_answers_K1_K2_K3():
    """zero-initialised final Linear ⇒ identity flow."""
    torch.manual_seed(0)
    for spec in (O.REAL_NVP_TOY, O.RQS_TOY):
        p = O.init_coupling_params(spec, seed=1, final_scale=0.0)
        x = torch.randn(64, 2, dtype=torch.float64) * 2
        lp = O.coupling_flow_log_prob(spec, p, x)
        np.testing.assert_allclose(lp.numpy(), O.log_prob_N01(x).numpy(), rtol=1e-12, atol=1e-12)
    spec = O.CouplingSpec(16, 2, [8], "rqs", 8, 0.0, 1.0, "uniform")
    p = O.init_coupling_params(spec, seed=1, final_scale=0.0)
    x = torch.rand(32, 16, dtype=torch.float64)
    np.testing.assert_allclose(O.coupling_flow_log_prob(spec, p, x).numpy(), 0.0, atol=1e-12)


def test_coupling_round_trip_and_logdet():
    for spec in (O.REAL_NVP_TOY, O.RQS_TOY, O.CouplingSpec(6, 3, [16], "affine"), O.CouplingSpec(6, 3, [16], "rqs", 5, -3.0, 3.0)):
        p = O.init_coupling_params(spec, seed=3, scale=0.5)
        z = torch.randn(50, spec.event_dim, dtype=torch.float64)
        x = O.coupling_flow_sample(spec, p, z)
        zb, ld = O.coupling_flow_to_base(spec, p, x)
        np.testing.assert_allclose(zb.numpy(), z.numpy(), atol=1e-9)
        # log-det vs autodiff Jacobian on a few samples
        for i in range(3):
            J = torch.autograd.functional.jacobian(lambda v: O.coupling_flow_to_base(spec, p, v[None])[0][0], x[i])
            np.testing.assert_allclose(torch.linalg.slogdet(J)[1].item(), ld[i].item(), atol=1e-8)


def test_made_autoregressive_property_and_round_trip():
    spec = O.MadeSpec(6, 2, [16, 16])
    p = O.init_made_params(spec, seed=2, scale=0.5)
    x = torch.randn(6, dtype=torch.float64)
    J = torch.autograd.functional.jacobian(lambda v: torch.cat(O.made_conditioner(spec, p[0], v[None])), x)
    J = J.reshape(2, 6, 6)  # [mu|alpha, out i, in j]
    for i in range(6):
        for j in range(i, 6):
            assert J[0, i, j] == 0 and J[1, i, j] == 0
    z = torch.randn(20, 6, dtype=torch.float64)
    xs = O.maf_sample(spec, p, z)
    lp = O.maf_log_prob(spec, p, xs)
    # change of variables: log p(x) = log N(z) - sum alpha; check through a jacobian
    Jx = torch.autograd.functional.jacobian(lambda v: O.maf_sample(spec, p, v[None])[0], z[0])
    np.testing.assert_allclose(lp[0].item(), (O.log_prob_N01(z[:1])[0] - torch.linalg.slogdet(Jx)[1]).item(), atol=1e-9)


# ---- device data step: Philox restatement (oracle/philox_oracle.py) ------------------------------
def test_philox4x32_10_known_answers():
    """Random123 kat_vectors for philox4x32-10 (Salmon et al. 2011): zero, all-ones and pi-digit inputs."""
    from oracle import philox_oracle as P

    kat = [
        ((0, 0, 0, 0), (0, 0), (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
        ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
        ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
         (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
    ]
    for ctr, key, want in kat:
        got = tuple(int(x) for x in P.philox4x32_10(ctr, key))
        assert got == want, (ctr, key, [hex(g) for g in got])


def test_density_streams_follow_the_reference_definitions():
    """Moments of the restated samplers against the closed forms implied by src/densities/*.py; the
    stream is a pure function of (seed, index): a batch taken at an offset equals that slice."""
    from oracle import philox_oracle as P

    n = 400_000
    x = P.density_sample(P.BANANA, 11, 0, n)       # x2~N(0,sqrt3), x1~N(x2^2/3-2, sqrt3/2)
    assert abs(x[:, 1].mean()) < 0.02 and abs(x[:, 1].var() - 3.0) < 0.05
    assert abs(x[:, 0].mean() + 1.0) < 0.02 and abs(x[:, 0].var() - 2.75) < 0.06
    x = P.density_sample(P.ENERGY1, 11, 0, n)      # x2~N(0,2); x1 = +-(2 - x2^2/32) + 0.3 n
    assert abs(x[:, 1].var() - 4.0) < 0.06 and abs(x[:, 0].mean()) < 0.02
    assert abs(np.abs(x[:, 0]).mean() - (2 - 4 / 32)) < 0.02
    x = P.density_sample(P.ENERGY2, 11, 0, n)      # x1~U(+-1.5pi), x2~N(sin(pi x1/2), .5)
    assert x[:, 0].min() >= -1.5 * np.pi and x[:, 0].max() < 1.5 * np.pi
    assert abs((x[:, 1] - np.sin(np.pi * x[:, 0] / 2)).var() - 0.25) < 0.01
    for kind, bump in ((P.ENERGY3, lambda z: -P.w2(z)), (P.ENERGY4, P.w3)):
        x = P.density_sample(kind, 11, 0, n)
        res = x[:, 1] - P.w1(x[:, 0])              # mixture of N(0,.5) and N(bump(x1), .5)
        want = 0.5 * bump(x[:, 0]).mean()
        assert abs(res.mean() - want) < 0.02
    x = P.density_sample(P.BLOBS, 11, 0, n)        # 8 blobs on a circle of radius 2, sigma .3
    assert abs(x.mean(0)).max() < 0.02 and abs(x.var(0) - (2.0 + 0.09)).max() < 0.05
    ang = np.round(np.arctan2(x[:, 1], x[:, 0]) / (np.pi / 4)).astype(int) % 8
    assert np.abs(np.bincount(ang, minlength=8) / n - 0.125).max() < 0.01
    a = P.density_sample(P.ENERGY4, 5, 0, 300)
    b = P.density_sample(P.ENERGY4, 5, 100, 50)
    assert np.array_equal(a[100:150], b)


def test_dequantize_and_base_noise_restatement():
    from oracle import philox_oracle as P

    b = np.random.default_rng(0).integers(0, 256, size=(7, 28, 28, 1), dtype=np.uint8)
    y = P.dequantize_u8(b, seed=9)
    assert y.dtype == np.float32 and y.shape == b.shape
    assert np.array_equal(np.floor(y.astype(np.float64) * 256).clip(0, 255).astype(np.uint8)[y < 1], b[y < 1])
    assert (y >= 0).all() and (y <= 1).all()
    assert np.array_equal(P.dequantize_u8(b.reshape(-1)[32:96], 9, offset=32), y.reshape(-1)[32:96])
    z = P.base_noise(0, 3, 0, 200_001)
    assert abs(z.mean()) < 0.01 and abs(z.var() - 1) < 0.02
    u = P.base_noise(1, 3, 0, 200_001)
    assert u.min() >= 0 and u.max() < 1 and abs(u.mean() - 0.5) < 0.005
    assert np.array_equal(P.base_noise(0, 3, 8, 40), z[8:48])


# ---------------------------------------------------------------------------------------------
# Rows #2-#4 and §8f-1: golden produced by executing the reference's own make_flow_model /
# make_conditioner / log_prob / sample (tests/golden/make_golden_flows.py) under numpy stand-ins
# for jax / haiku / distrax / TFP (tests/golden/ref_shim.py).  Pins the model WIRING (masks, flips,
# 3n+1, Flatten/Reshape, Inverse(Chain), bases, haiku names); the library primitives are restated.
# ---------------------------------------------------------------------------------------------
import sys

sys.path.insert(0, GOLDEN)
import mnist_params as MP  # noqa: E402


def unflatten_coupling(spec, flat):
    flat = T(flat)
    out, o = [], 0
    for _ in range(spec.num_layers):
        layer = []
        for m, n in spec.layer_shapes():
            W = flat[o : o + m * n].reshape(m, n); o += m * n
            b = flat[o : o + n]; o += n
            layer.append((W, b))
        out.append(layer)
    assert o == flat.numel()
    return out


def test_coupling_toy_oracle_matches_reference_make_flow_model():
    g = np.load(os.path.join(GOLDEN, "coupling_toy_golden.npz"))
    for key, spec in (("rnvp", O.REAL_NVP_TOY), ("maf", O.RQS_TOY)):
        params = unflatten_coupling(spec, g[f"{key}_flat"])
        lp = O.coupling_flow_log_prob(spec, params, T(g[f"{key}_x"]))
        np.testing.assert_allclose(lp.numpy(), g[f"{key}_log_prob"], rtol=1e-10, atol=1e-10)
        xs = O.coupling_flow_sample(spec, params, T(g[f"{key}_z"]))
        np.testing.assert_allclose(xs.numpy(), g[f"{key}_sample"], rtol=1e-9, atol=1e-9)


def test_haiku_names_are_the_reference_tree():
    """params_io.haiku_module_names must list exactly the modules the reference's make_flow_model created
    (the generator asserted its own tree against mnist_params.module_shapes and stored the names)."""
    from normalizing_flows_b200 import params_io

    g = np.load(os.path.join(GOLDEN, "coupling_toy_golden.npz"))
    ours = [n for layer in params_io.haiku_module_names(8, 2) for n in layer]
    assert ours == [str(s) for s in g["rnvp_names"]]


def mnist_golden_params():
    g = np.load(os.path.join(GOLDEN, "coupling_mnist_golden.npz"))
    tree = MP.make_tree(int(g["mnist_seed"]), 8, 784, [512, 512, 512], 19600, scale=float(g["mnist_scale"]))
    np.testing.assert_allclose(MP.tree_checksum(tree), g["mnist_checksum"], rtol=1e-12,
                               err_msg="numpy RNG stream differs from the one the golden file was made with")
    return g, MP.tree_to_flat(tree, 8, 784, [512, 512, 512], 19600)


def test_coupling_mnist_oracle_matches_reference_make_flow_model():
    g, flat = mnist_golden_params()
    spec = O.RQS_MNIST
    # float64 copies layer by layer (88 M parameters)
    params, o = [], 0
    for _ in range(spec.num_layers):
        layer = []
        for m, n in spec.layer_shapes():
            W = torch.from_numpy(flat[o : o + m * n].reshape(m, n)).double(); o += m * n
            b = torch.from_numpy(flat[o : o + n]).double(); o += n
            layer.append((W, b))
        params.append(layer)
    x = T(g["mnist_x"])  # [N, 28, 28, 1]
    lp = O.coupling_flow_log_prob(spec, params, x).numpy()
    ref = g["mnist_log_prob"]
    assert np.isneginf(ref[-1]) and np.isneginf(lp[-1])  # outside the Uniform(0,1) support
    np.testing.assert_allclose(lp[:-1], ref[:-1], rtol=1e-10, atol=1e-9)
    xs = O.coupling_flow_sample(spec, params, T(g["mnist_z"]))
    np.testing.assert_allclose(xs.numpy(), g["mnist_sample"].reshape(xs.shape), rtol=1e-9, atol=1e-9)


def test_bflow_oracle_matches_reference_script_prefix():
    g = np.load(os.path.join(GOLDEN, "bflow_golden.npz"))
    lp = O.bernstein_flow_log_prob(T(g["bflow_a"]), T(g["bflow_b"]), T(g["bflow_theta"]), T(g["bflow_alpha"]),
                                   T(g["bflow_beta"]), T(g["bflow_x"]))
    np.testing.assert_allclose(lp.numpy(), g["bflow_log_prob"], rtol=1e-9, atol=1e-9)
    X = T(g["bflow_X"])
    lp2 = O.bernstein_flow_log_prob(T(g["bflow_a"]), T(g["bflow_b"]), T(g["bflow_theta"]), T(g["bflow_alpha"]),
                                    T(g["bflow_beta"]), X).sum(-1)  # Block(ndims=1) + MVNDiag base (bflow.py:133-135)
    np.testing.assert_allclose(lp2.numpy(), g["bflow_mvn_log_prob"], rtol=1e-9, atol=1e-9)
    # SoftClip on its own (T2)
    y, ld = O.softclip_forward_and_log_det(T(g["bflow_T2_x"]))
    np.testing.assert_allclose(y.numpy(), g["bflow_T2_y"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(ld.numpy(), g["bflow_T2_ld"], rtol=1e-10, atol=1e-12)
