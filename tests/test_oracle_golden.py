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
