"""The reference-facing host API (distrax / real_nvp / BernsteinBijector mirrors) on the GPU,
checked against the CPU oracle.  These read like the reference's own call sites."""
import numpy as np
import pytest
import torch

from oracle import flows_oracle as O

pytestmark = pytest.mark.gpu


def scaled_err(got, ref):
    got = got.detach().double().cpu().numpy()
    ref = ref.detach().double().cpu().numpy() if isinstance(ref, torch.Tensor) else np.asarray(ref)
    return float((np.abs(got - ref) / np.maximum(1, np.abs(ref))).max())


def test_make_flow_model_real_nvp_toy():
    """real_nvp_toy_densities.py:74-136: log_prob.apply / sample.apply."""
    from normalizing_flows_b200 import flows

    params = O.init_coupling_params(O.REAL_NVP_TOY, seed=3, scale=0.3)
    model = flows.make_flow_model(event_shape=(2,), num_layers=8, hidden_sizes=[64, 64],
                                  params=O.flatten_layers(params).float())
    x = torch.randn(1000, 2, generator=torch.Generator().manual_seed(0)) * 2
    lp = model.log_prob(x.cuda())
    assert lp.shape == (1000,)
    assert scaled_err(lp, O.coupling_flow_log_prob(O.REAL_NVP_TOY, params, x.double())) < 1e-5
    z = torch.randn(500, 2, generator=torch.Generator().manual_seed(1))
    xs = model.sample_from_noise(z.cuda())
    assert scaled_err(xs, O.coupling_flow_sample(O.REAL_NVP_TOY, params, z.double())) < 2e-5
    s = model.sample(seed=42, sample_shape=(256,))
    assert s.shape == (256, 2) and torch.isfinite(s).all()
    # zero-initialised final Linear (the reference's init) => log_prob == Normal log-density (K1)
    m0 = flows.make_flow_model((2,), 8, [64, 64])
    assert scaled_err(m0.log_prob(x.cuda()), O.log_prob_N01(x.double())) < 5e-6


def test_chain_and_masked_coupling_protocol():
    """Chain.forward applies the last layer first; a single MaskedCoupling matches the oracle."""
    from normalizing_flows_b200 import distrax_api as distrax
    from normalizing_flows_b200 import flows

    sp = O.RQS_TOY
    params = O.init_coupling_params(sp, seed=4, scale=0.3)
    model = flows.make_flow_model((2,), 8, [64, 64], transform="rqs", params=O.flatten_layers(params).float())
    chain = model.bijector.bijector
    assert isinstance(chain, distrax.Chain) and chain._fused is not None
    x = torch.randn(300, 2, generator=torch.Generator().manual_seed(5)) * 3
    z, ld = model.bijector.inverse_and_log_det(x.cuda())  # Inverse(Chain).inverse = Chain.forward
    zr, ldr = O.coupling_flow_to_base(sp, params, x.double())
    assert scaled_err(z, zr) < 3e-5 and scaled_err(ld, ldr) < 3e-5
    for l in (0, 1):
        y, ldl = chain.bijectors[l].forward_and_log_det(x.cuda())
        yr, lr = O.masked_coupling(sp, params[l], l, x.double(), inverse=False)
        assert scaled_err(y, yr) < 2e-5 and scaled_err(ldl, lr) < 2e-5
        xi, ldi = chain.bijectors[l].inverse_and_log_det(y)
        assert scaled_err(xi, x) < 5e-5
    # un-fused composition (python loop over single-layer kernels) == fused stack
    loose = distrax.Chain(chain.bijectors)
    loose._fused = None
    z2, ld2 = loose.forward_and_log_det(x.cuda())
    assert scaled_err(z2, z) < 1e-5 and scaled_err(ld2, ld) < 1e-5
    # a Python conditioner callable + the reference's bijector_fn: composed step by step (conditioner call, bijector_fn(params),
    # the inner bijector's elementwise kernel, masking) — distrax.MaskedCoupling semantics, SURVEY A.3
    cond = lambda mx: torch.stack([0.3 * mx.sum(-1), 0.1 * mx.sum(-1)], -1)
    mc = distrax.MaskedCoupling(mask=torch.tensor([False, True]), conditioner=cond, bijector=distrax.scalar_affine_fn)
    xg = x.cuda()
    yg, ldg = mc.forward_and_log_det(xg)
    shift, ls = 0.3 * x[:, 1].double(), 0.1 * x[:, 1].double()
    yref = torch.stack([x[:, 0].double() * torch.exp(ls) + shift, x[:, 1].double()], -1)
    assert scaled_err(yg, yref) < 1e-5 and scaled_err(ldg, ls) < 1e-6
    xb, ldb = mc.inverse_and_log_det(yg)
    assert scaled_err(xb, x) < 1e-5 and scaled_err(ldb, -ls) < 1e-6
    with pytest.raises(ValueError):
        distrax.MaskedCoupling(mask=torch.tensor([0, 1]), conditioner=cond, bijector=distrax.scalar_affine_fn)  # mask must be bool


def test_real_nvp_functional_api():
    """real_nvp.py: forward_nvp / inverse_nvp / log_prob_nvp_chains / sample_nvp_chain / sgd_update."""
    from normalizing_flows_b200 import real_nvp as R

    p64, flips = O.init_nvp_chain_like(8, hidden=4, seed=9, scale=0.3)
    params = [[[W.float(), b.float()] for W, b in layer] for layer in p64]
    cfg = [(R.shift_and_log_scale_fn, f) for f in flips]
    y = torch.randn(128, 2, generator=torch.Generator().manual_seed(2)) * 1.5
    lp = R.log_prob_nvp_chains(params, cfg, R.log_prob_N01, y)
    assert scaled_err(lp, O.log_prob_nvp_chains(p64, flips, y.double())) < 1e-5
    x, ls = R.inverse_nvp(params[1], R.shift_and_log_scale_fn, y, flip=True)
    xr, lsr = O.inverse_nvp(p64[1], y.double(), True)
    assert ls.shape == (128, 1) and scaled_err(x, xr) < 1e-5 and scaled_err(ls, lsr) < 1e-5
    assert scaled_err(R.forward_nvp(params[1], R.shift_and_log_scale_fn, x, flip=True), y) < 1e-5
    assert scaled_err(R.log_prob_N01(y), O.log_prob_N01(y.double())) < 1e-6
    smp = R.sample_nvp_chain(params, cfg, lambda N, key: y, 128, None)
    assert scaled_err(smp, O.sample_nvp_chain(p64, flips, y.double())) < 2e-5
    # training: 20 SGD steps reproduce the oracle's autograd trajectory (real_nvp.py:295-298)
    tr = R.SGDTrainer(params, cfg, learning_rate=1e-2)
    flat = O.flatten_layers(p64).clone()
    for step in range(20):
        batch = torch.randn(128, 2, generator=torch.Generator().manual_seed(100 + step))
        nll, _ = tr.sgd_update(batch)

        def lp_fn(leaves):
            return O.log_prob_nvp_chains(_unflat(leaves[0]), flips, batch.double())

        loss_ref, (g,) = O.nll_and_grads(lp_fn, [flat])
        flat = O.sgd_step(flat, g, 1e-2)
        assert abs(nll.item() - loss_ref.item()) < 1e-4 * max(1, abs(loss_ref.item()))
    assert scaled_err(tr.flat, flat) < 1e-4


def _unflat(flat, L=8, H=4):
    out, o = [], 0
    for _ in range(L):
        W1 = flat[o:o + H].reshape(1, H); o += H
        b1 = flat[o:o + H]; o += H
        W2 = flat[o:o + 2 * H].reshape(H, 2); o += 2 * H
        b2 = flat[o:o + 2]; o += 2
        out.append([[W1, b1], [W2, b2]])
    return out


def test_bernstein_bijector_api():
    """BernsteinBijector.py:74-123 and the bflow.py demo parameters."""
    from normalizing_flows_b200 import BernsteinBijector as B

    theta = torch.tensor([-2.5, 2, 3, -7, -7, -7, -7, -7, 7])
    bij = B.BernsteinBijector(theta)
    thc = O.constrain_thetas(theta.double())
    assert scaled_err(bij.thetas, thc) < 1e-6
    x = torch.linspace(0, 1, 2000)
    y, ld = bij.forward_and_log_det(x)
    yr, lr = O.bernstein_forward_and_log_det(bij.thetas.double().cpu(), x.double())
    assert scaled_err(y, yr) < 1e-5 and scaled_err(ld, lr) < 1e-5
    assert scaled_err(bij.forward(x), yr) < 1e-5 and scaled_err(bij.forward_log_det_jacobian(x), lr) < 1e-5
    xi, zero = bij.inverse_and_log_det(y[:, None])
    assert zero == 0 and xi.shape == (1, 2000)
    # the reference's own acceptance figure: round-trip error of inverse(forward(x)) (bern_flow.py:79-81)
    assert float(((xi[0].cpu() - x) ** 2).sum()) < 1e-3
    grid = torch.linspace(-4, 4, 4001)
    lp = B.bernstein_flow_log_prob(2.0, -0.1, theta, 0.2, -2.5, grid)
    ref = O.bernstein_flow_log_prob(2.0, -0.1, theta.double(), 0.2, -2.5, grid.double())
    assert scaled_err(lp, ref) < 3e-5
    # it is a density: integrates to ~1 over the grid
    assert abs(torch.trapezoid(lp.exp().cpu().double(), grid.double()).item() - 1.0) < 2e-2
