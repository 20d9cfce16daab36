"""GPU: reverse mode of the BernsteinBijector (csrc/bernstein_grad.cu, SURVEY §8a row #7 for row #5).

Oracle: float64 autograd through oracle/flows_oracle.py's restatement of src/BernsteinBijector.py (the Beta-pdf
form the reference evaluates), i.e. what jax.grad returns in src/maf_mwe.py:143.  Tolerance: the kernels are
FP32; cotangents are compared at 2e-5·max(1,|ref|) per element for xbar / per-row thetabar and at
2e-5·max|ref| for the batch-summed thetabar (a sum of n FP32 terms, reduced in FP64).  For orders above 32
xbar is scaled by the size of the terms that are summed, not of their sum: dJ/du = n(n-1)·Σ Δ²θ_j b_{j,n-2}
has coefficients ~1e3·θ of alternating sign for the random thetas used here, so the FP32 sum (ours, and the
reference's own FP32 autodiff) cancels three to four digits; the bound is 2e-5·max(1, |ref|, |g|·n(n-1)·
Σ|Δ²θ_j| b_{j,n-2})."""
import numpy as np
import pytest
import torch

from oracle import flows_oracle as O

pytestmark = pytest.mark.gpu

TOL = 2e-5


LO, HI = np.float32(1e-7), np.float32(1.0) - np.float32(2.0 ** -23)  # the float32 clip bounds


def _rand_x(shape, g):
    """float32 uniforms with the (measure-zero) values exactly ON a clip bound moved inside: there jax splits
    the cotangent in half (tested on its own below) while torch.clamp, the oracle's clip, passes all of it."""
    x = torch.rand(shape, generator=g, dtype=torch.float64).float()
    return torch.where((x == float(LO)) | (x == float(HI)), torch.full_like(x, 0.5), x)


def _theta(order, rows=None, seed=0):
    g = torch.Generator().manual_seed(seed)
    raw = torch.randn((order,) if rows is None else (rows, order), generator=g, dtype=torch.float64)
    return O.constrain_thetas(raw)


def _ref(theta64, x64, ybar64, ldbar64):
    th = theta64.clone().requires_grad_(True)
    x = x64.clone().requires_grad_(True)
    y, ld = O.bernstein_forward_and_log_det(th, x)
    tot = 0.0
    if ybar64 is not None:
        tot = tot + (y * ybar64).sum()
    if ldbar64 is not None:
        tot = tot + (ld * ldbar64).sum()
    tot.backward()
    return x.grad, th.grad


def _xbar_scale(theta64, x64, ldbar64, ref):
    """max(1, |ref|, size of the cancelling dJ/du terms) for orders > 32 (see the module docstring)."""
    order = theta64.shape[-1]
    base = torch.clamp(ref.abs(), min=1.0)
    if order <= 32 or ldbar64 is None:
        return base
    n = order - 1
    d2 = (theta64[..., 2:] - 2 * theta64[..., 1:-1] + theta64[..., :-2]).abs()
    xc = O._bern_clip(x64)
    J = O.bernstein_poly_jac(theta64, xc)
    cond = (ldbar64 / J).abs() * n * (n - 1) * O.bernstein_poly(d2, xc)  # Σ|Δ²θ_j| b_{j,n-2}: a degree n-2 form
    return torch.maximum(base, cond)


def _check(got, ref, scale=None, what="", tol=TOL):
    got = got.double().cpu()
    denom = torch.clamp(ref.abs(), min=1.0) if scale is None else scale
    err = ((got - ref).abs() / denom).max().item()
    assert err <= tol, f"{what}: max scaled error {err:.3e}"


@pytest.mark.parametrize("order", [2, 3, 4, 5, 8, 9, 16, 17, 33, 64])
@pytest.mark.parametrize("n", [1, 1000, 100003])
def test_shared_theta_vjp_matches_float64_autograd(order, n):
    from normalizing_flows_b200 import ops

    g = torch.Generator().manual_seed(order * 7 + n)
    th = _theta(order, seed=order)
    x = _rand_x(n, g).double()
    yb = torch.randn(n, generator=g, dtype=torch.float64).float().double()
    lb = torch.randn(n, generator=g, dtype=torch.float64).float().double()
    xr, tr = _ref(th.float().double(), x, yb, lb)
    xbar, tbar = ops.bernstein_forward_and_log_det_vjp(th.float().cuda(), x.float().cuda(), yb.float().cuda(), lb.float().cuda())
    _check(xbar, xr, scale=_xbar_scale(th.float().double(), x, lb, xr), what="xbar")
    _check(tbar, tr, scale=max(1.0, tr.abs().max().item()), what="thetabar")
    # the same call again gives the same bits (fixed-order reduction)
    _, again = ops.bernstein_forward_and_log_det_vjp(th.float().cuda(), x.float().cuda(), yb.float().cuda(), lb.float().cuda())
    assert torch.equal(again, tbar)


@pytest.mark.parametrize("which", ["ybar", "ldbar"])
def test_single_cotangent_and_optional_outputs(which):
    from normalizing_flows_b200 import _lib, ops

    n, order = 5000, 9
    g = torch.Generator().manual_seed(3)
    th = _theta(order, seed=5).float()
    x = _rand_x(n, g)
    ct = torch.randn(n, generator=g)
    kw = {which: ct.cuda()}
    xr, tr = _ref(th.double(), x.double(), ct.double() if which == "ybar" else None, ct.double() if which == "ldbar" else None)
    xbar, tbar = ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda(), **kw)
    _check(xbar, xr, what="xbar")
    _check(tbar, tr, scale=max(1.0, tr.abs().max().item()), what="thetabar")
    only_x, none_t = ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda(), want_thetabar=False, **kw)
    assert none_t is None and torch.equal(only_x, xbar)
    none_x, only_t = ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda(), want_xbar=False, **kw)
    assert none_x is None and torch.equal(only_t, tbar)
    with pytest.raises(_lib.FlowCudaError):
        ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda())  # no cotangent at all


def test_clip_kills_the_input_gradient_outside_the_unit_interval():
    """jnp.clip(x, 1e-7, 1-1e-7) (BernsteinBijector.py:84-85): no gradient flows to x where the clip is active,
    theta still receives its gradient from the clipped point."""
    from normalizing_flows_b200 import ops

    th = _theta(9, seed=1).float()
    x = torch.tensor([-0.5, 0.0, 5e-8, 0.25, 0.999, 1.0, 1.5, 2e-7])
    yb, lb = torch.ones(8), torch.ones(8)
    xr, tr = _ref(th.double(), x.double(), yb.double(), lb.double())
    xbar, tbar = ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda(), yb.cuda(), lb.cuda())
    assert (xbar.cpu()[[0, 1, 2, 5, 6]] == 0).all() and (xbar.cpu()[[3, 4, 7]] != 0).all()
    _check(xbar, xr, what="xbar")
    _check(tbar, tr, scale=max(1.0, tr.abs().max().item()), what="thetabar")
    # exactly on a bound: clip = minimum(maximum(x, lo), hi), and lax.max / lax.min give each tied argument half
    # the cotangent, so jax.grad sees HALF the interior slope there
    xt = torch.tensor([float(LO), float(HI)])
    tie, _ = ops.bernstein_forward_and_log_det_vjp(th.cuda(), xt.cuda(), torch.ones(2).cuda(), torch.ones(2).cuda())
    inner = torch.tensor([float(np.nextafter(LO, np.float32(1))), float(np.nextafter(HI, np.float32(0)))])
    full, _ = ops.bernstein_forward_and_log_det_vjp(th.cuda(), inner.cuda(), torch.ones(2).cuda(), torch.ones(2).cuda())
    assert torch.allclose(tie, 0.5 * full, rtol=1e-4)


@pytest.mark.parametrize("order", [2, 4, 8, 9, 20, 64])
@pytest.mark.parametrize("shape", [(1,), (33,), (7, 28, 28, 1)])
def test_rows_theta_vjp_matches_float64_autograd(order, shape):
    """theta[n, M], one row per element (maf_mwe.py:103-107)."""
    from normalizing_flows_b200 import ops

    n = int(np.prod(shape))
    g = torch.Generator().manual_seed(order + n)
    th = _theta(order, rows=n, seed=order).float().reshape(*shape, order)
    x = _rand_x(shape, g)
    yb, lb = torch.randn(shape, generator=g), torch.randn(shape, generator=g)
    xr, tr = _ref(th.double(), x.double(), yb.double(), lb.double())
    xbar, tbar = ops.bernstein_forward_and_log_det_vjp(th.cuda(), x.cuda(), yb.cuda(), lb.cuda())
    assert tbar.shape == th.shape and xbar.shape == x.shape
    _check(xbar, xr, scale=_xbar_scale(th.double(), x.double(), lb.double(), xr), what="xbar")
    _check(tbar, tr, what="thetabar rows")


def test_empty_batch_and_bad_arguments():
    from normalizing_flows_b200 import _lib, ops

    th = _theta(9).float().cuda()
    e = torch.zeros(0, device="cuda")
    xbar, tbar = ops.bernstein_forward_and_log_det_vjp(th, e, e, e)
    assert xbar.shape == (0,) and (tbar == 0).all()
    with pytest.raises(_lib.FlowCudaError):
        ops.bernstein_forward_and_log_det_vjp(torch.zeros(65, device="cuda"), torch.rand(4, device="cuda"), torch.ones(4, device="cuda"))
    with pytest.raises(_lib.FlowCudaError):
        ops.bernstein_forward_and_log_det_vjp(th.cpu(), torch.rand(4), torch.ones(4))  # host tensors: no CPU path


@pytest.mark.parametrize("shape", [(9,), (5, 9), (3, 4, 8)])
def test_constrain_thetas_vjp(shape):
    from normalizing_flows_b200 import ops

    g = torch.Generator().manual_seed(11)
    raw = (torch.randn(shape, generator=g) * 3).float()
    tb = torch.randn(shape, generator=g).float()
    r = raw.double().requires_grad_(True)
    (O.constrain_thetas(r) * tb.double()).sum().backward()
    got = ops.bernstein_constrain_thetas_vjp(raw.cuda(), tb.cuda())
    _check(got, r.grad, what="rawbar")


def test_autograd_through_the_bijector_matches_the_reference_training_step():
    """The loss of src/maf_mwe.py:96-113 — thetas from a conditioner, one row per pixel, mean log-density — with
    torch autograd driving the two kernels; gradients w.r.t. the UNCONSTRAINED thetas and x against float64."""
    from normalizing_flows_b200 import BernsteinBijector as B

    g = torch.Generator().manual_seed(2)
    shape, order = (6, 28, 28, 1), 8
    raw = torch.randn(*shape, order, generator=g)
    x = _rand_x(shape, g)

    r64 = raw.double().requires_grad_(True)
    x64 = x.double().requires_grad_(True)
    y, ld = O.bernstein_forward_and_log_det(O.constrain_thetas(r64), x64)
    (ld.mean() + 0.1 * (y ** 2).mean()).backward()

    rg = raw.cuda().requires_grad_(True)
    xg = x.cuda().requires_grad_(True)
    bij = B.BernsteinBijector(rg)
    yg, ldg = bij.forward_and_log_det(xg)
    (ldg.mean() + 0.1 * (yg ** 2).mean()).backward()
    scale = r64.grad.abs().max().item()
    assert ((rg.grad.double().cpu() - r64.grad).abs().max().item()) <= TOL * max(scale, 1e-3) * 10
    assert ((xg.grad.double().cpu() - x64.grad).abs().max().item()) <= TOL * max(x64.grad.abs().max().item(), 1e-3) * 10

    # shared theta: forward_log_det_jacobian alone, gradient of the NLL w.r.t. the 9 raw parameters of bflow.py:24
    raw1 = torch.tensor([-2.5, 2, 3, -7, -7, -7, -7, -7, 7.0])
    xs = _rand_x(4096, g)
    r1 = raw1.double().requires_grad_(True)
    (-O.bernstein_forward_log_det_jacobian(O.constrain_thetas(r1), xs.double()).mean()).backward()
    rg1 = raw1.cuda().requires_grad_(True)
    (-B.BernsteinBijector(rg1).forward_log_det_jacobian(xs.cuda()).mean()).backward()
    assert (rg1.grad.double().cpu() - r1.grad).abs().max().item() <= TOL * max(1.0, r1.grad.abs().max().item())
    # no autograd bookkeeping when nothing asks for gradients
    out = B.BernsteinBijector(raw1).forward(xs.cuda())
    assert not out.requires_grad
