"""Mirror of the reference's src/BernsteinBijector.py on libflowcuda kernels: same names
(`constrain_thetas`, `get_bernstein_poly`, `get_bernstein_poly_jac`, `BernsteinBijector`),
same argument meaning, same quirks (the inverse returns [E, N] and log-det 0)."""
from __future__ import annotations

import torch

from . import ops
from .distrax_api import Bijector


class _ConstrainThetas(torch.autograd.Function):
    """constrain_thetas with its reverse mode (bg_constrain_vjp_kernel)."""

    @staticmethod
    def forward(ctx, raw):
        ctx.save_for_backward(raw)
        return ops.bernstein_constrain_thetas(raw)

    @staticmethod
    def backward(ctx, thetabar):
        (raw,) = ctx.saved_tensors
        return ops.bernstein_constrain_thetas_vjp(raw, thetabar.contiguous())


class _ForwardAndLogDet(torch.autograd.Function):
    """forward_and_log_det with its reverse mode (csrc/bernstein_grad.cu) — what jax.grad differentiates in
    src/maf_mwe.py:96-113,143.  Differentiable once (the backward is a kernel, not a graph)."""

    @staticmethod
    def forward(ctx, theta, x):
        ctx.save_for_backward(theta, x)
        y, ld = ops.bernstein_forward_and_log_det(theta, x)
        return y, ld

    @staticmethod
    def backward(ctx, ybar, ldbar):
        theta, x = ctx.saved_tensors
        need_t, need_x = ctx.needs_input_grad
        xbar, tbar = ops.bernstein_forward_and_log_det_vjp(theta, x, ybar, ldbar, want_xbar=need_x, want_thetabar=need_t)
        return tbar, xbar


def _tracked(*ts):
    return torch.is_grad_enabled() and any(isinstance(t, torch.Tensor) and t.requires_grad for t in ts)


def _fwd(theta, x):
    """(y, log_det); recorded for autograd when theta or x asks for gradients."""
    x = _cuda(x)
    if _tracked(theta, x):
        if theta.dim() > 1 and tuple(theta.shape[:-1]) != tuple(x.shape):
            theta = theta.expand(*x.shape, theta.shape[-1])
        return _ForwardAndLogDet.apply(theta.contiguous(), x.contiguous())
    return ops.bernstein_forward_and_log_det(theta, x)


def constrain_thetas(theta_unconstrained, range_min=None, range_max=None, fn=None):
    """BernsteinBijector.py:60-71 (range_min/range_max/fn are unused there too)."""
    raw = torch.as_tensor(theta_unconstrained, dtype=torch.float32).to("cuda")
    if _tracked(raw):
        return _ConstrainThetas.apply(raw.contiguous())
    return ops.bernstein_constrain_thetas(raw)


def get_bernstein_poly(theta):
    """BernsteinBijector.py:23-36: returns x -> mean_k Beta(k, M-k+1).pdf(x) * theta_k.

    Deviation: the kernel clips x to [1e-7, 1 - 1e-7] (the clip of `forward`, BernsteinBijector.py:81-84) here too; the
    reference's bare polynomial does not, and returns 0 / NaN from the Beta pdfs outside (0, 1)."""
    theta = torch.as_tensor(theta, dtype=torch.float32, device="cuda")

    def bernstein_poly(x):
        return ops.bernstein_forward_and_log_det(theta, _cuda(x), want_ld=False)[0]

    return bernstein_poly


def get_bernstein_poly_jac(theta):
    """BernsteinBijector.py:45-57: returns y -> sum_k Beta(k, M-k).pdf(y) * (theta_{k+1}-theta_k)."""
    theta = torch.as_tensor(theta, dtype=torch.float32, device="cuda")

    def bernstein_poly_jac(y):
        return torch.exp(ops.bernstein_forward_and_log_det(theta, _cuda(y), want_y=False)[1])

    return bernstein_poly_jac


def _cuda(x):
    return torch.as_tensor(x, dtype=torch.float32).to("cuda")


class BernsteinBijector(Bijector):
    """BernsteinBijector.py:74-123.  `thetas` are UNCONSTRAINED; __init__ constrains them."""

    def __init__(self, thetas):
        super().__init__(event_ndims_in=0)
        self.thetas = constrain_thetas(thetas)

    def forward(self, x):
        if _tracked(self.thetas, x):
            return _fwd(self.thetas, x)[0]
        return ops.bernstein_forward_and_log_det(self.thetas, _cuda(x), want_ld=False)[0]

    def forward_log_det_jacobian(self, x):
        if _tracked(self.thetas, x):
            return _fwd(self.thetas, x)[1]
        return ops.bernstein_forward_and_log_det(self.thetas, _cuda(x), want_y=False)[1]

    def forward_and_log_det(self, x):
        return _fwd(self.thetas, x)

    def inverse(self, y):
        return ops.bernstein_inverse(self.thetas, _cuda(y))

    def inverse_and_log_det(self, y):
        return self.inverse(y), 0  # reference behaviour (BernsteinBijector.py:119-123)


def bernstein_flow_log_prob(a, b, theta, alpha, beta, x, double_constrain: bool = True):
    """Transformed(Normal(0,1), Inverse(Chain([T4,T3,T2,T1]))).log_prob(x) of bflow.py:28-34,82-88
    in one fused kernel.  `theta` is unconstrained; bflow.py:30 constrains it before
    BernsteinBijector.__init__ constrains it again (`double_constrain`)."""
    th = constrain_thetas(theta)
    if double_constrain:
        th = ops.bernstein_constrain_thetas(th)
    scal = torch.tensor([float(a), float(b), float(alpha), float(beta)], dtype=torch.float32, device="cuda")
    x = _cuda(x)
    return ops.bernstein_flow_log_prob(th, scal, x.reshape(-1)).reshape(x.shape)
