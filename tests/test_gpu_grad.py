"""GPU: hand-written reverse mode of the toy MaskedCoupling flows (flowcuda_coupling_nll_grad) against
torch autograd through the float64 oracle — jax.value_and_grad(loss_fn) of
src/real_nvp_toy_densities.py:139-151,170 / src/maf_toy_densities.py:140-152,171."""
import numpy as np
import pytest
import torch

from oracle import flows_oracle as O
from test_gpu_parity import COUPLING_CASES, G, _spec_pair

pytestmark = pytest.mark.gpu


def oracle_loss_and_grad(ospec, params, x64):
    leaves = [t for layer in params for (W, b) in layer for t in (W, b)]

    def lp_fn(lv):
        it = iter(lv)
        p = [[(next(it), next(it)) for _ in layer] for layer in params]
        return O.coupling_flow_log_prob(ospec, p, x64)

    loss, grads = O.nll_and_grads(lp_fn, leaves)
    return loss, torch.cat([g.reshape(-1) for g in grads])


@pytest.mark.parametrize("case", ["realnvp_toy", "rqs_toy", "rqs_toy_k8", "affine_toy"])
@pytest.mark.parametrize("n", [1, 128, 1000, 40_000])
def test_coupling_nll_grad_matches_autograd(case, n):
    from normalizing_flows_b200 import ops

    ospec, scale, xs = COUPLING_CASES[case]
    params = O.init_coupling_params(ospec, seed=21, scale=scale)
    flat = O.flatten_layers(params)
    x = (torch.randn(n, 2, generator=torch.Generator().manual_seed(n + 3), dtype=torch.float64) * xs).float()
    loss_ref, grad_ref = oracle_loss_and_grad(ospec, params, x.double())
    spec = _spec_pair(ops, ospec)
    p = G(flat)
    loss, grad = ops.coupling_nll_grad(spec, p, x.cuda())
    assert abs(loss.item() - loss_ref.item()) <= 2e-5 * max(1.0, abs(loss_ref.item()))
    g, gr = grad.double().cpu().numpy(), grad_ref.numpy()
    err = np.abs(g - gr).max() / max(1.0, np.abs(gr).max())
    assert err <= 5e-5, err
    # deterministic: bit-identical on a second run
    loss2, grad2 = ops.coupling_nll_grad(spec, p, x.cuda())
    assert torch.equal(grad, grad2) and torch.equal(loss, loss2)
    # loss agrees with the forward kernel's log_prob
    lp, _, _ = ops.coupling_log_prob(spec, p, x.cuda())
    assert abs(loss.item() + lp.double().mean().item()) <= 2e-5 * max(1.0, abs(loss_ref.item()))


def test_adam_training_step_decreases_loss():
    """update() of src/real_nvp_toy_densities.py:164-173: grad -> optax.adam -> apply_updates, from the
    reference's zero-initialised final Linear."""
    from normalizing_flows_b200 import ops

    ospec = O.RQS_TOY
    params = O.init_coupling_params(ospec, seed=3, scale=0.3, final_scale=0.0)
    p = G(O.flatten_layers(params))
    spec = _spec_pair(ops, ospec)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    gen = torch.Generator().manual_seed(0)
    k = torch.randint(0, 8, (4096,), generator=gen).double() * (2 * np.pi / 8)
    x = (torch.stack([2 * torch.cos(k), 2 * torch.sin(k)], -1) + 0.3 * torch.randn(4096, 2, generator=gen, dtype=torch.float64)).float().cuda()
    losses = []
    for step in range(1, 31):
        loss, grad = ops.coupling_nll_grad(spec, p, x)
        ops.adam_step(p, grad, m, v, step, 1e-3)
        losses.append(loss.item())
    assert np.isfinite(losses).all() and losses[-1] < losses[0] - 0.05, losses


# ------------------------------------------------------------------ generic event size (row #4: maf_mnist.py training step)
DENSE_GRAD = {
    "rqs_d40_uniform": (O.CouplingSpec(40, 2, [32, 32, 32], "rqs", 8, 0.0, 1.0, "uniform"), 0.3, None, 77),
    "rqs_d7_k8": (O.CouplingSpec(7, 3, [24, 24], "rqs", 8, -3.0, 3.0), 0.3, 2.0, 130),
    "rqs_d9_k12": (O.CouplingSpec(9, 2, [20], "rqs", 12, -4.0, 4.0), 0.3, 2.0, 64),
    "affine_d6": (O.CouplingSpec(6, 3, [16], "affine"), 0.3, 1.5, 50),
    "affine_d300_splitk": (O.CouplingSpec(300, 2, [320, 64], "affine"), 0.2, 1.0, 1100),
    "rqs_d50_splitk": (O.CouplingSpec(50, 2, [300], "rqs", 8, -4.0, 4.0), 0.3, 2.0, 1500),
    "rqs_mnist_shape": (O.CouplingSpec(784, 2, [512, 512, 512], "rqs", 8, 0.0, 1.0, "uniform"), 0.1, None, 40),
}


@pytest.mark.parametrize("engine", ["fp32", "bf16x3"])
@pytest.mark.parametrize("case", sorted(DENSE_GRAD))
def test_dense_coupling_nll_grad_matches_autograd(case, engine):
    """Stored-activation reverse mode of MaskedCoupling flows of any event size (csrc/coupling_dense_grad.cu)
    vs autograd through the float64 oracle: loss 2e-5 relative, gradient 1e-4 of its largest entry,
    bit-identical on a second run.  Both GEMM engines: FP32 CUDA cores (mg_sgemm.cuh) and tcgen05 bf16x3
    (mgtc.cuh) — the north_star budget for the tensor-core path is 1e-3; the 3-pass split holds the FP32 one."""
    from normalizing_flows_b200 import _lib, ops

    ospec, scale, xs, n = DENSE_GRAD[case]
    params = O.init_coupling_params(ospec, seed=23, scale=scale)
    flat = O.flatten_layers(params)
    gen = torch.Generator().manual_seed(n)
    if xs is None:
        x = torch.rand(n, ospec.event_dim, generator=gen, dtype=torch.float64).float()
    else:
        x = (torch.randn(n, ospec.event_dim, generator=gen, dtype=torch.float64) * xs).float()
    loss_ref, grad_ref = oracle_loss_and_grad(ospec, params, x.double())
    spec = _spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3 if engine == "bf16x3" else 0)
    p = G(flat)
    loss, grad = ops.coupling_nll_grad(spec, p, x.cuda())
    loss2, grad2 = ops.coupling_nll_grad(spec, p, x.cuda())
    assert torch.equal(grad, grad2) and torch.equal(loss, loss2)
    assert abs(loss.item() - loss_ref.item()) <= 2e-5 * max(1.0, abs(loss_ref.item())), (loss.item(), loss_ref.item())
    g, gr = grad.double().cpu().numpy(), grad_ref.numpy()
    scale = max(1e-6, np.abs(gr).max())
    if engine == "fp32":
        assert np.abs(g - gr).max() / scale <= 1e-4
    else:
        # relu' is discontinuous: a hidden pre-activation within ~1e-5 of zero (rqs_d50_splitk has one at
        # 1.1e-6, unit 52 of the first Linear) can land on the other side in bf16x3 and moves ONE sample's
        # contribution to that unit's row.  So: 99.9 % of the entries within the FP32 bound, every entry
        # within the north_star budget for the tensor-core path (1e-3 absolute, here 2e-3 of the largest).
        assert np.quantile(np.abs(g - gr), 0.999) / scale <= 1e-4
        assert np.abs(g - gr).max() / scale <= 2e-3


@pytest.mark.parametrize("n", [130, 3000])
def test_toy_rqs_nll_grad_on_the_tensor_core_engine(n):
    """precision = BF16X3 routes even the event_dim = 2 flows through the stored-activation engine
    (K = 2 operand images, 12-bin compact adjoint): same loss and gradient as the float64 oracle."""
    from normalizing_flows_b200 import _lib, ops

    ospec, scale, xs = COUPLING_CASES["rqs_toy"]
    params = O.init_coupling_params(ospec, seed=21, scale=scale)
    x = (torch.randn(n, 2, generator=torch.Generator().manual_seed(n + 3), dtype=torch.float64) * xs).float()
    loss_ref, grad_ref = oracle_loss_and_grad(ospec, params, x.double())
    p = G(O.flatten_layers(params))
    loss, grad = ops.coupling_nll_grad(_spec_pair(ops, ospec, precision=_lib.PRECISION_BF16X3), p, x.cuda())
    assert abs(loss.item() - loss_ref.item()) <= 2e-5 * max(1.0, abs(loss_ref.item()))
    g, gr = grad.double().cpu().numpy(), grad_ref.numpy()
    scale_g = max(1e-6, np.abs(gr).max())
    assert np.quantile(np.abs(g - gr), 0.999) / scale_g <= 1e-4
    assert np.abs(g - gr).max() / scale_g <= 2e-3
    fused, gfused = ops.coupling_nll_grad(_spec_pair(ops, ospec), p, x.cuda())
    assert abs(fused.item() - loss.item()) <= 2e-5 * max(1.0, abs(fused.item()))


def test_layer_hook_reports_every_layer_and_changes_nothing():
    """flowcuda_coupling_nll_grad_layers: the per-flow-layer completion hook of data-parallel training.  One call per
    layer with that layer's slice of the flat gradient and a stream on which it is complete; loss and gradient are the
    bits of the plain entry point; both engines (fused toy kernels: all layers at once on the caller's stream)."""
    from normalizing_flows_b200 import _lib, ops
    from normalizing_flows_b200.data_parallel import LayerwiseAllReduce

    for case, prec in (("rqs_d40_uniform", _lib.PRECISION_BF16X3), ("rqs_d40_uniform", 0), ("rqs_mnist_shape", _lib.PRECISION_BF16X3)):
        ospec, scale, xs, n = DENSE_GRAD[case]
        flat = O.flatten_layers(O.init_coupling_params(ospec, seed=5, scale=scale))
        x = torch.rand(n, ospec.event_dim, generator=torch.Generator().manual_seed(1), dtype=torch.float64).float().cuda()
        spec = _spec_pair(ops, ospec, precision=prec)
        p = G(flat)
        loss0, grad0 = ops.coupling_nll_grad(spec, p, x)
        seen, sums = [], []

        def hook(layer, g, stream):
            assert isinstance(stream, torch.cuda.Stream)
            seen.append((layer, g.data_ptr(), g.numel()))
            with torch.cuda.stream(stream):  # ordered after the layer's last gradient kernel
                sums.append(g.double().sum())

        loss1, grad1 = ops.coupling_nll_grad(spec, p, x, on_layer=hook)
        torch.cuda.synchronize()
        per = p.numel() // ospec.num_layers
        assert sorted(set(s[0] for s in seen)) == list(range(ospec.num_layers))
        assert [s[0] for s in seen] == sorted(s[0] for s in seen)  # layers are reported in order
        for l in range(ospec.num_layers):  # a layer's slices are disjoint and cover exactly its parameters
            spans = sorted(((s[1] - grad1.data_ptr()) // 4, s[2]) for s in seen if s[0] == l)
            pos = l * per
            for lo, cnt in spans:
                assert lo == pos and cnt > 0
                pos += cnt
            assert pos == (l + 1) * per
        assert torch.equal(loss0, loss1) and torch.equal(grad0, grad1)
        for (l, ptr, cnt), sm in zip(seen, sums):  # every slice really was complete on the stream it was reported on
            lo = (ptr - grad1.data_ptr()) // 4
            assert sm.item() == grad1[lo:lo + cnt].double().sum().item()
        # single process: the all-reduce helper is a no-op and wait() returns at once
        ar = LayerwiseAllReduce()
        loss2, grad2 = ops.coupling_nll_grad(spec, p, x, on_layer=ar)
        ar.wait()
        assert torch.equal(grad0, grad2)

    def bad(layer, g, stream):
        raise ValueError("from the hook")

    with pytest.raises(ValueError):
        ops.coupling_nll_grad(spec, p, x, on_layer=bad)
    ospec, scale, xs = COUPLING_CASES["rqs_toy"]  # the fused toy kernels report all layers after the one kernel pair
    flat = O.flatten_layers(O.init_coupling_params(ospec, seed=5, scale=scale))
    calls = []
    ops.coupling_nll_grad(_spec_pair(ops, ospec), G(flat), torch.randn(300, 2).cuda(), on_layer=lambda l, g, st: calls.append(l))
    assert calls == list(range(ospec.num_layers))


def test_nll_grad_can_be_captured_in_a_cuda_graph():
    """flowcuda_coupling_nll_grad forks its weight-gradient jobs onto an internal stream and joins them by events, keeps
    no host state between calls and allocates nothing: a caller may capture it (torch.cuda.graph here, an XLA command
    buffer behind the FFI adapter) and replay it.  The replayed loss and gradient are the eager call's bits, also after
    the parameters have been updated in place between replays (tools/train_graph.py times the replayed step)."""
    from normalizing_flows_b200 import _lib, ops

    for case, prec in (("rqs_mnist_shape", _lib.PRECISION_BF16X3), ("rqs_d40_uniform", _lib.PRECISION_BF16X3), ("rqs_d7_k8", 0)):
        ospec, scale, xs, n = DENSE_GRAD[case]
        flat = O.flatten_layers(O.init_coupling_params(ospec, seed=5, scale=scale))
        g = torch.Generator().manual_seed(3)
        if ospec.base == "uniform":
            x = torch.rand(n, ospec.event_dim, generator=g, dtype=torch.float64).float().cuda()
        else:
            x = (torch.randn(n, ospec.event_dim, generator=g, dtype=torch.float64) * xs).float().cuda()
        spec = _spec_pair(ops, ospec, precision=prec)
        p = G(flat)
        loss0, grad0 = ops.coupling_nll_grad(spec, p, x)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph, stream=side):
            loss1, grad1 = ops.coupling_nll_grad(spec, p, x)
        torch.cuda.current_stream().wait_stream(side)
        for _ in range(2):
            grad1.zero_()
            graph.replay()
            torch.cuda.synchronize()
            assert torch.equal(loss0, loss1) and torch.equal(grad0, grad1), case
        p.add_(grad0, alpha=-1e-3)  # the graph reads the parameters where they live
        loss2, grad2 = ops.coupling_nll_grad(spec, p, x)
        graph.replay()
        torch.cuda.synchronize()
        assert torch.equal(loss2, loss1) and torch.equal(grad2, grad1), case
        assert not torch.equal(grad0, grad1)
