"""GPU: row #8 — the MADE-conditioned masked autoregressive affine flow (csrc/made.cu) against the
float64 oracle (oracle/flows_oracle.py:510-606; PARITY UNPINNED — the reference has no MADE code,
SURVEY App. A.7).  The density pass runs on the tensor cores (bf16x3): north_star tolerance 1e-3
absolute on log_prob, held here at 2e-4·max(1,|ref|); the sequential sampler is FP32."""
import numpy as np
import pytest
import torch

from oracle import flows_oracle as O
from test_gpu_parity import G, close

pytestmark = pytest.mark.gpu

CASES = {
    "d6": (O.MadeSpec(6, 2, [16, 16]), 0.5),
    "toy_d2": (O.MadeSpec(2, 8, [64, 64]), 0.3),
    "d50_ragged": (O.MadeSpec(50, 3, [40, 70]), 0.4),
    "d300_wide": (O.MadeSpec(300, 2, [512, 300]), 0.3),
    "d1": (O.MadeSpec(1, 2, [8, 8]), 0.5),
    "d32_fused": (O.MadeSpec(32, 4, [40, 40]), 0.4),   # D % 16 == 0: the epilogue writes the next layer's image
    "d160_fused": (O.MadeSpec(160, 3, [96, 200]), 0.3),
}


def _spec(ops, ospec):
    return ops.make_made_spec(ospec.event_dim, ospec.num_layers, ospec.hidden)


@pytest.mark.parametrize("case", sorted(CASES))
@pytest.mark.parametrize("n", [1, 77, 300])
def test_made_log_prob_parity(case, n):
    from normalizing_flows_b200 import ops

    ospec, scale = CASES[case]
    params = O.init_made_params(ospec, seed=3, scale=scale)
    flat = O.flatten_layers(params)
    spec = _spec(ops, ospec)
    assert ops.made_param_count(spec) == flat.numel()
    x = torch.randn(n, ospec.event_dim, generator=torch.Generator().manual_seed(n), dtype=torch.float64).float()
    lp, u, ld = ops.made_log_prob(spec, G(flat), x.cuda(), want_u=True, want_log_det=True)
    lpref = O.maf_log_prob(ospec, params, x.double())
    close(lp, lpref.numpy(), tol=2e-4, what="log_prob")
    # u and log_det from the oracle: log_prob = base(u) + log_det
    uref = _oracle_u(ospec, params, x.double())
    close(u, uref.numpy(), tol=2e-4, what="u")
    ldref = lpref - (-0.5 * uref * uref - 0.5 * O.LOG_2PI).sum(-1)
    close(ld, ldref.numpy(), tol=2e-4, what="log_det")


@pytest.mark.parametrize("case", sorted(CASES))
def test_made_log_prob_fp32_partner(case):
    """flowcuda_made_log_prob_fp32: the density pass with every product in FP32 (CUDA cores) against the float64 oracle
    at the FP32 tolerance (1e-5 scaled), and the tensor-core pass against IT inside its 1e-3 budget."""
    from normalizing_flows_b200 import ops
    from parity_util import hold

    ospec, scale = CASES[case]
    params = O.init_made_params(ospec, seed=3, scale=scale)
    flat = G(O.flatten_layers(params))
    spec = _spec(ops, ospec)
    n = 300
    x = torch.randn(n, ospec.event_dim, generator=torch.Generator().manual_seed(n), dtype=torch.float64).float()
    lp32, u32 = ops.made_log_prob_fp32(spec, flat, x.cuda(), want_u=True)
    hold(f"made_fp32_{case}_log_prob", lp32, O.maf_log_prob(ospec, params, x.double()), 1e-5, kind="toy")
    hold(f"made_fp32_{case}_u", u32, _oracle_u(ospec, params, x.double()), 1e-5, kind="toy")
    lp, _, _ = ops.made_log_prob(spec, flat, x.cuda())
    assert (lp - lp32).abs().max().item() <= 1e-3


def _oracle_u(ospec, params, x):
    v = x
    for l in reversed(range(ospec.num_layers)):
        vp = O._maf_perm(l, v)
        mu, alpha = O.made_conditioner(ospec, params[l], vp)
        v = O._maf_perm(l, (vp - mu) * torch.exp(-alpha))
    return v


@pytest.mark.parametrize("case", ["d6", "toy_d2", "d50_ragged", "d300_wide", "d32_fused"])
@pytest.mark.parametrize("n", [1, 13, 40])
def test_made_sample_parity(case, n):
    """Sequential sampler (FP32, incremental pre-activations) vs the oracle's D-pass loop."""
    from normalizing_flows_b200 import ops

    ospec, scale = CASES[case]
    params = O.init_made_params(ospec, seed=4, scale=scale)
    flat = O.flatten_layers(params)
    z = torch.randn(n, ospec.event_dim, generator=torch.Generator().manual_seed(7 + n), dtype=torch.float64).float()
    x = ops.made_sample(_spec(ops, ospec), G(flat), z.cuda())
    xref = O.maf_sample(ospec, params, z.double())
    close(x, xref.numpy(), tol=2e-5, what="sample")


def test_made_mnist_shape_log_prob_and_round_trip():
    """D=784, H=512 (the north_star MAF-MNIST shape; 2 flow layers to bound the CPU oracle):
    |log_prob - float64 oracle| <= 1e-3 absolute, and log_prob's u inverts the sampler."""
    from normalizing_flows_b200 import ops

    ospec = O.MadeSpec(784, 2, [512, 512])
    params = O.init_made_params(ospec, seed=5, scale=0.3)
    flat = G(O.flatten_layers(params))
    spec = _spec(ops, ospec)
    gen = torch.Generator().manual_seed(0)
    x = torch.randn(200, 784, generator=gen, dtype=torch.float64).float()
    lp, _, _ = ops.made_log_prob(spec, flat, x.cuda())
    ref = O.maf_log_prob(ospec, params, x.double()).numpy()
    err = np.abs(lp.double().cpu().numpy() - ref).max()
    assert err <= 1e-3, err
    z = torch.randn(24, 784, generator=gen, dtype=torch.float64).float()
    xs = ops.made_sample(spec, flat, z.cuda())
    xref = O.maf_sample(ospec, params, z.double())
    close(xs, xref.numpy(), tol=5e-5, what="sample d784")
    _, u, _ = ops.made_log_prob(spec, flat, xs, want_u=True)
    assert (u.cpu() - z).abs().max().item() <= 2e-3


def test_made_full_size_round_trip_and_autoregressive_property():
    """Size-independent properties at the bench shape (8 layers, D=784, batch 4096): sample then
    density pass returns the noise; perturbing x_j changes u_i only for i >= j in the layer's ordering."""
    from normalizing_flows_b200 import ops

    ospec = O.MadeSpec(784, 8, [512, 512])
    params = O.init_made_params(ospec, seed=6, dtype=torch.float32, scale=0.2)
    flat = O.flatten_layers(params).float().cuda()
    spec = _spec(ops, ospec)
    z = torch.randn(4096, 784, generator=torch.Generator().manual_seed(1)).cuda()
    x = ops.made_sample(spec, flat, z)
    assert torch.isfinite(x).all()
    lp, u, ld = ops.made_log_prob(spec, flat, x, want_u=True, want_log_det=True)
    assert (u - z).abs().max().item() <= 5e-3
    base = (-0.5 * u.double() ** 2 - 0.5 * O.LOG_2PI).sum(-1)
    assert ((lp.double() - base - ld.double()).abs() <= 1e-3 * base.abs().clamp(min=1)).all()
    # single layer: u_i must not depend on x_j for j > i (even layer = natural ordering)
    one = ops.make_made_spec(784, 1, [512, 512])
    xa = x[:64].clone()
    xb = xa.clone()
    xb[:, 400:] += 1.0
    _, ua, _ = ops.made_log_prob(one, flat, xa, want_u=True)
    _, ub, _ = ops.made_log_prob(one, flat, xb, want_u=True)
    assert torch.equal(ua[:, :400], ub[:, :400])
    assert (ua[:, 400:] - ub[:, 400:]).abs().min().item() > 0


def test_made_rejects_bad_arguments():
    from normalizing_flows_b200 import _lib, ops

    spec = ops.make_made_spec(6, 2, [16])
    flat = torch.zeros(ops.made_param_count(spec), device="cuda")
    with pytest.raises(_lib.FlowCudaError):
        ops.made_sample(spec, flat, torch.zeros(4, 6, device="cuda"))  # sampler needs two hidden layers
    with pytest.raises(_lib.FlowCudaError):
        ops.made_log_prob(spec, flat, torch.zeros(4, 6))  # CPU tensor: no CPU path
    assert ops.made_log_prob(spec, flat, torch.zeros(0, 6, device="cuda"))[0].numel() == 0


GRAD_CASES = {
    "d6": (O.MadeSpec(6, 2, [16, 16]), 0.5, 33),
    "d50_ragged": (O.MadeSpec(50, 3, [40, 70]), 0.4, 100),
    "d32": (O.MadeSpec(32, 4, [40, 40]), 0.4, 64),
    "toy_d2": (O.MadeSpec(2, 8, [64, 64]), 0.3, 257),
    "d784": (O.MadeSpec(784, 2, [512, 512]), 0.3, 48),
    "d50_splitk": (O.MadeSpec(50, 2, [40, 70]), 0.4, 1500),   # n >= 1024: split-K weight-gradient path
    "d300_splitk": (O.MadeSpec(300, 2, [200, 260]), 0.3, 1100),
}


@pytest.mark.parametrize("engine", ["fp32", "bf16x3"])
@pytest.mark.parametrize("case", sorted(GRAD_CASES))
def test_made_nll_grad_matches_float64_autograd(case, engine):
    """Hand-written reverse mode (csrc/made_grad.cu) vs autograd of the float64 oracle:
    loss within 1e-5 relative, every gradient entry within 1e-4 of the largest entry of its tensor,
    masked-out weights exactly zero; bit-identical when run twice.  Both GEMM engines (FP32 CUDA cores,
    tcgen05 bf16x3 — mgtc.cuh) are held to the same bound."""
    from normalizing_flows_b200 import _lib, ops

    ospec, scale, n = GRAD_CASES[case]
    params = O.init_made_params(ospec, seed=8, scale=scale)
    leaves = [t for layer in params for (W, b) in layer for t in (W, b)]
    x = torch.randn(n, ospec.event_dim, generator=torch.Generator().manual_seed(3), dtype=torch.float64).float().double()

    def lp_fn(lv):
        it = iter(lv)
        p = [[(next(it), next(it)) for _ in layer] for layer in params]
        return O.maf_log_prob(ospec, p, x)

    loss_ref, grads_ref = O.nll_and_grads(lp_fn, leaves)
    spec = ops.make_made_spec(ospec.event_dim, ospec.num_layers, ospec.hidden,
                              precision=_lib.PRECISION_BF16X3 if engine == "bf16x3" else _lib.PRECISION_FP32)
    flat = G(O.flatten_layers(params))
    loss, grad = ops.made_nll_grad(spec, flat, x.float().cuda())
    loss2, grad2 = ops.made_nll_grad(spec, flat, x.float().cuda())
    assert torch.equal(grad, grad2) and torch.equal(loss, loss2)
    assert abs(loss.item() - loss_ref.item()) <= 1e-5 * max(1.0, abs(loss_ref.item()))
    g = grad.double().cpu()
    off = 0
    masks = [m for _ in range(ospec.num_layers) for m in ospec.masks()]
    mi = 0
    for k, ref in enumerate(grads_ref):
        got = g[off:off + ref.numel()].reshape(ref.shape)
        off += ref.numel()
        tol = 1e-4 * max(ref.abs().max().item(), 1e-6)
        assert (got - ref).abs().max().item() <= tol, (case, k, (got - ref).abs().max().item(), tol)
        if ref.dim() == 2:
            assert (got[masks[mi] == 0] == 0).all()
            mi += 1
    assert off == g.numel()


def test_deep16_made_stack_d784():
    """configs[4] / SURVEY C5: all 16 MADE layers at D=784, H=512: |log_prob - float64 oracle| <= 1e-3 absolute."""
    from normalizing_flows_b200 import ops

    ospec = O.MadeSpec(784, 16, [512, 512])
    params = O.init_made_params(ospec, seed=9, scale=0.1)
    flat = G(O.flatten_layers(params))
    gen = torch.Generator().manual_seed(0)
    x = ((torch.randint(0, 256, (96, 784), generator=gen).double() + torch.rand(96, 784, generator=gen, dtype=torch.float64)) / 256).float()
    lp, _, _ = ops.made_log_prob(_spec(ops, ospec), flat, x.cuda())
    ref = O.maf_log_prob(ospec, params, x.double()).numpy()
    err = np.abs(lp.double().cpu().numpy() - ref).max()
    assert err <= 1e-3, err


def test_packed_weights_serving_path_is_bit_identical():
    """flowcuda_*_pack_weights + *_log_prob_packed (serving: parameters unchanged between calls): same bits as
    the repacking call, for several batch sizes against ONE packed workspace; a too-large batch is refused;
    stale images are the caller's responsibility (prepare again after an update) and are detectable."""
    from normalizing_flows_b200 import _lib, flows, ops

    ospec = O.MadeSpec(160, 3, [96, 200])
    params = O.init_made_params(ospec, seed=12, scale=0.3)
    flat = G(O.flatten_layers(params))
    spec = _spec(ops, ospec)
    pk = ops.made_pack_weights(spec, flat, 500)
    for n in (1, 130, 500):
        x = torch.randn(n, 160, generator=torch.Generator().manual_seed(n)).cuda()
        a = ops.made_log_prob(spec, flat, x, want_u=True, want_log_det=True)
        b = ops.made_log_prob(spec, flat, x, want_u=True, want_log_det=True, packed=pk)
        assert all(torch.equal(p, q) for p, q in zip(a, b))
    with pytest.raises(_lib.FlowCudaError):
        ops.made_log_prob(spec, flat, torch.zeros(501, 160, device="cuda"), packed=pk)
    # the host mirror: prepare() / invalidate()
    model = flows.make_maf_model((160,), 3, [96, 200], params=flat)
    x = torch.randn(77, 160, generator=torch.Generator().manual_seed(5)).cuda()
    ref = model.log_prob(x)
    assert torch.equal(model.prepare(128).log_prob(x), ref)
    model.bijector.params.mul_(1.5)                       # an update the library cannot see
    stale = model.log_prob(x)
    model.invalidate()
    fresh = model.log_prob(x)
    assert not torch.equal(stale, fresh)
    assert torch.equal(model.prepare(128).log_prob(x), fresh)
    # coupling stacks (bf16x3)
    cspec_o = O.CouplingSpec(40, 2, [32, 32, 32], "rqs", 8, 0.0, 1.0, "uniform")
    cp = G(O.flatten_layers(O.init_coupling_params(cspec_o, seed=3, scale=0.3)))
    cspec = ops.make_spec(40, 2, [32, 32, 32], 2, 8, 0.0, 1.0, _lib.BASE_UNIFORM01, _lib.PRECISION_BF16X3)
    cpk = ops.coupling_pack_weights(cspec, cp, 300)
    xc = torch.rand(300, 40, generator=torch.Generator().manual_seed(1)).cuda()
    a = ops.coupling_log_prob(cspec, cp, xc, want_z=True, want_log_det=True)
    b = ops.coupling_log_prob(cspec, cp, xc, want_z=True, want_log_det=True, packed=cpk)
    assert all(torch.equal(p, q) for p, q in zip(a, b))
    sa, _ = ops.coupling_sample(cspec, cp, xc[:50])
    sb, _ = ops.coupling_sample(cspec, cp, xc[:50], packed=cpk)
    assert torch.equal(sa, sb)
    with pytest.raises(_lib.FlowCudaError):
        ops.coupling_pack_weights(ops.make_spec(40, 2, [32, 32, 32], 2, 8, 0.0, 1.0, _lib.BASE_UNIFORM01, 0), cp, 10)


@pytest.mark.parametrize("engine", ["fp32", "bf16x3"])
def test_made_trainer_step_reduces_the_loss(engine):
    """DataParallelTrainer (value_and_grad -> [all-reduce] -> Adam, src/maf_mnist.py:155-166) on the MADE flow,
    single rank: the NLL of a fixed batch falls monotonically over a few steps, on both GEMM engines."""
    from normalizing_flows_b200 import _lib, data_parallel as dp, ops

    ospec = O.MadeSpec(32, 3, [48, 48])
    spec = ops.make_made_spec(32, 3, [48, 48], precision=_lib.PRECISION_BF16X3 if engine == "bf16x3" else 0)
    params = G(O.flatten_layers(O.init_made_params(ospec, seed=4, scale=0.1))).clone()
    x = (torch.randn(512, 32, generator=torch.Generator().manual_seed(0)) * 1.5 + 0.5).cuda()
    trainer = dp.DataParallelTrainer(params, dp.cuda_grad_fn(spec), dp.cuda_adam_fn, lr=2e-3)
    losses = [trainer.step(x).item() for _ in range(8)]
    assert all(np.isfinite(losses)) and all(b < a for a, b in zip(losses, losses[1:])), losses
    lp, _, _ = ops.made_log_prob(spec, params, x)
    assert abs(-lp.mean().item() - losses[-1]) < abs(losses[0] - losses[-1])  # the parameters really moved
