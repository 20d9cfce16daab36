"""GPU: the device data step (csrc/densities.cu) against oracle/philox_oracle.py.

The Philox word stream is integer work: every sample is reproduced from the same words, so the
float outputs must match the float64 restatement to FP32 rounding (1e-5·max(1,|ref|)); the
dequantised pixels, whose arithmetic is exact in FP32, must match BIT FOR BIT.  The distributions
themselves are the reference's (src/densities/*.py, src/maf_mnist.py:101-106) and are checked by
moments at full size."""
import numpy as np
import pytest
import torch

from oracle import philox_oracle as P
from test_gpu_parity import close

pytestmark = pytest.mark.gpu

KINDS = ["banana", "energy_1", "energy_2", "energy_3", "energy_4", "gaussian_blobs"]


@pytest.mark.parametrize("kind", KINDS)
@pytest.mark.parametrize("n,offset", [(1, 0), (1000, 0), (4097, 2 ** 33 + 5)])
def test_density_sample_matches_oracle(kind, n, offset):
    from normalizing_flows_b200 import ops

    x = ops.density_sample(kind, 1234567890123, n, offset=offset)
    ref = P.density_sample(ops.DENSITY_KINDS[kind], 1234567890123, offset, n)
    close(x, ref, tol=1e-5, what=kind)


def test_blobs_arguments_and_empty_batch():
    from normalizing_flows_b200 import _lib, ops

    x = ops.density_sample("gaussian_blobs", 7, 5000, num_blobs=5, r=3.5)
    close(x, P.density_sample(P.BLOBS, 7, 0, 5000, num_blobs=5, r=3.5), tol=1e-5, what="blobs5")
    assert ops.density_sample("banana", 1, 0).shape == (0, 2)
    with pytest.raises(_lib.FlowCudaError):
        ops.density_sample(9, 1, 8)
    with pytest.raises(_lib.FlowCudaError):
        ops.density_sample("gaussian_blobs", 1, 8, num_blobs=0)


def test_generators_mirror_the_reference_batching_and_shard_by_rank():
    """make_dataset_*(seed, batch_size, num_batches) yields [batch,2] batches; two ranks' slices
    concatenate to the single-rank batch (no host step, no exchange)."""
    from normalizing_flows_b200 import densities as D

    full = list(D.make_dataset_energy_2(3, batch_size=256, num_batches=3))
    assert len(full) == 3 and all(b.shape == (256, 2) and b.is_cuda for b in full)
    assert not torch.equal(full[0], full[1])
    r0 = list(D.make_dataset_energy_2(3, batch_size=256, num_batches=3, rank=0, world_size=2))
    r1 = list(D.make_dataset_energy_2(3, batch_size=256, num_batches=3, rank=1, world_size=2))
    for b, a0, a1 in zip(full, r0, r1):
        assert torch.equal(torch.cat([a0, a1]), b)
    again = list(D.make_dataset_energy_2(3, batch_size=256, num_batches=3))
    assert all(torch.equal(a, b) for a, b in zip(full, again))
    blobs = next(D.make_dataset_gaussian_blobs(1, batch_size=64, num_blobs=4, r=1.0))
    assert blobs.shape == (64, 2)


def test_full_size_moments():
    """2^22 samples per density (the bench batch): moments implied by the reference definitions."""
    from normalizing_flows_b200 import ops

    n = 1 << 22
    x = ops.density_sample("banana", 5, n).double()
    assert abs(x[:, 1].var().item() - 3.0) < 0.02 and abs(x[:, 0].mean().item() + 1.0) < 0.01
    assert abs(x[:, 0].var().item() - 2.75) < 0.03
    x = ops.density_sample("gaussian_blobs", 5, n).double()
    assert x.mean(0).abs().max().item() < 0.01 and (x.var(0) - 2.09).abs().max().item() < 0.02
    x = ops.density_sample("energy_2", 5, n).double()
    assert x[:, 0].min().item() >= -1.5 * np.pi and x[:, 0].max().item() < 1.5 * np.pi
    assert abs((x[:, 1] - torch.sin(np.pi * x[:, 0] / 2)).var().item() - 0.25) < 0.003
    assert torch.isfinite(x).all()


@pytest.mark.parametrize("shape,offset", [((3, 28, 28, 1), 0), ((1, 5), 0), ((64, 28, 28, 1), 784 * 64), ((0, 28, 28, 1), 0)])
def test_dequantize_is_bit_exact(shape, offset):
    from normalizing_flows_b200 import densities as D, ops

    b = np.random.default_rng(1).integers(0, 256, size=shape, dtype=np.uint8)
    got = ops.dequantize_u8(torch.from_numpy(b).cuda(), 42, offset)
    ref = P.dequantize_u8(b, 42, offset)
    assert got.shape == tuple(shape)
    assert np.array_equal(got.cpu().numpy(), ref)
    if b.size:
        via_host = D.prepare_data({"image": torch.from_numpy(b)}, 42, offset)
        assert torch.equal(via_host, got)
        plain = D.prepare_data({"image": torch.from_numpy(b)})
        assert np.array_equal(plain.cpu().numpy(), b.astype(np.float32) / 256)


def test_dequantize_rejects_bad_arguments():
    from normalizing_flows_b200 import _lib, ops

    with pytest.raises(_lib.FlowCudaError):
        ops.dequantize_u8(torch.zeros(16, dtype=torch.uint8), 1)  # host tensor: no CPU path
    with pytest.raises(_lib.FlowCudaError):
        ops.dequantize_u8(torch.zeros(16, dtype=torch.uint8, device="cuda"), 1, offset=3)


def test_base_noise_and_seeded_sample():
    from normalizing_flows_b200 import distrax_api as dx, ops

    z = ops.base_noise((1001, 2), 77)
    close(z, P.base_noise(0, 77, 0, 2002).reshape(1001, 2), tol=1e-5, what="normal")
    u = ops.base_noise((37,), 77, uniform=True, offset=8)
    assert np.array_equal(u.cpu().numpy(), P.base_noise(1, 77, 8, 37).astype(np.float32))
    # Transformed.sample(seed=int) draws its base noise from the same kernel
    base = dx.Independent(dx.Normal(torch.zeros(2), torch.ones(2)), 1)
    s = base.sample(seed=77, sample_shape=(1001,))
    assert torch.equal(s, z)
    big = ops.base_noise((1 << 22,), 3).double()
    assert abs(big.mean().item()) < 3e-3 and abs(big.var().item() - 1) < 5e-3


@pytest.mark.parametrize("n", [1, 2, 7, 100001])
def test_orthogonal_projection_2d(n):
    """src/rotation_flow.py:13-37: forward = x @ R, inverse = x @ R.T, 'log-det' = the constant 1 (the
    reference's behaviour, flagged); logdet='exact' gives zeros of the batch shape."""
    from normalizing_flows_b200.rotation_flow import OrthogonalProjection2D

    theta = np.pi / 4
    bij = OrthogonalProjection2D(theta=theta)
    x = torch.randn(n, 2, generator=torch.Generator().manual_seed(n)) * torch.tensor([1.5, 0.5])
    R = np.array([[np.cos(theta), -np.sin(theta)], [np.sin(theta), np.cos(theta)]]).T
    y, ld = bij.forward_and_log_det(x.cuda())
    assert ld == 1
    close(y, x.double().numpy() @ R, tol=1e-6, what="forward")
    xb, ld = bij.inverse_and_log_det(y)
    assert ld == 1
    close(xb, x.double().numpy(), tol=1e-6, what="round trip")
    close(bij.inverse(x.cuda()), x.double().numpy() @ R.T, tol=1e-6, what="inverse")
    exact = OrthogonalProjection2D(theta=theta, logdet="exact")
    _, ld0 = exact.forward_and_log_det(x.cuda())
    assert ld0.shape == (n,) and (ld0 == 0).all()
    assert tuple(bij.R.shape) == (2, 2) and abs(bij.R[0, 1].item() - np.sin(theta)) < 1e-7


def test_npz_params_drive_the_flow(tmp_path):
    """A haiku-keyed .npz written from known parameters reproduces the oracle's log_prob through
    flows.make_flow_model (params wire format)."""
    from normalizing_flows_b200 import flows, params_io as io
    from oracle import flows_oracle as O

    spec = O.RQS_TOY
    params = O.init_coupling_params(spec, seed=21, scale=0.3)
    names = io.haiku_module_names(spec.num_layers, len(spec.hidden))
    path = str(tmp_path / "toy.npz")
    np.savez(path, **{f"{mod}/{k}": v.numpy().astype(np.float32)
                      for layer, mods in zip(params, names) for (W, b), mod in zip(layer, mods)
                      for k, v in (("w", W), ("b", b))})
    flat = io.load_haiku_npz(path, spec.num_layers, spec.event_dim, spec.hidden, spec.cond_out)
    model = flows.make_flow_model((2,), spec.num_layers, spec.hidden, transform="rqs", num_bins=12,
                                  range_min=-8.0, range_max=8.0, params=flat)
    x = torch.randn(500, 2, generator=torch.Generator().manual_seed(2), dtype=torch.float64).float()
    ref = O.coupling_flow_log_prob(spec, params, x.double())
    close(model.log_prob(x.cuda()), ref.numpy(), tol=1e-5, what="log_prob from npz")
