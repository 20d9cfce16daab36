"""Host mirror of the reference's toy-density batch generators, with the draw on the device.

Same names, arguments and batch layout (`[batch_size, 2]`, columns `[x1, x2]`) as
`src/densities/banana.py:31-41`, `src/densities/energies.py:47-127` and
`src/densities/gaussian_blobs.py:34-59`; each generator yields `num_batches` CUDA tensors.  The
reference threads a haiku `PRNGSequence(seed)` through `tfd...sample`; here batch b is samples
`[b*batch_size, (b+1)*batch_size)` of the Philox stream keyed by `seed` (csrc/densities.cu), so a
data-parallel rank can take its own slice with `rank`/`world_size` and no host step is left in
the training loop.  `prepare_data` is `src/maf_mnist.py:101-106`.
"""
from __future__ import annotations

from typing import Dict, Iterator, Optional

import torch

from . import ops


def _batches(kind: str, seed: int, batch_size: int, num_batches: int, rank: int, world_size: int, device,
             **kw) -> Iterator[torch.Tensor]:
    if batch_size % world_size:
        raise ValueError(f"batch_size {batch_size} is not divisible by world_size {world_size}")
    local = batch_size // world_size
    for b in range(num_batches):
        yield ops.density_sample(kind, seed, local, offset=b * batch_size + rank * local, device=device, **kw)


def make_dataset_banana(seed: int, batch_size: int = 8, num_batches: int = 1, *, rank: int = 0, world_size: int = 1,
                        device=None):
    """x2 ~ N(0, sqrt 3), x1 ~ N(x2^2/3 - 2, sqrt(3)/2) (banana.py:31-41)."""
    return _batches("banana", seed, batch_size, num_batches, rank, world_size, device)


def make_dataset_energy_1(seed: int, batch_size: int = 8, num_batches: int = 1, *, rank: int = 0, world_size: int = 1,
                          device=None):
    return _batches("energy_1", seed, batch_size, num_batches, rank, world_size, device)


def make_dataset_energy_2(seed: int, batch_size: int = 8, num_batches: int = 1, *, rank: int = 0, world_size: int = 1,
                          device=None):
    return _batches("energy_2", seed, batch_size, num_batches, rank, world_size, device)


def make_dataset_energy_3(seed: int, batch_size: int = 8, num_batches: int = 1, *, rank: int = 0, world_size: int = 1,
                          device=None):
    return _batches("energy_3", seed, batch_size, num_batches, rank, world_size, device)


def make_dataset_energy_4(seed: int, batch_size: int = 8, num_batches: int = 1, *, rank: int = 0, world_size: int = 1,
                          device=None):
    return _batches("energy_4", seed, batch_size, num_batches, rank, world_size, device)


def make_dataset_gaussian_blobs(seed: int, batch_size: int = 8, num_batches: int = 1, num_blobs: int = 8,
                                r: float = 2.0, *, rank: int = 0, world_size: int = 1, device=None):
    return _batches("gaussian_blobs", seed, batch_size, num_batches, rank, world_size, device, num_blobs=num_blobs, r=r)


def prepare_data(batch: Dict[str, torch.Tensor], prng_key: Optional[int] = None, offset: int = 0) -> torch.Tensor:
    """`(image + U[0,1)) / 256`, or `image / 256` without a key (maf_mnist.py:101-106).
    `batch["image"]` is uint8, on the device or in (pinned) host memory; only the bytes cross PCIe —
    a quarter of the float32 image.  `prng_key` is an integer seed of the device Philox stream."""
    img = batch["image"]
    if img.dtype != torch.uint8:
        raise TypeError(f"prepare_data expects uint8 images, got {img.dtype}")
    if not img.is_cuda:
        img = img.contiguous().to("cuda", non_blocking=True)
    if prng_key is None:
        return img.to(torch.float32) * (1.0 / 256.0)
    return ops.dequantize_u8(img.contiguous(), int(prng_key), offset)
