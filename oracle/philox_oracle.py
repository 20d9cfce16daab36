"""TEST INFRASTRUCTURE ONLY — CPU restatement of the device data step.

Only tests/, __graft_entry__.smoke() and bench.py's CPU baseline may import this module.

What is restated:
  * Philox4x32-10 (Salmon, Moraes, Dror, Shaw 2011, "Parallel random numbers: as easy as 1, 2, 3"),
    pinned below against the Random123 known-answer vectors (tests/test_oracle_golden.py);
  * the reference's toy-density SAMPLERS as distributions — src/densities/banana.py:31-41,
    src/densities/energies.py:12-27,47-127, src/densities/gaussian_blobs.py:34-59 — and
    prepare_data of src/maf_mnist.py:101-106.
PARITY UNPINNED against the reference's random streams: JAX's threefry keys cannot be reproduced
here (JAX is not installable); the word-to-variate mapping below is this repo's own
and is what csrc/densities.cu must match.  The distributions themselves are checked by moments
against the reference's definitions (tests/test_gpu_densities.py).
"""
import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)
S32 = np.uint64(32)

BANANA, ENERGY1, ENERGY2, ENERGY3, ENERGY4, BLOBS = range(6)
TAG_DEQUANT, TAG_BASE = 0xDE0, 0xBA5E


def philox4x32_10(counter, key):
    """counter: four uint32 arrays (or ints), key: (k0, k1) ints -> four uint32 arrays."""
    c = [np.asarray(x, dtype=np.uint64) & MASK for x in counter]
    c = list(np.broadcast_arrays(*c))
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = M0 * c[0]
        p1 = M1 * c[2]
        n0 = ((p1 >> S32) ^ c[1] ^ np.uint64(k0)) & MASK
        n2 = ((p0 >> S32) ^ c[3] ^ np.uint64(k1)) & MASK
        c = [n0, p1 & MASK, n2, p0 & MASK]
        k0 = (k0 + W0) & 0xFFFFFFFF
        k1 = (k1 + W1) & 0xFFFFFFFF
    return [x.astype(np.uint32) for x in c]


def _key(seed):
    seed = int(seed) & 0xFFFFFFFFFFFFFFFF
    return seed & 0xFFFFFFFF, seed >> 32


def _ctr(idx):
    idx = np.asarray(idx, dtype=np.uint64)
    return idx & MASK, idx >> S32


def u01(r):
    """24 high bits -> [0,1)."""
    return (r >> np.uint32(8)).astype(np.float64) * 2.0 ** -24


def u01_open(r):
    return ((r >> np.uint32(8)).astype(np.float64) + 0.5) * 2.0 ** -24


def normal_pair(r0, r1):
    """Box-Muller: radius from r0, angle from r1."""
    rad = np.sqrt(-2.0 * np.log(u01_open(r0)))
    ang = 2.0 * np.pi * u01(r1)
    return rad * np.cos(ang), rad * np.sin(ang)


def w1(z):  # energies.py:12-13
    return np.sin(2 * np.pi * z / 4)


def w2(z):  # energies.py:16-18
    return 3 * np.exp(-0.5 * ((z - 1) / 0.6) ** 2)


def w3(z):  # energies.py:21-27
    return 3 / (1 + np.exp(-(z - 1) / 0.3))


def density_sample(kind, seed, offset, n, num_blobs=8, r=2.0):
    """float64 [n,2]: samples offset..offset+n of the density's stream; column order [x1, x2] as the
    reference stacks them."""
    idx = np.uint64(offset) + np.arange(n, dtype=np.uint64)
    lo, hi = _ctr(idx)
    w = philox4x32_10((lo, hi, kind, 0), _key(seed))
    n0, n1 = normal_pair(w[0], w[1])
    x1u = (u01(w[2]) - 0.5) * (3.0 * np.pi)  # Uniform(-3pi/2, 3pi/2)
    second = (w[3] >> np.uint32(31)) != 0    # Categorical([.5, .5])
    if kind == BANANA:       # banana.py:35-40
        x2 = np.sqrt(3.0) * n0
        x1 = x2 ** 2 / 3 - 2 + np.sqrt(3.0) / 2 * n1
    elif kind == ENERGY1:    # energies.py:47-66
        x2 = 2.0 * n0
        mu = 2 - x2 ** 2 / 32
        x1 = np.where(second, -mu, mu) + 0.3 * n1
    elif kind == ENERGY2:    # energies.py:69-82
        x1 = x1u
        x2 = w1(x1) + 0.5 * n0
    elif kind == ENERGY3:    # energies.py:85-104
        x1 = x1u
        m1 = w1(x1)
        x2 = np.where(second, m1 - w2(x1), m1) + 0.5 * n0
    elif kind == ENERGY4:    # energies.py:107-127
        x1 = x1u
        m1 = w1(x1)
        x2 = np.where(second, m1 + w3(x1), m1) + 0.5 * n0
    elif kind == BLOBS:      # gaussian_blobs.py:34-59
        c = ((w[3].astype(np.uint64) * np.uint64(num_blobs)) >> S32).astype(np.float64)
        th = 2.0 * np.pi * c / num_blobs
        x1 = r * np.cos(th) + 0.3 * n0
        x2 = r * np.sin(th) + 0.3 * n1
    else:
        raise ValueError(kind)
    return np.stack([x1, x2], axis=-1)


def _words(tag, seed, offset, count):
    """word e%4 of counter e/4 for elements offset..offset+count."""
    el = np.uint64(offset) + np.arange(count, dtype=np.uint64)
    ctr = el >> np.uint64(2)
    uniq, inv = np.unique(ctr, return_inverse=True)
    lo, hi = _ctr(uniq)
    w = np.stack(philox4x32_10((lo, hi, tag, 0), _key(seed)), axis=-1)  # [u,4]
    return w[inv], (el & np.uint64(3)).astype(np.int64), inv


def dequantize_u8(bytes_, seed, offset=0):
    """(byte + U[0,1)) / 256, maf_mnist.py:101-106; float32 arithmetic like the reference."""
    b = np.asarray(bytes_, dtype=np.uint8).reshape(-1)
    w, lane, _ = _words(TAG_DEQUANT, seed, offset, b.size)
    u = u01(w[np.arange(b.size), lane]).astype(np.float32)
    return ((b.astype(np.float32) + u) * np.float32(1 / 256)).reshape(np.shape(bytes_))


def base_noise(uniform, seed, offset, count):
    """float64 N(0,1) (words (0,1) and (2,3) pair up) or U[0,1)."""
    w, lane, _ = _words(TAG_BASE, seed, offset, count)
    if uniform:
        return u01(w[np.arange(count), lane])
    a0, a1 = normal_pair(w[:, 0], w[:, 1])
    b0, b1 = normal_pair(w[:, 2], w[:, 3])
    allv = np.stack([a0, a1, b0, b1], axis=-1)
    return allv[np.arange(count), lane]
