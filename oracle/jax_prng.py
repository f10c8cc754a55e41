"""NumPy restatement of the JAX sampler behind the reference's 128x8 golden vector.

TEST INFRASTRUCTURE ONLY (see oracle/nds_oracle.py).  The reference test
/root/reference/tests/test_crowding_distance.py:24-27 draws its input with

    key = jax.random.PRNGKey(314); x = jax.random.normal(key, (128, 8))

JAX is not installable here, so the input is regenerated from the published
algorithms JAX uses for that call (JAX 0.4.x defaults, the version range the
reference pins in pyproject.toml:34 -- ``jax_threefry_partitionable`` off):

  * Threefry-2x32, 20 rounds (Salmon et al., "Parallel random numbers: as easy
    as 1, 2, 3", SC'11), keyed with PRNGKey(seed) = (seed >> 32, seed & 0xffffffff);
    the counters 0..n-1 are split into two halves that form the two input words;
  * uniform floats from the top 23 bits (mantissa trick), mapped to
    [nextafter(-1, 0), 1);
  * normal = sqrt(2) * erfinv(u), with the single-precision polynomial of
    M. Giles, "Approximating the erfinv function" (GPU Computing Gems), which is
    what XLA's f32 ErfInv evaluates.

If this restatement were wrong in any step the 90 golden distances of the
reference test could not be reproduced; tests/test_oracle_golden.py checks them.
"""

from __future__ import annotations

import numpy as np

_ROT = ((13, 15, 26, 6), (17, 29, 16, 24))


def _rotl(x: np.ndarray, r: int) -> np.ndarray:
    return ((x << np.uint32(r)) | (x >> np.uint32(32 - r))).astype(np.uint32)


def threefry2x32(key: tuple[int, int], x0: np.ndarray, x1: np.ndarray):
    """Threefry-2x32-20 on two uint32 word arrays."""
    k0 = np.uint32(key[0])
    k1 = np.uint32(key[1])
    ks = (k0, k1, np.uint32(k0 ^ k1 ^ np.uint32(0x1BD11BDA)))
    x0 = (x0.astype(np.uint32) + ks[0]).astype(np.uint32)
    x1 = (x1.astype(np.uint32) + ks[1]).astype(np.uint32)
    for g in range(5):
        for r in _ROT[g % 2]:
            x0 = (x0 + x1).astype(np.uint32)
            x1 = _rotl(x1, r)
            x1 = x1 ^ x0
        x0 = (x0 + ks[(g + 1) % 3]).astype(np.uint32)
        x1 = (x1 + ks[(g + 2) % 3] + np.uint32(g + 1)).astype(np.uint32)
    return x0, x1


def prng_key(seed: int) -> tuple[int, int]:
    return ((seed >> 32) & 0xFFFFFFFF, seed & 0xFFFFFFFF)


def random_bits32(key: tuple[int, int], n: int) -> np.ndarray:
    """``threefry_random_bits`` for 32-bit words, non-partitionable layout."""
    count = np.arange(n, dtype=np.uint32)
    odd = n % 2
    if odd:
        count = np.concatenate([count, np.zeros(1, np.uint32)])
    half = count.size // 2
    with np.errstate(over="ignore"):
        y0, y1 = threefry2x32(key, count[:half], count[half:])
    out = np.concatenate([y0, y1])
    return out[:n]


def _erfinv_f32(x: np.ndarray) -> np.ndarray:
    """Giles' single-precision erfinv, evaluated in float32."""
    f = np.float32
    x = x.astype(np.float32)
    w = (-np.log1p((-x * x).astype(np.float32))).astype(np.float32)
    lt = w < f(5.0)
    w1 = (w - f(2.5)).astype(np.float32)
    w2 = (np.sqrt(np.maximum(w, f(0))).astype(np.float32) - f(3.0)).astype(np.float32)
    c1 = [2.81022636e-08, 3.43273939e-07, -3.5233877e-06, -4.39150654e-06, 0.00021858087,
          -0.00125372503, -0.00417768164, 0.246640727, 1.50140941]
    c2 = [-0.000200214257, 0.000100950558, 0.00134934322, -0.00367342844, 0.00573950773,
          -0.0076224613, 0.00943887047, 1.00167406, 2.83297682]
    ww = np.where(lt, w1, w2).astype(np.float32)
    p = np.where(lt, f(c1[0]), f(c2[0])).astype(np.float32)
    for a, b in zip(c1[1:], c2[1:]):
        p = (np.where(lt, f(a), f(b)) + p * ww).astype(np.float32)
    return (p * x).astype(np.float32)


def uniform_f32(key: tuple[int, int], n: int, minval: np.float32, maxval: np.float32) -> np.ndarray:
    bits = random_bits32(key, n)
    float_bits = (bits >> np.uint32(9)) | np.float32(1.0).view(np.uint32)
    floats = float_bits.view(np.float32) - np.float32(1.0)
    scaled = (floats * np.float32(maxval - minval) + minval).astype(np.float32)
    return np.maximum(minval, scaled)


def uniform_array_f32(seed: int, shape: tuple[int, ...], minval, maxval) -> np.ndarray:
    """``jax.random.uniform(jax.random.PRNGKey(seed), shape, minval=minval, maxval=maxval)`` in float32,
    with ``minval`` / ``maxval`` broadcast against ``shape`` (the call of tests/test_lsmop.py:10-14)."""
    n = int(np.prod(shape))
    bits = random_bits32(prng_key(seed), n)
    float_bits = (bits >> np.uint32(9)) | np.float32(1.0).view(np.uint32)
    floats = (float_bits.view(np.float32) - np.float32(1.0)).reshape(shape)
    lo = np.broadcast_to(np.asarray(minval, dtype=np.float32), shape)
    hi = np.broadcast_to(np.asarray(maxval, dtype=np.float32), shape)
    return np.maximum(lo, (floats * (hi - lo) + lo).astype(np.float32))


def normal_f32(seed: int, shape: tuple[int, ...]) -> np.ndarray:
    """``jax.random.normal(jax.random.PRNGKey(seed), shape)`` in float32."""
    n = int(np.prod(shape))
    lo = np.nextafter(np.float32(-1.0), np.float32(0.0), dtype=np.float32)
    hi = np.float32(1.0)
    u = uniform_f32(prng_key(seed), n, lo, hi)
    out = (np.float32(np.sqrt(2.0)) * _erfinv_f32(u)).astype(np.float32)
    return out.reshape(shape)


def split(key: tuple[int, int], num: int) -> list[tuple[int, int]]:
    """``jax.random.split(key, num)`` (threefry, non-partitionable layout): counters 0 .. 2 num - 1 are hashed
    in two halves and the 2 num output words, read in order, are the num new keys."""
    flat = random_bits32(key, 2 * num)
    return [(int(flat[2 * i]), int(flat[2 * i + 1])) for i in range(num)]


def randint_scalar(key: tuple[int, int], minval: int, maxval: int) -> int:
    """``jax.random.randint(key, (), minval, maxval)`` for int32 (jax/_src/random.py ``_randint``): two 32-bit
    draws from the two halves of ``split(key)``, combined as ((hi % span) * (2^32 % span) + lo % span) % span
    in uint32 arithmetic.  Restated from the published JAX 0.4.x source; no reference test pins it."""
    k1, k2 = split(key, 2)
    hi = int(random_bits32(k1, 1)[0])
    lo = int(random_bits32(k2, 1)[0])
    span = (maxval - minval) & 0xFFFFFFFF
    if maxval <= minval:
        span = 1
    mult = (1 << 16) % span
    mult = ((mult * mult) & 0xFFFFFFFF) % span  # == 2**32 % span while span <= 2**16 (uint32 product wraps beyond)
    off = (((hi % span) * mult) & 0xFFFFFFFF) + (lo % span)
    off = (off & 0xFFFFFFFF) % span
    return minval + off
