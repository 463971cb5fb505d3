"""CPU check of the arithmetic claim behind csrc/sr_math.cuh::lerp_rcp (IEEE double on the host behaves like the
device's __drcp_rn / __dmul_rn):  for 0 <= n < 2^31, 1 <= d < 2^31,
    floor(n / d) == trunc(fl(double(n) * R)),   R = fl(fl(1/d) * (1 + 2^-40)).
This replaces the integer division of Renderer.swift:492 on the GPU, so it has to hold for every input."""
import numpy as np


def quot(n, d):
    R = (1.0 / d.astype(np.float64)) * (1.0 + 2.0 ** -40)
    return np.trunc(n.astype(np.float64) * R).astype(np.int64)


def check(n, d):
    n = np.asarray(n, np.int64)
    d = np.asarray(d, np.int64)
    ok = (n >= 0) & (n < 2 ** 31) & (d >= 1) & (d < 2 ** 31)
    n, d = n[ok], d[ok]
    assert np.array_equal(quot(n, d), n // d)


def test_exhaustive_small():
    d = np.arange(1, 2049, dtype=np.int64)
    n = np.arange(0, 1 << 15, dtype=np.int64)
    check(np.repeat(n, d.size), np.tile(d, n.size))


def test_random_and_adversarial():
    rng = np.random.default_rng(7)
    for _ in range(20):
        d = rng.integers(1, 2 ** 31, size=400_000)
        k = rng.integers(0, 2 ** 31, size=400_000) // d
        for off in (-1, 0, 1):                       # exact multiples and their neighbours
            check(k * d + off, d)
        check(rng.integers(0, 2 ** 31, size=400_000), d)
        small = rng.integers(1, 5000, size=400_000)    # screen-sized divisors, largest numerators
        check(2 ** 31 - 1 - rng.integers(0, 10_000, size=400_000), small)
        kk = (2 ** 31 - 1) // small
        check(kk * small, small)
        check(kk * small - 1, small)
