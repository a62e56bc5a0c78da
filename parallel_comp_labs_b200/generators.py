"""Synthetic inputs in the reference's body format.

* :func:`uniform_bodies` -- the distributions of the reference's (commented-out)
  ``generate_data`` (lab_4_N_body_MPI/src/N_body.cpp:15-35): m ~ U[1e6,1e8],
  r ~ U[-100,100]^3, v ~ U[-10,10]^3, but drawn from a portable counter-based
  splitmix64 stream so every machine produces the same file.
* :func:`plummer_bodies` -- Plummer sphere (Aarseth-Henon-Wielen sampling), scale a=1,
  total mass M = 1/G (so G*M = 1), equal masses, radius cut 22.8a, centred on the
  centre of mass and momentum.

Both return (n,7) float64 arrays ``m x y z vx vy vz`` = the memory image of Body<double>
(lab_4 .../struct_Body.hpp:5-10).
"""
from __future__ import annotations

import numpy as np

_GOLD = np.uint64(0x9E3779B97F4A7C15)


def splitmix64(seed: int, counters: np.ndarray) -> np.ndarray:
    """splitmix64 output for state = seed + (counter+1)*golden (vectorised, wraps mod 2^64)."""
    with np.errstate(over="ignore"):
        z = np.uint64(seed) + (counters.astype(np.uint64) + np.uint64(1)) * _GOLD
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return z


def _u01(seed: int, counters: np.ndarray) -> np.ndarray:
    """Uniform doubles in [0,1) with 53 random bits."""
    return (splitmix64(seed, counters) >> np.uint64(11)).astype(np.float64) * (1.0 / (1 << 53))


def uniform_bodies(n: int, seed: int = 0x5EED0001) -> np.ndarray:
    c = np.arange(n, dtype=np.uint64) * np.uint64(7)
    out = np.empty((n, 7))
    out[:, 0] = 1e6 + (1e8 - 1e6) * _u01(seed, c)
    for k in range(3):
        out[:, 1 + k] = -1e2 + 2e2 * _u01(seed, c + np.uint64(1 + k))
        out[:, 4 + k] = -1e1 + 2e1 * _u01(seed, c + np.uint64(4 + k))
    return out


def plummer_bodies(n: int, seed: int = 0x5EED0004, G: float = 6.67e-11,
                   r_cut: float = 22.8) -> np.ndarray:
    M = 1.0 / G
    out = np.empty((n, 7))
    out[:, 0] = M / n
    idx = np.arange(n, dtype=np.uint64)
    # radii (stream 0), rejecting r > r_cut
    r = np.empty(n)
    todo = np.arange(n)
    attempt = 0
    while todo.size:
        x1 = _u01(seed, idx[todo] * np.uint64(64) + np.uint64(attempt))
        x1 = np.clip(x1, 1e-12, 1.0 - 1e-12)
        rr = 1.0 / np.sqrt(x1 ** (-2.0 / 3.0) - 1.0)
        ok = rr <= r_cut
        r[todo[ok]] = rr[ok]
        todo = todo[~ok]
        attempt += 1
    s2 = seed ^ 0xA5A5A5A5
    x2 = _u01(s2, idx * np.uint64(4))
    x3 = _u01(s2, idx * np.uint64(4) + np.uint64(1))
    z = (1.0 - 2.0 * x2) * r
    rxy = np.sqrt(np.maximum(r * r - z * z, 0.0))
    out[:, 1] = rxy * np.cos(2 * np.pi * x3)
    out[:, 2] = rxy * np.sin(2 * np.pi * x3)
    out[:, 3] = z
    # speed modulus q by von Neumann rejection: g(q) = q^2 (1-q^2)^{7/2}, max < 0.1
    q = np.empty(n)
    todo = np.arange(n)
    s3 = seed ^ 0x3C3C3C3C
    attempt = 0
    while todo.size:
        x4 = _u01(s3, idx[todo] * np.uint64(512) + np.uint64(2 * attempt))
        x5 = _u01(s3, idx[todo] * np.uint64(512) + np.uint64(2 * attempt + 1))
        ok = 0.1 * x5 < x4 * x4 * (1.0 - x4 * x4) ** 3.5
        q[todo[ok]] = x4[ok]
        todo = todo[~ok]
        attempt += 1
    ve = np.sqrt(2.0) * (1.0 + r * r) ** (-0.25)  # escape speed, G*M = a = 1
    v = q * ve
    x6 = _u01(s2, idx * np.uint64(4) + np.uint64(2))
    x7 = _u01(s2, idx * np.uint64(4) + np.uint64(3))
    vz = (1.0 - 2.0 * x6) * v
    vxy = np.sqrt(np.maximum(v * v - vz * vz, 0.0))
    out[:, 4] = vxy * np.cos(2 * np.pi * x7)
    out[:, 5] = vxy * np.sin(2 * np.pi * x7)
    out[:, 6] = vz
    out[:, 1:4] -= out[:, 1:4].mean(axis=0)
    out[:, 4:7] -= out[:, 4:7].mean(axis=0)
    return out
