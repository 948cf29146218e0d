"""CPU restatement of the device-side Gaussian beam generator (csrc/generate.cu, K9).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

The reference samples ``MultivariateNormal(mean, cov).sample((N,))`` per batch row from the
array framework's global generator (lynx/particles/particle_beam.py:143-170); a different
generator cannot reproduce those draws, so what is pinned here is

* the counter-based generator itself -- Philox4x32-10 (Salmon et al., SC'11; Random123
  v1.14 ``philox.h``; the same generator cuRAND and torch-CUDA use), checked against the
  published known-answer vectors of Random123's ``kat_vectors`` in
  ``tests/test_generate.py``;
* the layout of the stream (which counter word feeds which particle / coordinate) and the
  Box-Muller + affine map arithmetic, restated here in float64.
"""

import numpy as np

_M0 = np.uint64(0xD2511F53)
_M1 = np.uint64(0xCD9E8D57)
_W0 = 0x9E3779B9
_W1 = 0xBB67AE85
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(counter: np.ndarray, key) -> np.ndarray:
    """Philox4x32 with 10 rounds.  ``counter`` (..., 4) uint32, ``key`` (k0, k1)."""
    c = np.asarray(counter, dtype=np.uint64).copy()
    k0, k1 = int(key[0]) & 0xFFFFFFFF, int(key[1]) & 0xFFFFFFFF
    for _ in range(10):
        p0 = _M0 * c[..., 0]
        p1 = _M1 * c[..., 2]
        n0 = (p1 >> np.uint64(32)) ^ c[..., 1] ^ np.uint64(k0)
        n1 = p1 & _MASK
        n2 = (p0 >> np.uint64(32)) ^ c[..., 3] ^ np.uint64(k1)
        n3 = p0 & _MASK
        c = np.stack([n0, n1, n2, n3], axis=-1)
        k0 = (k0 + _W0) & 0xFFFFFFFF
        k1 = (k1 + _W1) & 0xFFFFFFFF
    return c.astype(np.uint32)


def _u01(x: np.ndarray) -> np.ndarray:
    """Top 24 bits -> (0, 1): ((x >> 8) + 0.5) * 2^-24 (exact in fp32, never 0 or 1)."""
    return ((x >> np.uint32(8)).astype(np.float64) + 0.5) * 2.0**-24


def standard_normals(n: int, b: int, seed: int, first: int = 0) -> np.ndarray:
    """(6, n) float64 standard normals of batch row ``b``: the stream of generate.cu --
    counter (group_lo, group_hi, b, j), key (seed_lo, seed_hi); words (0,1) -> particles
    0,1 of the group (r cos, r sin), words (2,3) -> particles 2,3."""
    assert first % 4 == 0
    groups = (n + 3) // 4
    g = np.arange(groups, dtype=np.uint64) + np.uint64(first // 4)
    z = np.empty((6, groups * 4))
    for j in range(6):
        ctr = np.stack([g & _MASK, g >> np.uint64(32), np.full_like(g, b), np.full_like(g, j)], -1)
        w = philox4x32_10(ctr, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))
        for pair in range(2):
            r = np.sqrt(-2.0 * np.log(_u01(w[:, 2 * pair])))
            th = 2.0 * np.pi * _u01(w[:, 2 * pair + 1])
            z[j, 2 * pair::4] = r * np.cos(th)
            z[j, 2 * pair + 1::4] = r * np.sin(th)
    return z[:, :n]


def beam_generate(mean: np.ndarray, sqrt_cov: np.ndarray, n: int, seed: int, first: int = 0):
    """``mean`` (B, 6), ``sqrt_cov`` (B, 6, 6) -> (B, 6, n) float64: mean + A z."""
    mean = np.asarray(mean, dtype=np.float64)
    A = np.asarray(sqrt_cov, dtype=np.float64)
    out = np.empty((mean.shape[0], 6, n))
    for b in range(mean.shape[0]):
        out[b] = mean[b][:, None] + A[b] @ standard_normals(n, b, seed, first)
    return out


def covariance_from_parameters(sigma, cor):
    """particle_beam.py:148-160: sigma (6,), cor (cor_x, cor_y, cor_s) -> cov (6, 6)."""
    cov = np.diag(np.asarray(sigma, dtype=np.float64) ** 2)
    for n_, (i, j) in enumerate(((0, 1), (2, 3), (4, 5))):
        cov[i, j] = cov[j, i] = cor[n_]
    return cov
