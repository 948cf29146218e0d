"""ASTRA distribution reader (lynx_b200/astra.py) -- CPU only.

The reference's sample file (tests/resources/ACHIP_EA1_2021.1351.001) is not in its checkout,
so the known answers here are closed forms: a particle that is ``z`` ahead of the reference
plane is moved back along its own straight line, ``x' = x - z px/pz``, ``y' = y - z py/pz``,
``s = -z |p| / (beta pz)``; ``delta = (gamma/gamma_ref - 1)/beta_ref``."""

import math

import numpy as np
import torch

import lynx_b200 as lx
from lynx_b200.astra import ELECTRON_MASS_EV, read_astra_beam

PZ_REF = 1.0e8  # eV/c
ROWS = [
    # x      y      z      px     py     pz      t    q[nC]  idx status
    (0.0,    0.0,   5.0,   0.0,   0.0,   PZ_REF, 0.0, -1e-3, 1, 5),   # reference particle
    (1e-4,  -2e-4,  0.0,   0.0,   0.0,   1.0e5,  0.0, -1e-3, 1, 5),   # on the plane, faster
    (3e-4,   1e-4,  1e-3,  2.0e4, -1.0e4, 0.0,   0.0, -2e-3, 1, 5),   # ahead, with a slope
    (9e-4,   9e-4, -2e-3,  0.0,   0.0,  -2.0e5,  0.0, -1e-3, 1, 3),   # behind, slower
    (5e-3,   5e-3,  0.0,   0.0,   0.0,   0.0,    0.0, -1e-3, 1, -1),  # lost: dropped
]


def _write(tmp_path):
    path = tmp_path / "beam.ini"
    np.savetxt(path, np.array(ROWS, dtype=np.float64), fmt="%.12e")
    return str(path)


def _gamma(p):
    return math.sqrt(1.0 + (p / ELECTRON_MASS_EV) ** 2)


def test_reader_known_answers(tmp_path):
    coords, energy, charges = read_astra_beam(_write(tmp_path))
    assert coords.shape == (4, 6) and charges.shape == (4,)
    g_ref = _gamma(PZ_REF)
    b_ref = math.sqrt(1.0 - 1.0 / g_ref ** 2)
    assert math.isclose(energy, g_ref * ELECTRON_MASS_EV, rel_tol=1e-14)
    np.testing.assert_allclose(charges, [1e-12, 1e-12, 2e-12, 1e-12], rtol=1e-12)
    # reference particle: the origin of every coordinate
    np.testing.assert_allclose(coords[0], 0.0, atol=1e-18)
    # on the plane: position unchanged, no slope, delta from the momentum offset
    pz = PZ_REF + 1.0e5
    want_delta = (_gamma(pz) / g_ref - 1.0) / b_ref
    np.testing.assert_allclose(coords[1], [1e-4, 0.0, -2e-4, 0.0, 0.0, want_delta], rtol=1e-12, atol=1e-20)
    # ahead of the plane by z with transverse momentum: straight line back to the plane
    z, px, py, pz = 1e-3, 2.0e4, -1.0e4, PZ_REF
    p = math.sqrt(px * px + py * py + pz * pz)
    beta = math.sqrt(1.0 - 1.0 / _gamma(p) ** 2)
    want = [3e-4 - z * px / pz, px / PZ_REF, 1e-4 - z * py / pz, py / PZ_REF,
            -z * p / (beta * pz), (_gamma(p) / g_ref - 1.0) / b_ref]
    np.testing.assert_allclose(coords[2], want, rtol=1e-9, atol=1e-18)
    # behind the plane: s > 0, slower: delta < 0
    assert coords[3, 4] > 0 and coords[3, 5] < 0
    np.testing.assert_allclose(coords[3, [0, 2]], [9e-4, 9e-4], rtol=1e-12)


def test_parameter_beam_from_astra(tmp_path):
    """parameter_beam.py:255-276: mean, unbiased covariance, summed charge."""
    path = _write(tmp_path)
    coords, energy, charges = read_astra_beam(path)
    beam = lx.ParameterBeam.from_astra(path, device="cpu")
    assert tuple(beam._mu.shape) == (1, 7) and tuple(beam._cov.shape) == (1, 7, 7)
    np.testing.assert_allclose(beam._mu[0, :6].numpy(), coords.mean(axis=0), rtol=1e-6, atol=1e-12)
    assert float(beam._mu[0, 6]) == 1.0
    np.testing.assert_allclose(beam._cov[0, :6, :6].numpy(), np.cov(coords.T), rtol=1e-6, atol=1e-20)
    assert bool((beam._cov[0, 6] == 0).all()) and bool((beam._cov[0, :, 6] == 0).all())
    np.testing.assert_allclose(beam.sigma_x.numpy(), [coords[:, 0].std(ddof=1)], rtol=1e-6)
    np.testing.assert_allclose(beam.energy.numpy(), [energy], rtol=1e-7)
    np.testing.assert_allclose(torch.as_tensor(beam.total_charge).numpy(), [5e-12], rtol=1e-6)
