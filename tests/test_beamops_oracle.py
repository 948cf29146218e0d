"""CPU: the restatements in oracle/beamops.py against the relational checks the reference's
own tests make for these operations (no golden vector exists for any of them):

* transformed_to: the transformed beam has the requested moments
  (tests/test_particle_beam.py:44-80);
* uniform_3d_ellipsoid: every particle inside the ellipsoid, momenta Gaussian with the
  requested sigmas (tests/test_particle_beam.py:108-146);
* Aperture: survivors are the masked particles in incoming order, rectangular vs elliptical
  (lynx/accelerator/aperture.py:80-97; tests/test_compare_ocelot.py:150-225 needs Ocelot);
* the plot_twiss loop: element by element equals the merged track.
"""

import numpy as np
import torch

from oracle import beamops as BO
from oracle import generate as G
from oracle import lattice as O


def _planes(n=20_000, seed=0):
    rng = np.random.default_rng(seed)
    sig = np.array([2e-4, 3e-5, 1.5e-4, 2e-5, 1e-5, 1e-3])
    mu = np.array([1e-4, -2e-6, 5e-5, 1e-6, 0.0, 0.0])
    return (mu[:, None] + sig[:, None] * rng.standard_normal((6, n))).astype(np.float32)


def test_transformed_to_reaches_requested_moments():
    x = _planes()
    mu, sd = BO.beam_stats(x)
    mu[4:] = 0.0  # particle_beam.py:675-676: s and p are not re-centred
    new_mu = np.array([1e-5, 1e-7, 2e-5, 2e-7, 0.0, 0.0], dtype=np.float32)
    new_sd = np.array([1.75e-7, 2e-7, 1.75e-7, 2e-7, 1e-6, 1e-6], dtype=np.float32)
    y = BO.transformed_to(x, mu, sd, new_mu, new_sd)
    assert y.dtype == np.float32 and y.shape == x.shape
    mu2, sd2 = BO.beam_stats(y)
    np.testing.assert_allclose(sd2, new_sd, rtol=1e-4)
    np.testing.assert_allclose(mu2[:4], new_mu[:4], rtol=1e-3, atol=1e-4 * new_sd[:4].max())
    # identity request: values move by at most a few ulp
    z = BO.transformed_to(x, mu, sd, mu, sd)
    assert np.max(np.abs(z - x) / (np.abs(x) + sd[:, None])) < 1e-6


def test_aperture_mask_and_order():
    x = _planes(5000, seed=3)
    for shape in ("rectangular", "elliptical"):
        kept, lost, _, _ = BO.aperture_track(x, 2e-4, 1.5e-4, shape)
        assert kept.shape[1] + lost.shape[1] == 5000
        assert 0 < kept.shape[1] < 5000
        if shape == "rectangular":
            assert np.all(np.abs(kept[0]) < 2e-4) and np.all(np.abs(kept[2]) < 1.5e-4)
        else:
            assert np.all((kept[0] / 2e-4) ** 2 + (kept[2] / 1.5e-4) ** 2 <= 1 + 1e-6)
        # order-preserving: kept is a subsequence of the incoming particles
        m = BO.aperture_mask(x[0], x[2], 2e-4, 1.5e-4, shape)
        np.testing.assert_array_equal(kept, x[:, m])
        np.testing.assert_array_equal(lost, x[:, ~m])
    # the ellipse is inside its bounding rectangle
    kr = BO.aperture_track(x, 2e-4, 1.5e-4, "rectangular")[0].shape[1]
    ke = BO.aperture_track(x, 2e-4, 1.5e-4, "elliptical")[0].shape[1]
    assert ke < kr
    # defaults (aperture.py:47-52): infinite half sizes keep everything; NaN is lost
    x[0, 7] = np.nan
    kept, lost, _, _ = BO.aperture_track(x, np.inf, np.inf, "rectangular")
    assert lost.shape[1] == 1 and np.isnan(lost[0, 0])
    q = np.arange(5000, dtype=np.float32)
    _, _, kq, lq = BO.aperture_track(x, np.inf, np.inf, "rectangular", charges=q)
    assert lq.tolist() == [7.0] and kq.shape == (4999,)


def test_ellipsoid_inside_uniform_and_sliceable():
    radii = np.array([[1e-3, 1e-4, 1e-5], [2e-3, 2e-4, 2e-5]])
    sigma = np.array([[2e-7, 3e-7, 1e-6], [1e-7, 2e-7, 2e-6]])
    n, seed = 60_000, 0xABCDEF12345
    x = BO.uniform_3d_ellipsoid(n, radii, sigma, seed)
    assert x.shape == (2, 6, n)
    for b in range(2):
        pos = x[b, 0::2]
        t = (pos[0] / radii[b, 0]) ** 2 + (pos[1] / radii[b, 1]) ** 2 + (pos[2] / radii[b, 2]) ** 2
        assert np.all(t < 1 + 1e-6)  # tests/test_particle_beam.py:139-141
        assert t.max() > 0.99  # ... and the ellipsoid is filled up to its surface
        # uniform in an ellipsoid: var = r^2 / 5 per axis, mean 0
        np.testing.assert_allclose(pos.var(axis=1), radii[b] ** 2 / 5, rtol=3e-2)
        assert np.all(np.abs(pos.mean(axis=1)) < 5 * radii[b] / np.sqrt(5 * n))
        # momenta: the from_parameters stream scaled by sigma (test_particle_beam.py:142-144)
        np.testing.assert_allclose(x[b, 1::2].std(axis=1), sigma[b], rtol=2e-2)
        z = G.standard_normals(n, b, seed)
        np.testing.assert_array_equal(x[b, 1], sigma[b, 0] * z[1])
    # a shard is a slice of the same stream
    part = BO.ellipsoid_positions(500, 1, radii[1], seed, first=1000)
    np.testing.assert_array_equal(part, x[1, 0::2, 1000:1500].astype(np.float32))
    # positions are float32 values
    assert np.array_equal(x[0, 0], x[0, 0].astype(np.float32).astype(np.float64))


def test_unit_interval_is_symmetric_and_open():
    w = np.array([0, 0x1FF, 0x200, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF], dtype=np.uint32)
    t = BO._unit(w).astype(np.float64)
    assert t[0] == -(1 - 2.0**-23) and t[-1] == 1 - 2.0**-23
    assert t[1] == t[0] and t[2] == t[0] + 2.0**-22
    assert t[3] == -(2.0**-23) and t[4] == 2.0**-23


def test_track_along_matches_merged_track():
    from helpers import oracle_elements
    from lynx_b200 import fixtures as F

    d = torch.float64
    spec = [s for s in F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)]
    els = oracle_elements(spec, d)
    beam = O.ParameterBeamO.from_parameters((1e-4, 0, -2e-4, 0),
                                            (*F.ASTRA_SIGMA[:4], 1e-6, 1e-6), F.ASTRA_ENERGY, dtype=d)
    beams = BO.track_along(els, beam)
    assert len(beams) == len(els)
    end = O.track(els, beam)
    torch.testing.assert_close(beams[-1].mu_, end.mu_, rtol=1e-10, atol=1e-16)
    torch.testing.assert_close(beams[-1].cov_, end.cov_, rtol=1e-9, atol=1e-24)
    # beta is positive everywhere and changes across the first quadrupole
    bx = torch.stack([b.beta("x") for b in beams])
    assert bool((bx > 0).all()) and not torch.allclose(bx[0], bx[-1])
