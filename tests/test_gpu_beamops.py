"""GPU parity of the beam-level kernels through the C ABI:
K10 ``lynx_beam_transform`` (transformed_to), K11 ``lynx_aperture_compact`` (Aperture.track),
K9b ``lynx_beam_generate_ellipsoid`` (uniform_3d_ellipsoid), K6-along
``lynx_parambeam_along`` (the plot_twiss loop) -- against oracle/beamops.py.

Bars: transform, aperture and the ellipsoid's spatial planes are byte-exact (explicit fp32
round-to-nearest operations on both sides); the ellipsoid's momentum planes carry K9's
tolerance (3e-5 sigma: SFU log / sqrt / sincos vs fp64); K6-along is within 2e-5 of the fp64
element-by-element oracle.
"""

import numpy as np
import pytest
import torch

from helpers import oracle_elements

pytestmark = pytest.mark.gpu

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import _lib as L  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from oracle import beamops as BO  # noqa: E402
from oracle import generate as G  # noqa: E402
from oracle import lattice as O  # noqa: E402


def _planes(n, seed=0, rows=1):
    rng = np.random.default_rng(seed)
    sig = np.array([2e-4, 3e-5, 1.5e-4, 2e-5, 1e-5, 1e-3])
    mu = np.array([1e-4, -2e-6, 5e-5, 1e-6, 0.0, 0.0])
    x = mu[None, :, None] + sig[None, :, None] * rng.standard_normal((rows, 6, n))
    return x.astype(np.float32)


def _to_device(x):
    rows, _, n = x.shape
    ld = (n + 3) & ~3
    soa = torch.zeros((rows, 6, ld), device="cuda")
    soa[:, :, :n] = torch.from_numpy(x).cuda()
    return soa, ld


# ------------------------------------------------------------------ K10 transformed_to
@pytest.mark.parametrize("n", [3, 4099, 100_003])
def test_transform_is_byte_exact(n):
    x = _planes(n, seed=n, rows=2)
    soa, ld = _to_device(x)
    tab = np.empty((2, 24), dtype=np.float32)
    new_mu = np.array([1e-5, 1e-7, 2e-5, 2e-7, 0.0, 0.0], dtype=np.float32)
    new_sd = np.array([1.75e-7, 2e-7, 1.75e-7, 2e-7, 1e-6, 1e-6], dtype=np.float32)
    want = np.empty_like(x)
    for r in range(2):
        mu, sd = BO.beam_stats(x[r])
        mu[4:] = 0.0
        tab[r] = np.concatenate([mu, sd, new_sd * (r + 1), new_mu * (1 - 2 * r)])
        want[r] = BO.transformed_to(x[r], mu, sd, new_mu * (1 - 2 * r), new_sd * (r + 1))
    out = torch.full_like(soa, float("nan"))
    t = torch.from_numpy(tab).cuda()
    L.check(L.lib().lynx_beam_transform(soa.data_ptr(), 6 * ld, ld, 2, n, t.data_ptr(),
                                        out.data_ptr(), ld, L.current_stream()))
    got = out.cpu().numpy()
    np.testing.assert_array_equal(got[:, :, :n], want)
    assert np.all(got[:, :, n:] == 0.0)
    # one beam, two targets (in_row_stride = 0)
    tab[1, :12] = tab[0, :12]
    t = torch.from_numpy(tab).cuda()
    L.check(L.lib().lynx_beam_transform(soa.data_ptr(), 0, ld, 2, n, t.data_ptr(),
                                        out.data_ptr(), ld, L.current_stream()))
    want1 = BO.transformed_to(x[0], tab[0, :6], tab[0, 6:12], tab[1, 18:], tab[1, 12:18])
    np.testing.assert_array_equal(out[1, :, :n].cpu().numpy(), want1)


def test_transformed_to_api_reaches_requested_moments():
    """tests/test_particle_beam.py:44-80 of the reference."""
    t = lambda v: torch.tensor([v])  # noqa: E731
    original = lx.ParticleBeam.from_parameters(seed=4)
    tb = original.transformed_to(
        mu_x=t(1e-5), mu_xp=t(1e-7), mu_y=t(2e-5), mu_yp=t(2e-7), sigma_x=t(1.75e-7),
        sigma_xp=t(2e-7), sigma_y=t(1.75e-7), sigma_yp=t(2e-7), sigma_s=t(0.000001),
        sigma_p=t(0.000001), energy=t(1e7), total_charge=t(1e-9))
    assert isinstance(tb, lx.ParticleBeam) and tb.num_particles == original.num_particles
    for name, want in (("mu_x", 1e-5), ("mu_xp", 1e-7), ("mu_y", 2e-5), ("mu_yp", 2e-7),
                       ("sigma_x", 1.75e-7), ("sigma_xp", 2e-7), ("sigma_y", 1.75e-7),
                       ("sigma_yp", 2e-7), ("sigma_s", 1e-6), ("sigma_p", 1e-6),
                       ("energy", 1e7), ("total_charge", 1e-9)):
        assert np.isclose(getattr(tb, name).cpu().numpy(), want, rtol=1e-4), name
    # the particle pass itself against the oracle, from the beam's own fp32 moments
    x = original.soa[0, :, : original.num_particles].cpu().numpy()
    om = np.array([float(getattr(original, f"mu_{c}")[0]) for c in ("x", "xp", "y", "yp")] + [0, 0],
                  dtype=np.float32)
    os_ = np.array([float(getattr(original, f"sigma_{c}")[0])
                    for c in ("x", "xp", "y", "yp", "s", "p")], dtype=np.float32)
    want = BO.transformed_to(x, om, os_, [1e-5, 1e-7, 2e-5, 2e-7, 0, 0],
                             [1.75e-7, 2e-7, 1.75e-7, 2e-7, 1e-6, 1e-6])
    np.testing.assert_array_equal(tb.soa[0, :, : tb.num_particles].cpu().numpy(), want)


# ------------------------------------------------------------------ K11 Aperture
def _compact(x, xm, ym, shape, charges=None):
    rows, _, n = x.shape
    soa, ld = _to_device(x)
    kept = torch.full_like(soa, float("nan"))
    lost = torch.full_like(soa, float("nan"))
    q = qk = ql = None
    if charges is not None:
        q = torch.from_numpy(charges).cuda().contiguous()
        qk = torch.full((rows, ld), float("nan"), device="cuda")
        ql = torch.full((rows, ld), float("nan"), device="cuda")
    lim = torch.tensor([xm, ym], dtype=torch.float32, device="cuda")
    counts = torch.zeros(rows, dtype=torch.int64, device="cuda")
    ws = torch.empty(max(L.lib().lynx_aperture_workspace_bytes(rows, n), 16), dtype=torch.uint8,
                     device="cuda")
    L.check(L.lib().lynx_aperture_compact(
        soa.data_ptr(), rows, n, ld, L.ptr(q), n if q is not None else 0, lim.data_ptr(),
        lim.data_ptr() + 4, 0, 0 if shape == "rectangular" else 1, kept.data_ptr(),
        lost.data_ptr(), ld, L.ptr(qk), L.ptr(ql), counts.data_ptr(), ws.data_ptr(),
        L.current_stream()))
    torch.cuda.synchronize()
    to = lambda t: None if t is None else t.cpu().numpy()  # noqa: E731
    return counts.cpu().numpy(), kept.cpu().numpy(), lost.cpu().numpy(), to(qk), to(ql)


@pytest.mark.parametrize("shape", ["rectangular", "elliptical"])
@pytest.mark.parametrize("n", [1, 5, 2048, 2049, 100_003])
def test_aperture_compaction_is_byte_exact(shape, n):
    x = _planes(n, seed=7 + n, rows=2)
    x[1] *= 0.5  # the second row keeps more
    if n > 100:
        x[0, 0, 11] = np.nan  # NaN coordinates are lost (every comparison fails)
    charges = np.arange(2 * n, dtype=np.float32).reshape(2, n)
    counts, kept, lost, qk, ql = _compact(x, 2e-4, 1.5e-4, shape, charges)
    for r in range(2):
        k, l_, kq, lq = BO.aperture_track(x[r], 2e-4, 1.5e-4, shape, charges[r])
        m = k.shape[1]
        assert counts[r] == m
        np.testing.assert_array_equal(kept[r, :, :m], k)
        np.testing.assert_array_equal(lost[r, :, : n - m], l_)
        np.testing.assert_array_equal(qk[r, :m], kq)
        np.testing.assert_array_equal(ql[r, : n - m], lq)


def test_aperture_all_or_nothing():
    x = _planes(10_000, seed=1)
    counts, kept, _, _, _ = _compact(x, np.inf, np.inf, "rectangular")
    assert counts[0] == 10_000
    np.testing.assert_array_equal(kept[0, :, :10_000], x[0])
    counts, _, lost, _, _ = _compact(x, 0.0, 0.0, "rectangular")
    assert counts[0] == 0
    np.testing.assert_array_equal(lost[0, :, :10_000], x[0])


def test_aperture_element_in_a_segment():
    """Drift -> Aperture -> Drift through Segment.track (the structure of
    tests/test_compare_ocelot.py:150-225) against the oracle: drift, mask, drift."""
    n = 50_001
    soa = F.gaussian_beam_soa(n, seed=12)
    charges = torch.linspace(1e-15, 2e-15, n).unsqueeze(0)
    for shape in ("rectangular", "elliptical"):
        beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0),
                               torch.tensor([F.ASTRA_ENERGY]), particle_charges=charges)
        ap = lx.Aperture(x_max=torch.tensor(2e-4), y_max=torch.tensor(2e-4), shape=shape,
                         name="ap")
        seg = lx.Segment([lx.Drift(torch.tensor([0.5])), ap, lx.Drift(torch.tensor([0.1]))])
        out = seg.track(beam)
        mid = lx.Drift(torch.tensor([0.5])).track(beam)
        xm = mid.soa[0, :, :n].cpu().numpy()
        k, l_, kq, lq = BO.aperture_track(xm, 2e-4, 2e-4, shape, charges[0].numpy())
        m = k.shape[1]
        assert 0 < m < n and out.num_particles == m
        assert ap.lost_particles.shape == (n - m, 7)
        np.testing.assert_array_equal(ap.lost_particles[:, :6].T.cpu().numpy(), l_)
        assert bool((ap.lost_particles[:, 6] == 1).all())
        np.testing.assert_array_equal(ap.lost_particle_charges.cpu().numpy(), lq)
        np.testing.assert_array_equal(out.particle_charges[0].cpu().numpy(), kq)
        # survivors through the last drift == drifting the oracle's survivors
        kb = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(k)).unsqueeze(0),
                             torch.tensor([F.ASTRA_ENERGY]))
        want = lx.Drift(torch.tensor([0.1])).track(kb)
        assert torch.equal(out.soa[0, :, :m], want.soa[0, :, :m])
        np.testing.assert_allclose(out.sigma_x.cpu().numpy(), want.sigma_x.cpu().numpy(), rtol=1e-6)
        # an inactive aperture is transparent; a ParameterBeam passes through
        ap.is_active = False
        assert seg.track(beam).num_particles == n
        ap.is_active = True
        pb = lx.ParameterBeam.from_parameters(energy=torch.tensor([F.ASTRA_ENERGY]))
        assert seg.track(pb)._mu.shape == (1, 7)
    # nothing survives -> Beam.empty (aperture.py:98-108)
    ap0 = lx.Aperture(x_max=torch.tensor(0.0), y_max=torch.tensor(0.0))
    assert lx.Segment([ap0, lx.Drift(torch.tensor([0.1]))]).track(beam) is lx.Beam.empty
    assert ap0.lost_particles.shape == (n, 7)


# ------------------------------------------------------------------ K9b uniform_3d_ellipsoid
def test_ellipsoid_matches_oracle_stream():
    radii = np.array([[1e-3, 1e-4, 1e-5], [2e-3, 2e-4, 2e-5]], dtype=np.float32)
    sigma = np.array([[2e-7, 3e-7, 1e-6], [1e-7, 2e-7, 2e-6]], dtype=np.float32)
    n, seed = 20_003, 0xFEED_0000_BEEF
    ld = (n + 3) & ~3

    def gen(n_, first=0):
        ld_ = (n_ + 3) & ~3
        soa = torch.full((2, 6, ld_), float("nan"), device="cuda")
        r, s = torch.from_numpy(radii).cuda(), torch.from_numpy(sigma).cuda()
        L.check(L.lib().lynx_beam_generate_ellipsoid(r.data_ptr(), s.data_ptr(), 2, n_, first,
                                                     seed, soa.data_ptr(), ld_,
                                                     L.current_stream()))
        return soa.cpu().numpy()

    got = gen(n)
    want = BO.uniform_3d_ellipsoid(n, radii, sigma, seed)
    np.testing.assert_array_equal(got[:, 0::2, :n], want[:, 0::2].astype(np.float32))
    err = np.abs(got[:, 1::2, :n] - want[:, 1::2])
    assert np.all(err <= 3e-5 * sigma[:, :, None]), (err / sigma[:, :, None]).max()
    assert np.all(got[:, :, n:ld] == 0.0)
    part = gen(1000, first=4096)
    np.testing.assert_array_equal(part[:, :, :1000], got[:, :, 4096:5096])


def test_uniform_3d_ellipsoid_api():
    """tests/test_particle_beam.py:108-146 of the reference."""
    radius_x, radius_y = torch.tensor([1e-3, 2e-3]), torch.tensor([1e-4, 2e-4])
    radius_s = torch.tensor([1e-5, 2e-5])
    sigma_xp, sigma_yp = torch.tensor([2e-7, 1e-7]), torch.tensor([3e-7, 2e-7])
    sigma_p = torch.tensor([0.000001, 0.000002])
    energy, total_charge = torch.tensor([1e7, 2e7]), torch.tensor([1e-9, 3e-9])
    n = 1_000_000
    beam = lx.ParticleBeam.uniform_3d_ellipsoid(
        num_particles=n, radius_x=radius_x, radius_y=radius_y, radius_s=radius_s,
        sigma_xp=sigma_xp, sigma_yp=sigma_yp, sigma_p=sigma_p, energy=energy,
        total_charge=total_charge, seed=31)
    assert beam.num_particles == n
    assert bool((beam.xs.abs().T.cpu() <= radius_x).all())
    assert bool((beam.ys.abs().T.cpu() <= radius_y).all())
    assert bool((beam.ss.abs().T.cpu() <= radius_s).all())
    t = (beam.xs.cpu() / radius_x[:, None]) ** 2 + (beam.ys.cpu() / radius_y[:, None]) ** 2 \
        + (beam.ss.cpu() / radius_s[:, None]) ** 2
    assert bool((t < 1 + 1e-6).all())
    assert torch.allclose(beam.sigma_xp.cpu(), sigma_xp, rtol=5e-3)
    assert torch.allclose(beam.sigma_yp.cpu(), sigma_yp, rtol=5e-3)
    assert torch.allclose(beam.sigma_p.cpu(), sigma_p, rtol=5e-3)
    assert torch.allclose(beam.sigma_x.cpu(), radius_x / 5**0.5, rtol=5e-3)  # waterbag
    assert torch.allclose(beam.energy.cpu(), energy)
    assert torch.allclose(beam.total_charge.cpu(), total_charge, rtol=1e-4)
    # the momentum planes are those of from_parameters under the same seed
    g = lx.ParticleBeam.from_parameters(
        num_particles=n, mu_xp=torch.zeros(2), mu_yp=torch.zeros(2), sigma_xp=sigma_xp,
        sigma_yp=sigma_yp, sigma_p=sigma_p, energy=energy, total_charge=total_charge, seed=31)
    for k in (1, 3, 5):
        assert torch.equal(beam.soa[:, k], g.soa[:, k])


# ------------------------------------------------------------------ K6-along / plot_twiss
def test_twiss_along_matches_element_by_element_oracle():
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    seg = F.build_segment(spec)
    t = lambda v: torch.tensor([v])  # noqa: E731
    pb = lx.ParameterBeam.from_parameters(
        mu_x=t(1e-4), mu_y=t(-2e-4), sigma_x=t(F.ASTRA_SIGMA[0]), sigma_xp=t(F.ASTRA_SIGMA[1]),
        sigma_y=t(F.ASTRA_SIGMA[2]), sigma_yp=t(F.ASTRA_SIGMA[3]), energy=t(F.ASTRA_ENERGY))
    s_pos, beams = seg.twiss_along(pb)
    d = torch.float64
    els = [e for e in oracle_elements(spec, d) if "length" in e.p and float(e.p["length"][0]) != 0]
    ob = O.ParameterBeamO.from_parameters((1e-4, 0, -2e-4, 0), (*F.ASTRA_SIGMA[:4], 1e-6, 1e-6),
                                          F.ASTRA_ENERGY, dtype=d)
    want = BO.track_along(els, ob)
    assert len(beams) == len(want) + 1 == len(s_pos)
    np.testing.assert_allclose(s_pos[-1], sum(float(e.p["length"][0]) for e in els), rtol=1e-6)
    for got, w in zip(beams[1:], want):
        sd = np.sqrt(np.abs(np.diag(w.cov_[0].numpy())))
        np.testing.assert_allclose(got._mu[0].cpu().numpy(), w.mu_[0].numpy(), rtol=2e-5,
                                   atol=2e-5 * sd.max())
        gc, wc = got._cov[0].cpu().numpy(), w.cov_[0].numpy()
        assert (np.abs(gc - wc) <= 3e-5 * (np.outer(sd, sd) + np.abs(wc))).all()
        # beta = sigma^2 / sqrt(sigma_x^2 sigma_xp^2 - sigma_xxp^2): the emittance is a
        # difference of nearly equal products once the beam is focused, so fp32 covariances
        # that are right to 3e-5 give a beta that is right to ~1e-3 (beam.py:262-279)
        np.testing.assert_allclose(got.beta_x.cpu().numpy(), w.beta("x").numpy(), rtol=3e-3)
        np.testing.assert_allclose(got.beta_y.cpu().numpy(), w.beta("y").numpy(), rtol=3e-3)
    # the last snapshot is the merged track's outgoing beam (to fp32 rounding)
    end = seg.track(pb)
    ec = end._cov[0].cpu().numpy()
    esd = np.sqrt(np.abs(np.diag(ec)))
    assert (np.abs(beams[-1]._cov[0].cpu().numpy() - ec) <= 1e-4 * np.outer(esd, esd)).all()
    s2, bx, by = seg.plot_twiss(pb)
    assert len(bx) == len(by) == len(s2) and all(v > 0 for v in bx + by)
    # a ParticleBeam is stepped element by element and agrees with the ParameterBeam betas
    beam = lx.ParticleBeam.from_parameters(
        num_particles=200_000, mu_x=t(1e-4), mu_y=t(-2e-4), sigma_x=t(F.ASTRA_SIGMA[0]),
        sigma_xp=t(F.ASTRA_SIGMA[1]), sigma_y=t(F.ASTRA_SIGMA[2]), sigma_yp=t(F.ASTRA_SIGMA[3]),
        energy=t(F.ASTRA_ENERGY), seed=5)
    _, bxp, _ = seg.plot_twiss(beam)
    np.testing.assert_allclose(bxp, bx, rtol=5e-2)


def test_split_matches_unsplit_track():
    """drift.py:71-78, quadrupole.py:99-110, horizontal_corrector.py:82-90: tracking the
    slices equals tracking the element."""
    q = lx.Quadrupole(torch.tensor([0.122]), k1=torch.tensor([7.5]))
    dr = lx.Drift(torch.tensor([0.43]))
    hc = lx.HorizontalCorrector(torch.tensor([0.02]), angle=torch.tensor([1e-3]))
    seg = lx.Segment([dr, q, hc])
    parts = seg.split(torch.tensor(0.05))
    assert len(parts) == 9 + 3 + 1
    np.testing.assert_allclose(sum(float(p.length) for p in parts), 0.43 + 0.122 + 0.02, rtol=1e-6)
    e = torch.tensor([F.ASTRA_ENERGY])
    a = seg.transfer_map(e).cpu().numpy()
    b = lx.Segment(parts).transfer_map(e).cpu().numpy()
    np.testing.assert_allclose(b, a, rtol=2e-5, atol=2e-7)


def test_aperture_element_on_a_batched_beam():
    """Three beam rows through one Aperture with a limit per row: the reference's mask indexing
    (aperture.py:89-97) flattens the batch -- the survivors of all rows, in row order, form ONE
    beam; the lost particles likewise."""
    n, R = 20_003, 3
    rows = np.stack([F.gaussian_beam_soa(n, seed=30 + r) for r in range(R)])
    aos = torch.from_numpy(np.stack([F.soa_to_aos7(rows[r]) for r in range(R)]))
    charges = torch.arange(R * n, dtype=torch.float32).reshape(R, n)
    beam = lx.ParticleBeam(aos, torch.full((R,), F.ASTRA_ENERGY), particle_charges=charges)
    xmax = torch.tensor([1e-4, 2e-4, 3e-4])
    for shape in ("rectangular", "elliptical"):
        ap = lx.Aperture(x_max=xmax, y_max=torch.tensor(2e-4), shape=shape, name="ap")
        out = ap.track(beam)
        kept, lost, kq, lq = [], [], [], []
        for r in range(R):
            k, l_, a, b = BO.aperture_track(rows[r], float(xmax[r]), 2e-4, shape, charges[r].numpy())
            kept.append(k)
            lost.append(l_)
            kq.append(a)
            lq.append(b)
        kept, lost = np.concatenate(kept, axis=1), np.concatenate(lost, axis=1)
        m = kept.shape[1]
        assert 0 < m < R * n and out.num_particles == m and out.soa.shape[0] == 6
        np.testing.assert_array_equal(out.soa[:, :m].cpu().numpy(), kept)
        np.testing.assert_array_equal(out.particle_charges.cpu().numpy(), np.concatenate(kq))
        assert ap.lost_particles.shape == (R * n - m, 7)
        np.testing.assert_array_equal(ap.lost_particles[:, :6].T.cpu().numpy(), lost)
        np.testing.assert_array_equal(ap.lost_particle_charges.cpu().numpy(), np.concatenate(lq))
