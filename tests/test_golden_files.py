"""Committed golden vectors tests/golden/{ares_ea,cavity,beam_ops}*.npz, generated from the fp64
oracle by tests/golden/make_golden.py.  (The vectors computed by the reference itself are
tests/golden/reference_*.npz, see tests/test_reference_golden.py.)

CPU: the oracle still reproduces them (guards the checker against drift).
GPU: the CUDA path, through the C ABI, against the fixed files (no oracle involved)."""

import os

import numpy as np
import pytest
import torch

from lynx_b200 import fixtures as F

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
QUADS = ("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")


def _load(name):
    return np.load(os.path.join(G, name))


def test_oracle_reproduces_golden_particles():
    from helpers import aos_from_soa, oracle_elements
    from oracle import lattice as O

    g = _load("ares_ea_particles.npz")
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    beam = O.ParticleBeamO(aos_from_soa(g["in_soa"], torch.float64),
                           torch.tensor([F.ASTRA_ENERGY], dtype=torch.float64))
    out = O.track(oracle_elements(spec, torch.float64), beam)
    np.testing.assert_allclose(out.particles[0, :, :6].T.numpy(), g["out_planes"], rtol=1e-12,
                               atol=1e-20)
    g4 = _load("cavity_dipole_line_particles.npz")
    beam = O.ParticleBeamO(aos_from_soa(g4["in_soa"], torch.float64),
                           torch.tensor([F.CFG4_ENERGY], dtype=torch.float64))
    out = O.track(oracle_elements(F.cavity_dipole_line(False), torch.float64), beam)
    np.testing.assert_allclose(out.particles[0, :, :6].T.numpy(), g4["out_planes"], rtol=1e-10,
                               atol=1e-20)


def test_oracle_reproduces_golden_beam_ops():
    from oracle import beamops as BO

    g = _load("beam_ops.npz")
    x = g["in_soa"]
    mu, sd = BO.beam_stats(x)
    mu[4:] = 0.0
    np.testing.assert_array_equal(mu, g["old_mu"])
    np.testing.assert_array_equal(sd, g["old_sigma"])
    y = BO.transformed_to(x, g["old_mu"], g["old_sigma"], g["new_mu"], g["new_sigma"])
    np.testing.assert_array_equal(y, g["transformed"])
    for shape, key in (("rectangular", "mask_rect"), ("elliptical", "mask_ellipse")):
        m = BO.aperture_mask(x[0], x[2], g["x_max"], g["y_max"], shape)
        np.testing.assert_array_equal(m, g[key])
    assert 0 < g["mask_ellipse"].sum() < g["mask_rect"].sum() < x.shape[1]
    ell = BO.uniform_3d_ellipsoid(1000, g["radii"], g["sigma3"], int(g["seed"]))
    np.testing.assert_array_equal(ell[:, 0::2].astype(np.float32), g["ellipsoid_pos"])
    np.testing.assert_allclose(ell[:, 1::2], g["ellipsoid_mom"], rtol=1e-12, atol=0)


def test_golden_files_are_small_and_complete():
    names = sorted(f for f in os.listdir(G) if f.endswith(".npz"))
    assert names == ["ares_ea_gradients.npz", "ares_ea_parameter_beam.npz",
                     "ares_ea_particles.npz", "beam_ops.npz", "cavity_dipole_line_particles.npz",
                     # computed by the reference itself (tests/golden/make_reference_golden.py)
                     "reference_ares_ea.npz", "reference_batched_particles.npz",
                     "reference_beam_ops.npz", "reference_bpm_line.npz",
                     "reference_cavity_line.npz", "reference_converters.npz", "reference_element_maps.npz",
                     "reference_gradients.npz",
                     "reference_parameter_beam.npz", "reference_quirks.npz", "reference_zoo.npz"]
    assert all(os.path.getsize(os.path.join(G, n)) < 300_000 for n in names)


def _rel(got, want, rtol, what):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    scale = np.sqrt(np.mean(want**2, axis=-1, keepdims=True)) + 1e-300
    err = np.abs(got - want) / scale
    assert err.max() <= rtol, f"{what}: worst err/rms {err.max():.3e} > {rtol}"


@pytest.mark.gpu
def test_gpu_particles_and_image_vs_golden():
    import lynx_b200 as lx

    g = _load("ares_ea_particles.npz")
    n = g["in_soa"].shape[1]
    beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(g["in_soa"])).unsqueeze(0),
                           torch.tensor([F.ASTRA_ENERGY]))
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    seg = F.build_segment(spec)
    out = seg.track(beam)
    # 2e-5 of the plane rms: fp32 kernel vs the fp64 golden (fp32 input rounding included)
    _rel(out.soa[0, :, :n].cpu().numpy(), g["out_planes"], 2e-5, "particles")
    np.testing.assert_allclose(out.sigma_x.item(), g["sigma"][0], rtol=2e-5)
    np.testing.assert_allclose(out.sigma_y.item(), g["sigma"][2], rtol=2e-5)
    tm = seg.transfer_map(torch.tensor([F.ASTRA_ENERGY]).cuda())[0].cpu().numpy()
    np.testing.assert_allclose(tm, g["transfer_map"], rtol=2e-5, atol=1e-9)

    small = {"resolution": (64, 48), "pixel_size": (4e-5, 4e-5), "binning": 1}
    seg_s = F.build_segment(F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3,
                                      screen_active=True, screen=small))
    assert seg_s.track(beam) is lx.Beam.empty
    img = seg_s.AREABSCR1.reading[0].cpu().numpy()
    want = g["image"]
    assert img.shape == want.shape and img.sum() == want.sum()
    # bins exact except particles within 1 ulp-ish of an edge (fp32 kernel coordinates vs
    # fp32-rounded fp64 oracle coordinates): at most a handful of unit moves
    assert np.abs(img - want).sum() <= 8

    g4 = _load("cavity_dipole_line_particles.npz")
    beam4 = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(g4["in_soa"])).unsqueeze(0),
                            torch.tensor([F.CFG4_ENERGY]))
    out4 = F.build_segment(F.cavity_dipole_line(False)).track(beam4)
    _rel(out4.soa[0, :, :g4["in_soa"].shape[1]].cpu().numpy(), g4["out_planes"], 5e-5,
         "cavity line particles")
    np.testing.assert_allclose(out4.energy.cpu().numpy(), g4["energy_out"], rtol=1e-6)


@pytest.mark.gpu
def test_gpu_parameter_beam_and_gradients_vs_golden():
    import lynx_b200 as lx

    g = _load("ares_ea_parameter_beam.npz")
    seg = F.build_segment(F.ares_ea())
    for i, q in enumerate(QUADS):
        getattr(seg, q).k1 = torch.from_numpy(g["k1"][:, i])
    seg.AREAMCVM1.angle = torch.from_numpy(g["angles"][:, 0])
    seg.AREAMCHM1.angle = torch.from_numpy(g["angles"][:, 1])
    pb = lx.ParameterBeam(torch.from_numpy(g["mu_in"]).float().unsqueeze(0),
                          torch.from_numpy(g["cov_in"]).float().unsqueeze(0),
                          torch.tensor([F.ASTRA_ENERGY]))
    out = seg.track(pb)
    np.testing.assert_allclose(out._mu.cpu().numpy()[:, :4], g["mu_out"][:, :4], rtol=3e-5,
                               atol=1e-9)
    d = np.sqrt(np.abs(np.einsum("bii->bi", g["cov_out"])))
    scale = d[:, :, None] * d[:, None, :] + 1e-300
    assert (np.abs(out._cov.cpu().numpy() - g["cov_out"]) / scale)[:, :6, :6].max() < 3e-5

    gg = _load("ares_ea_gradients.npz")
    seg = F.build_segment(F.ares_ea())
    k1 = torch.tensor(gg["k1"], requires_grad=True)
    ang = torch.tensor(gg["angles"], requires_grad=True)
    for i, q in enumerate(QUADS):
        getattr(seg, q).k1 = k1[:, i]
    seg.AREAMCVM1.angle, seg.AREAMCHM1.angle = ang[:, 0], ang[:, 1]
    beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(gg["in_soa"])).unsqueeze(0),
                           torch.tensor([F.ASTRA_ENERGY]))
    o = seg.track(beam)
    loss = (o.mu_x**2 + o.sigma_x**2 + o.mu_y**2 + o.sigma_y**2).sum() / 4
    loss.backward()
    np.testing.assert_allclose(loss.item(), gg["loss"], rtol=3e-5)
    for got, want, nm in ((k1.grad, gg["g_k1"], "g_k1"), (ang.grad, gg["g_angles"], "g_angles")):
        err = np.abs(got.numpy() - want).max() / np.abs(want).max()
        assert err < 3e-5, f"{nm}: {err:.3e}"


@pytest.mark.gpu
def test_gpu_beam_ops_vs_golden():
    """transformed_to, Aperture.track and uniform_3d_ellipsoid through the public API
    against the committed byte-level vectors (no oracle involved)."""
    import lynx_b200 as lx
    from lynx_b200 import _lib as L

    g = _load("beam_ops.npz")
    x = g["in_soa"]
    n = x.shape[1]
    charges = torch.from_numpy(g["charges"]).unsqueeze(0)
    beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(x)).unsqueeze(0),
                           torch.tensor([F.ASTRA_ENERGY]), particle_charges=charges)
    assert torch.equal(beam.soa[0, :, :n].cpu(), torch.from_numpy(x))
    # the beam's own moments are the fixture's (double sums, rounded to fp32)
    got_mu = np.array([float(getattr(beam, f"mu_{c}")[0]) for c in ("x", "xp", "y", "yp")])
    np.testing.assert_allclose(got_mu, g["old_mu"][:4], rtol=1e-6, atol=1e-12)
    t = lambda v: torch.tensor([float(v)])  # noqa: E731
    nm, ns = g["new_mu"], g["new_sigma"]
    tb = beam.transformed_to(mu_x=t(nm[0]), mu_xp=t(nm[1]), mu_y=t(nm[2]), mu_yp=t(nm[3]),
                             sigma_x=t(ns[0]), sigma_xp=t(ns[1]), sigma_y=t(ns[2]),
                             sigma_yp=t(ns[3]), sigma_s=t(ns[4]), sigma_p=t(ns[5]))
    got = tb.soa[0, :, :n].cpu().numpy()
    # the kernel itself is byte-exact for given moments (tests/test_gpu_beamops.py); here the
    # old moments come from the device's fused sums (fp32 partial sums shifted by the centre,
    # DESIGN.md K5), which agree with the fixture's float64 statistics to ~1e-7: allow
    # 1e-6 sigma_new + 2 ulp of the value, per plane
    tol = 1e-6 * ns[:, None] + 2.4e-7 * np.abs(g["transformed"])
    assert np.all(np.abs(got - g["transformed"]) <= tol)
    for shape, key in (("rectangular", "mask_rect"), ("elliptical", "mask_ellipse")):
        ap = lx.Aperture(x_max=torch.tensor(float(g["x_max"])), y_max=torch.tensor(float(g["y_max"])),
                         shape=shape)
        out = ap.track(beam)
        m = g[key]
        assert out.num_particles == int(m.sum())
        np.testing.assert_array_equal(out.soa[0, :, : out.num_particles].cpu().numpy(), x[:, m])
        np.testing.assert_array_equal(out.particle_charges[0].cpu().numpy(), g["charges"][m])
        np.testing.assert_array_equal(ap.lost_particles[:, :6].T.cpu().numpy(), x[:, ~m])
    r, s3 = torch.from_numpy(g["radii"]), torch.from_numpy(g["sigma3"])
    eb = lx.ParticleBeam.uniform_3d_ellipsoid(
        num_particles=1000, radius_x=r[:, 0], radius_y=r[:, 1], radius_s=r[:, 2],
        sigma_xp=s3[:, 0], sigma_yp=s3[:, 1], sigma_p=s3[:, 2], seed=int(g["seed"]))
    np.testing.assert_array_equal(eb.soa[:, 0::2, :1000].cpu().numpy(), g["ellipsoid_pos"])
    err = np.abs(eb.soa[:, 1::2, :1000].cpu().numpy() - g["ellipsoid_mom"])
    assert np.all(err <= 3e-5 * g["sigma3"][:, :, None])
    assert L.lib().lynx_launch_count() > 0
