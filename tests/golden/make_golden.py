"""Generate the committed golden vectors in tests/golden/*.npz FROM THE ORACLE.

    python tests/golden/make_golden.py

These vectors come from the CPU restatement in ``oracle/`` (fp64): they pin the oracle against
regressions and let the GPU tests compare against fixed files.  They are NOT outputs of the
reference -- those are tests/golden/reference_*.npz, written by
tests/golden/make_reference_golden.py (the reference run on torch through
tests/golden/reference_shim.py), which the fp32 oracle reproduces bit for bit
(tests/test_reference_golden.py).  The known answers from the reference's own tests /
notebooks (C2-C5) are checked in tests/test_oracle_golden.py.
"""

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

from helpers import aos_from_soa, oracle_elements  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from oracle import lattice as O  # noqa: E402
from oracle import screen_bin_edges, screen_reading_particles, segment_map  # noqa: E402

D = torch.float64
QUADS = ("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")
SMALL_SCREEN = {"resolution": (64, 48), "pixel_size": (4e-5, 4e-5), "binning": 1}


def ares_particles():
    n = 512
    soa = F.gaussian_beam_soa(n, seed=2001)
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    beam = O.ParticleBeamO(aos_from_soa(soa, D), torch.tensor([F.ASTRA_ENERGY], dtype=D))
    out = O.track(oracle_elements(spec, D), beam)
    tm = segment_map(oracle_elements(spec, D), torch.tensor([F.ASTRA_ENERGY], dtype=D))
    ex, ey = screen_bin_edges(SMALL_SCREEN["resolution"], SMALL_SCREEN["pixel_size"])
    planes32 = out.particles[0, :, :6].T.float()
    img = screen_reading_particles(planes32[0:1], planes32[2:3], ex, ey)
    np.savez(os.path.join(HERE, "ares_ea_particles.npz"), in_soa=soa,
             out_planes=out.particles[0, :, :6].T.numpy(), transfer_map=tm[0].numpy(),
             mu=np.array([out.mu(i).item() for i in range(6)]),
             sigma=np.array([out.sigma(i).item() for i in range(6)]),
             sigma_xxp=out.sigma_xxp.item(), sigma_yyp=out.sigma_yyp.item(),
             edges_x=ex.numpy(), edges_y=ey.numpy(), image=img[0].numpy())


def cavity_line_particles():
    n = 512
    soa = F.cfg4_beam_soa(n, seed=2004)
    spec = F.cavity_dipole_line(screen_active=False)
    beam = O.ParticleBeamO(aos_from_soa(soa, D), torch.tensor([F.CFG4_ENERGY], dtype=D))
    out = O.track(oracle_elements(spec, D), beam)
    np.savez(os.path.join(HERE, "cavity_dipole_line_particles.npz"), in_soa=soa,
             out_planes=out.particles[0, :, :6].T.numpy(), energy_out=out.energy.numpy())


def parameter_beam_settings():
    B = 32
    k1, ang = F.random_settings(B, seed=2003)
    ov = {(q, "k1"): k1[:, i] for i, q in enumerate(QUADS)}
    ov[("AREAMCVM1", "angle")] = ang[:, 0]
    ov[("AREAMCHM1", "angle")] = ang[:, 1]
    ob = O.ParameterBeamO.from_parameters(F.ASTRA_MU[:4], F.ASTRA_SIGMA, F.ASTRA_ENERGY, dtype=D)
    ob = O.ParameterBeamO(ob.mu_.expand(B, 7).clone(), ob.cov_.expand(B, 7, 7).clone(),
                          ob.energy.expand(B).clone())
    out = O.track(oracle_elements(F.ares_ea(), D, (B,), ov), ob)
    np.savez(os.path.join(HERE, "ares_ea_parameter_beam.npz"), k1=k1, angles=ang,
             mu_in=ob.mu_[0].numpy(), cov_in=ob.cov_[0].numpy(), mu_out=out.mu_.numpy(),
             cov_out=out.cov_.numpy())


def gradients():
    B, n = 8, 2048
    k1, ang = F.random_settings(B, seed=2005, k1_range=5.0)
    soa = F.gaussian_beam_soa(n, seed=2005)
    k1o = torch.tensor(k1, dtype=D, requires_grad=True)
    ango = torch.tensor(ang, dtype=D, requires_grad=True)
    ov = {(q, "k1"): k1o[:, i] for i, q in enumerate(QUADS)}
    ov[("AREAMCVM1", "angle")] = ango[:, 0]
    ov[("AREAMCHM1", "angle")] = ango[:, 1]
    beam = O.ParticleBeamO(aos_from_soa(soa, D).expand(B, -1, -1),
                           torch.full((B,), F.ASTRA_ENERGY, dtype=D))
    out = O.track(oracle_elements(F.ares_ea(), D, (B,), ov), beam)
    loss = (out.mu(0) ** 2 + out.sigma(0) ** 2 + out.mu(2) ** 2 + out.sigma(2) ** 2).sum() / 4
    loss.backward()
    np.savez(os.path.join(HERE, "ares_ea_gradients.npz"), k1=k1, angles=ang, in_soa=soa,
             loss=loss.item(), g_k1=k1o.grad.numpy(), g_angles=ango.grad.numpy())


def beam_ops():
    """Byte-level vectors of the beam-level operations (oracle/beamops.py, numpy fp32):
    transformed_to, Aperture.track (both shapes) and the ellipsoid sampler."""
    from oracle import beamops as BO

    n = 1500
    x = F.gaussian_beam_soa(n, seed=2006).astype(np.float32)
    mu, sd = BO.beam_stats(x)
    mu[4:] = 0.0
    new_mu = np.array([1e-5, 1e-7, 2e-5, 2e-7, 0.0, 0.0], dtype=np.float32)
    new_sd = np.array([1.75e-7, 2e-7, 1.75e-7, 2e-7, 1e-6, 1e-6], dtype=np.float32)
    y = BO.transformed_to(x, mu, sd, new_mu, new_sd)
    q = np.linspace(1e-15, 2e-15, n).astype(np.float32)
    xm, ym = np.float32(2e-4), np.float32(1.5e-4)
    mr = BO.aperture_mask(x[0], x[2], xm, ym, "rectangular")
    me = BO.aperture_mask(x[0], x[2], xm, ym, "elliptical")
    radii = np.array([[1e-3, 1e-4, 1e-5], [2e-3, 2e-4, 2e-5]], dtype=np.float32)
    sigma = np.array([[2e-7, 3e-7, 1e-6], [1e-7, 2e-7, 2e-6]], dtype=np.float32)
    seed = 0x2006_0000_BEA7
    ell = BO.uniform_3d_ellipsoid(1000, radii, sigma, seed)
    np.savez(os.path.join(HERE, "beam_ops.npz"), in_soa=x, old_mu=mu, old_sigma=sd, new_mu=new_mu,
             new_sigma=new_sd, transformed=y, charges=q, x_max=xm, y_max=ym, mask_rect=mr,
             mask_ellipse=me, radii=radii, sigma3=sigma, seed=np.uint64(seed),
             ellipsoid_pos=ell[:, 0::2].astype(np.float32), ellipsoid_mom=ell[:, 1::2])


if __name__ == "__main__":
    beam_ops()
    ares_particles()
    cavity_line_particles()
    parameter_beam_settings()
    gradients()
    for f in sorted(os.listdir(HERE)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(HERE, f)))
