"""Oracle: beams, element dispatch, Segment tracking, BPM / Screen readings.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Follows:
  lynx/accelerator/element.py:61-94        Element.track (mu' = R mu, cov' = R cov R^T,
                                           particles' = particles R^T)
  lynx/accelerator/segment.py:317-356      is_skippable / transfer_map / track grouping
  lynx/accelerator/custom_transfer_map.py:48-88
  lynx/accelerator/cavity.py:97-246        Cavity._track_beam (non-linear part)
  lynx/accelerator/bpm.py:48-58            BPM.track
  lynx/accelerator/screen.py:86-141,195-211  Screen.track / reading (ParticleBeam)
  lynx/particles/particle_beam.py:720-836  moments
  lynx/particles/parameter_beam.py:371-425 accessors
  lynx/particles/beam.py:241-310           emittance / twiss

Where the reference branches on ``any(...)`` over the WHOLE batch (cavity.py:128,164,
290; dipole.py:119) this oracle branches per batch entry when ``per_entry=True`` (the
default, = what a per-setting kernel does and what the reference does for B == 1) and
over the whole batch when ``per_entry=False`` (the reference as written for B > 1).
"""

import math
from dataclasses import dataclass, field
from typing import Optional

import torch
from scipy import constants

from . import maps as M
from .maps import ELECTRON_MASS_EV


@dataclass
class El:
    """One lattice element: ``kind`` + named parameter tensors of the batch shape."""

    kind: str
    name: str = ""
    p: dict = field(default_factory=dict)
    is_active: bool = False
    reading: Optional[torch.Tensor] = None  # BPM
    read_beam: object = None  # Screen


class ParticleBeamO:
    """particles (...,N,7) with constant-1 seventh column; particle_beam.py:24-45."""

    def __init__(self, particles, energy, particle_charges=None):
        assert particles.shape[-2] > 0 and particles.shape[-1] == 7
        self.particles = particles
        self.energy = energy
        self.particle_charges = particle_charges

    def col(self, i):
        return self.particles[..., i]

    def mu(self, i):  # particle_beam.py:736-738 et al.
        return self.col(i).mean(dim=-1)

    def sigma(self, i):  # :740-742 torch.std default = unbiased (N-1)
        return self.col(i).std(dim=-1)

    def cross(self, i, j):  # :822-836 biased mean of centred products
        a = self.col(i) - self.mu(i).unsqueeze(-1)
        b = self.col(j) - self.mu(j).unsqueeze(-1)
        return (a * b).mean(dim=-1)

    @property
    def sigma_xxp(self):
        return self.cross(0, 1)

    @property
    def sigma_yyp(self):
        return self.cross(2, 3)

    def emittance(self, plane):  # beam.py:262-270 / 290-297
        i = 0 if plane == "x" else 2
        v = self.sigma(i) ** 2 * self.sigma(i + 1) ** 2 - self.cross(i, i + 1) ** 2
        return torch.sqrt(torch.clamp_min(v, torch.finfo(v.dtype).tiny))


class ParameterBeamO:
    """mu (...,7), cov (...,7,7); parameter_beam.py:25-45."""

    def __init__(self, mu, cov, energy, total_charge=None):
        self.mu_, self.cov_, self.energy, self.total_charge = mu, cov, energy, total_charge

    @classmethod
    def from_parameters(cls, mu4, sig6, energy, cor=(0.0, 0.0, 0.0), dtype=torch.float32):
        """parameter_beam.py:47-146 for one batch entry (batch shape (1,))."""
        mu = torch.tensor([[*mu4, 0.0, 0.0, 1.0]], dtype=dtype)
        cov = torch.zeros(1, 7, 7, dtype=dtype)
        for i in range(6):
            cov[0, i, i] = torch.tensor(sig6[i], dtype=dtype) ** 2
        for n, (i, j) in enumerate(((0, 1), (2, 3), (4, 5))):
            cov[0, i, j] = cov[0, j, i] = cor[n]
        return cls(mu, cov, torch.tensor([energy], dtype=dtype))

    def mu(self, i):
        return self.mu_[..., i]

    def sigma(self, i):  # :375-417
        return torch.sqrt(torch.clamp_min(self.cov_[..., i, i], 1e-20))

    def emittance(self, plane):
        i = 0 if plane == "x" else 2
        v = self.sigma(i) ** 2 * self.sigma(i + 1) ** 2 - self.cov_[..., i, i + 1] ** 2
        return torch.sqrt(torch.clamp_min(v, torch.finfo(v.dtype).tiny))

    def beta(self, plane):  # beam.py:277-279
        i = 0 if plane == "x" else 2
        return self.sigma(i) ** 2 / self.emittance(plane)

    def alpha(self, plane):  # beam.py:282-284
        i = 0 if plane == "x" else 2
        return -self.cov_[..., i, i + 1] / self.emittance(plane)


EMPTY = "I'm an empty beam!"  # beam.py:12


def is_skippable(el: El) -> bool:
    """Per-class is_skippable: cavity.py:68-70, bpm.py:40-42, screen.py:82-84."""
    if el.kind == "cavity":
        return not bool((el.p["voltage"] != 0).any())
    if el.kind in ("bpm", "screen"):
        return not el.is_active
    return True


def element_map(el: El, energy: torch.Tensor, cavity_v0_as_drift=False) -> torch.Tensor:
    p, k = el.p, el.kind
    if k == "drift":
        return M.drift_map(p["length"], energy)
    if k == "quadrupole":
        return M.quadrupole_map(p["length"], p["k1"], p["misalignment"], p["tilt"], energy)
    if k in ("hcor", "vcor"):
        return M.corrector_map(p["length"], p["angle"], energy, k == "hcor")
    if k in ("dipole", "rbend"):
        e1, e2 = p["e1"], p["e2"]
        if k == "rbend":  # rbend.py:79-80
            e1, e2 = e1 + p["angle"] / 2, e2 + p["angle"] / 2
        return M.dipole_map(
            p["length"], p["angle"], e1, e2, p["tilt"],
            p["fringe_integral"], p["fringe_integral_exit"], p["gap"], energy,
        )
    if k == "cavity":
        R = M.cavity_rmatrix(p["length"], p["voltage"], p["phase"], p["frequency"], energy)
        if cavity_v0_as_drift:  # deviation from quirk B6 (0*inf NaN), see DESIGN.md
            z = p["voltage"] == 0
            if bool(z.any()):
                R = R.clone()
                R[z] = M.drift_map(p["length"], energy)[z]
        return R
    if k == "solenoid":
        return M.solenoid_map(p["length"], p["k"], p["misalignment"], energy)
    if k == "undulator":
        return M.undulator_map(p["length"], energy)
    if k == "custom":  # custom_transfer_map.py:87-88
        return p["transfer_map"]
    if k in ("marker", "bpm", "screen"):  # marker.py:33-36, bpm.py:44-47, screen.py:120-124
        return torch.eye(7, dtype=energy.dtype).repeat(*energy.shape, 1, 1)
    raise ValueError(k)


def segment_map(elements, energy, **kw) -> torch.Tensor:
    """segment.py:329-338 -- ordered product tm = R_e @ tm."""
    tm = torch.eye(7, dtype=energy.dtype).repeat(*energy.shape, 1, 1)
    for el in elements:
        tm = torch.matmul(element_map(el, energy, **kw), tm)
    return tm


def _apply(tm, beam):
    """element.py:71-92."""
    if isinstance(beam, ParameterBeamO):
        mu = torch.matmul(tm, beam.mu_.unsqueeze(-1)).squeeze(-1)
        cov = torch.matmul(tm, torch.matmul(beam.cov_, tm.transpose(-2, -1)))
        return ParameterBeamO(mu, cov, beam.energy, beam.total_charge)
    new = torch.matmul(beam.particles, tm.transpose(-2, -1))
    return ParticleBeamO(new, beam.energy, beam.particle_charges)


def _sel(cond_all, cond_entry, a, b, per_entry):
    """value = a where the branch is taken else b (whole batch or per entry)."""
    if per_entry:
        return torch.where(cond_entry, a, b)
    return a if bool(cond_all) else b


def cavity_track(el: El, beam, per_entry=True, cavity_v0_as_drift=False):
    """cavity.py:97-246, both beam kinds."""
    p = el.p
    L, V, f = p["length"], p["voltage"], p["frequency"]
    E = beam.energy
    one, zero = torch.ones_like(L), torch.zeros_like(L)

    mask = E != 0
    g0 = torch.where(mask, E / ELECTRON_MASS_EV, torch.full_like(L, 1e10))
    igamma2 = torch.where(mask, 1 / g0**2, zero)
    beta0 = torch.where(mask, torch.sqrt(1 - igamma2), one)
    phi = torch.deg2rad(p["phase"])

    # linear part (cavity.py:118); the per-entry variant of the :290 branch
    if per_entry:
        tm = _cavity_rmatrix_per_entry(L, V, p["phase"], f, E)
    else:
        tm = M.cavity_rmatrix(L, V, p["phase"], f, E)
    if cavity_v0_as_drift and bool((V == 0).any()):  # deviation from quirk B6
        tm = tm.clone()
        tm[V == 0] = M.drift_map(L, E)[V == 0]
    lin = _apply(tm, beam)
    dE = V * torch.cos(phi)

    T566 = 1.5 * L * igamma2 / beta0**3
    T556, T555 = zero.clone(), zero.clone()
    E1 = E + dE
    on = E1 > 0  # :128
    k = 2 * math.pi * f / constants.speed_of_light
    g1 = E1 / ELECTRON_MASS_EV
    beta1 = torch.sqrt(1 - 1 / g1**2)
    dgamma = V / ELECTRON_MASS_EV
    acc = dE > 0  # :164
    T566a = L * (beta0**3 * g0**3 - beta1**3 * g1**3) / (
        2 * beta0 * beta1**3 * g0 * (g0 - g1) * g1**3
    )
    T556a = (
        beta0 * k * L * dgamma * g0 * (beta1**3 * g1**3 + beta0 * (g0 - g1**3))
        * torch.sin(phi) / (beta1**3 * g1**3 * (g0 - g1) ** 2)
    )
    T555a = (
        beta0**2 * k**2 * L * dgamma / 2.0
        * (
            dgamma
            * (2 * g0 * g1**3 * (beta0 * beta1**3 - 1) + g0**2 + 3 * g1**2 - 2)
            / (beta1**3 * g1**3 * (g0 - g1) ** 3)
            * torch.sin(phi) ** 2
            - (g1 * g0 * (beta1 * beta0 - 1) + 1)
            / (beta1 * g1 * (g0 - g1) ** 2)
            * torch.cos(phi)
        )
    )
    T566 = _sel(acc.any(), acc, T566a, T566, per_entry)
    T556 = _sel(acc.any(), acc, T556a, T556, per_entry)
    T555 = _sel(acc.any(), acc, T555a, T555, per_entry)

    if isinstance(beam, ParameterBeamO):
        mu, cov = lin.mu_.clone(), lin.cov_.clone()
        mi, ci = beam.mu_, beam.cov_
        mu5 = mi[..., 5] * E * beta0 / (E1 * beta1) + V * beta0 / (E1 * beta1) * (
            torch.cos(-mi[..., 4] * beta0 * k + phi) - torch.cos(phi)
        )
        mu4 = mu[..., 4] + (
            T566 * mi[..., 5] ** 2 + T556 * mi[..., 4] * mi[..., 5] + T555 * mi[..., 4] ** 2
        )
        c44 = (  # overwritten, not added (quirk B5, cavity.py:208-217)
            T566 * ci[..., 5, 5] ** 2
            + T556 * ci[..., 4, 5] * ci[..., 5, 5]
            + T555 * ci[..., 4, 4] ** 2
        )
        if per_entry:
            mu[..., 5] = torch.where(on, mu5, mu[..., 5])
            mu[..., 4] = torch.where(on, mu4, mu[..., 4])
            cov[..., 5, 5] = torch.where(on, ci[..., 5, 5], cov[..., 5, 5])
            cov[..., 4, 4] = torch.where(on, c44, cov[..., 4, 4])
            cov[..., 4, 5] = torch.where(on, c44, cov[..., 4, 5])
            cov[..., 5, 4] = cov[..., 4, 5]
        elif bool(on.any()):
            mu[..., 5], mu[..., 4] = mu5, mu4
            cov[..., 5, 5] = ci[..., 5, 5]
            cov[..., 4, 4] = c44
            cov[..., 4, 5] = c44
            cov[..., 5, 4] = c44
        return ParameterBeamO(mu, cov, E1, beam.total_charge)

    out = lin.particles.clone()
    pi = beam.particles
    u = lambda t: t.unsqueeze(-1)  # noqa: E731
    p5 = pi[..., 5] * u(E) * u(beta0) / (u(E1) * u(beta1)) + u(V) * u(beta0) / (
        u(E1) * u(beta1)
    ) * (torch.cos(-1 * pi[..., 4] * u(beta0) * u(k) + u(phi)) - u(torch.cos(phi)))
    p4 = out[..., 4] + (
        u(T566) * pi[..., 5] ** 2 + u(T556) * pi[..., 4] * pi[..., 5] + u(T555) * pi[..., 4] ** 2
    )
    if per_entry:
        out[..., 5] = torch.where(u(on), p5, out[..., 5])
        out[..., 4] = torch.where(u(on), p4, out[..., 4])
    elif bool(on.any()):
        out[..., 5], out[..., 4] = p5, p4
    return ParticleBeamO(out, E1, beam.particle_charges)


def _cavity_rmatrix_per_entry(L, V, phase, f, E):
    """cavity.py:248-325 with the :290 branch decided per batch entry."""
    R = M.cavity_rmatrix(L, V, phase, f, E)
    off = ~((V != 0) & (E != 0))
    if bool(off.any()) and not bool(off.all()):
        # entries where the branch would NOT be taken alone: r56=0, beta0=beta1=1, r55_cor=0
        Ei = E / ELECTRON_MASS_EV
        Ef = (E + V * torch.cos(torch.deg2rad(phase))) / ELECTRON_MASS_EV
        R = R.clone()
        R[..., 4, 4] = torch.where(off, torch.ones_like(L), R[..., 4, 4])
        R[..., 4, 5] = torch.where(off, torch.zeros_like(L), R[..., 4, 5])
        R[..., 5, 5] = torch.where(off, Ei / Ef, R[..., 5, 5])
        k = 2 * math.pi * f / constants.speed_of_light
        r65 = k * torch.sin(torch.deg2rad(phase)) * V / (Ef * ELECTRON_MASS_EV)
        R[..., 5, 4] = torch.where(off, r65, R[..., 5, 4])
    return R


def screen_bin_edges(resolution, pixel_size, binning=1, dtype=torch.float32):
    """screen.py:86-88,105-118: linspace over the UN-binned extent."""
    res = torch.tensor(resolution, dtype=dtype)
    px = torch.tensor(pixel_size, dtype=dtype)
    b = torch.tensor(binning, dtype=dtype)
    eff = res / b
    ex = torch.linspace(-res[0] * px[0] / 2, res[0] * px[0] / 2, int(eff[0]) + 1)
    ey = torch.linspace(-res[1] * px[1] / 2, res[1] * px[1] / 2, int(eff[1]) + 1)
    return ex.to(dtype), ey.to(dtype)


def screen_reading_particles(xs, ys, edges_x, edges_y):
    """screen.py:195-211: per batch row histogramdd then flipud(hist.T) -> (B,H,W)."""
    imgs = []
    for x, y in zip(xs, ys):
        h, _ = torch.histogramdd(torch.stack((x, y)).T.contiguous(), bins=[edges_x, edges_y])
        imgs.append(torch.flipud(h.T))
    return torch.stack(imgs)


def screen_reading_parameter_beam(mu, cov, resolution, pixel_size, binning=1):
    """screen.py:157-194: bivariate-normal pdf on the grid of pixel left edges, meshgrid
    "ij", flipped along x (quirk B8) -> (B, W_eff, H_eff)."""
    from torch.distributions import MultivariateNormal

    dtype = mu.dtype
    res = torch.tensor(resolution, dtype=torch.float32)
    px = torch.tensor(pixel_size, dtype=torch.float32)
    left, right = -res[0] * px[0] / 2, res[0] * px[0] / 2
    bottom, top = -res[1] * px[1] / 2, res[1] * px[1] / 2
    hstep, vstep = px[0] * binning, px[1] * binning
    x, y = torch.meshgrid(torch.arange(left, right, hstep), torch.arange(bottom, top, vstep),
                          indexing="ij")
    pos = torch.dstack((x, y)).to(dtype)
    imgs = []
    for m, c in zip(mu, cov):
        tm = torch.stack([m[0], m[2]])
        tc = torch.stack([torch.stack([c[0, 0], c[0, 2]]), torch.stack([c[2, 0], c[2, 2]])])
        imgs.append(MultivariateNormal(loc=tm, covariance_matrix=tc).log_prob(pos).exp())
    return torch.flip(torch.stack(imgs), dims=[1])


def track(elements, beam, per_entry=True, cavity_v0_as_drift=False, screen_quirk_b1=True):
    """segment.py:340-356: merge runs of skippable elements, step through the rest."""
    run = []

    def flush(b):
        if run and b is not EMPTY:
            tm = segment_map(run, b.energy, cavity_v0_as_drift=cavity_v0_as_drift)
            b = _apply(tm, b)
        run.clear()
        return b

    for el in elements:
        if is_skippable(el):
            run.append(el)
            continue
        beam = flush(beam)
        if el.kind == "cavity":
            if beam is not EMPTY:
                beam = cavity_track(
                    el, beam, per_entry=per_entry, cavity_v0_as_drift=cavity_v0_as_drift
                )
        elif el.kind == "bpm":  # bpm.py:48-58
            if beam is EMPTY:
                el.reading = None
            else:
                el.reading = torch.stack([beam.mu(0), beam.mu(2)])
        elif el.kind == "screen":  # screen.py:126-141
            if beam is not EMPTY:
                mis = el.p["misalignment"]
                if isinstance(beam, ParameterBeamO):
                    mu = beam.mu_.clone()
                    mu[..., 0] -= mis[..., 0]
                    mu[..., 2] -= mis[..., 1]
                    el.read_beam = ParameterBeamO(mu, beam.cov_.clone(), beam.energy)
                else:
                    pt = beam.particles.clone()
                    pt[..., 0] -= mis[..., 0].unsqueeze(-1)
                    # quirk B1: y offset lands on column 1 (xp) as written, :134-135
                    pt[..., 1 if screen_quirk_b1 else 2] -= mis[..., 1].unsqueeze(-1)
                    el.read_beam = ParticleBeamO(pt, beam.energy, beam.particle_charges)
                beam = EMPTY
        else:
            raise ValueError(el.kind)
    return flush(beam)
