"""Beam state (host-side mirror of lynx/particles/*).

``ParticleBeam`` keeps the reference's constructor / property surface
(particle_beam.py:24-45, 720-843) but stores particles as SoA-6 device planes
``(*batch, 6, ld)``; ``.particles`` materialises the reference's AoS ``(*batch, N, 7)``
view on demand.  Moments come from the fused double-precision sums of the tracking
kernel (or one extra streaming pass when a beam was not produced by ``track``).
``ParameterBeam`` mirrors parameter_beam.py:25-45, 371-433.
"""

from typing import Optional

import torch

from . import _lib as L

ELECTRON_MASS_EV = 510998.95069  # lynx/particles/beam.py:7


class Beam:
    """Base class; ``Beam.empty`` is the reference's sentinel (beam.py:12)."""

    empty = "I'm an empty beam!"

    # ---- derived quantities shared by both beam kinds (beam.py:241-310)
    @property
    def relativistic_gamma(self):
        return self.energy / ELECTRON_MASS_EV

    @property
    def relativistic_beta(self):
        g = self.relativistic_gamma
        return torch.where(g.abs() > 0, torch.sqrt(1 - 1 / (g * g)), torch.ones_like(g))

    @property
    def emittance_x(self):
        v = self.sigma_x**2 * self.sigma_xp**2 - self.sigma_xxp**2
        return torch.sqrt(torch.clamp_min(v, torch.finfo(v.dtype).tiny))

    @property
    def emittance_y(self):
        v = self.sigma_y**2 * self.sigma_yp**2 - self.sigma_yyp**2
        return torch.sqrt(torch.clamp_min(v, torch.finfo(v.dtype).tiny))

    @property
    def normalized_emittance_x(self):
        return self.emittance_x * self.relativistic_beta * self.relativistic_gamma

    @property
    def normalized_emittance_y(self):
        return self.emittance_y * self.relativistic_beta * self.relativistic_gamma

    @property
    def beta_x(self):
        return self.sigma_x**2 / self.emittance_x

    @property
    def alpha_x(self):
        return -self.sigma_xxp / self.emittance_x

    @property
    def beta_y(self):
        return self.sigma_y**2 / self.emittance_y

    @property
    def alpha_y(self):
        return -self.sigma_yyp / self.emittance_y

    @property
    def parameters(self) -> dict:
        names = ["mu_x", "mu_xp", "mu_y", "mu_yp", "sigma_x", "sigma_xp", "sigma_y",
                 "sigma_yp", "sigma_s", "sigma_p", "energy"]
        return {n: getattr(self, n) for n in names}


_COORD = {"x": 0, "xp": 1, "y": 2, "yp": 3, "s": 4, "p": 5}


def _tri(i: int, j: int) -> int:
    return i * 6 - i * (i - 1) // 2 + (j - i)


class ParticleBeam(Beam):
    """Beam of individually tracked particles.

    :param particles: ``(*batch, N, 7)`` as in the reference (seventh column 1), on any
        device; converted once to the SoA-6 device layout.
    :param energy: ``(*batch)`` reference energy in eV.
    :param particle_charges: ``(*batch, N)``, passed through untouched.
    """

    def __init__(self, particles, energy, particle_charges=None, device=None,
                 dtype=torch.float32):
        assert dtype == torch.float32, "device path is fp32 (SURVEY.md section 5)"
        assert particles.shape[-2] > 0 and particles.shape[-1] == 7, (
            "Particle vectors must be 7-dimensional."
        )
        device = torch.device(device if device is not None else "cuda")
        aos = particles.to(device=device, dtype=torch.float32).contiguous()
        batch = tuple(aos.shape[:-2])
        N = aos.shape[-2]
        R = 1
        for s in batch:
            R *= s
        ld = (N + 3) & ~3
        soa = torch.empty((*batch, 6, ld), device=device, dtype=torch.float32)
        if ld != N:
            soa[..., N:] = 0
        L.check(L.lib().lynx_aos7_to_soa6(aos.data_ptr(), R, N, soa.data_ptr(), ld,
                                          L.current_stream()), "lynx_aos7_to_soa6")
        self._init_soa(soa, N, torch.as_tensor(energy, dtype=torch.float32, device=device),
                       particle_charges)

    def _init_soa(self, soa, N, energy, particle_charges):
        self._soa = soa
        self._n = int(N)
        self.energy = energy
        self.particle_charges = (
            particle_charges.to(device=soa.device, dtype=torch.float32)
            if particle_charges is not None
            else None
        )
        self._sums = None  # (*batch, 28) float64 raw moment sums

    @classmethod
    def from_soa(cls, soa, num_particles, energy, particle_charges=None, moment_sums=None):
        """Wrap existing SoA-6 planes ``(*batch, 6, ld)`` (no copy)."""
        self = cls.__new__(cls)
        self._init_soa(soa, num_particles, energy, particle_charges)
        self._sums = moment_sums
        return self

    @classmethod
    def from_parameters(cls, num_particles=None, mu_x=None, mu_y=None, mu_xp=None, mu_yp=None,
                        sigma_x=None, sigma_y=None, sigma_xp=None, sigma_yp=None, sigma_s=None,
                        sigma_p=None, cor_x=None, cor_y=None, cor_s=None, energy=None,
                        total_charge=None, device=None, dtype=torch.float32, seed=None,
                        first_particle=0):
        """particle_beam.py:47-170 (same arguments and defaults), sampled ON the device by
        ``lynx_beam_generate`` (csrc/generate.cu): the beam never exists on the host.

        ``seed`` (extra): 64-bit key of the counter-based generator; ``None`` draws one from
        torch's CPU generator, so ``torch.manual_seed`` makes beams reproducible.
        ``first_particle`` (extra): global index of this buffer's first particle, for
        generating one rank's shard of a larger beam (multiple of 4)."""
        assert dtype == torch.float32, "device path is fp32 (SURVEY.md section 5)"
        given = [a for a in (mu_x, mu_xp, mu_y, mu_yp, sigma_x, sigma_xp, sigma_y, sigma_yp,
                             sigma_s, sigma_p, cor_x, cor_y, cor_s, energy, total_charge)
                 if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        num_particles = int(num_particles) if num_particles is not None else 100_000
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float64).cpu() if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float64))
        zero = torch.zeros(shape, dtype=torch.float64)
        mean = torch.stack([d(mu_x, 0.0), d(mu_xp, 0.0), d(mu_y, 0.0), d(mu_yp, 0.0), zero, zero],
                           dim=-1)
        sig = [d(sigma_x, 175e-9), d(sigma_xp, 2e-7), d(sigma_y, 175e-9), d(sigma_yp, 2e-7),
               d(sigma_s, 1e-6), d(sigma_p, 1e-6)]
        cor = [d(cor_x, 0.0), d(cor_y, 0.0), d(cor_s, 0.0)]
        cov = torch.zeros(*shape, 6, 6, dtype=torch.float64)
        for i in range(6):
            cov[..., i, i] = sig[i] ** 2
        for n_, (i, j) in enumerate(((0, 1), (2, 3), (4, 5))):
            cov[..., i, j] = cov[..., j, i] = cor[n_]
        energy = d(energy, 1e8).float()
        total_charge = d(total_charge, 0.0).float()
        return cls._generate(mean, cov, num_particles, energy, total_charge, device, seed,
                             first_particle)

    @classmethod
    def _generate(cls, mean, cov, num_particles, energy, total_charge, device, seed,
                  first_particle=0):
        """mean (*shape, 6), cov (*shape, 6, 6) float64 on the host -> beam on ``device``."""
        device = torch.device(device if device is not None else "cuda")
        shape = tuple(mean.shape[:-1])
        B = 1
        for s_ in shape:
            B *= s_
        # a square root of cov in float64: Cholesky, or V sqrt(L) when only semi-definite
        # (sigma = 0 is legal: MultivariateNormal would reject it, the scale-tril form does not)
        root, info = torch.linalg.cholesky_ex(cov)
        bad = info.reshape(-1) != 0
        if bool(bad.any()):
            lam, vec = torch.linalg.eigh(cov)
            alt = vec * torch.sqrt(torch.clamp_min(lam, 0.0)).unsqueeze(-2)
            root = torch.where(bad.reshape(*shape, 1, 1), alt, root)
        if seed is None:
            seed = int(torch.randint(0, 2**62, (1,), dtype=torch.int64).item())
        N = num_particles
        ld = (N + 3) & ~3
        soa = torch.empty((*shape, 6, ld), device=device, dtype=torch.float32)
        mean_d = mean.reshape(B, 6).to(device=device, dtype=torch.float32).contiguous()
        root_d = root.reshape(B, 6, 6).to(device=device, dtype=torch.float32).contiguous()
        L.check(L.lib().lynx_beam_generate(mean_d.data_ptr(), root_d.data_ptr(), B, N,
                                           int(first_particle), seed, soa.data_ptr(), ld,
                                           L.current_stream()), "lynx_beam_generate")
        # particle_beam.py:137-142: equal charges; kept as a stride-0 view (0 bytes)
        charges = (total_charge.to(device) / N).unsqueeze(-1).expand(*shape, N)
        return cls.from_soa(soa, N, energy.to(device), charges)

    @classmethod
    def from_ocelot(cls, parray, device=None, dtype=torch.float32) -> "ParticleBeam":
        """particle_beam.py:545-561: an Ocelot ``ParticleArray`` (``rparticles`` (6, N), ``E`` in
        GeV, ``q_array``); duck-typed, Ocelot is not imported."""
        import numpy as np

        coords = np.asarray(parray.rparticles, dtype=np.float64)
        aos = torch.ones((1, coords.shape[1], 7), dtype=torch.float32)
        aos[0, :, :6] = torch.from_numpy(coords.T.copy()).to(torch.float32)
        charges = torch.as_tensor(np.asarray(parray.q_array), dtype=torch.float32).unsqueeze(0)
        return cls(aos, torch.tensor([1e9 * float(parray.E)], dtype=torch.float32), charges,
                   device=device)

    @classmethod
    def from_astra(cls, path: str, device=None, dtype=torch.float32) -> "ParticleBeam":
        """particle_beam.py:563-578: ASTRA distribution file (``lynx_b200/astra.py``)."""
        from .astra import read_astra_beam

        coords, energy, charges = read_astra_beam(path)
        aos = torch.ones((1, coords.shape[0], 7), dtype=torch.float32)
        aos[0, :, :6] = torch.from_numpy(coords).to(torch.float32)
        return cls(aos, torch.tensor([energy], dtype=torch.float32),
                   torch.from_numpy(charges).to(torch.float32).unsqueeze(0), device=device)

    @classmethod
    def from_twiss(cls, num_particles=None, beta_x=None, alpha_x=None, emittance_x=None,
                   beta_y=None, alpha_y=None, emittance_y=None, energy=None, sigma_s=None,
                   sigma_p=None, cor_s=None, total_charge=None, device=None,
                   dtype=torch.float32, seed=None, first_particle=0):
        """particle_beam.py:172-264 (same defaults, incl. num_particles = 1 000 000)."""
        given = [a for a in (beta_x, alpha_x, emittance_x, beta_y, alpha_y, emittance_y, energy,
                             sigma_s, sigma_p, cor_s, total_charge) if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float64).cpu() if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float64))
        beta_x, alpha_x, emittance_x = d(beta_x, 0.0), d(alpha_x, 0.0), d(emittance_x, 0.0)
        beta_y, alpha_y, emittance_y = d(beta_y, 0.0), d(alpha_y, 0.0), d(emittance_y, 0.0)
        sigma_x = torch.sqrt(beta_x * emittance_x)
        sigma_xp = torch.sqrt(emittance_x * (1 + alpha_x**2) / beta_x)
        sigma_y = torch.sqrt(beta_y * emittance_y)
        sigma_yp = torch.sqrt(emittance_y * (1 + alpha_y**2) / beta_y)
        zero = torch.zeros(shape, dtype=torch.float64)
        return cls.from_parameters(
            num_particles=num_particles if num_particles is not None else 1_000_000,
            mu_x=zero, mu_xp=zero, mu_y=zero, mu_yp=zero, sigma_x=sigma_x, sigma_xp=sigma_xp,
            sigma_y=sigma_y, sigma_yp=sigma_yp, sigma_s=d(sigma_s, 1e-6), sigma_p=d(sigma_p, 1e-6),
            energy=d(energy, 1e8), cor_s=d(cor_s, 0.0), cor_x=-emittance_x * alpha_x,
            cor_y=-emittance_y * alpha_y, total_charge=d(total_charge, 0.0), device=device,
            dtype=dtype, seed=seed, first_particle=first_particle)

    @classmethod
    def uniform_3d_ellipsoid(cls, num_particles=None, radius_x=None, radius_y=None,
                             radius_s=None, sigma_xp=None, sigma_yp=None, sigma_p=None,
                             energy=None, total_charge=None, device=None, dtype=torch.float32,
                             seed=None, first_particle=0):
        """particle_beam.py:266-385 (same arguments and defaults: as written, a ``None``
        sigma falls through to ``from_parameters``' defaults 2e-7 / 2e-7 / 1e-6, :363-374),
        generated ON the device by ``lynx_beam_generate_ellipsoid``: per-particle rejection
        from the bounding box instead of the reference's host loop over whole candidate
        batches; the momentum planes are those of ``from_parameters(seed=seed)``."""
        assert dtype == torch.float32, "device path is fp32 (SURVEY.md section 5)"
        given = [a for a in (radius_x, radius_y, radius_s, sigma_xp, sigma_yp, sigma_p, energy,
                             total_charge) if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        num_particles = int(num_particles) if num_particles is not None else 1_000_000
        device = torch.device(device if device is not None else "cuda")
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float32).cpu() if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float32))
        radii = torch.stack([d(radius_x, 1e-3), d(radius_y, 1e-3), d(radius_s, 1e-3)], dim=-1)
        sigma = torch.stack([d(sigma_xp, 2e-7), d(sigma_yp, 2e-7), d(sigma_p, 1e-6)], dim=-1)
        B = 1
        for s_ in shape:
            B *= s_
        if seed is None:
            seed = int(torch.randint(0, 2**62, (1,), dtype=torch.int64).item())
        N = num_particles
        ld = (N + 3) & ~3
        soa = torch.empty((*shape, 6, ld), device=device, dtype=torch.float32)
        radii_d = radii.reshape(B, 3).to(device).contiguous()
        sigma_d = sigma.reshape(B, 3).to(device).contiguous()
        L.check(L.lib().lynx_beam_generate_ellipsoid(
            radii_d.data_ptr(), sigma_d.data_ptr(), B, N, int(first_particle), seed,
            soa.data_ptr(), ld, L.current_stream()), "lynx_beam_generate_ellipsoid")
        charges = (d(total_charge, 0.0).to(device) / N).unsqueeze(-1).expand(*shape, N)
        return cls.from_soa(soa, N, d(energy, 1e8).to(device), charges)

    @classmethod
    def make_linspaced(cls, num_particles=None, mu_x=None, mu_y=None, mu_xp=None, mu_yp=None,
                       sigma_x=None, sigma_y=None, sigma_xp=None, sigma_yp=None, sigma_s=None,
                       sigma_p=None, energy=None, total_charge=None, device=None,
                       dtype=torch.float32):
        """particle_beam.py:387-503: every coordinate ``linspace(mu - sigma, mu + sigma, n)``
        (a handful of reference particles for trace plots; built with torch on the device,
        no kernel of its own)."""
        given = [a for a in (mu_x, mu_xp, mu_y, mu_yp, sigma_x, sigma_xp, sigma_y, sigma_yp,
                             sigma_s, sigma_p, energy, total_charge) if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        n = int(num_particles) if num_particles is not None else 10
        device = torch.device(device if device is not None else "cuda")
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float32).to(device) if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float32, device=device))
        mus = [d(mu_x, 0.0), d(mu_xp, 0.0), d(mu_y, 0.0), d(mu_yp, 0.0), d(None, 0.0), d(None, 0.0)]
        sgs = [d(sigma_x, 175e-9), d(sigma_xp, 2e-7), d(sigma_y, 175e-9), d(sigma_yp, 2e-7),
               d(sigma_s, 0.0), d(sigma_p, 0.0)]
        # torch.linspace(lo, hi, n) per beam row on the host, exactly as particle_beam.py:474-535
        # spells it (linspace steps symmetrically from both ends, so mu + sigma * linspace(-1, 1)
        # differs from it in the last bit; a handful of trace particles, not a hot path)
        ld = (n + 3) & ~3
        flat = int(torch.Size(shape).numel())
        host = torch.zeros((flat, 6, ld), dtype=torch.float32)
        for k in range(6):
            mu_k, sg_k = mus[k].reshape(-1).cpu(), sgs[k].reshape(-1).cpu()
            for r in range(flat):
                host[r, k, :n] = torch.linspace(float(mu_k[r] - sg_k[r]), float(mu_k[r] + sg_k[r]), n)
        soa = host.reshape(*shape, 6, ld).to(device)
        charges = (d(total_charge, 0.0) / n).unsqueeze(-1).expand(*shape, n)
        return cls.from_soa(soa, n, d(energy, 1e8), charges)

    def transformed_to(self, mu_x=None, mu_y=None, mu_xp=None, mu_yp=None, sigma_x=None,
                       sigma_y=None, sigma_xp=None, sigma_yp=None, sigma_s=None, sigma_p=None,
                       energy=None, total_charge=None, device=None, dtype=torch.float32):
        """particle_beam.py:580-715: the particle pass ``(x - old_mu) / old_sigma * new_sigma
        + new_mu`` (:696-702) is ONE kernel (``lynx_beam_transform``, 48 B/particle); the old
        moments come from the fused sums (one 24 B/particle pass if the beam has none yet)."""
        shape = self.batch_shape
        given = [a for a in (mu_x, mu_xp, mu_y, mu_yp, sigma_x, sigma_xp, sigma_y, sigma_yp,
                             sigma_s, sigma_p, energy, total_charge) if a is not None]
        assert all(tuple(a.shape) == tuple(shape) for a in given), (
            "Arguments must have the same shape.")
        dev = self._soa.device
        g = lambda a, cur: (torch.as_tensor(a, dtype=torch.float32).to(dev) if a is not None  # noqa: E731
                            else cur)
        # all twelve old moments from the 28 raw sums in a handful of vector ops (the same
        # arithmetic as the mu_* / sigma_* properties: double sums, N-1, rounded to fp32)
        sums = self.moment_sums()
        cnt = sums[..., 0:1]
        first = sums[..., 1:7]
        diag = sums[..., [7 + _tri(i, i) for i in range(6)]]
        old_mu = (first / cnt).to(torch.float32)
        old_mu[..., 4:] = 0.0  # particle_beam.py:675-676, 687-688: s and p are not re-centred
        old_sg = torch.sqrt(torch.clamp_min((diag - first**2 / cnt) / (cnt - 1), 0.0)).to(
            torch.float32)
        new_mu, new_sg = old_mu.clone(), old_sg.clone()
        for i, a in enumerate((mu_x, mu_xp, mu_y, mu_yp)):
            if a is not None:
                new_mu[..., i] = g(a, None)
        for i, a in enumerate((sigma_x, sigma_xp, sigma_y, sigma_yp, sigma_s, sigma_p)):
            if a is not None:
                new_sg[..., i] = g(a, None)
        table = torch.cat([old_mu, old_sg, new_sg, new_mu], dim=-1).contiguous()  # (*shape, 24)
        B = 1
        for s_ in shape:
            B *= s_
        soa = self._soa if self._soa.is_contiguous() else self._soa.contiguous()
        out = torch.empty_like(soa)
        L.check(L.lib().lynx_beam_transform(soa.data_ptr(), 6 * self.ld, self.ld, max(B, 1),
                                            self._n, table.data_ptr(), out.data_ptr(), self.ld,
                                            L.current_stream()), "lynx_beam_transform")
        pc = self.particle_charges
        if total_charge is not None and pc is not None:  # particle_beam.py:655-664
            tq = torch.as_tensor(total_charge, dtype=torch.float32).to(dev)
            pc = torch.ones_like(pc) * tq.reshape(*shape, 1) / pc.shape[-1]
        return ParticleBeam.from_soa(out, self._n, g(energy, self.energy), pc)

    # ---- storage views
    @property
    def soa(self) -> torch.Tensor:
        return self._soa

    @property
    def ld(self) -> int:
        return self._soa.shape[-1]

    @property
    def batch_shape(self):
        return tuple(self._soa.shape[:-2])

    @property
    def num_particles(self) -> int:
        return self._n

    def __len__(self) -> int:
        return self._n

    @property
    def particles(self) -> torch.Tensor:
        """AoS ``(*batch, N, 7)`` view of the reference (particle_beam.py:35-39)."""
        R = 1
        for s in self.batch_shape:
            R *= s
        aos = torch.empty((*self.batch_shape, self._n, 7), device=self._soa.device,
                          dtype=torch.float32)
        L.check(L.lib().lynx_soa6_to_aos7(self._soa.data_ptr(), R, self._n, self.ld,
                                          aos.data_ptr(), L.current_stream()),
                "lynx_soa6_to_aos7")
        return aos

    @property
    def total_charge(self):
        if self.particle_charges is None:
            return torch.zeros(self.batch_shape, device=self._soa.device)
        return self.particle_charges.sum(dim=-1)

    def _plane(self, k):
        return self._soa[..., k, : self._n]

    def _set_plane(self, k, value):
        """particle_beam.py:732-734 et al. (``particles[..., k] = value``): written into the
        SoA plane in place; cached moment sums are dropped."""
        if self._soa.stride(-1) == 0 or any(
                st == 0 and sz > 1 for st, sz in zip(self._soa.stride(), self._soa.shape)):
            self._soa = self._soa.contiguous()  # a broadcast (stride-0) beam gets its own storage
        self._soa[..., k, : self._n] = torch.as_tensor(value, dtype=torch.float32).to(
            self._soa.device)
        self._sums = None

    xs = property(lambda self: self._plane(0), lambda self, v: self._set_plane(0, v))
    xps = property(lambda self: self._plane(1), lambda self, v: self._set_plane(1, v))
    ys = property(lambda self: self._plane(2), lambda self, v: self._set_plane(2, v))
    yps = property(lambda self: self._plane(3), lambda self, v: self._set_plane(3, v))
    ss = property(lambda self: self._plane(4), lambda self, v: self._set_plane(4, v))
    ps = property(lambda self: self._plane(5), lambda self, v: self._set_plane(5, v))

    # ---- moments (particle_beam.py:736-836)
    def moment_sums(self) -> torch.Tensor:
        """``(*batch, 28)`` float64: N, sum v_i, sum v_i v_j (i<=j)."""
        if self._sums is None:
            from .segment import beam_moment_sums

            self._sums = beam_moment_sums(self)
        return self._sums

    def _mean(self, i):
        s = self.moment_sums()
        return (s[..., 1 + i] / s[..., 0]).to(torch.float32)

    def _std(self, i):  # torch.std default: unbiased, N-1 (particle_beam.py:742)
        s = self.moment_sums()
        n = s[..., 0]
        var = (s[..., 7 + _tri(i, i)] - s[..., 1 + i] ** 2 / n) / (n - 1)
        return torch.sqrt(torch.clamp_min(var, 0.0)).to(torch.float32)

    def _cross(self, i, j):  # biased mean of centred products (particle_beam.py:824-836)
        s = self.moment_sums()
        n = s[..., 0]
        return ((s[..., 7 + _tri(i, j)] - s[..., 1 + i] * s[..., 1 + j] / n) / n).to(
            torch.float32
        )

    sigma_xxp = property(lambda self: self._cross(0, 1))
    sigma_yyp = property(lambda self: self._cross(2, 3))

    def broadcast(self, shape: tuple) -> "ParticleBeam":
        """particle_beam.py:838-843 -- tiled by stride-0 views, nothing is copied."""
        soa = self._soa.expand(*shape, *self._soa.shape[-2:]) if self.batch_shape in ((), (1,)) \
            else self._soa.repeat(*shape, 1, 1)
        energy = self.energy.expand(shape) if self.energy.numel() == 1 else self.energy.repeat(shape)
        pc = self.particle_charges
        if pc is not None:
            pc = pc.expand(*shape, pc.shape[-1]) if pc.shape[:-1] in ((), (1,)) else pc.repeat(*shape, 1)
        return ParticleBeam.from_soa(soa, self._n, energy, pc)

    def __repr__(self) -> str:
        return f"ParticleBeam(n={self._n}, batch={self.batch_shape}, energy={self.energy})"


for _name, _i in _COORD.items():
    setattr(ParticleBeam, f"mu_{_name}", property(lambda self, i=_i: self._mean(i)))
    setattr(ParticleBeam, f"sigma_{_name}", property(lambda self, i=_i: self._std(i)))


class ParameterBeam(Beam):
    """Beam described by ``mu (*batch, 7)`` and ``cov (*batch, 7, 7)``
    (parameter_beam.py:25-45)."""

    def __init__(self, mu, cov, energy, total_charge=None, device=None, dtype=torch.float32):
        assert dtype == torch.float32
        device = torch.device(device if device is not None else "cuda")
        kw = {"device": device, "dtype": torch.float32}
        self._mu = torch.as_tensor(mu, **kw)
        self._cov = torch.as_tensor(cov, **kw)
        self.total_charge = torch.as_tensor(
            total_charge if total_charge is not None else [0.0], **kw
        )
        self.energy = torch.as_tensor(energy, **kw)

    @classmethod
    def from_parameters(cls, mu_x=None, mu_xp=None, mu_y=None, mu_yp=None, sigma_x=None,
                        sigma_xp=None, sigma_y=None, sigma_yp=None, sigma_s=None,
                        sigma_p=None, cor_x=None, cor_y=None, cor_s=None, energy=None,
                        total_charge=None, device=None, dtype=torch.float32):
        """parameter_beam.py:47-146 (same defaults)."""
        given = [a for a in (mu_x, mu_xp, mu_y, mu_yp, sigma_x, sigma_xp, sigma_y, sigma_yp,
                             sigma_s, sigma_p, cor_x, cor_y, cor_s, energy, total_charge)
                 if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float64).cpu() if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float64))
        mu_x, mu_xp, mu_y, mu_yp = d(mu_x, 0.0), d(mu_xp, 0.0), d(mu_y, 0.0), d(mu_yp, 0.0)
        sig = [d(sigma_x, 175e-9), d(sigma_xp, 2e-7), d(sigma_y, 175e-9), d(sigma_yp, 2e-7),
               d(sigma_s, 1e-6), d(sigma_p, 1e-6)]
        cor = [d(cor_x, 0.0), d(cor_y, 0.0), d(cor_s, 0.0)]
        mu = torch.stack([mu_x, mu_xp, mu_y, mu_yp, torch.zeros(shape, dtype=torch.float64),
                          torch.zeros(shape, dtype=torch.float64),
                          torch.ones(shape, dtype=torch.float64)], dim=-1).float()
        cov = torch.zeros(*shape, 7, 7)
        for i in range(6):
            cov[..., i, i] = (sig[i].float()) ** 2
        for n, (i, j) in enumerate(((0, 1), (2, 3), (4, 5))):
            cov[..., i, j] = cov[..., j, i] = cor[n].float()
        return cls(mu, cov, d(energy, 1e8).float(), d(total_charge, 0.0).float(), device=device)

    @classmethod
    def from_ocelot(cls, parray, device=None, dtype=torch.float32) -> "ParameterBeam":
        """parameter_beam.py:234-253: mean and ``np.cov`` of an Ocelot ``ParticleArray``."""
        import numpy as np

        coords = np.asarray(parray.rparticles, dtype=np.float64)
        mu = torch.ones(1, 7, dtype=torch.float32)
        mu[0, :6] = torch.from_numpy(coords.mean(axis=1)).to(torch.float32)
        cov = torch.zeros(1, 7, 7, dtype=torch.float32)
        cov[0, :6, :6] = torch.from_numpy(np.cov(coords)).to(torch.float32)
        return cls(mu, cov, torch.tensor([1e9 * float(parray.E)], dtype=torch.float32),
                   torch.tensor([float(np.sum(parray.q_array))], dtype=torch.float32), device=device)

    @classmethod
    def from_astra(cls, path: str, device=None, dtype=torch.float32) -> "ParameterBeam":
        """parameter_beam.py:255-276: mean and (unbiased, ``np.cov``) covariance of the
        ASTRA distribution, total charge = sum of the macro-particle charges."""
        import numpy as np

        from .astra import read_astra_beam

        coords, energy, charges = read_astra_beam(path)
        mu = torch.ones(1, 7, dtype=torch.float32)
        mu[0, :6] = torch.from_numpy(coords.mean(axis=0)).to(torch.float32)
        cov = torch.zeros(1, 7, 7, dtype=torch.float32)
        cov[0, :6, :6] = torch.from_numpy(np.cov(coords.T)).to(torch.float32)
        return cls(mu, cov, torch.tensor([energy], dtype=torch.float32),
                   torch.tensor([charges.sum()], dtype=torch.float32), device=device)

    @classmethod
    def from_twiss(cls, beta_x=None, alpha_x=None, emittance_x=None, beta_y=None, alpha_y=None,
                   emittance_y=None, sigma_s=None, sigma_p=None, cor_s=None, energy=None,
                   total_charge=None, device=None, dtype=torch.float32):
        """parameter_beam.py:147-233 (same defaults: beta 1, alpha 0, emittance 7.1971891e-13)."""
        given = [a for a in (beta_x, alpha_x, emittance_x, beta_y, alpha_y, emittance_y, sigma_s,
                             sigma_p, cor_s, energy, total_charge) if a is not None]
        shape = tuple(given[0].shape) if given else (1,)
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        d = lambda a, v: (torch.as_tensor(a, dtype=torch.float64).cpu() if a is not None  # noqa: E731
                          else torch.full(shape, v, dtype=torch.float64))
        beta_x, alpha_x, emittance_x = d(beta_x, 1.0), d(alpha_x, 0.0), d(emittance_x, 7.1971891e-13)
        beta_y, alpha_y, emittance_y = d(beta_y, 1.0), d(alpha_y, 0.0), d(emittance_y, 7.1971891e-13)
        assert bool((beta_x > 0).all()), "Beta function in x direction must be larger than 0 everywhere."
        assert bool((beta_y > 0).all()), "Beta function in y direction must be larger than 0 everywhere."
        return cls.from_parameters(
            sigma_x=torch.sqrt(emittance_x * beta_x),
            sigma_xp=torch.sqrt(emittance_x * (1 + alpha_x**2) / beta_x),
            sigma_y=torch.sqrt(emittance_y * beta_y),
            sigma_yp=torch.sqrt(emittance_y * (1 + alpha_y**2) / beta_y),
            sigma_s=d(sigma_s, 1e-6), sigma_p=d(sigma_p, 1e-6), energy=d(energy, 1e8),
            cor_s=d(cor_s, 0.0), cor_x=-emittance_x * alpha_x, cor_y=-emittance_y * alpha_y,
            total_charge=d(total_charge, 0.0), device=device, dtype=dtype)

    def transformed_to(self, mu_x=None, mu_xp=None, mu_y=None, mu_yp=None, sigma_x=None,
                       sigma_xp=None, sigma_y=None, sigma_yp=None, sigma_s=None, sigma_p=None,
                       energy=None, total_charge=None, device=None, dtype=torch.float32):
        """parameter_beam.py:278-369: a new beam from the requested moments, the rest taken
        from this beam (as written there, the correlations are NOT carried over)."""
        shape = tuple(self.mu_x.shape)
        given = [a for a in (mu_x, mu_xp, mu_y, mu_yp, sigma_x, sigma_xp, sigma_y, sigma_yp,
                             sigma_s, sigma_p, energy, total_charge) if a is not None]
        assert all(tuple(a.shape) == shape for a in given), "Arguments must have the same shape."
        device = device if device is not None else self._mu.device
        g = lambda a, cur: a if a is not None else cur  # noqa: E731
        return self.__class__.from_parameters(
            mu_x=g(mu_x, self.mu_x), mu_xp=g(mu_xp, self.mu_xp), mu_y=g(mu_y, self.mu_y),
            mu_yp=g(mu_yp, self.mu_yp), sigma_x=g(sigma_x, self.sigma_x),
            sigma_xp=g(sigma_xp, self.sigma_xp), sigma_y=g(sigma_y, self.sigma_y),
            sigma_yp=g(sigma_yp, self.sigma_yp), sigma_s=g(sigma_s, self.sigma_s),
            sigma_p=g(sigma_p, self.sigma_p), energy=g(energy, self.energy),
            total_charge=g(total_charge, self.total_charge), device=device, dtype=dtype)

    def _sigma(self, i):  # parameter_beam.py:375-417
        return torch.sqrt(torch.clamp_min(self._cov[..., i, i], 1e-20))

    sigma_xxp = property(lambda self: self._cov[..., 0, 1])
    sigma_yyp = property(lambda self: self._cov[..., 2, 3])

    def broadcast(self, shape: tuple) -> "ParameterBeam":
        """parameter_beam.py:427-433 (stride-0 views instead of ``repeat``)."""
        one = self._mu.shape[:-1] in ((), (1,))
        mu = self._mu.reshape(7).expand(*shape, 7) if one else self._mu.repeat(*shape, 1)
        cov = self._cov.reshape(7, 7).expand(*shape, 7, 7) if one else self._cov.repeat(*shape, 1, 1)
        e = self.energy.reshape(()).expand(shape) if one else self.energy.repeat(shape)
        q = self.total_charge.reshape(()).expand(shape) if one else self.total_charge.repeat(shape)
        out = ParameterBeam.__new__(ParameterBeam)
        out._mu, out._cov, out.energy, out.total_charge = mu, cov, e, q
        return out

    def __repr__(self) -> str:
        return f"ParameterBeam(mu_x={self.mu_x}, sigma_x={self.sigma_x}, energy={self.energy})"


for _name, _i in _COORD.items():
    setattr(ParameterBeam, f"mu_{_name}", property(lambda self, i=_i: self._mu[..., i]))
    setattr(ParameterBeam, f"sigma_{_name}", property(lambda self, i=_i: self._sigma(i)))
