"""GPU parity of the adjoint path (K3a streaming adjoint, K3b dual-number parameter
gradients, moment collapse, K6-bwd) through ``torch.autograd`` -> the C ABI
(``lynx_track_particles_bwd`` / ``lynx_parambeam_bwd``).

Gradient oracle: torch-fp64 autograd through the restatement in ``oracle/`` -- the
reference pins NO gradient value (tests/test_differentiable.py only asserts that a graph
exists), so this parity is "vs. the restatement", DESIGN.md 3.  Tolerance (north star:
gradients within 1e-5 relative in fp32): |g - g64| <= 1e-5 * scale + |g32 - g64| with
scale = max |g64| over the gradient tensor of that parameter and g32 the fp32 oracle's own
autograd result (its distance from fp64 is the reference's fp32 noise).  Where a looser
bound is used the test says why.
"""

import numpy as np
import pytest
import torch

from helpers import aos_from_soa, oracle_elements

pytestmark = pytest.mark.gpu

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from oracle import lattice as O  # noqa: E402

QUADS = ("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")


def _close(got, g64, g32=None, rtol=1e-5, what=""):
    got = got.detach().cpu().double().numpy()
    g64 = g64.detach().double().numpy()
    scale = np.abs(g64).max() + 1e-300
    tol = rtol * scale
    if g32 is not None:
        tol = tol + np.abs(g32.detach().double().numpy() - g64)
    err = np.abs(got - g64)
    assert (err <= tol).all(), f"{what}: worst err/scale {(err / scale).max():.3e} (rtol {rtol})"


def _moment_loss(beam):
    """gradientbased.ipynb cell 8: mse([mu_x, sigma_x, mu_y, sigma_y], 0), summed."""
    return (beam.mu_x**2 + beam.sigma_x**2 + beam.mu_y**2 + beam.sigma_y**2).sum() / 4


def _oracle_moment_loss(out):
    return (out.mu(0) ** 2 + out.sigma(0) ** 2 + out.mu(2) ** 2 + out.sigma(2) ** 2).sum() / 4


def _ares_oracle_grads(k1, ang, soa, dtype, loss_fn, want_beam_grad=False):
    B = k1.shape[0]
    k1o = torch.tensor(k1, dtype=dtype, requires_grad=True)
    ango = torch.tensor(ang, dtype=dtype, requires_grad=True)
    ov = {(q, "k1"): k1o[:, i] for i, q in enumerate(QUADS)}
    ov[("AREAMCVM1", "angle")] = ango[:, 0]
    ov[("AREAMCHM1", "angle")] = ango[:, 1]
    els = oracle_elements(F.ares_ea(), dtype, (B,), ov)
    p0 = aos_from_soa(soa, dtype).clone().requires_grad_(want_beam_grad)
    beam = O.ParticleBeamO(p0.expand(B, -1, -1), torch.full((B,), F.ASTRA_ENERGY, dtype=dtype))
    loss = loss_fn(O.track(els, beam))
    loss.backward()
    return k1o.grad, ango.grad, (p0.grad if want_beam_grad else None), loss.detach()


def _ares_gpu(k1, ang):
    seg = F.build_segment(F.ares_ea())
    k1g = torch.tensor(k1, requires_grad=True)
    angg = torch.tensor(ang, requires_grad=True)
    for i, q in enumerate(QUADS):
        getattr(seg, q).k1 = k1g[:, i]
    seg.AREAMCVM1.angle = angg[:, 0]
    seg.AREAMCHM1.angle = angg[:, 1]
    return seg, k1g, angg


@pytest.mark.parametrize("streaming", [False, True])
def test_ares_moment_loss_gradients(streaming):
    """Config 5 in small: one beam shared by B settings, loss on (mu, sigma) of the
    tracked beam, gradients w.r.t. 3 x k1 + 2 x angle per setting.  streaming=False takes
    the moment-collapse path, streaming=True (the beam also requires grad) the K3a pass."""
    B, n = 16, 50_000
    k1, ang = F.random_settings(B, seed=1005, k1_range=5.0)
    k1[0, :] = 0.0  # the ARES default k1 = 0 (tangent kept)
    soa = F.gaussian_beam_soa(n, seed=1005)
    seg, k1g, angg = _ares_gpu(k1, ang)
    aos = torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0)
    beam = lx.ParticleBeam(aos, torch.tensor([F.ASTRA_ENERGY]))
    if streaming:
        beam.soa.requires_grad_(True)
    loss = _moment_loss(seg.track(beam))
    loss.backward()
    g64 = _ares_oracle_grads(k1, ang, soa, torch.float64, _oracle_moment_loss, streaming)
    g32 = _ares_oracle_grads(k1, ang, soa, torch.float32, _oracle_moment_loss, streaming)
    np.testing.assert_allclose(loss.item(), g64[3].item(), rtol=2e-5)
    # k1 == 0 rows: the oracle (masked write, track_methods.py:67-68) has zero gradient
    # there by construction; compare those rows against a finite k1 -> 0 limit instead
    live = np.abs(k1).min(axis=1) > 0
    _close(k1g.grad[live], g64[0][live], g32[0][live], what="d/dk1")
    _close(angg.grad, g64[1], g32[1], what="d/dangle")
    assert torch.isfinite(k1g.grad).all() and (k1g.grad[0] != 0).any()
    if streaming:
        gin = beam.soa.grad[0, :, :n].T  # (n, 6)
        _close(gin, g64[2][0, :, :6], g32[2][0, :, :6], what="d/dparticles")


def test_k1_zero_tangent_matches_limit():
    """At k1 == 0 the kernel keeps d/dk1 (the reference's masked write would zero it):
    it must equal the oracle gradient at k1 = 1e-9."""
    n = 20_000
    soa = F.gaussian_beam_soa(n, seed=3)
    k1 = np.zeros((1, 3), dtype=np.float32)
    ang = np.zeros((1, 2), dtype=np.float32)
    seg, k1g, angg = _ares_gpu(k1, ang)
    beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0),
                           torch.tensor([F.ASTRA_ENERGY]))
    _moment_loss(seg.track(beam)).backward()
    g64 = _ares_oracle_grads(k1 + 1e-9, ang, soa, torch.float64, _oracle_moment_loss)
    _close(k1g.grad, g64[0], rtol=2e-5, what="d/dk1 at 0")


def test_per_particle_cotangent_and_all_columns():
    """A loss that touches every outgoing coordinate (g_out path of K3a) and parameters
    of every kind on the ARES line: lengths, k1, misalignment, tilt, corrector angles."""
    n = 10_000
    rng = np.random.default_rng(7)
    soa = F.gaussian_beam_soa(n, seed=11)
    wts = rng.standard_normal((n, 6))

    def build(dtype, gpu):
        mk = (lambda v: torch.tensor([v], dtype=dtype, requires_grad=True))
        P = {"L": mk(0.3), "Lq": mk(0.122), "k1": mk(4.2), "tilt": mk(0.3), "ang": mk(2e-3),
             "mis": torch.tensor([[1e-4, -2e-4]], dtype=dtype, requires_grad=True)}
        if gpu:
            els = [lx.Drift(P["L"]), lx.Quadrupole(P["Lq"], P["k1"], P["mis"], P["tilt"]),
                   lx.Drift(P["L"] * 2), lx.HorizontalCorrector(torch.tensor([0.02]), P["ang"]),
                   lx.Drift(torch.tensor([0.45]))]
            return lx.Segment(els), P
        E = O.El
        els = [E("drift", "d", {"length": P["L"]}),
               E("quadrupole", "q", {"length": P["Lq"], "k1": P["k1"], "misalignment": P["mis"],
                                     "tilt": P["tilt"]}),
               E("drift", "d", {"length": P["L"] * 2}),
               E("hcor", "h", {"length": torch.tensor([0.02], dtype=dtype), "angle": P["ang"]}),
               E("drift", "d", {"length": torch.tensor([0.45], dtype=dtype)})]
        return els, P

    seg, Pg = build(torch.float32, True)
    beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0),
                           torch.tensor([F.ASTRA_ENERGY]))
    out = seg.track(beam)
    w = torch.tensor(wts, dtype=torch.float32, device="cuda")
    (out.soa[0, :, :n].T * w).sum().backward()
    res = {}
    for d in (torch.float64, torch.float32):
        els, Po = build(d, False)
        ob = O.ParticleBeamO(aos_from_soa(soa, d), torch.tensor([F.ASTRA_ENERGY], dtype=d))
        o = O.track(els, ob)
        (o.particles[0, :, :6] * torch.tensor(wts, dtype=d)).sum().backward()
        res[d] = Po
    for k in Pg:
        _close(Pg[k].grad, res[torch.float64][k].grad, res[torch.float32][k].grad, what=k)


def test_bpm_reading_gradients():
    """SMATCH with active BPMs: loss on the three BPM readings and the final moments."""
    n = 20_000
    soa = F.gaussian_beam_soa(n, mu=(1e-4, 1e-5, -2e-4, 2e-5, 0, 0), sigma=(2e-4, 2e-5, 2e-4, 2e-5, 1e-6, 1e-6),
                              seed=21)
    angles = (1e-4, -2e-4, 1.5e-4)

    def run(dtype, gpu):
        a = torch.tensor(angles, dtype=dtype, requires_grad=True)
        Ld = torch.tensor([7.0], dtype=dtype, requires_grad=True)
        spec = F.smatch(bpms_active=True)
        if gpu:
            seg = F.build_segment(spec)
            seg.V7SMATCH.angle, seg.H10SMATCH.angle, seg.H12SMATCH.angle = a[0:1], a[1:2], a[2:3]
            seg.D4.length = Ld
            beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0),
                                   torch.tensor([1e8]))
            out = seg.track(beam)
            r = [seg.BPM1SMATCH.reading, seg.BPM6SMATCH.reading, seg.BPM13SMATCH.reading]
            loss = sum(((i + 1) * x**2).sum() for i, x in enumerate(r)) + out.sigma_x.sum() ** 2
        else:
            ov = {("V7SMATCH", "angle"): a[0:1], ("H10SMATCH", "angle"): a[1:2],
                  ("H12SMATCH", "angle"): a[2:3], ("D4", "length"): Ld}
            els = oracle_elements(spec, dtype, (1,), ov)
            ob = O.ParticleBeamO(aos_from_soa(soa, dtype), torch.tensor([1e8], dtype=dtype))
            out = O.track(els, ob)
            r = [e.reading for e in els if e.kind == "bpm"]
            loss = sum(((i + 1) * x**2).sum() for i, x in enumerate(r)) + out.sigma(0).sum() ** 2
        loss.backward()
        return a.grad, Ld.grad

    g = run(torch.float32, True)
    g64, g32 = run(torch.float64, False), run(torch.float32, False)
    _close(g[0], g64[0], g32[0], what="d/dangle via BPMs")
    _close(g[1], g64[1], g32[1], what="d/dlength via BPMs")


def test_cavity_dipole_line_gradients():
    """Config-4 lattice (active cavity + dipole + quads): gradients through the non-linear
    cavity step (multi-pass K3a), incl. the energy change feeding the downstream maps."""
    n = 20_000
    soa = F.cfg4_beam_soa(n, seed=31)
    spec = F.cavity_dipole_line(screen_active=False)
    names = {"V": ("C1", "voltage"), "phi": ("C1", "phase"), "k1": ("Q1", "k1"),
             "k2": ("Q2", "k1"), "ang": ("B1", "angle"), "Lc": ("C1", "length")}
    vals = {"V": 0.01815975e9, "phi": 10.0, "k1": 4.2, "k2": -4.2, "ang": 0.1, "Lc": 1.0377}

    def run(dtype, gpu):
        P = {k: torch.tensor([v], dtype=dtype, requires_grad=True) for k, v in vals.items()}
        if gpu:
            seg = F.build_segment(spec)
            for k, (el, attr) in names.items():
                setattr(getattr(seg, el), attr, P[k])
            beam = lx.ParticleBeam(torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0),
                                   torch.tensor([F.CFG4_ENERGY]))
            out = seg.track(beam, moments=2)
            loss = (out.sigma_x**2 + out.sigma_y**2 + 1e2 * out.sigma_s**2
                    + 1e-4 * out.sigma_p**2 + out.mu_x**2 + 1e-3 * out.mu_p).sum()
        else:
            ov = {names[k]: P[k] for k in P}
            els = oracle_elements(spec, dtype, (1,), ov)
            ob = O.ParticleBeamO(aos_from_soa(soa, dtype), torch.tensor([F.CFG4_ENERGY], dtype=dtype))
            out = O.track(els, ob)
            loss = (out.sigma(0) ** 2 + out.sigma(2) ** 2 + 1e2 * out.sigma(4) ** 2
                    + 1e-4 * out.sigma(5) ** 2 + out.mu(0) ** 2 + 1e-3 * out.mu(5)).sum()
        loss.backward()
        return {k: P[k].grad for k in P}, loss.detach()

    (g, lg), (g64, l64), (g32, _) = run(torch.float32, True), run(torch.float64, False), run(torch.float32, False)
    np.testing.assert_allclose(lg.item(), l64.item(), rtol=1e-4)
    for k in g:
        # 1e-4: the cavity's T5xx / r55 terms are differences of nearly equal quantities
        # (evaluated in double once per setting, but their fp32 inputs carry 6e-8)
        _close(g[k], g64[k], g32[k], rtol=1e-4, what=k)


def test_parameter_beam_gradients():
    """K6-bwd: ARES EA over B settings, loss on the outgoing (mu, sigma); gradients
    w.r.t. k1 / angles and w.r.t. the incoming mu / cov."""
    B = 64
    k1, ang = F.random_settings(B, seed=1003, k1_range=5.0)
    seg, k1g, angg = _ares_gpu(k1, ang)
    t = lambda v: torch.tensor([v])  # noqa: E731
    pb = lx.ParameterBeam.from_parameters(
        mu_x=t(F.ASTRA_MU[0]), mu_xp=t(F.ASTRA_MU[1]), sigma_x=t(F.ASTRA_SIGMA[0]),
        sigma_xp=t(F.ASTRA_SIGMA[1]), sigma_y=t(F.ASTRA_SIGMA[2]), sigma_yp=t(F.ASTRA_SIGMA[3]),
        energy=t(F.ASTRA_ENERGY))
    pb._mu.requires_grad_(True)
    pb._cov.requires_grad_(True)
    out = seg.track(pb)
    loss = (out.mu_x**2 + out.sigma_x**2 + out.mu_y**2 + out.sigma_y**2 + out._cov[..., 0, 1]).sum()
    loss.backward()

    def oracle(dtype):
        k1o = torch.tensor(k1, dtype=dtype, requires_grad=True)
        ango = torch.tensor(ang, dtype=dtype, requires_grad=True)
        ov = {(q, "k1"): k1o[:, i] for i, q in enumerate(QUADS)}
        ov[("AREAMCVM1", "angle")] = ango[:, 0]
        ov[("AREAMCHM1", "angle")] = ango[:, 1]
        els = oracle_elements(F.ares_ea(), dtype, (B,), ov)
        mu0 = pb._mu.detach().cpu().to(dtype).clone().requires_grad_(True)
        cov0 = pb._cov.detach().cpu().to(dtype).clone().requires_grad_(True)
        ob = O.ParameterBeamO(mu0.expand(B, 7), cov0.expand(B, 7, 7),
                              torch.full((B,), F.ASTRA_ENERGY, dtype=dtype))
        o = O.track(els, ob)
        (o.mu(0) ** 2 + o.sigma(0) ** 2 + o.mu(2) ** 2 + o.sigma(2) ** 2 + o.cov_[..., 0, 1]).sum().backward()
        return k1o.grad, ango.grad, mu0.grad, cov0.grad

    g64, g32 = oracle(torch.float64), oracle(torch.float32)
    _close(k1g.grad, g64[0], g32[0], what="pb d/dk1")
    _close(angg.grad, g64[1], g32[1], what="pb d/dangle")
    _close(pb._mu.grad[..., :6], g64[2][..., :6], g32[2][..., :6], what="pb d/dmu")
    _close(pb._cov.grad[..., :6, :6], g64[3][..., :6, :6], g32[3][..., :6, :6], what="pb d/dcov")


def test_parameter_beam_cavity_gradients():
    """K6-bwd through the ParameterBeam cavity branch (cavity.py:134-140, 202-218)."""
    spec = F.cavity_dipole_line(screen_active=False)
    vals = {"V": 0.01815975e9, "phi": 10.0, "k1": 4.2, "ang": 0.1}
    names = {"V": ("C1", "voltage"), "phi": ("C1", "phase"), "k1": ("Q1", "k1"), "ang": ("B1", "angle")}
    sx, sxp, _ = F.twiss_sigmas(5.91253677, 3.55631308, 3.494768647122823e-09)

    def run(dtype, gpu):
        P = {k: torch.tensor([v], dtype=dtype, requires_grad=True) for k, v in vals.items()}
        if gpu:
            seg = F.build_segment(spec)
            for k, (el, attr) in names.items():
                setattr(getattr(seg, el), attr, P[k])
            t = lambda v: torch.tensor([v])  # noqa: E731
            pb = lx.ParameterBeam.from_parameters(sigma_x=t(sx), sigma_xp=t(sxp), sigma_y=t(sx),
                                                  sigma_yp=t(sxp), sigma_s=t(1e-5), sigma_p=t(1e-4),
                                                  energy=t(F.CFG4_ENERGY))
            out = seg.track(pb)
            loss = (out.sigma_x**2 + out.sigma_y**2 + out._cov[..., 4, 5] + out._mu[..., 5] * 1e-3
                    + out._cov[..., 5, 5]).sum()
        else:
            els = oracle_elements(spec, dtype, (1,), {names[k]: P[k] for k in P})
            ob = O.ParameterBeamO.from_parameters((0, 0, 0, 0), (sx, sxp, sx, sxp, 1e-5, 1e-4),
                                                  F.CFG4_ENERGY, dtype=dtype)
            out = O.track(els, ob)
            loss = (out.sigma(0) ** 2 + out.sigma(2) ** 2 + out.cov_[..., 4, 5]
                    + out.mu_[..., 5] * 1e-3 + out.cov_[..., 5, 5]).sum()
        loss.backward()
        return {k: P[k].grad for k in P}

    g, g64, g32 = run(torch.float32, True), run(torch.float64, False), run(torch.float32, False)
    for k in g:
        _close(g[k], g64[k], g32[k], rtol=1e-4, what="pb cavity " + k)


def test_streaming_adjoint_large_beam_refilled_stages():
    """K3a at 4.2 M particles (every CTA runs ~14 tiles through its 4-stage ring, stages are
    refilled): with a per-particle cotangent and a zero moment cotangent the incoming
    cotangent of a cavity-free lattice is exactly g_in = T^T g_out per particle.  Checked
    against float64 with K1's transfer map, and bit-identical between two runs."""
    import bench
    import lynx_b200 as lx
    from lynx_b200 import fixtures as F

    n = 4_200_000
    dev = torch.device("cuda", 0)
    w = bench.GradWorkload(n, 1, dev, per_particle=True)
    w.g_mom.zero_()
    w.g_out[..., n:] = 0
    w.forward()
    w.backward()
    g1 = w.g_in.clone()
    gp1 = w.g_params.clone()
    w.backward()
    assert torch.equal(g1, w.g_in)
    torch.testing.assert_close(gp1, w.g_params, rtol=1e-5, atol=1e-9)
    k1, ang = F.random_settings(1, 1005, k1_range=5.0)
    seg = F.build_segment(F.ares_ea(k1=tuple(k1[0]), vkick=float(ang[0, 0]), hkick=float(ang[0, 1])))
    tm = seg.transfer_map(torch.tensor([F.ASTRA_ENERGY]).cuda())[0].cpu().double().numpy()
    go = w.g_out[0, :, :n].cpu().double().numpy()
    want = tm[:6, :6].T @ go
    got = g1[:, :n].cpu().numpy()
    scale = np.sqrt(np.mean(want**2, axis=1, keepdims=True))
    assert np.max(np.abs(got - want) / scale) < 2e-5
    # and the forward's outgoing particles of the same launch shape: x_out = T [x; 1]
    x7 = np.concatenate([w.soa[:, :n].cpu().double().numpy(), np.ones((1, n))])
    ref = (tm @ x7)[:6]
    out = w.out[0, :, :n].cpu().numpy()
    assert np.max(np.abs(out - ref) / np.sqrt(np.mean(ref**2, axis=1, keepdims=True))) < 2e-5
