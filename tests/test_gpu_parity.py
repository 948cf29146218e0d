"""GPU parity: the CUDA path (through the C ABI) against the restatement oracle on the
same seeded inputs.  Tolerances (north star): coordinates / moments within 1e-5
relative in fp32 -- measured per SURVEY 7.3-3 against the fp64 oracle with the fp32
oracle's own distance from it allowed on top; screen bins exact for the kernel's own
coordinates, and vs. the oracle exact except particles within a few ulp of an edge.
"""

import numpy as np
import pytest
import torch

from helpers import aos_from_soa, assert_close_scaled, oracle_elements

pytestmark = pytest.mark.gpu

import lynx_b200 as lx  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402
from oracle import lattice as O  # noqa: E402
from oracle import screen_bin_edges, screen_reading_particles, segment_map  # noqa: E402


def _oracle_track(spec, soa, energy, dtype, **kw):
    els = oracle_elements(spec, dtype)
    beam = O.ParticleBeamO(aos_from_soa(soa, dtype), torch.tensor([energy], dtype=dtype))
    out = O.track(els, beam, **kw)
    return els, out


def _planes(beam_o):
    return beam_o.particles[0, :, :6].T.numpy()


def _gpu_beam(soa, energy):
    t = torch.from_numpy(F.soa_to_aos7(soa)).unsqueeze(0)
    return lx.ParticleBeam(t, torch.tensor([energy]))


# ------------------------------------------------------------------ K1 maps
def _rand(rng, lo, hi, n):
    return torch.tensor(rng.uniform(lo, hi, n), dtype=torch.float32)


EPS32 = 2.0 ** -23


def _abs_product(els, energy64):
    """prod |R_e| in fp64: the scale in the standard forward error bound of an fp32 matrix
    product, |fl(R_E..R_1) - R_E..R_1| <= c*eps*|R_E|..|R_1| -- the un-cancelled size of
    each entry, which is what the reference's own fp32 result is accurate relative to."""
    from oracle import element_map

    acc = torch.eye(7, dtype=torch.float64).repeat(*energy64.shape, 1, 1)
    for e in els:
        acc = element_map(e, energy64).abs() @ acc
    return acc.numpy()


def _check_maps(got, els, energy):
    """Per entry: |got - truth| <= 1e-5 |truth| + |ref32 - truth| + 16 eps32 * scale, with
    scale = the un-cancelled magnitude (abs-product bound, or the matrix max for a single
    element whose map is itself a rotation sandwich)."""
    els64 = [_cast(e, torch.float64) for e in els]
    t64 = segment_map(els64, energy.double()).numpy()
    t32 = segment_map(els, energy).numpy()
    g = got.cpu().numpy().astype(np.float64)
    scale = np.maximum(_abs_product(els64, energy.double()),
                       np.abs(t64).max(axis=(-1, -2), keepdims=True))
    tol = 1e-5 * np.abs(t64) + np.abs(t32 - t64) + 16 * EPS32 * scale
    err = np.abs(g - t64)
    if not (err <= tol).all():
        b, i, j = np.argwhere(err > tol)[0]
        raise AssertionError(
            f"{els[0].kind}: entry [{b},{i},{j}] got {g[b, i, j]!r} truth {t64[b, i, j]!r} "
            f"ref32 {t32[b, i, j]!r}; params " +
            str({k: (v[b].tolist() if torch.is_tensor(v) else v) for k, v in els[0].p.items()}) +
            f" energy {energy[b].item()}")


def _cast(el, dtype):
    return O.El(el.kind, el.name, {k: (v.to(dtype) if torch.is_tensor(v) else v)
                                   for k, v in el.p.items()}, el.is_active)


def test_k1_element_maps_vs_oracle():
    rng = np.random.default_rng(11)
    B = 257
    energy = _rand(rng, 5e6, 2e8, B)
    z = torch.zeros(B)
    cases = []
    L_ = _rand(rng, 0.05, 2.0, B)
    cases.append(([lx.Drift(L_)], [O.El("drift", "d", {"length": L_})]))
    k1 = _rand(rng, -30, 30, B)
    k1[:5] = 0.0
    mis = torch.stack([_rand(rng, -1e-3, 1e-3, B), _rand(rng, -1e-3, 1e-3, B)], dim=-1)
    tilt = _rand(rng, -1.5, 1.5, B)
    tilt[::3] = 0.0
    cases.append(([lx.Quadrupole(L_, k1, mis, tilt)],
                  [O.El("quadrupole", "q", {"length": L_, "k1": k1, "misalignment": mis,
                                            "tilt": tilt})]))
    ang = _rand(rng, -6e-3, 6e-3, B)
    cases.append(([lx.HorizontalCorrector(L_, ang)], [O.El("hcor", "h", {"length": L_, "angle": ang})]))
    cases.append(([lx.VerticalCorrector(L_, ang)], [O.El("vcor", "v", {"length": L_, "angle": ang})]))
    bang = _rand(rng, -0.4, 0.4, B)
    e1, e2 = _rand(rng, -0.3, 0.3, B), _rand(rng, -0.3, 0.3, B)
    fi, fe, gap = _rand(rng, 0, 0.5, B), _rand(rng, 0, 0.5, B), _rand(rng, 0, 0.05, B)
    dp = {"length": L_, "angle": bang, "e1": e1, "e2": e2, "tilt": tilt, "fringe_integral": fi,
          "fringe_integral_exit": fe, "gap": gap}
    cases.append(([lx.Dipole(L_, bang, e1, e2, tilt, fi, fe, gap)], [O.El("dipole", "b", dp)]))
    cases.append(([lx.RBend(L_, bang, e1, e2, tilt, fi, fe, gap)], [O.El("rbend", "b", dp)]))
    ks = _rand(rng, -3, 3, B)
    ks[:3] = 0.0
    cases.append(([lx.Solenoid(L_, ks, mis)],
                  [O.El("solenoid", "s", {"length": L_, "k": ks, "misalignment": mis})]))
    cases.append(([lx.Undulator(L_)], [O.El("undulator", "u", {"length": L_})]))
    for gpu_els, o_els in cases:
        got = lx.Segment(gpu_els).transfer_map(energy.cuda())
        _check_maps(got, o_els, energy)


def test_k1_ares_segment_map_batched():
    """Segment.transfer_map over 4096 random settings (config 3 ranges)."""
    B = 4096
    k1, ang = F.random_settings(B)
    seg = F.build_segment(F.ares_ea())
    seg.AREAMQZM1.k1 = torch.from_numpy(k1[:, 0])
    seg.AREAMQZM2.k1 = torch.from_numpy(k1[:, 1])
    seg.AREAMQZM3.k1 = torch.from_numpy(k1[:, 2])
    seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0])
    seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1])
    energy = torch.full((B,), F.ASTRA_ENERGY)
    got = seg.transfer_map(energy.cuda())
    ov = {("AREAMQZM1", "k1"): k1[:, 0], ("AREAMQZM2", "k1"): k1[:, 1],
          ("AREAMQZM3", "k1"): k1[:, 2], ("AREAMCVM1", "angle"): ang[:, 0],
          ("AREAMCHM1", "angle"): ang[:, 1]}
    els = oracle_elements(F.ares_ea(), torch.float32, (B,), ov)
    _check_maps(got, els, energy)


# ------------------------------------------------------------------ K2 particles
@pytest.mark.parametrize("n", [1, 5, 1000, 100_003])
def test_ares_ea_particles(n):
    """Config 2 lattice (k1 = 5,-5,5; kicks 1e-3,-2e-3: tests/test_compare_ocelot.py:254-258),
    ragged particle counts."""
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    soa = F.gaussian_beam_soa(n, seed=1002)
    _, t64 = _oracle_track(spec, soa, F.ASTRA_ENERGY, torch.float64)
    _, t32 = _oracle_track(spec, soa, F.ASTRA_ENERGY, torch.float32)
    out = F.build_segment(spec).track(_gpu_beam(soa, F.ASTRA_ENERGY))
    assert out.particles.shape == (1, n, 7)
    got = out.soa[0, :, :n].cpu().numpy()
    if n >= 1000:
        assert_close_scaled(got, _planes(t64), _planes(t32), what="ARES EA particles")
    else:
        np.testing.assert_allclose(got, _planes(t64), rtol=2e-4, atol=1e-9)
    assert torch.all(out.particles[..., 6] == 1)
    np.testing.assert_allclose(out.energy.cpu().numpy(), [F.ASTRA_ENERGY], rtol=1e-7)


def test_smatch_bpms_and_moments():
    """Config 1: README SMATCH with the three BPMs active, 100k particles."""
    spec = F.smatch(bpms_active=True, angles=(1e-4, -2e-4, 1.5e-4))
    sig = (175e-9, 2e-7, 175e-9, 2e-7, 1e-6, 1e-6)
    soa = F.gaussian_beam_soa(100_000, mu=(0.0,) * 6, sigma=sig, seed=1001)
    els, t64 = _oracle_track(spec, soa, 1e8, torch.float64)
    seg = F.build_segment(spec)
    out = seg.track(_gpu_beam(soa, 1e8))
    _, t32 = _oracle_track(spec, soa, 1e8, torch.float32)
    assert_close_scaled(out.soa[0, :, :100_000].cpu().numpy(), _planes(t64), _planes(t32),
                        what="SMATCH particles")
    for name in ("BPM1SMATCH", "BPM6SMATCH", "BPM13SMATCH"):
        want = next(e for e in els if e.name == name).reading.numpy()  # (2, 1)
        got = getattr(seg, name).reading.cpu().numpy()
        assert got.shape == (2, 1)
        scale = np.abs(_planes(t64)[[0, 2]]).max()
        np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-6 * scale)
    for i, nm in enumerate(("x", "xp", "y", "yp", "s", "p")):
        np.testing.assert_allclose(getattr(out, f"mu_{nm}").cpu().numpy(), t64.mu(i).numpy(),
                                   rtol=1e-5, atol=1e-6 * t64.sigma(i).item())
        np.testing.assert_allclose(getattr(out, f"sigma_{nm}").cpu().numpy(),
                                   t64.sigma(i).numpy(), rtol=1e-5)
    np.testing.assert_allclose(out.sigma_xxp.cpu().numpy(), t64.sigma_xxp.numpy(), rtol=1e-5)
    np.testing.assert_allclose(out.sigma_yyp.cpu().numpy(), t64.sigma_yyp.numpy(), rtol=1e-5)


def test_screen_histogram_exact_and_vs_oracle():
    """Config 2 with AREABSCR1 active (tests/test_screen.py:78-98 geometry)."""
    n = 200_000
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3, screen_active=True,
                     screen=F.TEST_SCREEN)
    soa = F.gaussian_beam_soa(n, seed=1002)
    seg = F.build_segment(spec)
    beam = _gpu_beam(soa, F.ASTRA_ENERGY)
    assert seg.track(beam) is lx.Beam.empty
    img = seg.AREABSCR1.reading
    assert img.shape == (1, 2040, 2448)
    assert img.sum().item() <= n and img.sum().item() > 0.5 * n
    # (a) bit-exact vs torch.histogramdd on the kernel's OWN coordinates
    read = seg.AREABSCR1.get_read_beam()
    ex, ey = screen_bin_edges(F.TEST_SCREEN["resolution"], F.TEST_SCREEN["pixel_size"])
    own = screen_reading_particles(read.xs.cpu(), read.ys.cpu(), ex, ey)
    assert torch.equal(img.cpu(), own)
    # (b) vs the oracle end to end, particle by particle: a particle may sit in another bin than
    # the oracle's only if its coordinate is within the coordinate tolerance of that bin edge
    from test_gpu_parity_r2 import check_bins_particle_by_particle

    res = {}
    for d in (torch.float32, torch.float64):
        els, _ = _oracle_track(spec, soa, F.ASTRA_ENERGY, d)
        res[d] = els[-1].read_beam
    o32, o64 = res[torch.float32], res[torch.float64]
    moved = check_bins_particle_by_particle(
        read.xs.cpu().numpy()[0], read.ys.cpu().numpy()[0], o32.col(0)[0].numpy(),
        o32.col(2)[0].numpy(), o64.col(0)[0].numpy(), o64.col(2)[0].numpy(), ex, ey, "config 2 screen")
    want = screen_reading_particles(o32.col(0), o32.col(2), ex, ey)
    assert (img.cpu() - want).abs().sum().item() <= 2 * moved
    assert moved <= 1e-3 * n, moved


def test_screen_edge_semantics_device():
    """Left-closed bins, right-closed last bin, out-of-range dropped -- on the device."""
    scr = lx.Screen(resolution=torch.tensor([4.0, 2.0]), pixel_size=torch.tensor([1.0, 1.0]),
                    is_active=True, name="s")
    seg = lx.Segment([scr])
    xs = [-2.0, -1.0, 1.999, 2.0, 2.1, -2.1, 0.5]
    ys = [-1.0, 0.0, 0.999, 1.0, 0.0, 0.0, -0.5]
    p = torch.zeros(1, len(xs), 7)
    p[0, :, 0] = torch.tensor(xs)
    p[0, :, 2] = torch.tensor(ys)
    p[0, :, 6] = 1
    seg.track(lx.ParticleBeam(p, torch.tensor([1e8])))
    expect = torch.tensor([[[0.0, 1.0, 0.0, 2.0], [1.0, 0.0, 1.0, 0.0]]])
    assert torch.equal(scr.reading.cpu(), expect)


def test_cavity_dipole_line_particles():
    """Config 4 lattice: non-linear cavity + dipole, 50k particles."""
    n = 50_000
    spec = F.cavity_dipole_line(screen_active=False)
    soa = F.cfg4_beam_soa(n)
    _, t64 = _oracle_track(spec, soa, F.CFG4_ENERGY, torch.float64)
    _, t32 = _oracle_track(spec, soa, F.CFG4_ENERGY, torch.float32)
    out = F.build_segment(spec).track(_gpu_beam(soa, F.CFG4_ENERGY))
    assert_close_scaled(out.soa[0, :, :n].cpu().numpy(), _planes(t64), _planes(t32),
                        what="cavity line particles")
    np.testing.assert_allclose(out.energy.cpu().numpy(), t64.energy.numpy(), rtol=1e-6)


def test_vectorized_shapes_and_broadcast_equivalence():
    """tests/test_vectorized.py:54,88,324-366: batch dims (2,) and (3,2); broadcast then
    track == track."""
    n = 4096
    spec = F.ares_ea(k1=(4.2, -3.0, 1.0), vkick=2e-4)
    soa = F.gaussian_beam_soa(n, seed=5)
    seg = F.build_segment(spec)
    beam = _gpu_beam(soa, F.ASTRA_ENERGY)
    base = seg.track(beam)
    for shape in [(2,), (3, 2)]:
        out = seg.broadcast(shape).track(beam.broadcast(shape))
        assert out.particles.shape == (*shape, n, 7)
        assert out.mu_x.shape == shape and out.sigma_p.shape == shape
        assert out.energy.shape == shape
        flat = out.soa.reshape(-1, 6, out.ld)
        for r in range(flat.shape[0]):
            assert torch.equal(flat[r], base.soa[0])


def test_per_setting_particles():
    """Three settings with different k1, per-setting outputs match three oracle runs."""
    n = 3000
    soa = F.gaussian_beam_soa(n, seed=9)
    k1s = [(-5.0, 2.0, 7.5), (0.0, 0.0, 0.0), (12.0, -20.0, 3.0)]
    seg = F.build_segment(F.ares_ea())
    for i, nm in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
        getattr(seg, nm).k1 = torch.tensor([k[i] for k in k1s])
    out = seg.track(_gpu_beam(soa, F.ASTRA_ENERGY))
    assert out.particles.shape == (3, n, 7)
    for b, k1 in enumerate(k1s):
        _, t64 = _oracle_track(F.ares_ea(k1=k1), soa, F.ASTRA_ENERGY, torch.float64)
        _, t32 = _oracle_track(F.ares_ea(k1=k1), soa, F.ASTRA_ENERGY, torch.float32)
        assert_close_scaled(out.soa[b, :, :n].cpu().numpy(), _planes(t64), _planes(t32),
                            what=f"setting {b}")


# ------------------------------------------------------------------ K6 ParameterBeam
def test_parameter_beam_4096_settings():
    """Config 3: 4096 random settings x ParameterBeam through ARES EA."""
    B = 4096
    k1, ang = F.random_settings(B)
    seg = F.build_segment(F.ares_ea())
    names = ("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")
    for i, nm in enumerate(names):
        getattr(seg, nm).k1 = torch.from_numpy(k1[:, i])
    seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0])
    seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1])
    kw = dict(zip(("mu_x", "mu_xp", "mu_y", "mu_yp"), (torch.tensor([v]) for v in F.ASTRA_MU[:4])))
    kw.update(dict(zip(("sigma_x", "sigma_xp", "sigma_y", "sigma_yp", "sigma_s", "sigma_p"),
                       (torch.tensor([v]) for v in F.ASTRA_SIGMA))))
    beam = lx.ParameterBeam.from_parameters(energy=torch.tensor([F.ASTRA_ENERGY]), **kw)
    out = seg.track(beam)
    assert out._mu.shape == (B, 7) and out._cov.shape == (B, 7, 7)
    ov = {(nm, "k1"): k1[:, i] for i, nm in enumerate(names)}
    ov[("AREAMCVM1", "angle")] = ang[:, 0]
    ov[("AREAMCHM1", "angle")] = ang[:, 1]
    res = {}
    for d in (torch.float32, torch.float64):
        els = oracle_elements(F.ares_ea(), d, (B,), ov)
        pb = O.ParameterBeamO.from_parameters(F.ASTRA_MU[:4], F.ASTRA_SIGMA, F.ASTRA_ENERGY, dtype=d)
        pb = O.ParameterBeamO(pb.mu_.expand(B, 7).clone(), pb.cov_.expand(B, 7, 7).clone(),
                              pb.energy.expand(B).clone())
        res[d] = O.track(els, pb)
    t64, t32 = res[torch.float64], res[torch.float32]
    mu = out._mu.cpu().numpy().astype(np.float64)
    cov = out._cov.cpu().numpy().astype(np.float64)
    m64, m32 = t64.mu_.numpy(), t32.mu_.numpy().astype(np.float64)
    c64, c32 = t64.cov_.numpy(), t32.cov_.numpy().astype(np.float64)
    # un-cancelled scales: |M| |mu| and |M| |cov| |M|^T with |M| the abs-product bound
    els64 = oracle_elements(F.ares_ea(), torch.float64, (B,), ov)
    A = _abs_product(els64, torch.full((B,), F.ASTRA_ENERGY, dtype=torch.float64))
    pb0 = O.ParameterBeamO.from_parameters(F.ASTRA_MU[:4], F.ASTRA_SIGMA, F.ASTRA_ENERGY,
                                           dtype=torch.float64)
    mu_s = np.einsum("bij,j->bi", A, np.abs(pb0.mu_[0].numpy()))
    cov_s = np.einsum("bij,jk,blk->bil", A, np.abs(pb0.cov_[0].numpy()), A)
    sd = np.sqrt(np.abs(np.einsum("bii->bi", c64)))
    tol_mu = 1e-5 * (np.abs(m64) + sd) + np.abs(m32 - m64) + 16 * EPS32 * mu_s
    assert (np.abs(mu - m64) <= tol_mu).all()
    tol_c = (1e-5 * (sd[:, :, None] * sd[:, None, :] + np.abs(c64)) + np.abs(c32 - c64)
             + 32 * EPS32 * cov_s)
    bad = np.abs(cov - c64) > tol_c
    assert not bad.any(), (np.argwhere(bad)[0], cov[bad][0], c64[bad][0], c32[bad][0])


def test_parameter_beam_cavity_and_bpm():
    spec = F.cavity_dipole_line(screen_active=False)
    spec.insert(5, ("bpm", "BPMX", {"is_active": True}))
    sx, sxp, cx = F.twiss_sigmas(5.91253677, 3.55631308, 3.494768647122823e-09)
    sy, syp, cy = F.twiss_sigmas(5.91253677, 3.55631308, 3.497810737006068e-09)
    t = lambda v: torch.tensor([v])  # noqa: E731
    beam = lx.ParameterBeam.from_parameters(
        mu_x=t(1e-4), mu_xp=t(-2e-5), sigma_x=t(sx), sigma_xp=t(sxp), sigma_y=t(sy),
        sigma_yp=t(syp), cor_x=t(cx), cor_y=t(cy), energy=t(F.CFG4_ENERGY))
    seg = F.build_segment(spec)
    out = seg.track(beam)
    d = torch.float64
    els = oracle_elements(spec, d)
    pb = O.ParameterBeamO.from_parameters((1e-4, -2e-5, 0.0, 0.0), (sx, sxp, sy, syp, 1e-6, 1e-6),
                                          F.CFG4_ENERGY, cor=(cx, cy, 0.0), dtype=d)
    want = O.track(els, pb)
    sd = np.sqrt(np.abs(np.diag(want.cov_[0].numpy())))
    np.testing.assert_allclose(out._mu[0].cpu().numpy(), want.mu_[0].numpy(), rtol=2e-5,
                               atol=2e-5 * sd.max())
    got_c, want_c = out._cov[0].cpu().numpy(), want.cov_[0].numpy()
    tol = 3e-5 * (np.outer(sd, sd) + np.abs(want_c))
    assert (np.abs(got_c - want_c) <= tol).all(), np.abs(got_c - want_c).max()
    bp = next(e for e in els if e.name == "BPMX").reading.numpy()
    np.testing.assert_allclose(seg.BPMX.reading.cpu().numpy(), bp, rtol=2e-5, atol=1e-9)
    np.testing.assert_allclose(out.energy.cpu().numpy(), want.energy.numpy(), rtol=1e-6)


# ------------------------------------------------------------------ layout + host ABI
def test_aos_soa_roundtrip():
    for n in (1, 255, 256, 257, 10_001):
        p = torch.randn(2, n, 7)
        p[..., 6] = 1
        b = lx.ParticleBeam(p, torch.tensor([1e8, 1e8]))
        assert torch.equal(b.particles.cpu(), p)
        assert torch.equal(b.xs.cpu(), p[..., 0]) and torch.equal(b.ps.cpu(), p[..., 5])


def test_host_buffer_abi_matches_device_path():
    """lynx_host_track_particles with NumPy buffers == the torch-plumbed device path."""
    import ctypes as C

    from lynx_b200 import _lib as L
    from lynx_b200.program import Program, encode, layout, pack_params

    n = 10_001
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3, screen_active=True,
                     screen=F.TEST_SCREEN)
    seg = F.build_segment(spec)
    soa = F.gaussian_beam_soa(n, seed=3)
    seg.track(_gpu_beam(soa, F.ASTRA_ENERGY))
    want = seg.AREABSCR1.reading.cpu().numpy()
    enc = encode(seg.elements, (1,))
    prog = Program.get(enc)
    params = pack_params(enc, "cpu").numpy()
    ex, ey = seg.AREABSCR1.pixel_bin_edges
    ex, ey = ex.numpy(), ey.numpy()
    W, H = ex.size - 1, ey.size - 1
    img = np.empty((1, H, W), dtype=np.float32)
    mom = np.empty((1, 28), dtype=np.float64)
    e_out = np.empty(1, dtype=np.float32)
    e_in = np.array([F.ASTRA_ENERGY], dtype=np.float32)
    ctx = C.c_void_p()
    L.check(L.lib().lynx_host_ctx_create(0, C.byref(ctx)))
    vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    soa_c = np.ascontiguousarray(soa)
    L.check(L.lib().lynx_host_track_particles(
        ctx, prog.handle, vp(params), params.size, 1, vp(soa_c), 1, n, vp(e_in), None, None,
        vp(img), W, H, vp(ex), vp(ey), vp(mom), 2, vp(e_out)))
    assert np.array_equal(img, want)
    assert mom[0, 0] == n

    # persistent-image mode: only the lit rectangle crosses PCIe; the host image must stay
    # bit-identical to the dense transfer across calls with different settings
    L.check(L.lib().lynx_host_ctx_set_option(ctx, 1, 1))
    pimg = np.full((1, H, W), 7.0, dtype=np.float32)  # garbage: a new buffer is zero-filled
    # (k1, vkick, hkick, pixel-size scale): moving / growing / shrinking spots (the speculative
    # rectangle of csrc/host.cu misses and a second copy follows), the same spot twice (it
    # hits), a beam kicked off the screen (nothing lit), and a changed screen geometry (the
    # cached edge tables must be re-sent)
    cases = (((5.0, -5.0, 5.0), 1e-3, -2e-3, 1.0), ((2.0, -3.0, 1.0), -1e-3, 1e-3, 1.0),
             ((9.0, -9.0, 9.0), 2e-3, 2e-3, 1.0), ((9.0, -9.0, 9.0), 2e-3, 2e-3, 1.0),
             ((5.0, -5.0, 5.0), 1e-3, 0.3, 1.0), ((5.0, -5.0, 5.0), 1e-3, -2e-3, 1.0),
             ((5.0, -5.0, 5.0), 1e-3, -2e-3, 0.5), ((5.0, -5.0, 5.0), 1e-3, -2e-3, 1.0))
    for k1, vk, hk, px in cases:
        scr_k = dict(F.TEST_SCREEN)
        scr_k["pixel_size"] = tuple(v * px for v in F.TEST_SCREEN["pixel_size"])
        spec_k = F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=scr_k)
        seg_k = F.build_segment(spec_k)
        seg_k.track(_gpu_beam(soa, F.ASTRA_ENERGY))
        want_k = seg_k.AREABSCR1.reading.cpu().numpy()
        if hk > 0.1:
            assert want_k.sum() == 0  # off the screen
        params_k = pack_params(encode(seg_k.elements, (1,)), "cpu").numpy()
        ex_k, ey_k = (e.numpy() for e in seg_k.AREABSCR1.pixel_bin_edges)
        L.check(L.lib().lynx_host_track_particles(
            ctx, prog.handle, vp(params_k), params_k.size, 1, vp(soa_c), 1, n, vp(e_in), None,
            None, vp(pimg), W, H, vp(ex_k), vp(ey_k), None, 0, None))
        assert np.array_equal(pimg, want_k), (k1, vk, hk, px)
    L.lib().lynx_host_ctx_destroy(ctx)


@pytest.mark.parametrize("n", [131_075, 700_001])
def test_host_buffer_abi_chunked_overlap(n):
    """Above 128 K particles the host entry point sends the beam in 2 (>= 128 K) or 4 (>= 512 K)
    chunks and tracks each as it lands (csrc/host.cu): image and outgoing particles must be the
    device path's bit for bit, the moment sums equal up to the order of their additions."""
    import ctypes as C

    from lynx_b200 import _lib as L
    from lynx_b200.program import Program, encode, pack_params

    soa = F.gaussian_beam_soa(n, seed=5)
    e_in = np.array([F.ASTRA_ENERGY], dtype=np.float32)
    vp = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    ctx = C.c_void_p()
    L.check(L.lib().lynx_host_ctx_create(0, C.byref(ctx)))
    soa_c = np.ascontiguousarray(soa)
    for screen_on in (True, False):
        spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3, screen_active=screen_on,
                         screen=F.TEST_SCREEN)
        seg = F.build_segment(spec)
        out_dev = seg.track(_gpu_beam(soa, F.ASTRA_ENERGY))
        enc = encode(seg.elements, (1,))
        prog = Program.get(enc)
        params = pack_params(enc, "cpu").numpy()
        mom = np.empty((1, 28), dtype=np.float64)
        if screen_on:
            want = seg.AREABSCR1.reading.cpu().numpy()
            ex, ey = (e.numpy() for e in seg.AREABSCR1.pixel_bin_edges)
            W, H = ex.size - 1, ey.size - 1
            for persistent in (0, 1, 1):
                L.check(L.lib().lynx_host_ctx_set_option(ctx, 1, persistent))
                img = np.zeros((1, H, W), dtype=np.float32)
                L.check(L.lib().lynx_host_track_particles(
                    ctx, prog.handle, vp(params), params.size, 1, vp(soa_c), 1, n, vp(e_in), None,
                    None, vp(img), W, H, vp(ex), vp(ey), vp(mom), 2, None))
                assert np.array_equal(img, want)
                assert mom[0, 0] == n
        else:
            out = np.empty((1, 6, n), dtype=np.float32)
            L.check(L.lib().lynx_host_track_particles(
                ctx, prog.handle, vp(params), params.size, 1, vp(soa_c), 1, n, vp(e_in), vp(out),
                None, None, 0, 0, None, None, vp(mom), 2, None))
            assert np.array_equal(out[0], out_dev.soa[0, :, :n].cpu().numpy())
            sums = out_dev.moment_sums().cpu().numpy().reshape(28)
            assert mom[0, 0] == n == sums[0]
            diag = [7 + d for d in (0, 6, 11, 15, 18, 20)]  # sum v_i^2: no cancellation
            np.testing.assert_allclose(mom[0, diag], sums[diag], rtol=1e-5)
            sd = np.sqrt(sums[diag] / n)
            np.testing.assert_allclose(mom[0, 1:7] / n, sums[1:7] / n, atol=1e-5 * sd.max(), rtol=1e-4)
    L.lib().lynx_host_ctx_destroy(ctx)


def test_forward_moment_collapse_matches_streaming():
    """One beam shared by B settings, moments / BPM readings only: the collapse path
    (beam streamed once, sums pushed through each setting's map) must agree with the
    per-setting streaming pass and with the oracle."""
    B, n = 12, 40_000
    soa = F.gaussian_beam_soa(n, mu=(2e-4, 1e-5, -1e-4, -2e-5, 0, 0),
                              sigma=(2e-4, 2e-5, 3e-4, 1e-5, 1e-5, 1e-3), seed=91)
    beam = _gpu_beam(soa, 1e8)
    rng = np.random.default_rng(5)
    ang = torch.tensor(rng.uniform(-3e-4, 3e-4, (B, 3)), dtype=torch.float32)

    def seg():
        s = F.build_segment(F.smatch(bpms_active=True))
        s.V7SMATCH.angle, s.H10SMATCH.angle, s.H12SMATCH.angle = ang[:, 0], ang[:, 1], ang[:, 2]
        return s

    s1, s2 = seg(), seg()
    fast = s1.track(beam, particles=False, moments=2)   # collapse
    slow = s2.track(beam, particles=True, moments=2)    # streaming, per-setting output
    assert fast.soa is None and slow.soa.shape[0] == B
    for nm in ("mu_x", "mu_y", "mu_xp", "sigma_x", "sigma_y", "sigma_xp", "sigma_s", "sigma_p",
               "sigma_xxp", "sigma_yyp", "energy"):
        np.testing.assert_allclose(getattr(fast, nm).cpu().numpy(), getattr(slow, nm).cpu().numpy(),
                                   rtol=1e-5, atol=1e-12, err_msg=nm)
    for a, b in ((s1.BPM1SMATCH, s2.BPM1SMATCH), (s1.BPM6SMATCH, s2.BPM6SMATCH),
                 (s1.BPM13SMATCH, s2.BPM13SMATCH)):
        assert a.reading.shape == (2, B)
        np.testing.assert_allclose(a.reading.cpu().numpy(), b.reading.cpu().numpy(), rtol=1e-5,
                                   atol=1e-10)
    # oracle on setting 3
    d = torch.float64
    ov = {("V7SMATCH", "angle"): ang[3:4, 0], ("H10SMATCH", "angle"): ang[3:4, 1],
          ("H12SMATCH", "angle"): ang[3:4, 2]}
    els = oracle_elements(F.smatch(bpms_active=True), d, (1,), ov)
    out = O.track(els, O.ParticleBeamO(aos_from_soa(soa, d), torch.tensor([1e8], dtype=d)))
    np.testing.assert_allclose(fast.sigma_x[3].item(), out.sigma(0).item(), rtol=1e-5)
    np.testing.assert_allclose(fast.mu_x[3].item(), out.mu(0).item(), rtol=1e-5, atol=1e-10)
    np.testing.assert_allclose(s1.BPM13SMATCH.reading[:, 3].cpu().numpy(),
                               els[-1].reading[:, 0].numpy(), rtol=1e-5, atol=1e-10)


def test_parameter_beam_screen_reading():
    """K7: Screen.reading for a ParameterBeam (screen.py:157-194) vs the oracle's
    MultivariateNormal restatement: same (B, W, H) layout incl. the x flip (quirk B8);
    values within 2e-5 of the peak (fp32 exp of an fp32 quadratic form)."""
    B = 5
    scr_spec = {"resolution": (120, 90), "pixel_size": (2e-5, 3e-5), "binning": 1}
    k1, ang = F.random_settings(B, seed=17, k1_range=8.0, steer=2e-3)
    seg = F.build_segment(F.ares_ea(screen_active=True, screen=scr_spec))
    for i, nm in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
        getattr(seg, nm).k1 = torch.from_numpy(k1[:, i])
    seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0])
    seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1])
    t = lambda v: torch.tensor([v])  # noqa: E731
    pb = lx.ParameterBeam.from_parameters(
        mu_x=t(F.ASTRA_MU[0]), mu_y=t(F.ASTRA_MU[2]), sigma_x=t(F.ASTRA_SIGMA[0]),
        sigma_xp=t(F.ASTRA_SIGMA[1]), sigma_y=t(F.ASTRA_SIGMA[2]), sigma_yp=t(F.ASTRA_SIGMA[3]),
        energy=t(F.ASTRA_ENERGY))
    assert seg.track(pb) is lx.Beam.empty
    img = seg.AREABSCR1.reading
    assert img.shape == (B, 120, 90)
    rb = seg.AREABSCR1.get_read_beam()
    want = O.screen_reading_parameter_beam(rb._mu.cpu().double(), rb._cov.cpu().double(),
                                           scr_spec["resolution"], scr_spec["pixel_size"])
    got = img.cpu().double()
    peak = want.amax(dim=(1, 2), keepdim=True)
    assert (peak > 0).all()
    assert ((got - want).abs() / peak).max().item() < 2e-5
    # x is flipped as in the reference: the brightest column sits where the oracle has it
    assert torch.equal(got.sum(dim=2).argmax(dim=1), want.sum(dim=2).argmax(dim=1))


@pytest.mark.parametrize("case", ["centred", "corner", "pencil", "wide", "drifting", "odd_width", "three_settings"])
def test_histogram_tile_placements(case):
    """The shared-memory histogram tile of K2 (csrc/track_pipe.cu) is placed per CTA from the
    statistics of the CTA's first pipeline tile; wherever it lands -- on the beam, against an image
    border, on a beam narrower than a pixel or far wider than the tile, on a beam that moves away
    while the CTA works through its tiles -- the image must equal torch.histogramdd of the same
    coordinates bit for bit.  300 k particles so that every CTA gets a tile (>= ~1400 particles per
    CTA); 'odd_width' has W % 4 != 0, which takes the path without the tile (16-byte rows are
    needed for the bulk reduce)."""
    n, W, H = 300_000, 640, 480
    px = (1e-5, 1.25e-5)
    rng = np.random.default_rng(123)
    B = 3 if case == "three_settings" else 1
    mis = torch.zeros(B, 2)
    x = rng.normal(0.0, 40 * px[0], n)
    y = rng.normal(0.0, 25 * px[1], n)
    if case == "corner":  # the beam hangs over the lower left corner: tile clamped to the image
        x += -W / 2 * px[0] + 10 * px[0]
        y += -H / 2 * px[1] + 5 * px[1]
    elif case == "pencil":
        x, y = rng.normal(3.3 * px[0], 0.05 * px[0], n), rng.normal(-7.7 * px[1], 0.05 * px[1], n)
    elif case == "wide":
        x, y = rng.uniform(-W / 2 * px[0], W / 2 * px[0], n), rng.uniform(-H / 2 * px[1], H / 2 * px[1], n)
    elif case == "drifting":  # the centre moves by 8 sigma over the particle index
        x += np.linspace(-160, 160, n) * px[0]
    elif case == "odd_width":
        W = 642
    elif case == "three_settings":
        mis = torch.tensor([[0.0, 0.0], [1.2e-3, 0.0], [-2.0e-3, 0.0]])
    p = torch.zeros(1, n, 7)
    p[0, :, 0], p[0, :, 2], p[0, :, 6] = torch.from_numpy(x).float(), torch.from_numpy(y).float(), 1.0
    scr = lx.Screen(resolution=torch.tensor([float(W), float(H)]), pixel_size=torch.tensor(px),
                    misalignment=mis, is_active=True, name="s")
    seg = lx.Segment([lx.Drift(torch.zeros(B)), scr])
    beam = lx.ParticleBeam(p, torch.tensor([1e8]))
    assert seg.track(beam.broadcast((B,)) if B > 1 else beam) is lx.Beam.empty
    img = scr.reading.cpu()
    assert tuple(img.shape) == (B, H, W)
    ex, ey = screen_bin_edges((W, H), px)
    xs = p[0, :, 0].unsqueeze(0) - mis[:, :1]
    want = screen_reading_particles(xs, p[0, :, 2].unsqueeze(0).expand(B, -1), ex, ey)
    assert torch.equal(img, want), f"{case}: {(img != want).sum().item()} pixels differ"
    assert img.sum().item() > 0.3 * n * B


def test_screen_exchange_kernel_single_rank():
    """csrc/exchange.cu on one GPU (world = 1: the rank is its own only peer): the fused
    bounding-box all-reduce must reproduce the dense image of lynx_track_particles_fwd, clear
    what the previous step lit and nothing else, and alternate its two slots -- the same
    kernel, epoch messages and persistent-rectangle logic the multi-GPU path runs
    (tools/multi_gpu_check.py covers world > 1)."""
    import ctypes as C

    from lynx_b200 import _lib as L
    from lynx_b200.program import Program, encode, pack_params

    lib = L.lib()
    dev = torch.device("cuda")
    n = 200_003
    soa = torch.from_numpy(F.gaussian_beam_soa(n, seed=77))
    ld = (n + 3) & ~3
    soa_d = torch.zeros((6, ld), device=dev)
    soa_d[:, :n] = soa.to(dev)
    out = state = buf = None
    for step, (k1, vk, hk) in enumerate([((5.0, -5.0, 5.0), 1e-3, -2e-3), ((3.0, -6.0, 4.0), -5e-4, 1e-3),
                                         ((0.0, 0.0, 0.0), 0.0, 0.0), ((8.0, -9.0, 7.0), 2e-3, 2e-3)]):
        seg = F.build_segment(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True,
                                        screen=F.TEST_SCREEN))
        enc = encode(seg.elements, (1,))
        prog = Program.get(enc)
        params = pack_params(enc, dev)
        ex_d, ey_d = seg.AREABSCR1._device_edges(dev)
        W, H = ex_d.numel() - 1, ey_d.numel() - 1
        if buf is None:
            nbytes = lib.lynx_screen_symm_bytes(1, W, H)
            buf = torch.zeros(nbytes // 4, dtype=torch.float32, device=dev)
            out = torch.zeros((1, H, W), device=dev)
            state = torch.zeros(8, dtype=torch.int32, device=dev)
            peers = (C.c_void_p * 1)(buf.data_ptr())
        e_in = torch.tensor([F.ASTRA_ENERGY], device=dev)
        e_out = torch.empty(1, device=dev)
        ws = torch.empty(prog.workspace_bytes(1), dtype=torch.uint8, device=dev)
        ref = torch.empty((1, H, W), device=dev)

        def track(img_ptr, bbox_ptr):
            L.check(lib.lynx_track_particles_fwd_bbox(
                prog.handle, params.data_ptr(), 1, soa_d.data_ptr(), 1, ld, n, None, ld, None,
                img_ptr, W, H, ex_d.data_ptr(), ey_d.data_ptr(), None, 0, e_in.data_ptr(),
                e_out.data_ptr(), ws.data_ptr(), bbox_ptr, L.current_stream()))

        track(ref.data_ptr(), None)
        slot = step & 1
        img = lib.lynx_screen_symm_image(buf.data_ptr(), slot, 1, W, H)
        track(img, img - 64)
        L.check(lib.lynx_screen_allreduce(peers, 1, 0, slot, step + 1, 1, W, H, out.data_ptr(),
                                          state.data_ptr(), L.current_stream()))
        torch.cuda.synchronize()
        assert torch.equal(out, ref), f"step {step}"
        assert float(out.sum()) > 0.5 * n  # some settings steer part of the beam off the screen


@pytest.mark.parametrize("n", [20_001, 4_200_000])
def test_persistent_screen_image_matches_fresh_image(n):
    """lynx_track_particles_fwd_persistent: the image and its 8-int state are zeroed ONCE; every
    call clears only the rectangle the previous call lit.  After each call the image must be
    bit-identical to what lynx_track_particles_fwd writes into a fresh buffer -- for a moving /
    growing / shrinking spot, a beam kicked off the screen, the same spot twice, and (4.2 M
    particles) with the histogram replicas + fold pass in play."""
    from lynx_b200 import _lib as L
    from lynx_b200.program import Program, encode, pack_params

    lib = L.lib()
    soa_np = F.gaussian_beam_soa(n, seed=21)
    ld = (n + 3) & ~3
    soa = torch.zeros((6, ld), device="cuda")
    soa[:, :n] = torch.from_numpy(soa_np).cuda()
    e_in = torch.tensor([F.ASTRA_ENERGY], device="cuda")
    e_out = torch.empty(1, device="cuda")
    img_p = state = None
    cases = (((5.0, -5.0, 5.0), 1e-3, -2e-3), ((2.0, -3.0, 1.0), -1e-3, 1e-3),
             ((9.0, -9.0, 9.0), 2e-3, 2e-3), ((9.0, -9.0, 9.0), 2e-3, 2e-3),
             ((5.0, -5.0, 5.0), 1e-3, 0.3), ((5.0, -5.0, 5.0), 1e-3, -2e-3),
             ((0.0, 0.0, 0.0), 0.0, 0.0))
    for epoch, (k1, vk, hk) in enumerate(cases, start=7):
        seg = F.build_segment(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True,
                                        screen=F.TEST_SCREEN))
        enc = encode(seg.elements, (1,))
        prog = Program.get(enc)
        params = pack_params(enc, "cuda")
        ex, ey = seg.AREABSCR1._device_edges(torch.device("cuda"))
        W, H = ex.numel() - 1, ey.numel() - 1
        ws = torch.empty(prog.workspace_bytes(1), dtype=torch.uint8, device="cuda")
        fresh = torch.full((1, H, W), 3.0, device="cuda")
        L.check(lib.lynx_track_particles_fwd(
            prog.handle, params.data_ptr(), 1, soa.data_ptr(), 1, ld, n, None, ld, None,
            fresh.data_ptr(), W, H, ex.data_ptr(), ey.data_ptr(), None, 0, e_in.data_ptr(),
            e_out.data_ptr(), ws.data_ptr(), L.current_stream()))
        if img_p is None:
            img_p = torch.zeros((1, H, W), device="cuda")
            state = torch.zeros(8, dtype=torch.int32, device="cuda")
        L.check(lib.lynx_track_particles_fwd_persistent(
            prog.handle, params.data_ptr(), 1, soa.data_ptr(), 1, ld, n, None, ld, None,
            img_p.data_ptr(), W, H, ex.data_ptr(), ey.data_ptr(), None, 0, e_in.data_ptr(),
            e_out.data_ptr(), ws.data_ptr(), state.data_ptr(), epoch, L.current_stream()))
        assert torch.equal(img_p, fresh), (epoch, k1, vk, hk)
        lit = float(fresh.sum())
        assert lit == (0.0 if hk > 0.1 else lit) and lit <= n
    # two settings at once (the rectangle is the union over the settings' images)
    seg = F.build_segment(F.ares_ea(k1=(5.0, -5.0, 5.0), screen_active=True, screen=F.TEST_SCREEN))
    seg.AREAMCHM1.angle = torch.tensor([-2e-3, 1.5e-3])
    enc = encode(seg.elements, (1,))
    prog = Program.get(enc)
    params = pack_params(enc, "cuda")
    ws = torch.empty(prog.workspace_bytes(2), dtype=torch.uint8, device="cuda")
    e2, e2o = e_in.expand(2).contiguous(), torch.empty(2, device="cuda")
    img2 = torch.zeros((2, H, W), device="cuda")
    st2 = torch.zeros(8, dtype=torch.int32, device="cuda")
    for epoch, ang in enumerate(((-2e-3, 1.5e-3), (1e-3, -1e-3), (1e-3, -1e-3))):
        seg.AREAMCHM1.angle = torch.tensor(ang)
        params = pack_params(encode(seg.elements, (1,)), "cuda")
        fresh = torch.empty((2, H, W), device="cuda")
        L.check(lib.lynx_track_particles_fwd(
            prog.handle, params.data_ptr(), 2, soa.data_ptr(), 0, ld, n, None, ld, None,
            fresh.data_ptr(), W, H, ex.data_ptr(), ey.data_ptr(), None, 0, e2.data_ptr(),
            e2o.data_ptr(), ws.data_ptr(), L.current_stream()))
        L.check(lib.lynx_track_particles_fwd_persistent(
            prog.handle, params.data_ptr(), 2, soa.data_ptr(), 0, ld, n, None, ld, None,
            img2.data_ptr(), W, H, ex.data_ptr(), ey.data_ptr(), None, 0, e2.data_ptr(),
            e2o.data_ptr(), ws.data_ptr(), st2.data_ptr(), epoch, L.current_stream()))
        assert torch.equal(img2, fresh), ang
        assert float(fresh[0].sum()) > 0 and not torch.equal(fresh[0], fresh[1])


def test_large_beam_refilled_pipeline_stages():
    """4.2 M particles: every CTA of the pipelined K2 runs ~14 tiles through its 4-stage ring,
    so stages are refilled while other warps may still be reading them.  (Regression: the
    screen variant released a stage right after ISSUING its shared-memory loads; above ~1.3 M
    particles some threads then saw a later tile's particles, and the image -- not its sum --
    changed from run to run.)  The image must be bit-identical between runs, equal to an
    independent histogram (torch.histogramdd semantics, the oracle's) of the coordinates the
    particle-output variant produces, and those coordinates must equal M x + kick in fp64."""
    n = 4_200_000
    soa_np = F.gaussian_beam_soa(n, seed=33)
    beam = lx.ParticleBeam.from_soa(torch.from_numpy(soa_np).cuda().unsqueeze(0).contiguous(), n,
                                    torch.tensor([F.ASTRA_ENERGY], device="cuda"))
    kw = dict(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    seg_s = F.build_segment(F.ares_ea(screen_active=True, screen=F.TEST_SCREEN, **kw))
    imgs = []
    for _ in range(3):
        assert seg_s.track(beam) is lx.Beam.empty
        imgs.append(seg_s.AREABSCR1.reading.clone())
    assert torch.equal(imgs[0], imgs[1]) and torch.equal(imgs[0], imgs[2])
    seg_o = F.build_segment(F.ares_ea(**kw))
    out = seg_o.track(beam)
    out2 = seg_o.track(beam)
    assert torch.equal(out.soa, out2.soa)
    ex, ey = seg_s.AREABSCR1.pixel_bin_edges
    want = screen_reading_particles(out.xs.cpu(), out.ys.cpu(), ex, ey)
    assert torch.equal(imgs[0].cpu(), want)
    tm = seg_o.transfer_map(torch.tensor([F.ASTRA_ENERGY]).cuda())[0].cpu().double().numpy()
    x7 = np.concatenate([soa_np.astype(np.float64), np.ones((1, n))])
    ref = (tm @ x7)[:6]
    got = out.soa[0, :, :n].cpu().numpy()
    scale = np.sqrt(np.mean(ref**2, axis=1, keepdims=True))
    assert np.max(np.abs(got - ref) / scale) < 2e-5
    # the moments-only variant shares the early stage hand-back: its sums against float64
    # statistics of the particle-output variant's coordinates
    mo = seg_o.track(beam, particles=False, moments=2)
    g64 = got.astype(np.float64)
    np.testing.assert_allclose(mo.mu_x.cpu().numpy(), [g64[0].mean()], rtol=1e-5, atol=1e-9)
    np.testing.assert_allclose(mo.mu_y.cpu().numpy(), [g64[2].mean()], rtol=1e-5, atol=1e-9)
    for i, c in enumerate(("x", "xp", "y", "yp", "s", "p")):
        np.testing.assert_allclose(getattr(mo, f"sigma_{c}").cpu().numpy(), [g64[i].std(ddof=1)],
                                   rtol=1e-5)
    mo2 = seg_o.track(beam, particles=False, moments=2)
    assert torch.equal(mo.moment_sums(), mo2.moment_sums()) or torch.allclose(
        mo.moment_sums(), mo2.moment_sums(), rtol=1e-12)


def test_non_finite_particles_propagate_like_a_dense_product():
    """A dense 7x7 product (element.py:83-92) turns every coordinate of a particle into NaN as
    soon as one of them is inf / NaN (0 * inf).  The block-sparse application of the cavity's
    own map must keep that behaviour, and must not touch the other particles."""
    n = 4099
    soa = F.cfg4_beam_soa(n, seed=5).copy()
    soa[4, 17] = np.inf   # s of particle 17
    soa[0, 123] = np.nan  # x of particle 123
    # C1 alone, and C1 + D2: the particles meet the cavity's block-diagonal map first
    spec = F.cavity_dipole_line(screen_active=False)[3:4]
    out = F.build_segment(spec).track(_gpu_beam(soa, F.CFG4_ENERGY))
    got = out.soa[0, :, :n].cpu().numpy()
    # s = inf also poisons the cavity's own p update (cos(inf)); x = NaN does not: as in the
    # reference, p is overwritten from the INCOMING s and p (cavity.py:142-161), so it stays finite
    assert np.isnan(got[:, 17]).all()
    assert np.isnan(got[:5, 123]).all() and np.isfinite(got[5, 123])
    ok = np.ones(n, dtype=bool)
    ok[[17, 123]] = False
    assert np.isfinite(got[:, ok]).all()
    clean = F.cfg4_beam_soa(n, seed=5)
    ref = F.build_segment(spec).track(_gpu_beam(clean, F.CFG4_ENERGY)).soa[0, :, :n].cpu().numpy()
    np.testing.assert_array_equal(got[:, ok], ref[:, ok])
    assert got[5, 123] == ref[5, 123]
    spec2 = F.cavity_dipole_line(screen_active=False)[3:5]
    got2 = F.build_segment(spec2).track(_gpu_beam(soa, F.CFG4_ENERGY)).soa[0, :, :n].cpu().numpy()
    assert np.isnan(got2[:, [17, 123]]).all() and np.isfinite(got2[:, ok]).all()


@pytest.mark.gpu
def test_block_structured_stage_maps_match_the_dense_product():
    """K1 flags a stage map that is three 2x2 blocks with a zero kick column (drifts, untilted
    aligned quadrupoles: compose_common.cuh write_map) and K2 applies it with apply_block2.  The
    result must be what the dense product gives: within tolerance of the oracle, and for
    particles with a non-finite coordinate the oracle's pattern of NaN and +-inf (0 * inf in the
    rows outside the coordinate's block, an infinity inside it) with nobody else touched -- the
    finite particles' bits do not depend on their neighbours.  The same line with a corrector in
    it has a kick column and takes the dense path: same properties."""
    n = 5003
    base = [("drift", "D0", {"length": 0.5}), ("quadrupole", "Q1", {"length": 0.2, "k1": 4.2}),
            ("drift", "D1", {"length": 0.3}), ("quadrupole", "Q2", {"length": 0.2, "k1": -3.1}),
            ("drift", "D2", {"length": 0.7})]
    kicked = base[:2] + [("hcor", "H1", {"length": 0.02, "angle": 1.5e-4})] + base[2:]
    clean = F.cfg4_beam_soa(n, seed=11)
    dirty = clean.copy()
    bad = [40, 77, 4001, 4500]
    dirty[5, 40] = np.inf     # p: s', p' stay infinite, the transverse rows become NaN
    dirty[0, 77] = -np.inf    # x: x', xp' stay infinite, the other rows become NaN
    dirty[2, 4001] = np.nan   # y: everything NaN
    dirty[4, 4500] = np.inf   # s: s' infinite, p' = 0 * inf + p = NaN, the rest NaN
    ok = np.ones(n, dtype=bool)
    ok[bad] = False
    for spec in (base, kicked):
        _, t64 = _oracle_track(spec, clean, F.CFG4_ENERGY, torch.float64)
        _, t32 = _oracle_track(spec, clean, F.CFG4_ENERGY, torch.float32)
        ref = F.build_segment(spec).track(_gpu_beam(clean, F.CFG4_ENERGY)).soa[0, :, :n].cpu().numpy()
        assert_close_scaled(ref, _planes(t64), _planes(t32), what="block / dense stage")
        got = F.build_segment(spec).track(_gpu_beam(dirty, F.CFG4_ENERGY)).soa[0, :, :n].cpu().numpy()
        # the dense product of the SAME merged map (K1's), row by row in fp32
        M = F.build_segment(spec).transfer_map(torch.tensor([F.CFG4_ENERGY]).cuda())[0].cpu().numpy()
        hom = np.concatenate([dirty[:6, bad], np.ones((1, len(bad)), np.float32)]).astype(np.float32)
        with np.errstate(invalid="ignore", over="ignore"):
            want = np.zeros((6, len(bad)), np.float32)
            for i in range(6):
                acc = np.full(len(bad), M[i, 6], np.float32)
                for k in range(6):
                    acc = (M[i, k] * hom[k] + acc).astype(np.float32)
                want[i] = acc
        np.testing.assert_array_equal(np.isnan(got[:, bad]), np.isnan(want))
        np.testing.assert_array_equal(np.where(np.isinf(got[:, bad]), np.sign(got[:, bad]), 0.0),
                                      np.where(np.isinf(want), np.sign(want), 0.0))
        assert np.isinf(got[:, bad]).any() and np.isnan(got[:, bad]).any()
        np.testing.assert_array_equal(got[:, ok], ref[:, ok])
    # and through the screen: the two poisoned particles are not counted, everybody else is where
    # the clean beam puts them
    scr = dict(F.ARES_SCREEN)
    scr["is_active"] = True
    seg = F.build_segment(base + [("screen", "SCR", scr)])
    seg.track(_gpu_beam(clean, F.CFG4_ENERGY))
    img_clean = seg.SCR.reading.clone()
    seg2 = F.build_segment(base + [("screen", "SCR", scr)])
    seg2.track(_gpu_beam(dirty, F.CFG4_ENERGY))
    img_dirty = seg2.SCR.reading
    assert 0.0 <= float(img_clean.sum() - img_dirty.sum()) <= 4.0
    assert int((img_clean != img_dirty).sum()) <= 4


def test_bins_of_particles_placed_on_and_around_every_edge_are_exact():
    """The fast bin lookup trusts floor(t) only where frac(t) is farther than a margin from 0 and 1
    (track_common.cuh: n 2^-22 + the measured deviation of the table); everything closer is
    decided by the edge table.  Adversarial beam: a screen alone (identity map, coordinates reach
    the screen bit for bit), particles ON every edge of both axes, a few ulp to either side, and
    around the margin itself (edge +- 0.05 ... 2 margins) where the fast path takes over.  The
    image must equal torch.histogramdd's exactly.  (Every fp32 coordinate within 7 ulp of an edge
    is in the beam; that the margin bounds the rounding error of t for ANY table is the
    derivation in track_common.cuh, not something a test can show.)"""
    scr = dict(F.ARES_SCREEN)
    scr["is_active"] = True
    ex, ey = screen_bin_edges(scr["resolution"], scr["pixel_size"], scr.get("binning", 1))
    ex64, ey64 = ex.double().numpy(), ey.double().numpy()

    def around(e):
        e32 = e.astype(np.float32)
        n = len(e) - 1
        step = (e[-1] - e[0]) / n
        m = n * 2.0 ** -22 + 2.0 ** -20
        pts = [e32]
        for k in (1, 2, 3, 4, 5, 6, 7):
            up, dn = e32.copy(), e32.copy()
            for _ in range(k):
                up, dn = np.nextafter(up, np.float32(np.inf)), np.nextafter(dn, np.float32(-np.inf))
            pts += [up, dn]
        for f in (0.05, 0.1, 0.2, 0.3, 0.5, 0.9, 0.99, 1.0, 1.01, 1.1, 1.5, 2.0):
            pts += [(e + f * m * step).astype(np.float32), (e - f * m * step).astype(np.float32)]
        return np.concatenate(pts)

    xs, ys = around(ex64), around(ey64)
    rng = np.random.default_rng(5)
    # pair every x candidate with a random interior y and the other way round
    x = np.concatenate([xs, rng.uniform(ex64[3], ex64[-4], len(ys)).astype(np.float32)])
    y = np.concatenate([rng.uniform(ey64[3], ey64[-4], len(xs)).astype(np.float32), ys])
    n = len(x)
    soa = np.zeros((6, n), np.float32)
    soa[0], soa[2] = x, y
    seg = F.build_segment([("screen", "SCR", scr)])
    seg.track(_gpu_beam(soa, F.ASTRA_ENERGY))
    got = seg.SCR.reading[0].cpu()
    want = screen_reading_particles(torch.from_numpy(x)[None], torch.from_numpy(y)[None], ex, ey)[0]
    assert got.shape == want.shape
    assert float(want.sum()) > 0.9 * n
    diff = (got != want)
    assert int(diff.sum()) == 0, f"{int(diff.sum())} pixels differ, counts {float(got.sum())} vs {float(want.sum())}"
