"""The oracle against vectors computed by THE REFERENCE ITSELF (tests/golden/reference_*.npz,
written by tests/golden/make_reference_golden.py, which runs /root/reference on torch through
tests/golden/reference_shim.py).  CPU only; needs neither the reference nor the script.

This is what pins the oracle: in fp32 -- the reference's precision -- the
restatement must reproduce the reference bit for bit wherever it follows the same operation
order, and to a few ulp of the un-cancelled magnitude where DESIGN.md 1.4 / 1.5 say it does not.
"""

import os

import numpy as np
import pytest
import torch

from helpers import oracle_elements
from lynx_b200 import fixtures as F
from oracle import lattice as O
from oracle import element_map, screen_bin_edges, screen_reading_particles, segment_map
from oracle.lattice import El

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
f32 = torch.float32


def _load(name):
    return np.load(os.path.join(GOLDEN, name))


def t(v):
    return torch.tensor([v], dtype=f32)


def _el(kind, **p):
    q = {k: (torch.tensor([v], dtype=f32) if k == "misalignment" else t(v)) for k, v in p.items()}
    if kind in ("quadrupole", "solenoid") and "misalignment" not in q:
        q["misalignment"] = torch.zeros(1, 2)
    if kind == "quadrupole":
        q.setdefault("tilt", t(0.0))
    if kind in ("dipole", "rbend"):
        for k in ("e1", "e2", "tilt", "fringe_integral", "gap"):
            q.setdefault(k, t(0.0))
        q.setdefault("fringe_integral_exit", q["fringe_integral"])
    return El(kind, kind, q)


ELEMENT_CASES = {
    "drift": _el("drift", length=0.7),
    "quadrupole": _el("quadrupole", length=0.122, k1=7.3),
    "quadrupole_defocus": _el("quadrupole", length=0.122, k1=-11.0),
    "quadrupole_tilt_mis": _el("quadrupole", length=0.2, k1=4.2, tilt=0.3, misalignment=(1e-3, -2e-3)),
    "hcor": _el("hcor", length=0.02, angle=1.5e-3),
    "vcor": _el("vcor", length=0.02, angle=-2.5e-3),
    "dipole": _el("dipole", length=0.5, angle=0.1, e1=0.05, e2=0.02, fringe_integral=0.5, gap=0.03),
    "dipole_tilt": _el("dipole", length=0.3, angle=-0.2, tilt=0.4, fringe_integral=0.4,
                       fringe_integral_exit=0.3, gap=0.02),
    "rbend": _el("rbend", length=0.5, angle=0.1, fringe_integral=0.5, gap=0.03),
    "solenoid": _el("solenoid", length=0.5, k=0.3),
    "undulator": _el("undulator", length=0.25),
    "cavity": _el("cavity", length=1.0377, voltage=1.0e7, phase=30.0, frequency=1.3e9),
    "cavity_decel": _el("cavity", length=4.139, voltage=2.0e6, phase=160.0, frequency=2.998e9),
}


@pytest.mark.parametrize("name", sorted(ELEMENT_CASES))
def test_element_maps_are_the_reference_s(name):
    """every element's ``transfer_map`` at three energies."""
    g = _load("reference_element_maps.npz")
    for i, energy in enumerate(g["energy"]):
        got = element_map(ELEMENT_CASES[name], torch.tensor([energy], dtype=f32))[0].numpy()
        np.testing.assert_array_equal(got, g[name][i], err_msg=f"{name} at {energy:g} eV")


def _ares(g, dtype=f32, **kw):
    k1, vk, hk = tuple(g["k1"].tolist()), float(g["vkick"]), float(g["hkick"])
    return oracle_elements(F.ares_ea(k1=k1, vkick=vk, hkick=hk, **kw), dtype)


def test_ares_ea_map_and_particles_bit_for_bit():
    """a10 / a12 / a13: Segment.transfer_map and Segment.track of 4 001 particles."""
    g = _load("reference_ares_ea.npz")
    els, energy = _ares(g), torch.tensor(g["energy"], dtype=f32)
    np.testing.assert_array_equal(segment_map(els, energy)[0].numpy(), g["segment_map"])
    out = O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy))
    np.testing.assert_array_equal(out.particles.numpy(), g["particles_out"])
    moments = {"mu_x": out.mu(0), "mu_y": out.mu(2), "sigma_x": out.sigma(0), "sigma_xp": out.sigma(1),
               "sigma_y": out.sigma(2), "sigma_yp": out.sigma(3), "sigma_s": out.sigma(4),
               "sigma_p": out.sigma(5), "sigma_xxp": out.sigma_xxp, "sigma_yyp": out.sigma_yyp}
    for name, got in moments.items():  # a17: the same torch reductions on the same bits
        np.testing.assert_array_equal(got.numpy(), g["out_" + name], err_msg=name)


def test_ares_ea_parameter_beam():
    """a11: mu' = R mu, cov' = R cov R^T."""
    g = _load("reference_ares_ea.npz")
    els, energy = _ares(g), torch.tensor(g["energy"], dtype=f32)
    beam = O.ParameterBeamO(torch.tensor(g["pb_mu_in"]), torch.tensor(g["pb_cov_in"]), energy)
    out = O.track(els, beam)
    np.testing.assert_array_equal(out.mu_.numpy(), g["pb_mu_out"])
    np.testing.assert_array_equal(out.cov_.numpy(), g["pb_cov_out"])


def test_ares_ea_screen_reading():
    """a15 / a16: the 2-D histogram the reference's Screen.reading returns, every bin."""
    g = _load("reference_ares_ea.npz")
    screen = {"resolution": tuple(int(v) for v in g["screen_resolution"]),
              "pixel_size": tuple(float(v) for v in g["screen_pixel_size"]), "binning": 1}
    els = _ares(g, screen_active=True, screen=screen)
    energy = torch.tensor(g["energy"], dtype=f32)
    assert O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy)) is O.EMPTY
    seen = [e for e in els if e.kind == "screen"][0].read_beam
    ex, ey = screen_bin_edges(screen["resolution"], screen["pixel_size"], 1)
    img = screen_reading_particles(seen.col(0), seen.col(2), ex, ey)  # (B, N) rows
    want = g["screen_reading"]
    assert tuple(want.shape[-2:]) == (48, 64) and want.sum() > 0
    np.testing.assert_array_equal(np.asarray(img).reshape(want.shape), want)


def test_cavity_dipole_line_particles():
    """a8 / a9 / a7: the non-linear cavity pass, the dipole and the outgoing energy."""
    g = _load("reference_cavity_line.npz")
    els = oracle_elements(F.cavity_dipole_line(screen_active=False), f32)
    energy = torch.tensor(g["energy"], dtype=f32)
    out = O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy))
    np.testing.assert_array_equal(out.energy.numpy(), g["energy_out"])
    np.testing.assert_array_equal(out.particles.numpy(), g["particles_out"])


def test_bpm_readings():
    """a14: BPM.reading = (mu_x, mu_y) of the beam at the BPM."""
    g = _load("reference_bpm_line.npz")
    els = oracle_elements(F.smatch(bpms_active=True, angles=(1e-3, -5e-4, 2e-4)), f32)
    energy = torch.tensor(g["energy"], dtype=f32)
    out = O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy))
    np.testing.assert_array_equal(out.particles.numpy(), g["particles_out"])
    bpms = [e for e in els if e.kind == "bpm"]
    assert [b.name for b in bpms] == list(g["bpm_names"])
    for b, want in zip(bpms, g["bpm_readings"]):
        np.testing.assert_allclose(np.asarray(b.reading).reshape(want.shape), want, rtol=2e-6, atol=1e-12)


# ------------------------------------------------------------------ beam-level operations
def test_transformed_to_is_the_reference_s():
    """particle_beam.py:580-715: statistics of the beam (torch fp32 mean / unbiased std, the s
    and p means zeroed, :666-693), then ``(x - old_mu) / old_sigma * new_sigma + new_mu``."""
    from oracle import beamops as OB

    g = _load("reference_beam_ops.npz")
    full = torch.tensor(g["particles_in"])  # (1, N, 7)
    p = full[0, :, :6]
    target = dict(zip(g["transform_target_names"].tolist(), g["transform_target"].tolist()))
    # the beam's own statistics exactly as particle_beam.py:736-822 takes them: one strided
    # column of the (1, N, 7) array at a time
    old_mu = torch.stack([full[..., i].mean(dim=-1)[0] for i in range(6)])
    old_sigma = torch.stack([full[..., i].std(dim=-1)[0] for i in range(6)])
    old_mu[4:] = 0.0
    coords = ("x", "xp", "y", "yp", "s", "p")
    new_mu = np.array([target.get(f"mu_{c}", 0.0) for c in coords], np.float32)
    new_sigma = np.array([target[f"sigma_{c}"] for c in coords], np.float32)
    got = OB.transformed_to(p.numpy().T, old_mu.numpy(), old_sigma.numpy(), new_mu, new_sigma)
    np.testing.assert_array_equal(got.T, g["transformed"][0, :, :6])
    assert (g["transformed"][0, :, 6] == 1).all()


@pytest.mark.parametrize("shape", ["rectangular", "elliptical"])
def test_aperture_is_the_reference_s(shape):
    """aperture.py:69-109: survivors and lost particles, both in incoming order."""
    from oracle import beamops as OB

    g = _load("reference_beam_ops.npz")
    planes = g["particles_in"][0, :, :6].T
    kept, lost, _, _ = OB.aperture_track(planes, g["aperture_x_max"], g["aperture_y_max"], shape)
    np.testing.assert_array_equal(kept.T, g[f"aperture_{shape}_kept"][:, :6])
    np.testing.assert_array_equal(lost.T, g[f"aperture_{shape}_lost"][:, :6])
    assert 0 < kept.shape[1] < planes.shape[1]


def test_astra_reader_is_the_reference_s(tmp_path):
    """converters/astra.py:8-62 on a synthetic distribution with two lost particles: the
    coordinates, the reference energy and the charges of lynx_b200.astra.read_astra_beam."""
    from lynx_b200.astra import read_astra_beam

    g = _load("reference_beam_ops.npz")
    path = str(tmp_path / "beam.ini")
    np.savetxt(path, g["astra_rows"], fmt="%.12e")
    coords, energy, charges = read_astra_beam(path)
    assert coords.shape == g["astra_coords"].shape == (499, 6)
    np.testing.assert_allclose(coords, g["astra_coords"], rtol=1e-12, atol=1e-15)  # delta cancels to ~1e-16
    assert energy == float(g["astra_energy"])
    np.testing.assert_array_equal(charges, g["astra_charges"])


# ------------------------------------------------------------------ ParameterBeam
ACCESSOR_RTOL = 2e-6  # fp32 accessors; the host mirror builds mu / cov in float64 first


def _kw(g, prefix):
    return {k: torch.tensor([v], dtype=torch.float32)
            for k, v in zip(g[prefix + "_names"].tolist(), g[prefix + "_values"].tolist())}


def _accessors(beam, names):
    return np.array([float(torch.as_tensor(getattr(beam, n)).reshape(-1)[0]) for n in names])


def test_parameter_beam_host_mirror_is_the_reference_s():
    """lynx_b200.ParameterBeam (host logic, CPU tensors) against the reference's
    from_parameters / from_twiss / transformed_to and every accessor of beam.py:241-310 and
    parameter_beam.py:371-425 (a18)."""
    import lynx_b200 as lx

    g = _load("reference_parameter_beam.npz")
    names = g["accessor_names"].tolist()
    pb = lx.ParameterBeam.from_parameters(device="cpu", **_kw(g, "fp"))
    np.testing.assert_allclose(pb._mu.numpy(), g["fp_mu"], rtol=1e-7, atol=0)
    np.testing.assert_allclose(pb._cov.numpy(), g["fp_cov"], rtol=2e-7, atol=0)
    np.testing.assert_allclose(_accessors(pb, names), g["fp_accessors"], rtol=ACCESSOR_RTOL, err_msg=str(names))
    tw = lx.ParameterBeam.from_twiss(device="cpu", **_kw(g, "tw"))
    np.testing.assert_allclose(tw._mu.numpy(), g["tw_mu"], rtol=1e-7, atol=0)
    np.testing.assert_allclose(tw._cov.numpy(), g["tw_cov"], rtol=2e-7, atol=1e-30)
    np.testing.assert_allclose(_accessors(tw, names), g["tw_accessors"], rtol=ACCESSOR_RTOL, atol=1e-12, err_msg=str(names))
    t1 = lambda v: torch.tensor([v], dtype=torch.float32)  # noqa: E731
    moved = pb.transformed_to(mu_x=t1(1e-3), sigma_x=t1(3e-4), sigma_yp=t1(2e-6), energy=t1(9e7))
    np.testing.assert_allclose(moved._mu.numpy(), g["tr_mu"], rtol=1e-7, atol=0)
    np.testing.assert_allclose(moved._cov.numpy(), g["tr_cov"], rtol=2e-7, atol=0)
    np.testing.assert_array_equal(moved.energy.numpy(), g["tr_energy"])


def test_parameter_beam_screen_pdf_is_the_reference_s():
    """8(f) row 1 / screen.py:157-194: the pdf image, its (B, W, H) layout and x flip (B8)."""
    from oracle.lattice import ParameterBeamO, screen_reading_parameter_beam

    g = _load("reference_parameter_beam.npz")
    beam = ParameterBeamO(torch.tensor(g["fp_mu"]), torch.tensor(g["fp_cov"]), torch.tensor([1.2e8]))
    scr = El("screen", "scr", {"misalignment": torch.tensor([[1e-4, -5e-5]]), "resolution": (64, 48),
                               "pixel_size": (4e-5, 4e-5), "binning": 1}, is_active=True)
    els = [El("drift", "d", {"length": t(1.0)}), scr]
    assert O.track(els, beam) is O.EMPTY
    seen = scr.read_beam
    img = screen_reading_parameter_beam(seen.mu_, seen.cov_, (64, 48), (4e-5, 4e-5), 1)
    assert tuple(img.shape) == tuple(g["screen_pdf"].shape) == (1, 64, 48)
    np.testing.assert_array_equal(img.numpy(), g["screen_pdf"])


def test_batched_settings_are_the_reference_s():
    """Config 3 in small: four settings of three quadrupoles and two correctors (a19: elements
    broadcast, then per-setting parameters) -- mu and cov per setting, bit for bit."""
    g = _load("reference_parameter_beam.npz")
    B = g["batch_k1"].shape[0]
    over = {(q, "k1"): g["batch_k1"][:, i] for i, q in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3"))}
    over[("AREAMCVM1", "angle")] = g["batch_angles"][:, 0]
    over[("AREAMCHM1", "angle")] = g["batch_angles"][:, 1]
    els = oracle_elements(F.ares_ea(), f32, batch=(B,), overrides=over)
    beam = O.ParameterBeamO(torch.tensor(g["fp_mu"]).repeat(B, 1), torch.tensor(g["fp_cov"]).repeat(B, 1, 1),
                            torch.full((B,), 1.2e8))
    out = O.track(els, beam)
    np.testing.assert_array_equal(out.mu_.numpy(), g["batch_mu"])
    np.testing.assert_array_equal(out.cov_.numpy(), g["batch_cov"])


# ------------------------------------------------------------------ quirks kept on purpose
def test_screen_misalignment_quirk_b1_is_the_reference_s():
    """screen.py:134-135 subtracts the y misalignment from column 1 (xp), so the image moves in
    x only: the restatement with the quirk reproduces the reference's image bin for bin, the
    physically meant variant does not."""
    g = _load("reference_quirks.npz")
    k1, vk, hk = tuple(g["k1"].tolist()), float(g["vkick"]), float(g["hkick"])
    screen = {"resolution": (64, 48), "pixel_size": (4e-5, 4e-5), "binning": 1,
              "misalignment": tuple(float(v) for v in g["screen_misalignment"])}
    energy = torch.tensor(g["energy"], dtype=f32)
    ex, ey = screen_bin_edges(screen["resolution"], screen["pixel_size"], 1)
    images = {}
    for quirk in (True, False):
        els = oracle_elements(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=screen), f32)
        O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy), screen_quirk_b1=quirk)
        seen = [e for e in els if e.kind == "screen"][0].read_beam
        images[quirk] = screen_reading_particles(seen.col(0), seen.col(2), ex, ey).numpy()
    np.testing.assert_array_equal(images[True], g["screen_reading_misaligned"])
    assert (images[False] != g["screen_reading_misaligned"]).any()


def test_zero_voltage_cavity_quirk_b6_is_the_reference_s():
    """cavity.py:269: 0 * inf -- the reference's map of a V = 0 cavity holds NaN in R12 / R34.
    The restatement reproduces it as written; ``cavity_v0_as_drift`` (what the kernels do,
    DESIGN.md 1.4) gives the drift map those entries converge to."""
    g = _load("reference_quirks.npz")
    cav = _el("cavity", length=1.0, voltage=0.0, phase=0.0, frequency=1.3e9)
    e = torch.tensor([1e8], dtype=f32)
    as_written = element_map(cav, e)[0].numpy()
    want = g["cavity_v0_map"]
    assert np.isnan(want[0, 1]) and np.isnan(want[2, 3]) and np.isnan(want).sum() == 2
    np.testing.assert_array_equal(as_written, want)  # NaNs in the same places, the rest equal
    repaired = element_map(cav, e, cavity_v0_as_drift=True)[0].numpy()
    assert np.isfinite(repaired).all()
    mask = ~np.isnan(want)
    mask[4, 5] = False  # the drift carries R56 = -L / (beta gamma)^2 = -2.6e-5 here, the cavity formula 0
    np.testing.assert_allclose(repaired[mask], want[mask], rtol=1e-6, atol=1e-12)
    assert repaired[0, 1] == repaired[2, 3] == 1.0  # the drift's R12 = R34 = L
    assert -3e-5 < repaired[4, 5] < -2e-5 and want[4, 5] == 0.0


# ------------------------------------------------------------------ gradients
def _oracle_gradients(g, dtype, parameter_beam):
    k1 = torch.tensor(g["k1"], dtype=dtype, requires_grad=True)
    ang = torch.tensor(g["angles"], dtype=dtype, requires_grad=True)
    els = oracle_elements(F.ares_ea(), dtype)
    by_name = {e.name: e for e in els}
    for i, q in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
        by_name[q].p["k1"] = k1[i:i + 1]
    by_name["AREAMCVM1"].p["angle"], by_name["AREAMCHM1"].p["angle"] = ang[0:1], ang[1:2]
    energy = torch.tensor(g["energy"], dtype=dtype)
    if parameter_beam:
        beam = O.ParameterBeamO(torch.tensor(g["pb_mu_in"], dtype=dtype),
                                torch.tensor(g["pb_cov_in"], dtype=dtype), energy)
    else:
        beam = O.ParticleBeamO(torch.tensor(g["particles_in"], dtype=dtype), energy)
    o = O.track(els, beam)
    loss = (o.mu(0) ** 2 + o.sigma(0) ** 2 + o.mu(2) ** 2 + o.sigma(2) ** 2).sum() / 4
    loss.backward()
    return loss.item(), k1.grad.numpy(), ang.grad.numpy()


@pytest.mark.parametrize("kind", ["particles", "parameter"])
def test_gradients_are_the_reference_s(kind):
    """torch autograd through the reference (tests/golden/make_reference_golden.py::gradients)
    against torch autograd through the restatement -- the gradient oracle of the adjoint
    kernels (tests/test_gpu_adjoint.py).  fp32: same operations, same graph -> 1e-5; the fp64
    restatement bounds the fp32 noise of both."""
    g = _load("reference_gradients.npz")
    loss, g_k1, g_ang = _oracle_gradients(g, f32, kind == "parameter")
    np.testing.assert_allclose(loss, g[f"loss_{kind}"], rtol=1e-6)
    np.testing.assert_allclose(g_k1, g[f"g_k1_{kind}"], rtol=1e-5)
    np.testing.assert_allclose(g_ang, g[f"g_angles_{kind}"], rtol=1e-5)
    loss64, g_k1_64, g_ang_64 = _oracle_gradients(g, torch.float64, kind == "parameter")
    np.testing.assert_allclose(g[f"g_k1_{kind}"], g_k1_64, rtol=2e-3)  # fp32 cancellation in sigma^2
    np.testing.assert_allclose(g[f"g_angles_{kind}"], g_ang_64, rtol=1e-4)


def test_split_is_the_reference_s():
    """Element.split / Segment.split of the host mirror against the reference's slices.  The
    corrector is the exception: run on torch, the reference's ``remaining -= resolution``
    (horizontal_corrector.py:84-89) also shrinks ``self.length`` -- the two names alias -- so its
    slice angles come out as angle*l/(what is left); with immutable jnp arrays (the language the
    reference declares) ``-=`` rebinds and every slice gets angle*l/length, which is what
    lynx_b200 does.  The slice LENGTHS agree either way."""
    import lynx_b200 as lx

    g = _load("reference_quirks.npz")
    got = [float(e.length) for e in lx.Drift(length=t(2.0)).split(resolution=torch.tensor(0.3))]
    np.testing.assert_array_equal(np.array(got), g["split_drift"])
    got = [[float(e.length), float(e.k1)] for e in
           lx.Quadrupole(length=t(0.2), k1=t(4.2)).split(resolution=torch.tensor(0.03))]
    np.testing.assert_array_equal(np.array(got), g["split_quadrupole"])
    seg = lx.Segment([lx.Drift(length=t(0.6)), lx.Quadrupole(length=t(0.2), k1=t(0.0)), lx.Marker(),
                      lx.Drift(length=t(0.4))])
    got = [float(e.length) for e in seg.split(resolution=torch.tensor(0.25))]
    np.testing.assert_array_equal(np.array(got), g["split_segment"])
    hcor = lx.HorizontalCorrector(length=t(0.3), angle=t(3e-4))
    slices = hcor.split(resolution=torch.tensor(0.1))
    np.testing.assert_array_equal(np.array([float(e.length) for e in slices]), g["split_hcor"][:, 0])
    np.testing.assert_allclose(g["split_hcor"][:3, 1], [1e-4, 1.5e-4, 3e-4], rtol=1e-6)  # the aliasing
    np.testing.assert_allclose([float(e.angle) for e in slices[:3]], [1e-4] * 3, rtol=1e-6)
    assert float(hcor.length) == np.float32(0.3)  # and the element itself is left intact


# ------------------------------------------------------------------ lattice readers
@pytest.mark.skipif(not os.path.exists("/root/reference/tests/resources/Stage4v3_9.txt"),
                    reason="the reference's resource files are only in the build container")
def test_lattice_readers_are_the_reference_s():
    """lynx_b200's NX-tables and Bmad readers on the reference's resource files against what
    the reference's own readers return for them (reference_converters.npz): every element's
    name, class and length."""
    import lynx_b200 as lx

    g = _load("reference_converters.npz")
    seg = lx.Segment.from_nx_tables("/root/reference/tests/resources/Stage4v3_9.txt")
    assert [e.name for e in seg.elements] == g["nx_names"].tolist()
    assert [type(e).__name__ for e in seg.elements] == g["nx_classes"].tolist()
    np.testing.assert_array_equal(np.array([float(e.length) for e in seg.elements], np.float32), g["nx_lengths"])
    seg = lx.Segment.from_bmad("/root/reference/tests/resources/bmad_tutorial_lattice.bmad")
    assert [e.name for e in seg.elements] == g["bmad_names"].tolist()
    assert [type(e).__name__ for e in seg.elements] == g["bmad_classes"].tolist()
    np.testing.assert_array_equal(np.array([float(e.length) for e in seg.elements], np.float32), g["bmad_lengths"])
    np.testing.assert_array_equal(seg.b.e1.numpy(), g["bmad_b_e1"])
    np.testing.assert_array_equal(seg.q.k1.numpy(), g["bmad_q_k1"])


def test_element_zoo_is_the_reference_s():
    """Every map-bearing element kind with all optional parameters in one line (tilted and
    misaligned quadrupole, solenoid, both correctors, dipole with edges / fringe / tilt, rbend,
    undulator, cavity): 2 001 particles and a parameter beam, bit for bit."""
    g = _load("reference_zoo.npz")
    energy = torch.tensor(g["energy"], dtype=f32)
    els = oracle_elements(F.element_zoo(), f32)
    out = O.track(els, O.ParticleBeamO(torch.tensor(g["particles_in"]), energy))
    np.testing.assert_array_equal(out.energy.numpy(), g["energy_out"])
    np.testing.assert_array_equal(out.particles.numpy(), g["particles_out"])
    els = oracle_elements(F.element_zoo(), f32)
    pb = O.track(els, O.ParameterBeamO(torch.tensor(g["pb_mu_in"]), torch.tensor(g["pb_cov_in"]), energy))
    np.testing.assert_array_equal(pb.energy.numpy(), g["pb_energy_out"])
    np.testing.assert_array_equal(pb.mu_.numpy(), g["pb_mu_out"])
    np.testing.assert_array_equal(pb.cov_.numpy(), g["pb_cov_out"])


def test_screen_geometry_is_the_reference_s():
    """screen.py:86-118: bin edges (what the histogram kernel is handed), extent and the
    effective resolution / pixel size with 4x4 binning -- host mirror and oracle, bit for bit."""
    import lynx_b200 as lx

    g = _load("reference_quirks.npz")
    kw = dict(resolution=torch.tensor((2448, 2040)), pixel_size=torch.tensor((3.3198e-6, 2.4469e-6)),
              binning=torch.tensor(4), misalignment=torch.tensor([[0.0, 0.0]]))
    scr = lx.Screen(**kw)
    ex, ey = scr.pixel_bin_edges
    np.testing.assert_array_equal(ex.numpy(), g["screen_edges_x"])
    np.testing.assert_array_equal(ey.numpy(), g["screen_edges_y"])
    assert ex.numel() == 613 and ey.numel() == 511
    np.testing.assert_array_equal(np.array([float(v) for v in scr.extent], np.float32), g["screen_extent"])
    np.testing.assert_array_equal(scr.effective_resolution.numpy(), g["screen_effective_resolution"])
    np.testing.assert_array_equal(scr.effective_pixel_size.numpy(), g["screen_effective_pixel_size"])
    ox, oy = screen_bin_edges((2448, 2040), (3.3198e-6, 2.4469e-6), 4)
    np.testing.assert_array_equal(ox.numpy(), g["screen_edges_x"])
    np.testing.assert_array_equal(oy.numpy(), g["screen_edges_y"])


def test_batched_particle_settings_are_the_reference_s():
    """tests/test_dipole.py:25-44, tests/test_quadrupole.py:78-99 (config 5 in small): one beam
    broadcast over three settings.  Note the reference takes the tilt sandwich for the whole
    batch as soon as ANY setting is tilted (track_methods.py:100-104); here all three are."""
    g = _load("reference_batched_particles.npz")
    B = 3
    energy = torch.full((B,), float(g["energy"][0]), dtype=f32)
    p_in = torch.tensor(g["particles_in"]).repeat(B, 1, 1)
    zeros = torch.zeros(B)
    dip = El("dipole", "b", {"length": torch.full((B,), 0.5), "angle": torch.tensor(g["dipole_angles"]),
                             "e1": zeros, "e2": zeros, "tilt": zeros, "fringe_integral": zeros,
                             "fringe_integral_exit": zeros, "gap": zeros})
    drift = El("drift", "d", {"length": torch.full((B,), 0.5)})
    out = O.track([dip, drift], O.ParticleBeamO(p_in.clone(), energy))
    np.testing.assert_array_equal(out.particles.numpy(), g["dipole_out"])
    assert np.array_equal(g["dipole_out"][0], g["dipole_out"][2])
    quad = El("quadrupole", "q", {"length": torch.full((B,), 0.5), "k1": torch.ones(B),
                                  "tilt": torch.tensor(g["quad_tilts"]), "misalignment": torch.zeros(B, 2)})
    out = O.track([quad, drift], O.ParticleBeamO(p_in.clone(), energy))
    np.testing.assert_array_equal(out.particles.numpy(), g["quad_out"])
