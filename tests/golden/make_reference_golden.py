"""Golden vectors computed BY THE REFERENCE ITSELF (jank324/lynx in /root/reference), run on
torch through tests/golden/reference_shim.py (read that file's header for what the shim is).

    python tests/golden/make_reference_golden.py      # writes tests/golden/reference_*.npz

Inputs are deterministic (numpy seeds below) and stored next to the outputs, so the tests need
neither the reference nor this script.  Everything is fp32, as the reference computes.
"""

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "tests"), HERE):
    if p not in sys.path:
        sys.path.insert(0, p)

import reference_shim  # noqa: E402
from lynx_b200 import fixtures as F  # noqa: E402

lynx = reference_shim.install()
SMALL_SCREEN = {"resolution": (64, 48), "pixel_size": (4e-5, 4e-5), "binning": 1}


def t(v):
    return torch.tensor([v], dtype=torch.float32)


def build(spec):
    """Reference ``Segment`` from a lynx_b200.fixtures spec, through the reference's own
    constructors."""
    out = []
    for kind, name, p in spec:
        if kind == "marker":
            el = lynx.Marker(name=name)
        elif kind == "drift":
            el = lynx.Drift(length=t(p["length"]), name=name)
        elif kind == "quadrupole":
            kw = {}
            if "misalignment" in p:
                kw["misalignment"] = torch.tensor([p["misalignment"]], dtype=torch.float32)
            if "tilt" in p:
                kw["tilt"] = t(p["tilt"])
            el = lynx.Quadrupole(length=t(p["length"]), k1=t(p.get("k1", 0.0)), name=name, **kw)
        elif kind in ("hcor", "vcor"):
            cls = lynx.HorizontalCorrector if kind == "hcor" else lynx.VerticalCorrector
            el = cls(length=t(p["length"]), angle=t(p.get("angle", 0.0)), name=name)
        elif kind in ("dipole", "rbend"):
            cls = lynx.Dipole if kind == "dipole" else lynx.RBend
            el = cls(name=name, **{k: t(v) for k, v in p.items()})
        elif kind == "cavity":
            el = lynx.Cavity(length=t(p["length"]), voltage=t(p["voltage"]), phase=t(p["phase"]),
                             frequency=t(p["frequency"]), name=name)
        elif kind == "solenoid":
            el = lynx.Solenoid(length=t(p["length"]), k=t(p["k"]), name=name)
        elif kind == "undulator":
            el = lynx.Undulator(length=t(p["length"]), name=name)
        elif kind == "bpm":
            el = lynx.BPM(is_active=p.get("is_active", False), name=name)
        elif kind == "screen":
            el = lynx.Screen(resolution=torch.tensor(p["resolution"]),
                             pixel_size=torch.tensor(p["pixel_size"], dtype=torch.float32),
                             binning=torch.tensor(p.get("binning", 1)),
                             misalignment=torch.tensor([p.get("misalignment", (0.0, 0.0))],
                                                       dtype=torch.float32),
                             is_active=p.get("is_active", False), name=name)
        else:
            raise ValueError(kind)
        out.append(el)
    return lynx.Segment(out)


def seeded_particles(n, seed):
    """(1, n, 7) fp32 particles with the moments of the ASTRA beam the reference's tests use."""
    rng = np.random.default_rng(seed)
    a = np.ones((1, n, 7), dtype=np.float32)
    a[0, :, :6] = (rng.standard_normal((n, 6)) * np.array(F.ASTRA_SIGMA) + np.array(F.ASTRA_MU)).astype(np.float32)
    return torch.from_numpy(a)


def element_maps():
    """Every element's transfer_map at several energies / parameter sets (8a rows a1-a8)."""
    E = torch.tensor([1.0732e8, 6.0e6, 1.0e9], dtype=torch.float32)
    out = {"energy": E.numpy()}
    cases = {
        "drift": lambda e: lynx.Drift(length=t(0.7)).transfer_map(e),
        "quadrupole": lambda e: lynx.Quadrupole(length=t(0.122), k1=t(7.3)).transfer_map(e),
        "quadrupole_defocus": lambda e: lynx.Quadrupole(length=t(0.122), k1=t(-11.0)).transfer_map(e),
        "quadrupole_tilt_mis": lambda e: lynx.Quadrupole(
            length=t(0.2), k1=t(4.2), tilt=t(0.3),
            misalignment=torch.tensor([[1e-3, -2e-3]])).transfer_map(e),
        "hcor": lambda e: lynx.HorizontalCorrector(length=t(0.02), angle=t(1.5e-3)).transfer_map(e),
        "vcor": lambda e: lynx.VerticalCorrector(length=t(0.02), angle=t(-2.5e-3)).transfer_map(e),
        "dipole": lambda e: lynx.Dipole(length=t(0.5), angle=t(0.1), e1=t(0.05), e2=t(0.02),
                                        fringe_integral=t(0.5), gap=t(0.03)).transfer_map(e),
        "dipole_tilt": lambda e: lynx.Dipole(length=t(0.3), angle=t(-0.2), tilt=t(0.4),
                                             fringe_integral=t(0.4), fringe_integral_exit=t(0.3),
                                             gap=t(0.02)).transfer_map(e),
        "rbend": lambda e: lynx.RBend(length=t(0.5), angle=t(0.1), fringe_integral=t(0.5),
                                      gap=t(0.03)).transfer_map(e),
        "solenoid": lambda e: lynx.Solenoid(length=t(0.5), k=t(0.3)).transfer_map(e),
        "undulator": lambda e: lynx.Undulator(length=t(0.25)).transfer_map(e),
        "cavity": lambda e: lynx.Cavity(length=t(1.0377), voltage=t(1.0e7), phase=t(30.0),
                                        frequency=t(1.3e9)).transfer_map(e),
        "cavity_decel": lambda e: lynx.Cavity(length=t(4.139), voltage=t(2.0e6), phase=t(160.0),
                                              frequency=t(2.998e9)).transfer_map(e),
    }
    for name, fn in cases.items():
        out[name] = np.stack([fn(e.reshape(1))[0].numpy() for e in E])
    np.savez_compressed(os.path.join(HERE, "reference_element_maps.npz"), **out)
    return out


def ares_tracking():
    """ARES EA (C6): merged map, particles, parameter beam, BPM-free screen reading."""
    k1, vk, hk = (5.0, -5.0, 5.0), 1e-3, -2e-3
    energy = t(F.ASTRA_ENERGY)
    particles = seeded_particles(4001, seed=11)
    out = {"k1": np.array(k1, np.float32), "vkick": np.float32(vk), "hkick": np.float32(hk),
           "energy": energy.numpy(), "particles_in": particles.numpy()}
    seg = build(F.ares_ea(k1=k1, vkick=vk, hkick=hk))
    out["segment_map"] = seg.transfer_map(energy)[0].numpy()
    beam = lynx.ParticleBeam(particles.clone(), energy.clone())
    tracked = seg.track(beam)
    out["particles_out"] = tracked.particles.numpy()
    for name in ("mu_x", "mu_y", "sigma_x", "sigma_xp", "sigma_y", "sigma_yp", "sigma_s",
                 "sigma_p", "sigma_xxp", "sigma_yyp"):
        out["out_" + name] = getattr(tracked, name).numpy()
    # parameter beam with the same first and second moments as the particle set
    pb = lynx.ParameterBeam.from_parameters(
        mu_x=t(F.ASTRA_MU[0]), mu_xp=t(F.ASTRA_MU[1]), mu_y=t(F.ASTRA_MU[2]), mu_yp=t(F.ASTRA_MU[3]),
        sigma_x=t(F.ASTRA_SIGMA[0]), sigma_xp=t(F.ASTRA_SIGMA[1]), sigma_y=t(F.ASTRA_SIGMA[2]),
        sigma_yp=t(F.ASTRA_SIGMA[3]), sigma_s=t(F.ASTRA_SIGMA[4]), sigma_p=t(F.ASTRA_SIGMA[5]),
        energy=energy.clone())
    out["pb_mu_in"], out["pb_cov_in"] = pb._mu.numpy(), pb._cov.numpy()
    pb_out = seg.track(pb)
    out["pb_mu_out"], out["pb_cov_out"] = pb_out._mu.numpy(), pb_out._cov.numpy()
    # the screen, with a coarse camera so that the image stays small
    seg = build(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=SMALL_SCREEN))
    empty = seg.track(lynx.ParticleBeam(particles.clone(), energy.clone()))
    assert empty is lynx.ParticleBeam.empty or empty is lynx.particles.Beam.empty
    out["screen_reading"] = seg.AREABSCR1.reading.numpy()
    out["screen_resolution"] = np.array(SMALL_SCREEN["resolution"])
    out["screen_pixel_size"] = np.array(SMALL_SCREEN["pixel_size"], np.float32)
    np.savez_compressed(os.path.join(HERE, "reference_ares_ea.npz"), **out)
    return out


def cavity_line():
    """Cavity + dipole + quadrupoles (config 4 lattice): the non-linear cavity pass."""
    energy = t(6.0e6)
    particles = seeded_particles(2001, seed=12)
    seg = build(F.cavity_dipole_line(screen_active=False))
    beam = lynx.ParticleBeam(particles.clone(), energy.clone())
    tracked = seg.track(beam)
    out = {"energy": energy.numpy(), "particles_in": particles.numpy(),
           "particles_out": tracked.particles.numpy(), "energy_out": tracked.energy.numpy()}
    np.savez_compressed(os.path.join(HERE, "reference_cavity_line.npz"), **out)
    return out


def zoo():
    """Every map-bearing element kind with all its optional parameters, particles and a
    parameter beam (lynx_b200.fixtures.element_zoo)."""
    energy = t(2.0e7)
    particles = seeded_particles(2001, seed=19)
    seg = build(F.element_zoo())
    tracked = seg.track(lynx.ParticleBeam(particles.clone(), energy.clone()))
    out = {"energy": energy.numpy(), "particles_in": particles.numpy(),
           "particles_out": tracked.particles.numpy(), "energy_out": tracked.energy.numpy()}
    pb = lynx.ParameterBeam.from_parameters(
        mu_x=t(1e-4), mu_y=t(-2e-4), sigma_x=t(2e-4), sigma_xp=t(1e-5), sigma_y=t(1.5e-4),
        sigma_yp=t(8e-6), sigma_s=t(1e-5), sigma_p=t(2e-3), energy=energy.clone())
    out["pb_mu_in"], out["pb_cov_in"] = pb._mu.numpy(), pb._cov.numpy()
    res = build(F.element_zoo()).track(pb)
    out["pb_mu_out"], out["pb_cov_out"], out["pb_energy_out"] = res._mu.numpy(), res._cov.numpy(), res.energy.numpy()
    np.savez_compressed(os.path.join(HERE, "reference_zoo.npz"), **out)
    return out


def batched_particles():
    """tests/test_dipole.py:25-44 and tests/test_quadrupole.py:78-99 with seeded particles: one
    beam broadcast over three settings of a dipole / of a tilted quadrupole (config 5 in small)."""
    energy = t(1e9)
    particles = seeded_particles(1001, seed=20)
    out = {"energy": energy.numpy(), "particles_in": particles.numpy()}
    beam = lynx.ParticleBeam(particles.clone(), energy.clone()).broadcast((3,))
    seg = lynx.Segment([
        lynx.Dipole(length=torch.tensor([0.5, 0.5, 0.5]), angle=torch.tensor([0.1, 0.2, 0.1])),
        lynx.Drift(length=t(0.5)).broadcast((3,)),
    ])
    out["dipole_angles"] = np.array([0.1, 0.2, 0.1], np.float32)
    out["dipole_out"] = seg.track(beam).particles.numpy()
    tilts = torch.tensor([np.pi / 4, np.pi / 2, np.pi * 5 / 4], dtype=torch.float32)
    seg = lynx.Segment([
        lynx.Quadrupole(length=torch.tensor([0.5, 0.5, 0.5]), k1=torch.tensor([1.0, 1.0, 1.0]), tilt=tilts),
        lynx.Drift(length=t(0.5)).broadcast((3,)),
    ])
    beam = lynx.ParticleBeam(particles.clone(), energy.clone()).broadcast((3,))
    out["quad_tilts"] = tilts.numpy()
    out["quad_out"] = seg.track(beam).particles.numpy()
    np.savez_compressed(os.path.join(HERE, "reference_batched_particles.npz"), **out)
    return out


def bpm_line():
    """SMATCH-like line with active BPMs: readings (bpm.py:48-58)."""
    energy = t(F.ASTRA_ENERGY)
    particles = seeded_particles(2001, seed=13)
    seg = build(F.smatch(bpms_active=True, angles=(1e-3, -5e-4, 2e-4)))
    tracked = seg.track(lynx.ParticleBeam(particles.clone(), energy.clone()))
    out = {"energy": energy.numpy(), "particles_in": particles.numpy(),
           "particles_out": tracked.particles.numpy()}
    names = [e.name for e in seg.elements if isinstance(e, lynx.BPM)]
    out["bpm_names"] = np.array(names)
    out["bpm_readings"] = np.stack([getattr(seg, n).reading.numpy() for n in names])
    np.savez_compressed(os.path.join(HERE, "reference_bpm_line.npz"), **out)
    return out


def beam_ops():
    """transformed_to (particle_beam.py:580-715), Aperture.track (aperture.py:69-109) and the
    ASTRA reader (converters/astra.py:8-62)."""
    import tempfile

    energy = t(F.ASTRA_ENERGY)
    particles = seeded_particles(1501, seed=14)
    out = {"energy": energy.numpy(), "particles_in": particles.numpy()}
    beam = lynx.ParticleBeam(particles.clone(), energy.clone(),
                             particle_charges=torch.full((1, 1501), 1e-15))
    target = dict(mu_x=t(1e-4), mu_y=t(-2e-4), mu_xp=t(1e-6), mu_yp=t(-3e-6), sigma_x=t(3e-4),
                  sigma_y=t(1e-4), sigma_xp=t(5e-6), sigma_yp=t(2e-6), sigma_s=t(1e-5), sigma_p=t(1e-3))
    moved = beam.transformed_to(**target)
    out["transform_target"] = np.array([float(v) for v in target.values()], np.float32)
    out["transform_target_names"] = np.array(list(target))
    out["transformed"] = moved.particles.numpy()
    for shape in ("rectangular", "elliptical"):
        ap = lynx.Aperture(x_max=t(1.5e-4), y_max=t(2.0e-4), shape=shape, is_active=True)
        kept = ap.track(lynx.ParticleBeam(particles.clone(), energy.clone(),
                                          particle_charges=torch.full((1, 1501), 1e-15)))
        out[f"aperture_{shape}_kept"] = kept.particles.numpy()
        out[f"aperture_{shape}_lost"] = ap.lost_particles.numpy()
    out["aperture_x_max"], out["aperture_y_max"] = np.float32(1.5e-4), np.float32(2.0e-4)
    # ASTRA: a synthetic distribution file (the reference's sample is not in its checkout)
    from lynx.converters.astra import from_astrabeam

    rng = np.random.default_rng(15)
    n = 501
    rows = np.zeros((n, 10))
    rows[:, 0:2] = rng.normal(0, 1.7e-4, (n, 2))
    rows[:, 2] = rng.normal(0, 8e-6, n)
    rows[:, 3:5] = rng.normal(0, 4e2, (n, 2))
    rows[:, 5] = rng.normal(0, 2e5, n)
    rows[0, :6] = (0, 0, 2.0, 0, 0, 1.0732e8)
    rows[:, 7], rows[:, 8], rows[:, 9] = -2.5e-7, 1, 5
    rows[[17, 400], 9] = (-3, -1)  # two lost particles
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "beam.ini")
        np.savetxt(path, rows, fmt="%.12e")
        rows = np.loadtxt(path)  # what the file holds, to the printed precision
        coords, e_ref, charges = from_astrabeam(path)
    out["astra_rows"], out["astra_coords"] = rows, coords
    out["astra_energy"], out["astra_charges"] = np.float64(e_ref), charges
    np.savez_compressed(os.path.join(HERE, "reference_beam_ops.npz"), **out)
    return out


def quirks():
    """Behaviour the restatement reproduces on purpose: B1 -- a screen's y
    misalignment is subtracted from column 1 (xp), screen.py:134-135; B6 -- a cavity with zero
    voltage yields NaN, cavity.py:269; plus make_linspaced (particle_beam.py:385-470)."""
    k1, vk, hk = (5.0, -5.0, 5.0), 1e-3, -2e-3
    energy = t(F.ASTRA_ENERGY)
    particles = seeded_particles(2001, seed=17)
    screen = dict(SMALL_SCREEN, misalignment=(2e-4, -1.5e-4))
    seg = build(F.ares_ea(k1=k1, vkick=vk, hkick=hk, screen_active=True, screen=screen))
    seg.track(lynx.ParticleBeam(particles.clone(), energy.clone()))
    out = {"energy": energy.numpy(), "particles_in": particles.numpy(),
           "k1": np.array(k1, np.float32), "vkick": np.float32(vk), "hkick": np.float32(hk),
           "screen_misalignment": np.array(screen["misalignment"], np.float32),
           "screen_reading_misaligned": seg.AREABSCR1.reading.numpy()}
    cav = lynx.Cavity(length=t(1.0), voltage=t(0.0), phase=t(0.0), frequency=t(1.3e9))
    out["cavity_v0_map"] = cav.transfer_map(t(1e8))[0].numpy()
    lin = lynx.ParticleBeam.make_linspaced(
        num_particles=11, mu_x=t(1e-4), mu_y=t(-1e-4), sigma_x=t(2e-4), sigma_y=t(1e-4),
        sigma_xp=t(1e-5), sigma_yp=t(2e-5), sigma_s=t(1e-6), sigma_p=t(1e-3), energy=t(1e8))
    out["linspaced"] = lin.particles.numpy()
    # Screen geometry (screen.py:86-118) with 4x4 binning
    scr = lynx.Screen(resolution=torch.tensor((2448, 2040)), pixel_size=torch.tensor((3.3198e-6, 2.4469e-6)),
                      binning=torch.tensor(4), misalignment=torch.tensor([[0.0, 0.0]]))
    out["screen_edges_x"], out["screen_edges_y"] = (e.numpy() for e in scr.pixel_bin_edges)
    out["screen_extent"] = np.array([float(v) for v in scr.extent], np.float32)
    out["screen_effective_resolution"] = scr.effective_resolution.numpy()
    out["screen_effective_pixel_size"] = scr.effective_pixel_size.numpy()
    # split (drift.py:71-78, quadrupole.py:99-110, segment.py:310-315, horizontal_corrector.py:82-90)
    res = torch.tensor(0.3)
    out["split_drift"] = np.array([float(e.length) for e in lynx.Drift(length=t(2.0)).split(resolution=res)])
    out["split_quadrupole"] = np.array([[float(e.length), float(e.k1)] for e in lynx.Quadrupole(
        length=t(0.2), k1=t(4.2)).split(resolution=torch.tensor(0.03))])
    seg = lynx.Segment([lynx.Drift(length=t(0.6)), lynx.Quadrupole(length=t(0.2), k1=t(0.0)),
                        lynx.Marker(), lynx.Drift(length=t(0.4))])
    out["split_segment"] = np.array([float(e.length) for e in seg.split(resolution=torch.tensor(0.25))])
    # under torch, ``remaining = self.length; remaining -= resolution`` shrinks the corrector's
    # own length while it is being split (an aliasing that immutable jnp arrays do not have)
    out["split_hcor"] = np.array([[float(e.length), float(e.angle)] for e in lynx.HorizontalCorrector(
        length=t(0.3), angle=t(3e-4)).split(resolution=torch.tensor(0.1))])
    np.savez_compressed(os.path.join(HERE, "reference_quirks.npz"), **out)
    return out


def gradients():
    """torch autograd THROUGH THE REFERENCE (it runs on torch here): d loss / d k1, d angle for a
    particle beam and a parameter beam through ARES EA; loss as tests/test_differentiable.py
    builds them from mu / sigma."""
    energy = t(F.ASTRA_ENERGY)
    particles = seeded_particles(2001, seed=18)
    out = {"energy": energy.numpy(), "particles_in": particles.numpy()}
    k1_0, ang_0 = [5.0, -5.0, 5.0], [1e-3, -2e-3]

    def lattice():
        seg = build(F.ares_ea())
        k1 = torch.tensor(k1_0, requires_grad=True)
        ang = torch.tensor(ang_0, requires_grad=True)
        for i, q in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
            getattr(seg, q).k1 = k1[i:i + 1]
        seg.AREAMCVM1.angle, seg.AREAMCHM1.angle = ang[0:1], ang[1:2]
        return seg, k1, ang

    seg, k1, ang = lattice()
    o = seg.track(lynx.ParticleBeam(particles.clone(), energy.clone()))
    loss = (o.mu_x ** 2 + o.sigma_x ** 2 + o.mu_y ** 2 + o.sigma_y ** 2).sum() / 4
    loss.backward()
    out.update(k1=np.array(k1_0, np.float32), angles=np.array(ang_0, np.float32),
               loss_particles=np.float32(loss.item()), g_k1_particles=k1.grad.numpy(),
               g_angles_particles=ang.grad.numpy())
    seg, k1, ang = lattice()
    pb = lynx.ParameterBeam.from_parameters(
        mu_x=t(F.ASTRA_MU[0]), mu_xp=t(F.ASTRA_MU[1]), mu_y=t(F.ASTRA_MU[2]), mu_yp=t(F.ASTRA_MU[3]),
        sigma_x=t(F.ASTRA_SIGMA[0]), sigma_xp=t(F.ASTRA_SIGMA[1]), sigma_y=t(F.ASTRA_SIGMA[2]),
        sigma_yp=t(F.ASTRA_SIGMA[3]), sigma_s=t(F.ASTRA_SIGMA[4]), sigma_p=t(F.ASTRA_SIGMA[5]),
        energy=energy.clone())
    out["pb_mu_in"], out["pb_cov_in"] = pb._mu.numpy(), pb._cov.numpy()
    o = seg.track(pb)
    loss = (o.mu_x ** 2 + o.sigma_x ** 2 + o.mu_y ** 2 + o.sigma_y ** 2).sum() / 4
    loss.backward()
    out.update(loss_parameter=np.float32(loss.item()), g_k1_parameter=k1.grad.numpy(),
               g_angles_parameter=ang.grad.numpy())
    np.savez_compressed(os.path.join(HERE, "reference_gradients.npz"), **out)
    return out


def converters():
    """The reference's own lattice readers on its own resource files: element names, classes
    and lengths (converters/nxtables.py, converters/bmad.py)."""
    from pathlib import Path

    out = {}
    seg = lynx.Segment.from_nx_tables(Path("/root/reference/tests/resources/Stage4v3_9.txt"))
    out["nx_names"] = np.array([e.name for e in seg.elements])
    out["nx_classes"] = np.array([type(e).__name__ for e in seg.elements])
    out["nx_lengths"] = np.array([float(e.length) for e in seg.elements], np.float32)
    # (Segment.length itself raises here: the screens' default length has shape (), the other
    #  elements' (1,), and segment.py:323 stacks them)
    seg = lynx.Segment.from_bmad("/root/reference/tests/resources/bmad_tutorial_lattice.bmad")
    out["bmad_names"] = np.array([e.name for e in seg.elements])
    out["bmad_classes"] = np.array([type(e).__name__ for e in seg.elements])
    out["bmad_lengths"] = np.array([float(e.length) for e in seg.elements], np.float32)
    out["bmad_b_e1"], out["bmad_q_k1"] = seg.b.e1.numpy(), seg.q.k1.numpy()
    np.savez_compressed(os.path.join(HERE, "reference_converters.npz"), **out)
    return out


def parameter_beam():
    """ParameterBeam constructors / accessors (parameter_beam.py:47-433, beam.py:241-310), its
    Screen.reading (screen.py:157-194) and a batch of settings through ARES EA (config 3)."""
    out = {}
    args = dict(mu_x=t(1e-4), mu_xp=t(-2e-6), mu_y=t(-3e-4), mu_yp=t(1e-6), sigma_x=t(2e-4),
                sigma_xp=t(1e-5), sigma_y=t(1.5e-4), sigma_yp=t(8e-6), sigma_s=t(1e-5),
                sigma_p=t(2e-3), cor_x=t(1e-9), cor_y=t(-5e-10), cor_s=t(2e-9), energy=t(1.2e8),
                total_charge=t(1e-12))
    pb = lynx.ParameterBeam.from_parameters(**args)
    out["fp_names"] = np.array(list(args))
    out["fp_values"] = np.array([float(v) for v in args.values()], np.float64)
    out["fp_mu"], out["fp_cov"] = pb._mu.numpy(), pb._cov.numpy()
    accessors = ("mu_x", "mu_xp", "mu_y", "mu_yp", "mu_s", "mu_p", "sigma_x", "sigma_xp", "sigma_y",
                 "sigma_yp", "sigma_s", "sigma_p", "sigma_xxp", "sigma_yyp", "emittance_x",
                 "emittance_y", "normalized_emittance_x", "normalized_emittance_y", "beta_x",
                 "alpha_x", "beta_y", "alpha_y", "relativistic_gamma", "relativistic_beta")
    out["accessor_names"] = np.array(accessors)
    out["fp_accessors"] = np.array([float(getattr(pb, a)) for a in accessors], np.float64)
    tw = dict(beta_x=t(5.0), alpha_x=t(-0.5), emittance_x=t(2e-8), beta_y=t(3.0), alpha_y=t(0.2),
              emittance_y=t(1e-8), energy=t(1e8))
    ptw = lynx.ParameterBeam.from_twiss(**tw)
    out["tw_names"] = np.array(list(tw))
    out["tw_values"] = np.array([float(v) for v in tw.values()], np.float64)
    out["tw_mu"], out["tw_cov"] = ptw._mu.numpy(), ptw._cov.numpy()
    out["tw_accessors"] = np.array([float(getattr(ptw, a)) for a in accessors], np.float64)
    moved = pb.transformed_to(mu_x=t(1e-3), sigma_x=t(3e-4), sigma_yp=t(2e-6), energy=t(9e7))
    out["tr_mu"], out["tr_cov"], out["tr_energy"] = moved._mu.numpy(), moved._cov.numpy(), moved.energy.numpy()
    # Screen.reading of a ParameterBeam: pdf on the pixel grid
    scr = lynx.Screen(resolution=torch.tensor(SMALL_SCREEN["resolution"]),
                      pixel_size=torch.tensor(SMALL_SCREEN["pixel_size"]), binning=torch.tensor(1),
                      misalignment=torch.tensor([[1e-4, -5e-5]]), is_active=True, name="scr")
    seg = lynx.Segment([lynx.Drift(length=t(1.0)), scr])
    seg.track(lynx.ParameterBeam.from_parameters(**args))
    out["screen_pdf"] = seg.scr.reading.numpy()
    # four settings of the three quadrupoles and two correctors through ARES EA
    B = 4
    rng = np.random.default_rng(16)
    k1 = rng.uniform(-20, 20, (B, 3)).astype(np.float32)
    ang = rng.uniform(-3e-3, 3e-3, (B, 2)).astype(np.float32)
    seg = build(F.ares_ea()).broadcast((B,))
    for i, q in enumerate(("AREAMQZM1", "AREAMQZM2", "AREAMQZM3")):
        getattr(seg, q).k1 = torch.from_numpy(k1[:, i].copy())
    seg.AREAMCVM1.angle = torch.from_numpy(ang[:, 0].copy())
    seg.AREAMCHM1.angle = torch.from_numpy(ang[:, 1].copy())
    pbb = lynx.ParameterBeam.from_parameters(**{k: v for k, v in args.items()}).broadcast((B,))
    res = seg.track(pbb)
    out["batch_k1"], out["batch_angles"] = k1, ang
    out["batch_mu"], out["batch_cov"] = res._mu.numpy(), res._cov.numpy()
    np.savez_compressed(os.path.join(HERE, "reference_parameter_beam.npz"), **out)
    return out


if __name__ == "__main__":
    torch.manual_seed(0)
    for fn in (element_maps, zoo, batched_particles, ares_tracking, cavity_line, bpm_line, beam_ops, parameter_beam, quirks, gradients, converters):
        res = fn()
        print(fn.__name__, {k: getattr(v, "shape", v) for k, v in res.items()})
