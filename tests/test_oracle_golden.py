"""Pin the oracle against every known-answer constant the reference holds for the hot
path and against the relational checks of its own tests."""

import numpy as np
import torch

from helpers import aos_from_soa, oracle_elements
from lynx_b200 import fixtures as F
from oracle import El, corrector_map, drift_map, quadrupole_map, segment_map
from oracle.lattice import ParameterBeamO, ParticleBeamO, cavity_track, track


def _f(v, d=torch.float32):
    return torch.tensor([v], dtype=d)


def test_c3_corrector_map():
    """tests/test_vectorized.py:371-383: HorizontalCorrector(0.04, 1e-5) at 150 MeV."""
    tm = corrector_map(_f(0.04), _f(1e-5), _f(150e6), True)[0]
    assert abs(tm[0, 1].item() - 4.0e-02) < 1e-8 and abs(tm[2, 3].item() - 4.0e-02) < 1e-8
    assert abs(tm[1, 6].item() - 1.0e-05) < 1e-11
    assert abs(tm[4, 5].item() / -4.6422e-07 - 1) < 2e-5  # 5 printed digits


def test_c4_drift_r56():
    """docs/examples/optimize_speed.ipynb:287-300."""
    tm = drift_map(_f(0.3), _f(F.ASTRA_ENERGY))[0]
    assert abs(tm[0, 1].item() - 0.3) < 1e-7
    assert abs(tm[4, 5].item() / -6.8021e-06 - 1) < 2e-5


def _c5_lattice(d):
    els = [El("drift", "d", {"length": _f(0.3, d)})]
    z2 = torch.zeros(1, 2, dtype=d)
    q = lambda k: El("quadrupole", "q", {"length": _f(0.1, d), "k1": _f(k, d),  # noqa: E731
                                         "misalignment": z2, "tilt": _f(0.0, d)})
    dr = lambda: El("drift", "d", {"length": _f(0.2, d)})  # noqa: E731
    for _ in range(150):
        els += [q(4.2), dr(), q(-4.2), dr(), El("marker", "m"), q(0.0), dr()]
    return els


def test_c5_merged_1051_elements():
    """docs/examples/optimize_speed.ipynb:270-283.  The notebook's numbers are an fp32
    product of 1051 maps printed to 5 digits; fp64 differs from them by up to 1e-4
    relative, so that is the pin width."""
    golden = {(0, 0): 1.3122, (0, 1): -3.4577, (1, 0): 0.18828, (1, 1): 0.26594,
              (2, 2): 0.30360, (2, 3): -3.2559, (3, 2): 0.18828, (3, 3): 1.2746,
              (4, 5): -3.0678e-03}
    for d in (torch.float32, torch.float64):
        els = _c5_lattice(d)
        assert len(els) == 1051
        tm = segment_map(els, _f(F.ASTRA_ENERGY, d))[0]
        for (i, j), v in golden.items():
            assert abs(tm[i, j].item() / v - 1) < 2e-4, (d, i, j, tm[i, j].item(), v)


def test_c2_cavity_vs_bmad():
    """tests/test_compare_ocelot.py:633-653: twiss after the cavity, float64."""
    d = torch.float64
    sx, sxp, cx = F.twiss_sigmas(5.91253677, 3.55631308, 3.494768647122823e-09)
    sy, syp, cy = F.twiss_sigmas(5.91253677, 3.55631308, 3.497810737006068e-09)
    beam = ParameterBeamO.from_parameters((0, 0, 0, 0), (sx, sxp, sy, syp, 1e-6, 1e-6), 6e6,
                                          cor=(cx, cy, 0.0), dtype=d)
    cav = El("cavity", "c", {"length": _f(1.0377, d), "voltage": _f(0.01815975e9, d),
                             "phase": _f(0.0, d), "frequency": _f(1.3e9, d)})
    out = cavity_track(cav, beam)
    assert np.isclose(out.beta("x").item(), 0.23847352510683092)
    assert np.isclose(out.alpha("x").item(), -1.0160687592932345)
    assert np.isclose(out.beta("y").item(), 0.23847352512430994)
    assert np.isclose(out.alpha("y").item(), -1.0160687593664295)
    assert np.isclose(out.energy.item(), 6e6 + 0.01815975e9)


def test_relations_quadrupole_dipole():
    """tests/test_quadrupole.py:21 (k1=0 quad == drift), test_dipole.py:21 (angle=0
    dipole == drift), test_quadrupole.py:95-98 (tilt pi/4 == 5pi/4 != pi/2)."""
    d = torch.float64
    E = _f(1e8, d)
    z2 = torch.zeros(1, 2, dtype=d)
    dm = drift_map(_f(1.0, d), E)
    qm = quadrupole_map(_f(1.0, d), _f(0.0, d), z2, _f(0.0, d), E)
    assert torch.allclose(qm, dm, atol=1e-11)
    from oracle import dipole_map

    z = _f(0.0, d)
    bm = dipole_map(_f(1.0, d), z, z, z, z, z, z, z, E)
    assert torch.allclose(bm, dm, atol=1e-11)
    t = lambda a: quadrupole_map(_f(0.5, d), _f(3.0, d), z2, _f(a, d), E)  # noqa: E731
    assert torch.allclose(t(np.pi / 4), t(5 * np.pi / 4), atol=1e-12)
    assert not torch.allclose(t(np.pi / 4), t(np.pi / 2), atol=1e-6)


def test_merged_equals_stepwise_and_beam_kinds_agree():
    """tests/test_speed_optimizations.py:6-75 (merged == unmerged) and
    tests/test_compare_beam_type.py:115-139 (Parameter ~ Particle to 1e-2)."""
    d = torch.float64
    spec = F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3)
    els = oracle_elements(spec, d)
    soa = F.gaussian_beam_soa(20000, seed=7)
    pb = ParticleBeamO(aos_from_soa(soa, d), _f(F.ASTRA_ENERGY, d))
    merged = track(els, pb)
    step = pb
    for el in els:
        step = track([el], step)
    assert torch.allclose(merged.particles, step.particles, rtol=1e-9, atol=1e-15)
    qb = ParameterBeamO.from_parameters(F.ASTRA_MU[:4], F.ASTRA_SIGMA, F.ASTRA_ENERGY, dtype=d)
    qo = track(els, qb)
    for i in range(4):
        assert abs(qo.sigma(i).item() / merged.sigma(i).item() - 1) < 2e-2
        # the synthetic beam's sample mean differs from ASTRA_MU by ~sigma/sqrt(N)
        assert abs(qo.mu(i).item() - merged.mu(i).item()) < 5 * merged.sigma(i).item() / 100


def test_screen_edges_and_histogram_semantics():
    """left-closed bins, right-closed last bin, out-of-range dropped,
    image = flipud(hist.T) (screen.py:195-211)."""
    from oracle import screen_bin_edges, screen_reading_particles

    ex, ey = screen_bin_edges((4, 2), (1.0, 1.0))
    assert ex.tolist() == [-2.0, -1.0, 0.0, 1.0, 2.0] and ey.tolist() == [-1.0, 0.0, 1.0]
    xs = torch.tensor([[-2.0, -1.0, 1.999, 2.0, 2.1, -2.1, 0.5]])
    ys = torch.tensor([[-1.0, 0.0, 0.999, 1.0, 0.0, 0.0, -0.5]])
    img = screen_reading_particles(xs, ys, ex, ey)
    assert img.shape == (1, 2, 4)
    # row 0 = upper y bin [0,1]; row 1 = lower y bin [-1,0)
    expect = torch.tensor([[[0.0, 1.0, 0.0, 2.0], [1.0, 0.0, 1.0, 0.0]]])
    assert torch.equal(img, expect)
