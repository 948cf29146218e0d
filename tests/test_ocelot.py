"""Ocelot cell converter (lynx_b200/ocelot.py) -- CPU only, Ocelot not installed: stand-in
classes with Ocelot's names, attributes and defaults."""

import json
import os
import sys
import types

import pytest
import torch

import lynx_b200 as lx
from lynx_b200 import ocelot as OC

REF_LATTICE = "/root/reference/tests/resources/ARESlatticeStage3v1_9.py"
REF_JSON = "/root/reference/docs/examples/ARESlatticeStage3v1_9.json"


class _El:
    """Ocelot's ``Element`` constructor surface: every attribute a keyword with default 0."""

    defaults = {}

    def __init__(self, eid=None, **kw):
        self.id = eid
        for k, v in {"l": 0.0, **self.defaults, **kw}.items():
            setattr(self, k, v)


def _kind(name, base=_El, **defaults):
    return type(name, (base,), {"defaults": {**base.defaults, **defaults}})


Drift = _kind("Drift")
Quadrupole = _kind("Quadrupole", k1=0.0, k2=0.0, tilt=0.0)
Solenoid = _kind("Solenoid", k=0.0)
Hcor = _kind("Hcor", angle=0.0)
Vcor = _kind("Vcor", angle=0.0)
Bend = _kind("Bend", angle=0.0, k1=0.0, e1=0.0, e2=0.0, tilt=0.0, gap=0.0, fint=0.0, fintx=0.0)
SBend = _kind("SBend", Bend)
RBend = _kind("RBend", Bend)
Cavity = _kind("Cavity", v=0.0, phi=0.0, freq=0.0)
TDCavity = _kind("TDCavity", v=0.0, phi=0.0, freq=0.0)
Monitor = _kind("Monitor")
Marker = _kind("Marker")
Undulator = _kind("Undulator", lperiod=0.0, nperiods=0, Kx=0.0)
Aperture = _kind("Aperture", xmax=float("inf"), ymax=float("inf"), type="rect")
Sextupole = _kind("Sextupole", k2=0.0)
STAND_INS = dict(Drift=Drift, Quadrupole=Quadrupole, Solenoid=Solenoid, Hcor=Hcor, Vcor=Vcor,
                 Bend=Bend, SBend=SBend, RBend=RBend, Cavity=Cavity, TDCavity=TDCavity,
                 Monitor=Monitor, Marker=Marker, Undulator=Undulator, Aperture=Aperture)


def test_every_rule():
    """converters/ocelot.py:33-185, one element per rule."""
    cell = [
        Drift(l=0.5, eid="D1"),
        Quadrupole(l=0.122, k1=4.2, eid="Q1"),
        Solenoid(l=0.09, k=0.3, eid="S1"),
        Hcor(l=0.02, angle=1e-3, eid="H1"),
        Vcor(l=0.02, angle=-2e-3, eid="V1"),
        SBend(l=0.22, angle=0.1, e1=0.01, e2=0.02, tilt=0.3, fint=0.5, fintx=0.4, gap=0.03, eid="B1"),
        RBend(l=0.3, angle=0.2, e1=0.1, e2=0.1, eid="B2"),
        Cavity(l=4.139, v=0.018, freq=2.998e9, phi=-30.0, eid="C1"),
        TDCavity(l=1.0, v=0.001, freq=11.9952e9, phi=0.0, eid="C2"),
        Monitor(eid="AREABSCR1"),
        Monitor(eid="ARLIBPMG1"),
        Monitor(eid="ARLITORF1"),
        Marker(eid="M1"),
        Undulator(l=0.25, eid="U1"),
        Aperture(xmax=1e-3, ymax=2e-3, type="elip", eid="A1"),
        Sextupole(l=0.1, k2=3.0, eid="X1"),
    ]
    with pytest.warns(UserWarning):
        seg = lx.Segment.from_ocelot(cell, name="cell")
    kinds = [type(e) for e in seg.elements]
    assert kinds == [lx.Drift, lx.Quadrupole, lx.Solenoid, lx.HorizontalCorrector,
                     lx.VerticalCorrector, lx.Dipole, lx.Dipole, lx.Cavity, lx.Cavity, lx.Screen,
                     lx.BPM, lx.Marker, lx.Marker, lx.Undulator, lx.Aperture, lx.Drift]
    assert [e.name for e in seg.elements] == [e.id for e in cell]
    assert torch.equal(seg.Q1.k1, torch.tensor([4.2])) and torch.equal(seg.S1.k, torch.tensor([0.3]))
    assert torch.equal(seg.V1.angle, torch.tensor([-2e-3]))
    b1 = seg.B1
    for attr, want in (("angle", 0.1), ("e1", 0.01), ("e2", 0.02), ("tilt", 0.3),
                       ("fringe_integral", 0.5), ("fringe_integral_exit", 0.4), ("gap", 0.03)):
        assert torch.equal(getattr(b1, attr), torch.tensor([want])), attr
    assert type(seg.B2) is lx.Dipole and torch.equal(seg.B2.e1, torch.tensor([0.1]))  # Bend rule first
    assert torch.allclose(seg.C1.voltage, torch.tensor([0.018e9])) and torch.equal(seg.C1.phase, torch.tensor([-30.0]))
    assert seg.AREABSCR1.resolution.tolist() == [2448, 2040] and not seg.AREABSCR1.is_active
    assert seg.A1.shape == "elliptical" and seg.A1.is_active
    assert torch.equal(seg.X1.length, torch.tensor([0.1]))
    with pytest.warns(UserWarning):
        assert [e.id for e in OC.subcell(cell, "H1", "B1")] == ["H1", "V1", "B1"]
        lx.Segment.from_ocelot(cell[-1:])  # the unknown-kind warning


@pytest.mark.skipif(not (os.path.exists(REF_LATTICE) and os.path.exists(REF_JSON)),
                    reason="reference checkout not present")
def test_ares_cell_converts_to_the_reference_lattice_json(monkeypatch):
    """The reference's Ocelot description of ARES (tests/resources/ARESlatticeStage3v1_9.py,
    executed against the stand-in classes) converts to exactly the elements of the LatticeJSON
    file the reference generated from the same cell (docs/examples/ARESlatticeStage3v1_9.json)."""
    from lynx_b200 import latticejson as LJ

    fake = types.ModuleType("ocelot")
    fake.__dict__.update(STAND_INS)
    monkeypatch.setitem(sys.modules, "ocelot", fake)
    scope = {}
    exec(compile(open(REF_LATTICE).read(), REF_LATTICE, "exec"), scope)
    seg = lx.Segment.from_ocelot(scope["cell"], name="unnamed_element_1", warnings=False)
    want = json.load(open(REF_JSON))
    got = LJ.segment_to_dict(seg)
    assert got["lattices"] == want["lattices"]
    assert list(got["elements"]) == list(want["elements"])
    for key, (cls_name, feats) in want["elements"].items():
        g_cls, g_feats = got["elements"][key]
        assert g_cls == cls_name, key
        if cls_name == "Screen":
            # the file was written by a version whose Screen default misalignment was (1, 2);
            # the reference's code today (screen.py:71-76) and this package use (2,)
            assert feats.pop("misalignment") == [g_feats.pop("misalignment")], key
        assert g_feats == feats, (key, g_feats, feats)


class _ParticleArray:
    """Ocelot's ``ParticleArray`` surface: ``rparticles`` (6, N), ``E`` [GeV], ``q_array`` [C]."""

    def __init__(self, n, seed=3):
        import numpy as np

        rng = np.random.default_rng(seed)
        self.rparticles = rng.normal(0, 1e-4, (6, n)) * np.array([[1.0], [0.02], [1.0], [0.02], [0.05], [10.0]])
        self.E = 0.1073  # GeV
        self.q_array = np.full(n, 5e-18)


def test_parameter_beam_from_ocelot():
    """parameter_beam.py:234-253."""
    import numpy as np

    parray = _ParticleArray(5000)
    beam = lx.ParameterBeam.from_ocelot(parray, device="cpu")
    np.testing.assert_allclose(beam._mu[0, :6].numpy(), parray.rparticles.mean(axis=1), rtol=1e-6, atol=1e-12)
    np.testing.assert_allclose(beam._cov[0, :6, :6].numpy(), np.cov(parray.rparticles), rtol=1e-6, atol=1e-20)
    assert float(beam._mu[0, 6]) == 1.0 and float(beam._cov[0, 6, 6]) == 0.0
    np.testing.assert_allclose(beam.energy.numpy(), [1.073e8], rtol=1e-7)
    np.testing.assert_allclose(torch.as_tensor(beam.total_charge).numpy(), [5000 * 5e-18], rtol=1e-6)
