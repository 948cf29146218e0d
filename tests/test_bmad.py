"""Bmad lattice reader (lynx_b200/bmad.py) -- CPU only."""

import math
import os

import pytest
import torch

import lynx_b200 as lx
from lynx_b200 import bmad

REF_FILE = "/root/reference/tests/resources/bmad_tutorial_lattice.bmad"

TUTORIAL = """! Lattice file: simple.bmad
beginning[beta_a] = 10. ! m a-mode beta function
beginning[e_tot] = 10e6 ! eV
parameter[geometry] = open  ! Or closed

d: drift, L = 0.5
b: sbend, L = 0.5, g = 1, e1 = 0.1, dg = 0.001 ! g = 1/design_radius
q: quadrupole, L = 0.6, k1 = 0.23

lat: line = (d, b, q) ! List of lattice elements
use, lat ! Line used to construct the lattice
"""


def _check_tutorial(converted):
    """tests/test_bmad_conversion.py:6-33."""
    assert [e.name for e in converted.elements] == ["d", "b", "q"]
    assert isinstance(converted.d, lx.Drift) and converted.d.length == torch.tensor([0.5])
    assert isinstance(converted.b, lx.Dipole) and converted.b.length == torch.tensor([0.5])
    assert converted.b.e1 == torch.tensor([0.1])
    assert isinstance(converted.q, lx.Quadrupole) and converted.q.length == torch.tensor([0.6])
    assert converted.q.k1 == torch.tensor([0.23])


def test_bmad_tutorial(tmp_path):
    path = tmp_path / "simple.bmad"
    path.write_text(TUTORIAL)
    _check_tutorial(lx.Segment.from_bmad(str(path)))


@pytest.mark.skipif(not os.path.exists(REF_FILE), reason="reference checkout not present")
def test_bmad_tutorial_reference_resource():
    _check_tutorial(lx.Segment.from_bmad(REF_FILE))


def test_language_features(tmp_path, monkeypatch):
    """Variables and arithmetic, ``elem[attr]`` references, inheritance, wildcard attribute
    assignment, continuation lines, ``call`` with an environment variable in the path, nested
    lines with repetition and reversal, an unknown kind falling back to a drift."""
    monkeypatch.setenv("LYNX_TEST_LAT", str(tmp_path))
    (tmp_path / "magnets.bmad").write_text(
        "QF: Quadrupole, L = lq, k1 = kq      ! upper case is folded\n"
        "qd: qf, k1 = -qf[k1] * 2^2 / 4\n"
        "cav: lcavity, l = 1.0, rf_frequency = 2.856e9, voltage = 20e6, &\n"
        "     phi0 = 30 / 360\n")
    (tmp_path / "main.bmad").write_text(
        "lq = 0.1 + 2 * 0.05\n"
        "kq = sqrt(16) * cos(0) ! 4\n"
        "call, file = $LYNX_TEST_LAT/magnets.bmad\n"
        "d1: drift, l = lq / 2\n"
        "hk: hkicker, l = 0.05, kick = 1e-3\n"
        "sol: solenoid, l = 0.3, ks = 0.2\n"
        "col: ecollimator, l = 0, x_limit = 2e-3,\n"
        "     y_limit = 1e-3\n"
        "oct: octupole, l = 0.25\n"
        "m1: marker\n"
        "q*[tilt] = pi / 4\n"
        "cell: line = (qf, d1, qd)\n"
        "knob: overlay = {qf[k1]: 0.1*x}, var = {x}\n"
        "ring: line = (m1, 2*cell, -cell, hk, sol, cav, col, oct)\n"
        "use, ring\n")
    with pytest.warns(UserWarning, match="octupole"):
        seg = lx.Segment.from_bmad(str(tmp_path / "main.bmad"))
    assert seg.name == "ring"
    assert [e.name for e in seg.elements] == ["m1", "cell", "cell", "cell", "hk", "sol", "cav", "col", "oct"]
    cell, reversed_cell = seg.elements[1], seg.elements[3]
    assert [e.name for e in cell.elements] == ["qf", "d1", "qd"]
    assert [e.name for e in reversed_cell.elements] == ["qd", "d1", "qf"]
    qf, d1, qd = cell.elements
    assert torch.allclose(qf.length, torch.tensor([0.2])) and torch.allclose(qf.k1, torch.tensor([4.0]))
    assert torch.allclose(qd.length, torch.tensor([0.2])) and torch.allclose(qd.k1, torch.tensor([-4.0]))
    assert torch.allclose(qf.tilt, torch.tensor([math.pi / 4])) and torch.allclose(qd.tilt, qf.tilt)
    assert torch.allclose(d1.length, torch.tensor([0.1]))
    flat = {e.name: e for e in seg.flattened().elements}
    assert isinstance(flat["hk"], lx.HorizontalCorrector) and torch.allclose(flat["hk"].angle, torch.tensor([1e-3]))
    assert isinstance(flat["sol"], lx.Solenoid) and torch.allclose(flat["sol"].k, torch.tensor([0.2]))
    cav = flat["cav"]
    assert isinstance(cav, lx.Cavity) and torch.allclose(cav.phase, torch.tensor([-30.0]))
    assert torch.allclose(cav.voltage, torch.tensor([20e6])) and torch.allclose(cav.frequency, torch.tensor([2.856e9]))
    col = flat["col"]
    assert isinstance(col, lx.Aperture) and col.shape == "elliptical"
    assert torch.allclose(col.x_max, torch.tensor([2e-3])) and torch.allclose(col.y_max, torch.tensor([1e-3]))
    assert isinstance(flat["oct"], lx.Drift) and torch.allclose(flat["oct"].length, torch.tensor([0.25]))
    assert isinstance(flat["m1"], lx.Marker)
    assert torch.allclose(seg.length, torch.tensor([3 * 0.5 + 0.05 + 0.3 + 1.0 + 0.25]))


def test_expression_evaluator_is_closed():
    names = {"a": 2.0, "q": {"element_type": "quadrupole", "k1": 3.0}}
    assert bmad.evaluate("a * q[k1] ^ 2 - 1", names) == 17.0
    assert bmad.evaluate("-abs(-a) + twopi / pi", names) == 0.0
    assert bmad.evaluate("open", names) == "open"
    assert bmad.evaluate('"some text"', names) == "some text"
    # no attribute access, no calls outside the table, no names outside the namespace
    for text in ("__import__('os').getcwd()", "a.real", "print(1)", "undefined_name + 1"):
        assert bmad.evaluate(text, names) == text


def test_unknown_property_is_not_dropped_silently(tmp_path):
    path = tmp_path / "bad.bmad"
    path.write_text("q: quadrupole, l = 0.1, k1 = 1, b1_gradient = 3\nl1: line = (q)\nuse, l1\n")
    with pytest.raises(AssertionError, match="b1_gradient"):
        lx.Segment.from_bmad(str(path))
