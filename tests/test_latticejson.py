"""LatticeJSON load / save (lynx_b200/latticejson.py) -- CPU only.

The reference's example file docs/examples/ARESlatticeStage3v1_9.json is read where the
reference checkout exists (this container); it is never needed on the GPU box."""

import json
import os

import pytest
import torch

import lynx_b200 as lx
from lynx_b200 import fixtures as F
from lynx_b200 import latticejson as LJ

REF_JSON = "/root/reference/docs/examples/ARESlatticeStage3v1_9.json"


def _same_features(a, b):
    assert type(a) is type(b) and a.name == b.name
    for f in a.defining_features:
        attr = "_transfer_map" if f == "transfer_map" else f
        va, vb = getattr(a, attr), getattr(b, attr)
        if isinstance(va, torch.Tensor):
            assert torch.equal(va.float().reshape(-1), vb.float().reshape(-1)), (a.name, f)
        else:
            assert va == vb, (a.name, f)


def test_save_and_reload(tmp_path):
    """tests/test_lattice_json.py:6-31, with every element class and a nested segment."""
    one = torch.tensor([1.0])
    tm = torch.eye(7).reshape(1, 7, 7).clone()
    tm[0, 0, 1] = 0.3
    inner = lx.Segment([
        lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([4.2]),
                      misalignment=torch.tensor([[1e-4, -2e-4]]), tilt=torch.tensor([0.1]), name="Q1"),
        lx.Drift(length=torch.tensor([0.4]), name="D2"),
    ], name="inner_cell")
    original = lx.Segment([
        lx.Marker(name="start"),
        lx.Drift(length=torch.tensor([0.6]), name="D1"),
        inner,
        lx.HorizontalCorrector(length=torch.tensor([0.1]), angle=torch.tensor([1e-4]), name="H1"),
        lx.VerticalCorrector(length=torch.tensor([0.1]), angle=torch.tensor([-1e-4]), name="V1"),
        lx.Dipole(length=one, angle=torch.tensor([0.1]), e1=torch.tensor([0.05]), name="B1"),
        lx.RBend(length=one, angle=torch.tensor([0.2]), name="B2"),
        lx.Cavity(length=one, voltage=torch.tensor([1e6]), phase=torch.tensor([30.0]),
                  frequency=torch.tensor([1.3e9]), name="C1"),
        lx.Solenoid(length=one, k=torch.tensor([0.5]), name="S1"),
        lx.Undulator(length=one, name="U1"),
        lx.BPM(name="BPM1"),
        lx.Aperture(x_max=torch.tensor([1e-3]), y_max=torch.tensor([2e-3]), shape="elliptical",
                    is_active=True, name="A1"),
        lx.CustomTransferMap(tm, length=torch.tensor([0.3]), name="M1"),
        lx.Screen(resolution=torch.tensor((2448, 2040)), pixel_size=torch.tensor((3.5e-6, 2.5e-6)),
                  binning=torch.tensor(2), misalignment=torch.tensor([[1e-3, 0.0]]), is_active=True,
                  name="SCR1"),
    ], name="ARES_Segment")
    path = str(tmp_path / "lattice.json")
    original.to_lattice_json(path, title="ARES LatticeJSON", info="Save and reload test")
    reloaded = lx.Segment.from_lattice_json(path)

    assert original.name == reloaded.name
    assert len(original.elements) == len(reloaded.elements)
    assert torch.equal(original.length, reloaded.length)
    for a, b in zip(original.flattened().elements, reloaded.flattened().elements):
        _same_features(a, b)
    assert isinstance(reloaded.elements[2], lx.Segment) and reloaded.elements[2].name == "inner_cell"
    assert reloaded.SCR1.is_active is True and reloaded.A1.shape == "elliptical"

    doc = json.load(open(path))
    assert doc["version"] == "cheetah-0.6" and doc["root"] == "ARES_Segment"
    assert doc["title"] == "ARES LatticeJSON" and doc["info"] == "Save and reload test"
    assert doc["lattices"]["ARES_Segment"][2] == "inner_cell"
    assert doc["lattices"]["inner_cell"] == ["Q1", "D2"]
    assert doc["elements"]["SCR1"][1]["resolution"] == [2448, 2040]  # pixel counts stay integers
    # a second save of the reloaded segment writes the same bytes
    assert LJ.dumps(LJ.segment_to_dict(reloaded, "ARES LatticeJSON", "Save and reload test")) == open(path).read()


def test_default_title_and_root():
    """latticejson.py:84-92: the title falls back to the segment's name."""
    seg = lx.Segment([lx.Drift(length=torch.tensor([1.0]), name="D")], name="cell_7")
    doc = LJ.segment_to_dict(seg)
    assert doc["title"] == "cell_7" and doc["root"] == "cell_7"
    assert doc["info"] == "This is a placeholder lattice description"


def test_unknown_element_class_is_an_error():
    doc = {"root": "cell", "elements": {"X": ["Sextupole", {"length": [0.1]}]},
           "lattices": {"cell": ["X"]}}
    with pytest.raises(ValueError, match="Sextupole"):
        LJ.segment_from_dict(doc)


@pytest.mark.skipif(not os.path.exists(REF_JSON), reason="reference checkout not present")
def test_reference_example_file_loads_and_is_reproduced_byte_for_byte():
    """The reference's own 195-element ARES file: loads, its experimental-area subcell is the
    built-in fixture (C6) feature for feature, and saving it again writes the identical
    text."""
    text = open(REF_JSON).read()
    doc = json.loads(text)
    seg = lx.Segment.from_lattice_json(REF_JSON)
    assert len(seg.elements) == 195
    ea = seg.subcell("AREASOLA1", "AREABSCR1")
    fixture = F.build_segment(F.ares_ea())
    assert len(ea.elements) == len(fixture.elements) == 13
    for a, b in zip(ea.elements, fixture.elements):
        _same_features(a, b)
    assert LJ.dumps(LJ.segment_to_dict(seg, title=doc["title"], info=doc["info"])) == text
