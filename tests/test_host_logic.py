"""CPU: host-side logic -- element surface, program encoding, column layout, batch
shape inference, run-grouping decisions (no device calls)."""

import pytest
import torch

import lynx_b200 as lx
from lynx_b200 import _lib as L
from lynx_b200 import fixtures as F
from lynx_b200.program import encode, layout, pack_params


def test_reference_surface_names():
    for n in ("Segment", "Drift", "Quadrupole", "HorizontalCorrector", "VerticalCorrector",
              "Dipole", "RBend", "Cavity", "BPM", "Screen", "Marker", "CustomTransferMap",
              "Solenoid", "Undulator", "ParticleBeam", "ParameterBeam", "Beam"):
        assert hasattr(lx, n)
    assert lx.Beam.empty == "I'm an empty beam!"


def test_name_registry_and_subcell():
    seg = F.build_segment(F.ares_ea())
    assert seg.AREAMQZM2.name == "AREAMQZM2"
    seg.AREAMQZM2.k1 = torch.tensor([4.2])
    assert seg.elements[4].k1.item() == pytest.approx(4.2)
    sub = seg.subcell("AREAMQZM1", "AREAMCVM1")
    assert [e.name for e in sub.elements][0] == "AREAMQZM1"
    assert [e.name for e in sub.elements][-1] == "AREAMCVM1"
    assert len(sub.elements) == 5
    assert float(seg.length) == pytest.approx(2.0460, abs=1e-3)
    dup = lx.Segment([lx.Drift(torch.tensor([1.0]), name="d"), lx.Drift(torch.tensor([2.0]), name="d")])
    assert isinstance(dup.d, list) and len(dup.d) == 2


def test_encode_ares_ea_inactive_screen():
    seg = F.build_segment(F.ares_ea())
    enc = encode(seg.elements, (1,))
    # marker and inactive screen vanish; 12 ops remain? marker -> dropped, screen -> dropped
    assert [o[0] for o in enc.ops] == [L.OP_DRIFT, L.OP_QUAD, L.OP_DRIFT, L.OP_QUAD, L.OP_DRIFT,
                                       L.OP_VCOR, L.OP_DRIFT, L.OP_QUAD, L.OP_DRIFT, L.OP_HCOR,
                                       L.OP_DRIFT]
    assert enc.B == 1 and enc.screen_element is None and enc.bpm_elements == []
    off, stride, shared, varying, total = layout(enc)
    assert len(off) == len(enc.columns) == 6 * 1 + 3 * 5 + 2 * 2
    assert set(stride) == {0} and varying == []
    buf = pack_params(enc, "cpu")
    assert buf.numel() == total
    assert buf[off[enc.ops[1][1] + 1]].item() == 0.0  # k1 of AREAMQZM1


def test_encode_active_screen_is_terminal_and_bpm_slots():
    spec = F.smatch(bpms_active=True)
    spec.insert(4, ("screen", "S", {"resolution": (10, 10), "pixel_size": (1e-3, 1e-3),
                                    "is_active": True}))
    seg = F.build_segment(spec)
    enc = encode(seg.elements, (1,))
    assert enc.ops[-1][0] == L.OP_SCREEN
    assert [o[3] for o in enc.ops if o[0] == L.OP_BPM] == [0, 1]
    assert [e.name for e in enc.bpm_elements] == ["BPM1SMATCH", "BPM6SMATCH"]
    assert enc.screen_element.name == "S"


def test_per_setting_columns_and_batch_shape():
    seg = F.build_segment(F.ares_ea())
    seg.AREAMQZM1.k1 = torch.linspace(-30, 30, 12).reshape(3, 4)
    enc = encode(seg.elements, (1,))
    assert enc.batch_shape == (3, 4) and enc.B == 12
    off, stride, shared, varying, total = layout(enc)
    assert sum(stride) == 1
    k1_col = enc.ops[1][1] + 1
    assert stride[k1_col] == 1 and off[k1_col] % 4 == 0
    buf = pack_params(enc, "cpu")
    assert torch.equal(buf[off[k1_col]: off[k1_col] + 12], torch.linspace(-30, 30, 12))
    assert total == buf.numel()


def test_broadcast_is_stride0_and_skippability():
    q = lx.Quadrupole(torch.tensor([0.4]), torch.tensor([4.2]))
    qb = q.broadcast((3, 10))
    assert qb.length.shape == (3, 10) and qb.k1.shape == (3, 10)
    assert qb.misalignment.shape == (3, 10, 2)
    assert qb.k1.stride() == (0, 0) and bool((qb.k1 == 4.2).all())
    cav = lx.Cavity(torch.tensor([1.0]), torch.tensor([0.0]))
    assert cav.is_skippable and not cav.is_active
    cav.voltage = torch.tensor([1e6])
    assert not cav.is_skippable
    assert lx.BPM().is_skippable and not lx.BPM(is_active=True).is_skippable
    scr = lx.Screen(is_active=True)
    assert not scr.is_skippable
    seg = lx.Segment([q, cav])
    assert not seg.is_skippable
    ctm = lx.CustomTransferMap(torch.eye(7).unsqueeze(0), length=torch.tensor([0.4]))
    b = ctm.broadcast((3, 10))
    assert b._transfer_map.shape == (3, 10, 7, 7) and b.length.shape == (3, 10)


def test_screen_geometry_and_empty_reading():
    scr = lx.Screen(resolution=torch.tensor([2448.0, 2040.0]),
                    pixel_size=torch.tensor([3.3198e-6, 2.4469e-6]), name="s")
    ex, ey = scr.pixel_bin_edges
    assert ex.numel() == 2449 and ey.numel() == 2041
    assert scr.reading.shape == (2040, 2448) and float(scr.reading.sum()) == 0.0
    s2 = lx.Screen(resolution=torch.tensor([100.0, 100.0]), binning=torch.tensor(4.0))
    assert s2.pixel_bin_edges[0].numel() == 26


def test_track_requires_cuda():
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    seg = F.build_segment(F.smatch())
    with pytest.raises(Exception):
        lx.ParticleBeam(torch.ones(1, 8, 7), torch.tensor([1e8]))
    assert seg.track(lx.Beam.empty) is lx.Beam.empty
