"""NX-tables reader (lynx_b200/nxtables.py) -- CPU only.  The reference's resource file is
read where the reference checkout exists (this container), never on the GPU box."""

import os

import pytest
import torch

import lynx_b200 as lx

REF_FILE = "/root/reference/tests/resources/Stage4v3_9.txt"
HEADER = "NAME,CLASS,X_beam,Y_beam,Z_beam,Comments"


def _write(tmp_path, rows):
    path = tmp_path / "mini_stage.txt"
    path.write_text("\n".join([HEADER] + rows) + "\n")
    return path


def test_small_table(tmp_path):
    """Sorted by Z_beam, ignored classes dropped, the combined corrector split into its two
    coils, gaps filled with drifts named after their neighbours (nxtables.py:211-265)."""
    path = _write(tmp_path, [
        "ARMRMQZM1,MQZM,0,0,1.000,",
        "ARLISOLG1,SOLG,0,0,0.000,gun",
        "ARLIRSBG1,RSBG,0,0,0.050,ignored",
        "ARMRMCXG1,MCXG,0,0,0.500,",
        "ARMRBSCR1,BSCR,0,0,2.000,",
        "ARMRBPMG1,BPMG,0,0,2.000,same position as the screen",
    ])
    seg = lx.Segment.from_nx_tables(path)
    assert seg.name == "mini_stage"
    names = [e.name for e in seg.elements]
    assert names == ["ARLISOLG1", "DRIFT_ARLISOLG1_ARMRMCXG1", "ARMRMCHG1", "ARMRMCVG1",
                     "DRIFT_ARMRMCXG1_ARMRMQZM1", "ARMRMQZM1", "DRIFT_ARMRMQZM1_ARMRBSCR1",
                     "ARMRBSCR1", "ARMRBPMG1"]
    kinds = [type(e) for e in seg.elements]
    assert kinds == [lx.Marker, lx.Drift, lx.HorizontalCorrector, lx.VerticalCorrector, lx.Drift,
                     lx.Quadrupole, lx.Drift, lx.Screen, lx.BPM]
    d1, d2, d3 = seg.elements[1], seg.elements[4], seg.elements[6]
    assert torch.allclose(d1.length, torch.tensor([0.5 - 1e-4 / 2]))
    assert torch.allclose(d2.length, torch.tensor([0.5 - 1e-4 / 2 - 0.122 / 2]))
    assert torch.allclose(d3.length, torch.tensor([1.0 - 0.122 / 2]))
    assert torch.allclose(seg.length, torch.tensor([2.0]))  # centre of first to centre of last
    scr = seg.ARMRBSCR1
    assert scr.resolution.tolist() == [2448, 2040] and not scr.is_active


def test_unknown_class_and_overlap_are_errors(tmp_path):
    with pytest.raises(ValueError, match="unknown class XYZQ"):
        lx.Segment.from_nx_tables(_write(tmp_path, ["A,SOLG,0,0,0,", "B,XYZQ,0,0,1,"]))
    with pytest.raises(AssertionError, match="overlap"):
        lx.Segment.from_nx_tables(_write(tmp_path, ["ARMRMQZM1,MQZM,0,0,0.00,", "ARMRMQZM2,MQZM,0,0,0.10,"]))


@pytest.mark.skipif(not os.path.exists(REF_FILE), reason="reference checkout not present")
def test_reference_table_known_length():
    """tests/test_reading_nx_tables.py:4-25: plausible lattice, and the length the reference
    pinned when the reader was written."""
    seg = lx.Segment.from_nx_tables(REF_FILE)
    assert isinstance(seg, lx.Segment)
    assert len(seg.elements) > 1
    assert 40.0 < seg.length < 50.0
    assert seg.length == 44.2215
    assert not any(isinstance(e, lx.Segment) for e in seg.elements)  # flattened
