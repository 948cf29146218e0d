"""The reference's own tests that need no tracking, restated against ``lynx_b200`` (CPU).

Each test names the reference test it mirrors; the tracking ones live in the ``-m gpu``
files (tests/test_gpu_parity.py, tests/test_gpu_beamops.py)."""

import torch

import lynx_b200 as lx


def test_no_markers_left_after_removal():
    """tests/test_speed_optimizations.py:99-117."""
    segment = lx.Segment(elements=[
        lx.Drift(length=torch.tensor([0.6])),
        lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([4.2])),
        lx.Marker(),
        lx.Drift(length=torch.tensor([0.4])),
        lx.HorizontalCorrector(length=torch.tensor([0.1]), angle=torch.tensor([1e-4])),
        lx.Marker(),
    ])
    optimized = segment.without_inactive_markers()
    assert not any(isinstance(e, lx.Marker) for e in optimized.elements)
    assert len(optimized.elements) == 4


def test_inactive_magnet_is_replaced_by_drift():
    """tests/test_speed_optimizations.py:120-137."""
    segment = lx.Segment(elements=[
        lx.Drift(length=torch.tensor([0.6])),
        lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([0.0])),
        lx.Drift(length=torch.tensor([0.4])),
    ])
    optimized = segment.inactive_elements_as_drifts()
    assert all(isinstance(e, lx.Drift) for e in optimized.elements)
    assert torch.equal(optimized.elements[1].length, torch.tensor([0.2]))


def test_active_elements_not_replaced_by_drift():
    """tests/test_speed_optimizations.py:140-155."""
    segment = lx.Segment(elements=[
        lx.Drift(length=torch.tensor([0.6])),
        lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([4.2])),
        lx.Drift(length=torch.tensor([0.4])),
    ])
    optimized = segment.inactive_elements_as_drifts()
    assert isinstance(optimized.elements[1], lx.Quadrupole)


def test_except_for_keeps_named_elements():
    """segment.py ``except_for`` arguments of the three clean-up passes."""
    segment = lx.Segment(elements=[
        lx.Marker(name="keep_me"),
        lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([0.0]), name="Q_off"),
        lx.Marker(name="drop_me"),
    ])
    kept = segment.without_inactive_markers(except_for=["keep_me"])
    assert [e.name for e in kept.elements] == ["keep_me", "Q_off"]
    kept = segment.inactive_elements_as_drifts(except_for=["Q_off"])
    assert isinstance(kept.elements[1], lx.Quadrupole)


def test_segment_length_shape():
    """tests/test_vectorized.py:8-37 (one and two batch dimensions)."""
    segment = lx.Segment(elements=[
        lx.Drift(length=torch.tensor([0.6, 0.5])),
        lx.Quadrupole(length=torch.tensor([0.2, 0.25]), k1=torch.tensor([4.2, 4.2])),
        lx.Drift(length=torch.tensor([0.4, 0.3])),
    ])
    assert segment.length.shape == (2,)
    assert torch.allclose(segment.length, torch.tensor([1.2, 1.05]))
    segment = lx.Segment(elements=[
        lx.Drift(length=torch.tensor([[0.6, 0.5], [0.4, 0.3], [0.4, 0.3]])),
        lx.Quadrupole(length=torch.tensor([[0.2, 0.25], [0.3, 0.35], [0.3, 0.35]]),
                      k1=torch.tensor([[4.2, 4.2], [4.3, 4.3], [4.3, 4.3]])),
        lx.Drift(length=torch.tensor([[0.4, 0.3], [0.2, 0.1], [0.2, 0.1]])),
    ])
    assert segment.length.shape == (3, 2)


def test_screen_length_shape():
    """tests/test_vectorized.py:442-457."""
    screen = lx.Screen(misalignment=torch.tensor([[0.1, 0.2], [0.3, 0.4]]))
    assert screen.length.shape == screen.misalignment.shape[:-1]
    screen = lx.Screen(misalignment=torch.tensor([[0.1, 0.2]]))
    broadcast_screen = screen.broadcast((3, 10))
    assert broadcast_screen.length.shape == broadcast_screen.misalignment.shape[:-1]


def test_split_of_elements_without_a_slicing_rule_returns_the_element():
    """cavity.py:336-339, solenoid.py:122-124, dipole.py:196-199, undulator.py:73-75: the
    reference returns ``[self]`` for these (tests/test_split.py:44-130 are xfail there)."""
    one = torch.tensor([1.0])
    for el in (
        lx.Cavity(length=one, voltage=torch.tensor([1e6]), phase=torch.tensor([0.0]),
                  frequency=torch.tensor([1.3e9])),
        lx.Solenoid(length=one, k=torch.tensor([0.5])),
        lx.Dipole(length=one, angle=torch.tensor([0.1])),
        lx.Undulator(length=one),
    ):
        assert el.split(resolution=torch.tensor(0.1)) == [el]


def test_split_lengths_add_up():
    """tests/test_split.py:7-43: twenty 0.1 m slices of a 2 m drift, 0.01 m slices of a
    0.2 m quadrupole with the quadrupole's strength on every slice."""
    drift = lx.Drift(length=torch.tensor([2.0]))
    slices = drift.split(resolution=torch.tensor(0.1))
    assert all(isinstance(s, lx.Drift) for s in slices)
    assert torch.allclose(sum(s.length for s in slices), drift.length, atol=1e-6)
    quad = lx.Quadrupole(length=torch.tensor([0.2]), k1=torch.tensor([4.2]))
    slices = quad.split(resolution=torch.tensor(0.01))
    assert all(torch.equal(s.k1, quad.k1) for s in slices)
    assert torch.allclose(sum(s.length for s in slices), quad.length, atol=1e-6)
    hcor = lx.HorizontalCorrector(length=torch.tensor([0.3]), angle=torch.tensor([3e-4]))
    slices = hcor.split(resolution=torch.tensor(0.1))
    assert torch.allclose(sum(s.angle for s in slices), hcor.angle, atol=1e-9)


def test_lengthless_elements_have_zero_length():
    """tests/test_tracking_lengthless_elements.py: Marker / BPM / Screen segments have a
    defined (zero) length and are skippable unless active."""
    segment = lx.Segment(elements=[lx.Marker(), lx.BPM(), lx.Screen(), lx.Marker()])
    assert float(segment.length.sum()) == 0.0
    assert segment.is_skippable
    segment.elements[1].is_active = True
    assert not segment.is_skippable
