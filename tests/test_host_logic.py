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


def test_split_slices_follow_the_reference_loops():
    """drift.py:71-78, quadrupole.py:99-110, horizontal_corrector.py:82-90."""
    parts = lx.Drift(torch.tensor([0.43])).split(torch.tensor(0.05))
    assert len(parts) == 9 and all(isinstance(p, lx.Drift) for p in parts)
    assert float(sum(p.length for p in parts)) == pytest.approx(0.43, rel=1e-6)
    assert float(parts[-1].length) == pytest.approx(0.03, rel=1e-4)
    q = lx.Quadrupole(torch.tensor([0.122]), k1=torch.tensor([7.5]),
                      misalignment=torch.tensor([[1e-4, -2e-4]]))
    qs = q.split(torch.tensor(0.05))
    assert [round(float(p.length), 3) for p in qs] == [0.05, 0.05, 0.022]
    assert all(float(p.k1) == 7.5 and torch.equal(p.misalignment, q.misalignment) for p in qs)
    hc = lx.VerticalCorrector(torch.tensor([0.1]), angle=torch.tensor([2e-3]))
    hs = hc.split(torch.tensor(0.04))
    assert all(isinstance(p, lx.VerticalCorrector) for p in hs) and len(hs) == 3
    assert float(sum(p.angle for p in hs)) == pytest.approx(2e-3, rel=1e-5)  # kick is shared out
    # elements the reference does not split return themselves (dipole.py:196-199 et al.)
    d = lx.Dipole(torch.tensor([0.5]), angle=torch.tensor([0.1]))
    assert d.split(torch.tensor(0.1)) == [d]
    seg = lx.Segment([lx.Drift(torch.tensor([0.1])), d, lx.Marker(name="m")])
    assert len(seg.split(torch.tensor(0.05))) == 2 + 1 + 1


def test_aperture_surface():
    """aperture.py:24-68, 111-145."""
    ap = lx.Aperture(name="a")
    assert ap.is_active and not ap.is_skippable and ap.shape == "rectangular"
    assert torch.isinf(ap.x_max).all() and torch.isinf(ap.y_max).all()
    assert ap.lost_particles is None and ap._encode() is None
    ap.is_active = False
    assert ap.is_skippable
    pb = object.__new__(lx.ParameterBeam)
    assert ap.track(pb) is pb  # only ParticleBeams are cut (aperture.py:71-72)
    assert lx.Aperture().track(lx.Beam.empty) is lx.Beam.empty
    b = lx.Aperture(x_max=torch.tensor([1e-3]), y_max=torch.tensor([2e-3]), shape="elliptical")
    bb = b.broadcast((3,))
    assert bb.x_max.shape == (3,) and bb.shape == "elliptical" and bb.name == b.name
    assert b.defining_features == ["x_max", "y_max", "shape", "is_active"]
    # an inactive aperture vanishes from the encoded program; an active one is not an op either
    seg = lx.Segment([lx.Drift(torch.tensor([0.1])), ap, lx.Drift(torch.tensor([0.2]))])
    assert [o[0] for o in encode(seg.elements, (1,)).ops] == [L.OP_DRIFT, L.OP_DRIFT]


def test_new_entry_points_validate_arguments():
    """LYNX_E_BADARG before anything touches the device (no GPU needed)."""
    lib = L.lib()
    assert lib.lynx_beam_transform(None, 0, 4, 1, 4, None, None, 4, None) == L.E_BADARG
    assert lib.lynx_beam_transform(16, 0, 6, 1, 4, 16, 16, 8, None) == L.E_BADARG  # ld % 4
    assert lib.lynx_beam_transform(16, 0, 8, 1, 9, 16, 16, 8, None) == L.E_BADARG  # ld < N
    assert lib.lynx_beam_transform(20, 0, 8, 1, 8, 16, 16, 8, None) == L.E_BADARG  # alignment
    assert lib.lynx_beam_generate_ellipsoid(None, None, 1, 4, 0, 1, None, 4, None) == L.E_BADARG
    assert lib.lynx_beam_generate_ellipsoid(16, 16, 1, 8, 2, 1, 16, 8, None) == L.E_BADARG  # first % 4
    assert lib.lynx_aperture_workspace_bytes(1, 1) == 8 + 16  # one status word + the ticket block
    assert lib.lynx_aperture_workspace_bytes(2, 1025) == 2 * 2 * 8 + 16
    assert lib.lynx_aperture_workspace_bytes(0, 10) == 0
    args = [16, 1, 8, 8, None, 0, 16, 16, 0, 0, 16, None, 8, None, None, 16, 16, None]
    bad_shape = list(args)
    bad_shape[9] = 2
    assert lib.lynx_aperture_compact(*bad_shape) == L.E_BADARG
    bad_ld = list(args)
    bad_ld[12] = 6
    assert lib.lynx_aperture_compact(*bad_ld) == L.E_BADARG
    no_counts = list(args)
    no_counts[15] = None
    assert lib.lynx_aperture_compact(*no_counts) == L.E_BADARG
    assert lib.lynx_parambeam_along(None, None, 1, None, None, 0, None, None, None, None,
                                    None) == L.E_BADARG


def test_parameter_beam_from_twiss_and_transformed_to():
    """parameter_beam.py:147-233, 278-369 (host logic; constructed on the CPU here) with the
    reference's own checks: tests/test_parameter_beam.py from_twiss -> twiss, transformed_to."""
    t = lambda v: torch.tensor([v])  # noqa: E731
    beam = lx.ParameterBeam.from_twiss(
        beta_x=t(5.91253676811640894), alpha_x=t(3.55631307633660354),
        emittance_x=t(3.494768647122823e-09), beta_y=t(5.91253676811640982),
        alpha_y=t(1.0), emittance_y=t(3.497810737006068e-09), energy=t(6e6), device="cpu")
    assert beam._mu.shape == (1, 7) and beam._cov.shape == (1, 7, 7)
    assert beam.beta_x.item() == pytest.approx(5.91253676811640894, rel=1e-4)
    assert beam.alpha_x.item() == pytest.approx(3.55631307633660354, rel=1e-4)
    assert beam.emittance_x.item() == pytest.approx(3.494768647122823e-09, rel=1e-4)
    assert beam.beta_y.item() == pytest.approx(5.91253676811640982, rel=1e-4)
    assert beam.alpha_y.item() == pytest.approx(1.0, rel=1e-4)
    assert beam.energy.item() == pytest.approx(6e6)
    default = lx.ParameterBeam.from_twiss(device="cpu")
    assert default.beta_x.item() == pytest.approx(1.0, rel=1e-5)
    assert default.emittance_x.item() == pytest.approx(7.1971891e-13, rel=1e-4)
    with pytest.raises(AssertionError):
        lx.ParameterBeam.from_twiss(beta_x=t(-1.0), device="cpu")

    original = lx.ParameterBeam.from_parameters(device="cpu")
    tb = original.transformed_to(
        mu_x=t(1e-5), mu_xp=t(1e-7), mu_y=t(2e-5), mu_yp=t(2e-7), sigma_x=t(1.75e-7),
        sigma_xp=t(2e-7), sigma_y=t(1.75e-7), sigma_yp=t(2e-7), sigma_s=t(0.000001),
        sigma_p=t(0.000001), energy=t(1e7), total_charge=t(1e-9))
    assert isinstance(tb, lx.ParameterBeam)
    for name, want in (("mu_x", 1e-5), ("mu_xp", 1e-7), ("mu_y", 2e-5), ("mu_yp", 2e-7),
                       ("sigma_x", 1.75e-7), ("sigma_xp", 2e-7), ("sigma_y", 1.75e-7),
                       ("sigma_yp", 2e-7), ("sigma_s", 1e-6), ("sigma_p", 1e-6),
                       ("energy", 1e7), ("total_charge", 1e-9)):
        assert getattr(tb, name).item() == pytest.approx(want, rel=1e-5), name
    # only what is asked for changes
    tb2 = original.transformed_to(sigma_x=t(3e-6))
    assert tb2.sigma_x.item() == pytest.approx(3e-6, rel=1e-5)
    assert tb2.sigma_y.item() == pytest.approx(original.sigma_y.item(), rel=1e-6)
    assert tb2._mu.device.type == "cpu"


def test_particle_beam_coordinate_setters():
    """particle_beam.py:728-812: xs / xps / ... are settable (uniform_3d_ellipsoid itself
    assigns them, :381-383); the SoA plane is updated in place and cached moments dropped."""
    soa = torch.zeros((1, 6, 8))
    beam = lx.ParticleBeam.from_soa(soa, 6, torch.tensor([1e8]), moment_sums=torch.ones(1, 28))
    beam.xs = torch.arange(6.0).reshape(1, 6)
    beam.ps = torch.full((1, 6), 2.0)
    assert beam._sums is None
    assert torch.equal(beam.xs, torch.arange(6.0).reshape(1, 6))
    assert torch.equal(beam.soa[0, 0], torch.tensor([0.0, 1, 2, 3, 4, 5, 0, 0]))  # padding stays 0
    assert torch.equal(beam.soa[0, 5, :6], torch.full((6,), 2.0)) and float(beam.soa[0, 1].sum()) == 0
    # a broadcast beam does not write through to the beam it was broadcast from
    b2 = beam.broadcast((3,))
    b2.ys = torch.ones(3, 6)
    assert float(beam.ys.sum()) == 0.0 and float(b2.ys.sum()) == 18.0


def test_broadcast_shapes_matches_torch():
    """program.broadcast_shapes replaces torch.broadcast_shapes on the call path of
    Segment.track: same result (or the same RuntimeError) for every combination."""
    import itertools

    from lynx_b200.program import _batch_shape, broadcast_shapes

    cands = [(), (1,), (3,), (3, 1), (1, 4), (2, 3, 4), (1, 1), torch.Size([5, 1, 1]), (0,), (4,)]
    for r in range(0, 4):
        for combo in itertools.product(cands, repeat=r):
            try:
                want = tuple(torch.broadcast_shapes(*combo))
            except RuntimeError:
                with pytest.raises(RuntimeError):
                    broadcast_shapes(*combo)
                continue
            assert broadcast_shapes(*combo) == want
            if r:
                assert _batch_shape(list(combo)) == want


class _RecordingLib:
    """Stands in for the shared library so that the Python call path around the C ABI can be
    exercised without a GPU: records the arguments of every entry point, computes nothing."""

    def __init__(self, param_floats=0):
        self.calls = []
        self.param_floats = param_floats  # snapshot taken DURING the call: the buffer is freed after
        self.params = None

    def __getattr__(self, name):
        def fn(*args):
            import ctypes

            self.calls.append((name, args))
            if name.startswith("lynx_track_particles_fwd") and self.param_floats:
                self.params = list((ctypes.c_float * self.param_floats).from_address(args[1]))
            if name == "lynx_program_num_stages":
                return 1
            if name.endswith("workspace_bytes"):
                return 256
            return 0

        return fn


def test_track_call_path_arguments_without_gpu(monkeypatch):
    """Segment.track up to the C-ABI call, with the library replaced by a recorder: the
    packed parameter buffer, the batch size, the leading dimension and the particle count
    that reach ``lynx_track_particles_fwd`` are the ones the lattice and the beam define."""
    from lynx_b200 import program, segment

    fake = _RecordingLib()
    monkeypatch.setattr(L, "lib", lambda: fake)
    monkeypatch.setattr(L, "current_stream", lambda: 0)
    monkeypatch.setattr(segment, "_require_cuda", lambda d: None)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(program.Program, "_cache", {})

    seg = F.build_segment(F.ares_ea(k1=(5.0, -5.0, 5.0), vkick=1e-3, hkick=-2e-3))
    seg.AREAMQZM2.k1 = torch.linspace(-3.0, 3.0, 6).reshape(2, 3)
    n, ld = 1001, 1004
    beam = lx.ParticleBeam.from_soa(torch.zeros(1, 6, ld), n, torch.tensor([1e8]), torch.ones(n))
    enc = encode(seg.elements, (1,))
    enc.batch_shape = (2, 3)
    off, stride, shared, varying, total = layout(enc)
    fake.param_floats = total
    out = seg.track(beam)
    assert out.num_particles == n and tuple(out.soa.shape) == (2, 3, 6, ld)

    name, a = [c for c in fake.calls if c[0].startswith("lynx_track_particles_fwd")][-1]
    B, beam_stride, a_ld, a_n = a[2], a[4], a[5], a[6]
    assert (B, beam_stride, a_ld, a_n) == (6, 0, ld, n)  # one shared beam, six settings
    want = pack_params(enc, "cpu")
    assert fake.params == want.tolist()
    assert len(varying) == 1 and stride[varying[0]] == 1
    k = off[varying[0]]
    assert want[k:k + 6].tolist() == seg.AREAMQZM2.k1.reshape(-1).tolist()
