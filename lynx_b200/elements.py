"""Lattice elements: host-side mirror of lynx/accelerator/*.py.

Same class names, constructor arguments, settable attributes, ``transfer_map`` /
``track`` / ``broadcast`` / ``is_skippable`` / ``is_active`` as the reference.  Elements
only HOLD parameters (torch tensors, any device); all arithmetic happens in the CUDA
library: ``transfer_map`` runs the K1 composer, ``track`` the fused K2 / K6 kernels via
a one-element ``Segment``.
"""

from typing import Optional

import torch

from . import _lib as L
from .beams import Beam, ParameterBeam, ParticleBeam


class _NameGen:
    """lynx/utils.py:1-11."""

    def __init__(self, prefix):
        self._prefix, self._n = prefix, 0

    def __call__(self):
        name = f"{self._prefix}_{self._n}"
        self._n += 1
        return name


generate_unique_name = _NameGen("unnamed_element")


def _t(value, like=None, default=0.0) -> torch.Tensor:
    if value is None:
        return torch.full_like(like, default) if like is not None else torch.tensor([default])
    return torch.as_tensor(value, dtype=torch.float32)


class Element:
    """lynx/accelerator/element.py:23-149."""

    _params = ()  # names of the tensor attributes that ``broadcast`` tiles
    # attributes a track WRITES (readings, diagnostics): setting them does not change the lattice
    _outputs = frozenset(("reading", "cached_reading", "lost_particle_charges"))
    _ver = 0

    def __setattr__(self, name, value):
        # every (re)assignment of a public attribute -- a parameter tensor, is_active, binning ...
        # -- bumps the element's version; Segment.track keys its cache of the encoded lattice
        # (program handle + packed device parameters) on it and on the tensors' own in-place
        # version counters (_stamp below)
        object.__setattr__(self, name, value)
        if name[0] != "_" and name not in self._outputs or name == "_transfer_map":
            object.__setattr__(self, "_ver", self._ver + 1)

    def _stamp(self):
        """Changes whenever something ``_encode`` reads may have changed: an attribute was
        assigned, or a parameter tensor was modified in place (``Tensor._version``; writes
        through ``.data`` are invisible to it -- assign the attribute instead)."""
        d = self.__dict__
        return (id(self), self._ver) + tuple(d[n]._version for n in self._params if n in d)

    def __init__(self, name: Optional[str] = None) -> None:
        self.name = name if name is not None else generate_unique_name()
        if not hasattr(self, "length"):
            self.length = torch.zeros(1)

    # -- to be provided by subclasses
    def _encode(self):
        """(opcode, [column tensors], flags) for the C ABI, or None for identity."""
        raise NotImplementedError

    @property
    def is_skippable(self) -> bool:
        return True

    # -- reference surface
    def transfer_map(self, energy: torch.Tensor) -> torch.Tensor:
        """(..., 7, 7) map at ``energy`` (element.py:36-59), computed on the device."""
        from .segment import compose_transfer_map

        return compose_transfer_map([self], energy)

    def track(self, incoming):
        """element.py:61-94."""
        if incoming is Beam.empty:
            return incoming
        if not isinstance(incoming, (ParameterBeam, ParticleBeam)):
            raise TypeError(f"Parameter incoming is of invalid type {type(incoming)}")
        from .segment import track_elements

        return track_elements([self], incoming)

    def forward(self, incoming):
        return self.track(incoming)

    __call__ = forward

    def broadcast(self, shape: tuple) -> "Element":
        """Stride-0 broadcast of every parameter (e.g. drift.py:64-65 uses ``repeat``)."""
        new = self.__class__.__new__(self.__class__)
        new.__dict__.update(self.__dict__)
        for n in self._params:
            t = getattr(self, n)
            if n in ("misalignment",):
                v = t.reshape(2) if t.numel() == 2 else t
                setattr(new, n, v.expand(*shape, 2) if v.dim() == 1 else v.repeat(*shape, 1))
            else:
                v = t.reshape(()) if t.numel() == 1 else t
                setattr(new, n, v.expand(shape) if v.dim() == 0 else v.repeat(*shape))
        return new

    def split(self, resolution):
        return [self]

    @property
    def defining_features(self):
        return list(self._params)

    def __repr__(self) -> str:
        ps = ", ".join(f"{n}={getattr(self, n)!r}" for n in self._params)
        return f"{self.__class__.__name__}({ps}, name={self.name!r})"


def _split_lengths(length, resolution):
    """Slice lengths of the ``split`` loops (drift.py:71-78 et al.): ``min(resolution,
    remaining)`` while ``remaining > 0``."""
    remaining = torch.as_tensor(length, dtype=torch.float32).clone()
    resolution = torch.as_tensor(resolution, dtype=torch.float32)
    out = []
    while bool((remaining > 0).all()) and remaining.numel() > 0:
        out.append(torch.minimum(resolution, remaining))
        remaining = remaining - resolution
    return out


class Marker(Element):
    """marker.py:22-64."""

    def _encode(self):
        return None


class Drift(Element):
    """drift.py:23-88."""

    _params = ("length",)

    def __init__(self, length, name=None, device=None, dtype=torch.float32):
        self.length = _t(length)
        super().__init__(name)

    def _encode(self):
        return (L.OP_DRIFT, [self.length], 0)

    def split(self, resolution):  # drift.py:71-78
        return [Drift(l_) for l_ in _split_lengths(self.length, resolution)]


class Quadrupole(Element):
    """quadrupole.py:25-131."""

    _params = ("length", "k1", "misalignment", "tilt")

    def __init__(self, length, k1=None, misalignment=None, tilt=None, name=None, device=None,
                 dtype=torch.float32):
        self.length = _t(length)
        self.k1 = _t(k1, self.length)
        self.misalignment = (
            _t(misalignment) if misalignment is not None
            else torch.zeros((*self.length.shape, 2))
        )
        self.tilt = _t(tilt, self.length)
        super().__init__(name)

    @property
    def is_active(self) -> bool:
        return bool((self.k1 != 0).any())

    def _encode(self):
        m = self.misalignment
        return (L.OP_QUAD, [self.length, self.k1, m[..., 0], m[..., 1], self.tilt], 0)

    def split(self, resolution):  # quadrupole.py:99-110 (the tilt is dropped there, as written)
        return [Quadrupole(l_, self.k1, misalignment=self.misalignment)
                for l_ in _split_lengths(self.length, resolution)]


class HorizontalCorrector(Element):
    """horizontal_corrector.py:25-110."""

    _params = ("length", "angle")
    _opcode = L.OP_HCOR

    def __init__(self, length, angle=None, name=None, device=None, dtype=torch.float32):
        self.length = _t(length)
        self.angle = _t(angle, self.length)
        super().__init__(name)

    @property
    def is_active(self) -> bool:
        return bool((self.angle != 0).any())

    def _encode(self):
        return (self._opcode, [self.length, self.angle], 0)

    def split(self, resolution):  # horizontal_corrector.py:82-90, vertical_corrector.py:81-89
        return [self.__class__(l_, self.angle * l_ / self.length)
                for l_ in _split_lengths(self.length, resolution)]


class VerticalCorrector(HorizontalCorrector):
    """vertical_corrector.py:25-109."""

    _opcode = L.OP_VCOR


class Dipole(Element):
    """dipole.py:25-234."""

    _params = ("length", "angle", "e1", "e2", "tilt", "fringe_integral",
               "fringe_integral_exit", "gap")
    _flags = 0

    def __init__(self, length, angle=None, e1=None, e2=None, tilt=None, fringe_integral=None,
                 fringe_integral_exit=None, gap=None, name=None, device=None,
                 dtype=torch.float32):
        self.length = _t(length)
        self.angle = _t(angle, self.length)
        self.gap = _t(gap, self.length)
        self.tilt = _t(tilt, self.length)
        self.fringe_integral = _t(fringe_integral, self.length)
        self.fringe_integral_exit = (
            self.fringe_integral if fringe_integral_exit is None else _t(fringe_integral_exit)
        )
        self.e1 = _t(e1, self.length)
        self.e2 = _t(e2, self.length)
        super().__init__(name)

    @property
    def hx(self):
        return torch.where(self.length != 0, self.angle / self.length,
                           torch.zeros_like(self.length))

    @property
    def is_active(self) -> bool:
        return bool((self.angle != 0).any())

    def _encode(self):
        return (L.OP_DIPOLE,
                [self.length, self.angle, self.e1, self.e2, self.tilt, self.fringe_integral,
                 self.fringe_integral_exit, self.gap], self._flags)


class RBend(Dipole):
    """rbend.py:22-80: e1, e2 += angle/2 (applied on the device, flag LYNX_F_RBEND)."""

    _flags = L.F_RBEND


class Cavity(Element):
    """cavity.py:23-361."""

    _params = ("length", "voltage", "phase", "frequency")

    def __init__(self, length, voltage=None, phase=None, frequency=None, name=None,
                 device=None, dtype=torch.float32):
        self.length = _t(length)
        self.voltage = _t(voltage, self.length)
        self.phase = _t(phase, self.length)
        self.frequency = _t(frequency, self.length)
        super().__init__(name)

    @property
    def is_active(self) -> bool:
        return bool((self.voltage != 0).any())

    @property
    def is_skippable(self) -> bool:
        return not self.is_active

    def _encode(self):
        return (L.OP_CAVITY, [self.length, self.voltage, self.phase, self.frequency],
                L.F_ACTIVE if self.is_active else 0)


class Solenoid(Element):
    """solenoid.py:24-145."""

    _params = ("length", "k", "misalignment")

    def __init__(self, length=None, k=None, misalignment=None, name=None, device=None,
                 dtype=torch.float32):
        self.length = _t(length)
        self.k = _t(k, self.length)
        self.misalignment = (
            _t(misalignment) if misalignment is not None
            else torch.zeros((*self.length.shape[:-1], 2))  # solenoid.py:55-59, as written
        )
        super().__init__(name)

    @property
    def is_active(self) -> bool:
        return bool((self.k != 0).any())

    def _encode(self):
        m = self.misalignment
        return (L.OP_SOLENOID, [self.length, self.k, m[..., 0], m[..., 1]], 0)


class Undulator(Element):
    """undulator.py:22-95."""

    _params = ("length",)

    def __init__(self, length, is_active=False, name=None, device=None, dtype=torch.float32):
        self.length = _t(length)
        self.is_active = is_active
        super().__init__(name)

    def _encode(self):
        return (L.OP_UNDULATOR, [self.length], 0)


class CustomTransferMap(Element):
    """custom_transfer_map.py:22-116."""

    _params = ("length",)

    def _stamp(self):
        return super()._stamp() + (self._transfer_map._version,)

    def __init__(self, transfer_map, length=None, name=None, device=None, dtype=torch.float32):
        assert isinstance(transfer_map, torch.Tensor)
        assert transfer_map.shape[-2:] == (7, 7)
        self._transfer_map = transfer_map.to(torch.float32)
        self.length = _t(length) if length is not None else torch.zeros(transfer_map.shape[:-2])
        super().__init__(name)

    @classmethod
    def from_merging_elements(cls, elements, incoming_beam) -> "CustomTransferMap":
        """custom_transfer_map.py:48-85 -- one K1 launch instead of E matmuls."""
        assert all(e.is_skippable for e in elements), (
            "Combining the elements in a Segment that is not skippable will result in"
            " incorrect tracking results."
        )
        from .segment import compose_transfer_map

        tm = compose_transfer_map(elements, incoming_beam.energy)
        length = sum(e.length.to(tm.device) for e in elements)
        name = "combined_" + "_".join(e.name for e in elements)
        return cls(tm, length=length, name=name)

    def transfer_map(self, energy):
        return self._transfer_map

    @property
    def defining_features(self):
        # custom_transfer_map.py:108-109 lists "transfer_map" only, which loses the length in a
        # saved lattice; "length" is a constructor argument there too, so files stay loadable
        return ["length", "transfer_map"]

    def broadcast(self, shape):
        tm = self._transfer_map
        one = tm.shape[:-2] in ((), (1,))
        new = CustomTransferMap(
            tm.reshape(7, 7).expand(*shape, 7, 7) if one else tm.repeat(*shape, 1, 1),
            length=self.length.reshape(()).expand(shape) if self.length.numel() == 1
            else self.length.repeat(*shape),
            name=self.name,
        )
        return new

    def _encode(self):
        tm = self._transfer_map
        return (L.OP_CUSTOM, [tm[..., r, c] for r in range(6) for c in range(7)], 0)


class Aperture(Element):
    """aperture.py:24-153.  ``track`` for a ParticleBeam is an order-preserving stream
    compaction on the device (``lynx_aperture_compact``): the survivors form the outgoing
    beam, the rest is kept as ``lost_particles`` / ``lost_particle_charges`` exactly as the
    reference's boolean-mask indexing produces them (aperture.py:89-97)."""

    _params = ("x_max", "y_max")

    def __init__(self, x_max=None, y_max=None, shape="rectangular", is_active=True, name=None,
                 device=None, dtype=torch.float32):
        self.x_max = _t(x_max, default=float("inf"))
        self.y_max = _t(y_max, default=float("inf"))
        self.shape = shape
        self.is_active = is_active
        self._lost_beam = None  # SoA planes of the lost particles (device)
        self._lost_aos = None
        self.lost_particle_charges = None
        super().__init__(name)

    @property
    def lost_particles(self):
        """aperture.py:91: ``incoming.particles[~survived_mask]``, shape (n_lost, 7)."""
        if self._lost_aos is None and self._lost_beam is not None:
            self._lost_aos = self._lost_beam.particles.reshape(-1, 7)
        return self._lost_aos

    @property
    def is_skippable(self) -> bool:
        return not self.is_active

    def _encode(self):
        return None  # identity map (aperture.py:63-67); the cut itself is not a map

    def track(self, incoming):
        if incoming is Beam.empty:
            return incoming
        if not (isinstance(incoming, ParticleBeam) and self.is_active):
            return incoming  # aperture.py:71-72
        assert bool((self.x_max >= 0).all()) and bool((self.y_max >= 0).all())
        assert self.shape in ("rectangular", "elliptical"), f"Unknown aperture shape {self.shape}"
        soa = incoming.soa
        dev = soa.device
        if dev.type != "cuda":
            raise L.LynxError("lynx_b200 tracks on a CUDA device only (no CPU fallback)")
        R = 1
        for s_ in incoming.batch_shape:
            R *= s_
        lim_rows = max(self.x_max.numel(), self.y_max.numel())
        if lim_rows not in (1, R):
            raise ValueError(f"Aperture limits have {lim_rows} settings, the beam has {R} rows")
        soa = soa if soa.is_contiguous() else soa.contiguous()
        N, ld = incoming.num_particles, incoming.ld
        soa_r = soa.reshape(R, 6, ld)
        lim_stride = 1 if lim_rows > 1 else 0
        xm = self.x_max.reshape(-1).expand(lim_rows).to(device=dev, dtype=torch.float32).contiguous()
        ym = self.y_max.reshape(-1).expand(lim_rows).to(device=dev, dtype=torch.float32).contiguous()
        pc = incoming.particle_charges
        uniform_pc = pc is not None and pc.stride(-1) == 0
        q_in = (pc.expand(*incoming.batch_shape, N).reshape(R, N).contiguous()
                if (pc is not None and (not uniform_pc or R > 1)) else None)
        kept = torch.empty_like(soa_r)
        lost = torch.empty_like(soa_r)
        q_kept = torch.empty((R, ld), device=dev) if q_in is not None else None
        q_lost = torch.empty((R, ld), device=dev) if q_in is not None else None
        counts = torch.empty(R, dtype=torch.int64, device=dev)
        ws = torch.empty(max(L.lib().lynx_aperture_workspace_bytes(R, N), 16), dtype=torch.uint8,
                         device=dev)
        L.check(L.lib().lynx_aperture_compact(
            soa_r.data_ptr(), R, N, ld, L.ptr(q_in), N if q_in is not None else 0, xm.data_ptr(),
            ym.data_ptr(), lim_stride, 0 if self.shape == "rectangular" else 1, kept.data_ptr(),
            lost.data_ptr(), ld, L.ptr(q_kept), L.ptr(q_lost), counts.data_ptr(), ws.data_ptr(),
            L.current_stream()), "lynx_aperture_compact")
        cnt = counts.tolist()  # the boolean-mask indexing of the reference syncs here too
        if R > 1:
            return self._finish_flattened(incoming, kept, lost, q_kept, q_lost, cnt, N, pc)
        n = cnt[0]
        kept, lost = kept.reshape(soa.shape), lost.reshape(soa.shape)
        bshape = incoming.batch_shape

        def charges(buf, m):
            if pc is None:
                return None
            if uniform_pc:
                return pc[..., :1].expand(*bshape, m)
            return buf[0, :m].reshape(*bshape, m)

        # aperture.py:91-96: lost particles; the reference's (n_lost, 7) AoS view is built
        # on first access of ``lost_particles`` (it is a diagnostic, not part of the track)
        n_lost = N - n
        if n_lost > 0:
            lost[..., n_lost:(n_lost + 3) & ~3] = 0
            self._lost_beam = ParticleBeam.from_soa(lost, n_lost, incoming.energy, None)
            self._lost_aos = None
            self.lost_particle_charges = (
                charges(q_lost, n_lost).reshape(n_lost) if pc is not None else None)
        else:
            self._lost_beam = None
            self._lost_aos = torch.empty((0, 7), device=dev)
            self.lost_particle_charges = torch.empty(0, device=dev) if pc is not None else None
        if n == 0:
            return ParticleBeam.empty  # aperture.py:98-108
        kept[..., n:(n + 3) & ~3] = 0  # padding lanes [N, round4(N)) of the SoA contract
        return ParticleBeam.from_soa(kept, n, incoming.energy, charges(q_kept, n))

    def _finish_flattened(self, incoming, kept, lost, q_kept, q_lost, cnt, N, pc):
        """A beam with R > 1 rows: the reference's ``incoming.particles[survived_mask]``
        (aperture.py:89-97) indexes (R, N, 7) with an (R, N) mask, which FLATTENS the batch -- the
        outgoing beam is one row of all rows' survivors in row order (its energy keeps the
        incoming batch shape, as written there).  The rows were compacted on the device in one
        launch; here they are joined."""
        dev = kept.device

        def join(planes, q, counts):
            total = sum(counts)
            ld = (total + 3) & ~3
            out = torch.zeros((6, ld), device=dev)
            qo = torch.empty(total, device=dev) if q is not None else None
            o = 0
            for r, m in enumerate(counts):
                out[:, o:o + m] = planes[r, :, :m]
                if q is not None:
                    qo[o:o + m] = q[r, :m]
                o += m
            return out, qo, total

        k_soa, k_q, n = join(kept, q_kept, cnt)
        l_soa, l_q, n_lost = join(lost, q_lost, [N - m for m in cnt])
        if n_lost > 0:
            self._lost_beam = ParticleBeam.from_soa(l_soa, n_lost, incoming.energy, None)
            self._lost_aos = None
        else:
            self._lost_beam = None
            self._lost_aos = torch.empty((0, 7), device=dev)
        self.lost_particle_charges = l_q if pc is not None else None
        if n == 0:
            return ParticleBeam.empty
        return ParticleBeam.from_soa(k_soa, n, incoming.energy, k_q)

    def broadcast(self, shape):
        new = self.__class__(x_max=self.x_max.repeat(shape), y_max=self.y_max.repeat(shape),
                             shape=self.shape, is_active=self.is_active, name=self.name)
        new.length = self.length.repeat(shape)
        return new

    @property
    def defining_features(self):
        return ["x_max", "y_max", "shape", "is_active"]


class BPM(Element):
    """bpm.py:22-80."""

    def __init__(self, is_active: bool = False, name=None):
        self.is_active = is_active
        self.reading = None
        super().__init__(name)

    @property
    def is_skippable(self) -> bool:
        return not self.is_active

    def _encode(self):
        return (L.OP_BPM, [], 0) if self.is_active else None


class Screen(Element):
    """screen.py:24-271.  ``reading`` for a ParticleBeam is the fused device histogram."""

    _params = ("misalignment",)

    def __init__(self, resolution=None, pixel_size=None, binning=None, misalignment=None,
                 is_active: bool = False, name=None, device=None, dtype=torch.float32,
                 y_misalignment_fix: bool = False):
        self.resolution = _t(resolution) if resolution is not None else torch.tensor((1024.0, 1024.0))
        self.pixel_size = _t(pixel_size) if pixel_size is not None else torch.tensor((1e-3, 1e-3))
        self.binning = _t(binning) if binning is not None else torch.tensor(1.0)
        self.misalignment = _t(misalignment) if misalignment is not None else torch.tensor((0.0, 0.0))
        self.length = torch.zeros(self.misalignment.shape[:-1])
        self.is_active = is_active
        self.y_misalignment_fix = y_misalignment_fix
        self._read = None
        self.cached_reading = None
        self._edges = None
        super().__init__(name)

    @property
    def is_skippable(self) -> bool:
        return not self.is_active

    @property
    def defining_features(self):  # screen.py:253-261
        return ["resolution", "pixel_size", "binning", "misalignment", "is_active"]

    @property
    def effective_resolution(self):
        return self.resolution / self.binning

    @property
    def effective_pixel_size(self):
        return self.pixel_size * self.binning

    @property
    def extent(self):
        r, p = self.resolution.cpu(), self.pixel_size.cpu()
        return torch.stack([-r[0] * p[0] / 2, r[0] * p[0] / 2, -r[1] * p[1] / 2, r[1] * p[1] / 2])

    @property
    def pixel_bin_edges(self):
        """screen.py:105-118 -- torch CPU ``linspace`` in fp32 (host-side, 4.5k floats),
        the semantic original of the reference's call; uploaded once per geometry."""
        r, p = self.resolution.cpu().float(), self.pixel_size.cpu().float()
        eff = (self.resolution / self.binning).cpu()
        return (
            torch.linspace(-r[0] * p[0] / 2, r[0] * p[0] / 2, int(eff[0]) + 1),
            torch.linspace(-r[1] * p[1] / 2, r[1] * p[1] / 2, int(eff[1]) + 1),
        )

    def _device_edges(self, device):
        key = (tuple(self.resolution.tolist()), tuple(self.pixel_size.tolist()),
               float(self.binning), str(device))
        if self._edges is None or self._edges[0] != key:
            ex, ey = self.pixel_bin_edges
            self._edges = (key, ex.to(device), ey.to(device))
        return self._edges[1], self._edges[2]

    def _encode(self):
        if not self.is_active:
            return None
        m = self.misalignment
        return (L.OP_SCREEN, [m[..., 0], m[..., 1]],
                L.F_SCREEN_Y_FIX if self.y_misalignment_fix else 0)

    def set_read_beam(self, value) -> None:
        self._read = value
        self.cached_reading = None

    def get_read_beam(self):
        r = self._read
        return r() if callable(r) else r

    @property
    def reading(self) -> torch.Tensor:
        """screen.py:143-216: (..., H_eff, W_eff) image."""
        if self.cached_reading is not None:
            return self.cached_reading
        eff = self.effective_resolution
        shape = (*self.misalignment.shape[:-1], int(eff[1]), int(eff[0]))
        if self._read is None or self._read is Beam.empty:
            return torch.zeros(shape)
        beam = self.get_read_beam()
        if isinstance(beam, ParameterBeam):
            self.cached_reading = self._parameter_beam_image(beam)
            return self.cached_reading
        # a ParticleBeam set through set_read_beam: histogram its x, y as they are
        # (screen.py:195-211 reads read_beam.xs / ys; the misalignment was applied by track)
        from .segment import track_elements

        probe = Screen(resolution=self.resolution, pixel_size=self.pixel_size, binning=self.binning,
                       misalignment=torch.zeros_like(self.misalignment), is_active=True,
                       name=f"{self.name}_probe")
        track_elements([probe], beam)
        self.cached_reading = probe.cached_reading
        return self.cached_reading

    def _parameter_beam_image(self, beam: ParameterBeam) -> torch.Tensor:
        """screen.py:157-194 on the device (K7): (..., W_eff, H_eff) as the reference
        produces it (meshgrid "ij", flipped along x -- quirk B8)."""
        mu, cov = beam._mu.contiguous(), beam._cov.contiguous()
        device = mu.device
        ext = self.extent
        hstep = (self.pixel_size[0] * self.binning).cpu()
        vstep = (self.pixel_size[1] * self.binning).cpu()
        xs = torch.arange(ext[0], ext[1], hstep).to(device)  # :187-190 as data
        ys = torch.arange(ext[2], ext[3], vstep).to(device)
        bshape = tuple(mu.shape[:-1])
        B = 1
        for n in bshape:
            B *= n
        img = torch.empty((*bshape, xs.numel(), ys.numel()), dtype=torch.float32, device=device)
        L.check(L.lib().lynx_screen_pdf(mu.data_ptr(), cov.data_ptr(), B, xs.data_ptr(),
                                        ys.data_ptr(), xs.numel(), ys.numel(), img.data_ptr(),
                                        L.current_stream()), "lynx_screen_pdf")
        return img

    def broadcast(self, shape):
        new = super().broadcast(shape)
        new.length = self.length.reshape(()).expand(shape) if self.length.numel() == 1 \
            else self.length.repeat(*shape)
        return new
