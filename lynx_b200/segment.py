"""``Segment`` and the launch glue between the element objects and the C ABI.

Mirror of lynx/accelerator/segment.py:32-368 (container, name registry, subcell,
flattened, transfer_map, track, broadcast, the ``transfer_maps_merged`` family).
Plotting and file I/O are out of scope.
"""

from typing import Optional

import torch

from . import _lib as L
from . import autograd as AG
from .beams import Beam, ParameterBeam, ParticleBeam
from .elements import BPM, Aperture, CustomTransferMap, Drift, Element, Marker, Screen
from .program import Program, broadcast_shapes, encode, pack_params


def _flatten(elements):
    flat = []
    for e in elements:
        if isinstance(e, Segment):
            flat.extend(_flatten(e.elements))
        else:
            flat.append(e)
    return flat


def _numel(shape) -> int:
    n = 1
    for s in shape:
        n *= s
    return n


def _require_cuda(device):
    if device.type != "cuda":
        raise L.LynxError(
            "lynx_b200 tracks on a CUDA device only (sm_100a kernels, no CPU fallback); "
            f"got a beam on {device}"
        )


def _is_shared(t: torch.Tensor, nbatch_dims: int) -> bool:
    """True if the leading batch dims hold one logical row (size 1 or stride 0)."""
    return all(t.shape[d] == 1 or t.stride(d) == 0 for d in range(nbatch_dims))


# --------------------------------------------------------------------------- K1
def compose_transfer_map(elements, energy: torch.Tensor) -> Optional[torch.Tensor]:
    """Segment.transfer_map (segment.py:329-338): product of the elements' maps."""
    flat = _flatten(elements)
    # a lone active cavity: its own linear map (cavity.py:248-325, Cavity.transfer_map) is the
    # one stage K1 emits for it; any other non-skippable element has no map (segment.py:329-338)
    lone_cavity = len(flat) == 1 and type(flat[0]).__name__ == "Cavity"
    if not lone_cavity and not all(e.is_skippable for e in flat):
        return None
    energy = torch.as_tensor(energy, dtype=torch.float32)
    device = energy.device if energy.device.type == "cuda" else torch.device("cuda")
    _require_cuda(device)
    energy = energy.to(device)
    enc = encode(flat, energy.shape)
    prog = Program.get(enc)
    B = enc.B
    bshape = enc.batch_shape
    if prog.n_stages == 0:
        return torch.eye(7, device=device).expand(*bshape, 7, 7).clone()
    params = pack_params(enc, device)
    e_in = energy.expand(bshape).reshape(B).contiguous()
    rec = torch.empty(prog.workspace_bytes(B), dtype=torch.uint8, device=device)
    maps = torch.empty((B, prog.n_stages, 7, 7), dtype=torch.float32, device=device)
    L.check(L.lib().lynx_compose_maps(prog.handle, params.data_ptr(), B, e_in.data_ptr(), None,
                                      0, 0, rec.data_ptr(), maps.data_ptr(), None,
                                      L.current_stream()), "lynx_compose_maps")
    return maps[:, 0].reshape(*bshape, 7, 7)


# --------------------------------------------------------------------------- K2
def launch_track_particles(prog: Program, params, B, soa, shared, ld, N, energy_in, *,
                           want_out, moment_mode, screen: Optional[Screen], batch_shape):
    """One call of ``lynx_track_particles_fwd`` (through ``TrackParticlesFn`` when a
    gradient is wanted); returns the raw device outputs."""
    cfg = AG.TrackCfg(prog, B, shared, ld, N, energy_in, want_out, moment_mode, screen,
                      tuple(batch_shape))
    if AG.needs_grad(params, soa):
        return AG.TrackParticlesFn.apply(params, soa, cfg)
    return AG.track_particles_raw(cfg, params, soa)


def _prep_particle_beam(beam: ParticleBeam, batch_shape):
    soa = beam.soa
    nb = soa.dim() - 2
    if _is_shared(soa, nb):
        return soa[(0,) * nb].contiguous(), True
    return soa.expand(*batch_shape, *soa.shape[-2:]).contiguous(), False


# Encoded lattices of the last few Segment.track calls: (elements kept alive, Encoded, Program,
# packed device parameters).  encode + Program.get + pack_params are ~0.25 ms of Python and one
# H2D copy per call -- ten times the kernels of a 1 M-particle ARES step -- and a tracking loop
# calls track on an unchanged lattice most of the time.  A hit costs one _stamp() per element.
_LATTICE_CACHE = {}
_LATTICE_CACHE_MAX = 16


def _encoded_lattice(flat, energy_shape, beam_batch_shape, device):
    key = (tuple(e._stamp() for e in flat), tuple(energy_shape), tuple(beam_batch_shape), str(device))
    hit = _LATTICE_CACHE.get(key)
    if hit is not None:
        return hit[1], hit[2], hit[3]
    enc = encode(flat, energy_shape)
    bshape = enc.batch_shape
    if bshape != tuple(beam_batch_shape):
        bshape = broadcast_shapes(bshape, beam_batch_shape)
    enc.batch_shape = bshape
    prog = Program.get(enc)
    params = pack_params(enc, device)
    if not params.requires_grad:  # a differentiable packing belongs to its autograd graph
        if len(_LATTICE_CACHE) >= _LATTICE_CACHE_MAX:
            _LATTICE_CACHE.pop(next(iter(_LATTICE_CACHE)))
        _LATTICE_CACHE[key] = (list(flat), enc, prog, params)  # the elements stay alive: ids stay unique
    return enc, prog, params


def track_elements(elements, incoming, *, particles: bool = True, moments: int = 1):
    """Segment.track (segment.py:340-356) for a flat or nested element list.

    :param particles: write the outgoing particles (48 B/particle) or only the fused
        readings / moments (24 B/particle).
    :param moments: 0 none, 1 diagonal + xxp + yyp (all the ParticleBeam API exposes),
        2 all 21 second moments.
    """
    if incoming is Beam.empty:
        return incoming
    flat = _flatten(elements)
    if isinstance(incoming, ParameterBeam):
        return _track_parameter_beam(flat, incoming)
    if not isinstance(incoming, ParticleBeam):
        raise TypeError(f"Parameter incoming is of invalid type {type(incoming)}")
    # an active Aperture changes the number of particles: the lattice is cut there
    # (aperture.py:61-62 is_skippable = not is_active; segment.py:344-351 steps through it)
    cut = next((i for i, e in enumerate(flat) if isinstance(e, Aperture) and e.is_active), None)
    if cut is not None:
        beam = track_elements(flat[:cut], incoming, moments=moments) if cut else incoming
        beam = flat[cut].track(beam)
        return track_elements(flat[cut + 1:], beam, particles=particles, moments=moments)
    device = incoming.soa.device
    _require_cuda(device)
    enc, prog, params = _encoded_lattice(flat, incoming.energy.shape, incoming.batch_shape, device)
    bshape = enc.batch_shape
    B = enc.B
    e_in = incoming.energy.to(device).expand(bshape).reshape(B).contiguous()
    soa, shared = _prep_particle_beam(incoming, bshape)
    N, ld = incoming.num_particles, incoming.ld
    want_out = particles and not prog.terminal
    # behind a terminal screen there is no outgoing beam (Beam.empty, screen.py:139): nobody
    # reads its moments, and the moment-free screen variant of K2 is the fast one
    out, bpm, img, mom, e_out = launch_track_particles(
        prog, params, B, soa, shared, ld, N, e_in, want_out=want_out,
        moment_mode=0 if prog.terminal else moments, screen=enc.screen_element, batch_shape=bshape)
    for slot, el in enumerate(enc.bpm_elements):  # bpm.py:54
        el._bpm_sums = bpm[slot]  # additive over particle shards (sharding.py)
        el.reading = (bpm[slot] / N).to(torch.float32).T.reshape(2, *bshape)
    if enc.screen_element is not None:
        scr = enc.screen_element
        for el in flat[flat.index(scr) + 1:]:  # they see Beam.empty (screen.py:139, bpm.py:50-52)
            if type(el).__name__ == "BPM" and el.is_active:
                el.reading = None
        upto = flat[: flat.index(scr)]
        scr._read = lambda: _beam_at_screen(upto, incoming, scr)
        scr.cached_reading = img
        return Beam.empty
    if not want_out:
        return _MomentsOnly(mom, e_out, N)
    pc = incoming.particle_charges
    return ParticleBeam.from_soa(out, N, e_out, pc, moment_sums=mom)


class _MomentsOnly(ParticleBeam):
    """Outgoing beam of ``track(..., particles=False)``: moments and energy only."""

    def __init__(self, sums, energy, n):
        self._soa, self._n, self.energy, self.particle_charges, self._sums = None, n, energy, None, sums

    @property
    def batch_shape(self):
        return tuple(self.energy.shape)


def _beam_at_screen(upto, incoming, scr: Screen):
    """The beam ``Screen.track`` would have stashed (screen.py:126-139)."""
    beam = track_elements(upto, incoming, moments=0) if upto else incoming
    soa = beam.soa.clone()
    mis = scr.misalignment.to(soa.device)
    soa[..., 0, :] -= mis[..., 0].reshape(*mis.shape[:-1], 1)
    col = 2 if scr.y_misalignment_fix else 1  # quirk B1: y offset lands on xp as written
    soa[..., col, :] -= mis[..., 1].reshape(*mis.shape[:-1], 1)
    return ParticleBeam.from_soa(soa, beam.num_particles, beam.energy, beam.particle_charges)


def beam_moment_sums(beam: ParticleBeam) -> torch.Tensor:
    """One streaming pass (24 B/particle): all 28 raw moment sums of a beam."""
    bshape = beam.batch_shape
    B = max(_numel(bshape), 1)
    device = beam.soa.device
    _require_cuda(device)
    enc = encode([], ())
    enc.batch_shape = bshape
    prog = Program.get(enc)
    params = pack_params(enc, device)
    e_in = torch.as_tensor(beam.energy, device=device, dtype=torch.float32).expand(bshape).reshape(B).contiguous()
    soa, shared = _prep_particle_beam(beam, bshape)
    _, _, _, mom, _ = launch_track_particles(
        prog, params, B, soa, shared, beam.ld, beam.num_particles, e_in,
        want_out=False, moment_mode=2, screen=None, batch_shape=bshape)
    return mom


# --------------------------------------------------------------------------- K6
def _track_parameter_beam(flat, incoming: ParameterBeam):
    device = incoming._mu.device
    _require_cuda(device)
    enc, prog, params = _encoded_lattice(flat, incoming.energy.shape,
                                         tuple(incoming._mu.shape[:-1]), device)
    bshape = enc.batch_shape
    B = enc.B
    e_in = incoming.energy.to(device).expand(bshape).reshape(B).contiguous()
    mu, cov = incoming._mu, incoming._cov
    nb = mu.dim() - 1
    shared = _is_shared(mu, nb) and _is_shared(cov, nb)
    if shared:
        mu_in, cov_in = mu[(0,) * nb].contiguous(), cov[(0,) * nb].contiguous()
    else:
        mu_in = mu.expand(*bshape, 7).contiguous()
        cov_in = cov.expand(*bshape, 7, 7).contiguous()
    cfg = AG.ParamBeamCfg(prog, B, shared, e_in, tuple(bshape))
    if AG.needs_grad(params, mu_in, cov_in):
        mu_out, cov_out, bpm, e_out = AG.ParamBeamFn.apply(params, mu_in, cov_in, cfg)
    else:
        mu_out, cov_out, bpm, e_out = AG.parambeam_raw(cfg, params, mu_in, cov_in)
    for slot, el in enumerate(enc.bpm_elements):
        el.reading = bpm[slot].T.reshape(2, *bshape)
    out = ParameterBeam.__new__(ParameterBeam)
    out._mu, out._cov, out.energy = mu_out, cov_out, e_out
    out.total_charge = incoming.total_charge
    if enc.screen_element is not None:
        enc.screen_element.set_read_beam(out)
        return Beam.empty
    return out


def track_parameter_beam_along(flat, incoming: ParameterBeam):
    """``lynx_parambeam_along``: the beam after EVERY element of ``flat`` (each of which must
    encode to one op).  Returns ``(mu (E, *batch, 7), cov (E, *batch, 7, 7), energy (E, *batch))``."""
    device = incoming._mu.device
    _require_cuda(device)
    enc = encode(flat, incoming.energy.shape)
    assert len(enc.ops) == len(flat), "every element must encode to exactly one op"
    bshape = broadcast_shapes(enc.batch_shape, tuple(incoming._mu.shape[:-1]))
    enc.batch_shape = bshape
    B = enc.B
    prog = Program.get(enc)
    params = pack_params(enc, device)
    e_in = incoming.energy.to(device).expand(bshape).reshape(B).contiguous()
    mu, cov = incoming._mu, incoming._cov
    nb = mu.dim() - 1
    shared = _is_shared(mu, nb) and _is_shared(cov, nb)
    if shared:
        mu_in, cov_in = mu[(0,) * nb].contiguous(), cov[(0,) * nb].contiguous()
    else:
        mu_in = mu.expand(*bshape, 7).contiguous()
        cov_in = cov.expand(*bshape, 7, 7).contiguous()
    E = len(flat)
    s_mu = torch.empty((E, *bshape, 7), device=device)
    s_cov = torch.empty((E, *bshape, 7, 7), device=device)
    s_en = torch.empty((E, *bshape), device=device)
    L.check(L.lib().lynx_parambeam_along(
        prog.handle, params.data_ptr(), B, mu_in.data_ptr(), cov_in.data_ptr(),
        0 if shared else 1, e_in.data_ptr(), s_mu.data_ptr(), s_cov.data_ptr(), s_en.data_ptr(),
        L.current_stream()), "lynx_parambeam_along")
    return s_mu, s_cov, s_en


# --------------------------------------------------------------------------- Segment
class Segment(Element):
    """segment.py:32-368."""

    def __init__(self, elements, name: Optional[str] = None) -> None:
        super().__init__(name=name)
        self.elements = list(elements)
        for element in self.elements:  # name registry, segment.py:45-54
            if element.name in self.__dict__:
                cur = self.__dict__[element.name]
                if isinstance(cur, list):
                    cur.append(element)
                else:
                    self.__dict__[element.name] = [cur, element]
            else:
                self.__dict__[element.name] = element

    def subcell(self, start: str, end: str) -> "Segment":
        sub, inside = [], False
        for element in self.elements:
            if element.name == start:
                inside = True
            if inside:
                sub.append(element)
            if element.name == end:
                break
        return self.__class__(sub)

    @classmethod
    def from_lattice_json(cls, filepath: str) -> "Segment":
        """segment.py:220-228: load a LatticeJSON file (``lynx_b200/latticejson.py``)."""
        from . import latticejson

        return latticejson.load(filepath)

    def to_lattice_json(self, filepath: str, title: Optional[str] = None,
                        info: str = "This is a placeholder lattice description") -> None:
        """segment.py:230-247."""
        from . import latticejson

        latticejson.save(self, filepath, title, info)

    @classmethod
    def from_ocelot(cls, cell, name: Optional[str] = None, warnings: bool = True, device=None,
                    dtype=torch.float32, **kwargs) -> "Segment":
        """segment.py:248-279: list of Ocelot elements (``lynx_b200/ocelot.py``)."""
        from . import ocelot

        return cls([ocelot.convert_element(e, warnings=warnings) for e in cell], name=name, **kwargs)

    @classmethod
    def from_bmad(cls, bmad_lattice_file_path: str, environment_variables=None) -> "Segment":
        """segment.py:281-299: Bmad lattice file (``lynx_b200/bmad.py``)."""
        from . import bmad

        return bmad.read_bmad_lattice(bmad_lattice_file_path, environment_variables)

    @classmethod
    def from_nx_tables(cls, filepath) -> "Segment":
        """segment.py:301-313: ARES NX-tables component list (``lynx_b200/nxtables.py``)."""
        from . import nxtables

        return nxtables.read_nx_tables(filepath)

    def flattened(self) -> "Segment":
        return Segment(_flatten(self.elements), name=self.name)

    @property
    def is_skippable(self) -> bool:
        return all(e.is_skippable for e in self.elements)

    @property
    def length(self) -> torch.Tensor:
        """segment.py:321-327: the element lengths stacked and summed in one reduction (a
        running fp32 sum drifts by 3e-5 m over the 226 elements of the ARES NX table)."""
        if not self.elements:
            return torch.zeros(())
        lengths = torch.broadcast_tensors(*[e.length.cpu() for e in self.elements])
        return torch.stack(lengths, dim=-1).sum(dim=-1)

    @length.setter
    def length(self, value):  # Element.__init__ probes the attribute
        pass

    def transfer_map(self, energy):
        return compose_transfer_map(self.elements, energy)

    def track(self, incoming, **kw):
        return track_elements(self.elements, incoming, **kw)

    def broadcast(self, shape):
        return self.__class__([e.broadcast(shape) for e in self.elements], name=self.name)

    def split(self, resolution):
        return [s for e in self.elements for s in e.split(resolution)]

    def _encode(self):
        raise TypeError("nested Segments are flattened before encoding")

    # ---- speed optimisations of the reference (segment.py:84-218)
    def transfer_maps_merged(self, incoming_beam, except_for=None) -> "Segment":
        except_for = except_for or []
        merged, run = [], []
        beam = incoming_beam

        def close(beam):
            if len(run) == 1:
                merged.append(run[0])
                beam = run[0].track(beam)
            elif len(run) > 1:
                merged.append(CustomTransferMap.from_merging_elements(list(run), beam))
                beam = merged[-1].track(beam)
            run.clear()
            return beam

        for element in self.elements:
            if element.is_skippable and element.name not in except_for:
                run.append(element)
            else:
                beam = close(beam)
                merged.append(element)
                beam = element.track(beam)
        if run:
            merged.append(CustomTransferMap.from_merging_elements(list(run), beam))
        return Segment(merged, name=self.name)

    def without_inactive_markers(self, except_for=None) -> "Segment":
        except_for = except_for or []
        return Segment([e for e in self.elements
                        if not isinstance(e, Marker) or e.name in except_for], name=self.name)

    def without_inactive_zero_length_elements(self, except_for=None) -> "Segment":
        except_for = except_for or []
        return Segment([e for e in self.elements
                        if bool((e.length > 0.0).any()) or getattr(e, "is_active", False)
                        or e.name in except_for], name=self.name)

    def inactive_elements_as_drifts(self, except_for=None) -> "Segment":
        except_for = except_for or []
        return Segment([
            e if getattr(e, "is_active", False) or bool((e.length == 0.0).all())
            or e.name in except_for else Drift(e.length)
            for e in self.elements], name=self.name)

    def twiss_along(self, beam):
        """The data of ``plot_twiss`` (segment.py:490-518): the beam after every element of
        non-zero length, tracked element by element.  Returns ``(s_positions, beams)`` with
        ``beams[0]`` the incoming beam.  A ParameterBeam takes ONE launch for the whole
        lattice (``lynx_parambeam_along``, a snapshot per element); a ParticleBeam is stepped
        element by element like the reference."""
        flat = [e for e in _flatten(self.elements) if bool((e.length != 0).any())]
        s_positions, beams = [0.0], [beam]
        for e in flat:
            s_positions.append(s_positions[-1] + float(e.length.reshape(-1)[0]))
        if isinstance(beam, ParameterBeam) and flat:
            mu, cov, en = track_parameter_beam_along(flat, beam)
            for i in range(len(flat)):
                b = ParameterBeam.__new__(ParameterBeam)
                b._mu, b._cov, b.energy, b.total_charge = mu[i], cov[i], en[i], beam.total_charge
                beams.append(b)
        else:
            for e in flat:
                beams.append(e.track(beams[-1]))
        return s_positions, beams

    def plot_twiss(self, beam, ax=None):
        """segment.py:490-518.  Returns ``(s_positions, beta_x, beta_y)``; draws them when an
        axis is passed (matplotlib itself is outside this package's scope)."""
        s_positions, beams = self.twiss_along(beam)
        beta_x = [float(b.beta_x.reshape(-1)[0]) for b in beams]
        beta_y = [float(b.beta_y.reshape(-1)[0]) for b in beams]
        if ax is not None:
            ax.set_title("Twiss Parameters")
            ax.set_xlabel("s (m)")
            ax.set_ylabel(r"$\beta$ (m)")
            ax.plot(s_positions, beta_x, label=r"$\beta_x$", c="tab:red")
            ax.plot(s_positions, beta_y, label=r"$\beta_y$", c="tab:green")
            ax.legend()
        return s_positions, beta_x, beta_y

    def __repr__(self) -> str:
        return f"Segment(elements={self.elements!r}, name={self.name!r})"
