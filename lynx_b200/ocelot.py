"""Ocelot cells (lists of Ocelot elements) to a ``Segment`` (converters/ocelot.py:7-200).

Host logic only; a lattice data format in front of the tracking path.  Ocelot itself is not
imported: an element is recognised by the class names in its MRO, so real Ocelot objects and
any objects with the same class names and attributes convert alike.  The order of the rules
is the reference's: ``Bend`` comes before ``SBend`` / ``RBend``, which derive from it in
Ocelot, so every bend becomes a ``Dipole`` with its edge angles as given (the reference's
``RBend`` branch, ``converters/ocelot.py:100-114``, is unreachable for the same reason).
"""

import warnings as _warnings

import torch

# the default camera of an ARES screen station (converters/ocelot.py:146-152)
SCREEN_DEFAULTS = {"resolution": (2448, 2040), "pixel_size": (3.5488e-6, 2.5003e-6)}
_APERTURE_SHAPES = {"rect": "rectangular", "elip": "elliptical"}


def _t(value):
    return torch.tensor([value], dtype=torch.float32)


def convert_element(element, warnings: bool = True):
    import lynx_b200 as lx

    kinds = {c.__name__ for c in type(element).__mro__}
    name = element.id
    if "Drift" in kinds:
        return lx.Drift(length=_t(element.l), name=name)
    if "Quadrupole" in kinds:
        return lx.Quadrupole(length=_t(element.l), k1=_t(element.k1), name=name)
    if "Solenoid" in kinds:
        return lx.Solenoid(length=_t(element.l), k=_t(element.k), name=name)
    if "Hcor" in kinds:
        return lx.HorizontalCorrector(length=_t(element.l), angle=_t(element.angle), name=name)
    if "Vcor" in kinds:
        return lx.VerticalCorrector(length=_t(element.l), angle=_t(element.angle), name=name)
    if kinds & {"Bend", "SBend", "RBend"}:
        return lx.Dipole(length=_t(element.l), angle=_t(element.angle), e1=_t(element.e1),
                         e2=_t(element.e2), tilt=_t(element.tilt),
                         fringe_integral=_t(element.fint), fringe_integral_exit=_t(element.fintx),
                         gap=_t(element.gap), name=name)
    if kinds & {"Cavity", "TDCavity"}:  # Ocelot keeps the voltage in GV
        return lx.Cavity(length=_t(element.l), voltage=_t(element.v) * 1e9,
                         frequency=_t(element.freq), phase=_t(element.phi), name=name)
    if "Monitor" in kinds:
        if "BSC" in name:  # ARES naming: a screen station
            if warnings:
                _warnings.warn("Diagnostic screen was converted with default screen properties.")
            return lx.Screen(resolution=torch.tensor(SCREEN_DEFAULTS["resolution"]),
                             pixel_size=torch.tensor(SCREEN_DEFAULTS["pixel_size"]), name=name)
        return lx.BPM(name=name) if "BPM" in name else lx.Marker(name=name)
    if "Marker" in kinds:
        return lx.Marker(name=name)
    if "Undulator" in kinds:
        return lx.Undulator(length=_t(element.l), name=name)
    if "Aperture" in kinds:
        return lx.Aperture(x_max=_t(element.xmax), y_max=_t(element.ymax),
                           shape=_APERTURE_SHAPES[element.type], is_active=True, name=name)
    if warnings:
        _warnings.warn(f"Unknown element {name} of type {type(element)}, replacing with drift"
                       " section.")
    return lx.Drift(length=_t(element.l), name=name)


def subcell(cell: list, start: str, end: str) -> list:
    """Elements from the one whose id is ``start`` up to and including ``end``
    (converters/ocelot.py:188-200)."""
    out, inside = [], False
    for element in cell:
        inside = inside or element.id == start
        if inside:
            out.append(element)
        if element.id == end:
            break
    return out
