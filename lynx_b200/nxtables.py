"""ARES "NX tables" component lists (CSV) to a ``Segment`` (converters/nxtables.py:10-265).

Host logic only; one of the lattice data formats in front of the tracking path.  The file
lists every installed component with its device class and the longitudinal position of its
centre; components the model knows become elements, the gaps between them become drifts.
The device-class table below restates the reference's facility data (class -> element and
its fixed parameters); classes it ignores are ignored here, unknown ones are an error.
"""

import csv
from pathlib import Path

import torch

_CAM_A = {"resolution": (2448, 2040), "pixel_size": (3.5488e-6, 2.5003e-6)}  # nxtables.py:70-99
_CAM_B = {"resolution": (2464, 2056), "pixel_size": (0.00998e-3, 0.00715e-3)}  # :100-113
_OPEN = {"x_max": float("inf"), "y_max": float("inf")}  # sizes not known yet (:118-138)

# device class -> (element class, fixed parameters)
DEVICE_CLASSES = {
    "BSCX": ("Screen", {"resolution": (2464, 2056), "pixel_size": (0.00343e-3, 0.00247e-3)}),  # :63-69
    "BSCR": ("Screen", _CAM_A), "BSCM": ("Screen", _CAM_A), "BSCO": ("Screen", _CAM_A),
    "BSCA": ("Screen", _CAM_A), "BSCE": ("Screen", _CAM_B), "SCRD": ("Screen", _CAM_B),
    "BPMG": ("BPM", {}), "BPML": ("BPM", {}),  # :114-117
    "SLHG": ("Aperture", dict(_OPEN, shape="elliptical")),
    "SLHB": ("Aperture", dict(_OPEN, shape="rectangular")),
    "SLHS": ("Aperture", dict(_OPEN, shape="rectangular")),
    "MCHM": ("HorizontalCorrector", {"length": 0.02}),  # :139-142
    "MCVM": ("VerticalCorrector", {"length": 0.02}),
    "MBHL": ("Dipole", {"length": 0.322}), "MBHB": ("Dipole", {"length": 0.22}),  # :143-146
    "MBHO": ("Dipole", {"length": 0.43852543421396856, "angle": 0.8203047484373349,
                        "e2": -0.7504915783575616}),  # :147-153
    "MQZM": ("Quadrupole", {"length": 0.122}),  # :154-155
    "RSBL": ("Cavity", {"length": 4.139, "frequency": 2.998e9, "voltage": 76e6}),  # :156-162
    "RXBD": ("Cavity", {"length": 1.0, "frequency": 11.9952e9, "voltage": 0.0}),  # :163-169
    "UNDA": ("Undulator", {"length": 0.25}),  # :170-171
}
# positions worth keeping by name, no optics (nxtables.py:172-203)
MARKER_CLASSES = frozenset(
    "SOLG BCMG EOLG SOLS EOLS SOLA EOLA SOLT BSTB TORF EOLT SOLO EOLO SOLB EOLB ECHA MKBB MKBE "
    "MKPM EOLC SOLM EOLM SOLH BSCD STDE ECHS EOLH WINA LINA EOLX".split())
# components without a counterpart in the model (nxtables.py:25-51)
IGNORED_CLASSES = frozenset(
    "RSBG MSOB MSOH MSOG VVAG BSCL MIRA BAML SCRL TEMG FCNG SOLE EOLE MSOL BELS VVAF MIRM SCRY "
    "FPSA VPUL SOLC SCRE SOLX ICTB BSCS".split())
_COIL_LENGTH = 5e-05  # one coil of a combined corrector (nxtables.py:53-62)


def _build(cls_name, name, fixed):
    import lynx_b200 as lx

    kw = {}
    for key, value in fixed.items():
        if isinstance(value, str):
            kw[key] = value
        elif isinstance(value, tuple):
            kw[key] = torch.tensor(value)
        else:
            kw[key] = torch.tensor([value], dtype=torch.float32)
    if cls_name == "Screen":
        kw["binning"] = 1
    return getattr(lx, cls_name)(name=name, **kw)


def translate(device_class: str, name: str):
    """Element (or two-coil ``Segment``) for one row, ``None`` for an ignored component."""
    import lynx_b200 as lx

    if device_class in IGNORED_CLASSES:
        return None
    if device_class == "MCXG":  # combined corrector: ....X. -> ....H. and ....V.
        assert name[6] == "X"
        coil = torch.tensor([_COIL_LENGTH])
        return lx.Segment(elements=[
            lx.HorizontalCorrector(name=name[:6] + "H" + name[7:], length=coil.clone()),
            lx.VerticalCorrector(name=name[:6] + "V" + name[7:], length=coil.clone()),
        ], name=name)
    if device_class in DEVICE_CLASSES:
        cls_name, fixed = DEVICE_CLASSES[device_class]
        return _build(cls_name, name, fixed)
    if device_class in MARKER_CLASSES:
        return lx.Marker(name=name)
    raise ValueError(f"Encountered unknown class {device_class} for element {name}")


def read_nx_tables(filepath):
    """``Segment`` (flat) named after the file: elements in order of ``Z_beam`` with the gaps
    between neighbouring components filled by drifts (nxtables.py:211-265)."""
    import lynx_b200 as lx

    filepath = Path(filepath)
    with open(filepath, "r", newline="") as f:
        rows = list(csv.reader(f, delimiter=","))
    header, rows = rows[0], rows[1:]
    i_class, i_name, i_s = header.index("CLASS"), header.index("NAME"), header.index("Z_beam")
    placed = []
    for row in rows:
        element = translate(row[i_class], row[i_name])
        if element is not None:
            placed.append((float(row[i_s]), element))
    placed.sort(key=lambda item: item[0])  # stable: equal positions keep the file's order

    out = [placed[0][1]]
    for (s_prev, prev), (s_cur, cur) in zip(placed[:-1], placed[1:]):
        gap = (s_cur - s_prev) - prev.length / 2 - cur.length / 2  # centre to centre, fp32
        assert bool((gap >= 0.0).all()), f"Elements {prev.name} and {cur.name} overlap by {gap}."
        if bool((gap > 0.0).all()):
            out.append(lx.Drift(name=f"DRIFT_{prev.name}_{cur.name}", length=gap.reshape(1)))
        out.append(cur)
    return lx.Segment(elements=out, name=filepath.stem).flattened()
