"""LatticeJSON files (https://github.com/nobeam/latticejson) to and from ``Segment``.

Same on-disk format as the reference's ``latticejson.py:66-189`` ("cheetah-0.6": ``root``,
``elements`` = name -> [class, {feature: value}], ``lattices`` = name -> [member names], nested
segments as further lattices), so the reference's own files
(``docs/examples/ARESlatticeStage3v1_9.json``) load unchanged and files written here load in
the reference.  Host logic only -- this is one of the data formats either side of the tracking
path, not part of it.
"""

import json
from typing import Optional

import torch

VERSION = "cheetah-0.6"


_COUNTS = ("resolution", "binning")  # pixel counts: integer tensors in the reference's files


def _plain(value, feature=""):
    """Tensor features become (nested) lists, everything else is kept (latticejson.py:9-19)."""
    if not isinstance(value, torch.Tensor):
        return value
    value = value.detach().cpu()
    if feature in _COUNTS and bool((value == value.round()).all()):
        value = value.to(torch.int64)
    return value.tolist()


def _features(element) -> dict:
    out = {}
    for feature in element.defining_features:
        # CustomTransferMap keeps its matrix as ``_transfer_map`` (``transfer_map`` is the
        # method every element has); the file stores it under the constructor's name
        attr = "_transfer_map" if feature == "transfer_map" else feature
        out[feature] = _plain(getattr(element, attr), feature)
    return out


def segment_to_dict(segment, title: Optional[str] = None,
                    info: str = "This is a placeholder lattice description") -> dict:
    """The LatticeJSON dictionary of ``segment`` (latticejson.py:37-104)."""
    from .segment import Segment

    elements, lattices = {}, {}

    def walk(seg):
        members = []
        for el in seg.elements:
            if isinstance(el, Segment):
                walk(el)
            else:
                elements[el.name] = [el.__class__.__name__, _features(el)]
            members.append(el.name)
        lattices[seg.name] = members

    walk(segment)
    if title is None:
        title = segment.name if segment.name is not None else "Unnamed Lattice"
    return {
        "version": VERSION,
        "title": title,
        "info": info,
        "root": segment.name if segment.name is not None else "cell",
        "elements": elements,
        "lattices": lattices,
    }


def dumps(lattice: dict) -> str:
    """One line per element / lattice, as LatticeJSON files are usually laid out (the two
    outer levels are indented, everything below stays on its line)."""
    lines = ["{"]
    keys = list(lattice)
    for i, key in enumerate(keys):
        value = lattice[key]
        comma = "," if i + 1 < len(keys) else ""
        if isinstance(value, dict):
            lines.append(f"    {json.dumps(key)}: {{")
            inner = list(value.items())
            for j, (k, v) in enumerate(inner):
                sep = "," if j + 1 < len(inner) else ""
                lines.append(f"        {json.dumps(k)}: {json.dumps(v)}{sep}")
            lines.append(f"    }}{comma}")
        else:
            lines.append(f"    {json.dumps(key)}: {json.dumps(value)}{comma}")
    lines.append("}")
    return "\n".join(lines) + "\n"


def save(segment, filepath: str, title: Optional[str] = None,
         info: str = "This is a placeholder lattice description") -> None:
    with open(filepath, "w") as f:
        f.write(dumps(segment_to_dict(segment, title, info)))


def _feature(value):
    """latticejson.py:128-138: strings and bools stay, numbers and lists become tensors."""
    return value if isinstance(value, (str, bool)) else torch.tensor(value, dtype=torch.float32)


def segment_from_dict(lattice: dict, name: Optional[str] = None):
    """``Segment`` named ``name`` (default: the file's root) (latticejson.py:141-189)."""
    import lynx_b200 as lx

    def build(seg_name):
        members = []
        for member in lattice["lattices"][seg_name]:
            if member in lattice["lattices"]:
                members.append(build(member))
                continue
            cls_name, params = lattice["elements"][member]
            cls = getattr(lx, cls_name, None)
            if cls is None or not isinstance(cls, type) or not issubclass(cls, lx.Element):
                raise ValueError(f"{member}: unknown element class {cls_name!r}")
            members.append(cls(name=member, **{k: _feature(v) for k, v in params.items()}))
        return lx.Segment(elements=members, name=seg_name)

    return build(lattice["root"] if name is None else name)


def load(filepath: str):
    with open(filepath, "r") as f:
        return segment_from_dict(json.load(f))
